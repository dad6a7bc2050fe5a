"""GPU test of ``sigsvgd_median_bandwidth_static`` (the entry point for callers that may not allocate:
its histogram scratch is a file-scope device array inside the library) through the C ABI, against the
existing workspace form ``sigsvgd_median_bandwidth`` on the same arrays, on two calls in a row.

Collected last on purpose (``test_zz_``).  Runs alone in a fresh
process as well as after the other tests (no state is shared with them but the library itself)."""
import numpy as np
import pytest
import torch

from tests.util import TOL_H

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    from sigsvgd_b200 import _lib, ops

DEV = "cuda:0"
MAX_BATCH = 16  # SIGSVGD_STATIC_MAX_BATCH (include/sigsvgd_b200.h)


def _arrays(batch, count, stride, seed):
    """`batch` arrays of squared-distance-like floats (>= 0, repeated values, a few exact zeros)."""
    g = torch.Generator().manual_seed(seed)
    a = torch.full((batch, stride), -1.0)  # (the padding between the arrays must never be read)
    v = (torch.randn(batch, count, generator=g) ** 2) * (1.0 + torch.arange(batch)[:, None])
    v[:, ::97] = 0.0
    v[:, 1::5] = v[:, 0:1]
    a[:, :count] = v
    return a.to(DEV)


def _static(dist, count, n_log):
    batch, stride = dist.shape
    state = ops.new_state(torch.device(DEV), batch)
    ops.call("sigsvgd_median_bandwidth_static", ops._p(dist), count, batch, stride, n_log, 1.0, 1e-8, ops._p(state),
             ops._stream())
    return [ops.read_state(state, b) for b in range(batch)]


def _workspace_form(dist, count, n_log):
    """The existing entry point, one array at a time, with a caller-allocated workspace."""
    out = []
    nbytes = int(_lib.load().sigsvgd_select_workspace_bytes(count))
    for b in range(dist.shape[0]):
        state = ops.new_state(torch.device(DEV))
        ws = torch.empty(nbytes, dtype=torch.uint8, device=DEV)
        row = dist[b, :count].contiguous()
        ops.call("sigsvgd_median_bandwidth", ops._p(row), count, n_log, 1.0, 1e-8, ops._p(state), ops._p(ws), nbytes,
                 ops._stream())
        out.append(ops.read_state(state))
    return out


def _compare(dist, count, n_log):
    got, want = _static(dist, count, n_log), _workspace_form(dist, count, n_log)
    host = dist[:, :count].cpu().numpy()
    for b, (g, w) in enumerate(zip(got, want)):
        # exact order statistic: the same float, and torch.median's lower-median convention (math.py:28-34)
        assert g["median"] == w["median"] == float(np.sort(host[b])[(count - 1) // 2])
        assert abs(g["h"] - w["h"]) <= TOL_H * w["h"] and g["clamped"] == w["clamped"]
        assert abs(g["gamma"] - w["gamma"]) <= 2 * TOL_H * abs(w["gamma"])


def test_static_scratch_median_matches_workspace_form_twice_in_a_row():
    # two calls in a row on different data and different batch sizes: the second one finds the first
    # one's histograms in the shared array and must clear what it uses
    _compare(_arrays(3, 70001, 70016, seed=1), 70001, 300)
    _compare(_arrays(MAX_BATCH, 12345, 12352, seed=2), 12345, 111)
    _compare(_arrays(3, 70001, 70016, seed=1), 70001, 300)
    _compare(_arrays(1, 1, 4, seed=3), 1, 1)  # one element; stride is ignored for one array
    _compare(_arrays(2, 2, 2, seed=4), 2, 2)  # lower median of two = the smaller


def test_static_scratch_call_allocates_nothing_and_rejects_oversize():
    dist = _arrays(4, 50000, 50000, seed=5)
    state = ops.new_state(torch.device(DEV), 4)
    args = ("sigsvgd_median_bandwidth_static", ops._p(dist), 50000, 4, 50000, 200, 1.0, 1e-8, ops._p(state),
            ops._stream())
    ops.call(*args)  # (first call on this device: symbol lookup, lazy module load)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(torch.device(DEV))[0]
    for _ in range(3):
        ops.call(*args)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(torch.device(DEV))[0] == free0  # no cudaMalloc behind the call
    big = torch.zeros(MAX_BATCH + 1, 8, device=DEV)
    big_state = ops.new_state(torch.device(DEV), MAX_BATCH + 1)
    with pytest.raises(RuntimeError, match="SIGSVGD_STATIC_MAX_BATCH"):
        ops.call("sigsvgd_median_bandwidth_static", ops._p(big), 8, MAX_BATCH + 1, 8, 8, 1.0, 1e-8, ops._p(big_state),
                 ops._stream())
    with pytest.raises(ValueError, match="stride"):
        ops.call("sigsvgd_median_bandwidth_static", ops._p(big), 8, 2, 4, 8, 1.0, 1e-8, ops._p(big_state), ops._stream())
    # the failed calls left nothing behind: the next good call is still right
    _compare(_arrays(2, 999, 1000, seed=6), 999, 30)

"""world_size-2 ``gloo`` tests (CPU) of the multi-GPU orchestration in sigsvgd_b200/dist.py:
row-block sharding, the all-gather exchange, the histogram all-reduces of the exact GLOBAL
median, the tie / sum_kd reduction and the row offsets.  The arithmetic behind the backend
interface is the CPU checker backend (tests/checker_backend.py); the same orchestration code runs
with CudaBackend over NCCL on the GPU box (tests/test_gpu_dist.py)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from sigsvgd_b200.dist import phi_sharded, shard_range, shard_scenes


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _inputs(N, T, d, seed=0):
    g = torch.Generator().manual_seed(seed)
    X = torch.cumsum(0.1 * torch.randn(N, T, d, generator=g), 1).contiguous()
    G = torch.randn(N, T * d, generator=g)
    return X, G


def _worker(rank, world, port, N, T, d, depth, fixed_h, q, window=None):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from sigsvgd_b200 import ops
        from tests.checker_backend import CheckerBackend

        X, G = _inputs(N, T, d)
        M = N // world
        bw = ops.Bandwidth.fixed(fixed_h) if fixed_h else ops.Bandwidth.median()
        res = phi_sharded(X[rank * M:(rank + 1) * M], G[rank * M:(rank + 1) * M], depth=depth, bandwidth=bw,
                          want_k=True, backend=CheckerBackend(), window=window)
        q.put((rank, res.phi.numpy(), res.k_rows.numpy(), dict(res.state), res.rows))
    finally:
        dist.destroy_process_group()


def _run(world, N, T, d, depth, fixed_h=None, window=None):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, N, T, d, depth, fixed_h, q, window)) for r in range(world)]
    for p in procs:
        p.start()
    got = []
    import queue as _queue
    import time as _time

    deadline = _time.time() + 240
    while len(got) < world and _time.time() < deadline:
        try:
            got.append(q.get(timeout=1.0))
        except _queue.Empty:
            if any(p.exitcode not in (None, 0) for p in procs):
                for p in procs:
                    p.terminate()
                pytest.fail("a rank died: " + str([p.exitcode for p in procs]))
    assert len(got) == world, "ranks timed out"
    got.sort(key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    return got


@pytest.mark.parametrize("N,T,d,depth,window", [(24, 8, 2, 3, None), (30, 6, 3, 2, None), (24, 8, 2, 3, True)])
def test_sharded_update_matches_single_problem_oracle(N, T, d, depth, window):
    import oracle
    from tests.util import closed_form_with_ties

    got = _run(2, N, T, d, depth, window=window)
    X, G = _inputs(N, T, d)
    states = [g[3] for g in got]
    # identical global median / bandwidth on every rank, equal to the exact lower median
    assert states[0]["median"] == states[1]["median"] and states[0]["h"] == states[1]["h"]
    S = oracle.signature_np(X.numpy(), depth, True)
    Sc = (S - S.mean(0)).astype(np.float64)
    sq = oracle.pw_dist_sq(Sc, Sc)
    med = oracle.lower_median(sq)
    assert abs(states[0]["median"] - med) <= 2e-6 * med
    # ties found by the two ranks are disjoint row blocks and their count was all-reduced
    ties = [t for s in states for t in s["ties"]]
    assert states[0]["n_ties_all"] == states[1]["n_ties_all"] == len(ties) >= 1
    assert states[0]["coef"] == pytest.approx(states[1]["coef"], rel=1e-12)
    cf = closed_form_with_ties(X.double().numpy(), G.double().numpy(), depth, ties)
    phi = np.concatenate([g[1] for g in got], 0)
    K = np.concatenate([g[2] for g in got], 0)
    assert [g[4] for g in got] == [(0, N // 2), (N // 2, N)]
    assert np.abs(K - cf["K"]).max() < 1e-4
    assert np.abs(phi - cf["phi"]).max() / np.abs(cf["phi"]).max() < 1e-4


def test_sharded_update_fixed_bandwidth():
    from tests.util import closed_form_with_ties

    N, T, d, depth = 20, 6, 2, 3
    got = _run(2, N, T, d, depth, fixed_h=0.8)
    X, G = _inputs(N, T, d)
    cf = closed_form_with_ties(X.double().numpy(), G.double().numpy(), depth, [], h=0.8)
    phi = np.concatenate([g[1] for g in got], 0)
    assert np.abs(phi - cf["phi"]).max() / np.abs(cf["phi"]).max() < 1e-4
    assert all(g[3]["coef"] == 0.0 for g in got)


def test_shard_helpers():
    assert [shard_range(10, r, 4) for r in range(4)] == [(0, 3), (3, 6), (6, 8), (8, 10)]
    assert [shard_range(8, r, 8) for r in range(8)] == [(r, r + 1) for r in range(8)]
    assert shard_range(3, 3, 4) == (3, 3)  # empty shard
    cover = [i for r in range(8) for i in shard_scenes(1024, r, 8)]
    assert cover == list(range(1024)) and len(shard_scenes(1024, 0, 8)) == 128

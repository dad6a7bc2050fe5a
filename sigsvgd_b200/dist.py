"""Multi-GPU forms of the Stein update: one process per GPU, ``torch.distributed`` (NCCL over
NVLink) for the plumbing.  The reference has no distributed code at all (its only parallelism is
``mp.Pool`` over independent planning problems, examples/script_planning_robot.py:619-631); the
two modes below are the ones BASELINE.json:north_star names.

1. ``phi_sharded``  -- ONE swarm, particles sharded row-block over the ranks.  The path has one
   real exchange step: all-gather of signature features and scores; plus three 16 KiB histogram
   all-reduces for the exact global median and one all-reduce of (sum K*d, #ties) for the
   median-gradient term.  Every rank emits its rows of phi.
2. ``shard_scenes`` -- independent planning problems: split the batch, no communication.

The arithmetic goes through a *backend* object whose methods map 1:1 onto the C ABI
(``CudaBackend``; the only one this package ships).  The orchestration below touches the backend
only through that interface so the world_size-2 ``gloo`` tests can drive the same code on CPU
with a checker backend that lives under ``tests/``.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Tuple

import torch
import torch.distributed as dist

from . import _lib, ops


def shard_range(n: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous row block of rank ``rank``; the first ``n % world`` ranks get one extra row."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_scenes(n_scenes: int, rank: int, world: int) -> range:
    lo, hi = shard_range(n_scenes, rank, world)
    return range(lo, hi)


class CudaBackend:
    """Thin adapter over ``ops``' raw wrappers (each one C-ABI call)."""

    name = "cuda"

    def sig_fwd(self, x, depth):
        return ops.signature_fwd_raw(x, depth, True)

    def col_mean(self, s):
        return ops.col_mean(s)

    def pack_v(self, score, s, mean):
        return ops.pack_features(s, mean, score)

    def sqdist(self, a, b, D):
        return ops.sqdist_raw(a, b, D)

    def new_state(self, device):
        return ops.new_state(device)

    def new_hist(self, device):
        return torch.zeros(2048, dtype=torch.int64, device=device)

    def fixed_bandwidth(self, h, state):
        _lib.call("sigsvgd_fixed_bandwidth", float(h), state.data_ptr(), ops._stream())

    def select_begin(self, state, hist, k):
        _lib.call("sigsvgd_select_begin", state.data_ptr(), hist.data_ptr(), int(k), ops._stream())

    def select_hist(self, dist_loc, p, state, hist):
        _lib.call("sigsvgd_select_hist", dist_loc.data_ptr(), dist_loc.numel(), p, state.data_ptr(), hist.data_ptr(),
                  ops._stream())

    def select_scan(self, p, state, hist):
        _lib.call("sigsvgd_select_scan", p, state.data_ptr(), hist.data_ptr(), ops._stream())

    def select_finish(self, n_log, bw_scale, tol, state):
        _lib.call("sigsvgd_select_finish", n_log, float(bw_scale), float(tol), state.data_ptr(), ops._stream())

    def rbf_apply(self, k, state, row_offset):
        return ops.rbf_apply_raw(k, state, row_offset)

    def local_stats(self, state):
        """(sum_kd [1] float64, n_ties [1] int64) device tensors for the all-reduce."""
        return ops.state_field(state, "sum_kd").clone(), ops.state_field(state, "n_ties").to(torch.int64).clone()

    def median_coef(self, state, enable, sum_kd_all, n_ties_all):
        ops.median_coef_raw(state, enable, sum_kd_all, n_ties_all.to(torch.int32))

    def matmul_nn(self, a, v):
        return ops.matmul_nn_raw(a, v)

    def sig_grad(self, sc, sc_all, ks, rowsum, state, row_offset):
        return ops.sig_grad_raw(sc, sc_all, ks, rowsum, state, row_offset)

    def sig_bwd(self, x, sig, g, depth, add, alpha, beta):
        return ops.signature_bwd_raw(x, sig, g, depth, True, add, alpha, beta)


@dataclass
class ShardedPhi:
    phi: torch.Tensor  # [M, T, d] this rank's rows
    state: torch.Tensor  # bandwidth record (identical median / h on every rank)
    k_rows: Optional[torch.Tensor] = None  # [M, N] this rank's rows of k_xx (if requested)
    rows: Tuple[int, int] = (0, 0)


def _all_gather_rows(t: torch.Tensor, group) -> torch.Tensor:
    world = dist.get_world_size(group)
    out = torch.empty((world * t.shape[0],) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
    dist.all_gather_into_tensor(out, t.contiguous(), group=group)
    return out


def phi_sharded(x_local: torch.Tensor, score_local: torch.Tensor, depth: int = 3,
                bandwidth: Optional[ops.Bandwidth] = None, want_k: bool = False, center: bool = True,
                group=None, backend=None) -> ShardedPhi:
    """Row-block sharded Stein update.  ``x_local [M,T,d]`` / ``score_local [M,T*d]`` are this
    rank's particles; every rank must hold the same M (N = world * M).  SURVEY.md 8(e)."""
    be = backend or CudaBackend()
    bw = bandwidth or ops.Bandwidth.median()
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    M, T, d = x_local.shape
    N = world * M
    r0 = rank * M
    D = ops.sig_channels(d, depth)
    P = T * d
    x_local = x_local.contiguous()
    g_local = score_local.reshape(M, P).contiguous()

    s_local = be.sig_fwd(x_local, depth)
    if world > 1:
        # the one real exchange step of the path: [S | score] rows of every rank
        packed = torch.cat([s_local, g_local], dim=1)
        allp = _all_gather_rows(packed, group)
        s_all, g_all = allp[:, :D], allp[:, D:]
    else:
        s_all, g_all = s_local, g_local
    mean = be.col_mean(s_all) if center else None
    v_all = be.pack_v(g_all.contiguous(), s_all, mean)  # [N, C] = [score | S - mean | 1 | 0]
    f_all = v_all[:, P:]
    f_loc = v_all[r0:r0 + M, P:]
    k_rows = be.sqdist(f_loc, f_all, D)  # [M, N]

    state = be.new_state(x_local.device)
    if bw.mode == _lib.BW_FIXED:
        be.fixed_bandwidth(bw.h, state)
    else:
        hist = be.new_hist(x_local.device)
        be.select_begin(state, hist, (N * N - 1) // 2)
        for p in range(3):
            be.select_hist(k_rows, p, state, hist)
            if world > 1:
                dist.all_reduce(hist, op=dist.ReduceOp.SUM, group=group)
            be.select_scan(p, state, hist)
        be.select_finish(N, bw.bw_scale, bw.tol, state)
    be.rbf_apply(k_rows, state, r0)
    sum_kd, n_ties = be.local_stats(state)
    if world > 1:
        dist.all_reduce(sum_kd, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(n_ties, op=dist.ReduceOp.SUM, group=group)
    be.median_coef(state, bw.mode == _lib.BW_MEDIAN and bw.median_grad, sum_kd, n_ties)

    out = be.matmul_nn(k_rows, v_all)  # [M, C] = [K.score | K.S | rowsum | 0]
    g = be.sig_grad(f_loc, f_all, out[:, P:P + D], out[:, P + D], state, r0)
    phi = be.sig_bwd(x_local, s_local, g, depth, out, 1.0 / N, -1.0 / N)
    return ShardedPhi(phi, state, k_rows if want_k else None, (r0, r0 + M))


__all__ = ["phi_sharded", "shard_range", "shard_scenes", "CudaBackend", "ShardedPhi"]

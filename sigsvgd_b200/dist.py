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


@dataclass
class Features:
    """Operands of the two contractions, as one backend prepared them (opaque to the orchestration).

    ``sc_all``/``sc_loc``: 2-D fp32 views whose first D columns are the centred features of all
    particles / of this rank's rows; ``data``: whatever the backend's Gram and K.F kernels read."""

    sc_all: torch.Tensor
    sc_loc: torch.Tensor
    data: object
    pending: object = None  # exchanges still in flight (CudaBackend.prepare -> CudaBackend.gram)


class CudaBackend:
    """Thin adapter over ``ops``' raw wrappers (each one C-ABI call).

    ``tc=None`` picks the tcgen05 split-fp16 kernels when the swarm is large enough for them to
    pay (N >= 1024, like ``sigsvgd_phi``), else the exact-fp32 SIMT kernels."""

    name = "cuda"

    def __init__(self, tc: Optional[bool] = None):
        self.tc = tc

    def _use_tc(self, N: int) -> bool:
        if self.tc is None:
            return N >= 1024 and ops.tc_available()
        if self.tc and not ops.tc_available():
            raise RuntimeError("tcgen05 path requested but cuTensorMapEncodeTiled is unavailable")
        return bool(self.tc)

    def sig_fwd(self, x, depth):
        # with the prefix checkpoints of the time-parallel backward when the shape has them
        N, T, d = x.shape
        self._pref = None
        if N > 0 and ops.signature_prefix_floats(N, T, d, depth) > 0:
            sig, pref = ops.signature_fwd_prefix_raw(x, depth, True)
            self._pref = (sig.data_ptr(), pref)
            return sig
        return ops.signature_fwd_raw(x, depth, True)

    # ---- operand preparation / Gram / RBF / contraction ---------------------------------------
    def prepare(self, s_local, g_local, center, group, rank, world) -> Optional[Features]:
        """tcgen05 route: every rank packs the split-fp16 operand planes of ITS OWN rows only and the
        planes are what is exchanged (same bytes as the fp32 [S | score] rows they stand for), instead
        of every rank packing all N rows after the exchange.  The mean and the power-of-two scales
        that all ranks must share come from one all-gather of a (D + 2)-float statistics vector.
        Returns None when this backend wants the generic gather-then-pack route."""
        M, D = s_local.shape
        P = g_local.shape[1]
        N = world * M
        if not self._use_tc(N):
            return None
        dev = s_local.device
        Cp = (D + 1 + P + 63) // 64 * 64
        # 1. global statistics (column sums, max |s|, max |score|): one pass over this rank's rows into
        #    the library's 8-slot record, one all-gather of a (D + 2)-float vector per rank
        stats = torch.empty(8, D + 2, dtype=torch.float32, device=dev)
        ops.call("sigsvgd_tc_stats", ops._p(s_local), s_local.stride(0), M, D, ops._p(g_local), g_local.stride(0), P,
                 ops._p(stats), ops._stream())
        vec = torch.cat([stats[:, :D].sum(0), stats[:, D:].amax(0)])
        if world > 1:
            allv = torch.empty(world, D + 2, dtype=torch.float32, device=dev)
            dist.all_gather_into_tensor(allv, vec, group=group)
        else:
            allv = vec[None]
        rec = torch.zeros(8, D + 2, dtype=torch.float32, device=dev)  # every rank builds the same record
        if world <= 8:
            rec[:world] = allv
        else:
            rec[:7] = allv[:7]
            rec[7, :D] = allv[7:, :D].sum(0)
            rec[7, D:] = allv[7:, D:].amax(0)
        # 2. planes / norms / centred features of this rank's rows, packed in place into the buffers that
        #    will hold every rank's rows; the packing pass derives the mean (over all N particles) and the
        #    power-of-two scales from the record, exactly as the one-GPU update does
        r0 = rank * M
        scales = torch.zeros(8 + Cp, dtype=torch.float32, device=dev)
        fh = torch.empty(N, Cp, dtype=torch.float16, device=dev)
        fl = torch.empty(N, Cp, dtype=torch.float16, device=dev)
        norms = torch.empty(N, dtype=torch.float32, device=dev)
        sc_all = torch.empty(N, D, dtype=torch.float32, device=dev)
        ops.call("sigsvgd_tc_pack_from_stats", ops._p(g_local), g_local.stride(0), P, ops._p(s_local),
                 s_local.stride(0), ops._p(rec), N, M, D, int(bool(center)), ops._p(scales), ops._p(fh[r0:]),
                 ops._p(fl[r0:]), Cp, Cp, ops._p(norms[r0:]), ops._p(sc_all[r0:]), D, None, ops._stream())
        # 3. the one real exchange step of the path: all-gather of the operand planes (what the Gram strip
        #    waits for) and, behind them and asynchronous, of the centred fp32 features (needed only after
        #    the Gram -- median sample, tie partners -- so their transfer runs under it).
        #    (Measured on 8 GPUs, N = 65536: splitting the exchange into one broadcast per source rank
        #    and computing the Gram strip block by block as the blocks land was NOT faster -- 16
        #    collective launches cost the host more than the overlap gains: 7.31 vs 7.17 ms per update.)
        pending = None
        if world > 1:
            # (in place: every rank's rows already sit at their offset in the gathered buffer)
            works = [dist.all_gather_into_tensor(t, t[r0:r0 + M], group=group, async_op=True) for t in (fh, fl, norms)]
            sc_work = dist.all_gather_into_tensor(sc_all, sc_all[r0:r0 + M], group=group, async_op=True)
            pending = (works, sc_work)
        return Features(sc_all, sc_all[r0:r0 + M], ("tc", ops.Planes(fh, fl), norms, scales, D, P, r0, M), pending)

    def pack(self, s_all, g_all, center, r0, M) -> Features:
        N, D = s_all.shape
        P = g_all.shape[1]
        if self._use_tc(N):
            planes, norms, scales, sc = ops.tc_pack_fused_raw(s_all, g_all, center)
            return Features(sc, sc[r0:r0 + M], ("tc", planes, norms, scales, D, P, r0, M))
        mean = ops.col_mean(s_all) if center else None
        v = ops.pack_features(s_all, mean, g_all.contiguous())  # [N, C] = [score | S - mean | 1 | 0]
        return Features(v[:, P:], v[r0:r0 + M, P:], ("simt", v, D, P, r0, M))

    def gram(self, f: Features) -> torch.Tensor:
        if f.data[0] == "tc":
            _, planes, norms, scales, D, P, r0, M = f.data
            rows = ops.Planes(planes.hi[r0:r0 + M], planes.lo[r0:r0 + M])
            # (contiguous when N % 4 == 0: the flat select kernels and the TMA-fed contraction both take it)
            N = planes.hi.shape[0]
            out = torch.empty(M, N, dtype=torch.float32, device=planes.hi.device) if N % 4 == 0 else \
                ops.dist_buffer(M, N, planes.hi.device)
            if f.pending is None:
                return ops.tc_gram_raw(rows, norms[r0:r0 + M], planes, norms, D, scales, False, out=out)
            works, sc_work = f.pending
            for w in works:
                w.wait()  # the compute stream waits; the host does not
            ops.tc_gram_raw(rows, norms[r0:r0 + M], planes, norms, D, scales, False, out=out)
            sc_work.wait()
            f.pending = None
            return out
        _, v, D, P, r0, M = f.data
        return ops.sqdist_raw(v[r0:r0 + M, P:], v[:, P:], D)

    def rbf(self, k_rows, state, row_offset, f: Features, want_k: bool):
        """SIMT: exp in place, adding this strip's sum K*d / ties to ``state``.  tcgen05: nothing to
        do here -- the RBF map and its statistics run inside the contraction's operand path.
        Returns (contraction operand, k_xx rows or None)."""
        if f.data[0] == "tc":
            return k_rows, (ops.rbf_out_raw(k_rows, state) if want_k else None)
        k = ops.rbf_apply_raw(k_rows, state, row_offset)
        return k, (k if want_k else None)

    def contract(self, kmat, f: Features, state, row_offset):
        """-> (K.Sc [M,D], rowsum [M], K.score [M,P]) views of one [M, C] result.  After it ``state``
        holds this strip's sum K*d / ties on either path."""
        if f.data[0] == "tc":
            _, planes, norms, scales, D, P, r0, M = f.data
            out = ops.tc_kf_raw(kmat, planes, D + 1 + P, scales, state, _lib.KF_EXP, True, row_offset)
            return out[:, :D], out[:, D], out[:, D + 1:D + 1 + P]
        _, v, D, P, r0, M = f.data
        out = ops.matmul_nn_raw(kmat, v)
        return out[:, P:P + D], out[:, P + D], out[:, :P]

    # ---- bandwidth ------------------------------------------------------------------------------
    def new_state(self, device):
        return ops.new_state(device)

    def new_hist(self, device):
        return torch.zeros(2048, dtype=torch.int64, device=device)

    def fixed_bandwidth(self, h, state):
        ops.call("sigsvgd_fixed_bandwidth", float(h), ops._p(state), ops._stream())

    def select_begin(self, state, hist, k):
        ops.call("sigsvgd_select_begin", ops._p(state), ops._p(hist), int(k), ops._stream())

    def select_begin_window(self, state, hist, k_total, totals):
        """``select_begin`` for a select over window candidates with the rank taken relative to the
        all-rank totals ON THE DEVICE.  Returns a device int32 flag: 1 = the window held the rank."""
        ok = torch.zeros(1, dtype=torch.int32, device=state.device)
        ops.call("sigsvgd_select_begin_window", ops._p(state), ops._p(hist), int(k_total), ops._p(totals), ops._p(ok),
                 ops._stream())
        return ok

    def select_hist(self, dist_loc, p, state, hist):
        ops.call("sigsvgd_select_hist", ops._p(dist_loc), dist_loc.numel(), p, ops._p(state), ops._p(hist),
                  ops._stream())

    def select_scan(self, p, state, hist):
        ops.call("sigsvgd_select_scan", p, ops._p(state), ops._p(hist), ops._stream())

    def select_finish(self, n_log, bw_scale, tol, state):
        ops.call("sigsvgd_select_finish", n_log, float(bw_scale), float(tol), ops._p(state), ops._stream())

    # windowed first pass of the sharded median (one read of the strip instead of three)
    window_min_elems = 1 << 24  # strips smaller than this take the plain three passes

    def select_window(self, sc_all, state):
        """-> device tensor whose first two floats are [lo, hi]; identical on every rank."""
        lib = _lib.load()
        nbytes = lib.sigsvgd_select_window_workspace_bytes()
        ws = torch.empty(nbytes, dtype=torch.uint8, device=sc_all.device)
        ops.call("sigsvgd_select_window", ops._p(sc_all), sc_all.stride(0), sc_all.shape[0], sc_all.shape[1],
                  ops._p(ws), nbytes, ops._p(state), ops._stream())
        return ws[:8].view(torch.float32)

    def select_collect(self, dist_loc, bounds):
        """-> (candidates [cap] fp32 padded with a large value, totals int64[3] = below, kept, overflowed)."""
        lib = _lib.load()
        cap = max(int(lib.sigsvgd_select_collect_min_capacity()) * 64, dist_loc.numel() // 8)
        cand = torch.empty(cap, dtype=torch.float32, device=dist_loc.device)
        totals = torch.empty(3, dtype=torch.int64, device=dist_loc.device)
        ops.call("sigsvgd_select_collect", ops._p(dist_loc), dist_loc.numel(), ops._p(bounds), ops._p(cand),
                  cap, ops._p(totals), ops._stream())
        return cand, totals

    def local_stats(self, state):
        """[sum_kd, n_ties] as one float64 device tensor for the all-reduce."""
        return torch.cat([ops.state_field(state, "sum_kd"), ops.state_field(state, "n_ties").to(torch.float64)])

    def median_coef(self, state, enable, stats):
        ops.median_coef_raw(state, enable, stats[:1].contiguous(), stats[1:].to(torch.int32))

    def sig_grad(self, sc, sc_all, ks, rowsum, state, row_offset):
        return ops.sig_grad_raw(sc, sc_all, ks, rowsum, state, row_offset)

    def sig_bwd(self, x, sig, g, depth, add, alpha, beta):
        pref = getattr(self, "_pref", None)
        if pref is not None and pref[0] == sig.data_ptr():
            return ops.signature_bwd_prefix_raw(x, sig, pref[1], g, depth, True, add, alpha, beta)
        return ops.signature_bwd_raw(x, sig, g, depth, True, add, alpha, beta)


@dataclass
class ShardedPhi:
    phi: torch.Tensor  # [M, T, d] this rank's rows
    state: torch.Tensor  # bandwidth record (identical median / h on every rank)
    k_rows: Optional[torch.Tensor] = None  # [M, N] this rank's rows of k_xx (if requested)
    rows: Tuple[int, int] = (0, 0)
    window_ok: Optional[torch.Tensor] = None  # device flag of the windowed median (ShardedPlan only)


def _all_gather_rows(t: torch.Tensor, group) -> torch.Tensor:
    world = dist.get_world_size(group)
    out = torch.empty((world * t.shape[0],) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
    dist.all_gather_into_tensor(out, t.contiguous(), group=group)
    return out


def phi_sharded(x_local: torch.Tensor, score_local: torch.Tensor, depth: int = 3,
                bandwidth: Optional[ops.Bandwidth] = None, want_k: bool = False, center: bool = True,
                group=None, backend=None, window: Optional[bool] = None,
                timings: Optional[dict] = None, _defer_window_check: bool = False) -> ShardedPhi:
    """Row-block sharded Stein update.  ``x_local [M,T,d]`` / ``score_local [M,T*d]`` are this
    rank's particles; every rank must hold the same M (N = world * M).  SURVEY.md 8(e).
    ``window``: windowed first pass of the median (default: on for large strips).
    ``timings``: if a dict, filled with this rank's milliseconds per phase (CUDA events on the current
    stream, read after the step; the collectives are inside the phase that issues them)."""
    be = backend or CudaBackend()
    marks = []

    def mark(name):
        if timings is not None and x_local.is_cuda:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record(torch.cuda.current_stream(x_local.device))
            marks.append((name, ev))

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

    mark("start")
    s_local = be.sig_fwd(x_local, depth)
    mark("signature_fwd")
    feat = be.prepare(s_local, g_local, center, group, rank, world) if hasattr(be, "prepare") else None
    if feat is None:
        if world > 1:
            # the one real exchange step of the path: [S | score] rows of every rank
            allp = _all_gather_rows(torch.cat([s_local, g_local], dim=1), group)
            s_all, g_all = allp[:, :D], allp[:, D:]
        else:
            s_all, g_all = s_local, g_local
        feat = be.pack(s_all, g_all, center, r0, M)
    mark("pack_and_all_gather")
    k_rows = be.gram(feat)  # [M, N] squared distances of this rank's rows
    mark("gram_strip")

    state = be.new_state(x_local.device)
    window_ok = None  # device flag of an optimistic windowed select (checked once, at the end)
    if bw.mode == _lib.BW_FIXED:
        be.fixed_bandwidth(bw.h, state)
    else:
        # exact GLOBAL lower median: radix select (11/11/10 bits of the fp32 patterns), one 16 KiB
        # histogram all-reduce per pass; every rank ends with the same median and h
        hist = be.new_hist(x_local.device)
        k = (N * N - 1) // 2
        data = k_rows if k_rows.is_contiguous() else k_rows.contiguous()  # the select kernels take flat arrays
        if window is None:
            window = k_rows.numel() >= getattr(be, "window_min_elems", 1 << 62)
        if window:
            # one read of the strip: count the entries below a sampled window around the median and
            # compact the entries inside it; the select then runs over those candidates only.  Whether
            # the window held the median is decided on the device (no host read in the middle of the
            # step); in the ~1e-8 case that it did not, the step is repeated with the plain passes.
            bounds = be.select_window(feat.sc_all, state)
            data, totals = be.select_collect(data, bounds)
            if world > 1:
                dist.all_reduce(totals, op=dist.ReduceOp.SUM, group=group)
            window_ok = be.select_begin_window(state, hist, k, totals)
        else:
            be.select_begin(state, hist, k)
        for p in range(3):
            be.select_hist(data, p, state, hist)
            if world > 1:
                dist.all_reduce(hist, op=dist.ReduceOp.SUM, group=group)
            be.select_scan(p, state, hist)
        be.select_finish(N, bw.bw_scale, bw.tol, state)
    mark("global_median")
    kmat, k_out = be.rbf(k_rows, state, r0, feat, want_k)
    ks, rowsum, kg = be.contract(kmat, feat, state, r0)  # K.Sc [M,D], row sums [M], K.score [M,P]
    mark("kf_strip")
    stats = be.local_stats(state)
    if world > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.SUM, group=group)
    be.median_coef(state, bw.mode == _lib.BW_MEDIAN and bw.median_grad, stats)
    g = be.sig_grad(feat.sc_loc, feat.sc_all, ks, rowsum, state, r0)
    phi = be.sig_bwd(x_local, s_local, g, depth, kg, 1.0 / N, -1.0 / N)
    mark("coef_and_signature_bwd")
    if _defer_window_check:  # (ShardedPlan: the flag is read after the graph replay)
        return ShardedPhi(phi, state, k_out, (r0, r0 + M), window_ok)
    if window_ok is not None and not bool(window_ok.item()):  # identical on every rank (all-reduced totals)
        return phi_sharded(x_local, score_local, depth, bandwidth, want_k, center, group, backend, window=False,
                           timings=timings)
    if marks:
        marks[-1][1].synchronize()
        timings.clear()
        timings.update({b[0]: a[1].elapsed_time(b[1]) for a, b in zip(marks[:-1], marks[1:])})
    return ShardedPhi(phi, state, k_out, (r0, r0 + M))


class ShardedPlan:
    """``phi_sharded`` of one shard shape captured ONCE into a CUDA graph -- kernels AND the NCCL
    collectives -- and replayed: the sharded step is ~40 launches and 9 collectives, and at 8 ranks
    the host's enqueue time (not the GPUs) is a sixth of the step.  Every rank must build and call
    the plan in lock-step (the capture itself runs the collectives).

    The plan owns static buffers: pass this rank's rows to ``__call__`` (copied in) and read
    ``.phi`` / ``.state`` of the returned ``ShardedPhi``.  The windowed median's "window held the
    median" flag is read after the replay; in the ~1e-8 case that it did not, the step is repeated
    eagerly with the plain three passes.  Delete the plan before ``dist.destroy_process_group()``: a
    live graph that holds NCCL work keeps the communicator's teardown waiting."""

    def __init__(self, M: int, T: int, d: int, depth: int = 3, bandwidth: Optional[ops.Bandwidth] = None,
                 center: bool = True, group=None, device=None):
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.args = (depth, bandwidth, False, center, group)
        gen = torch.Generator().manual_seed(1234 + (dist.get_rank(group) if dist.is_initialized() else 0))
        self.x = torch.cumsum(0.05 * torch.randn(M, T, d, generator=gen), dim=1).to(dev)
        self.score = torch.randn(M, T * d, generator=gen).to(dev)
        self._stream = torch.cuda.Stream(device=dev)
        self._stream.wait_stream(torch.cuda.current_stream(dev))
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.stream(self._stream):
            for _ in range(2):  # warm-up outside the capture: communicator set-up, lazy initialisation
                phi_sharded(self.x, self.score, *self.args, _defer_window_check=True)
            self._stream.synchronize()
            with torch.cuda.graph(self.graph, stream=self._stream, capture_error_mode="thread_local"):
                self.res = phi_sharded(self.x, self.score, *self.args, _defer_window_check=True)
        torch.cuda.current_stream(dev).wait_stream(self._stream)

    def __call__(self, x_local: torch.Tensor, score_local: torch.Tensor) -> ShardedPhi:
        if x_local.data_ptr() != self.x.data_ptr():
            self.x.copy_(x_local.reshape(self.x.shape))
        if score_local.data_ptr() != self.score.data_ptr():
            self.score.copy_(score_local.reshape(self.score.shape))
        self.graph.replay()
        ok = self.res.window_ok
        if ok is not None and not bool(ok.item()):
            depth, bw, want_k, center, group = self.args
            return phi_sharded(self.x, self.score, depth, bw, want_k, center, group, window=False)
        return self.res


__all__ = ["phi_sharded", "shard_range", "shard_scenes", "CudaBackend", "ShardedPhi", "ShardedPlan"]

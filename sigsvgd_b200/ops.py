"""Torch-facing operators over the C ABI: device memory, streams and autograd plumbing only.

* ``signature``      -- drop-in for ``signatory.signature(path, depth, basepoint=True)``
                        (reference call sites src/kernels/_traj_kernels.py:125-126)
* ``sig_rbf_gram``   -- K = GaussianKernel(signature(X), signature(Y)) as ONE autograd node whose
                        backward is the closed form of ``autograd.grad(k_xx.sum(), X)``
                        (src/inference/svgd.py:43, src/inference/score.py:55) incl. the gradient
                        through ``torch.median`` (src/utils/math.py:31)
* ``rbf_gram``       -- the static RBF kernel on plain features (src/kernels/_kernels.py:78-115)
* ``phi``            -- fused Stein update (src/inference/svgd.py:36-44 + 82-83)

Everything computes in fp32 on the current CUDA device; other float dtypes are cast in and the
result is cast back.  No CPU path exists: CPU tensors raise.
"""
from __future__ import annotations

import math
import threading
from dataclasses import dataclass
from typing import Optional

import torch

from . import _lib


# ---------------------------------------------------------------------------------------------
# helpers
# ---------------------------------------------------------------------------------------------
# The C library launches on the CURRENT device and takes the stream as an argument, while the
# reference's only multi-GPU idiom is ``device=f"cuda:{i}"`` with no ``set_device`` anywhere
# (examples/script_planning_robot.py:486).  So every C call here runs with the device of its tensor
# arguments made current and on that device's current stream: ``_p`` notes the device of each pointer
# it hands out, ``_stream`` (always the last argument) and ``call`` use it.  Tensors on two different
# devices in one call raise.
_pending = threading.local()


def _note_device(dev) -> None:
    cur = getattr(_pending, "device", None)
    if cur is None:
        _pending.device = dev
    elif cur != dev:
        _pending.device = None
        raise RuntimeError(f"sigsvgd_b200: tensors of one call live on different devices ({cur} and {dev})")


def _p(t: Optional[torch.Tensor]):
    if t is None:
        return None
    if t.is_cuda:
        _note_device(t.device)
    return t.data_ptr()


def _stream():
    dev = getattr(_pending, "device", None)
    return torch.cuda.current_stream(dev).cuda_stream


def call(name: str, *args) -> None:
    dev = getattr(_pending, "device", None)
    _pending.device = None
    if dev is None or dev.index == torch.cuda.current_device():
        _lib.call(name, *args)
    else:
        with torch.cuda.device(dev):
            _lib.call(name, *args)


def _prep(t: torch.Tensor, what: str) -> torch.Tensor:
    if not isinstance(t, torch.Tensor):
        raise TypeError(f"{what}: expected a torch.Tensor, got {type(t)}")
    if not t.is_cuda:
        raise RuntimeError(
            f"{what}: tensor is on {t.device}; sigsvgd_b200 runs on CUDA (sm_100a) only and has no CPU fallback"
        )
    if not t.is_floating_point():
        raise TypeError(f"{what}: expected a floating-point tensor, got {t.dtype}")
    return t.detach().to(torch.float32).contiguous()


def sig_channels(d: int, depth: int) -> int:
    return sum(d ** k for k in range(1, depth + 1))


@dataclass
class Bandwidth:
    """How the RBF bandwidth is chosen (src/kernels/_kernels.py:106-109)."""

    mode: int = _lib.BW_MEDIAN  # BW_MEDIAN: src/utils/math.py:28-34 ; BW_FIXED: float(h)
    h: float = 1.0
    bw_scale: float = 1.0
    tol: float = 1.0e-8
    median_grad: bool = True  # differentiate through the median like torch.autograd does

    @staticmethod
    def median(bw_scale: float = 1.0, tol: float = 1.0e-8, median_grad: bool = True) -> "Bandwidth":
        return Bandwidth(_lib.BW_MEDIAN, 1.0, float(bw_scale), float(tol), bool(median_grad))

    @staticmethod
    def fixed(h: float) -> "Bandwidth":
        h = float(h)
        if not h > 0:
            raise ValueError(f"bandwidth must be > 0, got {h}")
        return Bandwidth(_lib.BW_FIXED, h, 1.0, 1.0e-8, False)


def new_state(device, batch: int = 1) -> torch.Tensor:
    """Device buffer holding ``batch`` sigsvgd_bw_state records (as int32 words)."""
    return torch.zeros(batch, _lib.STATE_BYTES // 4, dtype=torch.int32, device=device)


def state_field(state: torch.Tensor, name: str) -> torch.Tensor:
    """View of one field of every record, still on the device (no sync)."""
    if name in ("median", "h", "gamma", "coef"):
        off = {"median": 0, "h": 1, "gamma": 2, "coef": 3}[name]
        return state[:, off].view(torch.float32)
    if name == "sum_kd":
        return state[:, 4:6].contiguous().view(torch.float64).reshape(-1)
    if name in ("n_ties", "n_ties_all", "clamped"):
        return state[:, {"n_ties": 6, "n_ties_all": 7, "clamped": 8}[name]]
    if name == "ties":
        return state[:, _lib.OFF_TIES // 4:].reshape(state.shape[0], _lib.MAX_TIES, 2)
    raise KeyError(name)


def read_state(state: torch.Tensor, index: int = 0) -> dict:
    """Synchronising host copy of one record (diagnostics / tests)."""
    raw = bytes(state[index].cpu().numpy().tobytes())
    s = _lib.BwState.from_buffer_copy(raw)
    n = min(int(s.n_ties), _lib.MAX_TIES)
    return dict(median=s.median, h=s.h, gamma=s.gamma, coef=s.coef, sum_kd=s.sum_kd, n_ties=int(s.n_ties),
                n_ties_all=int(s.n_ties_all), clamped=int(s.clamped),
                ties=[(s.ties[2 * i], s.ties[2 * i + 1]) for i in range(n)])


# ---------------------------------------------------------------------------------------------
# raw (non-autograd) wrappers, one per C entry point
# ---------------------------------------------------------------------------------------------
def signature_fwd_raw(x: torch.Tensor, depth: int, basepoint: bool = True) -> torch.Tensor:
    N, T, d = x.shape
    sig = torch.empty(N, sig_channels(d, depth), dtype=torch.float32, device=x.device)
    call("sigsvgd_signature_fwd", _p(x), N, T, d, depth, int(basepoint), _p(sig), _stream())
    return sig


def signature_bwd_raw(x, sig, gsig, depth, basepoint=True, add=None, alpha=0.0, beta=1.0) -> torch.Tensor:
    N, T, d = x.shape
    out = torch.empty_like(x)
    ld_add = 0 if add is None else add.stride(0)
    call("sigsvgd_signature_bwd", _p(x), _p(sig), _p(gsig), N, T, d, depth, int(basepoint), _p(add), ld_add,
         float(alpha), float(beta), _p(out), _stream())
    return out


def signature_prefix_floats(N: int, T: int, d: int, depth: int) -> int:
    """Size of the prefix-checkpoint buffer of the time-parallel pair; 0 when the shape does not qualify."""
    return int(_lib.load().sigsvgd_signature_prefix_floats(N, T, d, depth))


def signature_fwd_prefix_raw(x: torch.Tensor, depth: int, basepoint: bool = True):
    """Forward that also checkpoints the signature at the ends of the first three time segments."""
    N, T, d = x.shape
    n = signature_prefix_floats(N, T, d, depth)
    if n <= 0:
        raise ValueError(f"no time-parallel signature pair for T={T} d={d} depth={depth}")
    sig = torch.empty(N, sig_channels(d, depth), dtype=torch.float32, device=x.device)
    pref = torch.empty(n, dtype=torch.float32, device=x.device)
    call("sigsvgd_signature_fwd_prefix", _p(x), N, T, d, depth, int(basepoint), _p(sig), _p(pref), _stream())
    return sig, pref


def signature_bwd_prefix_raw(x, sig, pref, gsig, depth, basepoint=True, add=None, alpha=0.0, beta=1.0) -> torch.Tensor:
    N, T, d = x.shape
    out = torch.empty_like(x)
    ld_add = 0 if add is None else add.stride(0)
    call("sigsvgd_signature_bwd_prefix", _p(x), _p(sig), _p(pref), _p(gsig), N, T, d, depth, int(basepoint), _p(add),
         ld_add, float(alpha), float(beta), _p(out), _stream())
    return out


def pack_features(s: torch.Tensor, mean: Optional[torch.Tensor], score: Optional[torch.Tensor] = None) -> torch.Tensor:
    """V = [score | s - mean | 1 | 0-pad]  (row stride a multiple of 8 floats)."""
    N, D = s.shape
    P = 0 if score is None else score.shape[1]
    C = (P + D + 1 + 7) // 8 * 8
    v = torch.empty(N, C, dtype=torch.float32, device=s.device)
    call("sigsvgd_pack_v", _p(score), P, _p(s), s.stride(0), _p(mean), N, D, _p(v), C, C, _stream())
    return v


def col_mean(s: torch.Tensor) -> torch.Tensor:
    N, D = s.shape
    part = torch.empty(64, D, dtype=torch.float32, device=s.device)
    mean = torch.empty(D, dtype=torch.float32, device=s.device)
    call("sigsvgd_col_mean", _p(s), s.stride(0), N, D, _p(part), _p(mean), _stream())
    return mean


def sqdist_raw(a: torch.Tensor, b: torch.Tensor, D: int, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """a, b: 2-D fp32 views whose first D columns are the features (row stride = .stride(0))."""
    M, N = a.shape[0], b.shape[0]
    if out is None:
        out = torch.empty(M, N, dtype=torch.float32, device=a.device)
    call("sigsvgd_sqdist", _p(a), a.stride(0), _p(b), b.stride(0), M, N, D, _p(out), out.stride(0), 0, _stream())
    return out


def bandwidth_raw(dist: torch.Tensor, bw: Bandwidth, n_log: int, state: Optional[torch.Tensor] = None) -> torch.Tensor:
    if state is None:
        state = new_state(dist.device)
    if bw.mode == _lib.BW_FIXED:
        call("sigsvgd_fixed_bandwidth", bw.h, _p(state), _stream())
    else:
        if not dist.is_contiguous():
            raise ValueError("median bandwidth needs a contiguous distance matrix")
        nbytes = _lib.load().sigsvgd_select_workspace_bytes(dist.numel())
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dist.device)
        call("sigsvgd_median_bandwidth", _p(dist), dist.numel(), n_log, bw.bw_scale, bw.tol, _p(state), _p(ws),
             nbytes, _stream())
    return state


def rbf_apply_raw(k: torch.Tensor, state: torch.Tensor, row_offset: int = 0, weight: Optional[torch.Tensor] = None):
    M, N = k.shape
    call("sigsvgd_rbf_apply", _p(k), k.stride(0), M, N, row_offset, _p(weight),
         0 if weight is None else weight.stride(0), _p(state), _stream())
    return k


def median_coef_raw(state, enable: bool, sum_kd_all=None, n_ties_all=None):
    call("sigsvgd_median_coef", _p(state), _p(sum_kd_all), _p(n_ties_all), int(enable), _stream())


def matmul_nn_raw(a: torch.Tensor, v: torch.Tensor) -> torch.Tensor:
    M, N = a.shape
    C = v.shape[1]
    out = torch.empty(M, C, dtype=torch.float32, device=a.device)
    call("sigsvgd_matmul_nn", _p(a), a.stride(0), _p(v), v.stride(0), M, N, C, _p(out), C, 0, _stream())
    return out


# ---- tensor-core building blocks (tcgen05 split-fp16 path) ----------------------------------
@dataclass
class Planes:
    """fp16 hi/lo planes of a scaled fp32 matrix (see include/sigsvgd_b200.h)."""

    hi: torch.Tensor
    lo: torch.Tensor

    @property
    def ld(self) -> int:
        return self.hi.stride(0)


def tc_available() -> bool:
    return bool(_lib.load().sigsvgd_tc_available())


def tc_pack_raw(s: torch.Tensor, mean: Optional[torch.Tensor], score: Optional[torch.Tensor] = None):
    """-> (F planes [N, Cp], norms [N], scales [8 + Cp], centred fp32 features [N, D])."""
    N, D = s.shape
    P = 0 if score is None else score.shape[1]
    Cp = (D + 1 + P + 63) // 64 * 64  # 128-byte rows: aligned TMA segments
    dev = s.device
    scales = torch.empty(8 + Cp, dtype=torch.float32, device=dev)
    call("sigsvgd_tc_scales", _p(s), s.stride(0), _p(mean), N, D, _p(score), 0 if score is None else score.stride(0), P,
         Cp, _p(scales), _stream())
    fh = torch.empty(N, Cp, dtype=torch.float16, device=dev)
    fl = torch.empty(N, Cp, dtype=torch.float16, device=dev)
    norms = torch.empty(N, dtype=torch.float32, device=dev)
    sc = torch.empty(N, D, dtype=torch.float32, device=dev)
    call("sigsvgd_tc_pack", _p(score), 0 if score is None else score.stride(0), P, _p(s), s.stride(0), _p(mean), N, D,
         _p(scales), _p(fh), _p(fl), Cp, Cp, _p(norms), _p(sc), D, _stream())
    return Planes(fh, fl), norms, scales, sc


def tc_pack_fused_raw(s: torch.Tensor, score: Optional[torch.Tensor] = None, center: bool = True):
    """Same outputs as ``col_mean`` + ``tc_pack_raw`` from one statistics pass and one packing pass."""
    N, D = s.shape
    P = 0 if score is None else score.shape[1]
    Cp = (D + 1 + P + 63) // 64 * 64
    dev = s.device
    scales = torch.empty(8 + Cp, dtype=torch.float32, device=dev)
    stats = torch.empty(8 * (D + 2), dtype=torch.float32, device=dev)
    fh = torch.empty(N, Cp, dtype=torch.float16, device=dev)
    fl = torch.empty(N, Cp, dtype=torch.float16, device=dev)
    norms = torch.empty(N, dtype=torch.float32, device=dev)
    sc = torch.empty(N, D, dtype=torch.float32, device=dev)
    call("sigsvgd_tc_pack_fused", _p(score), 0 if score is None else score.stride(0), P, _p(s), s.stride(0), N, D,
         int(center), _p(scales), _p(fh), _p(fl), Cp, Cp, _p(norms), _p(sc), D, None, _p(stats), _stream())
    return Planes(fh, fl), norms, scales, sc


def tc_median_workspace(M: int, N: int, device) -> torch.Tensor:
    nbytes = _lib.load().sigsvgd_tc_median_workspace_bytes(M, N)
    return torch.empty(nbytes, dtype=torch.uint8, device=device)


def tc_median_prepare_raw(a_c: torch.Tensor, b_c: torch.Tensor, ws: torch.Tensor, state: torch.Tensor) -> None:
    """a_c / b_c: centred fp32 features of the rows / columns of the distance matrix."""
    call("sigsvgd_tc_median_prepare", _p(a_c), a_c.stride(0), a_c.shape[0], _p(b_c), b_c.stride(0), b_c.shape[0],
         a_c.shape[1], _p(ws), ws.numel(), _p(state), _stream())


def tc_gram_raw(a: Planes, na: torch.Tensor, b: Planes, nb: torch.Tensor, D: int, scales: torch.Tensor,
                symmetric: bool = False, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    M, N = a.hi.shape[0], b.hi.shape[0]
    if out is None:
        out = torch.empty(M, N, dtype=torch.float32, device=a.hi.device)
    call("sigsvgd_tc_gram", _p(a.hi), _p(a.lo), a.ld, _p(na), _p(b.hi), _p(b.lo), b.ld, _p(nb), M, N, D, _p(scales),
         int(symmetric), _p(out), out.stride(0), _stream())
    return out


def dist_buffer(M: int, N: int, device) -> torch.Tensor:
    """[M, N] fp32 view whose rows are padded to whole 128-byte lines (what the TMA-fed kernels like)."""
    ld = (N + 31) // 32 * 32
    return torch.empty(M, ld, dtype=torch.float32, device=device)[:, :N]


def tc_median_select_raw(ws: torch.Tensor, dist: torch.Tensor, n_log: int, bw: "Bandwidth", state: torch.Tensor,
                         symmetric: bool = False) -> None:
    """``symmetric``: dist came from ``tc_gram_raw(..., symmetric=True)`` -> only the upper triangle is read."""
    M, N = dist.shape
    call("sigsvgd_tc_median_select", _p(ws), _p(dist), M, N, dist.stride(0), int(symmetric), n_log, bw.bw_scale,
         bw.tol, _p(state), _stream())


def tc_kf_raw(a: torch.Tensor, f: Planes, C: int, scales: torch.Tensor, state: Optional[torch.Tensor] = None,
              mode: int = _lib.KF_EXP, stats: bool = True, row_offset: int = 0) -> torch.Tensor:
    """out [M, Cp] = f(a)[M, K] . F[K, :C] with the RBF map fused into the operand path (columns >= C are
    not written).  ``a``: fp32 squared distances (``KF_EXP``) or kernel values (``KF_IDENTITY``); its row
    pitch must be a multiple of 4 floats."""
    M, K = a.shape
    if a.stride(1) != 1 or a.stride(0) % 4 or a.data_ptr() % 16:
        a = dist_buffer(M, K, a.device).copy_(a)
    out = torch.empty(M, f.hi.shape[1], dtype=torch.float32, device=a.device)
    call("sigsvgd_tc_kf", _p(a), a.stride(0), M, K, int(mode), _p(state), int(stats and mode == _lib.KF_EXP),
         row_offset, _p(f.hi), _p(f.lo), f.ld, C, _p(scales), _p(out), out.stride(0), _stream())
    return out


def rbf_out_raw(dist: torch.Tensor, state: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """K = exp(gamma * dist) as a contiguous fp32 matrix (no statistics)."""
    M, N = dist.shape
    if out is None:
        out = torch.empty(M, N, dtype=torch.float32, device=dist.device)
    call("sigsvgd_rbf_out", _p(dist), dist.stride(0), M, N, _p(state), _p(out), out.stride(0), _stream())
    return out


def opt_step_raw(x: torch.Tensor, direction: torch.Tensor, kind: int, lr: float, sign: float = 1.0,
                 m: Optional[torch.Tensor] = None, v: Optional[torch.Tensor] = None,
                 grad_out: Optional[torch.Tensor] = None, beta1: float = 0.0, beta2: float = 0.0, eps: float = 0.0,
                 c1: float = 0.0, c2: float = 1.0, first: bool = False) -> None:
    """One fused optimiser step IN PLACE on ``x`` (``sigsvgd_opt_step``; kinds ``_lib.OPT_*``)."""
    for t in (x, direction, m, v, grad_out):
        if t is not None and (not t.is_cuda or t.dtype != torch.float32 or not t.is_contiguous()
                              or t.numel() != x.numel()):
            raise ValueError("opt_step: contiguous float32 CUDA tensors of one size expected")
    call("sigsvgd_opt_step", _p(x), _p(direction), float(sign), _p(m), _p(v), _p(grad_out), x.numel(), int(kind),
         float(lr), float(beta1), float(beta2), float(eps), float(c1), float(c2), int(bool(first)), _stream())


def sig_grad_raw(sc, sc_all, ks, rowsum, state, row_offset: int = 0) -> torch.Tensor:
    """sc/sc_all/ks: 2-D views (features in the first D columns of ks.shape[1]); rowsum: 1-D view."""
    M, D = ks.shape
    g = torch.empty(M, D, dtype=torch.float32, device=ks.device)
    call("sigsvgd_sig_grad", _p(sc), sc.stride(0), _p(sc_all), sc_all.stride(0), _p(ks), ks.stride(0), _p(rowsum),
         rowsum.stride(0), M, D, row_offset, _p(state), _p(g), D, _stream())
    return g


def _swap_ties(state: torch.Tensor) -> torch.Tensor:
    """Record with (row, col) of every tie exchanged -- the second-argument view of the tie set."""
    s2 = state.clone()
    o = _lib.OFF_TIES // 4
    s2[:, o:] = state[:, o:].reshape(-1, _lib.MAX_TIES, 2).flip(-1).reshape(state.shape[0], -1)
    return s2


def _same_data(a: torch.Tensor, b: torch.Tensor) -> bool:
    return a.data_ptr() == b.data_ptr() and a.shape == b.shape and a.stride() == b.stride() and a.dtype == b.dtype


def _is_uniform_cotangent(g: torch.Tensor) -> bool:
    """True for the expanded scalar that ``k_xx.sum().backward()`` produces."""
    return all(s == 0 for s in g.stride()) or g.numel() == 1


def _sig_fwd_saving(x, depth, basepoint):
    """Forward for an autograd node: also the prefix checkpoints when the time-parallel backward
    exists for this shape (returns (sig, pref) with pref an empty tensor otherwise)."""
    N, T, d = x.shape
    if N > 0 and signature_prefix_floats(N, T, d, depth) > 0:
        return signature_fwd_prefix_raw(x, depth, basepoint)
    return signature_fwd_raw(x, depth, basepoint), torch.empty(0, dtype=torch.float32, device=x.device)


def _sig_bwd_saved(x, sig, pref, g, depth, basepoint):
    if pref.numel():
        return signature_bwd_prefix_raw(x, sig, pref, g, depth, basepoint)
    return signature_bwd_raw(x, sig, g, depth, basepoint)


# ---------------------------------------------------------------------------------------------
# autograd nodes
# ---------------------------------------------------------------------------------------------
class _Signature(torch.autograd.Function):
    @staticmethod
    def forward(ctx, path, depth, basepoint):
        x = _prep(path, "signature")
        if x.dim() != 3:
            raise ValueError(f"signature: expected [batch, length, channels], got {tuple(x.shape)}")
        sig, pref = _sig_fwd_saving(x, depth, basepoint)
        ctx.save_for_backward(x, sig, pref)
        ctx.depth, ctx.basepoint, ctx.in_dtype = depth, basepoint, path.dtype
        return sig.to(path.dtype)

    @staticmethod
    def backward(ctx, g):
        x, sig, pref = ctx.saved_tensors
        gx = _sig_bwd_saved(x, sig, pref, _prep(g, "signature.backward"), ctx.depth, ctx.basepoint)
        return gx.to(ctx.in_dtype), None, None


def signature(path, depth, stream=False, basepoint=False, inverse=False, initial=None, scalar_term=False):
    """``signatory.signature`` with the options the reference uses (path, depth, basepoint)."""
    if stream or inverse or initial is not None or scalar_term:
        raise NotImplementedError("sigsvgd_b200.signature supports (path, depth, basepoint) only")
    if isinstance(basepoint, torch.Tensor):
        bp = basepoint.to(path).reshape(-1, 1, path.shape[2]).expand(path.shape[0], 1, path.shape[2])
        path = torch.cat([bp, path], dim=1)
        basepoint = False
    return _Signature.apply(path, int(depth), bool(basepoint))


#: swarms at least this large take the tcgen05 kernels in the autograd nodes (like ``sigsvgd_phi``)
TC_MIN_ROWS = 256


def _use_tc(n_rows: int) -> bool:
    return n_rows >= TC_MIN_ROWS and tc_available()


def _square_buffer(M: int, N: int, device) -> torch.Tensor:
    """[M, N] fp32 for distances / kernel values that the TMA-fed kernels can read: contiguous when
    the pitch N is already a multiple of 4 floats, else a padded view."""
    return torch.empty(M, N, dtype=torch.float32, device=device) if N % 4 == 0 else dist_buffer(M, N, device)


class _GramCore:
    """Shared forward/backward of the two Gram nodes, operating on fp32 feature matrices.

    ``forward`` returns ``(K, fx, fy, state, extra)``: ``fx`` / ``fy`` are 2-D fp32 tensors whose first
    D columns are the centred features of the rows / columns, ``extra`` is ``()`` on the exact-fp32
    SIMT route and ``(f_hi, f_lo, scales, norms)`` -- the split-fp16 planes of ``[fy | 1]`` (and of
    ``fx`` above them when the two arguments differ) with their row norms -- on the tcgen05 route."""

    @staticmethod
    def forward(sx, sy, same, bw: Bandwidth, center: bool):
        N, D = sx.shape
        if _use_tc(N):
            return _GramCore._forward_tc(sx, sy, same, bw, center)
        mean = col_mean(sx) if center else None
        vx = pack_features(sx, mean)
        vy = vx if same else pack_features(sy, mean)
        K = sqdist_raw(vx, vy, D)
        state = bandwidth_raw(K, bw, N)
        rbf_apply_raw(K, state)
        median_coef_raw(state, bw.mode == _lib.BW_MEDIAN and bw.median_grad)
        return K, vx, vy, state, ()

    @staticmethod
    def _forward_tc(sx, sy, same, bw: Bandwidth, center: bool):
        """tcgen05 route: Gram on the tensor cores (symmetric when both arguments are the same data),
        windowed exact median, exp in place with the statistics of the median-gradient term."""
        N, D = sx.shape
        dev = sx.device
        rows = sx if same else torch.cat([sx, sy], dim=0)
        planes, norms, scales, sc = tc_pack_fused_raw(rows, None, center)
        if same:
            px = py = planes
            nx = ny = norms
            fx = fy = sc
        else:
            px, py = Planes(planes.hi[:N], planes.lo[:N]), Planes(planes.hi[N:], planes.lo[N:])
            nx, ny = norms[:N], norms[N:]
            fx, fy = sc[:N], sc[N:]
        K = tc_gram_raw(px, nx, py, ny, D, scales, symmetric=same, out=_square_buffer(N, N, dev))
        state = new_state(dev)
        if bw.mode == _lib.BW_FIXED:
            call("sigsvgd_fixed_bandwidth", bw.h, _p(state), _stream())
        else:
            ws = tc_median_workspace(N, N, dev)
            tc_median_prepare_raw(fx, fy, ws, state)
            tc_median_select_raw(ws, K, N, bw, state, symmetric=same)
        rbf_apply_raw(K, state)  # in place; accumulates sum K*d and the tie records
        median_coef_raw(state, bw.mode == _lib.BW_MEDIAN and bw.median_grad)
        return K, fx, fy, state, (planes.hi, planes.lo, scales, norms)

    @staticmethod
    def _contract(A, fy_planes: "Planes", scales, D):
        """[A . fy | rowsum(A)] on the tensor cores; A holds kernel values (identity operand map)."""
        out = tc_kf_raw(A, fy_planes, D + 1, scales, None, _lib.KF_IDENTITY)
        return out[:, :D], out[:, D]

    @staticmethod
    def backward(gK, K, vx, vy, state, same, D, need_y, bw: Bandwidth, extra=()):
        """Returns (g_x [N,D], g_y [N,D] or None): gradients wrt the feature matrices."""
        tc = len(extra) == 4
        N = K.shape[0]
        if tc:
            f_hi, f_lo, scales, norms = extra
            px = Planes(f_hi[:N], f_lo[:N])
            py = px if same else Planes(f_hi[N:], f_lo[N:])
        if _is_uniform_cotangent(gK):
            scale = gK[(0,) * gK.dim()].reshape(1).to(torch.float32)  # device scalar, applied at the end
            A, st = K, state
        else:
            # generic cotangent W: weighted kernel A = W o K with its own statistics
            W = _prep(gK, "gram.backward")
            scale = None
            if tc:
                # distances again from the planes (the forward's matrix now holds K)
                A = tc_gram_raw(px, norms[:N], py, norms[:N] if same else norms[N:], D, scales, symmetric=False,
                                out=_square_buffer(N, py.hi.shape[0], K.device))
            else:
                A = sqdist_raw(vx, vy, D)
            st = state.clone()
            st[:, 4:8] = 0  # sum_kd, n_ties, n_ties_all
            rbf_apply_raw(A, st, weight=W)
            median_coef_raw(st, bw.mode == _lib.BW_MEDIAN and bw.median_grad)
        if tc:
            ks, rowsum = _GramCore._contract(A, py, scales, D)
        else:
            out = matmul_nn_raw(A, vy)
            ks, rowsum = out[:, :D], out[:, D]
        gx = sig_grad_raw(vx, vy, ks, rowsum, st)
        gy = None
        if need_y:
            if same and scale is not None:
                gy = gx  # K and the tie set are bit-symmetric (see gemm_simt.cu / gram_tc.cu), so g^Y == g^X
            else:
                At = A.t().contiguous()
                if tc:
                    ks_t, rowsum_t = _GramCore._contract(At, px, scales, D)
                else:
                    out_t = matmul_nn_raw(At, vx)
                    ks_t, rowsum_t = out_t[:, :D], out_t[:, D]
                gy = sig_grad_raw(vy, vx, ks_t, rowsum_t, _swap_ties(st))
        if scale is not None:
            gx = gx * scale
            gy = None if gy is None else (gx if gy is gx else gy * scale)
        return gx, gy


class _SigRBFGram(torch.autograd.Function):
    """K[i,j] = exp(-|S(X_i) - S(Y_j)|^2 / (2 h^2)); PathSigKernel with a Gaussian static kernel."""

    @staticmethod
    def forward(ctx, X, Y, depth, bw, center):
        same = _same_data(X, Y)
        x = _prep(X, "PathSigKernel")
        ctx.set_materialize_grads(False)
        if same and not _use_tc(x.shape[0]) and x.dim() == 3 and x.shape[0] > 0:
            # Small swarm, same data on both sides (every call site of the reference): the regime is
            # launch-bound, so the forward is ONE call of the fused update with a zero score -- it yields
            # K and, at no extra launches, d(sum K)/dX, which is what every caller's backward asks for
            # (``autograd.grad(k_xx.sum(), X)``: svgd.py:43, score.py:55, trajectory_svgd.py:65).
            res = phi(x, torch.zeros(x.shape[0], x.shape[1] * x.shape[2], dtype=torch.float32, device=x.device),
                      depth=depth, bandwidth=bw, want_k=True, want_grad_k=True, center=center)
            ctx.save_for_backward(x, res.grad_k)
            ctx.meta = (depth, bw, True, X.dtype)
            ctx.fused, ctx.center = True, center
            ctx.bw_state = res.state
            K = res.k_xx
            return K if X.dtype == torch.float32 else K.to(X.dtype)
        y = x if same else _prep(Y, "PathSigKernel")
        sx, px = _sig_fwd_saving(x, depth, True)
        sy, py = (sx, px) if same else _sig_fwd_saving(y, depth, True)
        K, vx, vy, state, extra = _GramCore.forward(sx, sy, same, bw, center)
        ctx.save_for_backward(x, y, sx, sy, vx, vy, K, state, px, py, *extra)
        ctx.meta = (depth, bw, same, X.dtype)
        ctx.fused = False
        ctx.bw_state = state
        return K if X.dtype == torch.float32 else K.to(X.dtype)

    @staticmethod
    def backward(ctx, gK):
        if gK is None:
            return None, None, None, None, None
        depth, bw, same, in_dtype = ctx.meta
        need_x, need_y = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        if ctx.fused:
            x, grad_k = ctx.saved_tensors
            if _is_uniform_cotangent(gK):
                scale = gK[(0,) * gK.dim()].reshape(1).to(torch.float32)
                g = (grad_k.reshape(x.shape) * scale).to(in_dtype)
                # K and the tie set are bit-symmetric, so the second-argument gradient equals the first
                return (g if need_x else None), (g if need_y else None), None, None, None
            # generic cotangent: rebuild the pieces of the general route
            sx, px = _sig_fwd_saving(x, depth, True)
            K, vx, vy, state, extra = _GramCore.forward(sx, sx, True, bw, ctx.center)
            y, sy, py = x, sx, px
        else:
            x, y, sx, sy, vx, vy, K, state, px, py, *extra = ctx.saved_tensors
        D = sx.shape[1]
        gsx, gsy = _GramCore.backward(gK, K, vx, vy, state, same, D, need_y, bw, tuple(extra))
        gX = _sig_bwd_saved(x, sx, px, gsx, depth, True).to(in_dtype) if need_x else None
        gY = None
        if need_y:
            if gsy is gsx and gX is not None and same:
                gY = gX
            else:
                gY = _sig_bwd_saved(y, sy, py, gsy, depth, True).to(in_dtype)
        return gX, gY, None, None, None


class _RBFGram(torch.autograd.Function):
    """Static Gaussian kernel on [N, dim] features (src/kernels/_kernels.py:78-115)."""

    @staticmethod
    def forward(ctx, X, Y, bw, center):
        same = _same_data(X, Y)
        sx = _prep(X, "GaussianKernel")
        sy = sx if same else _prep(Y, "GaussianKernel")
        K, vx, vy, state, extra = _GramCore.forward(sx, sy, same, bw, center)
        ctx.save_for_backward(vx, vy, K, state, *extra)
        ctx.meta = (bw, same, X.dtype, sx.shape[1])
        ctx.set_materialize_grads(False)
        ctx.bw_state = state
        return K if X.dtype == torch.float32 else K.to(X.dtype)

    @staticmethod
    def backward(ctx, gK):
        if gK is None:
            return None, None, None, None
        vx, vy, K, state, *extra = ctx.saved_tensors
        bw, same, in_dtype, D = ctx.meta
        gsx, gsy = _GramCore.backward(gK, K, vx, vy, state, same, D, ctx.needs_input_grad[1], bw, tuple(extra))
        gX = gsx.to(in_dtype) if ctx.needs_input_grad[0] else None
        gY = gsy.to(in_dtype) if (ctx.needs_input_grad[1] and gsy is not None) else None
        return gX, gY, None, None


# ---------------------------------------------------------------------------------------------
# signature-PDE kernel: sigkernel.SigKernel.compute_Gram and its backward
# ---------------------------------------------------------------------------------------------
def _prep64(t: torch.Tensor, what: str) -> torch.Tensor:
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise RuntimeError(f"{what}: CUDA tensors only; sigsvgd_b200 has no CPU fallback")
    if t.dim() != 3:
        raise ValueError(f"{what}: expected [batch, length, dim] paths, got {tuple(t.shape)}")
    return t.detach().to(torch.float64).contiguous()


def sigpde_bandwidth_raw(x: torch.Tensor, y: torch.Tensor, bw_scale: float = 1.0, tol: float = 1e-8) -> torch.Tensor:
    """State record with the reference's default bandwidth of ``BatchGaussianKernel.Gram_matrix``: bw_median over
    all A*B*M*N point-pair distances, ln(A + 1).  x, y: float64 [A,M,D], [B,N,D]."""
    (A, M, D), (B, N, _) = x.shape, y.shape
    state = new_state(x.device)
    hist = torch.empty(2048, dtype=torch.int64, device=x.device)
    call("sigsvgd_sigpde_bandwidth", _p(x), _p(y), A, B, M, N, D, float(bw_scale), float(tol), _p(state), _p(hist),
         _stream())
    return state


def sigpde_gram_raw(x, y, order: int, kind: int = _lib.PDE_RBF, h: float = 1.0, state=None) -> torch.Tensor:
    (A, M, D), (B, N, D2) = x.shape, y.shape
    if D != D2:
        raise ValueError("sigpde_gram: paths of different dimension")
    k = torch.empty(A, B, dtype=torch.float64, device=x.device)
    call("sigsvgd_sigpde_gram", _p(x), _p(y), A, B, M, N, D, int(order), int(kind), float(h), _p(state), _p(k), _stream())
    return k


def sigpde_gram_bwd_raw(x, y, gout, order: int, kind: int = _lib.PDE_RBF, h: float = 1.0, state=None) -> torch.Tensor:
    (A, M, D), (B, N, _) = x.shape, y.shape
    nbytes = int(_lib.load().sigsvgd_sigpde_bwd_workspace_bytes(A, B, M, N, D, int(order)))
    ws = torch.empty(max(nbytes, 8), dtype=torch.uint8, device=x.device)
    gx = torch.empty_like(x)
    call("sigsvgd_sigpde_gram_bwd", _p(x), _p(y), A, B, M, N, D, int(order), int(kind), float(h), _p(state), _p(gout),
         _p(gx), _p(ws), ws.numel(), _stream())
    return gx


class _SigPdeGram(torch.autograd.Function):
    """k[a,b] = Goursat-PDE signature kernel of paths X_a, Y_b; gradient w.r.t. X only (as the package)."""

    @staticmethod
    def forward(ctx, X, Y, order, kind, h, bw_scale, tol):
        x, y = _prep64(X, "SigKernel.compute_Gram"), _prep64(Y, "SigKernel.compute_Gram")
        state = sigpde_bandwidth_raw(x, y, bw_scale, tol) if (h is None and kind == _lib.PDE_RBF) else None
        k = sigpde_gram_raw(x, y, order, kind, 1.0 if h is None else h, state)
        ctx.saved = (x, y, state)
        ctx.cfg = (order, kind, 1.0 if h is None else h, X.dtype)
        ctx.state = state
        return k.to(X.dtype)

    @staticmethod
    def backward(ctx, gout):
        x, y, state = ctx.saved
        order, kind, h, dtype = ctx.cfg
        gx = None
        if ctx.needs_input_grad[0]:
            gx = sigpde_gram_bwd_raw(x, y, gout.to(torch.float64).contiguous(), order, kind, h, state).to(dtype)
        return gx, None, None, None, None, None, None


def sigpde_gram(X, Y, order: int = 0, kind: int = _lib.PDE_RBF, h: Optional[float] = None, bw_scale: float = 1.0,
                tol: float = 1e-8) -> torch.Tensor:
    """Signature-PDE Gram matrix [A,B] of X [A,M,D] and Y [B,N,D] (any float dtype; computed in float64 as
    the reference does), differentiable w.r.t. X.  ``h=None`` (RBF): the reference's median heuristic over
    all point pairs; else ``exp(-d/h)`` (RBF) or ``<x,y> * h`` (linear)."""
    return _SigPdeGram.apply(X, Y, int(order), int(kind), None if h is None else float(h), float(bw_scale), float(tol))


def sig_rbf_gram(X, Y, depth: int = 3, bandwidth: Optional[Bandwidth] = None, center: bool = True) -> torch.Tensor:
    if X.shape != Y.shape:
        raise AssertionError("X and Y must have the same dimensions.")
    return _SigRBFGram.apply(X, Y, int(depth), bandwidth or Bandwidth.median(), bool(center))


def rbf_gram(X, Y, bandwidth: Optional[Bandwidth] = None, center: bool = True) -> torch.Tensor:
    if X.shape != Y.shape:
        raise AssertionError("X and Y must have the same dimensions.")
    return _RBFGram.apply(X, Y, bandwidth or Bandwidth.median(), bool(center))


def rbf_gram_with_grad(X, Y, bandwidth: Optional[Bandwidth] = None, center: bool = True):
    """(K, sum_j dK_ij/dx_i) -- the analytic-gradient branch of GaussianKernel
    (src/kernels/_kernels.py:111-113): d_K.sum(1) = -(1/h^2) sum_j K_ij (x_i - y_j), constant h."""
    bw = bandwidth or Bandwidth.median()
    same = _same_data(X, Y)
    sx = _prep(X, "GaussianKernel")
    sy = sx if same else _prep(Y, "GaussianKernel")
    D = sx.shape[1]
    bw_nograd = Bandwidth(bw.mode, bw.h, bw.bw_scale, bw.tol, False)
    K, vx, vy, state, extra = _GramCore.forward(sx, sy, same, bw_nograd, center)
    if len(extra) == 4:
        N = K.shape[0]
        py = Planes(extra[0][:N], extra[1][:N]) if same else Planes(extra[0][N:], extra[1][N:])
        ks, rowsum = _GramCore._contract(K, py, extra[2], D)
    else:
        out = matmul_nn_raw(K, vy)
        ks, rowsum = out[:, :D], out[:, D]
    dK = sig_grad_raw(vx, vy, ks, rowsum, state)
    return K.to(X.dtype), dK.to(X.dtype), state


# ---------------------------------------------------------------------------------------------
# fused Stein update
# ---------------------------------------------------------------------------------------------
_WS_CACHE: dict = {}
_HOST_CACHE: dict = {}  # phi_host: (device, shape) -> (state record, workspace)


def _workspace(device, nbytes: int) -> torch.Tensor:
    key = (device.index, torch.cuda.current_stream(device).cuda_stream)
    ws = _WS_CACHE.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = None
        _WS_CACHE.pop(key, None)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=device)
        _WS_CACHE[key] = ws
    return ws


@dataclass
class PhiResult:
    phi: torch.Tensor  # [.., N, T, d]  Stein direction (K.score - grad_k)/N ; reference velocity = -phi
    state: torch.Tensor  # device record(s): median, h, ties ...
    k_xx: Optional[torch.Tensor] = None
    grad_k: Optional[torch.Tensor] = None
    sig: Optional[torch.Tensor] = None

    @property
    def h(self) -> torch.Tensor:
        return state_field(self.state, "h")


def phi(X: torch.Tensor, score: torch.Tensor, depth: int = 3, bandwidth: Optional[Bandwidth] = None,
        want_k: bool = False, want_grad_k: bool = False, want_sig: bool = False, center: bool = True,
        flags: int = 0) -> PhiResult:
    """Fused update for one swarm ``X [N,T,d]`` or a batch of independent swarms ``X [B,N,T,d]``.

    ``score`` = grad_log_p (+ grad_log_prior), any shape that flattens to ``[.., N, T*d]``."""
    bw = bandwidth or Bandwidth.median()
    x = _prep(X, "phi")
    batched = x.dim() == 4
    if x.dim() not in (3, 4):
        raise ValueError(f"phi: X must be [N,T,d] or [B,N,T,d], got {tuple(x.shape)}")
    B = x.shape[0] if batched else 1
    N, T, d = x.shape[-3:]
    g = _prep(score, "phi").reshape(B, N, -1)
    if g.shape[-1] != T * d:
        raise ValueError(f"phi: score has {g.shape[-1]} columns per particle, expected T*d = {T * d}")
    lib = _lib.load()
    D = sig_channels(d, depth)
    dev = x.device
    out = torch.empty_like(x)
    k_out = torch.empty(B, N, N, dtype=torch.float32, device=dev) if want_k else None
    gk_out = torch.empty(B, N, T * d, dtype=torch.float32, device=dev) if want_grad_k else None
    sig_out = torch.empty(B, N, D, dtype=torch.float32, device=dev) if want_sig else None
    state = new_state(dev, B)
    nbytes = lib.sigsvgd_phi_workspace_bytes(B, N, T, d, depth, int(want_k))
    if nbytes == 0:
        raise ValueError("phi: bad shape")
    ws = _workspace(dev, nbytes)
    fl = int(flags) | (_lib.FLAG_MEDIAN_GRAD if (bw.mode == _lib.BW_MEDIAN and bw.median_grad) else 0)
    if not center:
        fl |= _lib.FLAG_NO_CENTER
    call("sigsvgd_phi", _p(x), _p(g), B, N, T, d, depth, bw.mode, bw.h, bw.bw_scale, bw.tol, fl, _p(out), _p(k_out),
         _p(gk_out), _p(sig_out), _p(state), _p(ws), ws.numel(), _stream())
    if not batched:
        k_out = None if k_out is None else k_out[0]
        gk_out = None if gk_out is None else gk_out[0]
        sig_out = None if sig_out is None else sig_out[0]
    return PhiResult(out.to(X.dtype), state, k_out, gk_out, sig_out)


class PhiPlan:
    """The fused update of one swarm shape, captured ONCE into a CUDA graph and replayed.

    Every launch of ``sigsvgd_phi`` is graph-capturable (the library neither allocates nor
    synchronises), so an optimisation loop that keeps its particles in one buffer -- the reference's
    ``SVGD.optimize`` (src/inference/svgd.py:136-159) -- can skip the per-step enqueue work and the
    launch-to-launch gaps: 278 -> 264 us per update at N=4096, 115 -> 102 us at N=256 on a B200.

    The plan owns static buffers: write the inputs into ``plan.x`` / ``plan.score`` (or pass tensors to
    ``__call__``, which copies them in) and read ``plan.phi`` / ``plan.state`` after the replay.
    """

    def __init__(self, N: int, T: int, d: int, depth: int = 3, bandwidth: Optional[Bandwidth] = None,
                 center: bool = True, flags: int = 0, device=None, batch: Optional[int] = None):
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        if dev.type != "cuda":
            raise ValueError("PhiPlan: a CUDA device is required (there is no CPU path)")
        bw = bandwidth or Bandwidth.median()
        B = 1 if batch is None else int(batch)
        lead = () if batch is None else (B,)
        lib = _lib.load()
        self.shape = (B, N, T, d, depth)
        # placeholder inputs for the warm-up / capture run: generic random walks (not a degenerate swarm)
        gen = torch.Generator().manual_seed(0)
        self.x = torch.cumsum(0.05 * torch.randn(*lead, N, T, d, generator=gen), dim=-2).to(dev)
        self.score = torch.randn(*lead, N, T * d, generator=gen).to(dev)
        self.phi = torch.empty_like(self.x)
        self.state = new_state(dev, B)
        nbytes = lib.sigsvgd_phi_workspace_bytes(B, N, T, d, depth, 0)
        if nbytes == 0:
            raise ValueError("PhiPlan: bad shape")
        self._ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)  # private: the graph bakes its address
        fl = int(flags) | (_lib.FLAG_MEDIAN_GRAD if (bw.mode == _lib.BW_MEDIAN and bw.median_grad) else 0)
        if not center:
            fl |= _lib.FLAG_NO_CENTER

        def enqueue():
            call("sigsvgd_phi", _p(self.x), _p(self.score), B, N, T, d, depth, bw.mode, bw.h, bw.bw_scale, bw.tol, fl,
                 _p(self.phi), None, None, None, _p(self.state), _p(self._ws), self._ws.numel(), _stream())

        self._stream = torch.cuda.Stream(device=dev)
        self._stream.wait_stream(torch.cuda.current_stream(dev))
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.stream(self._stream):
            enqueue()  # warm-up outside the capture: lazy driver-side initialisation (attributes, side stream)
            self._stream.synchronize()
            with torch.cuda.graph(self.graph, stream=self._stream, capture_error_mode="thread_local"):
                enqueue()
        torch.cuda.current_stream(dev).wait_stream(self._stream)

    def __call__(self, X: Optional[torch.Tensor] = None, score: Optional[torch.Tensor] = None) -> "PhiResult":
        if X is not None and X.data_ptr() != self.x.data_ptr():
            self.x.copy_(X.reshape(self.x.shape))
        if score is not None and score.data_ptr() != self.score.data_ptr():
            self.score.copy_(score.reshape(self.score.shape))
        self.graph.replay()
        return PhiResult(self.phi, self.state)


def phi_host(X: torch.Tensor, score: torch.Tensor, depth: int = 3, bandwidth: Optional[Bandwidth] = None,
             out: Optional[torch.Tensor] = None, center: bool = True, flags: int = 0,
             device: Optional[torch.device] = None) -> PhiResult:
    """Fused update for HOST-resident inputs: ``X [N,T,d]``, ``score [N,T*d]`` fp32 CPU tensors
    (pinned memory makes the copies asynchronous), result in ``out`` (CPU, pinned if given that
    way).  One ``sigsvgd_phi_host`` call: H2D of X, the score's H2D beside the front half of the
    update, chunked D2H of phi beside the signature backward.  The result is complete once the
    current stream of ``device`` has finished (``torch.cuda.current_stream().synchronize()``).  The
    returned ``state`` record is reused by the next ``phi_host`` call of the same shape on that device."""
    bw = bandwidth or Bandwidth.median()
    if X.is_cuda or score.is_cuda:
        raise RuntimeError("phi_host takes CPU tensors; use ops.phi for device-resident inputs")
    if X.dim() != 3:
        raise ValueError(f"phi_host: X must be [N,T,d], got {tuple(X.shape)}")
    dev = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
    x = X.detach().to(torch.float32).contiguous()
    N, T, d = x.shape
    g = score.detach().to(torch.float32).reshape(N, -1).contiguous()
    if g.shape[1] != T * d:
        raise ValueError(f"phi_host: score has {g.shape[1]} columns per particle, expected T*d = {T * d}")
    if out is None:
        out = torch.empty(N, T, d, dtype=torch.float32).pin_memory()
    if out.is_cuda or out.dtype != torch.float32 or not out.is_contiguous() or out.shape != x.shape:
        raise ValueError("phi_host: out must be a contiguous fp32 CPU tensor shaped like X")
    lib = _lib.load()
    key = (dev.index, N, T, d, depth)
    cached = _HOST_CACHE.get(key)
    if cached is None:
        nbytes = lib.sigsvgd_phi_host_workspace_bytes(N, T, d, depth)
        if nbytes == 0:
            raise ValueError("phi_host: bad shape")
        with torch.cuda.device(dev):
            # one state record and one workspace per (device, shape), reused by later calls: the pipeline
            # rewrites every field of the record, and calls on one device are stream-ordered
            cached = _HOST_CACHE[key] = (new_state(dev, 1), torch.empty(nbytes, dtype=torch.uint8, device=dev))
    state, ws = cached
    fl = int(flags) | (_lib.FLAG_MEDIAN_GRAD if (bw.mode == _lib.BW_MEDIAN and bw.median_grad) else 0)
    if not center:
        fl |= _lib.FLAG_NO_CENTER
    _note_device(dev)
    call("sigsvgd_phi_host", x.data_ptr(), g.data_ptr(), N, T, d, depth, bw.mode, bw.h, bw.bw_scale, bw.tol, fl,
         out.data_ptr(), _p(state), _p(ws), ws.numel(), _stream())
    res = PhiResult(out, state)
    res._keep = (x, g)  # the asynchronous copies read these host buffers until the stream drains
    return res


def log_np1(n: int) -> float:
    return math.log(n + 1.0)

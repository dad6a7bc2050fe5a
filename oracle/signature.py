"""ctypes binding of oracle/libsig_oracle.so + a torch.autograd wrapper.  TEST INFRASTRUCTURE.

Restates ``signatory.signature(path, depth, basepoint=True)`` as called at
/root/reference/src/kernels/_traj_kernels.py:125-126 (signatory==1.2.6.1.9.0, third party,
absent).  "parity unpinned" for the transform itself -- see sig_oracle.c.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def _lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "libsig_oracle.so")
        if not os.path.exists(path):
            subprocess.check_call(["make", "-C", _HERE, "-s"])
        lib = ctypes.CDLL(path)
        i64, i32, vp = ctypes.c_int64, ctypes.c_int, ctypes.c_void_p
        lib.sig_oracle_channels.restype = i64
        lib.sig_oracle_channels.argtypes = [i64, i32]
        lib.sig_oracle_max_threads.restype = i32
        for sfx in ("f32", "f64"):
            f = getattr(lib, f"sig_oracle_fwd_{sfx}")
            f.restype = i32
            f.argtypes = [vp, i64, i64, i64, i32, i32, vp]
            b = getattr(lib, f"sig_oracle_bwd_{sfx}")
            b.restype = i32
            b.argtypes = [vp, vp, vp, i64, i64, i64, i32, i32, vp]
        _LIB = lib
    return _LIB


def sig_channels(d: int, depth: int) -> int:
    return sum(d ** k for k in range(1, depth + 1))


def max_threads() -> int:
    return int(_lib().sig_oracle_max_threads())


def _sfx(dtype):
    if dtype == np.float32:
        return "f32"
    if dtype == np.float64:
        return "f64"
    raise TypeError(f"oracle signature supports float32/float64, got {dtype}")


def signature_np(x: np.ndarray, depth: int, basepoint: bool = True) -> np.ndarray:
    """x [N,T,d] -> [N,D]; D = sum_k d^k (levels 1..depth, level-major, row-major, no scalar)."""
    x = np.ascontiguousarray(x)
    assert x.ndim == 3
    N, T, d = x.shape
    out = np.empty((N, sig_channels(d, depth)), dtype=x.dtype)
    rc = getattr(_lib(), f"sig_oracle_fwd_{_sfx(x.dtype)}")(
        x.ctypes.data, N, T, d, depth, int(bool(basepoint)), out.ctypes.data
    )
    if rc != 0:
        raise ValueError("sig_oracle_fwd: bad arguments")
    return out


def signature_bwd_np(x, sig, gsig, depth: int, basepoint: bool = True) -> np.ndarray:
    x = np.ascontiguousarray(x)
    sig = np.ascontiguousarray(sig, dtype=x.dtype)
    gsig = np.ascontiguousarray(gsig, dtype=x.dtype)
    N, T, d = x.shape
    gx = np.empty_like(x)
    rc = getattr(_lib(), f"sig_oracle_bwd_{_sfx(x.dtype)}")(
        x.ctypes.data, sig.ctypes.data, gsig.ctypes.data, N, T, d, depth, int(bool(basepoint)), gx.ctypes.data
    )
    if rc != 0:
        raise ValueError("sig_oracle_bwd: bad arguments")
    return gx


try:  # torch wrapper (autograd-composable, CPU tensors only)
    import torch

    class OracleSignature(torch.autograd.Function):
        @staticmethod
        def forward(ctx, path, depth, basepoint):
            x = path.detach().contiguous()
            sig = torch.from_numpy(signature_np(x.numpy(), depth, basepoint))
            ctx.save_for_backward(x, sig)
            ctx.depth, ctx.basepoint = depth, basepoint
            return sig

        @staticmethod
        def backward(ctx, g):
            x, sig = ctx.saved_tensors
            gx = signature_bwd_np(x.numpy(), sig.numpy(), g.detach().contiguous().numpy(), ctx.depth, ctx.basepoint)
            return torch.from_numpy(gx), None, None

    def signature_torch(path, depth, stream=False, basepoint=False, inverse=False, initial=None, scalar_term=False):
        """Same call signature as ``signatory.signature`` (only the options the reference uses)."""
        if stream or inverse or initial is not None or scalar_term:
            raise NotImplementedError("oracle signature: only (path, depth, basepoint) are restated")
        if not isinstance(basepoint, bool):
            path = torch.cat([basepoint.to(path).expand(path.shape[0], 1, path.shape[2]), path], dim=1)
            basepoint = False
        return OracleSignature.apply(path, int(depth), bool(basepoint))

except ImportError:  # pragma: no cover
    OracleSignature = None
    signature_torch = None


def signature_pure_torch(path, depth, basepoint=True):
    """Slow, autograd-able, any-dtype Chen restatement in plain torch ops (independent of the C
    code; used to cross-check it).  S^k <- sum_{i=0..k} S^i (x) Delta^{(x)(k-i)}/(k-i)!"""
    import torch

    N, T, d = path.shape
    if basepoint:
        path = torch.cat([torch.zeros(N, 1, d, dtype=path.dtype), path], 1)
    inc = path[:, 1:] - path[:, :-1]
    S = [torch.zeros(N, d ** k, dtype=path.dtype) for k in range(1, depth + 1)]
    for t in range(inc.shape[1]):
        dl = inc[:, t]
        E = [dl]
        for k in range(2, depth + 1):
            E.append((E[-1][:, :, None] * dl[:, None, :]).reshape(N, -1) / k)
        newS = []
        for k in range(1, depth + 1):
            acc = S[k - 1] + E[k - 1]
            for i in range(1, k):
                acc = acc + (S[i - 1][:, :, None] * E[k - i - 1][:, None, :]).reshape(N, -1)
            newS.append(acc)
        S = newS
    return torch.cat(S, dim=1)

"""CPU restatement of the reference's signature-kernel SVGD update.  TEST INFRASTRUCTURE.

Two forms:

* ``phi_torch_literal`` -- the reference's own torch statements, line by line, with the oracle
  signature in place of ``signatory.signature`` (autograd through exp / median / addmm /
  signature, exactly the route ``SVGD._compute_kernel`` takes).  This is the timed CPU baseline
  (``cpu_baseline.kind == "port"``).
* ``phi_closed_form`` -- numpy, any dtype, the closed form of SURVEY.md Appendix A including the
  gradient through ``torch.median``; returns the tie set so tests can apply the parity protocol
  of SURVEY.md 8(a-4).

Golden vectors produced by the unmodified reference (``oracle/make_golden.py``) pin both.
"""
from __future__ import annotations

import math

import numpy as np

from .signature import signature_np, signature_bwd_np


# ---------------------------------------------------------------------------------------------
# numpy closed form
# ---------------------------------------------------------------------------------------------
def pw_dist_sq(m1: np.ndarray, m2: np.ndarray) -> np.ndarray:
    """/root/reference/src/utils/math.py:69-86 -- ||a||^2 + ||b||^2 - 2 a.b, clamped at 0."""
    n1 = (m1 * m1).sum(-1, keepdims=True)
    n2 = (m2 * m2).sum(-1, keepdims=True)
    res = n2.T + (-2.0) * (m1 @ m2.T)
    res = res + n1
    return np.maximum(res, 0)


def lower_median(a: np.ndarray):
    """torch.median on a flattened tensor = sorted element (n-1)//2 (lower median)."""
    flat = a.reshape(-1)
    k = (flat.size - 1) // 2
    return np.partition(flat, k)[k]


def bw_median(sq: np.ndarray, bw_scale: float = 1.0, tol: float = 1e-8):
    """/root/reference/src/utils/math.py:28-34 -- h = sqrt(median / ln(N+1)) * bw_scale, >= tol."""
    med = lower_median(sq)
    # NB: the reference evaluates ln(N+1) on a default-dtype (float32) 0-dim tensor even when the
    # distances are float64 (math.py:32) -- restated as such.
    L = sq.dtype.type(np.log(np.float32(sq.shape[0] + 1.0)))
    h = bw_scale * np.sqrt(med / L)
    return max(h, sq.dtype.type(tol)), med


def gaussian_gram(sx: np.ndarray, sy: np.ndarray, h=None, bw_scale: float = 1.0):
    """/root/reference/src/kernels/_kernels.py:102-110 -- returns (K, sq_dists, h, median)."""
    sq = pw_dist_sq(sx, sy)
    med = None
    if h is None:
        h, med = bw_median(sq, bw_scale)
    K = np.exp(-0.5 / h ** 2 * sq)
    return K, sq, h, med


def sig_space_grad(sx, sy, K, sq, h, med, median_grad=True, bw_scale: float = 1.0, tol: float = 1e-8):
    """d(sum K)/d sx, shape [N,D]  (SURVEY.md Appendix A step 6).

    constant-h part : g_i = -(1/h^2) sum_j K_ij (sx_i - sy_j)
    median part     : + (c/|T|) (sx_a - sy_b) for (a,b) in T = {d_ab == median},
                      c = sum_ij K_ij d_ij / (h^2 * median)   (= ln(N+1)/median^2 * sum K d for bw_scale 1)
    Returns (g, ties) with ties an int array [n_ties, 2].
    """
    g = -(sx * K.sum(1, keepdims=True) - K @ sy) / h ** 2
    ties = np.zeros((0, 2), dtype=np.int64)
    if median_grad and med is not None and h > tol and med > 0:
        ties = np.argwhere(sq == med)
        c = (K * sq).sum() / (h ** 2 * med)
        for a, b in ties:
            g[a] += (c / len(ties)) * (sx[a] - sy[b])
    return g, ties


def phi_closed_form(X, G, depth=3, h=None, bw_scale=1.0, median_grad=True):
    """Full update.  X [N,T,d], G [N,T*d] (= grad_log_p.flatten(1)).  Returns a dict with
    sig, sq, K, h, median, g (sig-space grad), grad_k [N,P], phi [N,T,d], ties.

    phi = (K G - grad_k) / N   -- north-star sign; the reference returns ``-phi``
    (/root/reference/src/inference/svgd.py:82-83)."""
    X = np.ascontiguousarray(X)
    N, T, d = X.shape
    S = signature_np(X, depth, True)
    K, sq, h_, med = gaussian_gram(S, S, h, bw_scale)
    g, ties = sig_space_grad(S, S, K, sq, h_, med, median_grad and h is None, bw_scale)
    grad_k = signature_bwd_np(X, S, g.astype(X.dtype), depth, True).reshape(N, -1)
    phi = (K @ G.reshape(N, -1) - grad_k) / N
    return dict(sig=S, sq=sq, K=K, h=h_, median=med, g=g, grad_k=grad_k, phi=phi.reshape(X.shape), ties=ties)


# ---------------------------------------------------------------------------------------------
# torch literal restatement (what the reference executes)
# ---------------------------------------------------------------------------------------------
def phi_torch_literal(X, G, depth=3, h=None, signature_fn=None):
    """Line-by-line torch restatement.  Returns dict(k_xx, grad_k, velocity, h).

    velocity is what ``SVGD._velocity`` returns, i.e. ``-phi``.
    """
    import torch

    if signature_fn is None:
        from .signature import signature_torch as signature_fn

    def _pw_dist_sq(mat1, mat2):  # math.py:81-86
        mat1_norm = mat1.pow(2).sum(dim=-1, keepdim=True)
        mat2_norm = mat2.pow(2).sum(dim=-1, keepdim=True)
        res = torch.addmm(mat2_norm.transpose(-2, -1), mat1, mat2.transpose(-2, -1), alpha=-2).add_(mat1_norm)
        return res.clamp(min=0)

    def _bw_median(sq_dists, bw_scale=1.0, tol=1.0e-8):  # math.py:28-34
        hh = torch.median(sq_dists)
        hh = hh / torch.tensor(sq_dists.shape[0] + 1.0).log()
        hh = bw_scale * hh.sqrt()
        return hh.clamp_min_(tol)

    Xr = X.detach().requires_grad_(True)  # svgd.py:41
    Y = Xr.detach()
    X_sig = signature_fn(Xr, depth, basepoint=True)  # _traj_kernels.py:125
    Y_sig = signature_fn(Y, depth, basepoint=True)  # _traj_kernels.py:126
    Xs, Ys = X_sig.flatten(1), Y_sig.flatten(1)  # _kernels.py:104
    sq_dists = _pw_dist_sq(Xs, Ys)  # _kernels.py:105
    hh = _bw_median(sq_dists) if h is None else float(h)  # _kernels.py:106-109
    k_xx = (-0.5 / hh ** 2 * sq_dists).exp()  # _kernels.py:110
    grad_k = torch.autograd.grad(k_xx.sum(), Xr)[0].flatten(1)  # svgd.py:43
    k_xx, grad_k = k_xx.detach(), grad_k.detach()
    score = G.flatten(1)  # svgd.py:69
    velocity = (k_xx @ score - grad_k) / X.shape[0]  # svgd.py:82
    velocity = -velocity.reshape(X.shape)  # svgd.py:83
    return dict(k_xx=k_xx, grad_k=grad_k, velocity=velocity, h=float(hh.detach()) if hasattr(hh, 'detach') else float(hh), sq_dists=sq_dists.detach())


def ln_np1(N: int) -> float:
    return math.log(N + 1.0)

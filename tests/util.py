"""Shared helpers for the parity tests (checker side only: may import oracle/)."""
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")

# Tolerances (north star: "max relative error 1e-4 on signatures, K and phi, identical median
# bandwidth"); all scored against the float64 oracle / float64 reference golden vectors.
TOL_SIG = 1e-5   # max|dS| / max|S|
TOL_H = 1e-5     # relative
TOL_K = 1e-4     # max|dK|, K in [0,1]
TOL_PHI = 1e-4   # max|dphi| / max|phi|


def load_golden(name):
    return dict(np.load(os.path.join(GOLDEN, name + ".npz")))


def rel_max(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))


def make_paths(kind, N, T, d, seed=0):
    """Synthetic generators of SURVEY.md 8(d) (numpy, float32)."""
    rng = np.random.default_rng(seed)
    if kind == "walk":  # config 4
        return np.cumsum(0.05 * rng.standard_normal((N, T, d)), axis=1).astype(np.float32)
    if kind == "field":  # config 1
        w = np.linspace(0, 1, T)[:, None]
        line = -1.0 + 2.0 * w * np.ones(d)
        return (line[None] + 0.4 * rng.standard_normal((N, T, d))).astype(np.float32)
    if kind == "panda":  # config 2
        b = np.array([2.9671, 1.8326, 2.9671, 0.0873, 2.9671, 3.8223, 2.9671])[:d]
        q0 = (rng.random((N, 1, d)) * 2 - 1) * b
        q1 = (rng.random((N, 1, d)) * 2 - 1) * b
        w = np.linspace(0, 1, T)[None, :, None]
        return (0.8 * (q0 + w * (q1 - q0)) + 0.2 * (2 * rng.random((N, T, d)) - 1) * b).astype(np.float32)
    raise ValueError(kind)


def closed_form_with_ties(X64, G64, depth, ties, h=None, bw_scale=1.0):
    """float64 closed form of the update, with the median-gradient pair(s) given by the
    implementation under test (parity protocol of SURVEY.md 8(a-4): the median VALUE is stable,
    WHICH pair attains it is not)."""
    import oracle

    N = X64.shape[0]
    S = oracle.signature_np(X64, depth, True)
    K, sq, h_, med = oracle.gaussian_gram(S, S, h, bw_scale)
    g = -(S * K.sum(1, keepdims=True) - K @ S) / h_ ** 2
    if h is None and len(ties) > 0 and med > 0:
        c = (K * sq).sum() / (h_ ** 2 * med)
        for a, b in ties:
            g[a] += (c / len(ties)) * (S[a] - S[b])
    grad_k = oracle.signature_bwd_np(X64, S, g, depth, True).reshape(N, -1)
    phi = (K @ G64.reshape(N, -1) - grad_k) / N
    return dict(sig=S, K=K, sq=sq, h=h_, median=med, g=g, grad_k=grad_k, phi=phi.reshape(X64.shape))

"""Known-answer tests that pin the CPU oracle of the signature transform.

The reference holds no golden vectors for ``signatory.signature`` and the package is absent
("parity unpinned"), so the restatement in oracle/sig_oracle.c is pinned by closed forms only:
KAT-1 level 1 = last point; KAT-2 level-2 closed form; KAT-3 shuffle identities; KAT-4 straight
line = tensor exponential; KAT-5 invariance to repeated points / re-parametrisation; KAT-6
channel counts; KAT-7 Chen's identity and time reversal; plus an independent pure-torch Chen
restatement (forward, and autograd for the backward) and fp64 finite differences.
"""
import math

import numpy as np
import pytest
import torch

import oracle
from oracle.signature import signature_pure_torch


def _paths(N, T, d, seed=0, dtype=np.float64):
    rng = np.random.default_rng(seed)
    return np.cumsum(0.3 * rng.standard_normal((N, T, d)), axis=1).astype(dtype)


def _levels(S, d, depth):
    out, o = [], 0
    for k in range(1, depth + 1):
        out.append(S[:, o:o + d ** k].reshape((S.shape[0],) + (d,) * k))
        o += d ** k
    return out


def _tensor_mul(A, B, d, depth):
    """Truncated tensor-algebra product of two group-like elements given WITHOUT scalar term."""
    N = A[0].shape[0]
    out = []
    for k in range(1, depth + 1):
        acc = A[k - 1] + B[k - 1]
        for i in range(1, k):
            a = A[i - 1].reshape(N, -1)
            b = B[k - i - 1].reshape(N, -1)
            acc = acc + (a[:, :, None] * b[:, None, :]).reshape(acc.shape)
        out.append(acc)
    return out


def test_kat6_channel_counts():
    assert oracle.sig_channels(7, 3) == 399
    assert oracle.sig_channels(7, 4) == 2800
    assert oracle.sig_channels(2, 3) == 14
    x = _paths(3, 5, 7)
    assert oracle.signature_np(x, 3).shape == (3, 399)


@pytest.mark.parametrize("d,depth", [(2, 3), (3, 4), (7, 3), (5, 2), (1, 5)])
def test_kat1_kat2_levels_closed_form(d, depth):
    x = _paths(6, 9, d, seed=1)
    S = oracle.signature_np(x, depth, True)
    lv = _levels(S, d, depth)
    # KAT-1: with a zero basepoint level 1 is the last point
    np.testing.assert_allclose(lv[0], x[:, -1], rtol=0, atol=1e-12)
    if depth >= 2:
        inc = np.diff(np.concatenate([np.zeros_like(x[:, :1]), x], 1), axis=1)
        pre = np.cumsum(inc, 1) - inc  # sum_{s<t} Delta_s
        s2 = np.einsum("nti,ntj->nij", pre, inc) + 0.5 * np.einsum("nti,ntj->nij", inc, inc)
        np.testing.assert_allclose(lv[1], s2, rtol=0, atol=1e-12)


def test_kat3_shuffle_identities():
    d, depth = 3, 3
    x = _paths(5, 11, d, seed=2)
    s1, s2, s3 = _levels(oracle.signature_np(x, depth, True), d, depth)
    np.testing.assert_allclose(np.einsum("ni,nj->nij", s1, s1), s2 + s2.transpose(0, 2, 1), atol=1e-12)
    lhs = np.einsum("ni,njk->nijk", s1, s2)
    rhs = s3 + s3.transpose(0, 2, 1, 3) + s3.transpose(0, 3, 1, 2)  # ijk + jik + jki
    np.testing.assert_allclose(lhs, rhs, atol=1e-12)


@pytest.mark.parametrize("d,depth", [(2, 5), (7, 3), (3, 4)])
def test_kat4_straight_line_is_tensor_exponential(d, depth):
    rng = np.random.default_rng(3)
    v = rng.standard_normal((4, d))
    T = 6
    x = v[:, None, :] * np.linspace(1.0 / T, 1.0, T)[None, :, None]  # zero basepoint -> v in T steps
    lv = _levels(oracle.signature_np(x, depth, True), d, depth)
    cur = v
    for k in range(1, depth + 1):
        np.testing.assert_allclose(lv[k - 1].reshape(4, -1), cur.reshape(4, -1) / math.factorial(k), atol=1e-12)
        cur = (cur.reshape(4, -1)[:, :, None] * v[:, None, :]).reshape(4, -1)


def test_kat5_repeated_point_and_reparametrisation():
    d, depth = 3, 4
    x = _paths(4, 7, d, seed=4)
    S = oracle.signature_np(x, depth, True)
    rep = np.concatenate([x[:, :3], x[:, 2:3], x[:, 3:]], 1)  # a repeated point = zero increment
    np.testing.assert_allclose(oracle.signature_np(rep, depth, True), S, atol=1e-13)
    mid = 0.5 * (x[:, 2:3] + x[:, 3:4])  # a point inserted ON a segment does not change the path
    ins = np.concatenate([x[:, :3], mid, x[:, 3:]], 1)
    np.testing.assert_allclose(oracle.signature_np(ins, depth, True), S, atol=1e-12)


def test_kat7_chen_and_time_reversal():
    d, depth = 3, 4
    x = _paths(3, 10, d, seed=5)
    S = _levels(oracle.signature_np(x, depth, True), d, depth)
    a = 4
    A = _levels(oracle.signature_np(x[:, :a], depth, True), d, depth)
    # second piece starts at x[a-1]: no basepoint, include the joining point
    B = _levels(oracle.signature_np(x[:, a - 1:], depth, False), d, depth)
    AB = _tensor_mul(A, B, d, depth)
    for k in range(depth):
        np.testing.assert_allclose(AB[k], S[k], atol=1e-12)
    # time reversal gives the tensor inverse: S(x) (x) S(reversed x) = 1
    full = np.concatenate([np.zeros_like(x[:, :1]), x], 1)
    R = _levels(oracle.signature_np(full[:, ::-1].copy(), depth, False), d, depth)
    F = _levels(oracle.signature_np(full, depth, False), d, depth)
    for lvl in _tensor_mul(F, R, d, depth):
        np.testing.assert_allclose(lvl, 0.0, atol=1e-12)


@pytest.mark.parametrize("N,T,d,depth,basepoint", [(5, 8, 2, 3, True), (3, 6, 7, 3, True), (4, 5, 3, 4, False),
                                                   (2, 1, 3, 3, True), (3, 2, 2, 6, False)])
def test_c_oracle_matches_independent_torch_restatement(N, T, d, depth, basepoint):
    x = _paths(N, T, d, seed=6)
    S = oracle.signature_np(x, depth, basepoint)
    xt = torch.from_numpy(x).requires_grad_(True)
    St = signature_pure_torch(xt, depth, basepoint)
    np.testing.assert_allclose(S, St.detach().numpy(), atol=1e-12)
    # backward: C reversible sweep vs autograd through the torch restatement
    g = np.random.default_rng(7).standard_normal(S.shape)
    (gx_t,) = torch.autograd.grad(St, xt, torch.from_numpy(g))
    gx = oracle.signature_bwd_np(x, S, g, depth, basepoint)
    np.testing.assert_allclose(gx, gx_t.numpy(), rtol=1e-9, atol=1e-10)


def test_backward_matches_fp64_finite_differences():
    N, T, d, depth = 2, 5, 3, 3
    x = _paths(N, T, d, seed=8)
    S = oracle.signature_np(x, depth, True)
    g = np.random.default_rng(9).standard_normal(S.shape)
    gx = oracle.signature_bwd_np(x, S, g, depth, True)
    eps = 1e-6
    for idx in [(0, 0, 0), (0, 2, 1), (1, 4, 2), (1, 3, 0)]:
        xp, xm = x.copy(), x.copy()
        xp[idx] += eps
        xm[idx] -= eps
        fd = ((oracle.signature_np(xp, depth, True) - oracle.signature_np(xm, depth, True)) * g).sum() / (2 * eps)
        assert abs(fd - gx[idx]) <= 1e-6 * max(1.0, abs(fd))


def test_kat8_fp32_vs_fp64_budget():
    """Sets the signature tolerance: the fp32 oracle itself agrees with fp64 to ~1e-6 of max|S|."""
    x = _paths(64, 64, 7, seed=10)
    x = (x * 0.15).astype(np.float32)
    s32 = oracle.signature_np(x, 3, True)
    s64 = oracle.signature_np(x.astype(np.float64), 3, True)
    assert np.abs(s32 - s64).max() / np.abs(s64).max() < 1e-5


def test_empty_batch_and_torch_wrapper():
    x = np.zeros((0, 4, 3))
    assert oracle.signature_np(x, 3).shape == (0, 39)
    xt = torch.randn(3, 6, 2, dtype=torch.float64, requires_grad=True)
    s = oracle.signature_torch(xt, 3, basepoint=True)
    s.sum().backward()
    assert xt.grad.shape == xt.shape
    with pytest.raises(NotImplementedError):
        oracle.signature_torch(xt, 3, stream=True)

"""GPU tests of the tcgen05 split-fp16 path (gram_tc.cu / kf3_tc.cu / planes.cu) through the C ABI: each
building block against float64 numpy, and the fused update forced onto the tensor-core path
against the CPU oracle.  Tolerances as in tests/util.py (fp32-grade: the path must be as accurate
as the fp32 SIMT kernels it replaces)."""
import numpy as np
import pytest
import torch

from tests.util import TOL_H, TOL_K, TOL_PHI, closed_form_with_ties, make_paths, rel_max

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    from sigsvgd_b200 import _lib, ops

DEV = "cuda:0"


def cu(a):
    return torch.as_tensor(np.ascontiguousarray(a), device=DEV)


def test_tc_available():
    assert ops.tc_available()


@pytest.mark.parametrize("M,N,D,scale", [(128, 256, 64, 1.0), (130, 70, 17, 1.0), (300, 517, 399, 1e-3),
                                         (1024, 1024, 399, 50.0), (257, 255, 65, 1.0), (64, 64, 14, 1.0)])
def test_tc_sqdist_vs_float64(M, N, D, scale):
    rng = np.random.default_rng(M + N + D)
    # columns with very different magnitudes, like signature levels 1..3
    colmag = np.exp(rng.uniform(-6, 0, size=D))
    a = (rng.standard_normal((M, D)) * colmag * scale).astype(np.float32)
    b = (rng.standard_normal((N, D)) * colmag * scale).astype(np.float32)
    allrows = cu(np.concatenate([a, b]))
    planes, norms, scales, sc = ops.tc_pack_raw(allrows, None)
    assert torch.equal(sc, allrows)
    pa = ops.Planes(planes.hi[:M], planes.lo[:M])
    pb = ops.Planes(planes.hi[M:], planes.lo[M:])
    got = ops.tc_gram_raw(pa, norms[:M], pb, norms[M:], D, scales, symmetric=False).cpu().numpy()
    a64, b64 = a.astype(np.float64), b.astype(np.float64)
    ref = np.maximum((a64 ** 2).sum(1)[:, None] + (b64 ** 2).sum(1)[None] - 2 * a64 @ b64.T, 0)
    assert got.shape == ref.shape
    assert np.abs(got - ref).max() <= 2e-6 * ref.max()
    assert (got >= 0).all()


@pytest.mark.parametrize("N,D,scale", [(64, 14, 1.0), (300, 39, 1.0), (1000, 399, 1e-2), (2048, 399, 30.0), (4099, 65, 1.0)])
def test_tc_gram_symmetric_and_single_pass_median(N, D, scale):
    rng = np.random.default_rng(N + D)
    colmag = np.exp(rng.uniform(-6, 0, size=D))
    s = (rng.standard_normal((N, D)) * colmag * scale).astype(np.float32)
    s[3] = s[7]  # duplicate point -> an exact zero distance off the diagonal
    mean = cu(s.mean(0).astype(np.float32))
    planes, norms, scales, sc = ops.tc_pack_raw(cu(s), mean)
    state = ops.new_state(torch.device(DEV))
    ws = ops.tc_median_workspace(N, N, DEV)
    ops.tc_median_prepare_raw(sc, sc, ws, state)
    d = ops.tc_gram_raw(planes, norms, planes, norms, D, scales, symmetric=True)
    ops.tc_median_select_raw(ws, d, N, ops.Bandwidth.median(), state)
    got = d.cpu().numpy()
    s64 = s.astype(np.float64)
    ref = np.maximum((s64 ** 2).sum(1)[:, None] + (s64 ** 2).sum(1)[None] - 2 * s64 @ s64.T, 0)
    assert np.abs(got - ref).max() <= 1e-6 * ref.max()
    assert np.array_equal(got, got.T)  # bit-symmetric
    assert (np.diag(got) == 0).all()
    rec = ops.read_state(state)
    want = np.sort(got.reshape(-1))[(N * N - 1) // 2]  # exact lower median of OUR matrix
    assert np.float32(rec["median"]) == want
    win = np.frombuffer(ws[:32].cpu().numpy().tobytes(), dtype=np.uint32)
    assert win[7] == 0, "window missed the median (fallback taken)"
    lo, hi = np.frombuffer(ws[:8].cpu().numpy().tobytes(), dtype=np.float32)
    assert lo <= want <= hi
    assert int(win[2]) == int(((got >= lo) & (got <= hi)).sum())  # candidates collected exactly once each
    # triangle scan: the same exact order statistic from the strict upper triangle only
    state_t = ops.new_state(torch.device(DEV))
    ops.tc_median_prepare_raw(sc, sc, ws, state_t)
    ops.tc_median_select_raw(ws, d, N, ops.Bandwidth.median(), state_t, symmetric=True)
    rec_t = ops.read_state(state_t)
    assert np.float32(rec_t["median"]) == want and rec_t["h"] == rec["h"]
    win_t = np.frombuffer(ws[:32].cpu().numpy().tobytes(), dtype=np.uint32)
    assert win_t[7] == 0, "window missed the median (fallback taken)"
    if N % 4 == 0 and lo > 0:
        iu = np.triu_indices(N, 1)
        assert int(win_t[2]) == int(((got[iu] >= lo) & (got[iu] <= hi)).sum())  # upper entries, once each
    # the non-symmetric kernel (strip of rows) gives the same numbers and the same median machinery
    M = min(N, 200)
    pa = ops.Planes(planes.hi[:M], planes.lo[:M])
    ws2 = ops.tc_median_workspace(M, N, DEV)
    state2 = ops.new_state(torch.device(DEV))
    ops.tc_median_prepare_raw(sc[:M], sc, ws2, state2)
    d2 = ops.tc_gram_raw(pa, norms[:M], planes, norms, D, scales, symmetric=False)
    ops.tc_median_select_raw(ws2, d2, N, ops.Bandwidth.median(), state2)
    g2 = d2.cpu().numpy()
    off = ~np.eye(N, dtype=bool)[:M]
    # (mirrored entries of the symmetric kernel sum the two correction products in the other order)
    assert np.abs(g2[off] - got[:M][off]).max() <= 1e-6 * ref.max()
    assert np.float32(ops.read_state(state2)["median"]) == np.sort(g2.reshape(-1))[(M * N - 1) // 2]


def test_single_pass_median_falls_back_when_window_misses():
    # half of the entries are exactly equal: the candidate list overflows -> full-matrix fallback
    N, D = 512, 8
    s = np.zeros((N, D), np.float32)
    s[: N // 4, 0] = np.random.default_rng(0).standard_normal(N // 4)
    planes, norms, scales, sc = ops.tc_pack_raw(cu(s), None)
    state = ops.new_state(torch.device(DEV))
    ws = ops.tc_median_workspace(N, N, DEV)
    ops.tc_median_prepare_raw(sc, sc, ws, state)
    d = ops.tc_gram_raw(planes, norms, planes, norms, D, scales, symmetric=True)
    ops.tc_median_select_raw(ws, d, N, ops.Bandwidth.median(), state)
    got = d.cpu().numpy()
    assert np.float32(ops.read_state(state)["median"]) == np.sort(got.reshape(-1))[(N * N - 1) // 2]


@pytest.mark.parametrize("M,N,D,P", [(128, 128, 14, 20), (130, 200, 39, 27), (512, 512, 399, 448), (300, 1000, 399, 448),
                                     (128, 4100, 399, 448), (1024, 2048, 399, 448)])
def test_tc_kf_fused_vs_float64(M, N, D, P):
    """K.[Sc | 1 | score] with the RBF map fused into the operand path (kf3_tc.cu): pair kernel for
    M > 128, single-CTA kernel otherwise; exp and identity modes; sum K*d and the tie records."""
    rng = np.random.default_rng(M * 3 + N)
    s = (rng.standard_normal((N, D)) * np.exp(rng.uniform(-5, 0, size=D))).astype(np.float32)
    g = (rng.standard_normal((N, P)) * 30).astype(np.float32)
    dist = (rng.random((M, N)) * 6).astype(np.float32)
    dist[0, :3] = 0
    mean = cu(s.mean(0).astype(np.float32))
    planes, norms, scales, sc = ops.tc_pack_raw(cu(s), mean, cu(g))
    state = ops.new_state(torch.device(DEV))
    _lib.call("sigsvgd_fixed_bandwidth", 1.3, state.data_ptr(), ops._stream())
    # plant a "median" so that the tie detection has something to find: two entries of the strip
    med = np.float32(2.5)
    dist[dist == med] = np.float32(2.25)
    dist[M // 2, N // 3] = med
    dist[M - 1, N - 1] = med
    ops.state_field(state, "median").fill_(float(med))
    d_dev = ops.dist_buffer(M, N, DEV).copy_(cu(dist))
    K64 = np.exp(-0.5 / np.float64(np.float32(1.3)) ** 2 * dist.astype(np.float64))
    assert np.abs(ops.rbf_out_raw(d_dev, state).cpu().numpy() - K64).max() <= 1e-6
    C = D + 1 + P
    out = ops.tc_kf_raw(d_dev, planes, C, scales, state, _lib.KF_EXP, True, row_offset=1000).cpu().numpy()
    assert np.array_equal(d_dev.cpu().numpy(), dist)  # the operand is read-only
    rec = ops.read_state(state)
    assert abs(rec["sum_kd"] - (K64 * dist).sum()) <= 1e-5 * (K64 * dist).sum()
    assert rec["n_ties"] == 2 and sorted(rec["ties"]) == sorted([(1000 + M // 2, N // 3), (1000 + M - 1, N - 1)])
    sc64 = s.astype(np.float64) - mean.cpu().numpy().astype(np.float64)
    F64 = np.concatenate([sc64, np.ones((N, 1)), g.astype(np.float64)], axis=1)
    ref = K64 @ F64
    for lo, hi in ((0, D), (D, D + 1), (D + 1, C)):
        blk = ref[:, lo:hi]
        # TMEM accumulation truncates (round toward zero): the bias grows with the chain length,
        # which the kernel bounds at 192 MMAs per partial sum
        assert np.abs(out[:, lo:hi] - blk).max() <= 1e-5 * np.abs(blk).max(), (lo, hi)
    # identity mode: the operand already holds kernel values (autograd backward with a saved K)
    k_dev = ops.dist_buffer(M, N, DEV).copy_(cu(K64.astype(np.float32)))
    out_i = ops.tc_kf_raw(k_dev, planes, C, scales, None, _lib.KF_IDENTITY).cpu().numpy()
    ref_i = K64.astype(np.float32).astype(np.float64) @ F64
    for lo, hi in ((0, D), (D, D + 1), (D + 1, C)):
        assert np.abs(out_i[:, lo:hi] - ref_i[:, lo:hi]).max() <= 1e-5 * np.abs(ref_i[:, lo:hi]).max(), (lo, hi)


@pytest.mark.parametrize("kind,N,T,d,depth", [("walk", 256, 64, 7, 3), ("panda", 256, 64, 7, 3), ("walk", 1000, 20, 7, 3),
                                              ("field", 100, 10, 2, 3), ("walk", 1024, 64, 7, 3)])
def test_phi_forced_tc_vs_oracle(kind, N, T, d, depth):
    X = make_paths(kind, N, T, d, seed=3)
    G = np.random.default_rng(4).standard_normal((N, T * d)).astype(np.float32)
    res = ops.phi(cu(X), cu(G), depth=depth, want_k=True, want_grad_k=True, flags=_lib.FLAG_FORCE_TC)
    rec = ops.read_state(res.state)
    cf = closed_form_with_ties(X.astype(np.float64), G.astype(np.float64), depth, rec["ties"])
    assert abs(rec["h"] - cf["h"]) <= TOL_H * cf["h"]
    assert np.abs(res.k_xx.cpu().numpy() - cf["K"]).max() <= TOL_K
    assert rel_max(res.grad_k.cpu().numpy(), cf["grad_k"]) <= TOL_PHI
    assert rel_max(res.phi.cpu().numpy(), cf["phi"]) <= TOL_PHI
    # the exact-fp32 SIMT path and the tensor-core path agree far inside the tolerance
    simt = ops.phi(cu(X), cu(G), depth=depth, want_k=True, flags=_lib.FLAG_FORCE_SIMT,
                   bandwidth=ops.Bandwidth.median(median_grad=False))
    tc = ops.phi(cu(X), cu(G), depth=depth, want_k=True, flags=_lib.FLAG_FORCE_TC,
                 bandwidth=ops.Bandwidth.median(median_grad=False))
    assert float((simt.k_xx - tc.k_xx).abs().max()) <= 2e-5
    assert rel_max(tc.phi.cpu().numpy(), simt.phi.cpu().numpy()) <= 2e-5
    # K not requested: same phi
    res2 = ops.phi(cu(X), cu(G), depth=depth, flags=_lib.FLAG_FORCE_TC)
    assert rel_max(res2.phi.cpu().numpy(), res.phi.cpu().numpy()) <= 1e-6

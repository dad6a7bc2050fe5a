"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle and the committed
golden vectors produced by the reference's own Python.  Run with  -m gpu  on a B200.

Tolerances are in tests/util.py (1e-5 signatures / h, 1e-4 K and phi, all vs float64)."""
import numpy as np
import pytest
import torch

import oracle
from tests.util import (TOL_H, TOL_K, TOL_PHI, TOL_SIG, closed_form_with_ties, load_golden, make_paths, rel_max)

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    from sigsvgd_b200 import _lib, ops
    from sigsvgd_b200.inference import FixedBandwidth, make_fused
    from tests import ref_mirror
    from tests.ref_mirror import ScoreEstimator

    SVGD = make_fused(ref_mirror.SVGD)  # the product's fused dispatch over the (restated) reference class
    from sigsvgd_b200.kernels import GaussianKernel, PathSigKernel, bw_median, pw_dist_sq

DEV = "cuda:0"


def cu(a):
    return torch.as_tensor(np.ascontiguousarray(a), device=DEV)


def _kernel_names(fn):
    """Names of the CUDA kernels `fn` launches (torch profiler / CUPTI); None if unavailable."""
    try:
        from torch.profiler import ProfilerActivity, profile

        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            fn()
            torch.cuda.synchronize()
        return [e.key for e in prof.key_averages()]
    except Exception:
        return None


# ---------------------------------------------------------------------------------------------
# K1 / K2 signature forward + VJP
# ---------------------------------------------------------------------------------------------
SIG_SHAPES = [
    # N, T, d, depth
    (1, 1, 7, 3), (3, 2, 7, 3), (17, 64, 7, 3), (130, 30, 2, 3), (16, 5, 1, 3), (33, 9, 3, 3), (20, 12, 4, 3),
    (9, 7, 5, 3), (11, 6, 6, 3), (40, 10, 8, 3), (100, 10, 2, 3), (5, 300, 7, 3),  # T > one staging chunk
    (5, 6, 2, 5), (8, 7, 5, 2), (4, 5, 3, 1), (3, 4, 10, 3), (3, 4, 1, 4), (3, 5, 9, 4),  # generic kernels
    # depth 4, 2 <= d <= 8: sig4 kernels (d^2 threads per particle; chunks of 8 steps; 256/d^2 particles per CTA)
    (7, 8, 3, 4), (6, 9, 7, 4), (2, 1, 7, 4), (11, 17, 7, 4), (13, 20, 2, 4), (9, 10, 4, 4), (6, 9, 5, 4),
    (15, 8, 6, 4), (6, 9, 8, 4), (4, 100, 7, 4),
]


@pytest.mark.parametrize("N,T,d,depth", SIG_SHAPES)
@pytest.mark.parametrize("basepoint", [True, False])
def test_signature_fwd_bwd_vs_oracle(N, T, d, depth, basepoint):
    if not basepoint and T < 2:
        pytest.skip("needs two points")
    x = make_paths("walk", N, T, d, seed=N + T) * 4.0
    x64 = x.astype(np.float64)
    ref = oracle.signature_np(x64, depth, basepoint)
    got = ops.signature_fwd_raw(cu(x), depth, basepoint).cpu().numpy()
    assert got.shape == ref.shape
    assert rel_max(got, ref) <= TOL_SIG
    rng = np.random.default_rng(1)
    g = rng.standard_normal(ref.shape)
    gx_ref = oracle.signature_bwd_np(x64, ref, g, depth, basepoint)
    gx = ops.signature_bwd_raw(cu(x), cu(got), cu(g.astype(np.float32)), depth, basepoint).cpu().numpy()
    assert rel_max(gx, gx_ref) <= 2e-5
    # fused epilogue  out = beta*gx + alpha*add  (row stride of add > T*d)
    add = rng.standard_normal((N, T * d + 3)).astype(np.float32)
    out = ops.signature_bwd_raw(cu(x), cu(got), cu(g.astype(np.float32)), depth, basepoint, add=cu(add), alpha=0.25,
                                beta=-0.5).cpu().numpy()
    want = -0.5 * gx_ref + 0.25 * add[:, :T * d].reshape(N, T, d)
    assert rel_max(out, want) <= 2e-5


def test_signature_depth4_full_size_rows_vs_oracle():
    """BASELINE.json configs[3] shape (4096, 64, 7, depth 4): every row computed, a row sample checked
    against the float64 oracle (forward and VJP)."""
    N, T, d, depth = 4096, 64, 7, 4
    x = make_paths("walk", N, T, d, seed=11)
    rows = np.arange(0, N, 97)
    sig = ops.signature_fwd_raw(cu(x), depth, True)
    assert sig.shape == (N, 2800)
    ref = oracle.signature_np(x[rows].astype(np.float64), depth, True)
    assert rel_max(sig[rows].cpu().numpy(), ref) <= TOL_SIG
    g = np.random.default_rng(2).standard_normal((N, 2800)).astype(np.float32)
    gx = ops.signature_bwd_raw(cu(x), sig, cu(g), depth, True).cpu().numpy()
    gx_ref = oracle.signature_bwd_np(x[rows].astype(np.float64), ref, g[rows].astype(np.float64), depth, True)
    assert rel_max(gx[rows], gx_ref) <= 2e-5


PREFIX_SHAPES = [(17, 64, 7), (130, 30, 2), (20, 12, 4), (40, 10, 8), (100, 10, 2), (33, 9, 4), (9, 11, 4), (16, 8, 1),
                 (6, 13, 8), (5, 300, 7), (21, 16, 3), (4097, 64, 7)]


@pytest.mark.parametrize("N,T,d", PREFIX_SHAPES)
@pytest.mark.parametrize("basepoint", [True, False])
def test_signature_time_parallel_pair_vs_oracle(N, T, d, basepoint):
    """Forward with prefix checkpoints + the backward that sweeps 4 time segments at once
    (sigsvgd_signature_fwd_prefix / _bwd_prefix) against the float64 oracle."""
    depth = 3
    L12 = d + d * d
    assert ops.signature_prefix_floats(N, T, d, depth) == N * 3 * L12
    x = make_paths("walk", N, T, d, seed=N + T) * 4.0
    x64 = x.astype(np.float64)
    ref = oracle.signature_np(x64, depth, basepoint)
    sig, pref = ops.signature_fwd_prefix_raw(cu(x), depth, basepoint)
    assert rel_max(sig.cpu().numpy(), ref) <= TOL_SIG
    # the checkpoints are levels 1-2 of the signatures of the truncated paths
    seg = -(-T // 4)
    pr = pref.cpu().numpy().reshape(N, 3, -1)
    for q in range(3):
        if (q + 1) * seg < T:
            want = oracle.signature_np(x64[:, :(q + 1) * seg], depth, basepoint)
            assert rel_max(pr[:, q], want[:, :L12]) <= TOL_SIG
    rng = np.random.default_rng(2)
    g = rng.standard_normal(ref.shape)
    gx_ref = oracle.signature_bwd_np(x64, ref, g, depth, basepoint)
    gx = ops.signature_bwd_prefix_raw(cu(x), sig, pref, cu(g.astype(np.float32)), depth, basepoint).cpu().numpy()
    assert rel_max(gx, gx_ref) <= 2e-5
    for pad in (4, 3):  # 16-byte rows of `add`: time-parallel kernel; otherwise the sequential one
        add = rng.standard_normal((N, T * d + pad)).astype(np.float32)
        out = ops.signature_bwd_prefix_raw(cu(x), sig, pref, cu(g.astype(np.float32)), depth, basepoint, add=cu(add),
                                           alpha=0.25, beta=-0.5).cpu().numpy()
        want = -0.5 * gx_ref + 0.25 * add[:, :T * d].reshape(N, T, d)
        assert rel_max(out, want) <= 2e-5


def test_signature_time_parallel_pair_availability():
    assert ops.signature_prefix_floats(8, 7, 4, 3) == 0      # fewer than 2 steps per segment
    assert ops.signature_prefix_floats(8, 9, 3, 3) == 0      # rows not a multiple of 16 bytes
    assert ops.signature_prefix_floats(8, 16, 4, 4) == 0     # depth 3 only
    with pytest.raises(ValueError):
        ops.signature_fwd_prefix_raw(torch.zeros(8, 7, 4, device=DEV), 3)


def test_signature_autograd_node_and_empty():
    x = torch.randn(6, 9, 3, device=DEV, requires_grad=True)
    s = ops.signature(x, 3, basepoint=True)
    (s * s).sum().backward()
    xr = x.detach().cpu().double().numpy()
    sr = oracle.signature_np(xr, 3, True)
    gr = oracle.signature_bwd_np(xr, sr, 2 * sr, 3, True)
    assert rel_max(x.grad.cpu().numpy(), gr) <= 2e-5
    empty = ops.signature_fwd_raw(torch.empty(0, 5, 3, device=DEV), 3, True)
    assert empty.shape == (0, 39)
    # tensor basepoint (signatory semantics): prepend it
    bp = torch.randn(6, 3, device=DEV)
    s2 = ops.signature(x.detach(), 3, basepoint=bp)
    s3 = oracle.signature_np(np.concatenate([bp.cpu().numpy()[:, None], x.detach().cpu().numpy()], 1).astype(np.float64),
                             3, False)
    assert rel_max(s2.cpu().numpy(), s3) <= TOL_SIG


def test_signature_errors():
    with pytest.raises(RuntimeError):
        ops.signature(torch.randn(2, 3, 2), 3, basepoint=True)  # CPU tensor: no fallback
    with pytest.raises(ValueError):
        ops.signature_fwd_raw(torch.randn(2, 1, 2, device=DEV), 3, False)  # one point, no basepoint
    with pytest.raises(ValueError):
        ops.signature_fwd_raw(torch.randn(2, 3, 2, device=DEV), 9, True)  # depth > 8


# ---------------------------------------------------------------------------------------------
# K3 / K4 / K5 pieces
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("M,N,D", [(1, 1, 1), (5, 7, 3), (64, 64, 14), (100, 100, 399), (130, 70, 17), (257, 255, 65)])
def test_sqdist_vs_oracle(M, N, D):
    rng = np.random.default_rng(M * N)
    a, b = rng.standard_normal((M, D)).astype(np.float32), rng.standard_normal((N, D)).astype(np.float32)
    got = pw_dist_sq(cu(a), cu(b)).cpu().numpy()
    ref = oracle.pw_dist_sq(a.astype(np.float64), b.astype(np.float64))
    assert np.abs(got - ref).max() <= 1e-5 * max(ref.max(), 1.0)
    assert (got >= 0).all()
    if M == N:
        sym = pw_dist_sq(cu(a), cu(a)).cpu().numpy()
        assert np.array_equal(sym, sym.T)  # bit-symmetric by construction


@pytest.mark.parametrize("n", [1, 2, 3, 7, 64, 1000, 4099, 65536 + 3])
def test_median_select_is_exact_lower_median(n):
    rng = np.random.default_rng(n)
    vals = (rng.standard_normal(n) ** 2).astype(np.float32)
    if n > 10:
        vals[rng.integers(0, n, n // 4)] = 0.0  # clamped zeros / duplicates
        vals[1::7] = vals[0]
    want = np.sort(vals)[(n - 1) // 2]
    state = ops.bandwidth_raw(cu(vals).reshape(1, -1), ops.Bandwidth.median(), n_log=10)
    rec = ops.read_state(state)
    assert np.float32(rec["median"]) == want  # bit exact
    h = np.float32(np.sqrt(np.float32(want) / np.float32(np.log(np.float32(11.0)))))
    h = max(h, np.float32(1e-8))
    assert abs(rec["h"] - h) <= 2e-7 * max(h, 1e-8)
    assert rec["clamped"] == int(want == 0)


def test_median_matches_torch_median_convention():
    torch.manual_seed(0)
    sq = torch.rand(64, 64, device=DEV)
    assert float(bw_median(sq)) == pytest.approx(float(torch.sqrt(torch.median(sq) / np.log(65.0))), rel=1e-6)


def test_gaussian_kernel_vs_reference_golden():
    g = load_golden("gaussian_20x5")
    A, B = cu(g["A"]), cu(g["B"])
    K, dK = GaussianKernel()(A, B)
    assert np.abs(K.cpu().numpy() - g["K_f64"]).max() <= TOL_K
    assert rel_max(dK.cpu().numpy(), g["dK_f64"]) <= TOL_PHI
    Kh, dKh = GaussianKernel()(A, B, h=0.7)
    assert np.abs(Kh.cpu().numpy() - g["Kh_f64"]).max() <= TOL_K
    assert rel_max(dKh.cpu().numpy(), g["dKh_f64"]) <= TOL_PHI
    K2 = GaussianKernel(bandwidth_fn=lambda _: 0.7)(A, B, compute_grad=False)
    assert np.abs(K2.cpu().numpy() - g["Kh_f64"]).max() <= TOL_K
    with pytest.raises(ValueError):
        GaussianKernel(bandwidth_fn=3.0)
    with pytest.raises(AssertionError):
        GaussianKernel()(A, B[:5])


# ---------------------------------------------------------------------------------------------
# full update against the reference's golden vectors
# ---------------------------------------------------------------------------------------------
GOLDEN_CASES = ["field_100x10x2_m3", "walk_64x16x7_m3", "maze_35x30x2_m3", "panda_48x12x7_m3", "walk_40x9x3_m4",
                "walk_33x8x5_m2"]


def _check_update(X, G, depth, res_phi, K, grad_k, rec, fixed_h=None):
    X64, G64 = X.astype(np.float64), G.astype(np.float64)
    ties = [] if fixed_h is not None else rec["ties"]
    cf = closed_form_with_ties(X64, G64, depth, ties, h=fixed_h)
    if fixed_h is None:
        assert abs(rec["h"] - cf["h"]) <= TOL_H * cf["h"]
        assert abs(rec["median"] - cf["median"]) <= TOL_H * cf["median"]
        assert rec["n_ties"] >= 1 and rec["n_ties"] == rec["n_ties_all"]
        for a, b in ties:  # each recorded pair really attains the median
            assert abs(cf["sq"][a, b] - cf["median"]) <= 1e-5 * cf["median"]
    if K is not None:
        assert np.abs(K - cf["K"]).max() <= TOL_K
    if grad_k is not None:
        assert rel_max(grad_k, cf["grad_k"]) <= TOL_PHI
    assert rel_max(res_phi, cf["phi"]) <= TOL_PHI
    return cf


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_phi_vs_reference_golden(name):
    g = load_golden(name)
    X, G, depth = g["X"], g["G"], int(g["depth"])
    res = ops.phi(cu(X), cu(G), depth=depth, want_k=True, want_grad_k=True, want_sig=True)
    rec = ops.read_state(res.state)
    # signature / distances / h / K: straight against the reference's float64 run
    assert rel_max(res.sig.cpu().numpy(), g["sig_f64"]) <= TOL_SIG
    assert abs(rec["h"] - float(g["h_f64"])) <= TOL_H * float(g["h_f64"])
    assert np.abs(res.k_xx.cpu().numpy() - g["k_xx_f64"]).max() <= TOL_K
    # phi / grad_k: rows outside both tie sets straight against the reference; all rows against the
    # closed form evaluated with OUR median pair (SURVEY.md 8(a-4) protocol)
    phi = res.phi.cpu().numpy()
    _check_update(X, G, depth, phi, res.k_xx.cpu().numpy(), res.grad_k.cpu().numpy(), rec)
    ref_vel = g["velocity_f64"]
    sq64 = g["sq_f64"]
    med64 = np.sort(sq64.reshape(-1))[(sq64.size - 1) // 2]
    ref_rows = set(np.argwhere(sq64 == med64)[:, 0].tolist())
    skip = sorted(ref_rows | {a for a, _ in rec["ties"]})
    keep = np.setdiff1d(np.arange(X.shape[0]), skip)
    assert len(keep) >= X.shape[0] - 4
    assert np.abs(-phi[keep] - ref_vel[keep]).max() <= TOL_PHI * np.abs(ref_vel[keep]).max()
    # fixed bandwidth (bandwidth_fn=lambda _: 1.5) has no median term: every row compares directly
    resf = ops.phi(cu(X), cu(G), depth=depth, bandwidth=ops.Bandwidth.fixed(1.5), want_k=True, want_grad_k=True)
    assert np.abs(resf.k_xx.cpu().numpy() - g["k_fix_f64"]).max() <= TOL_K
    if np.abs(g["gradk_fix_f64"]).max() < 1e-12:  # h=1.5 vs distances ~1e3 (panda): K == I, no repulsion
        assert np.abs(resf.grad_k.cpu().numpy()).max() < 1e-12
    else:
        assert rel_max(resf.grad_k.cpu().numpy(), g["gradk_fix_f64"]) <= TOL_PHI
    assert rel_max(-resf.phi.cpu().numpy(), g["vel_fix_f64"]) <= TOL_PHI


@pytest.mark.parametrize("kind,N,T,d,depth", [("walk", 256, 64, 7, 3), ("panda", 256, 64, 7, 3), ("field", 100, 10, 2, 3),
                                              ("walk", 35, 30, 2, 3), ("walk", 513, 20, 7, 3), ("walk", 64, 12, 3, 4),
                                              ("walk", 384, 64, 7, 4)])
@pytest.mark.parametrize("center", [True, False])
def test_phi_vs_oracle_seeded(kind, N, T, d, depth, center):
    X = make_paths(kind, N, T, d, seed=3)
    G = np.random.default_rng(4).standard_normal((N, T * d)).astype(np.float32)
    res = ops.phi(cu(X), cu(G), depth=depth, want_k=True, want_grad_k=True, center=center)
    rec = ops.read_state(res.state)
    if kind == "panda" and not center:
        pytest.skip("uncentred fp32 Gram of clustered joint-space paths is the reference's own weak spot")
    _check_update(X, G, depth, res.phi.cpu().numpy(), res.k_xx.cpu().numpy(), res.grad_k.cpu().numpy(), rec)
    # K not requested: same phi
    res2 = ops.phi(cu(X), cu(G), depth=depth, center=center)
    assert torch.equal(res2.phi, res.phi) or rel_max(res2.phi.cpu().numpy(), res.phi.cpu().numpy()) < 1e-6


def test_phi_median_grad_off_and_linearity():
    X = make_paths("walk", 128, 16, 7, seed=5)
    rng = np.random.default_rng(6)
    G1, G2 = rng.standard_normal((2, 128, 16 * 7)).astype(np.float32)
    bw = ops.Bandwidth.median(median_grad=False)
    r1 = ops.phi(cu(X), cu(G1), bandwidth=bw)
    cf = closed_form_with_ties(X.astype(np.float64), G1.astype(np.float64), 3, [])
    assert rel_max(r1.phi.cpu().numpy(), cf["phi"]) <= TOL_PHI
    # phi is affine in the score: phi(a G1 + b G2) - phi(0) = a (phi(G1)-phi(0)) + b (phi(G2)-phi(0))
    z = ops.phi(cu(X), cu(0 * G1), bandwidth=bw).phi
    r2 = ops.phi(cu(X), cu(G2), bandwidth=bw).phi
    r12 = ops.phi(cu(X), cu(0.5 * G1 - 2 * G2), bandwidth=bw).phi
    want = 0.5 * (r1.phi - z) - 2 * (r2 - z) + z
    assert rel_max(r12.cpu().numpy(), want.cpu().numpy()) <= 1e-5


def test_phi_batched_matches_loop():
    B, N, T, d = 5, 96, 16, 7  # N not a multiple of 128: exact-fp32 SIMT kernels, batched in the grid
    X = np.stack([make_paths("walk", N, T, d, seed=10 + b) for b in range(B)])
    G = np.random.default_rng(7).standard_normal((B, N, T * d)).astype(np.float32)
    res = ops.phi(cu(X), cu(G), want_k=True)
    for b in range(B):
        one = ops.phi(cu(X[b]), cu(G[b]), want_k=True)
        assert torch.equal(one.phi, res.phi[b])
        assert torch.equal(one.k_xx, res.k_xx[b])
        assert ops.read_state(one.state)["median"] == ops.read_state(res.state, b)["median"]


@pytest.mark.parametrize("B,N,T,d", [(6, 128, 64, 7), (3, 256, 20, 7), (9, 128, 10, 2)])
def test_phi_batched_scenes_on_tensor_cores(B, N, T, d):
    """Independent scenes whose particle count is a multiple of the 128-row tile run as ONE grouped
    tcgen05 Gram / K.F launch (BASELINE.json configs[4]); every scene keeps its own median."""
    X = np.stack([make_paths("walk", N, T, d, seed=40 + b) * (1.0 + 0.3 * b) for b in range(B)])
    G = np.random.default_rng(8).standard_normal((B, N, T * d)).astype(np.float32)
    res = ops.phi(cu(X), cu(G), want_k=True, want_grad_k=True)
    names = _kernel_names(lambda: ops.phi(cu(X), cu(G)))
    if names is not None and len(names):
        assert any("gram_kernel" in n for n in names) and any("kf_fused_kernel" in n for n in names), names
    nok = ops.phi(cu(X), cu(G))  # k_xx not requested: same phi
    assert rel_max(nok.phi.cpu().numpy(), res.phi.cpu().numpy()) <= 1e-6
    simt = ops.phi(cu(X), cu(G), want_k=True, flags=_lib.FLAG_FORCE_SIMT)
    for b in range(B):
        rec = ops.read_state(res.state, b)
        cf = closed_form_with_ties(X[b].astype(np.float64), G[b].astype(np.float64), 3, rec["ties"])
        assert abs(rec["h"] - cf["h"]) <= TOL_H * cf["h"]
        assert np.abs(res.k_xx[b].cpu().numpy() - cf["K"]).max() <= TOL_K
        assert rel_max(res.grad_k[b].cpu().numpy(), cf["grad_k"]) <= TOL_PHI
        assert rel_max(res.phi[b].cpu().numpy(), cf["phi"]) <= TOL_PHI
        assert float((simt.k_xx[b] - res.k_xx[b]).abs().max()) <= 2e-5
        assert ops.read_state(simt.state, b)["h"] == pytest.approx(rec["h"], rel=1e-6)


# ---------------------------------------------------------------------------------------------
# drop-in API: kernels / SVGD / ScoreEstimator mirrors
# ---------------------------------------------------------------------------------------------
def test_pathsig_kernel_autograd_route_matches_fused():
    g = load_golden("walk_64x16x7_m3")
    X, G = cu(g["X"]), cu(g["G"])
    kernel = PathSigKernel()
    assert kernel.analytic_grad is False and isinstance(kernel.static_kernel, GaussianKernel)
    # the reference's SVGD._compute_kernel statements (svgd.py:41-43)
    Xr = X.detach().requires_grad_(True)
    k_xx = kernel(Xr, Xr.detach(), compute_grad=False)
    grad_k = torch.autograd.grad(k_xx.sum(), Xr)[0].flatten(1)
    fused = ops.phi(X, G, want_k=True, want_grad_k=True)
    assert torch.equal(k_xx.detach(), fused.k_xx)
    assert rel_max(grad_k.cpu().numpy(), fused.grad_k.cpu().numpy()) <= 1e-6
    # both arguments carry grad (script_planning_robot.py:107): twice the one-sided gradient
    Xb = X.detach().requires_grad_(True)
    k2 = kernel(Xb, Xb, compute_grad=False)
    g2 = torch.autograd.grad(k2.sum(), Xb)[0].flatten(1)
    assert rel_max(g2.cpu().numpy(), 2 * fused.grad_k.cpu().numpy()) <= 1e-6
    # compute_grad=True: analytic signature-space gradient [N, D] (reference golden dK_sig)
    K3, dK = kernel(X, X.clone(), compute_grad=True)
    assert np.abs(K3.cpu().numpy() - g["k_xx_f64"]).max() <= TOL_K
    assert rel_max(dK.cpu().numpy(), g["dK_sig_f64"]) <= TOL_PHI
    # ignored arguments stay ignored: h and ref_vector change nothing
    k4 = kernel(X, X, ref_vector=X, h=123.0, compute_grad=False)
    assert torch.equal(k4, fused.k_xx)


def test_pathsig_kernel_two_sample_and_generic_cotangent():
    X = cu(make_paths("walk", 48, 12, 3, seed=8)).requires_grad_(True)
    Y = cu(make_paths("walk", 48, 12, 3, seed=9)).requires_grad_(True)
    W = torch.rand(48, 48, device=DEV)
    K = PathSigKernel()(X, Y, compute_grad=False)
    gX, gY = torch.autograd.grad((K * W).sum(), (X, Y))
    # float64 torch autograd through the oracle signature = what the reference would compute
    Xc, Yc = X.detach().cpu().double().requires_grad_(True), Y.detach().cpu().double().requires_grad_(True)
    sx, sy = oracle.signature_torch(Xc, 3, basepoint=True), oracle.signature_torch(Yc, 3, basepoint=True)
    sq = ((sx[:, None] - sy[None]) ** 2).sum(-1)
    h = torch.sqrt(torch.median(sq) / torch.tensor(49.0).log())
    Kc = torch.exp(-0.5 / h ** 2 * sq)
    gXc, gYc = torch.autograd.grad((Kc * W.cpu().double()).sum(), (Xc, Yc))
    assert np.abs(K.detach().cpu().numpy() - Kc.detach().numpy()).max() <= TOL_K
    assert rel_max(gX.cpu().numpy(), gXc.numpy()) <= TOL_PHI
    assert rel_max(gY.cpu().numpy(), gYc.numpy()) <= TOL_PHI


@pytest.mark.parametrize("N", [1024, 4096])
def test_reference_autograd_statements_run_on_tensor_cores(N):
    """The literal drop-in -- the reference's own statements at svgd.py:41-43 / score.py:54-55 --
    reaches the tcgen05 Gram and K.F kernels and agrees with the fused update."""
    X = cu(make_paths("walk", N, 64, 7, seed=21))
    G = cu(np.random.default_rng(22).standard_normal((N, 448)).astype(np.float32))
    kernel = PathSigKernel()
    out = {}

    def statements():
        Xr = X.detach().requires_grad_(True)
        k_xx = kernel(Xr, Xr.detach(), compute_grad=False)
        out["k_xx"], out["grad_k"] = k_xx.detach(), torch.autograd.grad(k_xx.sum(), Xr)[0].flatten(1)

    statements()
    names = _kernel_names(statements)
    if names is not None and len(names):
        assert any("gram_kernel" in n for n in names), names
        assert any("kf_fused_kernel" in n for n in names), names
        assert not any("sqdist_nt_kernel" in n or "gemm_nn_kernel" in n for n in names), names
    fused = ops.phi(X, G, want_k=True, want_grad_k=True)
    rf, ra = ops.read_state(fused.state), None
    assert float((out["k_xx"] - fused.k_xx).abs().max()) <= 2e-6
    # same median / h; the tie pair is the same because both routes select on the same Gram matrix
    assert rel_max(out["grad_k"].cpu().numpy(), fused.grad_k.cpu().numpy()) <= 1e-5
    # velocity = (k_xx @ score - grad_k) / N as the reference forms it (svgd.py:82), vs phi
    vel = (out["k_xx"].double() @ G.double() - out["grad_k"].double()) / N
    assert rel_max(vel.cpu().numpy(), fused.phi.reshape(N, -1).cpu().numpy()) <= 1e-5
    # both arguments carry grad (script_planning_robot.py:107): twice the one-sided gradient
    Xb = X.detach().requires_grad_(True)
    g2 = torch.autograd.grad(kernel(Xb, Xb, compute_grad=False).sum(), Xb)[0].flatten(1)
    assert rel_max(g2.cpu().numpy(), 2 * fused.grad_k.cpu().numpy()) <= 1e-5
    if N <= 1024:  # float64 closed form with the reported tie pair(s)
        cf = closed_form_with_ties(X.cpu().numpy().astype(np.float64), G.cpu().numpy().astype(np.float64), 3, rf["ties"])
        assert np.abs(out["k_xx"].cpu().numpy() - cf["K"]).max() <= TOL_K
        assert rel_max(out["grad_k"].cpu().numpy(), cf["grad_k"]) <= TOL_PHI


def test_two_sample_and_generic_cotangent_on_tensor_cores():
    N = 300  # >= ops.TC_MIN_ROWS: tcgen05 route, non-symmetric Gram, weighted kernel in the backward
    X = cu(make_paths("walk", N, 12, 3, seed=8)).requires_grad_(True)
    Y = cu(make_paths("walk", N, 12, 3, seed=9)).requires_grad_(True)
    W = torch.rand(N, N, device=DEV)
    K = PathSigKernel()(X, Y, compute_grad=False)
    gX, gY = torch.autograd.grad((K * W).sum(), (X, Y), retain_graph=True)
    Xc, Yc = X.detach().cpu().double().requires_grad_(True), Y.detach().cpu().double().requires_grad_(True)
    sx, sy = oracle.signature_torch(Xc, 3, basepoint=True), oracle.signature_torch(Yc, 3, basepoint=True)
    sq = ((sx[:, None] - sy[None]) ** 2).sum(-1)
    h = torch.sqrt(torch.median(sq) / torch.tensor(N + 1.0).log())
    Kc = torch.exp(-0.5 / h ** 2 * sq)
    gXc, gYc = torch.autograd.grad((Kc * W.cpu().double()).sum(), (Xc, Yc), retain_graph=True)
    assert np.abs(K.detach().cpu().numpy() - Kc.detach().numpy()).max() <= TOL_K
    assert rel_max(gX.cpu().numpy(), gXc.numpy()) <= TOL_PHI
    assert rel_max(gY.cpu().numpy(), gYc.numpy()) <= TOL_PHI
    # uniform cotangent, gradient wrt both (distinct) arguments
    gX1, gY1 = torch.autograd.grad(K.sum(), (X, Y))
    gX1c, gY1c = torch.autograd.grad(Kc.sum(), (Xc, Yc))
    assert rel_max(gX1.cpu().numpy(), gX1c.numpy()) <= TOL_PHI
    assert rel_max(gY1.cpu().numpy(), gY1c.numpy()) <= TOL_PHI


def test_gaussian_kernel_large_on_tensor_cores():
    rng = np.random.default_rng(5)
    A = (rng.standard_normal((512, 40)) * np.exp(rng.uniform(-3, 0, 40))).astype(np.float32)
    B = (rng.standard_normal((512, 40)) * np.exp(rng.uniform(-3, 0, 40))).astype(np.float32)
    K, dK = GaussianKernel()(cu(A), cu(B))
    A64, B64 = A.astype(np.float64), B.astype(np.float64)
    Kr, sq, h, med = oracle.gaussian_gram(A64, B64, None, 1.0)
    dKr = -(A64 * Kr.sum(1, keepdims=True) - Kr @ B64) / h ** 2  # _kernels.py:112-113, constant h
    assert np.abs(K.cpu().numpy() - Kr).max() <= TOL_K
    assert rel_max(dK.cpu().numpy(), dKr) <= TOL_PHI
    a_dev = cu(A)
    Ks, dKs = GaussianKernel()(a_dev, a_dev)  # same data: symmetric Gram, exact-zero diagonal distances
    Krs, _, hs, _ = oracle.gaussian_gram(A64, A64, None, 1.0)
    assert np.abs(Ks.cpu().numpy() - Krs).max() <= TOL_K and float(Ks.diagonal().min()) == 1.0


def test_svgd_mirror_velocity_and_optimize_vs_reference_golden():
    g = load_golden("optimize_24x8x2")
    centers = cu(g["centers"])

    def cost_fn(x, w_len=1.0):
        bumps = torch.exp(-((x[:, :, None, :] - centers) ** 2).sum(-1) / 0.5).sum((1, 2))
        length = ((x[:, 1:] - x[:, :-1]) ** 2).sum((1, 2))
        return bumps + w_len * length, {}

    kernel = PathSigKernel()
    est = ScoreEstimator(kernel, cost_fn, {"w_len": 0.5})
    assert est.score.__func__ is ScoreEstimator._svgd_ag_score
    sampler = SVGD(kernel=kernel, optimizer_class=None, lr=0.05)
    P = cu(g["X0"]).clone()
    data, opt_state = sampler.optimize(P, est.score, n_steps=5)
    assert opt_state is None and data["trace"].shape == g["trace_f64"].shape
    # WHICH pair attains the median is implementation dependent and the pair's
    # rows get a large push, so compare step 1 on all other rows
    ours = {a for a, _ in ops.read_state(ops.phi(cu(g["X0"]), cu(np.zeros((24, 16), np.float32))).state)["ties"]}
    theirs = set(oracle.phi_closed_form(g["X0"].astype(np.float64), np.zeros((24, 16)))["ties"][:, 0].tolist())
    keep = np.setdiff1d(np.arange(24), sorted(ours | theirs))
    assert rel_max(data["trace"][1].numpy()[keep], g["trace_f64"][1][keep]) <= 1e-4
    assert set(data[0].keys()) >= {"k_xx", "grad_k", "loss", "grad"}
    assert data[0]["k_xx"].device.type == "cpu"  # reference logging semantics (svgd.py:85-90)
    assert torch.equal(P.cpu(), data["trace"][-1])  # in-place write-back (svgd.py:157)
    # fused route (no score estimator): velocity = -phi
    X, G = cu(g["X0"]), torch.randn(24, 16, device=DEV)
    s2 = SVGD(kernel=PathSigKernel(static_kernel=GaussianKernel(bandwidth_fn=FixedBandwidth(1.5))), optimizer_class=None,
              lr=0.1)
    s2.log_level = "none"
    v, it = s2._velocity(X, G)
    cf = closed_form_with_ties(g["X0"].astype(np.float64), G.cpu().double().numpy(), 3, [], h=1.5)
    assert rel_max(-v.cpu().numpy(), cf["phi"]) <= TOL_PHI
    assert it["k_xx"] is None
    with pytest.raises(ValueError):
        SVGD(kernel=kernel)._velocity(X, None)


def test_adam_optimizer_state_roundtrip():
    X = cu(make_paths("walk", 32, 8, 2, seed=11))
    sampler = SVGD(kernel=PathSigKernel(), log_p=lambda x: -(x ** 2).sum((1, 2)), optimizer_class=torch.optim.Adam, lr=0.01)
    sampler.log_level = "device"
    d1, st = sampler.optimize(X.clone(), n_steps=2)
    d2, st2 = sampler.optimize(X.clone(), opt_state=st, n_steps=1)
    assert st2["state"][0]["step"] == 3 and d1["trace"].shape[0] == 3


# ---------------------------------------------------------------------------------------------
# BASELINE.json full size: size-independent properties
# ---------------------------------------------------------------------------------------------
def test_full_size_properties_n4096():
    N, T, d = 4096, 64, 7
    torch.manual_seed(0)
    X = torch.cumsum(0.05 * torch.randn(N, T, d, device=DEV), dim=1)
    G = torch.randn(N, T * d, device=DEV)
    res = ops.phi(X, G, want_k=True, want_sig=True, want_grad_k=True)
    K, S = res.k_xx, res.sig
    rec = ops.read_state(res.state)
    assert torch.equal(K, K.t())  # symmetric bit for bit
    assert float((K.diagonal() - 1).abs().max()) <= 1e-5
    assert float(K.min()) >= 0 and float(K.max()) <= 1.0
    # the bandwidth really is the lower median of all N^2 distances (checked with torch.sort on device)
    Sc = (S - S.mean(0)).double()
    sq = torch.cdist(Sc, Sc).pow(2)
    med = torch.sort(sq.reshape(-1)).values[(N * N - 1) // 2].item()
    assert abs(rec["median"] - med) <= 2e-5 * med
    h = (med / np.log(N + 1.0)) ** 0.5
    assert abs(rec["h"] - h) <= TOL_H * h
    assert float((K.double() - torch.exp(-0.5 * sq / h ** 2)).abs().max()) <= TOL_K
    # KAT-1: level 1 of the signature with basepoint is the last point; KAT-3 shuffle identity
    assert float((S[:, :d] - X[:, -1]).abs().max()) <= 1e-5
    S1, S2 = S[:, :d], S[:, d:d + d * d].reshape(N, d, d)
    assert float((S1[:, :, None] * S1[:, None, :] - (S2 + S2.transpose(1, 2))).abs().max()) <= 1e-5
    # phi = (K G - grad_k)/N with the returned pieces; repulsive term sums to ~0 for constant h part
    phi = (K @ G - res.grad_k) / N
    assert rel_max(res.phi.flatten(1).cpu().numpy(), phi.cpu().numpy()) <= 1e-4
    # float64 recomputation of the whole update on the device (torch as the checker)
    g_sig = -(Sc * K.double().sum(1, keepdim=True) - K.double() @ Sc) / h ** 2
    c = (K.double() * sq).sum() / (h ** 2 * med)
    for a, b in rec["ties"]:
        g_sig[a] += c / len(rec["ties"]) * (Sc[a] - Sc[b])
    gk = ops.signature_bwd_raw(X, S, g_sig.float().contiguous(), 3, True).flatten(1)
    assert rel_max(res.grad_k.cpu().numpy(), gk.cpu().numpy()) <= 1e-4

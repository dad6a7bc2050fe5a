"""The CPU oracle against the golden vectors produced by the UNMODIFIED reference Python
(oracle/make_golden.py -> tests/golden/*.npz), and -- when /root/reference is present, i.e. in the
build container only -- against the reference imported live.  CPU only."""
import numpy as np
import pytest
import torch

import oracle
from oracle import ref_shim
from tests.util import closed_form_with_ties, load_golden, rel_max

CASES = ["field_100x10x2_m3", "walk_64x16x7_m3", "maze_35x30x2_m3", "panda_48x12x7_m3", "walk_40x9x3_m4",
         "walk_33x8x5_m2"]


@pytest.mark.parametrize("name", CASES)
def test_closed_form_fp64_matches_reference_fp64(name):
    g = load_golden(name)
    depth = int(g["depth"])
    X, G = g["X"].astype(np.float64), g["G"].astype(np.float64)
    cf = oracle.phi_closed_form(X, G, depth)
    assert rel_max(cf["sig"], g["sig_f64"]) < 1e-12
    assert rel_max(cf["sq"], g["sq_f64"]) < 1e-10
    # the reference takes ln(N+1) in float32 even for float64 inputs (math.py:32): restated as such
    assert abs(cf["h"] - float(g["h_f64"])) / float(g["h_f64"]) < 1e-12
    assert np.abs(cf["K"] - g["k_xx_f64"]).max() < 1e-12
    # gradient through torch.median: WHICH entry attains the median is decided by the last bits of
    # the distance matrix ((a,b) vs (b,a): addmm is not bit-symmetric), so the closed form is
    # evaluated with the reference's own tie set (parity protocol of SURVEY.md 8(a-4))
    ties = np.argwhere(g["sq_f64"] == oracle.lower_median(g["sq_f64"]))
    assert 1 <= len(ties) <= 2
    cft = closed_form_with_ties(X, G, depth, ties)
    assert rel_max(cft["grad_k"], g["grad_k_f64"]) < 1e-9
    assert rel_max(-cft["phi"], g["velocity_f64"]) < 1e-9
    # and the oracle's own tie set is the same unordered pair
    assert {frozenset(map(int, t)) for t in cf["ties"]} == {frozenset(map(int, t)) for t in ties}
    # fixed bandwidth 1.5 (bandwidth_fn=lambda _: 1.5): no median term
    cfx = oracle.phi_closed_form(X, G, depth, h=1.5)
    assert np.abs(cfx["K"] - g["k_fix_f64"]).max() < 1e-12
    # (panda paths at h=1.5: off-diagonal K underflows, grad_k ~ 1e-13 is pure cancellation noise
    #  in both implementations -> absolute floor relative to the feature scale)
    floor = 1e-15 * np.abs(cf["sig"]).max()
    assert np.abs(cfx["grad_k"] - g["gradk_fix_f64"]).max() < 1e-9 * np.abs(g["gradk_fix_f64"]).max() + floor
    assert rel_max(-cfx["phi"], g["vel_fix_f64"]) < 1e-9


@pytest.mark.parametrize("name", CASES)
def test_analytic_sig_space_gradient_matches_reference(name):
    """PathSigKernel(compute_grad=True) returns GaussianKernel's d_K.sum(1) in signature space
    (_kernels.py:111-113): constant-h closed form, no median term."""
    g = load_golden(name)
    S = g["sig_f64"]
    K, sq, h, med = oracle.gaussian_gram(S, S)
    gs, _ = oracle.sig_space_grad(S, S, K, sq, h, med, median_grad=False)
    assert rel_max(gs, g["dK_sig_f64"]) < 1e-9


@pytest.mark.parametrize("name", CASES[:3])
def test_torch_literal_port_reproduces_reference_fp32(name):
    """The timed CPU baseline (bench.py cpu_baseline kind="port") is the reference's own torch
    statements; on the same inputs it must reproduce the reference's fp32 outputs (same library,
    same order of operations -> near bit-level; the median pair may differ only by fp32 ties)."""
    g = load_golden(name)
    X, G = torch.from_numpy(g["X"]), torch.from_numpy(g["G"])
    out = oracle.phi_torch_literal(X, G, depth=int(g["depth"]))
    assert abs(out["h"] - float(g["h_f32"])) <= 1e-6 * float(g["h_f32"])
    assert np.abs(out["k_xx"].numpy() - g["k_xx_f32"]).max() < 1e-5
    assert rel_max(out["velocity"].numpy(), g["velocity_f32"]) < 1e-4


def test_median_convention_is_lower_median():
    a = np.array([[4.0, 1.0], [3.0, 2.0]])
    assert oracle.lower_median(a) == 2.0 == float(torch.median(torch.from_numpy(a)))
    a = np.arange(9.0).reshape(3, 3)
    assert oracle.lower_median(a) == 4.0


def test_gaussian_kernel_golden():
    g = load_golden("gaussian_20x5")
    A, B = g["A"].astype(np.float64), g["B"].astype(np.float64)
    K, sq, h, med = oracle.gaussian_gram(A, B)
    assert np.abs(K - g["K_f64"]).max() < 1e-12
    dK, _ = oracle.sig_space_grad(A, B, K, sq, h, med, median_grad=False)
    assert rel_max(dK, g["dK_f64"]) < 1e-10
    Kh, sq, _, _ = oracle.gaussian_gram(A, B, h=0.7)
    assert np.abs(Kh - g["Kh_f64"]).max() < 1e-12


@pytest.mark.skipif(not ref_shim.reference_available(), reason="reference tree not present (GPU box)")
def test_live_reference_import_agrees_with_oracle():
    """Container-only: run the unmodified reference (src.kernels / src.inference) on fresh inputs."""
    ref_shim.import_reference()
    from src.inference import SVGD
    from src.kernels import PathSigKernel

    gen = torch.Generator().manual_seed(123)
    N, T, d = 30, 12, 3
    X = torch.cumsum(0.1 * torch.randn(N, T, d, generator=gen, dtype=torch.float64), 1)
    G = torch.randn(N, T * d, generator=gen, dtype=torch.float64)
    sampler = SVGD(kernel=PathSigKernel(), optimizer_class=None, lr=0.1)
    vel, it = sampler._velocity(X, G)
    cf = oracle.phi_closed_form(X.numpy(), G.numpy(), 3)
    assert np.abs(cf["K"] - it["k_xx"].numpy()).max() < 1e-12
    assert rel_max(-cf["phi"], vel.numpy()) < 1e-9
    torch.autograd.set_detect_anomaly(False)

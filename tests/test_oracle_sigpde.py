"""Known-answer tests of the signature-PDE oracle (oracle/sigpde_oracle.py).  The
arithmetic's home (``sigkernel`` @ 3b23739) is absent, so the restatement is pinned by closed forms."""
import numpy as np
import pytest
import torch
from scipy.special import i0

from oracle import sigpde_oracle as so


def _lines(u, v, M, N):
    X = np.linspace(0.0, 1.0, M)[None, :, None] * np.asarray(u, float)[None, None, :]
    Y = np.linspace(0.0, 1.0, N)[None, :, None] * np.asarray(v, float)[None, None, :]
    return X, Y


@pytest.mark.parametrize("u,v", [([1.0, 0.5], [0.3, 0.8]), ([2.0, -1.0, 0.5], [1.0, 0.2, 1.5]), ([0.4], [0.4])])
def test_linear_static_kernel_on_straight_lines_is_a_bessel_function(u, v):
    """k = sum_n <u,v>^n / (n!)^2 = I0(2 sqrt(<u,v>)) (J0 for a negative inner product)."""
    X, Y = _lines(u, v, 9, 7)
    uv = float(np.dot(u, v))
    want = i0(2.0 * np.sqrt(uv)) if uv >= 0 else None
    errs = []
    for order in (0, 2, 4):
        k = so.sigpde_gram(X, Y, order, so.LINEAR)[0, 0]
        errs.append(abs(k - want) / want)
    assert errs[2] < 5e-5 and errs[0] > 10 * errs[1] > 100 * errs[2]  # second-order convergence under refinement


def test_symmetry_and_resampling_invariance():
    rng = np.random.default_rng(0)
    X = np.cumsum(0.3 * rng.standard_normal((3, 6, 2)), 1)
    k = so.sigpde_gram(X, X, 2, so.RBF, h=1.3)
    assert np.allclose(k, k.T, rtol=1e-12) and (k > 0).all()
    # inserting the midpoints of a piecewise-linear path does not change the path: same limit
    mid = 0.5 * (X[:, 1:] + X[:, :-1])
    X2 = np.empty((3, 11, 2))
    X2[:, 0::2], X2[:, 1::2] = X, mid
    with_lin = so.sigpde_gram(X2, X2, 3, so.LINEAR), so.sigpde_gram(X, X, 4, so.LINEAR)
    assert np.allclose(*with_lin, rtol=2e-4)


def test_median_bandwidth_convention():
    rng = np.random.default_rng(1)
    X, Y = rng.standard_normal((4, 5, 2)), rng.standard_normal((3, 6, 2))
    _, dist = so.static_gram(X, Y, so.RBF, 1.0)
    h, med = so.bw_median_points(dist)
    ref = torch.median(torch.as_tensor(dist))  # lower median of all A*B*M*N entries
    assert med == float(ref) and h == pytest.approx(np.sqrt(med / np.log(4 + 1.0)), rel=1e-7)


@pytest.mark.parametrize("kind,h", [(so.RBF, 0.9), (so.LINEAR, 0.7)])
def test_backward_formula_approaches_the_exact_discrete_derivative(kind, h):
    rng = np.random.default_rng(2)
    X = np.cumsum(0.4 * rng.standard_normal((2, 5, 2)), 1)
    Y = np.cumsum(0.4 * rng.standard_normal((3, 4, 2)), 1)
    gout = rng.standard_normal((2, 3))
    errs = []
    for order in (0, 2, 4):
        Xt = torch.as_tensor(X, dtype=torch.float64).requires_grad_(True)
        k = so.sigpde_gram_torch(Xt, torch.as_tensor(Y), order, kind, h)
        assert np.allclose(k.detach().numpy(), so.sigpde_gram(X, Y, order, kind, h), rtol=1e-12)
        exact = torch.autograd.grad((k * torch.as_tensor(gout)).sum(), Xt)[0].numpy()
        got = so.sigpde_gram_bwd(X, Y, gout, order, kind, h)
        errs.append(np.abs(got - exact).max() / np.abs(exact).max())
    # first-order convergence (the formula pairs K at a cell's lower corner with K_rev at its upper corner)
    assert errs[1] < 0.4 * errs[0] and errs[2] < 0.3 * errs[1] and errs[2] < 3e-2


@pytest.mark.parametrize("name", ["maze", "small", "panda"])
def test_static_kernel_and_bandwidth_convention_vs_reference_golden(name):
    """tests/golden/sigpde_static.npz: the reference's own BatchGaussianKernel.Gram_matrix (median
    heuristic over the [A,B,M,N] tensor with ln(A + 1); ``exp(-dist / h)``), pushed through the solver."""
    from tests.util import load_golden

    g = load_golden("sigpde_static")
    X, Y, order = g[f"{name}_X"], g[f"{name}_Y"], int(g[f"{name}_order"])
    _, dist = so.static_gram(X, Y, so.RBF, 1.0)
    h, _ = so.bw_median_points(dist)
    assert h == pytest.approx(float(g[f"{name}_h_med"]), rel=1e-9)
    assert np.allclose(so.sigpde_gram(X, Y, order, so.RBF, h), g[f"{name}_k_med"], rtol=1e-9)
    assert np.allclose(so.sigpde_gram(X, Y, order, so.RBF, 0.75), g[f"{name}_k_fix"], rtol=1e-12)

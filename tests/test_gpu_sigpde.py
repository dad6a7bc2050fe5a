"""GPU parity tests of the signature-PDE kernel (sigpde.cu) through the C ABI against the
float64 oracle (oracle/sigpde_oracle.py), plus the reference-shaped wrappers (`SignatureKernel`,
`compat.PdeSigKernel`).  Tolerance: 1e-9 relative (both sides are float64; the bandwidth is a float32
record, so tests of the median route use 1e-6)."""
import numpy as np
import pytest
import torch
from scipy.special import i0

from oracle import sigpde_oracle as so
from tests.util import rel_max

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    from sigsvgd_b200 import _lib, compat, ops
    from sigsvgd_b200.inference import FixedBandwidth
    from sigsvgd_b200.kernels import SignatureKernel

DEV = "cuda:0"
TOL = 1e-9


def _paths(A, M, D, seed, step=0.3):
    return np.cumsum(step * np.random.default_rng(seed).standard_normal((A, M, D)), 1)


def cu(a):
    return torch.as_tensor(np.ascontiguousarray(a), device=DEV)


SHAPES = [  # A, B, M, N, D, order, kind, h
    (3, 4, 5, 7, 2, 0, so.RBF, 0.9), (2, 3, 6, 4, 3, 2, so.RBF, 1.7), (5, 2, 2, 2, 1, 3, so.RBF, 0.5),
    (2, 9, 40, 35, 2, 0, so.RBF, 2.0),    # more than one strip of 32 fine rows, more than 32 columns
    (3, 3, 12, 9, 7, 2, so.LINEAR, 0.05),  # 44 x 32 fine grid
    (35, 35, 30, 30, 2, 0, so.RBF, 32 ** 0.5),  # the particle-maze shapes (dust.py / particle_maze_config.yaml)
    (4, 5, 64, 64, 7, 0, so.RBF, 30.0),   # Panda-shaped paths
    (1, 17, 8, 8, 2, 1, so.LINEAR, 1.0),  # more pairs than warps in a block
]


@pytest.mark.parametrize("A,B,M,N,D,order,kind,h", SHAPES)
def test_gram_and_gradient_vs_oracle(A, B, M, N, D, order, kind, h):
    X, Y = _paths(A, M, D, 1), _paths(B, N, D, 2)
    want = so.sigpde_gram(X, Y, order, kind, h)
    got = ops.sigpde_gram_raw(cu(X), cu(Y), order, kind, h)
    assert got.shape == (A, B) and rel_max(got.cpu().numpy(), want) <= TOL
    gout = np.random.default_rng(3).standard_normal((A, B))
    gx_want = so.sigpde_gram_bwd(X, Y, gout, order, kind, h)
    gx = ops.sigpde_gram_bwd_raw(cu(X), cu(Y), cu(gout), order, kind, h)
    assert rel_max(gx.cpu().numpy(), gx_want) <= TOL
    # second call: the workspace and shared-memory state of the first leave nothing behind
    assert torch.equal(ops.sigpde_gram_raw(cu(X), cu(Y), order, kind, h), got)


def test_known_answer_linear_kernel_on_straight_lines():
    u, v = np.array([1.0, 0.5, -0.2]), np.array([0.3, 0.8, 0.4])
    X = np.linspace(0, 1, 17)[None, :, None] * u[None, None, :]
    Y = np.linspace(0, 1, 9)[None, :, None] * v[None, None, :]
    k = float(ops.sigpde_gram_raw(cu(X), cu(Y), 4, _lib.PDE_LINEAR, 1.0)[0, 0])
    assert abs(k - i0(2 * np.sqrt(u @ v))) <= 1e-5 * k


@pytest.mark.parametrize("A,B,M,N,D", [(4, 3, 5, 6, 2), (35, 35, 30, 30, 2), (6, 6, 64, 64, 7)])
def test_median_bandwidth_over_all_point_pairs(A, B, M, N, D):
    X, Y = _paths(A, M, D, 4), _paths(B, N, D, 5)
    _, dist = so.static_gram(X, Y, so.RBF, 1.0)
    h, med = so.bw_median_points(dist)
    state = ops.sigpde_bandwidth_raw(cu(X), cu(Y))
    rec = ops.read_state(state)
    assert rec["median"] == np.float32(med) and abs(rec["h"] - h) <= 1e-6 * h
    want = so.sigpde_gram(X, Y, 1, so.RBF, float(np.float32(rec["h"])))
    got = ops.sigpde_gram_raw(cu(X), cu(Y), 1, _lib.PDE_RBF, 1.0, state)
    assert rel_max(got.cpu().numpy(), want) <= TOL


@pytest.mark.parametrize("name", ["maze", "small", "panda"])
def test_vs_reference_golden_static_kernel(name):
    """The reference's own BatchGaussianKernel.Gram_matrix output (median heuristic and a constant
    bandwidth), pushed through the solver restatement: tests/golden/sigpde_static.npz."""
    from tests.util import load_golden

    g = load_golden("sigpde_static")
    X, Y, order = g[f"{name}_X"], g[f"{name}_Y"], int(g[f"{name}_order"])
    k_fix = ops.sigpde_gram(cu(X), cu(Y), order, _lib.PDE_RBF, 0.75)
    assert rel_max(k_fix.cpu().numpy(), g[f"{name}_k_fix"]) <= TOL
    k_med = ops.sigpde_gram(cu(X), cu(Y), order)  # h=None: the median heuristic over all point pairs
    assert rel_max(k_med.cpu().numpy(), g[f"{name}_k_med"]) <= 1e-6  # (the bandwidth record is float32)


def test_signature_kernel_class_runs_the_reference_statements():
    """score.py:64-76: k_xx = kernel(x, x.detach()); grad_k = ag(k_xx.sum(), x)[0] (fp32 in, fp32 out)."""
    X = _paths(12, 10, 2, 6).astype(np.float32)
    x = cu(X).requires_grad_(True)
    kernel = SignatureKernel(depth=2)  # median heuristic over all point pairs, dyadic_order = depth
    k_xx = kernel(x, x.detach())
    grad_k = torch.autograd.grad(k_xx.sum(), x)[0]
    assert k_xx.dtype == torch.float32 and grad_k.shape == x.shape
    X64 = X.astype(np.float64)
    h, _ = so.bw_median_points(so.static_gram(X64, X64, so.RBF, 1.0)[1])
    assert rel_max(k_xx.detach().cpu().numpy(), so.sigpde_gram(X64, X64, 2, so.RBF, h)) <= 1e-5
    assert rel_max(grad_k.cpu().numpy(), so.sigpde_gram_bwd(X64, X64, np.ones((12, 12)), 2, so.RBF, h)) <= 1e-5
    kf = SignatureKernel(bandwidth_fn=FixedBandwidth(0.8), depth=1)(x, x.detach())
    assert rel_max(kf.detach().cpu().numpy(), so.sigpde_gram(X64, X64, 1, so.RBF, 0.8)) <= 1e-5
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        kernel(x.cpu(), x.cpu())


def test_sigkernel_shaped_adapter_in_the_mpc_statements():
    """trajectory_svgd.py:55-69 at the maze shapes: compute_Gram(tau.double(), tau.detach().double(), sym=False)
    .float(), gradient reaching the actions through a rollout, .mean(0) over the action samples."""
    n_samples, n_pol, hz = 4, 35, 30
    g = torch.Generator().manual_seed(7)
    actions = torch.randn(n_samples, n_pol, hz, 2, generator=g).to(DEV).requires_grad_(True)
    traj = torch.cumsum(torch.cumsum(0.05 * actions, 2), 2)  # toy double integrator
    tau = traj.mean(0)
    kernel = compat.PdeSigKernel(compat.RBFKernel(sigma=(2 + hz) ** 0.5), dyadic_order=0)
    k_xx = kernel.compute_Gram(tau.double(), tau.detach().double(), sym=False).float()
    grad_k = torch.autograd.grad(k_xx.sum(), actions, retain_graph=True)[0].mean(0).flatten(1)
    assert k_xx.shape == (n_pol, n_pol) and grad_k.shape == (n_pol, hz * 2)
    # the same through the oracle's exact chain: d k / d tau pushed through the rollout by torch
    tau64 = tau.detach().double().cpu().numpy()
    want_k = so.sigpde_gram(tau64, tau64, 0, so.RBF, (2 + hz) ** 0.5)
    assert rel_max(k_xx.detach().cpu().numpy(), want_k) <= 1e-6
    gtau = so.sigpde_gram_bwd(tau64, tau64, np.ones((n_pol, n_pol)), 0, so.RBF, (2 + hz) ** 0.5)
    want = torch.autograd.grad(tau, actions, grad_outputs=cu(gtau).float())[0].mean(0).flatten(1)
    assert rel_max(grad_k.cpu().numpy(), want.cpu().numpy()) <= 1e-5


def test_errors():
    x = cu(_paths(2, 170, 2, 8))
    with pytest.raises(RuntimeError, match="shared memory"):
        ops.sigpde_gram_raw(x, x, 0, _lib.PDE_RBF, 1.0)  # 170 x 170 doubles per pair > 200 KB
    with pytest.raises(ValueError, match="bandwidth h must be > 0"):
        ops.sigpde_gram_raw(x[:, :4], x[:, :4], 0, _lib.PDE_RBF, 0.0)
    with pytest.raises(ValueError, match="bad shape"):
        ops.sigpde_gram_raw(x[:, :1], x[:, :4], 0, _lib.PDE_RBF, 1.0)

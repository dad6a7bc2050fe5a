"""GPU parity of the fused occupancy / self-collision network query through the C ABI
(``sigsvgd_mlp_pack`` / ``sigsvgd_mlp_query``) against the golden vectors of the unmodified reference classes
and against the float64 oracle.

Tolerance: the kernel multiplies fp16 operands and accumulates in fp32 (the reference: fp32 throughout).  With
200-wide layers the logit carries a relative rounding error of about 2^-11; stated bar: probabilities within
2e-3 absolute, gradients within 1 % of the largest gradient entry of the batch (measured: see
profiles/README.md)."""
import os

import numpy as np
import pytest
import torch

from oracle import mlp_oracle

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    from sigsvgd_b200 import occupancy

DEV = "cuda:0"
GOLD = os.path.join(os.path.dirname(__file__), "golden", "occupancy_mlp.npz")
TOL_PROB, TOL_GRAD = 2e-3, 1e-2


def _net(g, name):
    n = sum(1 for k in g.files if k.startswith(f"{name}_w"))
    return [torch.tensor(g[f"{name}_w{i}"]) for i in range(n)], [torch.tensor(g[f"{name}_b{i}"]) for i in range(n)]


def _check(prob, grad, want_prob, want_grad):
    assert np.isfinite(prob).all() and np.isfinite(grad).all()
    assert np.abs(prob - want_prob).max() <= TOL_PROB
    assert np.abs(grad - want_grad).max() <= TOL_GRAD * np.abs(want_grad).max()


@pytest.mark.parametrize("name", ["occ", "col"])
def test_golden_reference_networks(name):
    g = np.load(GOLD)
    ws, bs = _net(g, name)
    f = occupancy.FusedOccupancy(ws, bs, DEV)
    x = torch.tensor(g[f"{name}_x"], device=DEV)
    prob, grad = f.query_raw(x)
    _check(prob.cpu().numpy(), grad.cpu().numpy(), g[f"{name}_y"][:, 0], g[f"{name}_grad"])
    # forward only (no gradient buffer): the same probabilities, bit for bit
    prob2, none = f.query_raw(x, want_grad=False)
    assert none is None and torch.equal(prob, prob2)


@pytest.mark.parametrize("in_dim,hidden,n_mid,P", [(3, 200, 3, 1), (3, 200, 3, 127), (3, 200, 3, 129), (7, 200, 3, 40000),
                                                   (8, 208, 4, 1000), (1, 16, 1, 300), (5, 97, 2, 513)])
def test_against_oracle_shapes(in_dim, hidden, n_mid, P):
    rng = np.random.default_rng(in_dim * 1000 + hidden + n_mid)
    dims = [in_dim] + [hidden] * (n_mid + 1) + [1]
    ws = [rng.standard_normal((o, i)) * (1.5 / np.sqrt(i)) for i, o in zip(dims[:-1], dims[1:])]
    bs = [0.3 * rng.standard_normal(o) for o in dims[1:]]
    x = rng.standard_normal((P, in_dim)).astype(np.float32)
    f = occupancy.FusedOccupancy([torch.tensor(w, dtype=torch.float32) for w in ws],
                                 [torch.tensor(b, dtype=torch.float32) for b in bs], DEV)
    prob, grad = f.query_raw(torch.tensor(x, device=DEV))
    ws32 = [w.astype(np.float32) for w in ws]
    bs32 = [b.astype(np.float32) for b in bs]
    want_prob, want_grad = mlp_oracle.mlp_query(x, ws32, bs32)
    _check(prob.cpu().numpy(), grad.cpu().numpy(), want_prob, want_grad)


def test_drop_in_module_call_and_autograd():
    torch.manual_seed(0)
    layers = [torch.nn.Linear(3, 200), torch.nn.ReLU()]
    for _ in range(3):
        layers += [torch.nn.Linear(200, 200), torch.nn.ReLU()]
    layers += [torch.nn.Linear(200, 1), torch.nn.Sigmoid()]

    class Net(torch.nn.Module):  # the shape of ContinuousOccupancyMap (continuous_occupancy_map.py:10-35)
        def __init__(self):
            super().__init__()
            self.layers = torch.nn.Sequential(*layers)

        def forward(self, *args):
            return self.layers(*args)

    net = Net().to(DEV)
    with torch.no_grad():
        net.layers[-2].weight.mul_(30.0)
    f = occupancy.fuse(net)
    # called as batch_cost_function calls it: [n_body_points, batch * timesteps, 3] -> [..., 1], then reductions
    pts = torch.randn(54, 600, 3, device=DEV, requires_grad=True)
    w = torch.rand(54, 600, device=DEV)
    cost_ref = (net(pts).squeeze(-1) * w).sum()
    (g_ref,) = torch.autograd.grad(cost_ref, pts)
    out = f(pts)
    assert out.shape == (54, 600, 1)
    cost = (out.squeeze(-1) * w).sum()
    (g,) = torch.autograd.grad(cost, pts)
    assert abs(float(cost) - float(cost_ref)) <= 1e-3 * abs(float(cost_ref))
    assert float((g - g_ref).abs().max()) <= TOL_GRAD * float(g_ref.abs().max())
    # no gradient requested: no gradient buffer is formed
    with torch.no_grad():
        assert float((f(pts) - net(pts)).abs().max()) <= TOL_PROB


def test_errors():
    w = lambda o, i: torch.zeros(o, i)  # noqa: E731
    b = lambda o: torch.zeros(o)  # noqa: E731
    with pytest.raises(ValueError, match="unsupported shape"):
        occupancy.FusedOccupancy([w(300, 3), w(300, 300), w(1, 300)], [b(300), b(300), b(1)], DEV)
    with pytest.raises(ValueError, match="unsupported shape"):
        occupancy.FusedOccupancy([w(20, 9), w(20, 20), w(1, 20)], [b(20), b(20), b(1)], DEV)
    f = occupancy.FusedOccupancy([w(20, 3), w(20, 20), w(1, 20)], [b(20), b(20), b(1)], DEV)
    with pytest.raises(ValueError, match="points"):
        f.query_raw(torch.zeros(4, 5, device=DEV))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        f.query_raw(torch.zeros(4, 3))
    prob, grad = f.query_raw(torch.zeros(0, 3, device=DEV))
    assert prob.numel() == 0 and grad.numel() == 0
    prob, _ = f.query_raw(torch.randn(10, 3, device=DEV))
    assert torch.allclose(prob, torch.full_like(prob, 0.5))  # all-zero weights

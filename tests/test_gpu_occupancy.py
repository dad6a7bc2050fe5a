"""GPU parity of the fused occupancy / self-collision network query through the C ABI
(``sigsvgd_mlp_pack`` / ``sigsvgd_mlp_query``) against the golden vectors of the unmodified reference classes
and against the float64 oracle.

The networks are piecewise linear up to the sigmoid: the Jacobian is constant on each ReLU pattern and jumps
between them, so a gradient can only be compared where both sides see the same pattern.

* precise mode (fp16 hi + lo planes, 3 MMAs per product = the reference's fp32 GEMMs to fp32 accuracy):
  probabilities within 2e-5; gradients within 2e-4 of the batch's largest gradient entry at every point, except
  at most max(2, P / 5000) points whose pre-activations lie within fp32 rounding of zero (the reference's own
  fp32 result is arbitrary there: it differs from ITS float64 evaluation at the same rate).
* fast mode (hi planes only): compared with the float64 restatement of that arithmetic contract
  (``mlp_oracle.mlp_query_fp16_contract``: operands rounded to half precision, exact accumulation): 92 % of
  the probabilities within 2e-5 and all within 1e-3 (an activation that lands on a half-precision rounding
  boundary may round the other way after fp32 accumulation), gradients within 2e-4 at all but 1 % of the
  points; and with the reference network itself: probabilities within 2e-3; the median gradient deviation below
  1e-3 of the largest entry, at most 10 % of the points on a neighbouring linear piece (deviation > 1 %)."""
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
TOL_PROB, TOL_GRAD = 2e-5, 2e-4
TOL_PROB_FAST = 2e-3


def _net(g, name):
    n = sum(1 for k in g.files if k.startswith(f"{name}_w"))
    return [torch.tensor(g[f"{name}_w{i}"]) for i in range(n)], [torch.tensor(g[f"{name}_b{i}"]) for i in range(n)]


def _dev(grad, want_grad):
    return np.abs(grad - want_grad).max(axis=1) / np.abs(want_grad).max()


def _check_tight(prob, grad, want_prob, want_grad):
    assert np.isfinite(prob).all() and np.isfinite(grad).all()
    assert np.abs(prob - want_prob).max() <= TOL_PROB
    off = int((_dev(grad, want_grad) > TOL_GRAD).sum())
    assert off <= max(2, len(prob) // 5000), off


def _check_contract(prob, grad, want_prob, want_grad):
    assert np.isfinite(prob).all() and np.isfinite(grad).all()
    dp = np.abs(prob - want_prob)
    assert dp.max() <= 1e-3 and (dp > TOL_PROB).mean() <= 0.08
    off = int((_dev(grad, want_grad) > TOL_GRAD).sum())
    assert off <= max(2, len(prob) // 100), off


def _check_fast_vs_reference(prob, grad, want_prob, want_grad):
    assert np.abs(prob - want_prob).max() <= TOL_PROB_FAST
    dev = _dev(grad, want_grad)
    assert np.median(dev) <= 1e-3 and (dev > 1e-2).mean() <= 0.10


@pytest.mark.parametrize("name", ["occ", "col"])
def test_golden_reference_networks(name):
    g = np.load(GOLD)
    ws, bs = _net(g, name)
    f = occupancy.FusedOccupancy(ws, bs, DEV)
    x = torch.tensor(g[f"{name}_x"], device=DEV)
    prob, grad = f.query_raw(x)
    _check_tight(prob.cpu().numpy(), grad.cpu().numpy(), g[f"{name}_y"][:, 0], g[f"{name}_grad"])
    # forward only (no gradient buffer): the same probabilities, bit for bit
    prob2, none = f.query_raw(x, want_grad=False)
    assert none is None and torch.equal(prob, prob2)
    # fast mode: exactly its stated arithmetic, and close to the reference
    prob_f, grad_f = f.query_raw(x, precise=False)
    w_np, b_np = [w.numpy() for w in ws], [b.numpy() for b in bs]
    c_prob, c_grad = mlp_oracle.mlp_query_fp16_contract(g[f"{name}_x"], w_np, b_np)
    _check_contract(prob_f.cpu().numpy(), grad_f.cpu().numpy(), c_prob, c_grad)
    _check_fast_vs_reference(prob_f.cpu().numpy(), grad_f.cpu().numpy(), g[f"{name}_y"][:, 0], g[f"{name}_grad"])


@pytest.mark.parametrize("precise", [True, False])
@pytest.mark.parametrize("in_dim,hidden,n_mid,P", [(3, 200, 3, 1), (3, 200, 3, 127), (3, 200, 3, 129), (7, 200, 3, 40000),
                                                   (8, 208, 4, 1000), (1, 16, 1, 300), (5, 97, 2, 513)])
def test_against_oracle_shapes(in_dim, hidden, n_mid, P, precise):
    rng = np.random.default_rng(in_dim * 1000 + hidden + n_mid)
    dims = [in_dim] + [hidden] * (n_mid + 1) + [1]
    ws = [(rng.standard_normal((o, i)) * (1.5 / np.sqrt(i))).astype(np.float32) for i, o in zip(dims[:-1], dims[1:])]
    bs = [(0.3 * rng.standard_normal(o)).astype(np.float32) for o in dims[1:]]
    x = rng.standard_normal((P, in_dim)).astype(np.float32)
    f = occupancy.FusedOccupancy([torch.tensor(w) for w in ws], [torch.tensor(b) for b in bs], DEV, precise=precise)
    prob, grad = f.query_raw(torch.tensor(x, device=DEV))
    if precise:
        _check_tight(prob.cpu().numpy(), grad.cpu().numpy(), *mlp_oracle.mlp_query(x, ws, bs))
    else:
        _check_contract(prob.cpu().numpy(), grad.cpu().numpy(), *mlp_oracle.mlp_query_fp16_contract(x, ws, bs))


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
    assert abs(float(cost) - float(cost_ref)) <= 1e-5 * abs(float(cost_ref))
    dev = (g - g_ref).abs().amax(-1).flatten() / g_ref.abs().max()
    assert int((dev > TOL_GRAD).sum()) <= max(2, dev.numel() // 5000)
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

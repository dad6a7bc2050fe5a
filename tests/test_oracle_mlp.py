"""The numpy restatement of the occupancy / self-collision networks (oracle/mlp_oracle.py) against the golden
vectors the UNMODIFIED reference classes produced (oracle/make_golden_mlp.py), and against torch autograd."""
import os

import numpy as np
import torch

from oracle import mlp_oracle

GOLD = os.path.join(os.path.dirname(__file__), "golden", "occupancy_mlp.npz")


def _net(g, name):
    n = sum(1 for k in g.files if k.startswith(f"{name}_w"))
    return [g[f"{name}_w{i}"] for i in range(n)], [g[f"{name}_b{i}"] for i in range(n)]


def test_oracle_matches_reference_networks():
    g = np.load(GOLD)
    for name in ("occ", "col"):
        ws, bs = _net(g, name)
        assert [w.shape[0] for w in ws] == [200, 200, 200, 200, 1]  # continuous_occupancy_map.py:11 defaults
        prob, grad = mlp_oracle.mlp_query(g[f"{name}_x"], ws, bs)
        # the reference ran float32; the restatement float64
        assert np.abs(prob - g[f"{name}_y"][:, 0]).max() < 2e-6
        assert np.abs(grad - g[f"{name}_grad"]).max() < 2e-6 * np.abs(grad).max() + 1e-9


def test_oracle_gradient_is_the_autograd_gradient():
    rng = np.random.default_rng(0)
    dims = [5, 33, 33, 33, 1]
    ws = [rng.standard_normal((o, i)) for i, o in zip(dims[:-1], dims[1:])]
    bs = [rng.standard_normal(o) for o in dims[1:]]
    x = torch.tensor(rng.standard_normal((64, 5)), requires_grad=True)
    h = x
    for i, (w, b) in enumerate(zip(ws, bs)):
        h = h @ torch.tensor(w).T + torch.tensor(b)
        h = torch.relu(h) if i + 1 < len(ws) else torch.sigmoid(h)
    (gx,) = torch.autograd.grad(h.sum(), x)
    prob, grad = mlp_oracle.mlp_query(x.detach().numpy(), ws, bs)
    np.testing.assert_allclose(prob, h.detach().numpy()[:, 0], rtol=1e-12, atol=1e-14)
    np.testing.assert_allclose(grad, gx.numpy(), rtol=1e-10, atol=1e-13)


def test_linear_stack_reads_the_reference_layout():
    from sigsvgd_b200 import occupancy

    layers = [torch.nn.Linear(3, 16), torch.nn.ReLU(), torch.nn.Linear(16, 16), torch.nn.ReLU(), torch.nn.Linear(16, 1),
              torch.nn.Sigmoid()]

    class Net(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.layers = torch.nn.Sequential(*layers)

    ws, bs = occupancy.linear_stack(Net())
    assert [tuple(w.shape) for w in ws] == [(16, 3), (16, 16), (1, 16)] and len(bs) == 3
    import pytest

    with pytest.raises(ValueError):
        occupancy.linear_stack(torch.nn.Sequential(torch.nn.Linear(3, 4), torch.nn.Tanh(), torch.nn.Linear(4, 1)))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        occupancy.fuse(Net())

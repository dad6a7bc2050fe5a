"""GPU parity of the serial-chain forward kinematics through the C ABI (``sigsvgd_chain_fk`` /
``sigsvgd_chain_fk_bwd``) on the reference's Panda chain (tests/golden/panda_chain.npz) against the float64 oracle.
Tolerances: positions 2e-6 m (fp32 sin/cos along nine frames, arm reach ~1 m); gradients 1e-5 of the largest
entry."""
import os

import numpy as np
import pytest
import torch

from oracle import fk_oracle

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    from sigsvgd_b200 import kinematics

DEV = "cuda:0"
GOLD = os.path.join(os.path.dirname(__file__), "golden", "panda_chain.npz")


def _chain():
    g = np.load(GOLD)
    return g, kinematics.SerialChain(g["xyz"], g["rpy"], g["axis"], g["dof"], list(g["link_names"]))


def test_positions_match_the_oracle():
    g, sc = _chain()
    xs = sc(torch.tensor(g["q"], dtype=torch.float32, device=DEV))
    assert xs.shape == (9, 64, 3)
    assert np.abs(xs.cpu().numpy() - g["xs"]).max() <= 2e-6
    # the layout of qs_to_joints_xs with leading batch dimensions, and a choice of links
    q = torch.tensor(g["q"], dtype=torch.float32, device=DEV).reshape(4, 16, 7)
    sel = sc(q, links=["panda_link8", "panda_hand"])
    assert sel.shape == (2, 4, 16, 3) and torch.equal(sel.reshape(2, 64, 3), xs[[7, 8]])


def test_gradient_matches_autograd_of_the_oracle():
    g, sc = _chain()
    rng = np.random.default_rng(3)
    q64 = rng.uniform(g["limits_lo"], g["limits_hi"], size=(3000, 7))
    w = rng.standard_normal((9, 3000, 3))
    qt = torch.tensor(q64, requires_grad=True)
    ref = fk_oracle.chain_fk_torch(g["xyz"], g["rpy"], g["axis"], g["dof"], qt)
    (g_ref,) = torch.autograd.grad((ref * torch.tensor(w)).sum(), qt)
    q = torch.tensor(q64, dtype=torch.float32, device=DEV, requires_grad=True)
    xs = sc(q)
    assert float((xs.detach().cpu().double() - ref.detach()).abs().max()) <= 2e-6
    (gq,) = torch.autograd.grad((xs * torch.tensor(w, dtype=torch.float32, device=DEV)).sum(), q)
    assert float((gq.cpu().double() - g_ref).abs().max()) <= 1e-5 * float(g_ref.abs().max())


def test_other_chains_and_errors():
    # a planar two-link arm about y with a fixed tool offset, second joint driven by column 0 and first by column 1
    sc = kinematics.SerialChain([[0, 0, 0], [1, 0, 0], [0.5, 0, 0]], [[0, 0, 0]] * 3, [[0, 1, 0], [0, 1, 0], [0, 0, 0]],
                                [1, 0, -1])
    q = torch.tensor([[0.3, -0.2], [1.0, 0.5]], device=DEV)
    want = fk_oracle.chain_fk(sc.xyz, [[0, 0, 0]] * 3, [[0, 1, 0], [0, 1, 0], [0, 0, 1]], [1, 0, -1], q.cpu().numpy())
    assert np.abs(sc(q).cpu().numpy() - want).max() <= 1e-6
    assert sc(torch.zeros(0, 2, device=DEV)).shape == (3, 0, 3)
    with pytest.raises(ValueError, match="configurations"):
        sc(torch.zeros(4, 3, device=DEV))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        sc(torch.zeros(4, 2))
    with pytest.raises(ValueError, match="joints"):
        big = kinematics.SerialChain(np.zeros((17, 3)), np.zeros((17, 3)), np.tile([0, 0, 1.0], (17, 1)), np.arange(17) % 4)
        big(torch.zeros(2, 4, device=DEV))

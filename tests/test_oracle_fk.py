"""Known answers for the forward-kinematics oracle (oracle/fk_oracle.py) on the reference's Panda chain
(tests/golden/panda_chain.npz, extracted from robot_resources/panda/urdf/panda.urdf by oracle/make_golden_fk.py),
and the host side of ``sigsvgd_b200.kinematics`` (URDF parsing; no GPU)."""
import os

import numpy as np
import pytest
import torch

from oracle import fk_oracle

GOLD = os.path.join(os.path.dirname(__file__), "golden", "panda_chain.npz")
URDF = "/root/reference/robot_resources/panda/urdf/panda.urdf"


def _chain():
    g = np.load(GOLD)
    return g, (g["xyz"], g["rpy"], g["axis"], g["dof"])


def test_franka_flange_at_zero_configuration():
    g, ch = _chain()
    names = list(g["link_names"])
    xs = fk_oracle.chain_fk(*ch, np.zeros((1, 7)))
    # published kinematics of the arm: flange 0.088 m ahead, 0.333 + 0.316 + 0.384 - 0.107 m up
    np.testing.assert_allclose(xs[names.index("panda_link8"), 0], [0.088, 0.0, 0.926], atol=1e-9)
    np.testing.assert_allclose(xs[names.index("panda_link1"), 0], [0.0, 0.0, 0.333], atol=1e-12)
    np.testing.assert_allclose(xs, g["xs"][:, :1], atol=1e-12)


def test_invariants_in_every_configuration():
    g, ch = _chain()
    xs = fk_oracle.chain_fk(*ch, g["q"])
    # consecutive frames stay at the URDF offsets' distances
    d = np.linalg.norm(np.diff(np.concatenate([np.zeros((1,) + xs.shape[1:]), xs]), axis=0), axis=-1)
    np.testing.assert_allclose(d, np.broadcast_to(np.linalg.norm(ch[0], axis=1)[:, None], d.shape), atol=1e-12)
    # a link's position does not depend on its own joint (nor on later ones)
    q2 = g["q"].copy()
    q2[:, 3:] += 0.7
    xs2 = fk_oracle.chain_fk(*ch, q2)
    np.testing.assert_allclose(xs2[:4], xs[:4], atol=1e-12)  # links 1..4 are moved by joints 1..3 only
    assert np.abs(xs2[5:] - xs[5:]).max() > 1e-2


def test_torch_recursion_is_the_same_function():
    g, ch = _chain()
    q = torch.tensor(g["q"], requires_grad=True)
    xs = fk_oracle.chain_fk_torch(*ch, q)
    np.testing.assert_allclose(xs.detach().numpy(), g["xs"], atol=1e-12)
    (gq,) = torch.autograd.grad(xs[-1, :, 2].sum(), q)  # height of the hand
    eps = 1e-6
    qp = g["q"].copy()
    qp[:, 1] += eps
    fd = (fk_oracle.chain_fk(*ch, qp)[-1, :, 2] - g["xs"][-1, :, 2]) / eps
    np.testing.assert_allclose(gq[:, 1].numpy(), fd, atol=1e-5)
    assert np.abs(gq[:, 6].numpy()).max() < 1e-12  # the last joint only spins the flange about its own axis


@pytest.mark.skipif(not os.path.exists(URDF), reason="reference tree absent")
def test_serial_chain_from_the_reference_urdf():
    from sigsvgd_b200 import kinematics

    g, ch = _chain()
    sc = kinematics.SerialChain.from_urdf(URDF, "panda_hand", [f"panda_joint{i}" for i in range(1, 8)])
    assert sc.link_names == list(g["link_names"]) and sc.n_joints == 9 and sc.n_dof == 7
    np.testing.assert_allclose(sc.xyz, g["xyz"], atol=1e-7)
    np.testing.assert_array_equal(sc.dof, g["dof"])
    np.testing.assert_allclose(sc.rot.reshape(-1, 3, 3), np.stack([fk_oracle.rpy_matrix(*r) for r in g["rpy"]]), atol=1e-7)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        sc(torch.zeros(2, 7))

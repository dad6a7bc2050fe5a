"""The natural-cubic-spline matrices (sigsvgd_b200/spline.py; reference: examples/script_planning_robot.py:118-127
through torchcubicspline, absent here) against scipy's natural CubicSpline -- the natural spline through given
knots is unique, so any correct implementation agrees to rounding."""
import numpy as np
import pytest
import torch
from scipy.interpolate import CubicSpline

from sigsvgd_b200 import spline


@pytest.mark.parametrize("K,T", [(2, 5), (3, 50), (5, 200), (7, 100), (12, 33)])
def test_matches_scipy_natural_spline(K, T):
    rng = np.random.default_rng(K * 100 + T)
    knots = rng.standard_normal((4, K, 7))
    t = np.linspace(0, 1, T)
    ref = CubicSpline(np.linspace(0, 1, K), knots, axis=1, bc_type="natural")
    kt = torch.tensor(knots)
    np.testing.assert_allclose(spline.create_spline_trajectory(kt, T).numpy(), ref(t), atol=1e-12)
    np.testing.assert_allclose(spline.spline_derivative(kt, T, 1).numpy(), ref(t, 1), atol=1e-10)
    np.testing.assert_allclose(spline.spline_derivative(kt, T, 2).numpy(), ref(t, 2), atol=1e-9)


def test_interpolates_knots_and_is_differentiable():
    knots = torch.randn(3, 5, 2, dtype=torch.float64, requires_grad=True)
    traj = spline.create_spline_trajectory(knots, 9)  # times 0, 1/8, ..., 1 contain the knots 0, 1/4, ..., 1
    assert torch.allclose(traj[:, ::2], knots)
    (g,) = torch.autograd.grad(traj.sum(), knots)
    s0, _, _ = spline.natural_spline_matrices(5, 9, dtype=torch.float64)
    assert torch.allclose(g, s0.sum(0)[None, :, None].expand_as(g))
    with pytest.raises(ValueError):
        spline.spline_derivative(knots, 9, 3)

"""Natural cubic spline through equally spaced knots, as a fixed linear map (the first
statement of the robot cost).

The reference builds its trajectories with ``torchcubicspline`` (absent here):
``create_spline_trajectory(knots, timesteps)`` = ``NaturalCubicSpline(natural_cubic_spline_coeffs(linspace(0, 1,
K), knots)).evaluate(linspace(0, 1, timesteps))`` (examples/script_planning_robot.py:118-127), and the
curvature term takes ``spline.derivative(t, order=1 | 2)`` (:321-328).  The natural spline through K knots is
linear in the knot values, so value and derivatives at fixed times are three ``[timesteps, K]`` matrices,
formed once per (K, timesteps) in float64 on the host (a K x K tridiagonal solve) and applied on the device
as one small matrix product per call -- differentiable w.r.t. the knots by torch's own matmul."""
from __future__ import annotations

import functools

import numpy as np
import torch


@functools.lru_cache(maxsize=64)
def _matrices_np(n_knots: int, timesteps: int):
    K = int(n_knots)
    if K < 2 or timesteps < 1:
        raise ValueError("natural_spline: at least two knots and one evaluation time")
    h = 1.0 / (K - 1)
    # second derivatives m at the knots: m_0 = m_{K-1} = 0, h (m_{i-1} + 4 m_i + m_{i+1}) / 6 = (y_{i+1} - 2 y_i + y_{i-1}) / h
    A = np.zeros((K, K))
    B = np.zeros((K, K))
    A[0, 0] = A[-1, -1] = 1.0
    for i in range(1, K - 1):
        A[i, i - 1], A[i, i], A[i, i + 1] = h / 6, 4 * h / 6, h / 6
        B[i, i - 1], B[i, i], B[i, i + 1] = 1 / h, -2 / h, 1 / h
    Mm = np.linalg.solve(A, B)  # m = Mm @ y
    t = np.linspace(0.0, 1.0, timesteps)
    idx = np.minimum((t * (K - 1)).astype(int), K - 2)  # interval of each time; t = 1 belongs to the last one
    a = (t - idx * h)  # distance from the left knot
    b = h - a
    eye = np.eye(K)
    yl, yr, ml, mr = eye[idx], eye[idx + 1], Mm[idx], Mm[idx + 1]
    a_, b_ = a[:, None], b[:, None]
    s0 = ml * b_ ** 3 / (6 * h) + mr * a_ ** 3 / (6 * h) + (yl - ml * h * h / 6) * b_ / h + (yr - mr * h * h / 6) * a_ / h
    s1 = -ml * b_ ** 2 / (2 * h) + mr * a_ ** 2 / (2 * h) - (yl - ml * h * h / 6) / h + (yr - mr * h * h / 6) / h
    s2 = ml * b_ / h + mr * a_ / h
    return s0, s1, s2


def natural_spline_matrices(n_knots: int, timesteps: int, device=None, dtype=torch.float32):
    """(S0, S1, S2), each ``[timesteps, n_knots]``: value, first and second derivative at ``linspace(0, 1,
    timesteps)`` of the natural cubic spline through knots at ``linspace(0, 1, n_knots)``."""
    return tuple(torch.tensor(m, dtype=dtype, device=device) for m in _matrices_np(int(n_knots), int(timesteps)))


def create_spline_trajectory(knots: torch.Tensor, timesteps: int = 100) -> torch.Tensor:
    """``[..., K, d] -> [..., timesteps, d]`` (the reference function of the same name)."""
    s0, _, _ = natural_spline_matrices(knots.shape[-2], timesteps, knots.device, knots.dtype)
    return s0 @ knots


def spline_derivative(knots: torch.Tensor, timesteps: int, order: int = 1) -> torch.Tensor:
    """``NaturalCubicSpline.derivative(linspace(0, 1, timesteps), order)`` for order 1 or 2."""
    if order not in (1, 2):
        raise ValueError("spline_derivative: order 1 or 2")
    return natural_spline_matrices(knots.shape[-2], timesteps, knots.device, knots.dtype)[order] @ knots


__all__ = ["natural_spline_matrices", "create_spline_trajectory", "spline_derivative"]

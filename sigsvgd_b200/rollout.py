"""Stein-MPC rollout of the particle model on the device.

``particle_rollout`` is one CUDA launch for what ``DuSt._rollout`` (src/controllers/dust.py:293-300) does
with a Python loop over ``ParticleModel.step`` (src/models/particle.py:123-173), and one more for the
gradient w.r.t. the actions that ``TrajectorySVGD`` asks autograd for (src/inference/trajectory_svgd.py:
55-69).  ``rollout_with(model)`` builds the call from an object with the reference model's attributes, so a
controller that keeps the reference's ``_rollout`` signature can delegate to it::

    class FusedDuSt(DuSt):
        def _rollout(self, init_state, base_actions, model, params_dist):
            ...                                   # the reference's own parameter sampling, unchanged
            states = rollout_with(model)(init_state, actions, params_dict)

There is no CPU path."""
from __future__ import annotations

import math
from typing import Optional

import torch

from . import ops


class _ParticleRollout(torch.autograd.Function):
    @staticmethod
    def forward(ctx, init_state, actions, mass, cfg):
        dt, max_acc, max_speed, grid, inv_cell, off, can_crash = cfg
        if not actions.is_cuda:
            raise RuntimeError("particle_rollout: CUDA tensors only; sigsvgd_b200 has no CPU fallback")
        R, H, da = actions.shape
        if da != 2 or init_state.numel() != 4:
            raise ValueError("particle_rollout: actions [R, H, 2] and a 4-dimensional initial state expected")
        dev = actions.device
        a = actions.detach().to(torch.float32).contiguous()
        s0 = init_state.detach().to(device=dev, dtype=torch.float32).contiguous()
        m_t, m_s = None, 1.0
        if isinstance(mass, torch.Tensor) and mass.numel() > 1:
            m_t = mass.detach().to(device=dev, dtype=torch.float32).reshape(-1).contiguous()
            if m_t.numel() != R:
                raise ValueError("particle_rollout: one mass per rollout expected")
        else:
            m_s = float(mass)
        states = torch.empty(R, H + 1, 4, dtype=torch.float32, device=dev)
        flags = torch.empty(R, H, dtype=torch.uint8, device=dev)
        gx, gy = (grid.shape if grid is not None else (0, 0))
        ops.call("sigsvgd_particle_rollout", ops._p(s0), ops._p(a), ops._p(m_t), m_s, R, H, float(dt), float(max_acc),
                 float(max_speed), ops._p(grid), gx, gy, float(inv_cell), float(off[0]), float(off[1]), int(bool(can_crash)),
                 ops._p(states), ops._p(flags), ops._stream())
        ctx.saved = (flags, m_t, m_s, float(dt), actions.dtype)
        return states.to(actions.dtype)

    @staticmethod
    def backward(ctx, gstates):
        flags, m_t, m_s, dt, dtype = ctx.saved
        R, H = flags.shape
        g = gstates.to(torch.float32).contiguous()
        ga = torch.empty(R, H, 2, dtype=torch.float32, device=g.device)
        ops.call("sigsvgd_particle_rollout_bwd", ops._p(m_t), m_s, R, H, dt, ops._p(g), ops._p(flags), ops._p(ga),
                 ops._stream())
        return None, ga.to(dtype), None, None


def particle_rollout(init_state: torch.Tensor, actions: torch.Tensor, mass=1.0, dt: float = 0.05,
                     max_accel: float = math.inf, max_speed: float = math.inf, grid: Optional[torch.Tensor] = None,
                     cell_size: float = 1.0, c_offset=(0, 0), can_crash: bool = False) -> torch.Tensor:
    """states [R, H+1, 4] of the particle model for actions [R, H, 2] (differentiable w.r.t. the actions).
    ``grid``: uint8 occupancy ``[gx, gy]`` on the actions' device (``ObstacleMap.map != 0``), indexed as
    ``floor(pos / cell_size + c_offset)``."""
    if grid is not None and (grid.dtype != torch.uint8 or not grid.is_contiguous() or grid.device != actions.device):
        raise ValueError("particle_rollout: grid must be a contiguous uint8 tensor on the actions' device")
    # (the reference multiplies by the Python float 1 / cell_size, rounded to float32 by torch)
    cfg = (dt, max_accel, max_speed, grid, 1.0 / cell_size, c_offset, can_crash and grid is not None)
    return _ParticleRollout.apply(init_state, actions, mass, cfg)


def rollout_with(model):
    """``f(init_state, actions [R,H,2], params_dict) -> states [R,H+1,4]`` for an object with the reference
    ``ParticleModel``'s attributes (``dt``, ``params_dict``, ``control_type``, ``deterministic``, ``can_crash``,
    ``with_obstacle``, ``obst_map``, the name-mangled speed / acceleration bounds).  The occupancy grid is
    uploaded once."""
    if getattr(model, "control_type", "acceleration") != "acceleration" or not getattr(model, "deterministic", True):
        raise NotImplementedError("rollout_with: acceleration control without control noise only")
    cache = {}

    def bound(name):
        v = getattr(model, f"_ParticleModel__{name}", None)
        return math.inf if v is None else float(v)

    def run(init_state, actions, params_dict=None):
        dev = actions.device
        grid, cell, off = None, 1.0, (0, 0)
        if getattr(model, "with_obstacle", False) and getattr(model, "can_crash", False):
            om = model.obst_map
            if dev not in cache:
                cache[dev] = (om.map != 0).to(device=dev, dtype=torch.uint8).contiguous()
            grid, cell, off = cache[dev], float(om.cell_size), tuple(float(v) for v in om.c_offset)
        mass = (params_dict or {}).get("mass", model.params_dict["mass"])
        return particle_rollout(init_state, actions, mass, model.dt, bound("max_acc"), bound("max_speed"), grid, cell,
                                off, grid is not None)

    return run


__all__ = ["particle_rollout", "rollout_with"]

"""Golden vectors of the Stein-MPC rollout, produced in THIS container by the UNMODIFIED
reference model:

    python -m oracle.make_golden_rollout        # from the repo root; needs /root/reference

``src.models.ParticleModel`` with ``examples/particle_maze_config.yaml`` (acceleration control, obstacle
grid ``sm_grid_4x4``, ``can_crash``, speed clamp 5) stepped exactly as ``DuSt._rollout`` does
(src/controllers/dust.py:293-300: ``states = cat([states, model.step(states[:, t], actions[:, t], params)])``),
with per-rollout masses; plus ``autograd.grad`` of a random linear functional of the states w.r.t. the
actions.  The occupancy grid itself is stored (bit-packed) so that the GPU box needs no reference code.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
OUT = os.path.join(ROOT, "tests", "golden", "rollout_maze.npz")


def main():
    sys.path.insert(0, ROOT)
    from oracle import make_golden_mpc

    torch.manual_seed(0)
    np.random.seed(0)
    _, simulator, _, state, _, exp_hp = make_golden_mpc.build_controller({"tau": [], "k_xx": [], "dk_dtau": []})
    torch.autograd.set_detect_anomaly(False)
    R, H = 350, exp_hp["horizon"]
    g = torch.Generator().manual_seed(5)
    out = {}
    for name, scale in (("free", 10.0), ("wild", 60.0)):  # "wild": crashes into obstacles and hits the speed clamp
        actions = (scale * torch.randn(R, H, 2, generator=g)).requires_grad_(True)
        mass = 2.0 + 0.1 * torch.randn(R, 1, generator=g)
        params = {"mass": mass}
        states = state.expand(R, 1, -1).clone()
        for t in range(H):
            states = torch.cat([states, simulator.step(states[:, t], actions[:, t], params).unsqueeze(1)], dim=1)
        w = torch.randn(R, H + 1, 4, generator=g)
        (ga,) = torch.autograd.grad((states * w).sum(), actions)
        out[f"{name}_actions"], out[f"{name}_mass"] = actions.detach().numpy(), mass.numpy()
        out[f"{name}_states"], out[f"{name}_w"], out[f"{name}_grad_actions"] = states.detach().numpy(), w.numpy(), ga.numpy()
    om = simulator.obst_map
    grid = om.map.numpy()
    assert set(np.unique(grid)) <= {0.0, 1.0}
    out["grid_bits"], out["grid_shape"] = np.packbits(grid.astype(np.uint8)), np.array(grid.shape)
    out["cell_size"], out["c_offset"] = np.float64(om.cell_size), om.c_offset.numpy()
    out["init_state"], out["dt"] = state.numpy(), np.float64(simulator.dt)
    out["max_speed"] = np.float64(5.0)
    np.savez_compressed(OUT, **out)
    print({k: np.shape(v) for k, v in out.items()}, "->", OUT, "crashed steps:",
          int(grid[np.clip(np.floor(out["wild_states"][..., 0] / om.cell_size + 200).astype(int), 0, 399),
                   np.clip(np.floor(out["wild_states"][..., 1] / om.cell_size + 200).astype(int), 0, 399)].sum()))


if __name__ == "__main__":
    main()

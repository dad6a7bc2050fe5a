"""numpy restatement (float32, the reference's operation order) of ``ParticleModel.step`` iterated as
``DuSt._rollout`` does (src/models/particle.py:123-173, src/controllers/dust.py:293-300,
src/utils/obstacle_map.py:68-97) and of its gradient w.r.t. the actions.  TEST INFRASTRUCTURE ONLY.
Pinned by tests/golden/rollout_maze.npz (the unmodified reference model, oracle/make_golden_rollout.py)."""
import numpy as np


def particle_rollout_np(init_state, actions, mass, dt, max_acc=np.inf, max_speed=np.inf, grid=None, cell_size=1.0,
                        c_offset=(0, 0), can_crash=False):
    a_all = np.asarray(actions, np.float32)
    R, H, _ = a_all.shape
    m = np.broadcast_to(np.asarray(mass, np.float32).reshape(-1, 1), (R, 1))
    s = np.broadcast_to(np.asarray(init_state, np.float32), (R, 4)).copy()
    states, flags = [s.copy()], np.zeros((R, H, 5), bool)  # crash, clamp vx, vy, ax, ay
    dt32, inv = np.float32(dt), np.float32(1.0 / cell_size)
    for t in range(H):
        acc = a_all[:, t] / m
        flags[:, t, 3:] = (acc < -max_acc) | (acc > max_acc)
        acc = np.clip(acc, -max_acc, max_acc).astype(np.float32)
        f = np.ones((R, 1), np.float32)
        if can_crash and grid is not None:
            idx = np.floor(s[:, :2] * inv + np.asarray(c_offset, np.float32)).astype(np.int64)
            ix, iy = np.clip(idx[:, 0], 0, grid.shape[0] - 1), np.clip(idx[:, 1], 0, grid.shape[1] - 1)
            hit = grid[ix, iy] != 0
            flags[:, t, 0] = hit
            f[hit] = 0.0
        xdot = np.concatenate([s[:, 2:], acc], 1)
        s = s + (xdot * dt32) * f
        flags[:, t, 1:3] = (s[:, 2:] < -max_speed) | (s[:, 2:] > max_speed)
        s[:, 2:] = np.clip(s[:, 2:], -max_speed, max_speed)
        states.append(s.copy())
    return np.stack(states, 1), flags


def particle_rollout_bwd_np(grad_states, flags, mass, dt):
    g = np.asarray(grad_states, np.float64)
    R, H = flags.shape[:2]
    m = np.broadcast_to(np.asarray(mass, np.float64).reshape(-1, 1), (R, 1))
    a = g[:, H].copy()
    ga = np.zeros((R, H, 2))
    for t in range(H - 1, -1, -1):
        k = np.where(flags[:, t, 0:1], 0.0, dt)
        gv = np.where(flags[:, t, 1:3], 0.0, a[:, 2:])
        ga[:, t] = np.where(flags[:, t, 3:5], 0.0, gv * k / m)
        a = np.concatenate([a[:, :2], a[:, :2] * k + gv], 1) + g[:, t]
    return ga

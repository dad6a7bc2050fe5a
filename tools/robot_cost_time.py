"""The reference's robot-planning cost and its gradient w.r.t. the knots, assembled twice on one GPU:

  torch arm   the statements of ``batch_cost_function`` (examples/script_planning_robot.py:248-356) with the
              forward kinematics as batched torch statements (what differentiable_robot_model runs per joint)
              and the two networks as ``nn.Sequential`` modules, differentiated by autograd (:107);
  fused arm   the SAME statements with ``exp_datapack.robot.qs_to_joints_xs``, ``exp_datapack.occmap`` and
              ``exp_datapack.self_collision_pred`` replaced by this package's drop-ins
              (``kinematics.SerialChain``, ``occupancy.fuse``): INTEGRATION.md route (F).

Both arms take the spline from ``sigsvgd_b200.spline`` (torchcubicspline is absent).  Random-initialised
networks of the reference's architecture (the trained weights are a download), the Panda chain of the
reference's URDF (tests/golden/panda_chain.npz).  MEASUREMENT / TEST INFRASTRUCTURE: restating the reference's
cost here is what lets the two arms be compared; the product does not contain it.
Usage: python tools/robot_cost_time.py [particles] [timesteps]"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from sigsvgd_b200 import kinematics, occupancy, spline  # noqa: E402
from sigsvgd_b200.kinematics import _rpy_matrix  # noqa: E402


def _net(in_dim, seed, dev):
    torch.manual_seed(seed)
    layers = [torch.nn.Linear(in_dim, 200), torch.nn.ReLU()]
    for _ in range(3):
        layers += [torch.nn.Linear(200, 200), torch.nn.ReLU()]
    layers += [torch.nn.Linear(200, 1), torch.nn.Sigmoid()]
    net = torch.nn.Sequential(*layers).to(dev)
    with torch.no_grad():
        net[-2].weight.mul_(30.0)
    return net


def _torch_fk(g, dev):
    xyz = [torch.tensor(v, dtype=torch.float32, device=dev) for v in g["xyz"]]
    rot = [torch.tensor(_rpy_matrix(*r), dtype=torch.float32, device=dev) for r in g["rpy"]]
    Ks = []
    for a in g["axis"]:
        a = a / max(np.linalg.norm(a), 1e-30)
        Ks.append(torch.tensor([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]], dtype=torch.float32, device=dev))
    eye = torch.eye(3, device=dev)

    def fk(q):  # [B, 7] -> [9, B, 3]
        B = q.shape[0]
        R, p, out = eye.expand(B, 3, 3), torch.zeros(B, 3, device=dev), []
        for j in range(len(xyz)):
            p = p + R @ xyz[j]
            R = R @ rot[j]
            out.append(p)
            if g["dof"][j] >= 0:
                ang = q[:, int(g["dof"][j])][:, None, None]
                R = R @ (eye + torch.sin(ang) * Ks[j] + (1 - torch.cos(ang)) * (Ks[j] @ Ks[j]))
        return torch.stack(out)

    return fk


def cost_fn(fk, occmap, self_collision_pred, timesteps):
    """script_planning_robot.py:248-356 with the default weights (all 1), ``use_ee_for_traj_dist=False``,
    ``optimise_ee_curvature=True``."""

    def cost(x, start_pose, target_pose):
        batch = x.shape[0]
        knots = torch.cat((start_pose.repeat(batch, 1, 1), x, target_pose.repeat(batch, 1, 1)), 1)  # :262-264
        qs = spline.create_spline_trajectory(knots, timesteps)  # :265
        shp = qs.shape
        xs = fk(qs.reshape(-1, shp[-1]))  # :269  [n_links, batch * timesteps, 3]
        ee_xs = xs[-1].reshape(shp[0], shp[1], -1)  # :271-272
        q_weights = torch.linspace(1.0, 0.7, 7, device=x.device)[None, None, :]  # :279
        qs_dist = torch.linalg.norm(q_weights * (qs[:, 1:, :] - qs[:, :-1, :]), dim=-1).sum(1)  # :281-283
        ee_dist = torch.linalg.norm(ee_xs[:, 1:, :] - ee_xs[:, :-1, :], dim=-1).sum(1)  # :284
        traj_dist = qs_dist + 1.0 * ee_dist  # :285
        n_pts = 10  # :300
        inter = torch.arange(0, 1, 1 / n_pts, device=x.device)  # create_body_points, :240-244
        body_points = (xs[:-1, None] + inter[:, None, None] * xs[1:, None]).flatten(0, 1)
        collision_prob = occmap(body_points).squeeze(-1)  # :302
        collision_prob = (collision_prob.sum(0) / (n_pts - 1)).reshape(shp[0], -1).sum(-1)  # :309-314
        self_collision_prob = self_collision_pred(qs).squeeze(-1).sum(1)  # :316-318
        t50 = 50  # :321
        d1, d2 = spline.spline_derivative(ee_xs, t50, 1), spline.spline_derivative(ee_xs, t50, 2)  # :323-325
        curv = (torch.cross(d1, d2, dim=2).norm(dim=2) / (d1.norm(dim=2) ** 3)).mean()  # :326-329
        return collision_prob + self_collision_prob + traj_dist + curv  # :334-339

    return cost


def timed(fn, n=10, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


def run(dev, particles=256, timesteps=200, knots=3, precise=True):
    g = np.load(os.path.join(ROOT, "tests", "golden", "panda_chain.npz"))
    lo = torch.tensor(g["limits_lo"], dtype=torch.float32, device=dev)
    hi = torch.tensor(g["limits_hi"], dtype=torch.float32, device=dev)
    occ, col = _net(3, 5, dev), _net(7, 6, dev)
    chain = kinematics.SerialChain(g["xyz"], g["rpy"], g["axis"], g["dof"], list(g["link_names"]))
    gen = torch.Generator(device=dev).manual_seed(0)
    x = (lo + torch.rand(particles, knots, 7, device=dev, generator=gen) * (hi - lo)) * 0.5  # script_planning_robot.py:380-383
    start, target = (0.3 * lo)[None, None, :], (0.3 * hi)[None, None, :]
    cost_torch = cost_fn(_torch_fk(g, dev), occ, col, timesteps)
    cost_fused = cost_fn(chain, occupancy.fuse(occ, precise=precise), occupancy.fuse(col, precise=precise), timesteps)

    def grad_of(cost):
        def f():
            xx = x.detach().requires_grad_(True)
            c = cost(xx, start, target)
            return c.detach(), torch.autograd.grad(-c.sum(), xx)[0]  # :107 (the likelihood is exp(-cost))
        return f

    ct, gt = grad_of(cost_torch)()
    cf, gf = grad_of(cost_fused)()
    return {"particles": particles, "knots": knots, "timesteps": timesteps, "configurations": particles * timesteps,
            "occupancy_points": particles * timesteps * 80, "mode": "precise" if precise else "fast",
            "ms_torch_statements": timed(grad_of(cost_torch), 4, 1), "ms_fused_drop_ins": timed(grad_of(cost_fused)),
            "max_rel_dev_cost": float(((cf - ct).abs() / ct.abs()).max()),
            "max_dev_grad_rel_to_max_grad": float((gf - gt).abs().max() / gt.abs().max()),
            "median_dev_grad_rel_to_max_grad": float((gf - gt).abs().median() / gt.abs().max())}


def run_cpu_sample(particles=32, timesteps=200, knots=3, repeats=2):
    """The torch arm on the host cores (the reference's own CPU path for the cost) on a bounded sample of the
    workload: ``particles`` of the 256, same spline steps and body points."""
    import time

    dev = torch.device("cpu")
    g = np.load(os.path.join(ROOT, "tests", "golden", "panda_chain.npz"))
    lo = torch.tensor(g["limits_lo"], dtype=torch.float32)
    hi = torch.tensor(g["limits_hi"], dtype=torch.float32)
    occ, col = _net(3, 5, dev), _net(7, 6, dev)
    gen = torch.Generator().manual_seed(0)
    x = (lo + torch.rand(particles, knots, 7, generator=gen) * (hi - lo)) * 0.5
    start, target = (0.3 * lo)[None, None, :], (0.3 * hi)[None, None, :]
    cost = cost_fn(_torch_fk(g, dev), occ, col, timesteps)

    def f():
        xx = x.detach().requires_grad_(True)
        c = cost(xx, start, target)
        return torch.autograd.grad(-c.sum(), xx)[0]

    f()
    t0 = time.perf_counter()
    for _ in range(repeats):
        f()
    ms = (time.perf_counter() - t0) / repeats * 1e3
    return {"particles": particles, "ms": ms, "ms_scaled_to_256_particles": ms * 256 / particles,
            "cores": torch.get_num_threads(), "kind": "the reference's torch statements on the host cores"}


if __name__ == "__main__":
    P = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    T = int(sys.argv[2]) if len(sys.argv) > 2 else 200
    for precise in (True, False):
        out = run(torch.device("cuda:0"), P, T, precise=precise)
        out["speedup"] = out["ms_torch_statements"] / out["ms_fused_drop_ins"]
        print(json.dumps(out))
    print(json.dumps({"cpu_sample": run_cpu_sample()}))

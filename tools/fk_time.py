"""Timing of the serial-chain forward kinematics against the same recursion as batched torch
statements + autograd on the same GPU (what differentiable_robot_model does per joint).
Usage: python tools/fk_time.py [B]"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from sigsvgd_b200 import kinematics  # noqa: E402
from sigsvgd_b200.kinematics import _rpy_matrix  # noqa: E402


def timed(fn, n=20, warm=3):
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


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 256 * 200
    dev = torch.device("cuda:0")
    g = np.load(os.path.join(ROOT, "tests", "golden", "panda_chain.npz"))
    sc = kinematics.SerialChain(g["xyz"], g["rpy"], g["axis"], g["dof"], list(g["link_names"]))
    lo, hi = torch.tensor(g["limits_lo"], dtype=torch.float32, device=dev), torch.tensor(g["limits_hi"], dtype=torch.float32, device=dev)
    q = lo + torch.rand(B, 7, device=dev) * (hi - lo)
    w = torch.randn(9, B, 3, device=dev)
    xyz = [torch.tensor(v, dtype=torch.float32, device=dev) for v in g["xyz"]]
    rot = [torch.tensor(_rpy_matrix(*r), dtype=torch.float32, device=dev) for r in g["rpy"]]
    Ks = []
    for a in g["axis"]:
        a = a / max(np.linalg.norm(a), 1e-30)
        Ks.append(torch.tensor([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]], dtype=torch.float32, device=dev))
    eye = torch.eye(3, device=dev)

    def torch_fk(qq):
        R = eye.expand(B, 3, 3)
        p = torch.zeros(B, 3, device=dev)
        out = []
        for j in range(len(xyz)):
            p = p + R @ xyz[j]
            R = R @ rot[j]
            out.append(p)
            if g["dof"][j] >= 0:
                ang = qq[:, int(g["dof"][j])][:, None, None]
                R = R @ (eye + torch.sin(ang) * Ks[j] + (1 - torch.cos(ang)) * (Ks[j] @ Ks[j]))
        return torch.stack(out)

    def run_torch():
        qq = q.detach().requires_grad_(True)
        xs = torch_fk(qq)
        return xs, torch.autograd.grad((xs * w).sum(), qq)[0]

    def run_fused():
        qq = q.detach().requires_grad_(True)
        xs = sc(qq)
        return xs, torch.autograd.grad((xs * w).sum(), qq)[0]

    xt, gt = run_torch()
    xf, gf = run_fused()
    print(json.dumps({"configurations": B, "links": 9, "ms_kernels_fwd_bwd": timed(run_fused), "ms_torch_fwd_autograd": timed(run_torch, 5, 2),
                      "dev_positions": float((xf - xt).abs().max()), "dev_grad_rel": float((gf - gt).abs().max() / gt.abs().max())}))


if __name__ == "__main__":
    main()

"""Golden vectors of the occupancy / self-collision networks, produced in THIS container by
the UNMODIFIED reference classes:

    python -m oracle.make_golden_mlp        # from the repo root; needs /root/reference

``ContinuousOccupancyMap()`` (3 -> 200 x 4 -> 1) and ``ContinuousSelfCollisionPredictor(n_dimension=7)`` with
torch's default initialisation under ``torch.manual_seed`` (the trained weights are a download), evaluated in
float32 on CPU at random points; plus ``autograd.grad(out.sum(), points)``.  ``pytorch_lightning`` (imported by
both files for their trainers, absent here) is replaced by a permissive stand-in: the network classes are plain
``nn.Module``s and do not touch it."""
from __future__ import annotations

import importlib
import importlib.util
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
OUT = os.path.join(ROOT, "tests", "golden", "occupancy_mlp.npz")


def reference_networks():
    sys.path.insert(0, ROOT)
    from oracle import ref_shim

    ref_shim.install_stubs()
    # (the predictor's file also imports PandaRobot / get_project_root for its training script: stand-ins, this
    #  process imports nothing else from the reference)
    for name in ("pytorch_lightning", "pytorch_lightning.callbacks", "pytorch_lightning.callbacks.early_stopping",
                 "src", "src.models", "src.models.robot", "src.models.robot.robot_simulator", "src.utils",
                 "src.utils.helper"):
        if name not in sys.modules:
            sys.modules[name] = ref_shim._Permissive(name)
    # the two files are loaded by path: ``import src.models`` pulls in pybullet (robot simulator), absent here
    mods = []
    for stem in ("continuous_occupancy_map", "continuous_self_collision_pred"):
        path = os.path.join(ref_shim.REFERENCE_ROOT, "src", "models", "robot_learning", stem + ".py")
        spec = importlib.util.spec_from_file_location("_ref_" + stem, path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        mods.append(mod)
    return mods[0].ContinuousOccupancyMap, mods[1].ContinuousSelfCollisionPredictor


def main():
    Occ, Col = reference_networks()
    out = {}
    g = torch.Generator().manual_seed(11)
    for name, make, dim, scale in (("occ", lambda: Occ(), 3, 1.0), ("col", lambda: Col(n_dimension=7), 7, 2.0)):
        torch.manual_seed(3 if name == "occ" else 4)
        net = make()
        # spread the logits (default init gives |logit| << 1): scale the last layer, keep everything else as drawn
        with torch.no_grad():
            net.layers[-2].weight.mul_(40.0)
        x = (scale * torch.randn(700, dim, generator=g)).requires_grad_(True)
        y = net(x)
        (gx,) = torch.autograd.grad(y.sum(), x)
        lin = [m for m in net.layers if isinstance(m, torch.nn.Linear)]
        for i, m in enumerate(lin):
            out[f"{name}_w{i}"], out[f"{name}_b{i}"] = m.weight.detach().numpy(), m.bias.detach().numpy()
        out[f"{name}_x"], out[f"{name}_y"], out[f"{name}_grad"] = x.detach().numpy(), y.detach().numpy(), gx.numpy()
    np.savez_compressed(OUT, **out)
    print({k: np.shape(v) for k, v in out.items()}, "->", OUT)


if __name__ == "__main__":
    main()

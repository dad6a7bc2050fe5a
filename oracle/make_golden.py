"""Generate tests/golden/*.npz by running the UNMODIFIED reference Python in this container.

    python -m oracle.make_golden            # from the repo root; needs /root/reference

What is real reference code here: ``src.kernels.PathSigKernel``, ``src.kernels.GaussianKernel``,
``src.utils.math.pw_dist_sq`` / ``bw_median``, ``src.inference.SVGD._compute_kernel`` /
``_velocity`` / ``optimize``, ``src.inference.ScoreEstimator``.  What is NOT: ``signatory``
(absent, third party) -- ``signatory.signature`` is the oracle restatement
(oracle/sig_oracle.c), so these vectors pin everything downstream of the signature, and the
signature only through the closed-form KATs in tests/test_oracle_signature.py.

Each case stores float32 inputs, the reference's float32 outputs and the reference's float64
outputs on the same inputs (the "truth" both the fp32 reference and the CUDA path are scored
against).
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
OUT = os.path.join(ROOT, "tests", "golden")


def _paths(kind, N, T, d, gen):
    if kind == "walk":  # SURVEY.md 8(d) config 4 generator
        return torch.cumsum(0.05 * torch.randn(N, T, d, generator=gen), dim=1)
    if kind == "field":  # config 1: straight line start->target + 0.4 randn
        start, target = torch.full((d,), -1.0), torch.full((d,), 1.0)
        w = torch.linspace(0, 1, T)[:, None]
        return (start + w * (target - start))[None] + 0.4 * torch.randn(N, T, d, generator=gen)
    if kind == "panda":  # config 2: uniform in symmetric joint limits (tests/test_smoothbox.py:5-6)
        b = torch.tensor([2.9671, 1.8326, 2.9671, 0.0873, 2.9671, 3.8223, 2.9671])[:d]
        u = torch.rand(N, T, d, generator=gen)
        q0, q1 = (torch.rand(2, N, 1, d, generator=gen) * 2 - 1) * b
        w = torch.linspace(0, 1, T)[None, :, None]
        return 0.8 * (q0 + w * (q1 - q0)) + 0.2 * (2 * u - 1) * b
    raise ValueError(kind)


CASES = [
    # name, kind, N, T, d, depth
    ("field_100x10x2_m3", "field", 100, 10, 2, 3),
    ("walk_64x16x7_m3", "walk", 64, 16, 7, 3),
    ("maze_35x30x2_m3", "walk", 35, 30, 2, 3),
    ("panda_48x12x7_m3", "panda", 48, 12, 7, 3),
    ("walk_40x9x3_m4", "walk", 40, 9, 3, 4),
    ("walk_33x8x5_m2", "walk", 33, 8, 5, 2),
]


def main():
    sys.path.insert(0, ROOT)
    from oracle import ref_shim

    ref_shim.import_reference()
    from src.kernels import PathSigKernel, GaussianKernel
    from src.inference import SVGD, ScoreEstimator
    from src.utils.math import pw_dist_sq, bw_median
    import signatory  # the stub -> oracle restatement

    os.makedirs(OUT, exist_ok=True)
    for name, kind, N, T, d, depth in CASES:
        gen = torch.Generator().manual_seed(0)
        X = _paths(kind, N, T, d, gen).float().contiguous()
        G = torch.randn(N, T * d, generator=gen).float()
        out = {"X": X.numpy(), "G": G.numpy(), "depth": np.int64(depth)}
        for tag, dt in (("f32", torch.float32), ("f64", torch.float64)):
            Xd, Gd = X.to(dt), G.to(dt)
            kernel = PathSigKernel()
            # the reference hard-codes depth=3 as a default argument; other depths are reached
            # through kwargs of the kernel call -> wrap so SVGD passes it along
            call = (lambda k, dep: (lambda A, B, **kw: k(A, B, depth=dep, **kw)))(kernel, depth)
            call.analytic_grad = False
            sig = signatory.signature(Xd, depth, basepoint=True)
            sq = pw_dist_sq(sig, sig)
            h = bw_median(sq)
            sampler = SVGD(kernel=call, optimizer_class=None, lr=0.1)
            k_xx, grad_k = sampler._compute_kernel(Xd)
            vel, _ = sampler._velocity(Xd, Gd)
            K2, dK_sig = kernel(Xd, Xd.clone(), depth=depth, compute_grad=True)
            # fixed bandwidth (examples: bandwidth_fn=lambda _: const)
            kfix = PathSigKernel(static_kernel=GaussianKernel(bandwidth_fn=lambda _: 1.5))
            cfix = (lambda k, dep: (lambda A, B, **kw: k(A, B, depth=dep, **kw)))(kfix, depth)
            cfix.analytic_grad = False
            sfix = SVGD(kernel=cfix, optimizer_class=None, lr=0.1)
            k_fix, gk_fix = sfix._compute_kernel(Xd)
            v_fix, _ = sfix._velocity(Xd, Gd)
            assert torch.equal(K2, k_xx)
            for key, val in dict(sig=sig, sq=sq, h=h, k_xx=k_xx, grad_k=grad_k, velocity=vel, dK_sig=dK_sig,
                                 k_fix=k_fix, gradk_fix=gk_fix, vel_fix=v_fix).items():
                out[f"{key}_{tag}"] = val.detach().numpy()
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
        print(name, "h32", float(out["h_f32"]), "h64", float(out["h_f64"]),
              "rel err fp32 ref vs fp64 ref (velocity):",
              float(np.abs(out["velocity_f32"] - out["velocity_f64"]).max() / np.abs(out["velocity_f64"]).max()))

    # ------------------------------------------------------------------ static RBF kernel alone
    gen = torch.Generator().manual_seed(1)
    A = torch.randn(20, 5, generator=gen)
    B = torch.randn(20, 5, generator=gen)
    out = {"A": A.numpy(), "B": B.numpy()}
    for tag, dt in (("f32", torch.float32), ("f64", torch.float64)):
        K, dK = GaussianKernel()(A.to(dt), B.to(dt))
        Kh, dKh = GaussianKernel()(A.to(dt), B.to(dt), h=0.7)
        out.update({f"K_{tag}": K.numpy(), f"dK_{tag}": dK.numpy(), f"Kh_{tag}": Kh.numpy(), f"dKh_{tag}": dKh.numpy()})
    np.savez_compressed(os.path.join(OUT, "gaussian_20x5.npz"), **out)

    # ------------------------------------------------------------------ optimize() trace
    # ScoreEstimator routes PathSigKernel to _svgd_ag_score (score.py:19-25); SVGD.optimize with
    # optimizer_class=None is plain gradient descent X <- X - lr * velocity (svgd.py:115).
    gen = torch.Generator().manual_seed(2)
    N, T, d = 24, 8, 2
    X0 = _paths("field", N, T, d, gen).float()
    centers = torch.randn(3, d, generator=gen)

    def cost_fn(x, w_len=1.0):  # spline-free cost in the style of script_planning_obstacle_field.py:120-126
        bumps = torch.exp(-((x[:, :, None, :] - centers) ** 2).sum(-1) / 0.5).sum((1, 2))
        length = ((x[:, 1:] - x[:, :-1]) ** 2).sum((1, 2))
        return bumps + w_len * length, {}

    out = {"X0": X0.numpy(), "centers": centers.numpy()}
    for tag, dt in (("f32", torch.float32), ("f64", torch.float64)):
        kernel = PathSigKernel()
        est = ScoreEstimator(kernel, cost_fn, {"w_len": 0.5})
        sampler = SVGD(kernel=kernel, optimizer_class=None, lr=0.05)
        P = X0.to(dt).clone()
        data, _ = sampler.optimize(P, est.score, n_steps=5)
        out[f"trace_{tag}"] = data["trace"].numpy()
        out[f"k_xx_last_{tag}"] = data[4]["k_xx"].numpy()
        out[f"grad_k_last_{tag}"] = data[4]["grad_k"].numpy()
    np.savez_compressed(os.path.join(OUT, "optimize_24x8x2.npz"), **out)
    print("written to", OUT, sorted(os.listdir(OUT)))


if __name__ == "__main__":
    main()

"""Golden vectors for the signature-PDE leg, produced in THIS container:

    python -m oracle.make_golden_sigpde        # from the repo root; needs /root/reference

What is real reference code here: ``src.kernels._traj_kernels.BatchGaussianKernel.Gram_matrix`` (the static
kernel the reference hands to ``sigkernel.SigKernel``, incl. its default ``bw_median`` bandwidth over the
[A,B,M,N] tensor) run UNMODIFIED in float64.  What is NOT: the Goursat solver (``sigkernel``, third party,
absent) -- ``k_from_G`` is the oracle's restatement of the package's recursion applied to the reference's
own ``Gram_matrix`` output, so these vectors pin the static kernel and the bandwidth convention; the
solver is pinned by the closed forms of tests/test_oracle_sigpde.py.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
OUT = os.path.join(ROOT, "tests", "golden", "sigpde_static.npz")

CASES = [("maze", 35, 35, 30, 30, 2, 0), ("small", 4, 3, 5, 6, 2, 2), ("panda", 6, 5, 16, 12, 7, 1)]


def main():
    sys.path.insert(0, ROOT)
    from oracle import ref_shim
    from oracle import sigpde_oracle as so

    ref_shim.import_reference()
    from src.kernels._traj_kernels import BatchGaussianKernel

    out = {}
    for name, A, B, M, N, D, order in CASES:
        g = torch.Generator().manual_seed(len(name) + A)
        X = torch.cumsum(0.2 * torch.randn(A, M, D, generator=g, dtype=torch.float64), 1)
        Y = torch.cumsum(0.2 * torch.randn(B, N, D, generator=g, dtype=torch.float64), 1)
        G_med = BatchGaussianKernel().Gram_matrix(X, Y)                               # median heuristic
        G_fix = BatchGaussianKernel(bandwidth_fn=lambda _: 0.75).Gram_matrix(X, Y)    # constant bandwidth
        out[f"{name}_X"], out[f"{name}_Y"], out[f"{name}_order"] = X.numpy(), Y.numpy(), np.int64(order)
        for tag, G in (("med", G_med), ("fix", G_fix)):
            out[f"{name}_k_{tag}"] = so._solve(so.increments(G.numpy()), order)[..., -1, -1]
        # the bandwidth the reference used, recovered from its own output at one entry
        d00 = float(((X[0, 1] - Y[0, 2]) ** 2).sum())
        out[f"{name}_h_med"] = np.float64(-d00 / np.log(float(G_med[0, 0, 1, 2])))
    np.savez_compressed(OUT, **out)
    print({k: np.shape(v) for k, v in out.items()}, "->", OUT)


if __name__ == "__main__":
    main()

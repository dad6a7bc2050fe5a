"""Stand-ins for the two third-party modules the reference imports on this path, backed by the
CUDA library, so the reference's UNMODIFIED files run on the new kernels:

* ``signatory``  (pinned signatory==1.2.6.1.9.0, setup.py:58): ``signature(path, depth,
  basepoint=...)`` -> ``ops.signature`` (call sites src/kernels/_traj_kernels.py:125-126);
* ``sigkernel``  (setup.py:71): ``SigKernel(static_kernel, dyadic_order).compute_Gram(X, Y,
  sym=False)`` and ``RBFKernel(sigma)``.  ``src/inference/trajectory_svgd.py:9`` imports
  ``SigKernel`` and its branch at :55-69 is the only one through which the unmodified Stein-MPC
  controller (src/controllers/dust.py) reaches a signature kernel.  Here ``compute_Gram`` returns
  the north-star kernel -- truncated signature + RBF with the median heuristic -- as an autograd
  node; the Goursat-PDE kernel of the real package is SURVEY.md 8(f)-1 ("next").

``install()`` registers both in ``sys.modules`` (only if the real packages are absent, unless
``force=True``).
"""
from __future__ import annotations

import sys
import types

import torch

from . import ops


class RBFKernel:
    """``sigkernel.RBFKernel(sigma)``; sigma is kept for API compatibility.  The truncated-signature
    adapter uses the median heuristic on signature features unless ``fixed=True``."""

    def __init__(self, sigma: float = 1.0, fixed: bool = False):
        self.sigma = float(sigma)
        self.fixed = bool(fixed)


class SigKernel:
    """``sigkernel.SigKernel(static_kernel, dyadic_order)`` shaped adapter."""

    def __init__(self, static_kernel=None, dyadic_order: int = 0, depth: int = 3):
        self.static_kernel = static_kernel
        self.dyadic_order = dyadic_order
        self.depth = int(depth)

    def _bandwidth(self):
        sk = self.static_kernel
        if isinstance(sk, RBFKernel) and sk.fixed:
            return ops.Bandwidth.fixed(sk.sigma)
        return ops.Bandwidth.median()

    def compute_Gram(self, X: torch.Tensor, Y: torch.Tensor, sym: bool = False) -> torch.Tensor:
        """[A, M, d] x [A, M, d] -> [A, A]; differentiable wrt X (and Y)."""
        return ops.sig_rbf_gram(X, Y, self.depth, self._bandwidth())

    def compute_kernel(self, X: torch.Tensor, Y: torch.Tensor) -> torch.Tensor:
        """k(X_i, Y_i) for paired batches -> [A] (diagonal of the Gram matrix)."""
        return self.compute_Gram(X, Y).diagonal()


def signature(path, depth, stream=False, basepoint=False, inverse=False, initial=None, scalar_term=False):
    return ops.signature(path, depth, stream=stream, basepoint=basepoint, inverse=inverse, initial=initial,
                         scalar_term=scalar_term)


def signature_channels(channels: int, depth: int, scalar_term: bool = False) -> int:
    return ops.sig_channels(channels, depth) + (1 if scalar_term else 0)


def install(force: bool = False) -> dict:
    """Register ``signatory`` / ``sigkernel`` stand-ins.  Returns {name: module} of what was set."""
    done = {}
    for name, attrs in (
        ("signatory", dict(signature=signature, signature_channels=signature_channels, __version__="sigsvgd_b200")),
        ("sigkernel", dict(SigKernel=SigKernel, RBFKernel=RBFKernel, __version__="sigsvgd_b200")),
    ):
        if not force:
            try:
                __import__(name)
                if not getattr(sys.modules[name], "__sigsvgd_b200__", False):
                    continue  # the real package is present: leave it alone
            except ImportError:
                pass
        mod = types.ModuleType(name)
        mod.__dict__.update(attrs)
        mod.__sigsvgd_b200__ = True
        sys.modules[name] = mod
        done[name] = mod
    return done


__all__ = ["SigKernel", "RBFKernel", "signature", "signature_channels", "install"]

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

``install()`` registers ``signatory`` in ``sys.modules`` (only if the real package is absent, unless
``force=True``).  The ``sigkernel`` stand-in computes a DIFFERENT kernel from the package whose name it
takes (truncated signature + RBF, not the Goursat-PDE signature kernel), so it is registered only on
request -- ``install(sigkernel_truncated=True)`` -- and says so with a warning.
"""
from __future__ import annotations

import sys
import types
import warnings

import torch

from . import ops


_SUBSTITUTION = ("sigsvgd_b200.compat.SigKernel is NOT the Goursat-PDE signature kernel of the `sigkernel` package: "
                 "compute_Gram returns exp(-|S(x) - S(y)|^2 / (2 h^2)) on depth-%d truncated signatures, h from the median "
                 "heuristic (RBFKernel(sigma, fixed=True): h = sigma).  `dyadic_order` has no meaning here.")


class RBFKernel:
    """``sigkernel.RBFKernel(sigma)`` shaped.  In the real package sigma is the width of a static kernel
    between path POINTS; the truncated-signature adapter has no such kernel, so sigma is only used
    when the caller opts in with ``fixed=True`` (then it is the bandwidth h on signature features);
    otherwise the bandwidth is the median heuristic, as in the reference's ``PathSigKernel``."""

    def __init__(self, sigma: float = 1.0, fixed: bool = False):
        self.sigma = float(sigma)
        self.fixed = bool(fixed)


class SigKernel:
    """``sigkernel.SigKernel(static_kernel, dyadic_order)`` shaped adapter."""

    _warned = False

    def __init__(self, static_kernel=None, dyadic_order: int = 0, depth: int = 3):
        self.static_kernel = static_kernel
        self.dyadic_order = dyadic_order
        self.depth = int(depth)
        if not SigKernel._warned:
            SigKernel._warned = True
            warnings.warn(_SUBSTITUTION % self.depth, stacklevel=2)

    def _bandwidth(self):
        sk = self.static_kernel
        if isinstance(sk, RBFKernel) and sk.fixed:
            return ops.Bandwidth.fixed(sk.sigma)
        return ops.Bandwidth.median()

    def compute_Gram(self, X: torch.Tensor, Y: torch.Tensor, sym: bool = False) -> torch.Tensor:
        """[A, M, d] x [A, M, d] -> [A, A]; differentiable wrt X (and Y).

        The reference's only call site passes ``tau.double()`` and ``tau.detach().double()``
        (src/inference/trajectory_svgd.py:60-62): two tensors with equal VALUES.  For small swarms
        (the MPC regime, launch-bound) that is detected -- one tiny comparison -- and the symmetric
        single-call path is taken instead of the two-sample one."""
        if not Y.requires_grad and X.shape == Y.shape and X.data_ptr() != Y.data_ptr() and X.is_cuda:
            if self.same_inputs or (X.numel() <= self._equal_check_max and bool(torch.equal(X.detach(), Y))):
                Y = X.detach()
        return ops.sig_rbf_gram(X, Y, self.depth, self._bandwidth())

    _equal_check_max = 1 << 16
    #: set to True to promise that ``compute_Gram(X, Y)`` is only ever called with ``Y`` holding the
    #: values of ``X`` (the reference's call site): skips the comparison and its host synchronisation,
    #: which also makes a whole MPC step capturable into a CUDA graph
    same_inputs = False

    def compute_kernel(self, X: torch.Tensor, Y: torch.Tensor) -> torch.Tensor:
        """k(X_i, Y_i) for paired batches -> [A] (diagonal of the Gram matrix)."""
        return self.compute_Gram(X, Y).diagonal()


class LinearKernel:
    """``sigkernel.LinearKernel`` shaped: static kernel ``<x_s, y_t> * scale``."""

    def __init__(self, scale: float = 1.0):
        self.scale = float(scale)


class PdeSigKernel:
    """``sigkernel.SigKernel(static_kernel, dyadic_order)`` computing the package's OWN kernel: the
    Goursat-PDE signature kernel (``ops.sigpde_gram``), in float64 on the GPU.
    ``static_kernel``: ``RBFKernel(sigma)`` (``exp(-|x - y|^2 / sigma)``, the package's convention),
    ``LinearKernel(scale)``, or the reference's ``BatchGaussianKernel`` (``sigsvgd_b200.kernels``)."""

    def __init__(self, static_kernel=None, dyadic_order: int = 0, _naive_solver: bool = False):
        if _naive_solver:
            raise NotImplementedError("PdeSigKernel: only the package's default (second-order) solver is built")
        self.static_kernel = static_kernel if static_kernel is not None else RBFKernel(1.0)
        self.dyadic_order = int(dyadic_order)

    def _spec(self):
        from . import _lib

        sk = self.static_kernel
        if isinstance(sk, LinearKernel):
            return _lib.PDE_LINEAR, sk.scale
        if isinstance(sk, RBFKernel):
            return _lib.PDE_RBF, sk.sigma
        if hasattr(sk, "bandwidth_spec"):
            return _lib.PDE_RBF, sk.bandwidth_spec()
        raise TypeError(f"PdeSigKernel: unsupported static kernel {type(sk).__name__}")

    def compute_Gram(self, X: torch.Tensor, Y: torch.Tensor, sym: bool = False) -> torch.Tensor:
        """[A,M,D] x [B,N,D] -> [A,B] in X's dtype; gradient w.r.t. X (trajectory_svgd.py:55-69)."""
        kind, h = self._spec()
        return ops.sigpde_gram(X, Y, self.dyadic_order, kind, h)

    def compute_kernel(self, X: torch.Tensor, Y: torch.Tensor) -> torch.Tensor:
        return self.compute_Gram(X, Y).diagonal()


def signature(path, depth, stream=False, basepoint=False, inverse=False, initial=None, scalar_term=False):
    return ops.signature(path, depth, stream=stream, basepoint=basepoint, inverse=inverse, initial=initial,
                         scalar_term=scalar_term)


def signature_channels(channels: int, depth: int, scalar_term: bool = False) -> int:
    return ops.sig_channels(channels, depth) + (1 if scalar_term else 0)


def install(force: bool = False, sigkernel_truncated: bool = False, sigkernel_pde: bool = False) -> dict:
    """Register the ``signatory`` stand-in and, on request, a module named ``sigkernel``: with
    ``sigkernel_pde=True`` its ``SigKernel`` is ``PdeSigKernel`` (the package's own Goursat-PDE kernel on the
    CUDA library); with ``sigkernel_truncated=True`` it is the truncated-signature RBF kernel of the north
    star (see the module docstring; warns).  Returns {name: module} of what was set."""
    done = {}
    wanted = [("signatory", dict(signature=signature, signature_channels=signature_channels, __version__="sigsvgd_b200"))]
    if sigkernel_pde:
        wanted.append(("sigkernel", dict(SigKernel=PdeSigKernel, RBFKernel=RBFKernel, LinearKernel=LinearKernel,
                                         __version__="sigsvgd_b200-pde")))
    elif sigkernel_truncated:
        warnings.warn("registering a module named `sigkernel`: " + _SUBSTITUTION % 3, stacklevel=2)
        wanted.append(("sigkernel", dict(SigKernel=SigKernel, RBFKernel=RBFKernel, __version__="sigsvgd_b200-truncated",
                                         __doc__=_SUBSTITUTION % 3)))
    for name, attrs in wanted:
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


__all__ = ["SigKernel", "PdeSigKernel", "RBFKernel", "LinearKernel", "signature", "signature_channels", "install"]

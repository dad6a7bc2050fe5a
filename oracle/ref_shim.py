"""Import the UNMODIFIED reference Python (``/root/reference/src``) in the build container.

TEST INFRASTRUCTURE, container-only: ``/root/reference`` does not exist on the GPU box, so
nothing in ``-m gpu`` tests, ``smoke()`` or ``bench.py`` may import this module.  It is used by
``oracle/make_golden.py`` (golden-vector generation) and by the ``not gpu`` tests that are
skipped when the reference tree is absent.

Recipe: register stand-ins for the third-party modules the reference
imports at module-import time but that are absent here, then ``import src...``:

* ``signatory``  -> ``signature`` = the oracle restatement (or any callable passed in);
* ``sigkernel``  -> classes ``SigKernel`` / ``RBFKernel`` (used only in isinstance checks unless
  the caller passes real implementations);
* ``matplotlib*``, ``gpytorch.priors`` -> permissive dummies (never called on the hot path).
"""
from __future__ import annotations

import importlib
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("SIGSVGD_REFERENCE_ROOT", "/root/reference")


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "src", "kernels"))


class _Permissive(types.ModuleType):
    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        obj = type(name, (), {"__init__": lambda self, *a, **k: None, "__call__": lambda self, *a, **k: None})
        setattr(self, name, obj)
        return obj


def install_stubs(signature_fn=None, sigkernel_module=None):
    """Idempotent.  Returns the dict of modules that were registered."""
    if signature_fn is None:
        from .signature import signature_torch as signature_fn
    reg = {}
    sig = types.ModuleType("signatory")
    sig.signature = signature_fn
    sig.__version__ = "1.2.6-oracle-stub"
    reg["signatory"] = sig
    if sigkernel_module is None:
        sk = types.ModuleType("sigkernel")

        class SigKernel:  # noqa: D401 - stand-in used for isinstance checks only
            def __init__(self, static_kernel=None, dyadic_order=0):
                self.static_kernel, self.dyadic_order = static_kernel, dyadic_order

            def compute_Gram(self, X, Y, sym=False):
                raise NotImplementedError("sigkernel PDE solver is absent (SURVEY.md 8(f)-1)")

        class RBFKernel:
            def __init__(self, sigma=1.0):
                self.sigma = sigma

        sk.SigKernel, sk.RBFKernel = SigKernel, RBFKernel
        sigkernel_module = sk
    reg["sigkernel"] = sigkernel_module
    for name in ("matplotlib", "matplotlib.pyplot", "matplotlib.cm", "matplotlib.animation", "matplotlib.patches",
                 "matplotlib.colors", "matplotlib.collections", "gpytorch", "gpytorch.priors", "seaborn", "altair"):
        if name not in sys.modules or isinstance(sys.modules[name], _Permissive):
            try:
                if name.split(".")[0] in ("matplotlib", "gpytorch", "seaborn", "altair"):
                    importlib.import_module(name)
                    continue
            except Exception:
                pass
            reg[name] = _Permissive(name)
    for k, v in reg.items():
        sys.modules[k] = v
    for name in list(reg):
        if "." in name:
            parent, child = name.rsplit(".", 1)
            if parent in sys.modules:
                setattr(sys.modules[parent], child, sys.modules[name])
    return reg


def import_reference(signature_fn=None, sigkernel_module=None):
    """Returns the reference's ``src`` package (kernels, inference, utils imported)."""
    if not reference_available():
        raise ImportError(f"reference tree not found at {REFERENCE_ROOT}")
    install_stubs(signature_fn, sigkernel_module)
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import torch

    src = importlib.import_module("src")
    importlib.import_module("src.kernels")
    importlib.import_module("src.inference")  # NB: turns autograd anomaly detection ON (mpf.py:9)
    torch.autograd.set_detect_anomaly(False)
    return src

"""Fused dispatch for the reference's samplers (``src/inference``), which stay UNCHANGED.

The reference's ``SVGD._velocity`` (src/inference/svgd.py:46-91) builds the Stein direction from
``k_xx``, ``grad_k`` and the score in a dozen torch statements, and copies ``k_xx`` to the host every
step.  ``make_fused(SVGD)`` returns a subclass of whichever ``SVGD`` class the caller imports that
overrides that one method: when the kernel is this package's ``PathSigKernel`` over a
``GaussianKernel`` the whole update (signature -> Gram -> median bandwidth -> exp -> K.[score|S|1] ->
signature VJP -> phi) is ONE call into the CUDA library (``ops.phi`` / a replayed ``ops.PhiPlan``);
every other case falls through to the base class untouched.  ``step`` / ``optimize`` /
``ScoreEstimator`` / ``TrajectorySVGD`` are the reference's own code.

    from src.inference import SVGD                    # the unmodified reference
    from sigsvgd_b200.inference import make_fused
    from sigsvgd_b200.kernels import PathSigKernel
    sampler = make_fused(SVGD)(kernel=PathSigKernel(), optimizer_class=torch.optim.Adam, lr=1e-2)

or ``install()`` to swap the classes inside an already imported ``src.inference`` in place.
"""
from __future__ import annotations

import sys
from typing import Tuple

import torch
import torch.autograd as autograd

from . import ops
from .kernels import GaussianKernel, PathSigKernel, bw_median


class FixedBandwidth:
    """``bandwidth_fn`` returning a constant, recognisable by the fused path.  Equivalent to the
    reference idiom ``bandwidth_fn=lambda _: h`` (script_planning_obstacle_field.py:321-322)."""

    def __init__(self, h: float):
        self.h = float(h)

    def __call__(self, _sq_dists=None) -> float:
        return self.h


def _fused_spec(kernel):
    """Bandwidth spec if ``kernel`` can take the fused path, else None."""
    if not isinstance(kernel, PathSigKernel) or not isinstance(kernel.static_kernel, GaussianKernel):
        return None
    fn = kernel.static_kernel.get_bandwidth
    if fn is bw_median:
        return ops.Bandwidth.median()
    if isinstance(fn, FixedBandwidth):
        return ops.Bandwidth.fixed(fn.h)
    return None


def _target_score(sampler, X, grad_log_p, kwargs):
    """(score [N, T*d], loss) as the base class's ``_velocity`` defines them: the caller's
    ``grad_log_p`` or the autograd of ``log_p``, plus the gradient of ``log_prior`` when there is one."""

    def grad_of(fn):
        Xg = X.detach().requires_grad_(True)
        total = fn(Xg).sum()
        return autograd.grad(total, Xg)[0].detach().flatten(1), total.detach()

    if grad_log_p is None:
        score, log_lik = grad_of(sampler.log_p)
        loss = -log_lik
    else:
        score = grad_log_p.flatten(1)
        loss = kwargs["loss"].sum() if "loss" in kwargs else grad_log_p.norm()
    if getattr(sampler, "log_prior", None) is not None:
        score = score + grad_of(sampler.log_prior)[0]
    return score, loss


def _plain(x) -> bool:
    return isinstance(x, (int, float)) and not isinstance(x, bool)


def _fused_optimizer_step(optimizer, X: torch.Tensor, grad: torch.Tensor) -> bool:
    """``optimizer.step()`` for ``X.grad = grad`` as ONE library launch (``sigsvgd_opt_step``) when the
    optimizer is a stock ``torch.optim.Adam`` / ``torch.optim.SGD`` over ``[X]`` with the flags the
    reference uses (src/inference/svgd.py:131: ``optimizer_class(params=[X], **opt_args)``); the state
    dictionary keeps torch's own layout, so ``state_dict()`` / ``load_state_dict`` work unchanged.
    Returns False (nothing done) for anything else."""
    from . import _lib

    if not (X.is_cuda and X.dtype == torch.float32 and X.is_contiguous() and grad.is_contiguous()):
        return False
    groups = optimizer.param_groups
    if len(groups) != 1 or len(groups[0]["params"]) != 1 or groups[0]["params"][0] is not X:
        return False
    g = groups[0]
    if g.get("weight_decay", 0) != 0 or g.get("maximize", False) or g.get("differentiable", False) \
            or g.get("capturable", False) or g.get("fused", None) or not _plain(g["lr"]):
        return False
    if type(optimizer) is torch.optim.Adam:
        if g.get("amsgrad", False) or not all(_plain(b) for b in g["betas"]):
            return False
        st = optimizer.state[X]
        if len(st) == 0:
            st["step"] = torch.tensor(0.0, dtype=torch.float32)
            st["exp_avg"] = torch.zeros_like(X, memory_format=torch.preserve_format)
            st["exp_avg_sq"] = torch.zeros_like(X, memory_format=torch.preserve_format)
        if st["step"].is_cuda or not st["exp_avg"].is_contiguous() or not st["exp_avg_sq"].is_contiguous():
            return False
        st["step"] += 1
        t = float(st["step"])
        b1, b2 = g["betas"]
        ops.opt_step_raw(X, grad, _lib.OPT_ADAM, g["lr"], 1.0, st["exp_avg"], st["exp_avg_sq"], None, b1, b2, g["eps"],
                         g["lr"] / (1.0 - b1 ** t), (1.0 - b2 ** t) ** 0.5)
        return True
    if type(optimizer) is torch.optim.SGD:
        if g.get("nesterov", False) or g.get("dampening", 0) != 0 or not _plain(g.get("momentum", 0)):
            return False
        mom = g.get("momentum", 0)
        buf, first = None, False
        if mom != 0:
            st = optimizer.state[X]
            buf = st.get("momentum_buffer")
            if buf is None:
                buf, first = torch.empty_like(X), True
                st["momentum_buffer"] = buf
            elif not buf.is_contiguous():
                return False
        ops.opt_step_raw(X, grad, _lib.OPT_SGD, g["lr"], 1.0, buf, None, None, mom, first=first)
        return True
    return False


def make_fused(base_cls):
    """Subclass of ``base_cls`` (an ``SVGD`` class with the reference's interface) whose ``_velocity``
    takes the fused CUDA path when it can.  Class attributes of the result:

    * ``log_level``: ``"full"`` (reference behaviour: every entry of the step's dictionary copied to
      the CPU), ``"device"`` (same entries, left on the device: no blocking D2H of the [N,N] matrix),
      ``"none"`` (``k_xx`` / ``grad_k`` are not materialised at all; the update replays a CUDA graph);
    * ``fused_optimizer``: with ``log_level`` other than "full", the particle update of ``step`` (stock
      ``torch.optim.Adam`` / ``SGD`` over ``[X]``, or the base class's hand-written Adagrad) runs as one
      ``sigsvgd_opt_step`` launch; the optimizer's state dictionary keeps torch's layout;
    * ``depth``: signature depth of the fused path (the reference hard-wires PathSigKernel's default 3);
    * ``last_state``: device record of the last fused update (median, h, tie pairs)."""
    root = next(c for c in reversed(base_cls.__mro__) if "_velocity" in vars(c))
    # classes that re-define how the kernel terms are obtained (TrajectorySVGD) are never short-cut
    fusable = base_cls._velocity is root._velocity and base_cls._compute_kernel is root._compute_kernel

    class Fused(base_cls):
        log_level = "full"
        depth = 3
        use_graph = True
        fused_optimizer = True  # Adam / SGD / Adagrad update as one library launch (log_level != "full")
        last_state = None

        def _velocity(self, X: torch.Tensor, grad_log_p: torch.Tensor, **kwargs) -> Tuple[torch.Tensor, dict]:
            spec = _fused_spec(getattr(self, "kernel", None)) if fusable else None
            given = "k_xx" in kwargs and "grad_k" in kwargs  # caller-supplied kernel terms (svgd.py:54-58)
            no_target = self.log_p is None and grad_log_p is None  # the base class raises for this
            if spec is None or given or no_target or X.dim() != 3:
                return super()._velocity(X, grad_log_p, **kwargs)
            score, loss = _target_score(self, X, grad_log_p, kwargs)
            want = self.log_level != "none"
            if not want and X.is_cuda and self.use_graph:
                # nothing but phi is asked for: replay the update from its CUDA graph (one plan per swarm
                # shape and bandwidth rule; the inputs are copied into the plan's buffers)
                plans = self.__dict__.setdefault("_plans", {})
                key = (tuple(X.shape), X.device, self.depth, spec.mode, spec.h, spec.bw_scale, spec.tol,
                       spec.median_grad)
                plan = plans.get(key)
                if plan is None:
                    plan = plans[key] = ops.PhiPlan(X.shape[0], X.shape[1], X.shape[2], self.depth, bandwidth=spec,
                                                    device=X.device)
                res = plan(X.detach(), score.detach())
                phi, k_xx, grad_k = res.phi.to(X.dtype), None, None
            else:
                res = ops.phi(X.detach(), score.detach(), depth=self.depth, bandwidth=spec, want_k=want,
                              want_grad_k=want)
                phi, k_xx, grad_k = res.phi, res.k_xx, res.grad_k
            self.last_state = res.state
            velocity = -phi.reshape(X.shape)  # the reference's sign: X.grad of a minimising optimizer
            iter_dict = {"k_xx": k_xx, "grad_k": grad_k, "loss": loss}
            iter_dict.update(kwargs)
            to_host = self.log_level == "full"
            iter_dict = {k: ((v.detach().cpu() if to_host else v.detach()) if hasattr(v, "detach") else v)
                         for k, v in iter_dict.items()}
            return velocity, iter_dict

        # ---- the loop without host copies -----------------------------------------------------
        # The base class's ``step`` / ``optimize`` copy the step's gradient and the particle trace to
        # the CPU every iteration (svgd.py:101-113, 143-152): at N = 4096 that is more time than the
        # update itself.  With ``log_level`` other than "full" the same update rules run with the
        # records left on the device; "full" is the base class's own code.
        def step(self, X, grad_log_p=None, optimizer=None, **kwargs):
            if self.log_level == "full":
                return super().step(X, grad_log_p, optimizer, **kwargs)
            grad, iter_dict = self._velocity(X, grad_log_p, **kwargs)
            fuse = self.fused_optimizer and X.is_cuda and X.dtype == torch.float32
            if isinstance(optimizer, torch.optim.Optimizer):
                X.grad = grad
                if not (fuse and _fused_optimizer_step(optimizer, X, grad.contiguous())):
                    optimizer.step()
            elif fuse and X.is_contiguous():
                # the base class's hand-written update rules (svgd.py:110-115) as one launch; X is a new
                # tensor as in the base class
                from . import _lib

                X, grad = X.detach().clone(), grad.contiguous()
                if self.opt_adagrad:
                    if not isinstance(self.opt_inertia, torch.Tensor):
                        self.opt_inertia = torch.full_like(X, float(self.opt_inertia))
                    logged = torch.empty_like(X)
                    ops.opt_step_raw(X, grad, _lib.OPT_ADAGRAD, self.opt_args["lr"], 1.0, None, self.opt_inertia, logged)
                    grad = logged
                else:
                    ops.opt_step_raw(X, grad, _lib.OPT_SGD, self.opt_args["lr"])
            else:
                if self.opt_adagrad:  # the base class's hand-written Adagrad, on the device
                    self.opt_inertia = self.opt_inertia + grad * grad
                    grad = grad * torch.rsqrt(self.opt_inertia + 1e-12)
                X = X - self.opt_args["lr"] * grad
            iter_dict["grad"] = grad
            return X, iter_dict

        def optimize(self, particles, score_estimator=None, opt_state=None, n_steps=100, debug=False,
                     callback_func=None, **kwargs):
            if self.log_level == "full":
                return super().optimize(particles, score_estimator, opt_state, n_steps, debug, callback_func,
                                        **kwargs)
            X = particles.detach()
            optimizer = None
            if self.optimizer_class is not None:
                optimizer = self.optimizer_class(params=[X], **self.opt_args)
                if opt_state is not None:
                    optimizer.load_state_dict(opt_state)
            # trace: "device" -> pinned host buffer filled by asynchronous copies; "none" -> not kept
            trace = None
            if self.log_level == "device":
                trace = torch.empty((n_steps + 1,) + tuple(X.shape), dtype=X.dtype, pin_memory=X.is_cuda)
                trace[0].copy_(X, non_blocking=True)
            data_dict, grad_log_p = {}, None
            for i in range(n_steps):
                if score_estimator is not None:
                    X.requires_grad_(True)
                    grad_log_p, score_dict = score_estimator(X)
                    kwargs.update(score_dict)
                    if optimizer is None:
                        X = X.detach()  # (no autograd chain through the hand-written update)
                X, data_dict[i] = self.step(X, grad_log_p, optimizer, **kwargs)
                if trace is not None:
                    trace[i + 1].copy_(X.detach(), non_blocking=True)
                if callback_func is not None:
                    callback_func(X)
            if trace is not None and X.is_cuda:
                torch.cuda.current_stream(X.device).synchronize()
            data_dict["trace"] = trace
            particles[:] = X.detach()
            return data_dict, (optimizer.state_dict() if optimizer is not None else None)

    Fused.__name__ = base_cls.__name__
    Fused.__qualname__ = base_cls.__qualname__
    Fused.__doc__ = (base_cls.__doc__ or "") + "\n\n(with sigsvgd_b200's fused CUDA update behind ``_velocity``)"
    Fused.__fused_base__ = base_cls
    return Fused


def install(module=None) -> dict:
    """Swap ``SVGD`` (and its subclasses ``TrajectorySVGD`` / ``ScaledSVGD`` re-based on the fused class
    are left alone: they override the kernel terms) inside the reference's ``src.inference`` package,
    so that code which does ``from src.inference import SVGD`` afterwards gets the fused class.
    Returns ``{name: class}`` of what was replaced."""
    if module is None:
        module = sys.modules.get("src.inference")
        if module is None:
            raise ImportError("src.inference is not imported: put the reference on sys.path and import it first")
    done = {}
    base = getattr(module, "SVGD", None)
    if base is not None and not hasattr(base, "__fused_base__"):
        fused = make_fused(base)
        setattr(module, "SVGD", fused)
        sub = sys.modules.get(module.__name__ + ".svgd")
        if sub is not None and getattr(sub, "SVGD", None) is base:
            setattr(sub, "SVGD", fused)
        done["SVGD"] = fused
    return done


__all__ = ["FixedBandwidth", "make_fused", "install"]

"""Host-side mirror of the reference's ``src/inference`` interface for the hot path.

``SVGD`` (src/inference/svgd.py:11-159), ``ScoreEstimator`` (src/inference/score.py:6-76) and
``TrajectorySVGD`` (src/inference/trajectory_svgd.py:13-84): same constructor / method
signatures, same returned dictionaries, same exceptions.  What differs is where the arithmetic
runs: when the kernel is this package's ``PathSigKernel`` over a ``GaussianKernel`` the whole
update (signature -> Gram -> median bandwidth -> exp -> K.[score|S|1] -> signature VJP -> phi) is one
call into the CUDA library (``ops.phi``), otherwise the kernel's own autograd node is used.

The reference's unmodified ``SVGD`` also works with this package's kernels (its
``autograd.grad(k_xx.sum(), X)`` lands in the fused backward); this mirror additionally avoids
materialising / copying ``k_xx`` when the caller does not log it (``log_level``).
"""
from __future__ import annotations

from typing import Callable, Tuple

import torch
import torch.autograd as autograd
import torch.optim as optim

from . import ops
from .kernels import BaseKernel, GaussianKernel, PathSigKernel, bw_median


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


class SVGD:
    """Stein variational gradient descent with custom kernels (src/inference/svgd.py:11-159)."""

    #: "full"   -- reference behaviour: k_xx, grad_k, loss ... copied to the CPU every step
    #: "device" -- same entries, left on the device (no blocking D2H of the [N,N] matrix)
    #: "none"   -- k_xx / grad_k are not materialised at all (entries are None)
    log_level = "full"
    #: signature depth used by the fused path (the reference hard-wires PathSigKernel's default 3)
    depth = 3

    def __init__(self, kernel: BaseKernel = None, log_p: Callable = None, log_prior: Callable = None,
                 bw_scale: float = 1.0, optimizer_class: optim.Optimizer = optim.Adam,
                 adaptive_gradient: bool = False, **opt_args):
        self.kernel = GaussianKernel() if kernel is None else kernel
        self.log_p = log_p
        self.log_prior = log_prior
        self.bw_scale = bw_scale  # stored, never forwarded -- as in the reference (svgd.py:30)
        self.optimizer_class = optimizer_class
        self.opt_args = opt_args
        self.opt_adagrad = adaptive_gradient
        self.opt_inertia = 0
        self.last_state = None  # device record of the last fused update (median, h, ties)

    # ------------------------------------------------------------------------------ kernel terms
    def _compute_kernel(self, X: torch.Tensor, **kwargs):
        """svgd.py:36-44."""
        if hasattr(self.kernel, "analytic_grad") and self.kernel.analytic_grad:
            k_xx, grad_k = self.kernel(X, X)
        else:
            X = X.detach().requires_grad_(True)
            k_xx = self.kernel(X, X.detach(), compute_grad=False)
            grad_k = autograd.grad(k_xx.sum(), X)[0].flatten(1)
        return k_xx.detach(), grad_k.detach()

    def _score(self, X, grad_log_p, kwargs):
        """svgd.py:62-80: score = grad_log_p (+ grad log_prior); returns (score, loss)."""
        if grad_log_p is None:
            Xg = X.detach().requires_grad_(True)
            log_lik = self.log_p(Xg).sum()
            score = autograd.grad(log_lik, Xg)[0].flatten(1)
            loss = -log_lik.detach()
        else:
            score = grad_log_p.flatten(1)
            loss = kwargs["loss"].sum() if "loss" in kwargs else grad_log_p.norm()
        if self.log_prior is not None:
            Xg = X.detach().requires_grad_(True)
            log_prior_grad = autograd.grad(self.log_prior(Xg).sum(), Xg)[0]
            score = score + log_prior_grad.detach().flatten(1)
        return score, loss

    def _velocity(self, X: torch.Tensor, grad_log_p: torch.Tensor, **kwargs) -> Tuple[torch.Tensor, dict]:
        """svgd.py:46-91.  Returns (-phi reshaped like X, iter_dict)."""
        if self.log_p is None and grad_log_p is None:
            raise ValueError(
                """SVGD needs a function to evaluate the log probability of the target
                distribution or an estimate of the gradient for every particle.""",
            )
        N = X.shape[0]
        spec = _fused_spec(self.kernel) if not ("k_xx" in kwargs and "grad_k" in kwargs) else None
        if spec is not None and X.dim() == 3:
            score, loss = self._score(X, grad_log_p, kwargs)
            want = self.log_level != "none"
            if not want and X.is_cuda and getattr(self, "use_graph", True):
                # nothing but phi is asked for: replay the update from its CUDA graph (ops.PhiPlan; one
                # plan per swarm shape and bandwidth rule, inputs copied into the plan's buffers)
                plans = self.__dict__.setdefault("_plans", {})
                key = (tuple(X.shape), X.device, self.depth, spec.mode, spec.h, spec.bw_scale, spec.tol, spec.median_grad)
                plan = plans.get(key)
                if plan is None:
                    plan = plans[key] = ops.PhiPlan(X.shape[0], X.shape[1], X.shape[2], self.depth, bandwidth=spec,
                                                    device=X.device)
                res = plan(X.detach(), score.detach())
                res = ops.PhiResult(res.phi.to(X.dtype), res.state)
            else:
                res = ops.phi(X.detach(), score.detach(), depth=self.depth, bandwidth=spec, want_k=want,
                              want_grad_k=want)
            self.last_state = res.state
            velocity = -res.phi.reshape(X.shape)
            k_xx, grad_k = res.k_xx, res.grad_k
        else:
            if "k_xx" in kwargs and "grad_k" in kwargs:
                k_xx, grad_k = kwargs["k_xx"], kwargs["grad_k"]
                if len(grad_k.shape) > 1:
                    grad_k = grad_k.flatten(1)
            else:
                k_xx, grad_k = self._compute_kernel(X, **kwargs)
            score, loss = self._score(X, grad_log_p, kwargs)
            if k_xx.is_cuda:
                ks = ops.matmul_nn_raw(ops._prep(k_xx, "k_xx"), ops._prep(score, "score")).to(score.dtype)
            else:
                raise RuntimeError("sigsvgd_b200.SVGD: k_xx is not a CUDA tensor; there is no CPU path")
            velocity = (ks - grad_k) / N
            velocity = -velocity.reshape(X.shape)
        iter_dict = {"k_xx": k_xx, "grad_k": grad_k, "loss": loss}
        iter_dict.update(kwargs)
        if self.log_level == "full":
            iter_dict = {k: v.detach().cpu() if hasattr(v, "detach") else v for k, v in iter_dict.items()}
        else:
            iter_dict = {k: v.detach() if hasattr(v, "detach") else v for k, v in iter_dict.items()}
        return velocity, iter_dict

    # ------------------------------------------------------------------------------ stepping
    def step(self, X: torch.Tensor, grad_log_p: torch.Tensor = None, optimizer: optim.Optimizer = None, **kwargs):
        """svgd.py:93-116."""
        to_log = (lambda t: t.cpu()) if self.log_level == "full" else (lambda t: t.detach())

        def closure():
            optimizer.zero_grad()
            X.grad, iter_dict = self._velocity(X, grad_log_p, **kwargs)
            iter_dict["grad"] = to_log(X.grad)
            return iter_dict

        if isinstance(optimizer, torch.optim.Optimizer):
            iter_dict = optimizer.step(closure)
        else:
            grad, iter_dict = self._velocity(X, grad_log_p, **kwargs)
            if self.opt_adagrad:
                self.opt_inertia += grad ** 2
                grad = grad / torch.sqrt(self.opt_inertia + 1e-12)
            iter_dict["grad"] = to_log(grad)
            X = X - self.opt_args["lr"] * grad
        return X, iter_dict

    def optimize(self, particles: torch.Tensor, score_estimator: Callable = None, opt_state: dict = None,
                 n_steps: int = 100, debug: bool = False, callback_func=None, **kwargs) -> tuple:
        """svgd.py:118-159.  ``data_dict[i]`` per step, ``data_dict["trace"]`` = [n_steps+1, ...]."""
        X = particles.detach()
        if self.optimizer_class is not None:
            optimizer = self.optimizer_class(params=[X], **self.opt_args)
            if opt_state is not None:
                optimizer.load_state_dict(opt_state)
        else:
            optimizer = None
        grad_log_p = None
        data_dict = {}
        if debug:
            from tqdm import trange

            iterator = trange(n_steps, position=0, leave=True)
        else:
            iterator = range(n_steps)
        trace = [X.clone().unsqueeze(0)]
        for i in iterator:
            if score_estimator is not None:
                X.requires_grad_(True)
                grad_log_p, score_dict = score_estimator(X)
                kwargs.update(score_dict)
            X, data_dict[i] = self.step(X, grad_log_p, optimizer, **kwargs)
            trace.append(X.detach().clone().unsqueeze(0))
            if debug:
                iterator.set_postfix(loss=data_dict[i]["loss"].norm(), refresh=False)
            if callback_func is not None:
                callback_func(X)
        # one host copy at the end instead of the reference's per-step cat (same tensor)
        data_dict["trace"] = torch.cat(trace, dim=0).cpu()
        particles[:] = X.detach()
        opt_state = optimizer.state_dict() if optimizer is not None else None
        return data_dict, opt_state


class ScoreEstimator:
    """src/inference/score.py:6-76 -- cost -> grad_log_p plus the kernel terms."""

    def __init__(self, kernel, cost_fn, cost_fn_params, scheduler=None, ctx={"device": "cpu"}):
        self.ctx = ctx
        self.kernel = kernel
        self.cost_fn = cost_fn
        self.cost_fn_params = cost_fn_params
        self.scheduler = (lambda: 1) if scheduler is None else scheduler
        if isinstance(self.kernel, BaseKernel):
            self.score = self._svgd_score if self.kernel.analytic_grad is True else self._svgd_ag_score
        elif hasattr(self.kernel, "compute_Gram") or type(self.kernel).__name__ == "SignatureKernel":
            self.score = self._pathsig_score

    def sgd_score(self, x):
        cost, cost_dict = self.cost_fn(x, **self.cost_fn_params)
        grad_log_p = autograd.grad(-cost.sum(), x, retain_graph=True)[0]
        k_xx = torch.eye(x.shape[0], **self.ctx)
        grad_k = torch.zeros_like(grad_log_p, **self.ctx)
        return grad_log_p, {"k_xx": k_xx, "grad_k": grad_k, "loss": cost, **cost_dict}

    def _svgd_score(self, x):
        cost, cost_dict = self.cost_fn(x, **self.cost_fn_params)
        grad_log_p = autograd.grad(-cost.sum(), x, retain_graph=True)[0]
        k_xx, grad_k = self.kernel(x, x, compute_grad=True)
        return grad_log_p, {"k_xx": k_xx, "grad_k": self.scheduler() * grad_k, "loss": cost, **cost_dict}

    def _svgd_ag_score(self, x):
        cost, cost_dict = self.cost_fn(x, **self.cost_fn_params)
        grad_log_p = autograd.grad(-cost.sum(), x, retain_graph=True)[0]
        k_xx = self.kernel(x, x.detach(), compute_grad=False)
        grad_k = autograd.grad(k_xx.sum(), x)[0]
        return grad_log_p, {"k_xx": k_xx, "grad_k": self.scheduler() * grad_k, "loss": cost, **cost_dict}

    def _pathsig_score(self, x):
        cost, cost_dict = self.cost_fn(x, **self.cost_fn_params)
        grad_log_p = autograd.grad(-cost.sum(), x, retain_graph=True)[0]
        k_xx = self.kernel(x, x.detach())
        grad_k = autograd.grad(k_xx.sum(), x)[0]
        return grad_log_p, {"k_xx": k_xx.detach(), "grad_k": self.scheduler() * grad_k.detach(), "loss": cost,
                            **cost_dict}


class TrajectorySVGD(SVGD):
    """src/inference/trajectory_svgd.py:13-84 -- the sampler the Stein-MPC controller builds
    (src/controllers/dust.py:143-166).  Only the ``SigKernel`` branch (:55-69) yields a correctly
    shaped ``grad_k`` in the reference; it is mirrored with
    ``compat.SigKernel`` whose ``compute_Gram`` is this package's autograd node."""

    def __init__(self, kernel: BaseKernel, log_p: Callable = None, log_prior: Callable = None, bw_scale: float = 1.0,
                 gradient_mask=None, optimizer_class: optim.Optimizer = optim.Adam, **opt_args):
        super().__init__(kernel, log_p, log_prior, bw_scale, optimizer_class, **opt_args)
        self.gradient_mask = gradient_mask

    def _compute_kernel(self, X, **kwargs):
        from .compat import SigKernel

        def _tau():
            tau = kwargs["trajectories"][..., 1:, :2]  # x, y positions from t+1 on
            return tau.mean(0) if kwargs["sample_shape"] else tau

        if isinstance(self.kernel, PathSigKernel):  # trajectory_svgd.py:48-54 (shape quirk kept)
            tau = _tau()
            k_xx, grad_k = self.kernel(tau, tau, X, compute_grad=True)
            grad_k = grad_k.detach().flatten(1)
        elif isinstance(self.kernel, SigKernel):  # trajectory_svgd.py:55-69
            tau = _tau()
            k_xx = self.kernel.compute_Gram(tau.double(), tau.detach().double(), sym=False).float()
            grad_k = autograd.grad(k_xx.sum(), kwargs["actions"])[0]
            grad_k = grad_k.mean(0).flatten(1) if kwargs["sample_shape"] else grad_k.flatten(1)
        elif hasattr(self.kernel, "analytic_grad") and self.kernel.analytic_grad:
            k_xx, grad_k = self.kernel(X, X)
            grad_k = grad_k.sum(1)
        else:
            X = X.detach().requires_grad_(True)
            k_xx = self.kernel(X, X.detach(), compute_grad=False)
            grad_k = autograd.grad(-k_xx.sum(), X)[0]
        return k_xx.detach(), grad_k.detach()

    def _velocity(self, X: torch.Tensor, grad_log_p: torch.Tensor, **kwargs) -> Tuple[torch.Tensor, dict]:
        if not ("k_xx" in kwargs and "grad_k" in kwargs):
            k_xx, grad_k = self._compute_kernel(X, **kwargs)
            kwargs = dict(kwargs, k_xx=k_xx, grad_k=grad_k)
        velocity, iter_dict = super()._velocity(X, grad_log_p, **kwargs)
        return velocity * self.gradient_mask, iter_dict


__all__ = ["SVGD", "ScoreEstimator", "TrajectorySVGD", "FixedBandwidth"]

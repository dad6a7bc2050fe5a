"""TEST INFRASTRUCTURE: a restatement of the reference's ``src/inference`` samplers --
``SVGD`` (src/inference/svgd.py:11-159), ``ScoreEstimator`` (src/inference/score.py:6-76),
``TrajectorySVGD`` (src/inference/trajectory_svgd.py:13-84) -- for the GPU box, where
``/root/reference`` does not exist.  Statement for statement the reference's logic on stock torch;
nothing here is product code.  The product is ``sigsvgd_b200.inference.make_fused``, which the tests
apply to these classes exactly as a user applies it to the reference's own (``tests/
test_reference_wiring.py`` does the latter in the build container)."""
from __future__ import annotations

from typing import Callable, Tuple

import torch
import torch.autograd as autograd
import torch.optim as optim

from sigsvgd_b200.kernels import BaseKernel, GaussianKernel, PathSigKernel


class SVGD:
    """Stein variational gradient descent with custom kernels (src/inference/svgd.py:11-159)."""

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
        if "k_xx" in kwargs and "grad_k" in kwargs:
            k_xx, grad_k = kwargs["k_xx"], kwargs["grad_k"]
            if len(grad_k.shape) > 1:
                grad_k = grad_k.flatten(1)
        else:
            k_xx, grad_k = self._compute_kernel(X, **kwargs)
        score, loss = self._score(X, grad_log_p, kwargs)
        velocity = (k_xx @ score - grad_k) / X.shape[0]
        velocity = -velocity.reshape(X.shape)
        iter_dict = {"k_xx": k_xx, "grad_k": grad_k, "loss": loss}
        iter_dict.update(kwargs)
        iter_dict = {k: v.detach().cpu() if hasattr(v, "detach") else v for k, v in iter_dict.items()}
        return velocity, iter_dict

    # ------------------------------------------------------------------------------ stepping
    def step(self, X: torch.Tensor, grad_log_p: torch.Tensor = None, optimizer: optim.Optimizer = None, **kwargs):
        """svgd.py:93-116."""
        to_log = lambda t: t.cpu()  # noqa: E731

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
        data_dict["trace"] = X.clone().unsqueeze(0).cpu()
        for i in iterator:
            if score_estimator is not None:
                X.requires_grad_(True)
                grad_log_p, score_dict = score_estimator(X)
                kwargs.update(score_dict)
            X, data_dict[i] = self.step(X, grad_log_p, optimizer, **kwargs)
            data_dict["trace"] = torch.cat([data_dict["trace"], X.detach().unsqueeze(0).cpu()], dim=0)
            if debug:
                iterator.set_postfix(loss=data_dict[i]["loss"].norm(), refresh=False)
            if callback_func is not None:
                callback_func(X)
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
        from sigsvgd_b200.compat import SigKernel

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
        velocity, iter_dict = super()._velocity(X, grad_log_p, **kwargs)
        return velocity * self.gradient_mask, iter_dict


__all__ = ["SVGD", "ScoreEstimator", "TrajectorySVGD"]

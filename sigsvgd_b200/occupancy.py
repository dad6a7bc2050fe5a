"""Learned occupancy / self-collision networks of the robot-planning cost on the device.

The reference's cost (examples/script_planning_robot.py:248-356) spends its time in two small networks,
``ContinuousOccupancyMap`` (src/models/robot_learning/continuous_occupancy_map.py:10-35: Linear(3, 200),
ReLU, 3 x [Linear(200, 200), ReLU], Linear(200, 1), Sigmoid) evaluated at batch x timesteps x body-point
positions (``exp_datapack.occmap(body_points)``, :302) and the same stack over joint configurations
(``exp_datapack.self_collision_pred(qs)``, :317; continuous_self_collision_pred.py:11-36), and in torch's
autograd through them (``ag(-cost.sum(), x)``, :107).

``fuse(module)`` returns a callable with the module's call signature -- ``f(points[..., in]) -> [..., 1]`` --
that answers from ONE CUDA launch computing the probabilities and their gradient w.r.t. the points (the
hidden-to-hidden layers as tcgen05 GEMMs); the gradient w.r.t. the network's weights is not formed (the
planner holds them fixed).  A maintainer's two lines::

    exp_datapack.occmap = sigsvgd_b200.occupancy.fuse(exp_datapack.occmap)
    exp_datapack.self_collision_pred = sigsvgd_b200.occupancy.fuse(exp_datapack.self_collision_pred)

Arithmetic: ``precise=True`` (default) holds every operand as fp16 hi + lo planes and spends 3 MMAs per product:
the reference's fp32 GEMMs to fp32 accuracy, so that the ReLU patterns (and the piecewise-constant Jacobians)
agree with the reference's.  ``precise=False`` uses the hi planes only (about 1.5x faster): probabilities to
~5e-4, and roughly 1 point in 20 has a pre-activation within fp16 rounding of zero and takes the neighbouring
linear piece's gradient.  tests/test_gpu_occupancy.py states both tolerances.  There is no CPU path."""
from __future__ import annotations

from typing import List, Tuple

import torch

from . import _lib, ops

MAX_IN, MAX_HIDDEN, MAX_MID = 8, 208, 4


def linear_stack(module) -> Tuple[List[torch.Tensor], List[torch.Tensor]]:
    """(weights, biases) of a module shaped like the reference's networks: ``module.layers`` (or the module
    itself) iterates over Linear / ReLU pairs and ends in Linear(H, 1), Sigmoid."""
    layers = list(getattr(module, "layers", module))
    lin = [m for m in layers if isinstance(m, torch.nn.Linear)]
    kinds = [type(m).__name__ for m in layers]
    want = ["Linear", "ReLU"] * (len(lin) - 1) + ["Linear", "Sigmoid"]
    if len(lin) < 3 or kinds != want:
        raise ValueError(f"fuse: expected Linear/ReLU ... Linear/Sigmoid with at least three Linear layers, got {kinds}")
    return [m.weight for m in lin], [m.bias for m in lin]


class _Query(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x2d, fused):
        prob, grad = fused.query_raw(x2d, want_grad=ctx.needs_input_grad[0])
        ctx.grad = grad
        return prob

    @staticmethod
    def backward(ctx, gout):
        if ctx.grad is None:
            return None, None
        return ctx.grad * gout.to(torch.float32).unsqueeze(-1), None


class FusedOccupancy:
    """Packed weights of one network plus the query call.  ``repack()`` after the weights change."""

    def __init__(self, weights, biases, device=None, precise: bool = True):
        dev = torch.device(device) if device is not None else weights[0].device
        if dev.type != "cuda":
            raise RuntimeError("FusedOccupancy: CUDA only; sigsvgd_b200 has no CPU fallback")
        self.device, self.precise = dev, bool(precise)
        self._src = (list(weights), list(biases))
        self.repack()

    def repack(self):
        ws, bs = self._src
        f32 = lambda t: t.detach().to(device=self.device, dtype=torch.float32).contiguous()  # noqa: E731
        hidden, in_dim = ws[0].shape
        n_mid = len(ws) - 2
        if ws[-1].shape != (1, hidden) or any(w.shape != (hidden, hidden) for w in ws[1:-1]):
            raise ValueError("FusedOccupancy: layers must be (in -> H), (H -> H) ..., (H -> 1)")
        if any(b is None for b in bs):
            raise ValueError("FusedOccupancy: every Linear layer must have a bias")
        self.in_dim, self.hidden, self.n_mid = int(in_dim), int(hidden), int(n_mid)
        nbytes = int(_lib.load().sigsvgd_mlp_packed_bytes(self.in_dim, self.hidden, self.n_mid))
        if nbytes == 0:
            raise ValueError(f"FusedOccupancy: unsupported shape in={in_dim} (<= {MAX_IN}), hidden={hidden} "
                             f"(<= {MAX_HIDDEN}), hidden-to-hidden layers={n_mid} (1 .. {MAX_MID})")
        with torch.cuda.device(self.device):
            w_mid = torch.stack([f32(w) for w in ws[1:-1]]).contiguous()
            b_mid = torch.stack([f32(b) for b in bs[1:-1]]).contiguous()
            w0, b0, wl, bl = f32(ws[0]), f32(bs[0]), f32(ws[-1]).reshape(-1), f32(bs[-1]).reshape(-1)
            self.packed = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
            ops.call("sigsvgd_mlp_pack", ops._p(w0), ops._p(b0), ops._p(w_mid), ops._p(b_mid), ops._p(wl), ops._p(bl),
                     self.in_dim, self.hidden, self.n_mid, ops._p(self.packed), nbytes, ops._stream())
        return self

    def query_raw(self, x2d: torch.Tensor, want_grad: bool = True, precise=None):
        """x [P, in] float32 CUDA -> (prob [P], d prob / d x [P, in] or None)."""
        precise = self.precise if precise is None else bool(precise)
        if not x2d.is_cuda:
            raise RuntimeError("FusedOccupancy: CUDA tensors only; sigsvgd_b200 has no CPU fallback")
        if x2d.dim() != 2 or x2d.shape[1] != self.in_dim:
            raise ValueError(f"FusedOccupancy: points [P, {self.in_dim}] expected, got {tuple(x2d.shape)}")
        x = x2d.detach().to(torch.float32).contiguous()
        P = x.shape[0]
        with torch.cuda.device(x.device):
            prob = torch.empty(P, dtype=torch.float32, device=x.device)
            grad = torch.empty(P, self.in_dim, dtype=torch.float32, device=x.device) if want_grad else None
            ops.call("sigsvgd_mlp_query", ops._p(self.packed), self.in_dim, self.hidden, self.n_mid, int(precise), ops._p(x), P,
                     ops._p(prob), ops._p(grad), ops._stream())
        return prob, grad

    def __call__(self, x: torch.Tensor) -> torch.Tensor:
        lead = x.shape[:-1]
        out = _Query.apply(x.reshape(-1, x.shape[-1]), self)
        return out.reshape(*lead, 1).to(x.dtype)


def fuse(module, device=None, precise: bool = True) -> FusedOccupancy:
    """Drop-in for ``module.__call__`` (a ``ContinuousOccupancyMap`` / ``ContinuousSelfCollisionPredictor`` or any
    module with the same ``layers`` stack)."""
    ws, bs = linear_stack(module)
    return FusedOccupancy(ws, bs, device, precise)


__all__ = ["FusedOccupancy", "fuse", "linear_stack"]

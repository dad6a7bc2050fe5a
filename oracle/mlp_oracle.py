"""CPU restatement (numpy, float64) of the robot-planning cost's two networks and of the input gradient
autograd forms through them.  TEST INFRASTRUCTURE ONLY -- nothing under ``sigsvgd_b200/``
imports this.

Follows (relative to the reference root):

* ``src/models/robot_learning/continuous_occupancy_map.py:10-35`` ``ContinuousOccupancyMap``: for
  ``i in range(n_hidden_layers)``: ``Linear(n_dimension, hidden)`` + ReLU (i = 0), ``Linear(hidden, hidden)`` + ReLU,
  ``Linear(hidden, 1)`` + Sigmoid (last); ``forward = self.layers``;
* ``src/models/robot_learning/continuous_self_collision_pred.py:11-36``: the same stack over ``n_dimension`` = dof;
* ``examples/script_planning_robot.py:300-318`` (the calls) and ``:107`` (``ag(-cost.sum(), x)``: the gradient
  w.r.t. the query points is what the planner needs).

Pinned by ``tests/golden/occupancy_mlp.npz``: outputs and autograd gradients of the UNMODIFIED reference
classes, produced here by ``oracle/make_golden_mlp.py`` (random-initialised weights with ``torch.manual_seed``:
the trained weights are downloaded by the reference's Makefile and are not available offline)."""
from __future__ import annotations

import numpy as np


def mlp_forward(x, weights, biases):
    """prob [P] and the per-layer pre-activation signs, for x [P, in]; weights[i] [out, in] (nn.Linear layout)."""
    h = np.asarray(x, np.float64)
    masks = []
    for i, (w, b) in enumerate(zip(weights, biases)):
        z = h @ np.asarray(w, np.float64).T + np.asarray(b, np.float64)
        if i + 1 < len(weights):
            masks.append(z > 0)
            h = np.where(z > 0, z, 0.0)
        else:
            h = z
    logit = h[:, 0]
    return 1.0 / (1.0 + np.exp(-logit)), masks


def mlp_query(x, weights, biases):
    """(prob [P], d prob_p / d x_p [P, in])."""
    prob, masks = mlp_forward(x, weights, biases)
    g = np.broadcast_to(np.asarray(weights[-1], np.float64)[0], (len(prob), weights[-1].shape[1])).copy()  # d logit / d h_last
    for w, m in zip(reversed(weights[:-1]), reversed(masks)):
        g = (g * m) @ np.asarray(w, np.float64)
    return prob, g * (prob * (1.0 - prob))[:, None]


def _h(a):
    """round to half precision (the kernel's operand format), back in float64"""
    return np.asarray(a, np.float64).astype(np.float16).astype(np.float64)


def mlp_query_fp16_contract(x, weights, biases):
    """The arithmetic contract of the CUDA kernel (sigsvgd_b200/csrc/mlp_tc.cu), restated in float64: the
    hidden-to-hidden layers multiply HALF-PRECISION operands (weights rounded once, activations and the
    back-propagated rows rounded per layer, activations clamped to 65504) and accumulate exactly; the first
    and last layers and the biases are float32 data used as they are.  Same return as ``mlp_query``.
    The ReLU pattern is the one THIS forward pass finds: where a pre-activation lies within the half-precision
    rounding error of zero it can differ from the float32 network's, and the gradient then belongs to the
    neighbouring linear piece."""
    ws = [np.asarray(w, np.float64) for w in weights]
    bs = [np.asarray(b, np.float64) for b in biases]
    h = np.minimum(np.maximum(np.asarray(x, np.float64) @ ws[0].T + bs[0], 0.0), 65504.0)
    masks = [h > 0]
    for w, b in zip(ws[1:-1], bs[1:-1]):
        h = np.minimum(np.maximum(_h(h) @ _h(w).T + b, 0.0), 65504.0)
        masks.append(h > 0)
    logit = h @ ws[-1][0] + bs[-1][0]
    prob = 1.0 / (1.0 + np.exp(-logit))
    g = np.where(masks[-1], ws[-1][0][None, :], 0.0)
    for w, m in zip(reversed(ws[1:-1]), reversed(masks[:-1])):
        g = np.where(m, _h(g) @ _h(w), 0.0)
    return prob, (g @ ws[0]) * (prob * (1.0 - prob))[:, None]

"""GPU tests of the particle-model rollout kernels (rollout.cu) through the C ABI:
bit-identical states against the golden vectors of the UNMODIFIED reference ParticleModel stepped as
DuSt._rollout does (tests/golden/rollout_maze.npz), gradients against its autograd, and the chain
rollout -> signature-PDE kernel -> gradient w.r.t. the actions (trajectory_svgd.py:55-69)."""
import math

import numpy as np
import pytest
import torch

from oracle import rollout_oracle as ro
from tests.test_oracle_rollout import golden_case
from tests.util import rel_max

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    from sigsvgd_b200 import compat
    from sigsvgd_b200.rollout import particle_rollout, rollout_with

DEV = "cuda:0"


def cu(a, dtype=None):
    t = torch.as_tensor(np.ascontiguousarray(a), device=DEV)
    return t if dtype is None else t.to(dtype)


@pytest.mark.parametrize("name", ["free", "wild"])
def test_rollout_vs_reference_golden(name):
    g, grid, kw = golden_case(name)
    actions = cu(g[f"{name}_actions"]).requires_grad_(True)
    states = particle_rollout(cu(g["init_state"]), actions, cu(g[f"{name}_mass"]), kw["dt"], math.inf, kw["max_speed"],
                              cu(grid, torch.uint8), kw["cell_size"], kw["c_offset"], True)
    assert np.array_equal(states.detach().cpu().numpy(), g[f"{name}_states"])  # bit for bit
    (ga,) = torch.autograd.grad((states * cu(g[f"{name}_w"])).sum(), actions)
    assert rel_max(ga.cpu().numpy(), g[f"{name}_grad_actions"]) <= 1e-6
    # second call: nothing left behind
    again = particle_rollout(cu(g["init_state"]), actions, cu(g[f"{name}_mass"]), kw["dt"], math.inf, kw["max_speed"],
                             cu(grid, torch.uint8), kw["cell_size"], kw["c_offset"], True)
    assert torch.equal(again, states)


def test_rollout_bounds_scalar_mass_no_grid_vs_oracle():
    rng = np.random.default_rng(3)
    R, H = 77, 13
    actions = (20 * rng.standard_normal((R, H, 2))).astype(np.float32)
    init = np.array([0.3, -0.2, 1.0, -2.0], np.float32)
    want, flags = ro.particle_rollout_np(init, actions, 1.7, 0.05, max_acc=8.0, max_speed=3.0)
    a = cu(actions).requires_grad_(True)
    got = particle_rollout(cu(init), a, 1.7, 0.05, 8.0, 3.0)
    assert np.array_equal(got.detach().cpu().numpy(), want) and flags[..., 1:].any()
    w = rng.standard_normal(want.shape).astype(np.float32)
    (ga,) = torch.autograd.grad((got * cu(w)).sum(), a)
    assert rel_max(ga.cpu().numpy(), ro.particle_rollout_bwd_np(w, flags, 1.7, 0.05)) <= 1e-6
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        particle_rollout(torch.as_tensor(init), torch.as_tensor(actions), 1.7, 0.05)
    with pytest.raises(ValueError, match="mass must be > 0"):
        particle_rollout(cu(init), a, 0.0, 0.05)


def test_rollout_with_a_reference_shaped_model_and_the_kernel_gradient_chain():
    """The statements of TrajectorySVGD's SigKernel branch over device rollouts: tau = mean rollout positions,
    k_xx = compute_Gram(tau.double(), tau.detach().double()).float(), grad wrt the ACTIONS through the rollout."""
    g, grid, kw = golden_case("free")

    class ObstMap:  # the attributes rollout_with reads from the reference's ObstacleMap / ParticleModel
        map, cell_size, c_offset = torch.as_tensor(grid, dtype=torch.float32), kw["cell_size"], torch.tensor(kw["c_offset"])

    class Model:
        dt, params_dict, control_type, deterministic = kw["dt"], {"mass": 2.0}, "acceleration", True
        can_crash = with_obstacle = True
        obst_map = ObstMap()
        _ParticleModel__max_speed, _ParticleModel__max_acc = torch.tensor(kw["max_speed"]), torch.tensor(float("inf"))

    n_samples, n_pol, hz = 10, 35, 30
    actions = cu(g["free_actions"]).reshape(n_samples, n_pol, hz, 2).clone().requires_grad_(True)
    run = rollout_with(Model())
    states = run(cu(g["init_state"]), actions.flatten(end_dim=-3), {"mass": cu(g["free_mass"])})
    assert np.array_equal(states.detach().cpu().numpy(), g["free_states"])
    tau = states.reshape(n_samples, n_pol, hz + 1, 4)[..., 1:, :2].mean(0)
    kernel = compat.PdeSigKernel(compat.RBFKernel(sigma=(2 + hz) ** 0.5), dyadic_order=1)
    k_xx = kernel.compute_Gram(tau.double(), tau.detach().double(), sym=False).float()
    grad_k = torch.autograd.grad(k_xx.sum(), actions)[0].mean(0).flatten(1)
    assert grad_k.shape == (n_pol, hz * 2) and torch.isfinite(grad_k).all() and float(grad_k.abs().max()) > 0
    # the same gradient with torch's own rollout graph (double integrator statements of particle.py:155-172)
    a2 = cu(g["free_actions"]).reshape(n_samples, n_pol, hz, 2).clone().requires_grad_(True)
    flat, m = a2.flatten(end_dim=-3), cu(g["free_mass"])
    s = cu(g["init_state"]).expand(flat.shape[0], -1)
    seq = [s]
    fl = ro.particle_rollout_np(g["init_state"], g["free_actions"], g["free_mass"], grid=grid, **kw)[1]
    for t in range(hz):
        f = cu(1.0 - fl[:, t, 0:1].astype(np.float32))
        nxt = s + torch.cat([s[:, 2:], flat[:, t] / m], 1) * kw["dt"] * f
        s = torch.cat([nxt[:, :2], nxt[:, 2:].clamp(-kw["max_speed"], kw["max_speed"])], 1)
        seq.append(s)
    tau2 = torch.stack(seq, 1).reshape(n_samples, n_pol, hz + 1, 4)[..., 1:, :2].mean(0)
    k2 = kernel.compute_Gram(tau2.double(), tau2.detach().double(), sym=False).float()
    grad2 = torch.autograd.grad(k2.sum(), a2)[0].mean(0).flatten(1)
    assert rel_max(grad_k.cpu().numpy(), grad2.cpu().numpy()) <= 1e-5

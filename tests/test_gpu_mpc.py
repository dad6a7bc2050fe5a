"""GPU tests of the Stein-MPC leg (BASELINE.json configs[2]): the ``sigkernel``-shaped
adapter ``compat.SigKernel`` behind the ``SigKernel`` branch of ``TrajectorySVGD``
(src/inference/trajectory_svgd.py:55-69), at the particle-maze shapes (35 policies x 30 steps x 2,
10 action samples), against

* the golden vectors produced by the UNMODIFIED ``DuSt`` + ``ParticleModel`` controller in the build
  container (oracle/make_golden_mpc.py -> tests/golden/mpc_maze_35x30x2.npz), and
* float64 torch autograd through the oracle signature for the chain back to ``actions``.
"""
import warnings

import numpy as np
import pytest
import torch

import oracle
from tests.util import TOL_H, TOL_K, TOL_PHI, closed_form_with_ties, load_golden, rel_max

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    from sigsvgd_b200 import compat, ops
    from sigsvgd_b200.inference import make_fused
    from tests import ref_mirror

DEV = "cuda:0"


def _kernel(**kw):
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        return compat.SigKernel(compat.RBFKernel(**kw), dyadic_order=3)


@pytest.mark.parametrize("call", range(6))
def test_compute_gram_vs_the_unmodified_controller(call):
    g = load_golden("mpc_maze_35x30x2")
    tau = torch.as_tensor(g["tau"][call], device=DEV).requires_grad_(True)  # fp32 [35,30,2], as DuSt produced it
    # the reference's statements (trajectory_svgd.py:60-63): double in, float out, sym=False
    k_xx = _kernel(sigma=32 ** 0.5).compute_Gram(tau.double(), tau.detach().double(), sym=False).float()
    assert k_xx.dtype == torch.float32 and k_xx.shape == (35, 35)
    assert np.abs(k_xx.detach().cpu().numpy() - g["k_xx"][call]).max() <= TOL_K
    dk = torch.autograd.grad(k_xx.sum(), tau)[0].cpu().numpy()
    # parity protocol of SURVEY.md 8(a-4): rows of the median pair are compared with the closed form
    # evaluated on the pair THIS implementation reports, every other row with the controller's own numbers
    rec = ops.read_state(k_xx.grad_fn.next_functions[0][0].bw_state) if hasattr(k_xx.grad_fn.next_functions[0][0], "bw_state") else None
    X64 = g["tau"][call].astype(np.float64)
    ties = rec["ties"] if rec else []
    cf = closed_form_with_ties(X64, np.zeros((35, 60)), 3, ties)
    assert rel_max(dk.reshape(35, -1), cf["grad_k"]) <= TOL_PHI
    tie_rows = sorted({a for a, _ in ties})
    ref_rows = [r for r in range(35) if r not in tie_rows]
    gold = g["dk_dtau"][call].reshape(35, -1)
    # rows outside BOTH tie sets (the golden's own pair shows as its two largest-deviation rows)
    dev = np.abs(dk.reshape(35, -1) - gold).max(1) / np.abs(gold).max()
    assert np.sort(dev[ref_rows])[:-2].max() <= TOL_PHI


def _rollout(actions, x0, dt=0.015):
    """Toy stand-in of ParticleModel.step under acceleration control (double integrator):
    actions [S,P,H,2] -> trajectories [S,P,H+1,4] = (x, y, vx, vy), differentiable wrt actions."""
    vel = torch.cumsum(actions * dt, dim=-2) + x0[2:]
    pos = torch.cumsum(vel * dt, dim=-2) + x0[:2]
    first = x0.expand(actions.shape[0], actions.shape[1], 1, 4)
    return torch.cat([first, torch.cat([pos, vel], dim=-1)], dim=-2)


def test_trajectory_svgd_sigkernel_branch_reaches_actions_through_the_rollout():
    S, P, H = 10, 35, 30
    gen = torch.Generator().manual_seed(4)
    pol = (5.0 * torch.randn(P, H, 2, generator=gen)).to(DEV)
    actions = (pol[None] + 5.0 * torch.randn(S, P, H, 2, generator=gen).to(DEV)).requires_grad_(True)
    x0 = torch.tensor([-1.8, -1.8, 0.0, 0.0], device=DEV)
    score = torch.randn(P, H * 2, generator=gen).to(DEV)
    mask = torch.ones(P, H, 2, device=DEV)
    mask[-5:] = 0.0  # the five primitive policies are not optimised (dust.py:141-142)
    for fixed in (True, False):
        sampler = make_fused(ref_mirror.TrajectorySVGD)(kernel=_kernel(sigma=0.9, fixed=fixed), gradient_mask=mask,
                                                        optimizer_class=None, lr=1.0)
        # (a fresh rollout graph per call, like DuSt's score_estimator: dust.py:496-505)
        kw = lambda: dict(trajectories=_rollout(actions, x0), actions=actions, sample_shape=(S,))  # noqa: E731
        k_xx, grad_k = sampler._compute_kernel(pol, **kw())
        assert k_xx.shape == (P, P) and grad_k.shape == (P, H * 2) and k_xx.dtype == torch.float32
        # float64 on the CPU: the same statements with the oracle signature
        a64 = actions.detach().cpu().double().requires_grad_(True)
        tau64 = _rollout(a64, x0.cpu().double())[..., 1:, :2].mean(0)
        s1 = oracle.signature_torch(tau64, 3, basepoint=True)
        s2 = s1.detach()
        sq = ((s1[:, None] - s2[None]) ** 2).sum(-1)
        h = torch.tensor(0.9, dtype=torch.float64) if fixed else torch.sqrt(torch.median(sq) / torch.tensor(P + 1.0).log())
        K64 = torch.exp(-0.5 / h ** 2 * sq)
        gk64 = torch.autograd.grad(K64.sum(), a64)[0].mean(0).flatten(1)
        assert np.abs(k_xx.cpu().numpy() - K64.detach().numpy()).max() <= TOL_K
        if fixed:
            assert rel_max(grad_k.cpu().numpy(), gk64.numpy()) <= TOL_PHI
        else:  # the median pair may differ between the fp32 and the fp64 matrix: all other rows agree
            dev = np.abs(grad_k.cpu().numpy() - gk64.numpy()).max(1) / np.abs(gk64.numpy()).max()
            assert np.sort(dev)[:-4].max() <= TOL_PHI
        vel, it = sampler._velocity(pol, score, **kw())
        want = -((k_xx @ score - grad_k) / P).reshape(P, H, 2) * mask
        assert rel_max(vel.cpu().numpy(), want.cpu().numpy()) <= 1e-6
        assert float(vel[-5:].abs().max()) == 0.0 and set(("k_xx", "grad_k", "loss", "actions")) <= set(it)

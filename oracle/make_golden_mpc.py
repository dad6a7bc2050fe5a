"""Golden vectors of the Stein-MPC leg (BASELINE.json configs[2]), produced in THIS
container by the UNMODIFIED reference controller:

    python -m oracle.make_golden_mpc        # from the repo root; needs /root/reference

``src.controllers.DuSt`` + ``src.models.ParticleModel`` + ``examples/particle_maze_config.yaml`` run three
control steps (``opt_steps = 2``) through the ``SigKernel`` branch of ``TrajectorySVGD``
(src/inference/trajectory_svgd.py:55-69).  ``sigkernel`` is an absent third-party package; the
stand-in registered here has ``compute_Gram(X, Y)`` = the reference's OWN
``src.kernels.PathSigKernel()(X, Y, compute_grad=False)`` (truncated signature + RBF with the median
heuristic, the north-star kernel), with ``signatory.signature`` = the oracle restatement.  So every
statement that runs is reference code except the signature transform.

Stored per ``compute_Gram`` call (3 control steps x 2 optimiser steps): ``tau`` [35,30,2] (fp32 mean
rollout positions the kernel sees), ``k_xx`` [35,35], ``dk_dtau`` = d(k_xx.sum())/d tau in float64, and
per optimiser step ``grad_k`` [35,60] as the sampler used it; per control step the chosen ``a_seq``.
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
OUT = os.path.join(ROOT, "tests", "golden", "mpc_maze_35x30x2.npz")


def build_controller(record):
    """The harness of SURVEY.md 8(d) config 3.  Returns (controller, simulator, system, state, prior)."""
    sys.path.insert(0, ROOT)
    from oracle import ref_shim

    holder = {}
    sk = types.ModuleType("sigkernel")

    class RBFKernel:
        def __init__(self, sigma=1.0):
            self.sigma = sigma

    class SigKernel:
        def __init__(self, static_kernel=None, dyadic_order=0):
            self.static_kernel, self.dyadic_order = static_kernel, dyadic_order

        def compute_Gram(self, X, Y, sym=False):
            kernel = holder["PathSigKernel"]()
            # side computation for the fixture: gradient of sum(K) wrt the paths themselves
            X2 = X.detach().clone().requires_grad_(True)
            K2 = kernel(X2, X2.detach(), compute_grad=False)
            g = torch.autograd.grad(K2.sum(), X2)[0]
            record["tau"].append(X.detach().float().numpy().copy())
            record["k_xx"].append(K2.detach().numpy().copy())
            record["dk_dtau"].append(g.numpy().copy())
            return kernel(X, Y, compute_grad=False)

    sk.SigKernel, sk.RBFKernel = SigKernel, RBFKernel
    ref_shim.import_reference(sigkernel_module=sk)
    # (if the reference was imported earlier in this process, its isinstance check still points at the
    # stand-in class of that import: trajectory_svgd.py:9 binds the name at import time)
    tsvgd = sys.modules.get("src.inference.trajectory_svgd")
    if tsvgd is not None:
        tsvgd.SigKernel = SigKernel
    import yaml
    from src.controllers import DuSt
    from src.kernels import PathSigKernel
    from src.models import ParticleModel

    holder["PathSigKernel"] = PathSigKernel
    cfg = yaml.safe_load(open(os.path.join(ref_shim.REFERENCE_ROOT, "examples", "particle_maze_config.yaml")))
    exp_hp, env_hp = cfg["exp_params"], cfg["env_params"]
    prior = torch.distributions.Normal(exp_hp["dyn_prior_arg1"], exp_hp["dyn_prior_arg2"])
    kw = {"uncertain_params": ["mass"], "mass": prior.mean}
    system, simulator = ParticleModel(**env_hp, **kw), ParticleModel(**env_hp, **kw)
    prim = torch.zeros(5, exp_hp["horizon"], 2)  # examples/script_control_particle_maze.py:197-201
    prim[1], prim[2] = -10.0, 10.0
    prim[3, ..., :] = torch.tensor([-10.0, 10.0])
    prim[4, ..., :] = torch.tensor([10.0, -10.0])
    kernel = SigKernel(RBFKernel(sigma=(exp_hp["ctrl_dim"] + exp_hp["horizon"]) ** 0.5), exp_hp["dyadic_order"])
    controller = DuSt(
        observation_space=simulator.observation_space, action_space=simulator.action_space,
        hz_len=exp_hp["horizon"], n_pol=exp_hp["n_policies"], n_action_samples=exp_hp["action_samples"],
        n_params_samples=exp_hp["params_samples"], pol_mean=None,
        pol_cov=torch.eye(exp_hp["ctrl_dim"]) * exp_hp["ctrl_sigma"] ** 2, pol_hyper_prior=True,
        stein_sampler=exp_hp["stein_sampler"], kernel=kernel, temperature=exp_hp["alpha"],
        params_log_space=exp_hp["mpf_log_space"], inst_cost_fn=simulator.default_inst_cost,
        term_cost_fn=simulator.default_term_cost, action_primitives=prim, optimizer_class=torch.optim.Adam,
        lr=exp_hp["learning_rate"])
    state = torch.tensor(env_hp["init_state"], dtype=torch.float32)
    return controller, simulator, system, state, prior, exp_hp


def main():
    torch.manual_seed(0)
    np.random.seed(0)
    record = {"tau": [], "k_xx": [], "dk_dtau": []}
    controller, simulator, system, state, prior, exp_hp = build_controller(record)
    torch.autograd.set_detect_anomaly(False)
    a_seqs, grad_ks = [], []
    for _ in range(3):
        a_seq, data = controller.forward(state, simulator, prior, opt_steps=exp_hp["opt_steps"])
        a_seqs.append(a_seq.numpy().copy())
        for i in range(exp_hp["opt_steps"]):
            grad_ks.append(data[i]["grad_k"].numpy().copy())
        state = system.step(state, a_seq[0].squeeze())
    out = dict(tau=np.stack(record["tau"]), k_xx=np.stack(record["k_xx"]), dk_dtau=np.stack(record["dk_dtau"]),
               grad_k=np.stack(grad_ks), a_seq=np.stack(a_seqs))
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    np.savez_compressed(OUT, **out)
    print({k: v.shape for k, v in out.items()}, "->", OUT)


if __name__ == "__main__":
    main()

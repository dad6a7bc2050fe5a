"""CPU wiring tests of ``sigsvgd_b200.inference.make_fused`` over the UNMODIFIED reference classes
(container only: needs /root/reference; skipped on the GPU box).

INTEGRATION.md routes (A)+(C): the reference's own ``SVGD.optimize`` / ``step`` / ``ScoreEstimator``
run unchanged, only ``_velocity`` is re-routed to the fused update.  There is no GPU here, so the one
CUDA call (``ops.phi``) is replaced by the checker (the float64 closed form of the oracle): what is
tested is the plumbing -- arguments, signs, dictionaries, optimizer state, trace -- not arithmetic.
"""
import numpy as np
import pytest
import torch

import oracle
from oracle import ref_shim
from tests.util import rel_max

pytestmark = pytest.mark.skipif(not ref_shim.reference_available(), reason="reference tree not present (GPU box)")


class _CheckerResult:
    def __init__(self, phi, k_xx, grad_k):
        self.phi, self.k_xx, self.grad_k, self.state = phi, k_xx, grad_k, None


def _checker_phi(X, score, depth=3, bandwidth=None, want_k=False, want_grad_k=False, **_):
    h = None if bandwidth is None or bandwidth.mode == 0 else bandwidth.h
    cf = oracle.phi_closed_form(X.detach().double().numpy(), score.detach().double().numpy(), depth, h=h)
    as_t = lambda a: torch.as_tensor(np.asarray(a)).to(X.dtype)  # noqa: E731
    return _CheckerResult(as_t(cf["phi"]), as_t(cf["K"]) if want_k else None,
                          as_t(cf["grad_k"]) if want_grad_k else None)


@pytest.fixture()
def ref(monkeypatch):
    ref_shim.import_reference()
    import src.inference as ref_inf
    import src.kernels as ref_kern

    from sigsvgd_b200 import inference as fused_mod

    monkeypatch.setattr(fused_mod.ops, "phi", _checker_phi)
    yield ref_inf, ref_kern, fused_mod
    torch.autograd.set_detect_anomaly(False)


def _inputs(N=24, T=8, d=2, seed=3):
    gen = torch.Generator().manual_seed(seed)
    X = torch.cumsum(0.2 * torch.randn(N, T, d, generator=gen, dtype=torch.float64), 1)
    centers = torch.randn(3, T, d, generator=gen, dtype=torch.float64)

    def log_p(x):  # mixture-free toy target: pull towards the nearest of three anchor paths
        d2 = ((x[:, None] - centers[None]) ** 2).sum((-1, -2))
        return -0.5 * d2.min(1).values

    return X, log_p


def test_fused_subclass_of_the_unmodified_reference_svgd_tracks_the_reference(ref):
    ref_inf, ref_kern, fused_mod = ref
    from sigsvgd_b200.kernels import PathSigKernel

    X0, log_p = _inputs()
    Fused = fused_mod.make_fused(ref_inf.SVGD)
    assert issubclass(Fused, ref_inf.SVGD) and Fused.__fused_base__ is ref_inf.SVGD
    # every method but the three re-routed ones is literally the reference's function object
    assert Fused._compute_kernel is ref_inf.SVGD._compute_kernel
    # one step with the median heuristic: identical (which pair attains the median is decided by the
    # last bits of the distance matrix -- SURVEY.md 8(a-4) -- so longer runs of the two arithmetic
    # routes drift apart chaotically; the multi-step comparison below uses a constant bandwidth)
    v_ref, it_ref = ref_inf.SVGD(kernel=ref_kern.PathSigKernel(), log_p=log_p, optimizer_class=None, lr=0.1)._velocity(X0, None)
    v_fus, it_fus = Fused(kernel=PathSigKernel(), log_p=log_p, optimizer_class=None, lr=0.1)._velocity(X0, None)
    assert rel_max(v_fus.numpy(), v_ref.numpy()) < 1e-9 and set(it_fus) == set(it_ref)
    assert rel_max(it_fus["grad_k"].numpy(), it_ref["grad_k"].numpy()) < 1e-9
    from sigsvgd_b200.inference import FixedBandwidth
    from sigsvgd_b200.kernels import GaussianKernel

    runs = {}
    for name, cls, kernel in (
            ("reference", ref_inf.SVGD, ref_kern.PathSigKernel(static_kernel=ref_kern.GaussianKernel(bandwidth_fn=lambda _: 0.8))),
            ("fused", Fused, PathSigKernel(static_kernel=GaussianKernel(bandwidth_fn=FixedBandwidth(0.8))))):
        # (SGD with momentum: Adam's first steps are lr * sign(g), which turns rounding noise on a
        # near-zero gradient component into an O(lr) difference and says nothing about the wiring)
        sampler = cls(kernel=kernel, log_p=log_p, optimizer_class=torch.optim.SGD, lr=0.05, momentum=0.9)
        p = X0.clone()
        data, opt_state = sampler.optimize(p, n_steps=4)
        runs[name] = (p, data, opt_state)
    pr, dr, sr = runs["reference"]
    pf, df, sf = runs["fused"]
    assert rel_max(df["trace"].numpy(), dr["trace"].numpy()) < 1e-8
    assert torch.equal(pf, df["trace"][-1]) and df["trace"].shape == (5,) + tuple(X0.shape)
    for i in range(4):
        assert set(df[i]) == set(dr[i])  # k_xx, grad_k, loss, grad
        assert np.abs(df[i]["k_xx"].numpy() - dr[i]["k_xx"].numpy()).max() < 1e-10
        assert rel_max(df[i]["grad"].numpy(), dr[i]["grad"].numpy()) < 1e-7
        assert float(df[i]["loss"]) == pytest.approx(float(dr[i]["loss"]), rel=1e-9)
    assert rel_max(sf["state"][0]["momentum_buffer"].numpy(), sr["state"][0]["momentum_buffer"].numpy()) < 1e-7


def test_loop_without_host_copies_gives_the_same_particles(ref):
    ref_inf, _, fused_mod = ref
    from sigsvgd_b200.inference import FixedBandwidth
    from sigsvgd_b200.kernels import GaussianKernel, PathSigKernel

    X0, log_p = _inputs(seed=5)
    Fused = fused_mod.make_fused(ref_inf.SVGD)
    out = {}
    for level in ("full", "device", "none"):
        for opt in (torch.optim.SGD, None):
            s = Fused(kernel=PathSigKernel(static_kernel=GaussianKernel(bandwidth_fn=FixedBandwidth(0.8))), log_p=log_p,
                      optimizer_class=opt, lr=0.05, adaptive_gradient=opt is None)
            s.log_level, s.use_graph = level, False
            p = X0.clone()
            data, _ = s.optimize(p, n_steps=3)
            out[(level, opt)] = (p, data)
    for opt in (torch.optim.SGD, None):
        ref_p, ref_d = out[("full", opt)]
        for level in ("device", "none"):
            p, d = out[(level, opt)]
            assert rel_max(p.numpy(), ref_p.numpy()) < 1e-12
            assert (d[0]["k_xx"] is None) == (level == "none")
            if level == "device":
                assert rel_max(d["trace"].numpy(), ref_d["trace"].numpy()) < 1e-12
            else:
                assert d["trace"] is None


def test_unfusable_cases_fall_through_to_the_reference(ref):
    ref_inf, ref_kern, fused_mod = ref
    X0, log_p = _inputs(N=12)
    Fused = fused_mod.make_fused(ref_inf.SVGD)
    # the reference's own kernel object: not this package's PathSigKernel -> base class path
    s = Fused(kernel=ref_kern.PathSigKernel(), log_p=log_p, optimizer_class=None, lr=0.1)
    v1, _ = s._velocity(X0, None)
    v2, _ = ref_inf.SVGD(kernel=ref_kern.PathSigKernel(), log_p=log_p, optimizer_class=None, lr=0.1)._velocity(X0, None)
    assert torch.equal(v1, v2)
    # caller-supplied kernel terms (svgd.py:54-58) and the missing-target error come from the base class
    from sigsvgd_b200.kernels import PathSigKernel

    s = Fused(kernel=PathSigKernel(), optimizer_class=None, lr=0.1)
    K, gk, G = torch.eye(12, dtype=torch.float64), torch.zeros(12, 16, dtype=torch.float64), torch.ones(12, 16, dtype=torch.float64)
    v, it = s._velocity(X0, G, k_xx=K, grad_k=gk)
    assert torch.allclose(v.flatten(1), -G / 12)
    with pytest.raises(ValueError, match="log probability"):
        s._velocity(X0, None)
    # TrajectorySVGD re-defines the kernel terms: never short-cut
    TF = fused_mod.make_fused(ref_inf.TrajectorySVGD)
    assert TF._compute_kernel is ref_inf.TrajectorySVGD._compute_kernel


def test_install_swaps_the_class_in_place(ref):
    ref_inf, _, fused_mod = ref
    base = ref_inf.SVGD
    try:
        done = fused_mod.install(ref_inf)
        assert done["SVGD"].__fused_base__ is base and ref_inf.SVGD is done["SVGD"]
        assert fused_mod.install(ref_inf) == {}  # idempotent
    finally:
        ref_inf.SVGD = base
        import sys

        sub = sys.modules.get("src.inference.svgd")
        if sub is not None:
            sub.SVGD = base


def test_mpc_golden_is_reproducible_from_the_unmodified_controller():
    """One control step of the harness that produced tests/golden/mpc_maze_35x30x2.npz (the unmodified
    DuSt + ParticleModel + particle_maze_config.yaml through the SigKernel branch)."""
    from oracle import make_golden_mpc
    from tests.util import load_golden

    torch.manual_seed(0)
    np.random.seed(0)
    record = {"tau": [], "k_xx": [], "dk_dtau": []}
    controller, simulator, system, state, prior, exp_hp = make_golden_mpc.build_controller(record)
    torch.autograd.set_detect_anomaly(False)
    a_seq, data = controller.forward(state, simulator, prior, opt_steps=exp_hp["opt_steps"])
    g = load_golden("mpc_maze_35x30x2")
    assert data[0]["k_xx"].shape == (35, 35) and data[0]["grad_k"].shape == (35, 60)
    assert np.abs(record["k_xx"][0] - g["k_xx"][0]).max() < 1e-9
    assert np.abs(record["tau"][1] - g["tau"][1]).max() < 1e-5
    assert rel_max(data[1]["grad_k"].numpy(), g["grad_k"][1]) < 1e-4
    assert rel_max(a_seq.numpy(), g["a_seq"][0]) < 1e-4


def test_unmodified_reference_signature_kernel_over_the_oracle_solver():
    """src/kernels/_traj_kernels.py:198-206 ``SignatureKernel`` UNMODIFIED, with a ``sigkernel`` stand-in whose
    ``compute_Gram`` = the reference's own ``static_kernel.Gram_matrix`` + the oracle's Goursat recursion:
    what ``sigsvgd_b200.kernels.SignatureKernel`` must reproduce (checked on the GPU against the same
    oracle functions in tests/test_gpu_sigpde.py)."""
    import sys
    import types

    from oracle import sigpde_oracle as so

    sk = types.ModuleType("sigkernel")

    class SigKernel:
        def __init__(self, static_kernel, dyadic_order=0):
            self.static_kernel, self.dyadic_order = static_kernel, dyadic_order

        def compute_Gram(self, X, Y, sym=False):
            G = self.static_kernel.Gram_matrix(X, Y).numpy()
            return torch.as_tensor(so._solve(so.increments(G), self.dyadic_order)[..., -1, -1])

    sk.SigKernel, sk.RBFKernel = SigKernel, object
    ref_shim.import_reference(sigkernel_module=sk)
    import src.kernels._traj_kernels as tk

    old = tk.sigkernel
    tk.sigkernel = sk
    try:
        X, _ = _inputs(N=6, T=7, d=2, seed=11)
        X = X.float()
        k_ref = tk.SignatureKernel(depth=2)(X, X)  # median heuristic, dyadic_order = depth
        assert k_ref.dtype == torch.float32 and k_ref.shape == (6, 6)
        X64 = X.double().numpy()
        h, _ = so.bw_median_points(so.static_gram(X64, X64, so.RBF, 1.0)[1])
        assert rel_max(k_ref.numpy(), so.sigpde_gram(X64, X64, 2, so.RBF, h)) < 1e-6
    finally:
        tk.sigkernel = old


def test_rollout_with_reads_the_unmodified_particle_model(monkeypatch):
    """``rollout_with(model)`` over the reference's own ParticleModel + ObstacleMap objects, against the
    unmodified ``DuSt._rollout`` (dust.py:250-309); the CUDA launch is replaced by the numpy checker."""
    from oracle import make_golden_mpc
    from oracle import rollout_oracle as ro
    from sigsvgd_b200 import rollout as rmod

    torch.manual_seed(0)
    controller, simulator, _, state, prior, _ = make_golden_mpc.build_controller({"tau": [], "k_xx": [], "dk_dtau": []})
    torch.autograd.set_detect_anomaly(False)

    def checker(init_state, actions, mass=1.0, dt=0.05, max_accel=float("inf"), max_speed=float("inf"), grid=None,
                cell_size=1.0, c_offset=(0, 0), can_crash=False):
        m = mass.numpy() if isinstance(mass, torch.Tensor) else mass
        s, _ = ro.particle_rollout_np(init_state.numpy(), actions.detach().numpy(), m, dt, max_accel, max_speed,
                                      None if grid is None else grid.numpy(), cell_size, c_offset, can_crash)
        return torch.as_tensor(s)

    monkeypatch.setattr(rmod, "particle_rollout", checker)
    base_actions = controller._sample_actions().detach()
    ref_states, ref_actions, params = controller._rollout(state, base_actions, simulator, prior)
    got = rmod.rollout_with(simulator)(state, base_actions.flatten(end_dim=-3), params)
    assert torch.equal(got.reshape(ref_states.shape), ref_states)

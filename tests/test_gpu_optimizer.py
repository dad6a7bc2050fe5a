"""GPU tests of the fused optimiser step (``sigsvgd_opt_step``) through the C ABI:
against ``torch.optim.Adam`` / ``torch.optim.SGD`` and the reference's hand-written Adagrad
(src/inference/svgd.py:93-116), and inside the fused sampler's loop."""
import numpy as np
import pytest
import torch

from tests.util import make_paths, rel_max

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    from sigsvgd_b200 import _lib, ops
    from sigsvgd_b200.inference import FixedBandwidth, _fused_optimizer_step, make_fused
    from sigsvgd_b200.kernels import GaussianKernel, PathSigKernel
    from tests import ref_mirror

DEV = "cuda:0"
TOL = 2e-6  # relative, float32 update rules evaluated in a different association


def _grads(shape, steps, seed=0):
    g = torch.Generator().manual_seed(seed)
    return [torch.randn(*shape, generator=g).to(DEV) * (0.1 + i) for i in range(steps)]


@pytest.mark.parametrize("cls,kw", [("Adam", dict(lr=0.05)), ("Adam", dict(lr=1.0, betas=(0.8, 0.95), eps=1e-6)),
                                    ("SGD", dict(lr=0.1)), ("SGD", dict(lr=0.05, momentum=0.9))])
def test_fused_step_matches_torch_optimizer_and_keeps_its_state_layout(cls, kw):
    shape, steps = (130, 9, 7), 6
    x0 = torch.randn(*shape, generator=torch.Generator().manual_seed(1)).to(DEV)
    xr, xf = x0.clone(), x0.clone()
    opt_r = getattr(torch.optim, cls)(params=[xr], **kw)
    opt_f = getattr(torch.optim, cls)(params=[xf], **kw)
    for g in _grads(shape, steps):
        xr.grad = g.clone()
        opt_r.step()
        assert _fused_optimizer_step(opt_f, xf, g)
    assert rel_max(xf.cpu().numpy(), xr.cpu().numpy()) <= TOL
    sr, sf = opt_r.state_dict()["state"].get(0, {}), opt_f.state_dict()["state"].get(0, {})
    assert set(sr) == set(sf)
    for k in sr:
        a, b = (torch.as_tensor(v).float().cpu().numpy() for v in (sf[k], sr[k]))
        assert rel_max(a, b) <= TOL, k
    # the fused state loads into a stock optimizer, which carries on from it
    opt_c = getattr(torch.optim, cls)(params=[xf], **kw)
    opt_c.load_state_dict(opt_f.state_dict())
    g = _grads(shape, 1, seed=9)[0]
    xr.grad, xf.grad = g.clone(), g.clone()
    opt_r.step()
    opt_c.step()
    assert rel_max(xf.cpu().numpy(), xr.cpu().numpy()) <= TOL


def test_unsupported_optimizers_are_left_to_torch():
    x = torch.zeros(4, 3, 2, device=DEV)
    g = torch.ones_like(x)
    assert not _fused_optimizer_step(torch.optim.Adam([x], lr=0.1, amsgrad=True), x, g)
    assert not _fused_optimizer_step(torch.optim.Adam([x], lr=0.1, weight_decay=0.1), x, g)
    assert not _fused_optimizer_step(torch.optim.SGD([x], lr=0.1, momentum=0.9, nesterov=True), x, g)
    assert not _fused_optimizer_step(torch.optim.AdamW([x], lr=0.1), x, g)
    assert not _fused_optimizer_step(torch.optim.Adam([x, torch.zeros(2, device=DEV)], lr=0.1), x, g)
    assert float(x.abs().max()) == 0.0
    with pytest.raises(ValueError, match="Adam needs"):
        ops.opt_step_raw(x, g, _lib.OPT_ADAM, 0.1)
    with pytest.raises(ValueError, match="unknown optimiser"):
        ops.opt_step_raw(x, g, 7, 0.1)


def test_adagrad_rule_of_the_reference():
    shape = (50, 8, 3)
    x = torch.randn(*shape, generator=torch.Generator().manual_seed(2)).to(DEV)
    ref, inertia_ref = x.double().cpu().numpy(), np.zeros(shape)
    inertia, logged = torch.zeros_like(x), torch.empty_like(x)
    for phi in _grads(shape, 5, seed=3):
        ops.opt_step_raw(x, phi, _lib.OPT_ADAGRAD, 0.3, -1.0, None, inertia, logged)  # grad = -phi
        g = -phi.double().cpu().numpy()
        inertia_ref += g ** 2  # svgd.py:112-115
        gs = g / np.sqrt(inertia_ref + 1e-12)
        ref = ref - 0.3 * gs
        assert rel_max(logged.cpu().numpy(), gs) <= TOL
    assert rel_max(x.cpu().numpy(), ref) <= TOL and rel_max(inertia.cpu().numpy(), inertia_ref) <= TOL


@pytest.mark.parametrize("opt", ["Adam", "SGD", None])
def test_sampler_loop_with_and_without_the_fused_step(opt):
    N, T, d = 256, 16, 3
    X0 = torch.as_tensor(make_paths("walk", N, T, d, seed=5), device=DEV)
    centers = torch.randn(3, T, d, generator=torch.Generator().manual_seed(6)).to(DEV)

    def log_p(x):
        return -0.5 * ((x[:, None] - centers[None]) ** 2).sum((-1, -2)).min(1).values

    Fused = make_fused(ref_mirror.SVGD)
    out = {}
    for fused in (True, False):
        s = Fused(kernel=PathSigKernel(static_kernel=GaussianKernel(bandwidth_fn=FixedBandwidth(0.8))), log_p=log_p,
                  optimizer_class=getattr(torch.optim, opt) if opt else None, lr=0.05, adaptive_gradient=opt is None)
        s.log_level, s.fused_optimizer = "none", fused
        p = X0.clone()
        data, state = s.optimize(p, n_steps=4)
        out[fused] = (p, data, state)
    assert rel_max(out[True][0].cpu().numpy(), out[False][0].cpu().numpy()) <= 1e-5
    assert rel_max(out[True][1][3]["grad"].cpu().numpy(), out[False][1][3]["grad"].cpu().numpy()) <= 1e-4
    if opt:
        assert set(out[True][2]["state"].get(0, {})) == set(out[False][2]["state"].get(0, {}))

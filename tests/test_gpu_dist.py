"""GPU tests of the row-block sharded update (sigsvgd_b200/dist.py) with the CUDA
backend: world_size 1 in-process (both kernel families), and world_size 2 over NCCL when the box
has two GPUs (spawned ranks; skipped otherwise).  Checked against the float64 closed form with
the tie pair(s) the implementation reports (parity protocol of SURVEY.md 8(a-4))."""
import os
import socket

import numpy as np
import pytest
import torch

from tests.util import TOL_H, TOL_K, TOL_PHI, closed_form_with_ties, make_paths, rel_max

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    from sigsvgd_b200 import ops
    from sigsvgd_b200.dist import CudaBackend, phi_sharded


def _inputs(N, T, d, seed=0):
    X = make_paths("walk", N, T, d, seed)
    G = np.random.default_rng(seed + 1).standard_normal((N, T * d)).astype(np.float32)
    return X, G


def _check(X, G, depth, phi, K, recs):
    ties = [t for r in recs for t in r["ties"]]
    cf = closed_form_with_ties(X.astype(np.float64), G.astype(np.float64), depth, ties)
    for r in recs:
        assert abs(r["h"] - cf["h"]) <= TOL_H * cf["h"]
        assert r["n_ties_all"] == len(ties)
    assert np.abs(K - cf["K"]).max() <= TOL_K
    assert rel_max(phi, cf["phi"]) <= TOL_PHI


@pytest.mark.parametrize("tc", [False, True])
@pytest.mark.parametrize("N,T,d,depth", [(96, 16, 3, 3), (512, 64, 7, 3)])
def test_world1_matches_closed_form(tc, N, T, d, depth):
    X, G = _inputs(N, T, d)
    dev = torch.device("cuda:0")
    res = phi_sharded(torch.as_tensor(X, device=dev), torch.as_tensor(G, device=dev), depth=depth, want_k=True,
                      backend=CudaBackend(tc=tc))
    _check(X, G, depth, res.phi.cpu().numpy(), res.k_rows.cpu().numpy(), [ops.read_state(res.state)])


@pytest.mark.parametrize("N", [1100, 3000])
def test_world1_windowed_median_is_exact(N):
    """The windowed first pass (count below + compact the window, then select among the candidates)
    must land on the same exact order statistic as the three full passes."""
    T, d, depth = 16, 3, 3
    X, G = _inputs(N, T, d, seed=9)
    dev = torch.device("cuda:0")
    x, g = torch.as_tensor(X, device=dev), torch.as_tensor(G, device=dev)
    a = phi_sharded(x, g, depth=depth, want_k=True, window=True)
    b = phi_sharded(x, g, depth=depth, want_k=True, window=False)
    ra, rb = ops.read_state(a.state), ops.read_state(b.state)
    assert abs(ra["median"] - rb["median"]) <= 1e-6 * rb["median"]
    _check(X, G, depth, a.phi.cpu().numpy(), a.k_rows.cpu().numpy(), [ra])
    # the collect pass itself: counts and candidates against numpy on the same strip
    be = CudaBackend()
    s = ops.signature_fwd_raw(x, depth, True)
    feat = be.pack(s, g, True, 0, N)
    dmat = be.gram(feat)
    st = be.new_state(dev)
    bounds = be.select_window(feat.sc_all, st)
    cand, totals = be.select_collect(dmat, bounds)
    lo, hi = (float(v) for v in bounds.cpu())
    dn = dmat.cpu().numpy().reshape(-1)
    below, kept, over = (int(v) for v in totals.cpu())
    assert over == 0 and below == int((dn < lo).sum()) and kept == int(((dn >= lo) & (dn <= hi)).sum())
    cn = np.sort(cand.cpu().numpy())
    assert np.array_equal(cn[:kept], np.sort(dn[(dn >= lo) & (dn <= hi)])) and (cn[kept:] > hi).all()
    k = (N * N - 1) // 2
    assert below <= k < below + kept, "sampled window missed the median"


def test_world1_default_backend_agrees_with_fused_update():
    N, T, d, depth = 1536, 64, 7, 3  # >= 1024: both take the tcgen05 kernels
    X, G = _inputs(N, T, d, seed=3)
    dev = torch.device("cuda:0")
    x, g = torch.as_tensor(X, device=dev), torch.as_tensor(G, device=dev)
    a = phi_sharded(x, g, depth=depth)
    b = ops.phi(x, g, depth=depth)
    ra, rb = ops.read_state(a.state), ops.read_state(b.state)
    assert abs(ra["h"] - rb["h"]) <= TOL_H * rb["h"]
    # rows that carry a median-gradient push may differ (different tie pairs): compare the others
    skip = {t[0] for t in ra["ties"]} | {t[0] for t in rb["ties"]}
    keep = np.array([i for i in range(N) if i not in skip])
    pa, pb = a.phi.cpu().numpy()[keep], b.phi.cpu().numpy()[keep]
    assert rel_max(pa, pb) <= TOL_PHI


def test_full_size_two_routes_agree():
    """N = 16384 (BASELINE configs[3] range; too large for the CPU oracle in a test): the fused
    update and the strip pipeline are independent routes -- symmetric vs non-symmetric Gram,
    single-pass window median vs sampled window + radix select -- and must give the same median,
    bandwidth and phi rows."""
    N, T, d, depth = 16384, 64, 7, 3
    g = torch.Generator().manual_seed(11)
    dev = torch.device("cuda:0")
    x = torch.cumsum(0.05 * torch.randn(N, T, d, generator=g), dim=1).to(dev)
    s = torch.randn(N, T * d, generator=g).to(dev)
    a = ops.phi(x, s, depth=depth)
    b = phi_sharded(x, s, depth=depth)
    ra, rb = ops.read_state(a.state), ops.read_state(b.state)
    assert abs(ra["median"] - rb["median"]) <= 1e-6 * rb["median"]
    assert abs(ra["h"] - rb["h"]) <= 1e-6 * rb["h"]
    keep = torch.ones(N, dtype=torch.bool, device=dev)
    keep[sorted({t[0] for t in ra["ties"]} | {t[0] for t in rb["ties"]})] = False
    diff = (a.phi[keep] - b.phi[keep]).abs().max().item() / b.phi[keep].abs().max().item()
    assert diff <= 1e-5
    assert bool(torch.isfinite(a.phi).all()) and bool(torch.isfinite(b.phi).all())
    ops._WS_CACHE.clear()
    torch.cuda.empty_cache()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _nccl_worker(rank, world, port, N, T, d, depth, tc, q):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        X, G = _inputs(N, T, d)
        M = N // world
        res = phi_sharded(torch.as_tensor(X[rank * M:(rank + 1) * M], device=dev),
                          torch.as_tensor(G[rank * M:(rank + 1) * M], device=dev), depth=depth, want_k=True,
                          backend=CudaBackend(tc=tc), window=tc)
        torch.cuda.synchronize()
        out = [rank, res.phi.cpu().numpy(), res.k_rows.cpu().numpy(), ops.read_state(res.state), None]
        if tc:
            # the same step (kernels + NCCL collectives) replayed from one CUDA graph
            from sigsvgd_b200.dist import ShardedPlan

            plan = ShardedPlan(M, T, d, depth)
            for _ in range(2):
                rg = plan(torch.as_tensor(X[rank * M:(rank + 1) * M], device=dev),
                          torch.as_tensor(G[rank * M:(rank + 1) * M], device=dev))
            torch.cuda.synchronize()
            out[4] = (rg.phi.cpu().numpy(), ops.read_state(rg.state))
            del plan, rg  # (a live graph holding NCCL work must go before the process group does)
        q.put(tuple(out))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("tc,N", [(False, 256), (True, 2048)])
def test_world2_nccl(tc, N):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp

    T, d, depth, world = 64, 7, 3, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_nccl_worker, args=(r, world, port, N, T, d, depth, tc, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = sorted((q.get(timeout=300) for _ in range(world)), key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    X, G = _inputs(N, T, d)
    recs = [g[3] for g in got]
    assert recs[0]["median"] == recs[1]["median"] and recs[0]["h"] == recs[1]["h"]
    _check(X, G, depth, np.concatenate([g[1] for g in got]), np.concatenate([g[2] for g in got]), recs)
    if tc:
        grecs = [g[4][1] for g in got]
        # (two runs of the step: the column statistics are accumulated with floating-point atomics, so the
        # centred features -- and with them the entry that is the median -- can differ in the last bits)
        assert grecs[0]["median"] == grecs[1]["median"] and grecs[0]["h"] == grecs[1]["h"]
        assert abs(grecs[0]["median"] - recs[0]["median"]) <= 1e-6 * recs[0]["median"]
        assert abs(grecs[1]["h"] - recs[1]["h"]) <= 1e-6 * recs[1]["h"]
        skip = {t[0] for r in recs + grecs for t in r["ties"]}
        keep = np.array([i for i in range(N) if i not in skip])
        phi_e = np.concatenate([g[1] for g in got]).reshape(N, -1)[keep]
        phi_g = np.concatenate([g[4][0] for g in got]).reshape(N, -1)[keep]
        assert rel_max(phi_g, phi_e) <= 1e-5


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_calls_follow_the_device_of_their_tensors():
    """The reference's multi-GPU idiom is ``device=f"cuda:{i}"`` with no ``set_device``
    (examples/script_planning_robot.py:486): tensors on cuda:1 while cuda:0 is the current device."""
    from sigsvgd_b200.kernels import PathSigKernel

    assert torch.cuda.current_device() == 0
    X, G = _inputs(300, 16, 3, seed=2)
    d0, d1 = torch.device("cuda:0"), torch.device("cuda:1")
    r0 = ops.phi(torch.as_tensor(X, device=d0), torch.as_tensor(G, device=d0), want_k=True)
    r1 = ops.phi(torch.as_tensor(X, device=d1), torch.as_tensor(G, device=d1), want_k=True)
    assert r1.phi.device == d1 and torch.cuda.current_device() == 0
    # (same arithmetic on both devices; the column statistics are accumulated with floating-point
    # atomics, so the last bits may differ from launch to launch)
    assert rel_max(r1.phi.cpu().numpy(), r0.phi.cpu().numpy()) < 1e-5
    assert float((r0.k_xx.cpu() - r1.k_xx.cpu()).abs().max()) < 1e-6
    x1 = torch.as_tensor(X, device=d1).requires_grad_(True)
    k = PathSigKernel()(x1, x1.detach(), compute_grad=False)
    gk = torch.autograd.grad(k.sum(), x1)[0]
    assert gk.device == d1 and torch.isfinite(gk).all()
    with pytest.raises(RuntimeError, match="different devices"):
        ops.phi(torch.as_tensor(X, device=d0), torch.as_tensor(G, device=d1))
    r0b = ops.phi(torch.as_tensor(X, device=d0), torch.as_tensor(G, device=d0))  # the guard left no stale state
    assert rel_max(r0b.phi.cpu().numpy(), r0.phi.cpu().numpy()) < 1e-5

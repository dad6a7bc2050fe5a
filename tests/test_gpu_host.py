"""GPU tests of the host-input entry point sigsvgd_phi_host (H2D of X, score copy beside the front
half, chunked D2H of phi): same numbers as the device-resident sigsvgd_phi, on both kernel
families, and stable under back-to-back reuse of the workspace and the internal copy streams."""
import numpy as np
import pytest
import torch

from tests.util import TOL_H, TOL_PHI, closed_form_with_ties, make_paths, rel_max

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    from sigsvgd_b200 import ops


@pytest.mark.parametrize("N,T,d", [(300, 16, 3), (2048, 64, 7), (1100, 10, 2)])
def test_phi_host_matches_device_path_and_oracle(N, T, d):
    depth = 3
    X = make_paths("walk", N, T, d, seed=5)
    G = np.random.default_rng(6).standard_normal((N, T * d)).astype(np.float32)
    Xh, Gh = torch.from_numpy(X).pin_memory(), torch.from_numpy(G).pin_memory()
    out = torch.empty_like(Xh).pin_memory()
    dev = torch.device("cuda:0")
    ref = ops.phi(Xh.to(dev), Gh.to(dev), depth=depth)
    rr = ops.read_state(ref.state)
    for rep in range(3):  # back-to-back calls reuse the workspace, streams and events
        out.fill_(float("nan"))
        res = ops.phi_host(Xh, Gh, depth=depth, out=out)
        torch.cuda.current_stream().synchronize()
        rec = ops.read_state(res.state)
        # not bit-equal: the column means are accumulated with float atomics (order varies per launch)
        assert abs(rec["median"] - rr["median"]) <= 1e-6 * rr["median"] and abs(rec["h"] - rr["h"]) <= 1e-6 * rr["h"]
        skip = {t[0] for t in rec["ties"]} | {t[0] for t in rr["ties"]}
        keep = np.array([i for i in range(N) if i not in skip])
        assert rel_max(out.numpy()[keep], ref.phi.cpu().numpy()[keep]) <= 1e-5
    cf = closed_form_with_ties(X.astype(np.float64), G.astype(np.float64), depth, rec["ties"])
    assert abs(rec["h"] - cf["h"]) <= TOL_H * cf["h"]
    assert rel_max(out.numpy(), cf["phi"]) <= TOL_PHI


def test_phi_host_rejects_device_tensors():
    x = torch.zeros(8, 4, 2, device="cuda:0")
    with pytest.raises(RuntimeError):
        ops.phi_host(x, torch.zeros(8, 8, device="cuda:0"))


def test_phi_plan_graph_replay_matches_direct_call():
    """ops.PhiPlan (sigsvgd_phi captured into a CUDA graph) must track new inputs on every replay and
    agree with the call-by-call update, for the tcgen05 path and the SIMT path."""
    for N, T, d in [(512, 64, 7), (96, 16, 3)]:
        plan = ops.PhiPlan(N, T, d, 3, device="cuda:0")
        for seed in (1, 2, 3):
            g = torch.Generator().manual_seed(seed)
            x = torch.cumsum(0.05 * torch.randn(N, T, d, generator=g), dim=1).cuda()
            s = torch.randn(N, T * d, generator=g).cuda()
            want = ops.phi(x, s, depth=3)
            got = plan(x, s)
            torch.cuda.synchronize()
            ra, rb = ops.read_state(got.state), ops.read_state(want.state)
            assert abs(ra["h"] - rb["h"]) <= 1e-6 * rb["h"]
            assert ra["n_ties_all"] == rb["n_ties_all"]
            skip = {t[0] for t in ra["ties"]} | {t[0] for t in rb["ties"]}
            keep = torch.ones(N, dtype=torch.bool, device="cuda:0")
            keep[sorted(skip)] = False
            den = want.phi[keep].abs().max().item()
            assert (got.phi[keep] - want.phi[keep]).abs().max().item() <= 1e-5 * den
        # in-place use: write into the plan's buffers, replay without arguments
        plan.x.mul_(1.5)
        got = plan()
        want = ops.phi(plan.x.clone(), plan.score.clone(), depth=3)
        torch.cuda.synchronize()
        assert abs(ops.read_state(got.state)["h"] - ops.read_state(want.state)["h"]) <= 1e-6 * ops.read_state(want.state)["h"]


def test_phi_plan_batch_of_scenes():
    """A plan over a batch of independent swarms (B > 1: per-scene medians, SIMT kernels)."""
    B, N, T, d = 3, 48, 12, 2
    g = torch.Generator().manual_seed(7)
    x = torch.cumsum(0.1 * torch.randn(B, N, T, d, generator=g), dim=2).cuda()
    s = torch.randn(B, N, T * d, generator=g).cuda()
    plan = ops.PhiPlan(N, T, d, 3, device="cuda:0", batch=B)
    got = plan(x, s)
    want = ops.phi(x, s, depth=3)
    torch.cuda.synchronize()
    for b in range(B):
        ra, rb = ops.read_state(got.state, b), ops.read_state(want.state, b)
        assert abs(ra["h"] - rb["h"]) <= 1e-6 * rb["h"]
        skip = {t[0] for t in ra["ties"]} | {t[0] for t in rb["ties"]}
        keep = torch.ones(N, dtype=torch.bool, device="cuda:0")
        keep[sorted(skip)] = False
        den = want.phi[b][keep].abs().max().item()
        assert (got.phi[b][keep] - want.phi[b][keep]).abs().max().item() <= 1e-5 * den


def test_two_host_threads_on_one_device():
    """sigsvgd_phi shares one side stream and event set per device; the library serialises the enqueue
    phase per device (include/sigsvgd_b200.h, "host threads"), so concurrent callers on their own streams
    get the results of sequential calls."""
    import threading

    N, T, d = 512, 16, 3
    gen = torch.Generator().manual_seed(21)
    inputs = [(torch.cumsum(0.1 * torch.randn(N, T, d, generator=gen), 1).to("cuda:0"),
               torch.randn(N, T * d, generator=gen).to("cuda:0")) for _ in range(2)]
    want = [ops.phi(x, g).phi.clone() for x, g in inputs]
    torch.cuda.synchronize()
    got, errs = [None, None], []

    def work(i):
        try:
            st = torch.cuda.Stream()
            with torch.cuda.stream(st):
                for _ in range(25):
                    r = ops.phi(*inputs[i])
                got[i] = r.phi.clone()
            st.synchronize()
        except Exception as e:  # noqa: BLE001
            errs.append(e)

    threads = [threading.Thread(target=work, args=(i,)) for i in range(2)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errs, errs
    for i in range(2):
        assert float((got[i] - want[i]).abs().max() / want[i].abs().max()) < 1e-5

"""Per-kernel CUDA-event timings of the phi-update pieces (development aid; run under gpurun)."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sigsvgd_b200 import _lib, ops

N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
T, d, depth = 64, 7, 3
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
X = torch.cumsum(0.05 * torch.randn(N, T, d, generator=g), dim=1).to(dev)
G = torch.randn(N, T * d, generator=g).to(dev)
flush = torch.empty(64 * 1024 * 1024, dtype=torch.float32, device=dev)


def timeit(fn, n=20, warm=3, do_flush=True):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        if do_flush:
            flush.fill_(1.0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


D = ops.sig_channels(d, depth)
P = T * d
res = {}
sig = ops.signature_fwd_raw(X, depth, True)
res["sig_fwd"] = timeit(lambda: ops.signature_fwd_raw(X, depth, True))
mean = ops.col_mean(sig)
res["col_mean"] = timeit(lambda: ops.col_mean(sig))
planes, norms, scales, sc = ops.tc_pack_raw(sig, mean, G)
res["tc_pack(+scales)"] = timeit(lambda: ops.tc_pack_raw(sig, mean, G))
dist = ops.dist_buffer(N, N, dev)
state = ops.new_state(dev)
ws = ops.tc_median_workspace(N, N, dev)
ops.tc_median_prepare_raw(sc, sc, ws, state)
res["median_prepare"] = timeit(lambda: ops.tc_median_prepare_raw(sig, sig, ws, state))
res["tc_gram_sym"] = timeit(lambda: ops.tc_gram_raw(planes, norms, planes, norms, D, scales, True, dist))
res["tc_gram_sym_noflush"] = timeit(lambda: ops.tc_gram_raw(planes, norms, planes, norms, D, scales, True, dist), do_flush=False)
res["tc_gram_nonsym"] = timeit(lambda: ops.tc_gram_raw(planes, norms, planes, norms, D, scales, False, dist))
ops.tc_gram_raw(planes, norms, planes, norms, D, scales, True, dist)
def _sel():
    ops.tc_median_prepare_raw(sig, sig, ws, state)
    ops.tc_median_select_raw(ws, dist, N, ops.Bandwidth.median(), state, True)
res["prepare+select_sym"] = timeit(_sel)
_sel()
print("median", ops.read_state(state)["median"], "h", ops.read_state(state)["h"], file=sys.stderr)
out = ops.tc_kf_raw(dist, planes, D + 1 + P, scales, state)
res["tc_kf_fused"] = timeit(lambda: ops.tc_kf_raw(dist, planes, D + 1 + P, scales, state))
res["tc_kf_fused_noflush"] = timeit(lambda: ops.tc_kf_raw(dist, planes, D + 1 + P, scales, state), do_flush=False)
res["tc_kf_fused_nostats"] = timeit(lambda: ops.tc_kf_raw(dist, planes, D + 1 + P, scales, state, stats=False))
kmat = ops.rbf_out_raw(dist, state, ops.dist_buffer(N, N, dev))
res["rbf_out"] = timeit(lambda: ops.rbf_out_raw(dist, state, kmat))
res["tc_kf_identity"] = timeit(lambda: ops.tc_kf_raw(kmat, planes, D + 1 + P, scales, None, _lib.KF_IDENTITY))
gs = ops.sig_grad_raw(sc, sc, out[:, :D], out[:, D], state)
res["sig_grad"] = timeit(lambda: ops.sig_grad_raw(sc, sc, out[:, :D], out[:, D], state))
res["sig_bwd"] = timeit(lambda: ops.signature_bwd_raw(X, sig, gs, depth, True, add=out[:, D + 1:], alpha=1.0 / N, beta=-1.0 / N))
res["phi_total"] = timeit(lambda: ops.phi(X, G, depth=depth))
res["phi_total_noflush"] = timeit(lambda: ops.phi(X, G, depth=depth), do_flush=False)
res["phi_total_want_k"] = timeit(lambda: ops.phi(X, G, depth=depth, want_k=True))
plan = ops.PhiPlan(N, T, d, depth, device=dev)
plan.x.copy_(X); plan.score.copy_(G)
res["phi_plan"] = timeit(lambda: plan())
fl_gram, fl_kv = 2.0 * N * N * D, 2.0 * N * N * (D + P)
res["TFLOPs_alg_gram"] = fl_gram / res["tc_gram_sym"] / 1e6
res["TFLOPs_alg_kf"] = fl_kv / res["tc_kf_fused"] / 1e6
print(json.dumps({"N": N, "us": res}, indent=1))

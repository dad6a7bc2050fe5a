"""Per-phase timings of ONE rank's work in the row-block sharded update, without the collectives
(development aid; run under gpurun on one GPU):  python tools/strip_pieces.py N world"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sigsvgd_b200 import _lib, ops
from sigsvgd_b200.dist import CudaBackend

N = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
world = int(sys.argv[2]) if len(sys.argv) > 2 else 8
M = N // world
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
X = torch.cumsum(0.05 * torch.randn(N, 64, 7, generator=g), dim=1).to(dev)
G = torch.randn(N, 448, generator=g).to(dev)
be = CudaBackend()


def t(fn, n=3):
    out = fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): out = fn()
    e1.record(); torch.cuda.synchronize()
    return out, e0.elapsed_time(e1) / n


res = {}
s_loc, res["sig_fwd_rows"] = t(lambda: be.sig_fwd(X[:M], 3))
s_all = be.sig_fwd(X, 3)
allp = torch.cat([s_all, G], dim=1)
s_v, g_v = allp[:, :399], allp[:, 399:]
_, res["cat_local"] = t(lambda: torch.cat([s_loc, G[:M]], dim=1))
feat, res["pack_all"] = t(lambda: be.pack(s_v, g_v, True, 0, M))
k_rows, res["gram_strip"] = t(lambda: be.gram(feat))
state = be.new_state(dev); hist = be.new_hist(dev)
be.select_begin(state, hist, (N * N - 1) // 2)
for p in range(3):
    _, res[f"hist_pass{p}"] = t(lambda: be.select_hist(k_rows, p, state, hist), n=1)
    hist.zero_(); be.select_hist(k_rows, p, state, hist)
    # emulate the all-reduce over `world` identical strips
    hist.mul_(world)
    be.select_scan(p, state, hist)
be.select_finish(N, 1.0, 1e-8, state)
bounds, res["select_window"] = t(lambda: be.select_window(feat.sc_all, be.new_state(dev)))
(cand, totals), res["select_collect"] = t(lambda: be.select_collect(k_rows, bounds))
res["collect_totals"] = [int(v) for v in totals.cpu()]
res["collect_cap"] = cand.numel()
_, res["hist_pass_on_cand"] = t(lambda: be.select_hist(cand, 1, state, hist), n=1)
d_copy = k_rows.clone()
def rbf():
    k_rows.copy_(d_copy)
    return be.rbf(k_rows, state, 0, feat, False)
_, res["copy_strip"] = t(lambda: k_rows.copy_(d_copy))
kmat, res["rbf_strip(+copy)"] = t(rbf)
stats = be.local_stats(state)
be.median_coef(state, True, stats)
(ks, rowsum, kg), res["kf_strip"] = t(lambda: be.contract(kmat, feat))
gg, res["sig_grad"] = t(lambda: be.sig_grad(feat.sc_loc, feat.sc_all, ks, rowsum, state, 0))
_, res["sig_bwd_rows"] = t(lambda: be.sig_bwd(X[:M], s_loc, gg, 3, kg, 1.0 / N, -1.0 / N))
res["median"] = ops.read_state(state)["median"]
print(json.dumps({"N": N, "world": world, "M": M, "ms": res}, indent=1))

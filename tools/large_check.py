"""Full-size cross-check (development aid; run under gpurun): the fused single-GPU update
(window median, symmetric Gram) against the strip pipeline of dist.phi_sharded at world 1
(non-symmetric Gram, 3-pass u64 histogram median) -- two independent routes to the same numbers."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sigsvgd_b200 import ops
from sigsvgd_b200.dist import phi_sharded

dev = torch.device("cuda:0")
res = {}
for N in [int(a) for a in sys.argv[1:]] or [16384, 65536]:
    g = torch.Generator().manual_seed(7)
    X = torch.cumsum(0.05 * torch.randn(N, 64, 7, generator=g), dim=1).to(dev)
    G = torch.randn(N, 448, generator=g).to(dev)
    r = {}
    try:
        def t(fn, n=3):
            fn(); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(n): out = fn()
            e1.record(); torch.cuda.synchronize()
            return out, e0.elapsed_time(e1) / n
        a, r["fused_ms"] = t(lambda: ops.phi(X, G, depth=3))
        ra = ops.read_state(a.state)
        pa = a.phi.clone()
        del a; ops._WS_CACHE.clear(); torch.cuda.empty_cache()
        b, r["strip_ms"] = t(lambda: phi_sharded(X, G, depth=3))
        rb = ops.read_state(b.state)
        r["median_fused"], r["median_strip"] = ra["median"], rb["median"]
        r["h_fused"], r["h_strip"] = ra["h"], rb["h"]
        r["ties_fused"], r["ties_strip"] = ra["ties"], rb["ties"]
        skip = sorted({t_[0] for t_ in ra["ties"]} | {t_[0] for t_ in rb["ties"]})
        keep = torch.ones(N, dtype=torch.bool, device=dev); keep[skip] = False
        d = (pa[keep] - b.phi[keep]).abs().max().item()
        r["phi_rel_diff"] = d / b.phi[keep].abs().max().item()
        r["finite"] = bool(torch.isfinite(pa).all().item() and torch.isfinite(b.phi).all().item())
        del b
    except Exception as e:
        r["error"] = f"{type(e).__name__}: {e}"
    ops._WS_CACHE.clear(); torch.cuda.empty_cache()
    res[N] = r
    print(json.dumps({N: r}), flush=True)

"""Timings of the other BASELINE.json configs (development aid; run under gpurun)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sigsvgd_b200 import ops

dev = torch.device("cuda:0")


def t(fn, n=20, warm=5):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


res = {}
g = torch.Generator().manual_seed(0)
for name, shape in [("field_100x10x2", (100, 10, 2)), ("panda_256x64x7", (256, 64, 7)), ("walk_1024x64x7", (1024, 64, 7)),
                    ("walk_2048x64x7", (2048, 64, 7)), ("walk_4096x64x7", (4096, 64, 7)), ("walk_8192x64x7", (8192, 64, 7)),
                    ("walk_16384x64x7", (16384, 64, 7))]:
    N, T, d = shape
    X = torch.cumsum(0.05 * torch.randn(N, T, d, generator=g), dim=1).to(dev)
    G = torch.randn(N, T * d, generator=g).to(dev)
    res[name + "_us"] = t(lambda: ops.phi(X, G, depth=3))
    if N <= 2048:
        res[name + "_force_tc_us"] = t(lambda: ops.phi(X, G, depth=3, flags=8)) if N >= 128 else None
        res[name + "_force_simt_us"] = t(lambda: ops.phi(X, G, depth=3, flags=4))
B, N, T, d = 1024, 128, 64, 7
X = torch.cumsum(0.05 * torch.randn(B, N, T, d, generator=g), dim=2).to(dev)
G = torch.randn(B, N, T * d, generator=g).to(dev)
res["scenes_1024x128x64x7_us"] = t(lambda: ops.phi(X, G, depth=3), n=5, warm=2)
res["scenes_128x128x64x7_us (one GPU's share of 8)"] = t(lambda: ops.phi(X[:128], G[:128], depth=3), n=5, warm=2)
print(json.dumps(res, indent=1))

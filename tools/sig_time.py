"""Times the sequential and the time-parallel signature pairs (CUDA events, L2 flushed)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch

from sigsvgd_b200 import ops

dev = torch.device("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(fn, reps=20):
    ts = []
    for _ in range(reps + 3):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts = sorted(ts[3:])
    return ts[len(ts) // 2]


out = {}
for N in [int(a) for a in sys.argv[1:]] or [256, 1024, 4096]:
    T, d, depth = 64, 7, 3
    x = torch.cumsum(0.05 * torch.randn(N, T, d, device=dev), 1)
    sig, pref = ops.signature_fwd_prefix_raw(x, depth)
    g = torch.randn_like(sig)
    add = torch.randn(N, T * d, device=dev)
    out[f"N{N}"] = {
        "fwd": timeit(lambda: ops.signature_fwd_raw(x, depth)),
        "fwd_prefix": timeit(lambda: ops.signature_fwd_prefix_raw(x, depth)),
        "bwd": timeit(lambda: ops.signature_bwd_raw(x, sig, g, depth, add=add, alpha=1.0, beta=-1.0)),
        "bwd_prefix": timeit(lambda: ops.signature_bwd_prefix_raw(x, sig, pref, g, depth, add=add, alpha=1.0, beta=-1.0)),
    }
print(json.dumps(out, indent=1))

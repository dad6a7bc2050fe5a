"""Experiment: one phi-update replayed from a CUDA graph vs enqueued call by call (CUDA events, L2 flushed)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sigsvgd_b200 import ops  # noqa: E402

dev = torch.device("cuda:0")
N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
T, d, depth = 64, 7, 3
x = torch.cumsum(0.05 * torch.randn(N, T, d, device=dev), 1)
g = torch.randn(N, T * d, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(fn, reps=30):
    ts = []
    for _ in range(reps + 5):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts = sorted(ts[5:])
    return ts[len(ts) // 2]


def timeit_stream(fn, reps=30):
    """back to back, no flush: the launch queue stays full"""
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / reps


res = ops.phi(x, g, depth)
torch.cuda.synchronize()
side = torch.cuda.Stream()
graph = torch.cuda.CUDAGraph()
with torch.cuda.stream(side):
    ops.phi(x, g, depth)
    torch.cuda.synchronize()
    with torch.cuda.graph(graph, stream=side):
        res_g = ops.phi(x, g, depth)
torch.cuda.synchronize()
graph.replay()
torch.cuda.synchronize()
diff = (res_g.phi - res.phi).abs().max().item() / res.phi.abs().max().item()
print(json.dumps({
    "N": N, "rel_diff_graph_vs_direct": diff,
    "direct_flushed_us": timeit(lambda: ops.phi(x, g, depth)),
    "graph_flushed_us": timeit(graph.replay),
    "direct_back_to_back_us": timeit_stream(lambda: ops.phi(x, g, depth)),
    "graph_back_to_back_us": timeit_stream(graph.replay),
}, indent=1))

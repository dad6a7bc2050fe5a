"""Timings of the signature-PDE kernel at the reference's configurations (development aid; run under gpurun)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sigsvgd_b200 import _lib, ops

dev = torch.device("cuda:0")


def t(fn, n=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


res = {}
for name, A, M, D, order in [("maze_35x30x2_o0", 35, 30, 2, 0), ("maze_35x30x2_o3", 35, 30, 2, 3),
                             ("field_100x10x2_o3", 100, 10, 2, 3), ("panda_256x64x7_o0", 256, 64, 7, 0),
                             ("panda_256x64x7_o2", 256, 64, 7, 2)]:
    g = torch.Generator().manual_seed(0)
    x = torch.cumsum(0.1 * torch.randn(A, M, D, generator=g, dtype=torch.float64), 1).to(dev)
    gout = torch.ones(A, A, dtype=torch.float64, device=dev)
    st = ops.sigpde_bandwidth_raw(x, x)
    cells = A * A * ((M - 1) << order) ** 2
    f = t(lambda: ops.sigpde_gram_raw(x, x, order, _lib.PDE_RBF, 1.0, st))
    b = t(lambda: ops.sigpde_gram_bwd_raw(x, x, gout, order, _lib.PDE_RBF, 1.0, st))
    res[name] = {"bandwidth_ms": t(lambda: ops.sigpde_bandwidth_raw(x, x)), "gram_ms": f, "gram_bwd_ms": b,
                 "fine_cells": cells, "gram_gcells_per_s": cells / f / 1e6, "bwd_gcells_per_s": 2 * cells / b / 1e6}
print(json.dumps(res, indent=1))

"""Kernel table (CUPTI) of the batched-scenes update, 1024 x 128 particles (development aid)."""
import os, re, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.autograd import DeviceType
from torch.profiler import ProfilerActivity, profile
from sigsvgd_b200 import ops
B, N = int(sys.argv[1]) if len(sys.argv) > 1 else 1024, 128
g = torch.Generator().manual_seed(0)
x = torch.cumsum(0.05 * torch.randn(B, N, 64, 7, generator=g), 2).cuda()
s = torch.randn(B, N, 448, generator=g).cuda()
for _ in range(3):
    ops.phi(x, s)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(3):
        ops.phi(x, s)
    torch.cuda.synchronize()
tab = {}
for e in prof.events():
    if e.device_type == DeviceType.CUDA:
        m = re.search(r"(\w+_kernel)", e.name)
        k = m.group(1) if m else e.name[:40]
        tab[k] = tab.get(k, 0.0) + e.device_time / 3.0
for k, v in sorted(tab.items(), key=lambda kv: -kv[1]):
    print(f"{k:40s} {v:9.1f} us")
print("total", sum(tab.values()))

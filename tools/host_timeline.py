"""Device timeline (CUPTI, via the torch profiler) of one sigsvgd_phi_host call: kernels and copies with
their start offsets -- shows what overlaps what across PCIe (development aid; run under gpurun)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.autograd import DeviceType
from torch.profiler import ProfilerActivity, profile

from sigsvgd_b200 import ops

N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
X = torch.cumsum(0.05 * torch.randn(N, 64, 7, generator=g), dim=1).pin_memory()
G = torch.randn(N, 448, generator=g).pin_memory()
out = torch.empty_like(X).pin_memory()
for _ in range(3):
    ops.phi_host(X, G, out=out, device=dev)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    ops.phi_host(X, G, out=out, device=dev)
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == DeviceType.CUDA]
t0 = min(e.time_range.start for e in evs)
for e in sorted(evs, key=lambda e: e.time_range.start):
    print(f"{e.time_range.start - t0:9.1f} us  +{e.time_range.end - e.time_range.start:8.1f}  {e.name[:70]}")
print("total", max(e.time_range.end for e in evs) - t0)

"""Device timeline (CUPTI via the torch profiler) of one replay of the captured update: kernel start
offsets, durations and the gaps between consecutive kernels (development aid; run under gpurun)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.autograd import DeviceType
from torch.profiler import ProfilerActivity, profile
from sigsvgd_b200 import ops

N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
plan = ops.PhiPlan(N, 64, 7, 3)
for _ in range(5):
    plan()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(3):
        plan()
    torch.cuda.synchronize()
evs = sorted([e for e in prof.events() if e.device_type == DeviceType.CUDA], key=lambda e: e.time_range.start)
evs = evs[len(evs) * 2 // 3:]  # the last replay
t0 = evs[0].time_range.start
prev_end = None
for e in evs:
    gap = "" if prev_end is None else f"gap {e.time_range.start - prev_end:6.1f}"
    print(f"{e.time_range.start - t0:8.1f} us  +{e.time_range.end - e.time_range.start:7.1f}  {gap:12s} {e.name[:60]}")
    prev_end = max(prev_end or 0, e.time_range.end)
print("total", max(e.time_range.end for e in evs) - t0)

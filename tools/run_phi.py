"""A few phi-updates at one size (for ncu launch lists; run under gpurun)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sigsvgd_b200 import _lib, ops

N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 3
flags = int(sys.argv[3]) if len(sys.argv) > 3 else 0
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
X = torch.cumsum(0.05 * torch.randn(N, 64, 7, generator=g), dim=1).to(dev)
G = torch.randn(N, 448, generator=g).to(dev)
for _ in range(iters):
    r = ops.phi(X, G, depth=3, flags=flags)
torch.cuda.synchronize()
print("ok", float(r.phi.abs().max()), ops.read_state(r.state)["h"])

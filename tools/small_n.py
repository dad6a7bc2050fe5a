import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sigsvgd_b200 import ops
N = int(sys.argv[1]); flags = int(sys.argv[2]) if len(sys.argv) > 2 else 0
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
X = torch.cumsum(0.05 * torch.randn(N, 64, 7, generator=g), dim=1).to(dev)
G = torch.randn(N, 448, generator=g).to(dev)
for _ in range(6):
    ops.phi(X, G, depth=3, flags=flags)
torch.cuda.synchronize()

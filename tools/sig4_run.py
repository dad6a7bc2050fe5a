"""Launch the depth-4 signature kernels a few times (target of ncu captures; development aid)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sigsvgd_b200 import ops
N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
dev = torch.device("cuda:0")
X = torch.cumsum(0.05 * torch.randn(N, 64, 7, device=dev), dim=1)
for _ in range(3):
    s = ops.signature_fwd_raw(X, 4, True)
    g = torch.randn_like(s)
    ops.signature_bwd_raw(X, s, g, 4, True)
torch.cuda.synchronize()

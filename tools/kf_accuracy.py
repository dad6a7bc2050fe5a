"""Error of the fused K.F kernel against float64 at N = 4096 (chain-length study; development aid)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from sigsvgd_b200 import _lib, ops
dev = "cuda:0"
N, D, P = 4096, 399, 448
rng = np.random.default_rng(0)
s = (rng.standard_normal((N, D)) * np.exp(rng.uniform(-5, 0, size=D))).astype(np.float32)
g = (rng.standard_normal((N, P)) * 30).astype(np.float32)
dist = (rng.random((N, N)) * 6).astype(np.float32)
cu = lambda a: torch.as_tensor(a, device=dev)
mean = cu(s.mean(0).astype(np.float32))
planes, norms, scales, sc = ops.tc_pack_raw(cu(s), mean, cu(g))
state = ops.new_state(torch.device(dev))
_lib.call("sigsvgd_fixed_bandwidth", 1.3, state.data_ptr(), ops._stream())
d_dev = ops.dist_buffer(N, N, dev).copy_(cu(dist))
out = ops.tc_kf_raw(d_dev, planes, D + 1 + P, scales, state, _lib.KF_EXP, True).cpu().numpy().astype(np.float64)
K64 = torch.exp(-0.5 / float(np.float32(1.3)) ** 2 * cu(dist).double())
F64 = torch.cat([cu(s).double() - mean.double(), torch.ones(N, 1, device=dev, dtype=torch.float64), cu(g).double()], 1)
ref = (K64 @ F64).cpu().numpy()
for name, lo, hi in (("K.Sc", 0, D), ("rowsum", D, D + 1), ("K.score", D + 1, D + 1 + P)):
    e = out[:, lo:hi] - ref[:, lo:hi]
    print(f"{name:8s} max|err|/max|ref| {np.abs(e).max() / np.abs(ref[:, lo:hi]).max():.2e}   mean signed err / mean |ref| "
          f"{e.mean() / np.abs(ref[:, lo:hi]).mean():+.2e}")

"""Eager vs graph-replayed sharded step on every rank (development aid; torchrun under gpurun)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from sigsvgd_b200 import ops
from sigsvgd_b200.dist import ShardedPlan, phi_sharded

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl")
dev = torch.device("cuda", torch.cuda.current_device())
N = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
M = N // world
g = torch.Generator().manual_seed(100 + rank)
x = torch.cumsum(0.05 * torch.randn(M, 64, 7, generator=g), dim=1).to(dev)
s = torch.randn(M, 448, generator=g).to(dev)
e = phi_sharded(x, s)
plan = ShardedPlan(M, 64, 7, 3)
for it in range(3):
    r = plan(x, s)
    torch.cuda.synchronize()
    dv = (r.phi - e.phi).abs().amax(dim=(1, 2)) / e.phi.abs().max()
    rec_e, rec_g = ops.read_state(e.state), ops.read_state(r.state)
    print(f"rank {rank} it {it}: max dev {float(dv.max()):.3e} rows>1e-5 {int((dv > 1e-5).sum())} of {M}; "
          f"first bad {int(torch.nonzero(dv > 1e-5)[0]) if (dv > 1e-5).any() else -1}; "
          f"eager med/h/skd/nt {rec_e['median']:.6f} {rec_e['h']:.6f} {rec_e.get('sum_kd')} {rec_e['n_ties']} {rec_e.get('coef')} | "
          f"graph {rec_g['median']:.6f} {rec_g['h']:.6f} {rec_g.get('sum_kd')} {rec_g['n_ties']} {rec_g.get('coef')}", flush=True)
dist.destroy_process_group()

"""Depth-4 leg of BASELINE.json configs[3]: time the update at N in {4096, 16384} and check a row
sample of phi against the float64 closed form (development aid; run under gpurun)."""
import json
import re
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import oracle
from sigsvgd_b200 import ops

dev = torch.device("cuda:0")
depth = int(sys.argv[2]) if len(sys.argv) > 2 else 4
res = {}
for N in [int(a) for a in (sys.argv[1].split(",") if len(sys.argv) > 1 else ["4096"])]:
    g = torch.Generator().manual_seed(0)
    X = torch.cumsum(0.05 * torch.randn(N, 64, 7, generator=g), dim=1)
    G = torch.randn(N, 448, generator=g)
    Xd, Gd = X.to(dev), G.to(dev)
    for _ in range(2):
        r = ops.phi(Xd, Gd, depth=depth, want_sig=(N <= 4096))
    torch.cuda.synchronize()
    ts = []
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); r = ops.phi(Xd, Gd, depth=depth, want_sig=(N <= 4096)); e1.record()
        torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); s = ops.signature_fwd_raw(Xd, depth, True); e1.record(); torch.cuda.synchronize()
    t_fwd = e0.elapsed_time(e1)
    gs = torch.randn_like(s)
    e0.record(); ops.signature_bwd_raw(Xd, s, gs, depth, True); e1.record(); torch.cuda.synchronize()
    t_bwd = e0.elapsed_time(e1)
    out = {"ms_per_update": sorted(ts)[2], "sig_fwd_ms": t_fwd, "sig_bwd_ms": t_bwd}
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(5):
            ops.phi(Xd, Gd, depth=depth)
        torch.cuda.synchronize()
    tab = {}
    for ev in prof.events():
        if ev.device_type == torch.autograd.DeviceType.CUDA:
            m = re.search(r"(\w+_kernel)", ev.name)
            k = m.group(1) if m else ev.name[:40]
            tab[k] = tab.get(k, 0.0) + ev.device_time / 5.0
    out["kernels_us"] = {k: round(v, 1) for k, v in sorted(tab.items(), key=lambda kv: -kv[1])}
    if N <= 4096:
        t0 = time.time()
        rec = ops.read_state(r.state)
        cf = oracle.phi_closed_form(X.double().numpy(), G.double().numpy(), depth)
        ties = {a for a, _ in rec["ties"]} | {int(a) for a, _ in cf["ties"]}
        rows = [i for i in range(0, N, 7) if i not in ties]
        phi = r.phi.cpu().numpy()
        out.update({"h_rel": abs(rec["h"] - cf["h"]) / cf["h"],
                    "sig_rel": float(np.abs(r.sig.cpu().numpy() - cf["sig"]).max() / np.abs(cf["sig"]).max()),
                    "phi_rel_rows": float(np.abs(phi[rows] - cf["phi"][rows]).max() / np.abs(cf["phi"]).max()),
                    "rows": len(rows), "oracle_s": time.time() - t0})
    res[N] = out
    del Xd, Gd, r
    ops._WS_CACHE.clear(); torch.cuda.empty_cache()
print(json.dumps(res, indent=1))

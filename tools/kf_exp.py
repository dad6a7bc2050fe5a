"""K.F kernel experiments (development aid; run under gpurun)."""
import os, sys, json, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if len(sys.argv) > 1 and sys.argv[1] == "child":
    import torch
    from sigsvgd_b200 import ops
    N = int(sys.argv[2]); dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(0)
    X = torch.cumsum(0.05 * torch.randn(N, 64, 7, generator=g), dim=1).to(dev)
    G = torch.randn(N, 448, generator=g).to(dev)
    sig = ops.signature_fwd_raw(X, 3, True)
    planes, norms, scales, sc = ops.tc_pack_raw(sig, ops.col_mean(sig), G)
    d = ops.tc_gram_raw(planes, norms, planes, norms, 399, scales, True)
    st = ops.bandwidth_raw(d.clone(), ops.Bandwidth.median(), N)
    kpl = ops.tc_rbf_raw(d, st)
    flush = torch.empty(64 * 1024 * 1024, dtype=torch.float32, device=dev)
    fn = lambda: ops.tc_matmul_raw(kpl, planes, N, 848, scales)
    for _ in range(3): fn()
    torch.cuda.synchronize(); ts = []
    for _ in range(10):
        flush.fill_(1.0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort(); print(json.dumps({"us": ts[len(ts) // 2]}))
else:
    N = sys.argv[1] if len(sys.argv) > 1 else "4096"
    for env in [{"SIGSVGD_KF_EQUAL_TILES": str(v)} for v in (0, 1, 0, 1)]:
        e = dict(os.environ); e.update(env); e["SIGSVGD_KF_DEBUG"] = "1"
        r = subprocess.run([sys.executable, __file__, "child", N], env=e, capture_output=True, text=True)
        dbg = [l for l in r.stderr.splitlines() if l.startswith("kf:")]
        print(env, r.stdout.strip(), dbg[:1], flush=True)

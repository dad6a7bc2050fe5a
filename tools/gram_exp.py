"""Timing experiments on the Gram kernel (SIGSVGD_GRAM_DEBUG bit flags); development aid."""
import os, subprocess, sys, json
if len(sys.argv) > 1 and sys.argv[1] == "child":
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import torch
    from sigsvgd_b200 import ops
    N = 4096
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(0)
    X = torch.cumsum(0.05 * torch.randn(N, 64, 7, generator=g), dim=1).to(dev)
    G = torch.randn(N, 448, generator=g).to(dev)
    sig = ops.signature_fwd_raw(X, 3, True)
    planes, norms, scales, sc = ops.tc_pack_raw(sig, ops.col_mean(sig), G)
    dist = torch.empty(N, N, device=dev)
    state = ops.new_state(dev)
    ws = ops.tc_median_workspace(N, N, dev)
    flush = torch.empty(64 * 1024 * 1024, dtype=torch.float32, device=dev)
    def t(fn, n=10):
        for _ in range(3): fn()
        torch.cuda.synchronize(); ts = []
        for _ in range(n):
            flush.fill_(1.0)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1) * 1e3)
        return sorted(ts)[len(ts) // 2]
    ops.tc_median_prepare_raw(sc, sc, ws, state)
    def win():
        ops.tc_gram_raw(planes, norms, planes, norms, 399, scales, True, dist)
    print(json.dumps({"dbg": os.environ.get("SIGSVGD_GRAM_DEBUG", "0"),
                      "sym_win_us": t(win),
                      "sym_nowin_us": t(lambda: ops.tc_gram_raw(planes, norms, planes, norms, 399, scales, True, dist)),
                      "nonsym_nowin_us": t(lambda: ops.tc_gram_raw(planes, norms, planes, norms, 399, scales, False, dist))}))
else:
    for dbg in sys.argv[1:] or ["0", "1", "16", "4", "36", "12", "44", "61"]:
        env = dict(os.environ, SIGSVGD_GRAM_DEBUG=dbg)
        subprocess.run([sys.executable, __file__, "child"], env=env)

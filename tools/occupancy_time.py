"""Timing of the fused occupancy-network query against torch fp32 + autograd on the same GPU.
Usage: python tools/occupancy_time.py [P] [in_dim] [--profile]"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sigsvgd_b200 import occupancy  # noqa: E402


def timed(fn, n=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    P = int(args[0]) if len(args) > 0 else 256 * 200 * 60
    in_dim = int(args[1]) if len(args) > 1 else 3
    dev = torch.device("cuda:0")
    torch.manual_seed(5)
    layers = [torch.nn.Linear(in_dim, 200), torch.nn.ReLU()]
    for _ in range(3):
        layers += [torch.nn.Linear(200, 200), torch.nn.ReLU()]
    layers += [torch.nn.Linear(200, 1), torch.nn.Sigmoid()]
    net = torch.nn.Sequential(*layers).to(dev)
    with torch.no_grad():
        net[-2].weight.mul_(30.0)
    f = occupancy.fuse(net)
    x = torch.randn(P, in_dim, device=dev)

    def torch_query():
        xr = x.detach().requires_grad_(True)
        y = net(xr)
        return y, torch.autograd.grad(y.sum(), xr)[0]

    y_t, g_t = torch_query()
    y_t, g_t = y_t.detach(), g_t.detach()
    ms_t = timed(torch_query, 4, 2)
    flops = 2.0 * P * (200 * in_dim + 3 * 200 * 200 + 200) * 2
    out = {"points": P, "in_dim": in_dim, "ms_torch_fp32": ms_t}
    for name, precise in (("precise", True), ("fast", False)):
        prob, grad = f.query_raw(x, precise=precise)
        ms = timed(lambda: f.query_raw(x, precise=precise))
        ms_f = timed(lambda: f.query_raw(x, want_grad=False, precise=precise))
        dev = (grad - g_t).abs().amax(1) / g_t.abs().max()
        out[name] = {"ms": ms, "ms_fwd_only": ms_f, "speedup_vs_torch": ms_t / ms, "useful_tflops": flops / ms / 1e9,
                     "dev_prob": float((prob - y_t[:, 0]).abs().max()), "dev_grad_median": float(dev.median()),
                     "frac_points_grad_dev_above_1e-3": float((dev > 1e-3).float().mean())}
        if "--profile" in sys.argv:
            import ctypes

            from sigsvgd_b200 import _lib
            buf = (ctypes.c_ulonglong * 8)()
            _lib.load().sigsvgd_mlp_profile_enable(1)
            _lib.load().sigsvgd_mlp_profile_cycles(buf, 1)
            f.query_raw(x, precise=precise)
            torch.cuda.synchronize()
            _lib.load().sigsvgd_mlp_profile_cycles(buf, 1)
            _lib.load().sigsvgd_mlp_profile_enable(0)
            tot = float(sum(buf[:6])) or 1.0
            names = ["first_layer", "barrier", "mma_issue", "mma_wait", "epilogue", "output"]
            out[name]["thread0_cycle_share"] = {n: round(buf[i] / tot, 3) for i, n in enumerate(names)}
            out[name]["share_waiting_for_weights"] = round(buf[6] / tot, 3)
            out[name]["thread0_cycles_per_tile"] = tot / (148 * ((P + 127) // 128 / 148))
    print(json.dumps(out))


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Benchmark of the signature-kernel SVGD phi-update (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--particles 4096]

One "step" = one phi-update of a swarm of `--particles` 7-DoF trajectories (64 waypoints,
signature depth 3): X [N,64,7], score [N,448] -> phi [N,64,7].  Prints ONE JSON line.

* value      : phi-updates/s with inputs resident in HBM (CUDA events, max over ranks)
* e2e        : the same through the public API with HOST (pinned) buffers, H2D of X and score and
               D2H of phi inside the timed region
* roofline   : dominant kernel (the K.[score|S|1] contraction) timed alone with CUDA events
* cpu_baseline: the oracle port of the reference's CPU path on this box's host cores (bounded sample)
* --impl reference : the reference's CPU implementation (oracle port: the reference's own torch
               statements + the C/OpenMP signature restatement; `signatory` itself is absent)

N > 1 (torchrun), two things are measured and both are in the line:
* headline `value` ("scaling": "weak"): every rank updates its OWN independent swarm of
  `--particles` particles -- no communication (BASELINE.json north_star: "batches of independent
  planning problems shard with no communication"; also the reference's only multi-GPU mode,
  examples/script_planning_robot.py:619-631); value = ranks x updates / max-over-ranks time.
* `sharded_swarm`: ONE swarm sharded row-block over the ranks (all-gather of signatures + scores
  over NCCL, histogram all-reduces for the exact global median), at `--particles` and at
  `--large-particles` (default 65536, the north star's 1 -> 8 GPU scaling point; the same keys are
  printed at N = 1 from the single-GPU path so the speed-up can be read across runs).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

T_STEPS, D_CH, DEPTH = 64, 7, 3
# kernels of libsigsvgd_b200.so launched by one sigsvgd_phi call on the tcgen05 path (sig fwd + stats,
# window prepare, pack, Gram, window select, rbf planes, K.F, sig bwd); memset nodes
# are not counted
LAUNCHES_PER_UPDATE = 8
METRIC = "svgd_phi_updates_per_sec"
UNIT = "phi-updates/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--particles", type=int, default=4096)
    ap.add_argument("--large-particles", type=int, default=65536,
                    help="swarm size of the extra row-block-sharded measurement (0 = skip)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true",
                    help="enqueue every update call by call instead of replaying ops.PhiPlan's CUDA graph")
    ap.add_argument("--cpu-steps", type=int, default=3)
    return ap.parse_args()


def make_inputs(N, seed=0):
    import torch

    g = torch.Generator().manual_seed(seed)
    X = torch.cumsum(0.05 * torch.randn(N, T_STEPS, D_CH, generator=g), dim=1).contiguous()
    G = torch.randn(N, T_STEPS * D_CH, generator=g)
    return X, G


# ------------------------------------------------------------------------------------------------
# clocks
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i",
                 str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
                for name, col in (("hw_slowdown", 5), ("hw_thermal_slowdown", 6), ("sw_thermal_slowdown", 7),
                                  ("sw_power_cap", 8)):
                    if r[col].lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                continue
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle port of the reference's CPU path
# ------------------------------------------------------------------------------------------------
def cpu_phi_updates(N, steps, warmup=1):
    """Time `steps` phi-updates of the oracle port on the host cores.  Returns (updates/s, cores)."""
    import torch

    import oracle
    from oracle.signature import max_threads

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.autograd.set_detect_anomaly(False)
    X, G = make_inputs(N)
    for _ in range(warmup):
        oracle.phi_torch_literal(X, G, depth=DEPTH)
    t0 = time.perf_counter()
    for _ in range(steps):
        oracle.phi_torch_literal(X, G, depth=DEPTH)
    dt = (time.perf_counter() - t0) / steps
    return 1.0 / dt, min(cores, max(max_threads(), torch.get_num_threads())), dt


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    N = args.particles
    steps = max(1, min(args.steps, 5))
    val, cores, dt = cpu_phi_updates(N, steps, warmup=min(args.warmup, 1))
    sample = f"{steps} full phi-updates at N={N}, T={T_STEPS}, d={D_CH}, depth={DEPTH} (oracle port, all host threads)"
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": min(args.warmup, 1), "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(N), "particles": N, "waypoints": T_STEPS, "channels": D_CH,
                   "depth": DEPTH, "note": "host CPU arm: one swarm at a time on all host threads, whatever --gpus is"},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def workload_name(N):
    return (f"BASELINE.json configs[3] at N={N}: synthetic random-walk 7-D trajectories x 64 waypoints, "
            f"signature depth 3, median-heuristic RBF, phi = (K.score - grad_k)/N")


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist

    from sigsvgd_b200 import _lib, ops
    from sigsvgd_b200.dist import phi_sharded

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback for --impl ours)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        # NCCL logs ("NCCL version ...", warnings) go to STDOUT by default; stdout carries the JSON line only
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=dev)
    _lib.load()
    N = args.particles
    # headline: one independent swarm of N particles per rank (rank r's swarm has seed r)
    Xh, Gh = make_inputs(N, seed=rank)
    Xh, Gh = Xh.pin_memory(), Gh.pin_memory()
    X, G = Xh.to(dev), Gh.to(dev)
    phi_host = torch.empty_like(Xh).pin_memory()
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)  # > 126 MB L2
    M = N

    # inputs resident in HBM: the plan's static buffers; one CUDA-graph replay per update
    plan = None if args.no_graph else ops.PhiPlan(N, T_STEPS, D_CH, DEPTH, device=dev)
    if plan is not None:
        plan.x.copy_(X)
        plan.score.copy_(G)

    def step_device():
        if plan is not None:
            return plan().phi
        return ops.phi(X, G, depth=DEPTH).phi

    def step_e2e():
        # the public host-input call: pinned X, score -> pinned phi (H2D and D2H inside the call)
        return ops.phi_host(Xh, Gh, depth=DEPTH, out=phi_host, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        evs = []
        for _ in range(steps):
            flush.fill_(1.0)  # L2 flush between timed iterations (outside the events)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            evs.append((e0, e1))
        barrier()
        ms = sum(a.elapsed_time(b) for a, b in evs)
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()) / steps

    def sharded_swarm(Ns, steps, warmup):
        """ONE swarm of Ns particles, row-block over the ranks (single-GPU fused path at world 1)."""
        try:
            if Ns % world:
                return {"particles": Ns, "error": f"not divisible by {world} ranks"}
            Ms = Ns // world
            # only this rank's rows are generated (rows are i.i.d. random walks)
            g = torch.Generator().manual_seed(1234 + rank)
            xs = torch.cumsum(0.05 * torch.randn(Ms, T_STEPS, D_CH, generator=g), dim=1).to(dev)
            gs = torch.randn(Ms, T_STEPS * D_CH, generator=g).to(dev)
            if world == 1:
                fn = lambda: ops.phi(xs, gs, depth=DEPTH).phi  # noqa: E731
            else:
                fn = lambda: phi_sharded(xs, gs, depth=DEPTH).phi  # noqa: E731
            ms = timed(fn, steps, warmup)
            out = fn()
            ok = bool(torch.isfinite(out).all().item())
            del xs, gs, out
            ops._WS_CACHE.clear()
            torch.cuda.empty_cache()
            D = ops.sig_channels(D_CH, DEPTH)
            flops = 2.0 * Ns * Ns * (2 * D + T_STEPS * D_CH)
            return {"particles": Ns, "ms_per_step": ms, "value": 1e3 / ms, "unit": UNIT, "steps": steps,
                    "warmup": warmup, "finite": ok, "algorithmic_tflops": flops / (ms * 1e-3) / 1e12,
                    "exchange": ("none (one GPU)" if world == 1 else
                                 "NCCL all-gather of [S|score] rows + 3 histogram all-reduces + 1 stats all-reduce")}
        except Exception as e:  # the headline must survive a failure of the extra measurement
            torch.cuda.empty_cache()
            return {"particles": Ns, "error": f"{type(e).__name__}: {e}"[:300]}

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ms_dev = timed(step_device, args.steps, max(args.warmup, 3))
    clocks = sampler.stop() if rank == 0 else None
    ms_e2e = timed(step_e2e, args.steps, max(args.warmup, 3))

    # ---- roofline of the dominant kernel (K.[score|S|1] contraction), timed alone -------------
    roof = None
    if rank == 0:
        D = ops.sig_channels(D_CH, DEPTH)
        P = T_STEPS * D_CH
        Xa, Ga = make_inputs(N)
        s_all = ops.signature_fwd_raw(Xa.to(dev), DEPTH, True)
        use_tc = ops.tc_available() and N >= 256  # what the timed step actually ran
        if use_tc:
            # the tcgen05 contraction K.[S - mean | 1 | score] exactly as sigsvgd_phi launches it
            planes, norms, scales, sc = ops.tc_pack_raw(s_all, ops.col_mean(s_all), Ga.to(dev))
            d_rows = ops.tc_gram_raw(planes, norms, planes, norms, D, scales, True, out=ops.dist_buffer(N, N, dev))
            st = ops.bandwidth_raw(d_rows.contiguous(), ops.Bandwidth.median(), N)
            dominant = lambda: ops.tc_kf_raw(d_rows, planes, D + 1 + P, scales, st)  # noqa: E731
            kname = "kf_fused_kernel (tcgen05 cta_group::2 split-fp16, exp(gamma d).[Sc|1|score], RBF map fused)"
        else:
            v_all = ops.pack_features(s_all, ops.col_mean(s_all), Ga.to(dev))
            k_rows = ops.sqdist_raw(v_all[:, P:], v_all[:, P:], D)
            st = ops.bandwidth_raw(k_rows.clone(), ops.Bandwidth.median(), N)
            ops.rbf_apply_raw(k_rows, st)
            dominant = lambda: ops.matmul_nn_raw(k_rows, v_all)  # noqa: E731
            kname = "gemm_nn_simt_kernel (fp32 SIMT, K.[score|S|1])"
        for _ in range(3):
            dominant()
        torch.cuda.synchronize()
        evs = []
        for _ in range(10):
            flush.fill_(1.0)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            dominant()
            e1.record()
            evs.append((e0, e1))
        torch.cuda.synchronize()
        ms_k = sum(a.elapsed_time(b) for a, b in evs) / len(evs)
        flops = 2.0 * M * N * (P + D)  # algorithmic: K.score and K.S (the row-sum column is free)
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("bf16_tflops", 1590.0))
        # DRAM bytes of one launch of this kernel at this shape, from the committed ncu --set full capture
        traffic = None
        try:
            tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))["kf_kernel"]
            if use_tc and N == 4096:
                traffic = tj["dram_bytes_per_launch"]
        except Exception:
            pass
        roof = {"bound": "tensor", "kernel": kname, "mma_passes_per_fp32_product": 3 if use_tc else None,
                "achieved": flops / (ms_k * 1e-3) / 1e12,
                "peak": peak, "peak_source": ("MEASURED_PEAKS.json bf16_tflops (burst; fp16 and bf16 share the tcgen05 kind::f16 rate)"
                                if peaks else "fallback 1590 (B200_PROFILING.md)"),
                "unit": "TFLOP/s", "frac": flops / (ms_k * 1e-3) / 1e12 / peak, "traffic": traffic,
                "traffic_unit": "DRAM bytes per launch (ncu, profiles/traffic.json); operand planes are 82 MB",
                "ms_per_launch": ms_k, "algorithmic_flops_per_launch": flops,
                "step_algorithmic_flops": 2.0 * M * N * (2 * D + P),
                "step_frac_of_peak": 2.0 * M * N * (2 * D + P) / (ms_dev * 1e-3) / 1e12 / peak}
        if use_tc:
            # fp32-grade products cost three fp16 MMA passes (hi.hi, hi.lo, lo.hi): executed tensor flops
            # are 3x the algorithmic ones, so `frac` cannot exceed 1/3
            roof["frac_cap"] = 1.0 / 3.0
            roof["executed_frac_of_peak"] = 3.0 * roof["frac"]

    # ---- one swarm sharded row-block over the ranks (the path's NCCL exchange) ------------------
    sharded = []
    if world > 1:
        sharded.append(sharded_swarm(N, min(args.steps, 20), 3))
    if args.large_particles > 0:
        sharded.append(sharded_swarm(args.large_particles, 3, 3))

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        val, cores, dt = cpu_phi_updates(N, args.cpu_steps, warmup=1)
        cpu = {"value": val, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": f"{args.cpu_steps} full phi-updates at N={N} (oracle port of the reference's torch CPU path, "
                         f"C/OpenMP signature; signatory absent)", "ms_per_step": dt * 1e3}

    if rank == 0:
        line = {
            "metric": METRIC, "value": world * 1e3 / ms_dev, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_dev, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(N), "particles": N, "waypoints": T_STEPS, "channels": D_CH,
                       "depth": DEPTH,
                       "sharding": ("none" if world == 1 else
                                    f"replicas: {world} independent swarms of {N} particles, one per GPU, no communication "
                                    "(north_star: independent planning problems; reference: "
                                    "examples/script_planning_robot.py:619-631); see sharded_swarm for ONE swarm "
                                    "sharded row-block over NCCL"),
                       "l2": "256 MB buffer written between timed iterations", "k_xx_materialised_to_caller": False,
                       "device_step": ("ops.phi, enqueued call by call" if args.no_graph else
                                       "ops.PhiPlan: sigsvgd_phi captured once into a CUDA graph, one replay per update")},
            "clocks": clocks,
            "e2e": {"value": world * 1e3 / ms_e2e, "unit": UNIT,
                    "h2d_bytes_per_step": world * int(Xh.numel() + Gh.numel()) * 4,
                    "d2h_bytes_per_step": world * int(phi_host.numel()) * 4, "ms_per_step": ms_e2e},
            "gpu_launches": world * LAUNCHES_PER_UPDATE * args.steps,
            "sharded_swarm": sharded,
            "roofline": roof,
            "cpu_baseline": cpu,
        }
        emit(line)
    if world > 1:
        dist.destroy_process_group()


_JSON_FD = None


def _reserve_stdout():
    """stdout carries the JSON line and nothing else: libraries that write to file descriptor 1 from
    C (NCCL prints its version banner there regardless of NCCL_DEBUG_FILE) are sent to stderr, and the
    line is written to the saved descriptor."""
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    args = parse()
    _reserve_stdout()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Benchmark of the signature-kernel SVGD phi-update (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--particles 4096]

One "step" = one phi-update of a swarm of `--particles` 7-DoF trajectories (64 waypoints,
signature depth 3): X [N,64,7], score [N,448] -> phi [N,64,7].  Prints ONE JSON line.

* value      : phi-updates/s with inputs resident in HBM (CUDA events, max over ranks); also
               `ms_per_step_with_input_copy` (X and score copied into the plan's buffers every step)
* e2e        : the same through the public API with HOST (pinned) buffers, H2D of X and score and
               D2H of phi inside the timed region
* roofline   : dominant kernel (the fused exp(gamma d).[Sc|1|score] contraction) timed alone with CUDA
               events; `other_kernels`: the Gram (tensor), the signature forward / VJP and the median (HBM)
* gpu_launches : kernels of one update COUNTED by the profiler (CUPTI) x steps x ranks
* cpu_baseline: the oracle port of the reference's CPU path on this box's host cores (bounded sample)
* --impl reference : the reference's CPU implementation (oracle port: the reference's own torch
               statements + the C/OpenMP signature restatement; `signatory` itself is absent)

Extras in the same line (skipped with --skip-extras):
* `sharded_swarm`: ONE swarm sharded row-block over the ranks (all-gather of signatures + scores over
  NCCL, histogram all-reduces for the exact global median), at `--particles` and at `--large-particles`
  (default 65536, the north star's 1 -> 8 GPU scaling point), >= 10 timed steps, with a cross-rank PARITY
  record (median / h identical on every rank and equal to the one-GPU route's, max deviation of phi over
  a row sample) and the one-GPU time of the same swarm measured in the SAME run -> `sharded_efficiency`;
* `batched_scenes`: BASELINE.json configs[4], 1024 scenes x 128 particles split over the ranks;
* `mpc_step`: the "ms/MPC step" half of the metric (configs[2]) at the particle-maze shapes;
* `depth4_swarm`: configs[3]'s depth-4 end (N = 4096, D = 2800 signature channels), same update call;
* `sigpde`: the signature-PDE Gram matrix / gradient at the maze and Panda shapes, fp64.

N > 1 (torchrun): headline `value` ("scaling": "weak") = every rank updates its OWN independent swarm
of `--particles` particles, no communication (north_star: "batches of independent planning problems
shard with no communication"; the reference's only multi-GPU mode, examples/script_planning_robot.py:
619-631); value = ranks x updates / max-over-ranks time.
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
    ap.add_argument("--skip-extras", action="store_true",
                    help="headline, e2e and rooflines only (no sharded swarm / batched scenes / MPC step)")
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
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    val, cores, dt = cpu_phi_updates(N, steps, warmup=warmup)
    sample = (f"{steps} full phi-updates (after {warmup} warm-up) at N={N}, T={T_STEPS}, d={D_CH}, depth={DEPTH} "
              f"(oracle port, all host threads)")
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(N), "particles": N, "waypoints": T_STEPS, "channels": D_CH,
                   "depth": DEPTH},
        "note": "host CPU arm: one swarm at a time on all host threads, whatever --gpus is",
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
def count_kernels(fn):
    """Kernels launched by one call of `fn`, counted by the profiler (CUPTI), not assumed."""
    import torch

    try:
        from torch.profiler import ProfilerActivity, profile

        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            fn()
            torch.cuda.synchronize()
        from torch.autograd import DeviceType

        n, names = 0, {}
        for e in prof.events():
            if e.device_type != DeviceType.CUDA or e.name.lower().startswith(("memcpy", "memset")):
                continue
            n += 1
            key = e.name.replace("sigsvgd::<unnamed>::", "").replace("void ", "")[:48]
            names[key] = names.get(key, 0) + 1
        return (n, names) if n > 0 else (None, {})
    except Exception:
        return None, {}


def run_ours(args):
    import torch
    import torch.distributed as dist

    from sigsvgd_b200 import _lib, ops
    from sigsvgd_b200.dist import ShardedPlan, phi_sharded, shard_scenes

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
    D = ops.sig_channels(D_CH, DEPTH)
    P = T_STEPS * D_CH
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak_tf = float(peaks.get("bf16_tflops", 1590.0))
    peak_hbm = float(peaks.get("hbm_gbs", 6550.0))
    peak_src = "MEASURED_PEAKS.json" if peaks else "fallback (B200_PROFILING.md)"
    # headline: one independent swarm of N particles per rank (rank r's swarm has seed r)
    Xh, Gh = make_inputs(N, seed=rank)
    Xh, Gh = Xh.pin_memory(), Gh.pin_memory()
    X, G = Xh.to(dev), Gh.to(dev)
    phi_host = torch.empty_like(Xh).pin_memory()
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)  # > 126 MB L2

    # inputs resident in HBM: the plan's static buffers; one CUDA-graph replay per update
    plan = None if args.no_graph else ops.PhiPlan(N, T_STEPS, D_CH, DEPTH, device=dev)
    if plan is not None:
        plan.x.copy_(X)
        plan.score.copy_(G)

    def step_device():
        if plan is not None:
            return plan().phi
        return ops.phi(X, G, depth=DEPTH).phi

    def step_device_copy_in():
        # what an optimisation loop that owns its particles pays: inputs copied into the plan each step
        if plan is not None:
            return plan(X, G).phi
        return ops.phi(X, G, depth=DEPTH).phi

    def step_e2e():
        # the public host-input call: pinned X, score -> pinned phi (H2D and D2H inside the call)
        return ops.phi_host(Xh, Gh, depth=DEPTH, out=phi_host, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup, reduce_max=True):
        for _ in range(warmup):
            fn()
        if reduce_max:
            barrier()
        else:
            torch.cuda.synchronize()
        evs = []
        for _ in range(steps):
            flush.fill_(1.0)  # L2 flush between timed iterations (outside the events)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            evs.append((e0, e1))
        if reduce_max:
            barrier()
        else:
            torch.cuda.synchronize()
        ms = sum(a.elapsed_time(b) for a, b in evs)
        if reduce_max and world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms / steps


    def rows_of_rank(Ns, r):
        """Rows of rank r of the sharded swarm of Ns particles (rows are i.i.d. random walks)."""
        Ms = Ns // world
        g = torch.Generator().manual_seed(1234 + r)
        xs = torch.cumsum(0.05 * torch.randn(Ms, T_STEPS, D_CH, generator=g), dim=1)
        gs = torch.randn(Ms, T_STEPS * D_CH, generator=g)
        return xs, gs

    def sharded_swarm(Ns, steps, warmup):
        """ONE swarm of Ns particles, row-block over the ranks (single-GPU fused path at world 1), with
        a cross-rank parity record and, at world > 1, the SAME-RUN one-GPU time of the same swarm."""
        try:
            if Ns % world:
                return {"particles": Ns, "error": f"not divisible by {world} ranks"}
            xs, gs = (t.to(dev) for t in rows_of_rank(Ns, rank))
            if world == 1:
                fn = lambda: ops.phi(xs, gs, depth=DEPTH)  # noqa: E731
            else:
                fn = lambda: phi_sharded(xs, gs, depth=DEPTH)  # noqa: E731
            ms = timed(lambda: fn().phi, steps, warmup)
            res = fn()
            out = {"particles": Ns, "ms_per_step": ms, "value": 1e3 / ms, "unit": UNIT, "steps": steps, "warmup": warmup,
                   "algorithmic_tflops": 2.0 * Ns * Ns * (2 * D + P) / (ms * 1e-3) / 1e12,
                   "exchange": ("none (one GPU)" if world == 1 else
                                "NCCL all-gather of the fp16 operand planes [S|1|score] + centred features of every rank's rows, 3 histogram all-reduces, 2 small all-reduces")}
            if world > 1:
                phases = {}
                phi_sharded(xs, gs, depth=DEPTH, timings=phases)
                out["phase_ms_rank0"] = {k: round(v, 3) for k, v in phases.items()}
                out["device_step"] = "phi_sharded, enqueued call by call"
                try:
                    # the same step (kernels and collectives) replayed from one CUDA graph
                    plan = ShardedPlan(xs.shape[0], T_STEPS, D_CH, DEPTH)
                    ms_g = timed(lambda: plan(xs, gs).phi, steps, warmup)
                    res_g = plan(xs, gs)
                    # (run-to-run: the column statistics use floating-point atomics -> last-bit differences)
                    med_g, med_e = float(ops.state_field(res_g.state, "median")[0]), float(ops.state_field(res.state, "median")[0])
                    same = abs(med_g - med_e) <= 1e-6 * abs(med_e)
                    # (rows of the recorded median pairs differ from run to run once more pairs tie than the
                    # record holds -- include/sigsvgd_b200.h SIGSVGD_MAX_TIES -- so rows are counted, not maxed)
                    dev_rows = (res_g.phi - res.phi).abs().amax(dim=(1, 2)) / res.phi.abs().max()
                    n_bad = int((dev_rows > 1e-5).sum())
                    out["ms_per_step_call_by_call"] = ms
                    out["graph_replay"] = {"ms_per_step": ms_g, "same_median_as_eager": same,
                                           "rows_differing_from_eager_above_1e-5": n_bad, "rows": int(dev_rows.numel()),
                                           "median_dev_phi_rows": float(dev_rows.median())}
                    if same and n_bad <= 4 * _lib.MAX_TIES and ms_g < ms:
                        ms = ms_g
                        out.update({"ms_per_step": ms, "value": 1e3 / ms, "device_step": "ShardedPlan: the sharded step "
                                    "(kernels and NCCL collectives) captured once into a CUDA graph, one replay per update",
                                    "algorithmic_tflops": 2.0 * Ns * Ns * (2 * D + P) / (ms * 1e-3) / 1e12})
                    del plan, res_g
                except Exception as e:
                    out["graph_replay"] = {"error": f"{type(e).__name__}: {e}"[:300]}
                # ---- parity record: identical bandwidth on every rank, phi rows against the fused one-GPU update
                rec = torch.stack([ops.state_field(res.state, "median")[0], ops.state_field(res.state, "h")[0]]).double()
                allrec = [torch.empty_like(rec) for _ in range(world)]
                dist.all_gather(allrec, rec)
                nsamp = min(64, xs.shape[0])
                mine = res.phi[:nsamp].contiguous()
                rows = [torch.empty_like(mine) for _ in range(world)]
                dist.all_gather(rows, mine)
                # rows that carry a recorded median pair on this rank (global indices; -1 = empty slot)
                nt = int(ops.state_field(res.state, "n_ties")[0].clamp(max=_lib.MAX_TIES))
                tie_mine = torch.full((_lib.MAX_TIES,), -1, dtype=torch.int64, device=dev)
                tie_mine[:nt] = ops.state_field(res.state, "ties")[0, :nt, 0].to(torch.int64)
                tie_all = [torch.empty_like(tie_mine) for _ in range(world)]
                dist.all_gather(tie_all, tie_mine)
                del res
                ms1 = None
                parity = {"ranks_agree_on_median_and_h": bool(all(torch.equal(a, allrec[0]) for a in allrec))}
                if rank == 0:
                    full = [rows_of_rank(Ns, r) for r in range(world)]
                    xa = torch.cat([f[0] for f in full]).to(dev)
                    ga = torch.cat([f[1] for f in full]).to(dev)
                    one = ops.phi(xa, ga, depth=DEPTH)
                    Ms = Ns // world
                    ref_rows = torch.stack([one.phi[r * Ms:r * Ms + nsamp] for r in range(world)])
                    got_rows = torch.stack(rows)
                    scale = float(one.phi.abs().max())
                    # rows of the median pair differ when the two routes pick another of two equal-to-rounding
                    # candidates; they are reported, not hidden
                    dev_rows = (got_rows - ref_rows).abs().amax(dim=(2, 3)) / scale
                    # the same with the rows of the recorded median pairs of either route left out
                    tie_rows = {a for a, _ in ops.read_state(one.state)["ties"]} | \
                        {int(v) for v in torch.cat(tie_all).cpu().tolist() if v >= 0}
                    keep = torch.tensor([[(r * Ms + i) not in tie_rows for i in range(nsamp)] for r in range(world)],
                                        device=dev)
                    dev_keep = torch.where(keep, dev_rows, torch.zeros_like(dev_rows))
                    parity.update({
                        "max_rel_dev_phi_outside_tie_rows": float(dev_keep.max()),
                        "rows_above_1e-4_outside_tie_rows": int((dev_keep > 1e-4).sum()),
                        "sampled_rows_that_are_tie_rows": int((~keep).sum()),
                    })
                    parity.update({
                        "one_gpu_median": float(ops.state_field(one.state, "median")[0]),
                        "sharded_median": float(allrec[0][0]),
                        "h_rel_diff": abs(float(ops.state_field(one.state, "h")[0]) - float(allrec[0][1])) /
                                      float(ops.state_field(one.state, "h")[0]),
                        "rows_compared": int(world * nsamp),
                        "max_rel_dev_phi": float(dev_rows.max()),
                        "rows_above_1e-4": int((dev_rows > 1e-4).sum()),
                        "one_gpu_tie_rows": sorted({a for a, _ in ops.read_state(one.state)["ties"]}),
                    })
                    # same-run one-GPU time of the same swarm (rank 0 alone; the others wait at the barrier)
                    ms1 = timed(lambda: ops.phi(xa, ga, depth=DEPTH).phi, min(5, steps), 2, reduce_max=False)
                    del xa, ga, one, ref_rows, got_rows
                dist.barrier()
                out["parity"] = parity
                if ms1 is not None:
                    out["one_gpu_ms_same_run"] = ms1
                    out["speedup_vs_one_gpu"] = ms1 / ms
                    out["sharded_efficiency"] = ms1 / ms / world
            else:
                out["finite"] = bool(torch.isfinite(res.phi).all().item())
                del res
            del xs, gs
            ops._WS_CACHE.clear()
            torch.cuda.empty_cache()
            return out
        except Exception as e:  # the headline must survive a failure of the extra measurement
            torch.cuda.empty_cache()
            return {"particles": Ns, "error": f"{type(e).__name__}: {e}"[:300]}

    def batched_scenes(n_scenes, n_part, steps, warmup):
        """BASELINE.json configs[4]: independent planning problems split over the ranks, no
        communication; each rank updates its scenes in ONE batched call (grouped tcgen05 launches)."""
        try:
            mine = shard_scenes(n_scenes, rank, world)
            g = torch.Generator().manual_seed(77 + rank)
            xs = torch.cumsum(0.05 * torch.randn(len(mine), n_part, T_STEPS, D_CH, generator=g), dim=2).to(dev)
            gs = torch.randn(len(mine), n_part, T_STEPS * D_CH, generator=g).to(dev)
            ms = timed(lambda: ops.phi(xs, gs, depth=DEPTH).phi, steps, warmup)
            ok = bool(torch.isfinite(ops.phi(xs, gs, depth=DEPTH).phi).all().item())
            del xs, gs
            ops._WS_CACHE.clear()
            torch.cuda.empty_cache()
            return {"scenes": n_scenes, "particles_per_scene": n_part, "scenes_per_rank": len(mine), "ms_per_step": ms,
                    "scene_updates_per_s": n_scenes / (ms * 1e-3), "steps": steps, "finite": ok,
                    "sharding": f"shard_scenes over {world} rank(s), no communication"}
        except Exception as e:
            torch.cuda.empty_cache()
            return {"scenes": n_scenes, "error": f"{type(e).__name__}: {e}"[:300]}

    def depth4_swarm(Ns, steps, warmup):
        """BASELINE.json configs[3], depth 4: the same update call with 2800 signature channels (the
        dedicated depth-4 signature kernels, tcgen05 Gram / K.F over D = 2800)."""
        try:
            xs, gs = (t.to(dev) for t in make_inputs(Ns, seed=5))
            ms = timed(lambda: ops.phi(xs, gs, depth=4).phi, steps, warmup, reduce_max=False)
            res = ops.phi(xs, gs, depth=4)
            rec = ops.read_state(res.state)
            ok = bool(torch.isfinite(res.phi).all().item())
            del xs, gs, res
            ops._WS_CACHE.clear()
            torch.cuda.empty_cache()
            return {"particles": Ns, "depth": 4, "sig_channels": ops.sig_channels(D_CH, 4), "ms_per_step": ms,
                    "updates_per_s": 1e3 / ms, "steps": steps, "finite": ok, "median": rec["median"], "h": rec["h"],
                    "n_ties": rec["n_ties"]}
        except Exception as e:
            torch.cuda.empty_cache()
            return {"particles": Ns, "depth": 4, "error": f"{type(e).__name__}: {e}"[:300]}

    def sigpde_leg(steps, warmup):
        """the signature-PDE Gram matrix and its gradient (what the reference's example
        scripts call through ``sigkernel``) at the particle-maze shapes (35 x 30 x 2, dyadic order 3,
        examples/particle_maze_config.yaml:20) and at Panda-shaped paths (256 x 64 x 7, order 0), fp64."""
        try:
            out = {}
            for name, A, M, Dp, order in (("maze_35x30x2_order3", 35, 30, 2, 3), ("panda_256x64x7_order0", 256, 64, 7, 0)):
                g = torch.Generator().manual_seed(3)
                xp = torch.cumsum(0.1 * torch.randn(A, M, Dp, generator=g, dtype=torch.float64), 1).to(dev)
                go = torch.ones(A, A, dtype=torch.float64, device=dev)
                st = ops.sigpde_bandwidth_raw(xp, xp)
                cells = A * A * ((M - 1) << order) ** 2
                ms_bw = timed(lambda: ops.sigpde_bandwidth_raw(xp, xp), steps, warmup, reduce_max=False)
                ms_f = timed(lambda: ops.sigpde_gram_raw(xp, xp, order, _lib.PDE_RBF, 1.0, st), steps, warmup, reduce_max=False)
                ms_b = timed(lambda: ops.sigpde_gram_bwd_raw(xp, xp, go, order, _lib.PDE_RBF, 1.0, st), steps, warmup,
                             reduce_max=False)
                out[name] = {"median_bandwidth_ms": ms_bw, "gram_ms": ms_f, "gram_bwd_ms": ms_b, "fine_cells": cells,
                             "gram_gcells_per_s": cells / ms_f / 1e6, "dtype": "f64"}
            return out
        except Exception as e:
            torch.cuda.empty_cache()
            return {"error": f"{type(e).__name__}: {e}"[:300]}

    def occupancy_leg(steps, warmup):
        """the learned occupancy network of the robot-planning cost (3 -> 200 x 4 -> 1,
        continuous_occupancy_map.py:10-35) queried WITH its input gradient at 256 particles x 200 spline steps x
        60 body points (script_planning_robot.py:300-318 at BASELINE's Panda particle count) and the
        self-collision network (7 -> 200 x 4 -> 1) at 256 x 200 configurations: one fused tcgen05 launch each,
        against the same query as torch fp32 statements + autograd on the same GPU.  Random-initialised weights
        (the trained ones are a download)."""
        try:
            from sigsvgd_b200 import occupancy

            out = {}
            for name, in_dim, P in (("occupancy_3d_256x200x60", 3, 256 * 200 * 60), ("self_collision_7d_256x200", 7, 256 * 200)):
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
                ms_t = timed(torch_query, max(2, steps // 2), 2, reduce_max=False)
                flops = 2.0 * P * (200 * in_dim + 3 * 200 * 200 + 200) * 2  # forward + input gradient
                out[name] = {"points": P, "ms_torch_fp32_forward_autograd": ms_t}
                for mode, precise in (("precise", True), ("fast", False)):
                    prob, grad = f.query_raw(x, precise=precise)
                    ms = timed(lambda: f.query_raw(x, precise=precise), steps, warmup, reduce_max=False)
                    ms_fwd = timed(lambda: f.query_raw(x, want_grad=False, precise=precise), steps, warmup, reduce_max=False)
                    dev_g = (grad - g_t).abs().amax(1) / g_t.abs().max()
                    mmas = 3 if precise else 1
                    out[name][mode] = {
                        "ms_prob_and_grad": ms, "ms_prob_only": ms_fwd, "speedup_vs_torch": ms_t / ms,
                        "useful_tflops": flops / ms / 1e9,
                        "tensor_tflops_issued": 2.0 * P * 6 * 208 * 208 * mmas / ms / 1e9,
                        "frac_of_bf16_peak_issued": 2.0 * P * 6 * 208 * 208 * mmas / ms / 1e9 / peak_tf,
                        "max_abs_dev_prob_vs_torch_fp32": float((prob - y_t[:, 0]).abs().max()),
                        "median_dev_grad_rel_to_max_grad": float(dev_g.median()),
                        "frac_points_grad_dev_above_1e-3": float((dev_g > 1e-3).float().mean()),
                        "arithmetic": ("fp16 hi + lo planes, 3 MMAs per product, fp32 accumulate" if precise else
                                       "fp16 operands, 1 MMA per product, fp32 accumulate")}
                prob = grad = None
                del net, f, x, prob, grad, y_t, g_t
            torch.cuda.empty_cache()
            return out
        except Exception as e:
            torch.cuda.empty_cache()
            return {"error": f"{type(e).__name__}: {e}"[:300]}

    def robot_cost_leg():
        """SURVEY.md 8(f)-4 assembled: the reference's robot-planning cost and its gradient w.r.t. the knots
        (examples/script_planning_robot.py:248-356, :107) at BASELINE's 256 Panda particles, 3 knots, 200 spline
        steps -- as torch statements, and with this package's drop-ins for the forward kinematics and the two
        networks (tools/robot_cost_time.py holds both arms)."""
        try:
            sys.path.insert(0, os.path.join(ROOT, "tools"))
            import robot_cost_time

            out = {}
            for mode, precise in (("precise", True), ("fast", False)):
                r = robot_cost_time.run(dev, 256, 200, precise=precise)
                r["speedup_vs_torch"] = r["ms_torch_statements"] / r["ms_fused_drop_ins"]
                out[mode] = r
            if world == 1:
                out["cpu_sample"] = robot_cost_time.run_cpu_sample()
            torch.cuda.empty_cache()
            return out
        except Exception as e:
            torch.cuda.empty_cache()
            return {"error": f"{type(e).__name__}: {e}"[:300]}

    def mpc_step(steps, warmup):
        """"ms/MPC step" half of the metric (BASELINE.json configs[2]): one `stein_sampler.optimize(...,
        opt_steps=2)` of the particle-maze controller at its shapes (35 policies x 30 steps x 2 controls,
        10 action samples; src/controllers/dust.py:520-522, examples/particle_maze_config.yaml:8-20)
        through the `SigKernel` branch of TrajectorySVGD (trajectory_svgd.py:55-69).  The rollout is the
        particle model's (`sigsvgd_particle_rollout`, bit-identical to the reference model on its golden
        vectors) with a block obstacle grid and a quadratic cost to the target; the kernel statements are the
        reference's."""
        try:
            import warnings

            from sigsvgd_b200 import compat
            from sigsvgd_b200.rollout import particle_rollout

            S, Pn, H, dt = 10, 35, 30, 0.015
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                kernel = compat.SigKernel(compat.RBFKernel(sigma=32 ** 0.5), dyadic_order=3)

            def make(device, kern, capturable=False):
                gen = torch.Generator().manual_seed(5)
                pol = (5.0 * torch.randn(Pn, H, 2, generator=gen)).to(device)
                noise = (5.0 * torch.randn(2, S, Pn, H, 2, generator=gen)).to(device)
                mask = torch.ones(Pn, H, 2, device=device)
                mask[-5:] = 0.0
                target = torch.tensor([1.8, 1.8], device=device)
                grid = torch.zeros(400, 400, dtype=torch.uint8, device=device)  # 4 x 4 blocks, as sm_grid_4x4
                for gi in range(4):
                    for gj in range(4):
                        grid[70 + 80 * gi:100 + 80 * gi, 70 + 80 * gj:100 + 80 * gj] = 1

                def rollout(x0, acts):
                    if device.type == "cuda":
                        return particle_rollout(x0, acts, 2.0, dt, float("inf"), 5.0, grid, 0.01, (200, 200), True)
                    s_ = x0.expand(acts.shape[0], 1, -1).clone()  # the CPU arm: the reference's statements per step
                    for t in range(H):
                        cur = s_[:, t]
                        idx = (cur[:, :2] * 100.0 + 200).floor().long().clamp(0, 399)
                        crash = grid[idx[:, 0], idx[:, 1]].float().unsqueeze(-1)
                        nxt = cur + torch.cat((cur[:, 2:], acts[:, t] / 2.0), -1) * dt * (1 - crash)
                        nxt = torch.cat((nxt[:, :2], nxt[:, 2:].clamp(-5.0, 5.0)), -1)
                        s_ = torch.cat([s_, nxt.unsqueeze(1)], dim=1)
                    return s_

                def optimize(x0, opt_steps=2):
                    p = pol.clone()
                    opt = torch.optim.Adam([p], lr=1.0, capturable=capturable)
                    cost = None
                    for i in range(opt_steps):
                        actions = (p.detach()[None] + noise[i]).requires_grad_(True)
                        # rollouts: DuSt._rollout over ParticleModel.step (mass 2, dt 0.015, speed clamp 5, crash
                        # mask from an occupancy grid) as one launch, differentiable w.r.t. the actions
                        states = rollout(x0, actions.flatten(end_dim=-3)).reshape(S, Pn, H + 1, 4)
                        pos = states[..., 1:, :2]
                        cost = ((pos - target) ** 2).sum((-1, -2)) + 0.2 * (actions ** 2).sum((-1, -2))
                        score = torch.autograd.grad(-cost.sum(), actions, retain_graph=True)[0].mean(0).flatten(1)
                        tau = pos.mean(0)
                        k_xx = kern.compute_Gram(tau.double(), tau.detach().double(), sym=False).float()
                        grad_k = torch.autograd.grad(k_xx.sum(), actions)[0].mean(0).flatten(1)
                        velocity = -((k_xx.detach() @ score - grad_k) / Pn).reshape(Pn, H, 2) * mask
                        p.grad = velocity
                        opt.step()
                    return torch.index_select(p.detach(), 0, cost.mean(0).argmin().reshape(1))[0]  # (no host sync)

                return optimize

            opt_gpu = make(dev, kernel)
            x0_host = torch.tensor([-1.8, -1.8, 0.0, 0.0]).pin_memory()
            a_host = torch.empty(H, 2).pin_memory()

            def e2e():
                x0 = x0_host.to(dev, non_blocking=True)
                a_host.copy_(opt_gpu(x0), non_blocking=True)

            ms_e2e = timed(e2e, steps, warmup, reduce_max=False)
            tau = torch.cumsum(0.05 * torch.randn(Pn, H, 2, generator=torch.Generator().manual_seed(6)), 1).to(dev)

            def kernel_only():
                t = tau.detach().requires_grad_(True)
                k = kernel.compute_Gram(t.double(), t.detach().double(), sym=False).float()
                return torch.autograd.grad(k.sum(), t)[0]

            ms_kernel = timed(kernel_only, steps, warmup, reduce_max=False)
            # the same step captured ONCE into a CUDA graph and replayed (every launch of the library is
            # capturable; the adapter is told that both arguments hold the same values, which removes its
            # one host synchronisation)
            ms_graph = None
            try:
                kernel.same_inputs = True
                opt_cap = make(dev, kernel, capturable=True)
                x0_dev = x0_host.to(dev)
                side = torch.cuda.Stream(device=dev)
                side.wait_stream(torch.cuda.current_stream(dev))
                with torch.cuda.stream(side):
                    for _ in range(3):
                        a_static = opt_cap(x0_dev)
                    side.synchronize()
                    graph = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(graph, stream=side):
                        a_static = opt_cap(x0_dev)
                torch.cuda.current_stream(dev).wait_stream(side)

                def replay():
                    x0_dev.copy_(x0_host, non_blocking=True)
                    graph.replay()
                    a_host.copy_(a_static, non_blocking=True)

                ms_graph = timed(replay, steps, warmup, reduce_max=False)
            except Exception as e:  # capture is an extra; the eager numbers stand without it
                ms_graph = f"capture failed: {type(e).__name__}: {e}"[:200]
            finally:
                kernel.same_inputs = False
            out = {"ms_per_mpc_step_e2e": ms_e2e, "ms_per_mpc_step_cuda_graph_replay": ms_graph,
                   "ms_kernel_terms_only": ms_kernel, "opt_steps": 2,
                   "shapes": "35 policies x 30 x 2, 10 action samples, signature depth 3 (D = 14)",
                   "h2d_bytes_per_step": 16, "d2h_bytes_per_step": H * 2 * 4,
                   "rollout": "sigsvgd_particle_rollout: the reference's ParticleModel.step loop (acceleration control, "
                              "mass 2, dt 0.015, speed clamp 5, crash mask from a 4 x 4 block occupancy grid) as one launch; "
                              "quadratic cost to the target"}
            # the same step with the reference example's own kernel: the Goursat-PDE signature kernel at
            # the yaml's dyadic order 3 (script_control_particle_maze.py:42-44)
            try:
                opt_pde = make(dev, compat.PdeSigKernel(compat.RBFKernel(sigma=32 ** 0.5), dyadic_order=3))

                def e2e_pde():
                    x0 = x0_host.to(dev, non_blocking=True)
                    a_host.copy_(opt_pde(x0), non_blocking=True)

                out["ms_per_mpc_step_e2e_pde_kernel_order3"] = timed(e2e_pde, steps, warmup, reduce_max=False)
            except Exception as e:
                out["ms_per_mpc_step_e2e_pde_kernel_order3"] = f"{type(e).__name__}: {e}"[:200]
            # the rollout itself: the reference's loop over ParticleModel.step as torch
            # statements vs the one-launch kernel pair, forward + gradient w.r.t. the actions, 350 x 30
            try:
                from sigsvgd_b200.rollout import particle_rollout

                gen = torch.Generator().manual_seed(9)
                acts = (10 * torch.randn(350, 30, 2, generator=gen)).to(dev).requires_grad_(True)
                grid = torch.zeros(400, 400, dtype=torch.uint8, device=dev)
                grid[150:250:20, :] = 1
                s0 = torch.tensor([-1.8, -1.8, 0.0, 0.0], device=dev)
                wgt = torch.randn(350, 31, 4, generator=gen).to(dev)

                def fused():
                    st_ = particle_rollout(s0, acts, 2.0, 0.015, float("inf"), 5.0, grid, 0.01, (200, 200), True)
                    return torch.autograd.grad((st_ * wgt).sum(), acts)[0]

                def looped():  # particle.py:155-172 per step, dust.py:293-300
                    s_ = s0.expand(350, 1, -1).clone()
                    for t in range(30):
                        cur = s_[:, t]
                        idx = (cur[:, :2] * 100.0 + 200).floor().long().clamp(0, 399)
                        mask = grid[idx[:, 0], idx[:, 1]].float().unsqueeze(-1)
                        nxt = cur + torch.cat((cur[:, 2:], acts[:, t] / 2.0), -1) * 0.015 * (1 - mask)
                        nxt = torch.cat((nxt[:, :2], nxt[:, 2:].clamp(-5.0, 5.0)), -1)
                        s_ = torch.cat([s_, nxt.unsqueeze(1)], dim=1)
                    return torch.autograd.grad((s_ * wgt).sum(), acts)[0]

                same = float((fused() - looped()).abs().max() / looped().abs().max())
                out["rollout_350x30_fwd_bwd_ms"] = {
                    "kernel_pair": timed(fused, steps, warmup, reduce_max=False),
                    "torch_statements_per_step": timed(looped, steps, warmup, reduce_max=False),
                    "max_rel_dev_of_the_gradients": same}
            except Exception as e:
                out["rollout_350x30_fwd_bwd_ms"] = {"error": f"{type(e).__name__}: {e}"[:200]}
            if rank == 0 and world == 1 and not args.no_cpu_baseline:
                import oracle

                class CpuKernel:  # the same statements on the host cores: oracle signature + torch (the reference's path)
                    def compute_Gram(self, Xp, Yp, sym=False):
                        sx = oracle.signature_torch(Xp, 3, basepoint=True)
                        sy = oracle.signature_torch(Yp, 3, basepoint=True)
                        sq = (sx.pow(2).sum(1, keepdim=True) + sy.pow(2).sum(1)[None] - 2 * sx @ sy.t()).clamp(min=0)
                        h = (torch.median(sq) / torch.tensor(sq.shape[0] + 1.0).log()).sqrt().clamp_min(1e-8)
                        return torch.exp(-0.5 / h ** 2 * sq)

                opt_cpu = make(torch.device("cpu"), CpuKernel())
                x0c = torch.tensor([-1.8, -1.8, 0.0, 0.0])
                opt_cpu(x0c)
                t0 = time.perf_counter()
                for _ in range(5):
                    opt_cpu(x0c)
                out["cpu_port_ms_per_mpc_step"] = (time.perf_counter() - t0) / 5 * 1e3
            return out
        except Exception as e:
            return {"error": f"{type(e).__name__}: {e}"[:300]}

    # ---- headline ----------------------------------------------------------------------------
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ms_dev = timed(step_device, args.steps, max(args.warmup, 3))
    clocks = sampler.stop() if rank == 0 else None
    ms_copy = timed(step_device_copy_in, args.steps, 3)
    ms_e2e = timed(step_e2e, args.steps, max(args.warmup, 3))
    n_launch, launch_names = count_kernels(step_device)  # counted on one untimed update
    launches_note = "kernels of one update counted by the profiler (CUPTI) x steps x ranks"
    if n_launch is None:
        n_launch, launches_note = 7, "profiler unavailable: 7 = kernels sigsvgd_phi enqueues on this path"

    # ---- rooflines: every kernel of the update timed alone (CUDA events, L2 flushed) ------------
    roof = None
    if rank == 0:
        use_tc = ops.tc_available() and N >= 256  # what the timed step actually ran

        def alone(fn, n=10):
            return timed(fn, n, 3, reduce_max=False)

        Xa, Ga = make_inputs(N)
        Xa, Ga = Xa.to(dev), Ga.to(dev)
        kernels = []
        sig, pref = ops.signature_fwd_prefix_raw(Xa, DEPTH, True)
        ms_k1 = alone(lambda: ops.signature_fwd_prefix_raw(Xa, DEPTH, True))
        by_k1 = 4.0 * N * (P + D)
        kernels.append({"kernel": "sig3_fwd_kernel (signature forward, K1)", "bound": "hbm", "ms_per_launch": ms_k1,
                        "algorithmic_bytes": by_k1, "achieved": by_k1 / (ms_k1 * 1e-3) / 1e9, "peak": peak_hbm,
                        "unit": "GB/s", "frac": by_k1 / (ms_k1 * 1e-3) / 1e9 / peak_hbm})
        if use_tc:
            planes, norms, scales, sc = ops.tc_pack_fused_raw(sig, Ga, True)
            d_rows = ops.tc_gram_raw(planes, norms, planes, norms, D, scales, True, out=ops.dist_buffer(N, N, dev))
            ms_gram = alone(lambda: ops.tc_gram_raw(planes, norms, planes, norms, D, scales, True, out=d_rows))
            fl_gram = 2.0 * N * N * D  # algorithmic: the full distance matrix (the kernel computes the upper triangle)
            kernels.append({"kernel": "gram_kernel (tcgen05 split-fp16, symmetric squared distances, K3)", "bound": "tensor",
                            "ms_per_launch": ms_gram, "algorithmic_flops": fl_gram,
                            "achieved": fl_gram / (ms_gram * 1e-3) / 1e12, "peak": peak_tf, "unit": "TFLOP/s",
                            "frac": fl_gram / (ms_gram * 1e-3) / 1e12 / peak_tf,
                            "executed_mma_flops": 3 * fl_gram / 2 * (1 + 1.0 / (N / 128))})
            st = ops.new_state(dev)
            ws = ops.tc_median_workspace(N, N, dev)

            def median():
                ops.tc_median_prepare_raw(sig, sig, ws, st)
                ops.tc_median_select_raw(ws, d_rows, N, ops.Bandwidth.median(), st, symmetric=True)

            ms_med = alone(median)
            by_med = 2.0 * N * N  # the upper triangle of the fp32 matrix, read once
            kernels.append({"kernel": "window_prepare + window_select (exact median, K4)", "bound": "hbm",
                            "ms_per_launch": ms_med, "algorithmic_bytes": by_med,
                            "achieved": by_med / (ms_med * 1e-3) / 1e9, "peak": peak_hbm, "unit": "GB/s",
                            "frac": by_med / (ms_med * 1e-3) / 1e9 / peak_hbm})
            dominant = lambda: ops.tc_kf_raw(d_rows, planes, D + 1 + P, scales, st)  # noqa: E731
            kname = "kf_fused_kernel (tcgen05 cta_group::2 split-fp16, exp(gamma d).[Sc|1|score], RBF map in the operand path, K5)"
            out_kf = dominant()
            gsig = ops.sig_grad_raw(sc, sc, out_kf[:, :D], out_kf[:, D], st)
            add = out_kf[:, D + 1:]
        else:
            v_all = ops.pack_features(sig, ops.col_mean(sig), Ga)
            k_rows = ops.sqdist_raw(v_all[:, P:], v_all[:, P:], D)
            st = ops.bandwidth_raw(k_rows.clone(), ops.Bandwidth.median(), N)
            ops.rbf_apply_raw(k_rows, st)
            dominant = lambda: ops.matmul_nn_raw(k_rows, v_all)  # noqa: E731
            kname = "gemm_nn_simt_kernel (fp32 SIMT, K.[score|S|1])"
            out_kf = dominant()
            gsig = ops.sig_grad_raw(v_all[:, P:], v_all[:, P:], out_kf[:, P:P + D], out_kf[:, P + D], st)
            add = out_kf[:, :P]
        ms_k = alone(dominant)
        flops = 2.0 * N * N * (P + D)  # algorithmic: K.score and K.S (the row-sum column is free)
        ms_k2 = alone(lambda: ops.signature_bwd_prefix_raw(Xa, sig, pref, gsig, DEPTH, True, add=add, alpha=1.0 / N,
                                                          beta=-1.0 / N))
        by_k2 = 4.0 * N * (3 * P + 2 * D)
        kernels.append({"kernel": "sig3_bwd_seg_kernel (signature VJP + phi epilogue, K2)", "bound": "hbm",
                        "ms_per_launch": ms_k2, "algorithmic_bytes": by_k2, "achieved": by_k2 / (ms_k2 * 1e-3) / 1e9,
                        "peak": peak_hbm, "unit": "GB/s", "frac": by_k2 / (ms_k2 * 1e-3) / 1e9 / peak_hbm})
        traffic, traffic_src = None, None
        try:
            tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))["kf_kernel"]
            if use_tc and N == 4096:
                traffic, traffic_src = tj["dram_bytes_per_launch"], "NOT measured by this run: " + tj.get("source", "profiles/")
        except Exception:
            pass
        roof = {"bound": "tensor", "kernel": kname, "mma_passes_per_fp32_product": 3 if use_tc else None,
                "achieved": flops / (ms_k * 1e-3) / 1e12, "peak": peak_tf,
                "peak_source": peak_src + " bf16_tflops (burst; fp16 and bf16 share the tcgen05 kind::f16 rate)",
                "unit": "TFLOP/s", "frac": flops / (ms_k * 1e-3) / 1e12 / peak_tf, "traffic": traffic,
                "traffic_source": traffic_src, "traffic_unit": "DRAM bytes per launch; algorithmic operand bytes are 96 MB (fp32 distances 67, fp16 planes 15, fp32 output 15)",
                "ms_per_launch": ms_k, "algorithmic_flops_per_launch": flops,
                "step_algorithmic_flops": 2.0 * N * N * (2 * D + P),
                "step_frac_of_peak": 2.0 * N * N * (2 * D + P) / (ms_dev * 1e-3) / 1e12 / peak_tf,
                "other_kernels": kernels}
        if use_tc:
            # fp32-grade products cost three fp16 MMA passes (hi.hi, hi.lo, lo.hi): executed tensor flops
            # are 3x the algorithmic ones, so `frac` cannot exceed 1/3
            roof["frac_cap"] = 1.0 / 3.0
            roof["executed_frac_of_peak"] = 3.0 * roof["frac"]
        del Xa, Ga
        torch.cuda.empty_cache()

    # ---- extras: one swarm sharded over the ranks, batched scenes, the MPC step -----------------
    sharded, scenes, mpc, deep, pde, occ, rcost = [], None, None, None, None, None, None
    if not args.skip_extras:
        if world > 1:
            sharded.append(sharded_swarm(N, min(args.steps, 20), 3))
        if args.large_particles > 0:
            sharded.append(sharded_swarm(args.large_particles, 10, 3))
        scenes = batched_scenes(1024, 128, 10, 3)
        if rank == 0:
            mpc = mpc_step(20, 5)
            deep = depth4_swarm(N, 10, 3)
            pde = sigpde_leg(5, 2)
            occ = occupancy_leg(10, 3)
            rcost = robot_cost_leg()

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
            "ms_per_step_with_input_copy": ms_copy,
            "e2e": {"value": world * 1e3 / ms_e2e, "unit": UNIT,
                    "h2d_bytes_per_step": world * int(Xh.numel() + Gh.numel()) * 4,
                    "d2h_bytes_per_step": world * int(phi_host.numel()) * 4, "ms_per_step": ms_e2e},
            "gpu_launches": world * n_launch * args.steps,
            "gpu_launches_note": launches_note, "kernels_per_update": launch_names,
            "sharded_swarm": sharded,
            "batched_scenes": scenes,
            "mpc_step": mpc,
            "depth4_swarm": deep,
            "sigpde": pde,
            "occupancy_query": occ,
            "robot_cost": rcost,
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

#!/bin/bash
# ncu --set full of the non-tensor kernels of the update (one launch each, after warm-up)
set -x
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'sig3_fwd_kernel|sig3_bwd_seg|rbf_planes|window_select|pack_stats' -s 10 -c 5 -o gpurun_out/prof_rest -f python bench.py --steps 3 --warmup 3 --no-cpu-baseline --large-particles 0 --no-graph > gpurun_out/ncu_rest.log 2>&1
tail -n 2 gpurun_out/ncu_rest.log

#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'window_select' -s 4 -c 1 -o gpurun_out/wsel -f python bench.py --steps 3 --warmup 3 --no-cpu-baseline --large-particles 0 > gpurun_out/ncu_wsel.log 2>&1
tail -n 3 gpurun_out/ncu_wsel.log

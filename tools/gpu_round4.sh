#!/bin/bash
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
python tools/bench_pieces.py 4096 > gpurun_out/pieces.json 2>&1; cat gpurun_out/pieces.json
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'sig3_' -s 4 -c 2 \
    -o gpurun_out/prof_sig -f python bench.py --steps 3 --warmup 3 --no-cpu-baseline --large-particles 0 > gpurun_out/ncu_sig.log 2>&1

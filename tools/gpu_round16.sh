#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -n 3 gpurun_out/pytest_gpu.log
timeout 300 python tools/bench_pieces.py 4096 > gpurun_out/pieces.json 2>&1; grep -E "select_sym|rbf|pack|gram|phi_total" gpurun_out/pieces.json
timeout 600 python bench.py --steps 30 --no-cpu-baseline --large-particles 0 > gpurun_out/bench1.json 2> gpurun_out/bench1.err; python -c "
import json; j=json.load(open('gpurun_out/bench1.json')); print(j['ms_per_step'], j['e2e']['ms_per_step'], j['roofline']['ms_per_launch'])"

#!/bin/bash
# signature time-parallel pair: tests first, then the update's timings
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "signature" > gpurun_out/pytest_sig.log 2>&1; tail -n 5 gpurun_out/pytest_sig.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -n 3 gpurun_out/pytest_gpu.log
timeout 300 python tools/bench_pieces.py 4096 > gpurun_out/pieces.json 2>&1; grep -E "tc_matmul|phi_total|sig_" gpurun_out/pieces.json
timeout 300 python tools/configs_time.py > gpurun_out/configs_time.log 2>&1; tail -n 12 gpurun_out/configs_time.log
timeout 600 python bench.py --steps 30 --no-cpu-baseline --large-particles 0 > gpurun_out/bench1.json 2> gpurun_out/bench1.err; python -c "
import json; j=json.load(open('gpurun_out/bench1.json')); print(j['ms_per_step'], j['e2e']['ms_per_step'], j['roofline']['ms_per_launch'])"

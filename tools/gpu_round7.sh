#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_tc.py -x -q > gpurun_out/pytest_tc.log 2>&1; tail -n 3 gpurun_out/pytest_tc.log
timeout 300 python tools/kf_exp.py 4096 2>&1 | tee gpurun_out/kf_exp.log
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -n 3 gpurun_out/pytest_gpu.log
timeout 300 python tools/bench_pieces.py 4096 > gpurun_out/pieces.json 2>&1; grep -E "tc_matmul|phi_total" gpurun_out/pieces.json

#!/bin/bash
set -x
mkdir -p gpurun_out
python -m pytest tests/test_gpu_dist.py -x -q > gpurun_out/pytest_dist.log 2>&1; tail -5 gpurun_out/pytest_dist.log
timeout 600 python tools/large_check.py 8192 32768 65536 > gpurun_out/large_check.log 2>&1; cat gpurun_out/large_check.log
timeout 900 python bench.py --steps 20 > gpurun_out/bench1.json 2> gpurun_out/bench1.err; cat gpurun_out/bench1.json; tail -3 gpurun_out/bench1.err

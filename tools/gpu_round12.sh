#!/bin/bash
set -x
mkdir -p gpurun_out
SIGSVGD_KF_PAIRS=1 timeout 300 python -m pytest tests/test_gpu_tc.py -x -q > gpurun_out/pytest_tc_pairs.log 2>&1; tail -n 15 gpurun_out/pytest_tc_pairs.log
timeout 300 python tools/kf_exp.py 4096 2>&1 | tee gpurun_out/kf_exp.log
timeout 300 python tools/kf_exp.py 16384 2>&1 | tee -a gpurun_out/kf_exp.log

#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "signature" > gpurun_out/pytest_sig.log 2>&1; tail -n 3 gpurun_out/pytest_sig.log
timeout 300 python tools/sig_time.py > gpurun_out/sig_time.json 2>&1; cat gpurun_out/sig_time.json

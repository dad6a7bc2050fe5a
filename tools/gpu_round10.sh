#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -n 5 gpurun_out/pytest_gpu.log
timeout 300 python tools/strip_pieces.py 65536 8 > gpurun_out/strip_pieces.json 2>&1; cat gpurun_out/strip_pieces.json
timeout 600 python tools/large_check.py 65536 > gpurun_out/large_check.log 2>&1; cat gpurun_out/large_check.log

#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 600 python tools/kf_exp.py 4096 > gpurun_out/kf_exp.log 2>&1; cat gpurun_out/kf_exp.log
SIGSVGD_KF_EQUAL_TILES=1 timeout 600 python -m pytest tests/test_gpu_tc.py tests/test_gpu_parity.py -m gpu -x -q -k "tc or phi or full_size" > gpurun_out/pytest_eq.log 2>&1; tail -n 3 gpurun_out/pytest_eq.log
SIGSVGD_KF_EQUAL_TILES=1 timeout 600 python bench.py --steps 30 --no-cpu-baseline --large-particles 0 2>/dev/null | python -c "
import json,sys; j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(j['ms_per_step'], j['e2e']['ms_per_step'], j['roofline']['ms_per_launch'])"

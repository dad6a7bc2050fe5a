#!/bin/bash
# One GPU call: parity tests, smoke, bench, ncu launch list, ncu --set full of the tensor-core kernels.
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; cat gpurun_out/bench.json
python tools/bench_pieces.py 4096 > gpurun_out/pieces.json 2>&1; cat gpurun_out/pieces.json
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --large-particles 0 > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 3 --warmup 3 --no-cpu-baseline --large-particles 0 > gpurun_out/ncu_launches.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'kf_pair|gram_kernel' -s 6 -c 4 --launch-skip-before-match 0 \
    -o gpurun_out/prof_tc -f python bench.py --steps 3 --warmup 3 --no-cpu-baseline --large-particles 0 > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out

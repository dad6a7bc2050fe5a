#!/bin/bash
set -x
mkdir -p gpurun_out
nvidia-smi -L
python -m pytest tests/test_gpu_dist.py -x -q > gpurun_out/pytest_dist2.log 2>&1; tail -5 gpurun_out/pytest_dist2.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench2.json 2> gpurun_out/bench2.err; cat gpurun_out/bench2.json; tail -5 gpurun_out/bench2.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --impl reference --steps 2 --warmup 1 > gpurun_out/bench2_ref.json 2> gpurun_out/bench2_ref.err; cat gpurun_out/bench2_ref.json

#!/bin/bash
set -x
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/bench8.json 2> gpurun_out/bench8.err; cat gpurun_out/bench8.json; tail -n 5 gpurun_out/bench8.err

#!/bin/bash
# 2-GPU box: dist tests over NCCL, the multi-rank bench line (stdout must be ONE JSON line), the reference arm
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_dist.py tests/test_gpu_host.py -m gpu -x -q > gpurun_out/pytest_dist2.log 2>&1; tail -n 3 gpurun_out/pytest_dist2.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench2.json 2> gpurun_out/bench2.err; wc -l gpurun_out/bench2.json; python -c "
import json; j=json.load(open('gpurun_out/bench2.json')); print(j['value'], j['ms_per_step'], j['e2e']['value'], j['sharded_swarm'])"; tail -n 3 gpurun_out/bench2.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/bench2_ref.json 2> gpurun_out/bench2_ref.err; cat gpurun_out/bench2_ref.json | cut -c1-400

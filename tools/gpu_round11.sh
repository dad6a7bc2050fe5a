#!/bin/bash
set -x
mkdir -p gpurun_out
python -m pytest tests/test_gpu_dist.py -x -q > gpurun_out/pytest_dist2.log 2>&1; tail -n 3 gpurun_out/pytest_dist2.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench2.json 2> gpurun_out/bench2.err; python -c "
import json; j=json.load(open('gpurun_out/bench2.json')); print(j['value'], j['e2e']['value'], j['sharded_swarm'])"; tail -n 3 gpurun_out/bench2.err

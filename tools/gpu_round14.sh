#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'sig3_bwd_seg' -s 6 -c 1 -o gpurun_out/sigbwd -f python tools/sig_time.py 4096 > gpurun_out/ncu_sig.log 2>&1
tail -n 3 gpurun_out/ncu_sig.log
ls -la gpurun_out/*.ncu-rep

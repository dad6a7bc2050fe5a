#!/bin/bash
# Evidence of one round on one B200 (run under gpurun from the repo root): GPU tests, smoke, the bench
# line, the ncu launch list of the same bench command, one `ncu --set full` capture of every kernel of
# the update, and the depth-4 / piece timings.  Everything lands in gpurun_out/<tag>_*.
tag=${1:-r02}
o=gpurun_out
mkdir -p $o
timeout 900 python -m pytest tests -m gpu -q > $o/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $o/${tag}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $o/${tag}_smoke.log 2>&1; tail -1 $o/${tag}_smoke.log
timeout 900 python bench.py > $o/${tag}_bench_1gpu.json 2> $o/${tag}_bench_1gpu.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $o/${tag}_bench_reference.json 2> $o/${tag}_bench_reference.err; echo "reference rc=$?"
timeout 300 python bench.py --steps 2 --warmup 1 --skip-extras --no-cpu-baseline > /dev/null 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/${tag}_ncu_launches.csv \
    python bench.py --steps 2 --warmup 1 --skip-extras --no-cpu-baseline > $o/${tag}_ncu_launches.log 2>&1
python tools/ncu_summary.py $o/${tag}_ncu_launches.csv > $o/${tag}_ncu_launches_summary.txt 2>&1
timeout 120 python tools/run_phi.py 4096 3 > /dev/null 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -s 16 -c 8 -o $o/${tag}_full \
    python tools/run_phi.py 4096 3 > $o/${tag}_ncu_full.log 2>&1
timeout 300 python tools/depth4_check.py 4096,16384 4 > $o/${tag}_depth4.json 2> $o/${tag}_depth4.err
timeout 300 python tools/bench_pieces.py 4096 > $o/${tag}_pieces.json 2> $o/${tag}_pieces.err
timeout 120 python tools/host_timeline.py 4096 > $o/${tag}_host_timeline.txt 2>&1
ls -la $o | tail -20

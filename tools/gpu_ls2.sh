#!/bin/bash
# Two-halves LS kernel: parity (plain tests) and A/B bench against the one-set kernel.
mkdir -p gpurun_out
SV_LS_HALVES=2 timeout 1200 python -m pytest tests -m gpu -q --timeout 300 -x -k "plain" > gpurun_out/pytest_ls2.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_ls2.log
tail -5 gpurun_out/pytest_ls2.log
SV_LS_HALVES=2 timeout 600 python bench.py --kernel ls --steps 5 --no-cpu > gpurun_out/bench_ls2.log 2>&1; tail -c 1500 gpurun_out/bench_ls2.log
timeout 600 python bench.py --kernel ls --steps 5 --no-cpu > gpurun_out/bench_ls1.log 2>&1; tail -c 1500 gpurun_out/bench_ls1.log
CMD="python bench.py --kernel ls --steps 1 --warmup 1 --no-cpu"
SV_LS_HALVES=2 timeout 900 ncu --set full --clock-control none --import-source on -k regex:sv_ls -s 1 -c 1 -o gpurun_out/prof_ls2 $CMD > gpurun_out/ncu_ls2.log 2>&1
tail -2 gpurun_out/ncu_ls2.log

#!/bin/bash
# LS kernel iteration: parity (plain tests), bench both workloads, one ncu --set full capture with source.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --timeout 300 -x -k "plain" > gpurun_out/pytest_ls.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_ls.log
tail -5 gpurun_out/pytest_ls.log
timeout 600 python bench.py --kernel ls --steps 5 --no-cpu > gpurun_out/bench_ls.log 2>&1; tail -c 1500 gpurun_out/bench_ls.log
timeout 600 python bench.py --kernel ls --workload profile2405 --steps 3 --no-cpu > gpurun_out/bench_ls_profile.log 2>&1; tail -c 700 gpurun_out/bench_ls_profile.log
CMD="python bench.py --kernel ls --steps 1 --warmup 1 --no-cpu"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sv_ls -s 1 -c 1 -o gpurun_out/prof_ls $CMD > gpurun_out/ncu_ls.log 2>&1
tail -2 gpurun_out/ncu_ls.log

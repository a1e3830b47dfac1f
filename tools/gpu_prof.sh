#!/bin/bash
# usage: gpu_prof.sh <workload>  -- parity (plain tests), bench, ncu --set full capture of the LS kernel
W=${1:-synthetic2405}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --timeout 300 -x -k "plain" > gpurun_out/pytest_ls.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_ls.log
tail -3 gpurun_out/pytest_ls.log
CMD="python bench.py --kernel ls --workload $W --steps 3 --warmup 1 --no-cpu"
timeout 600 $CMD > gpurun_out/bench_$W.log 2>&1; grep -o '"relax_per_clk_per_sm": [0-9.]*' gpurun_out/bench_$W.log; grep -o '"step_ms": \[[^]]*\]' gpurun_out/bench_$W.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sv_ls -s 1 -c 1 -o gpurun_out/prof_$W $CMD > gpurun_out/ncu_$W.log 2>&1
tail -1 gpurun_out/ncu_$W.log

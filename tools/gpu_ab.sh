#!/bin/bash
# LS kernel A/B: plain parity tests + bench of both 2405-state workloads (no ncu).
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --timeout 300 -x -k "plain" > gpurun_out/pytest_ls.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_ls.log
tail -3 gpurun_out/pytest_ls.log
timeout 600 python bench.py --kernel ls --steps 5 --no-cpu > gpurun_out/bench_ls.log 2>&1; grep -o '"step_ms": \[[^]]*\]' gpurun_out/bench_ls.log
timeout 600 python bench.py --kernel ls --workload profile2405 --steps 3 --no-cpu > gpurun_out/bench_ls_profile.log 2>&1; grep -o '"step_ms": \[[^]]*\]' gpurun_out/bench_ls_profile.log

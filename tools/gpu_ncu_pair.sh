#!/bin/bash
# ncu --set full of K3'' on both 2405-state models (launch 2 = real model, warm; launch 8 = synthetic)
mkdir -p gpurun_out
CMD="python tools/bench_pair.py"
timeout 600 $CMD > gpurun_out/pair_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pair_rows -o gpurun_out/prof_pair $CMD > gpurun_out/ncu_pair.log 2>&1
tail -2 gpurun_out/ncu_pair.log

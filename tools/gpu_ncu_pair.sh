#!/bin/bash
mkdir -p gpurun_out
CMD="python tools/bench_spec.py 2405"
timeout 600 $CMD > gpurun_out/spec_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pair_rows -s 2 -c 2 -o gpurun_out/prof_pair $CMD > gpurun_out/ncu_pair.log 2>&1
tail -2 gpurun_out/ncu_pair.log

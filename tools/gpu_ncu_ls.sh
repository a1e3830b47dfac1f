#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --kernel ls --steps 1 --no-cpu"
timeout 600 $CMD > gpurun_out/plain_ls.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:sv_ls -s 3 -c 1 -o gpurun_out/prof_ls $CMD > gpurun_out/ncu_ls.log 2>&1
tail -3 gpurun_out/ncu_ls.log

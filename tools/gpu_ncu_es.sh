#!/bin/bash
# ncu --set full of the ES kernel on the latency configs (configs[1] then configs[3])
mkdir -p gpurun_out
CMD="python tools/bench_configs.py"
timeout 600 $CMD > gpurun_out/configs.log 2>&1; cat gpurun_out/configs.log | cut -c1-400
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sv_es -s 3 -c 1 -o gpurun_out/prof_es $CMD > gpurun_out/ncu_es.log 2>&1
tail -2 gpurun_out/ncu_es.log

#!/bin/bash
mkdir -p gpurun_out
for h in 0 1 2; do
SV_LS_MGR_HANDICAP=$h timeout 600 python bench.py --kernel ls --steps 3 --no-cpu 2>/dev/null | grep -o '"step_ms": \[[^]]*\]' | sed "s/^/handicap $h /"
done

#!/bin/bash
mkdir -p gpurun_out
for it in 1000000 6000000 30000000; do
SV_LS_SA_ITERS=$it timeout 900 python bench.py --kernel ls --steps 3 --no-cpu > gpurun_out/bench_sa_$it.log 2>&1
python - <<PY
import json
l=[x for x in open('gpurun_out/bench_sa_$it.log') if x.startswith('{')][-1]; j=json.loads(l)
print($it, j['ms_per_step'], j['kernel']['ls_gather_wavefronts'], j['roofline']['relax_per_clk_per_sm'])
PY
done

#!/bin/bash
mkdir -p gpurun_out
for kb in 0 56 72 110; do
SV_K4P_SMEM_KB=$kb timeout 600 python tools/bench_spec.py 2405 quick > gpurun_out/p.json 2> gpurun_out/p.err; python - <<PY
import json; d=json.load(open("gpurun_out/p.json")); print("smem_kb=$kb cluster", round(d["run_ms_level2_implicit_cluster"],2), "step", round(d["run_ms_level2_implicit"],2), d["bits_equal_level2_implicit_cluster"])
PY
tail -2 gpurun_out/p.err
done

#!/bin/bash
mkdir -p gpurun_out
for g in 1 2 4; do for r in 1 4; do
SV_IMPLICIT_GROUPS=$g SV_IMPLICIT_ROWS=$r timeout 600 python tools/bench_spec.py 2405 quick > gpurun_out/i2.json 2> gpurun_out/i2.err; python - <<PY
import json; d=json.load(open("gpurun_out/i2.json")); print("groups=$g rows=$r", {k:round(v,2) for k,v in d.items() if k.startswith("run_ms")})
PY
done; done
SV_IMPLICIT_GROUPS=1 SV_IMPLICIT_ROWS=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:implicit -c 60 --csv --log-file gpurun_out/i2_launches.csv python tools/bench_spec.py 2405 quick > gpurun_out/i2_ncu.log 2>&1
tail -25 gpurun_out/i2_launches.csv | cut -d, -f5,12- | cut -c1-200

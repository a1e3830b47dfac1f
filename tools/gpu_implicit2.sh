#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_spec.py -m gpu -x -q --timeout 600 2>&1 | tail -3
for g in 1 4; do for r in 1 2 4; do
SV_IMPLICIT_GROUPS=$g SV_IMPLICIT_ROWS=$r timeout 600 python tools/bench_spec.py 2405 quick > gpurun_out/i2.json 2> gpurun_out/i2.err; python - <<PY
import json; d=json.load(open("gpurun_out/i2.json")); print("groups=$g rows=$r", {k:round(v,2) for k,v in d.items() if k.startswith("run_ms")})
PY
done; done

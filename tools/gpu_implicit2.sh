#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_spec.py -m gpu -x -q --timeout 600 2>&1 | tail -3
SV_IMPLICIT_THREADS=64 SV_IMPLICIT_ROWS=4 timeout 600 python -m pytest tests/test_gpu_spec.py -m gpu -x -q -k implicit 2>&1 | tail -2
for t in 64 128; do for g in 1 4; do for r in 1 2 4; do
SV_IMPLICIT_THREADS=$t SV_IMPLICIT_GROUPS=$g SV_IMPLICIT_ROWS=$r timeout 600 python tools/bench_spec.py 2405 quick > gpurun_out/i2.json 2> gpurun_out/i2.err; python - <<PY
import json; d=json.load(open("gpurun_out/i2.json")); print("threads=$t groups=$g rows=$r", round(d["run_ms_level2_implicit"],2))
PY
done; done; done

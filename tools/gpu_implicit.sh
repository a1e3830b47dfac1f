#!/bin/bash
# K4i / K4p check: spec parity suite, then the config-3 numbers for each register blocking.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_spec.py -m gpu -x -q --timeout 600 2>&1 | tail -15 > gpurun_out/i_pytest.log; cat gpurun_out/i_pytest.log
for r in 1 2 4; do
SV_IMPLICIT_ROWS=$r timeout 600 python tools/bench_spec.py 2405 quick > gpurun_out/i_bench_spec_2405_r$r.json 2> gpurun_out/i_bench_spec.err; python - <<PY
import json; d=json.load(open("gpurun_out/i_bench_spec_2405_r$r.json")); print("rows=$r", {k:v for k,v in d.items() if "implicit" in k or k=="run_ms_level2"})
PY
tail -3 gpurun_out/i_bench_spec.err
done
SV_IMPLICIT_ROWS=4 timeout 600 python -m pytest tests/test_gpu_spec.py -m gpu -x -q -k implicit 2>&1 | tail -3

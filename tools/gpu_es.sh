#!/bin/bash
# ES kernel iteration: plain parity tests + the latency configs.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --timeout 300 -x -k "plain or harness" > gpurun_out/pytest_es.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_es.log
tail -3 gpurun_out/pytest_es.log
timeout 600 python tools/bench_configs.py > gpurun_out/configs.log 2>&1; python - <<'PY'
import json
for l in open('gpurun_out/configs.log'):
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l); print(d['config'][:52], 'kernel', d['kernel'], 'ms', round(d['ms_per_call_host_to_host'],3), 'us/step', round(d['us_per_step_critical_path'],3))
PY

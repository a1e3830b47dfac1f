#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --timeout 600 -k spec > gpurun_out/pytest_spec.log 2>&1; tail -1 gpurun_out/pytest_spec.log
timeout 300 python tools/bench_pair.py sparse > gpurun_out/bench_lhs.json 2> gpurun_out/bench_lhs.err; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_lhs.json').read().strip().splitlines()[-1])
for k,v in d.items(): print(k, round(v['ms'],3), 'ms', round(v['frac_of_hbm'],3))
PY

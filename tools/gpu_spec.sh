#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 600 -x -k "spec" > gpurun_out/pytest_spec.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_spec.log
tail -6 gpurun_out/pytest_spec.log
timeout 900 python tools/bench_spec.py 2405 > gpurun_out/bench_spec_2405.json 2> gpurun_out/bench_spec.err; cat gpurun_out/bench_spec_2405.json; tail -3 gpurun_out/bench_spec.err

#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 600 -x -k "spec" > gpurun_out/pytest_spec.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_spec.log
tail -4 gpurun_out/pytest_spec.log
timeout 600 python tools/bench_pair.py > gpurun_out/bench_pair.json 2> gpurun_out/bench_pair.err; cat gpurun_out/bench_pair.json; tail -3 gpurun_out/bench_pair.err

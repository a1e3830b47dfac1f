#!/bin/bash
# Full GPU pass: parity suite, smoke, default bench (+ reference arm), spec numbers, ncu launch lists.
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; tail -c 600 gpurun_out/bench_default.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2>&1; tail -c 400 gpurun_out/bench_reference.json
timeout 900 python tools/bench_spec.py 2405 > gpurun_out/bench_spec_2405.json 2> gpurun_out/bench_spec.err; cat gpurun_out/bench_spec_2405.json; tail -3 gpurun_out/bench_spec.err
CMD="python bench.py --steps 2 --no-cpu"
timeout 600 $CMD > gpurun_out/plain_launch.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_bench.csv $CMD > gpurun_out/ncu_l.log 2>&1

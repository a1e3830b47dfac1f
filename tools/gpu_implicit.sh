#!/bin/bash
# K4i check: spec parity suite, then the config-3 numbers.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_spec.py -m gpu -x -q --timeout 600 2>&1 | tail -15 > gpurun_out/i_pytest.log; cat gpurun_out/i_pytest.log
timeout 600 python tools/bench_spec.py 2405 > gpurun_out/i_bench_spec_2405.json 2> gpurun_out/i_bench_spec.err; cat gpurun_out/i_bench_spec_2405.json; tail -3 gpurun_out/i_bench_spec.err

#!/bin/bash
# Round-end pass: parity suite, smoke, default bench + reference arm, spec numbers, launch lists, ncu of K4i.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/f_pytest.log 2>&1; tail -3 gpurun_out/f_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/f_smoke.log 2>&1; tail -1 gpurun_out/f_smoke.log
timeout 600 python bench.py > gpurun_out/f_bench_default.json 2> gpurun_out/f_bench_default.err; tail -c 300 gpurun_out/f_bench_default.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/f_bench_reference.json 2> gpurun_out/f_ref.err; tail -c 300 gpurun_out/f_bench_reference.json
timeout 600 python tools/bench_spec.py 2405 > gpurun_out/f_bench_spec_2405.json 2> gpurun_out/f_spec.err; tail -c 200 gpurun_out/f_bench_spec_2405.json
CMD="python bench.py --steps 2 --no-cpu"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/f_launches_bench.csv $CMD > gpurun_out/f_ncu_l.log 2>&1
SV_IMPLICIT_GROUPS=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:implicit_pair_step -s 300 -c 1 -f -o gpurun_out/f_k4i python tools/bench_spec.py 2405 quick > gpurun_out/f_ncu_k4i.log 2>&1
ls -la gpurun_out/f_k4i.ncu-rep

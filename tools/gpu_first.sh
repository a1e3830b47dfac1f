#!/bin/bash
# First GPU pass: probe, parity tests, smoke, bench (generic + ES kernels).
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 300 python - > gpurun_out/probe.txt 2>&1 <<'PY'
import ctypes as C
from spec_viterbi_b200 import _lib
lib=_lib.lib()
names=["2 FADD","2 FMNMX","2 FMNMX3","2 FADD+FMNMX3","4 FADD+FMNMX3","LDS+2FADD+FMNMX"]
for mix in range(6):
    r=C.c_double(); mhz=C.c_double()
    rc=lib.sv_probe_fp32(0,mix,20000,C.byref(r),C.byref(mhz))
    print(mix,names[mix],rc,"lane-instr/s=%.4g"%r.value,"clk MHz=%.0f"%mhz.value,"per clk per SM=%.1f"%(r.value/148/(mhz.value*1e6) if mhz.value else 0))
PY
timeout 1500 python -m pytest tests -m gpu -q --timeout 180 > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1
echo "smoke exit $?" >> gpurun_out/smoke.log
timeout 300 python bench.py --workload small --steps 3 --no-cpu > gpurun_out/bench_small.log 2>&1
timeout 600 python bench.py --kernel generic --steps 3 --no-cpu > gpurun_out/bench_generic.log 2>&1
timeout 600 python bench.py --kernel es --steps 5 > gpurun_out/bench_es.log 2>&1
timeout 600 python bench.py --workload profile2405 --steps 3 --no-cpu > gpurun_out/bench_profile.log 2>&1
tail -5 gpurun_out/pytest_gpu.log; cat gpurun_out/probe.txt; tail -3 gpurun_out/smoke.log; tail -2 gpurun_out/bench_*.log

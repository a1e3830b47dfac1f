#!/bin/bash
mkdir -p gpurun_out
timeout 120 python - > gpurun_out/tmem_probe.txt 2>&1 <<'PY'
import ctypes as C
from spec_viterbi_b200 import _lib
lib=_lib.lib()
err=C.c_uint(); bw=C.c_double()
rc=lib.sv_probe_tmem(0, 20000, C.byref(err), C.byref(bw))
print("rc",rc,"errors",err.value,"tcgen05.ld bytes/clk/SM %.1f"%bw.value, lib.sv_last_error())
PY
cat gpurun_out/tmem_probe.txt
CMD="python bench.py --kernel es --steps 1 --no-cpu"
timeout 600 $CMD > gpurun_out/plain_r1a.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_r1a.csv $CMD > gpurun_out/ncu_launches.log 2>&1
timeout 600 $CMD > gpurun_out/plain_r1a2.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:sv_es -s 3 -c 1 -o gpurun_out/prof_r1a_es $CMD > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log; ls -la gpurun_out

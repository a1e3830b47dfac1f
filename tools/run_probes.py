#!/usr/bin/env python
"""Runs the micro-benchmark library (include/sv_probe_b200.h) and prints the measured ceilings the roofline
figures use: FP32 pipe rates, tensor-memory read rate, shared-memory pipe under the lockstep kernel's pattern."""
import ctypes as C
import json
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from spec_viterbi_b200 import _lib  # noqa: E402


def main():
    pl = _lib.probe_lib()
    v, w = C.c_double(), C.c_double()
    out = {"fp32": {}, "smem": {}}
    names = ["FADD", "FMNMX", "FMNMX3", "2 FADD + FMNMX3", "4 FADD + FMNMX3 (oracle-order pair)", "LDS.32 + 2 FADD + FMNMX",
             "FADD2 (packed f32x2, 2 adds each)", "2 FADD2 + FMNMX3 (oracle-order pair, packed)"]
    sms = None
    for mix, nm in enumerate(names):
        rc = pl.sv_probe_fp32(0, mix, 20000, C.byref(v), C.byref(w))
        if rc != 0:
            raise SystemExit(pl.sv_probe_last_error().decode())
        out["fp32"][nm] = {"lane_instr_per_s": v.value, "sm_clock_mhz": w.value}
    import torch
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    for nm, d in out["fp32"].items():
        d["lanes_per_clk_per_sm"] = d["lane_instr_per_s"] / (sms * d["sm_clock_mhz"] * 1e6)
    err = C.c_uint()
    pl.sv_probe_tmem(0, 20000, C.byref(err), C.byref(v))
    out["tmem"] = {"errors": err.value, "bytes_per_clk_per_sm": v.value}
    for mode, nm in ((1, "3 LDS.128 : 1 STS.128 (lockstep kernel's ratio)"),
                     (2, "the lockstep inner loop (gathers + stores + 8 FADD + 2 FMNMX3 per load)")):
        rc = pl.sv_probe_smem(0, mode, 30000, C.byref(v), C.byref(w))
        if rc != 0:
            raise SystemExit(pl.sv_probe_last_error().decode())
        out["smem"][nm] = {"bytes_per_clk_per_sm": v.value, "lane_loads_per_clk_per_sm": w.value}
    out["sms"] = sms
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()

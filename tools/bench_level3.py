#!/usr/bin/env python
"""Level-3 precompute (K3'' level-2 prefix table + K3' sparse-LHS level): 500.chmm, 8000 handlers of
502 x 502 floats = 8.06 GB, built into an already allocated table (wall clock, min of 3); run of
emit_50_3500_20 at level 3 checked against the plain path (argmin equal)."""
import json, sys, time
from pathlib import Path
import numpy as np
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from spec_viterbi_b200 import api, _lib  # noqa: E402

def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "500"
    hmm = api.read_HMM(str(ROOT / "tests/golden/models" / f"{name}.chmmb"))
    n, E = hmm.states_num, hmm.emit_num
    m = api.Model(hmm)
    sp = api.Spec(m, 3)
    ptr, tb, te = sp.table()
    ts = []
    for _ in range(3):
        t0 = time.perf_counter()
        sp2 = api.Spec(m, 3, _lib.SPEC_AUTO, tb, te, d_table=ptr)
        ts.append((time.perf_counter() - t0) * 1e3)
        sp2.close()
    gb = E ** 3 * n * n * 4 / 1e9
    seqs = api.read_emit_seq(str(ROOT / "tests/golden/seqs/emit_50_3500_20.essb"))
    syms, offs = api.pack_sequences(seqs)
    t0 = time.perf_counter()
    out, best, arg = sp.run_batch(syms, offs)
    run_ms = (time.perf_counter() - t0) * 1e3
    pout, _, _ = m.viterbi_batch(syms, offs)
    print(json.dumps({"model": name, "states": n, "level": 3, "handlers": E ** 3, "table_GB": gb,
                      "precompute_ms_into_caller_table": min(ts), "write_GBs": gb / (min(ts) * 1e-3),
                      "run_ms_level3": run_ms, "argmin_equal_plain": bool(np.array_equal(out.argmin(1), pout.argmin(1))),
                      "max_rel_vs_plain": float(np.max(np.abs(out - pout) / np.abs(pout)))}))

if __name__ == "__main__":
    main()

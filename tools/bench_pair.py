#!/usr/bin/env python
"""K3'' (pair_rows_kernel) alone: level-2 table of the real 2405.chmm and of the synthetic 2405-state
model, built into an already allocated table (wall clock around the synchronous C-ABI call, min of 5)."""
import json, sys, time
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from spec_viterbi_b200 import api, _lib, generators  # noqa: E402

def main():
    method = {"pair": _lib.SPEC_PAIR_ROWS, "sparse": _lib.SPEC_SPARSE_LHS}[sys.argv[1] if len(sys.argv) > 1 else "pair"]
    peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text()) if (ROOT / "MEASURED_PEAKS.json").exists() else {}
    hbm = peaks.get("hbm_gbs", 6546.2)
    res = {}
    for label in ("2405.chmm", "synthetic2405"):
        if label == "2405.chmm":
            hmm = api.read_HMM(str(ROOT / "tests/golden/models/2405.chmmb"))
        else:
            hmm = generators.random_hmm(states_num=2405, trans_per_state=3, num_non_zero_start_probs=2, seed=2405)
        n, E = hmm.states_num, hmm.emit_num
        m = api.Model(hmm)
        sp = api.Spec(m, 2, method)
        ptr, tb, te = sp.table()
        ts = []
        for _ in range(5):
            t0 = time.perf_counter()
            sp2 = api.Spec(m, 2, method, tb, te, d_table=ptr)
            ts.append((time.perf_counter() - t0) * 1e3)
            sp2.close()
        gb = E * E * n * n * 4 / 1e9
        res[label] = {"ms": min(ts), "all_ms": ts, "table_GB": gb, "write_GBs": gb / (min(ts) * 1e-3),
                      "frac_of_hbm": gb / (min(ts) * 1e-3) / hbm}
        sp.close()
        m.close() if hasattr(m, "close") else None
    print(json.dumps(res))

if __name__ == "__main__":
    main()

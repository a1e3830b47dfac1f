#!/usr/bin/env python
"""BASELINE.json configs 2 and 4 (the latency cases): few sequences, one persistent launch each.
Prints one JSON line per config: kernel chosen, ms per call (host buffers in, vectors out), us per time step."""
import json, sys, time
from pathlib import Path
import numpy as np
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from spec_viterbi_b200 import api, _lib  # noqa: E402

def run(cfg, model, seqfile, kernel=None):
    hmm = api.read_HMM(str(ROOT / "tests/golden/models" / f"{model}.chmmb"))
    seqs = api.read_emit_seq(str(ROOT / "tests/golden/seqs" / f"{seqfile}.essb"))
    syms, offs = api.pack_sequences(seqs)
    m = api.Model(hmm)
    if kernel is not None:
        m.set_kernel(kernel)
    ts = []
    for _ in range(7):
        t0 = time.perf_counter()
        out, best, arg = m.viterbi_batch(syms, offs)
        ts.append((time.perf_counter() - t0) * 1e3)
    ms = float(np.median(ts[2:]))
    lens = np.diff(offs.astype(np.int64))
    n = hmm.states_num
    print(json.dumps({"config": cfg, "model": model, "states": int(n), "nnz": int(m.info().nnz), "sequences": len(seqs),
                      "longest": int(lens.max()), "total_symbols": int(lens.sum()), "kernel": int(m.info().kernel), "seqs_per_cta": int(m.info().chain_seqs_per_cta or m.info().es_seqs_per_cta),
                      "ms_per_call_host_to_host": ms, "us_per_step_critical_path": ms * 1e3 / float(lens.max()),
                      "state_updates_per_s": float(lens.sum()) * n / (ms * 1e-3)}), flush=True)
    m.close()

if __name__ == "__main__":
    run("configs[1]: 1200.chmm x emit_3_3500_20, plain", "1200", "emit_3_3500_20")
    run("configs[3]: 2405.chmm x covid-19 (16 ragged sequences, longest 7096), plain", "2405", "covid-19")
    # the 1/2/4-way split question: the chain kernel on a quarter / half / the whole of the work per CTA
    run("600.chmm (128-thread CTA: a quarter of 2405's positions per sequence)", "600", "emit_3_3500_20")
    run("1200.chmm (256-thread CTA: half)", "1200", "emit_3_3500_20")
    run("2405.chmm (512-thread CTA)", "2405", "emit_3_3500_20")
    run("configs[1] on the edge-stationary kernel (round-1 choice)", "1200", "emit_3_3500_20", _lib.KERNEL_ES)
    run("configs[3] on the edge-stationary kernel (round-1 choice)", "2405", "covid-19", _lib.KERNEL_ES)
    run("configs[1] on the generic kernel", "1200", "emit_3_3500_20", _lib.KERNEL_GENERIC)
    run("configs[3] on the lockstep kernel", "2405", "covid-19", _lib.KERNEL_LS)

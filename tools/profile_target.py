#!/usr/bin/env python
"""A minimal launch target for ncu: builds one workload and runs its step kernel a few times.
usage: tools/profile_target.py [synthetic2405|profile2405|<model>.chmmb] [nseq] [T] [kernel: auto|es|ls|chain|generic] [reps]
(ncu replays every captured kernel ~40 times: keep nseq*T small enough, or capture with -k/-s/-c.)"""
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from spec_viterbi_b200 import api, generators, _lib  # noqa: E402


def main():
    wl = sys.argv[1] if len(sys.argv) > 1 else "synthetic2405"
    nseq = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
    T = int(sys.argv[3]) if len(sys.argv) > 3 else 7000
    kernel = sys.argv[4] if len(sys.argv) > 4 else "auto"
    reps = int(sys.argv[5]) if len(sys.argv) > 5 else 3
    if wl == "synthetic2405":
        hmm = generators.random_hmm(states_num=2405, seed=2405)
    elif wl == "profile2405":
        hmm = api.read_HMM(str(ROOT / "tests/golden/models/2405.chmmb"))
    else:
        hmm = api.read_HMM(str(ROOT / "tests/golden/models" / wl))
    syms, offs = generators.random_sequences(nseq, T, emit_range=int(hmm.emit_num), seed=7000, dtype=np.uint8)
    m = api.Model(hmm)
    m.set_kernel({"auto": _lib.KERNEL_AUTO, "es": _lib.KERNEL_ES, "ls": _lib.KERNEL_LS, "chain": _lib.KERNEL_CHAIN,
                  "generic": _lib.KERNEL_GENERIC}[kernel])
    m.prepare()
    seqs = m.upload(syms, offs)
    import torch
    out = torch.empty((nseq, hmm.states_num), dtype=torch.float32, device="cuda")
    for _ in range(reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        m.viterbi_device(seqs, out.data_ptr(), 0, 0, 0)
        torch.cuda.synchronize()
        print(f"{wl} nseq={nseq} T={T} kernel={m.info().kernel} q={m.info().chain_seqs_per_cta or m.info().ls_seqs_per_cta}: "
              f"{(time.perf_counter() - t0) * 1e3:.3f} ms", flush=True)


if __name__ == "__main__":
    main()

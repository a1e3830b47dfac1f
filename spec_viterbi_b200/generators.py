"""Seeded, parameterised restatements of the reference's random data generators.

Reference: chmm_files/chmm_gen.py (random sparse HMM: `trans_per_state` distinct uniform
destinations per source, every probability list = randrange(100) draws divided by their sum
and truncated to 6 characters, so exact zeros -- i.e. +inf weights -- occur) and
ess_files/ess_gen.py (uniform symbols).  The reference scripts hard-code their parameters
and do not seed `random`; here both are arguments, and a list whose draws sum to 0 is redrawn.
Generation is vectorised with numpy's PCG64, so the streams differ from CPython's `random`
but the distributions are the same.
"""
from __future__ import annotations

import numpy as np

from .api import HMM, to_modified_prob

RNG_RANGE = 100  # chmm_gen.py:16


def _probability_rows(rng: np.random.Generator, rows: int, length: int) -> np.ndarray:
    """rows x length probabilities: integer draws / row sum, truncated to 6 characters."""
    draws = rng.integers(0, RNG_RANGE, size=(rows, length))
    sums = draws.sum(axis=1)
    while (sums == 0).any():
        bad = np.flatnonzero(sums == 0)
        draws[bad] = rng.integers(0, RNG_RANGE, size=(bad.size, length))
        sums = draws.sum(axis=1)
    exact = draws / sums[:, None]
    # str(float)[:6] of a value in [0, 1]: "0.dddd" keeps 4 decimals (truncation, not rounding);
    # values printed in exponent form (< 1e-4, e.g. "5e-05") do not occur: the smallest
    # non-zero value is 1/(99*length) > 1e-4 for length <= 100.  "1.0" stays 1.0.
    trunc = np.floor(exact * 1e4 + 1e-9) / 1e4
    return trunc.astype(np.float32)  # the fp32 value the reader parses from that text


def random_hmm(states_num: int = 2405, trans_per_state: int = 3, num_non_zero_start_probs: int = 2,
               emit_range: int = 20, seed: int = 2405) -> HMM:
    rng = np.random.default_rng(seed)
    start_p = _probability_rows(rng, 1, num_non_zero_start_probs)[0]
    emis_p = _probability_rows(rng, states_num, emit_range)  # [state][symbol], as in the file
    # distinct destinations per source
    dst = np.empty((states_num, trans_per_state), dtype=np.uint64)
    for s in range(states_num):
        dst[s] = rng.choice(states_num, size=trans_per_state, replace=False)
    trans_p = _probability_rows(rng, states_num, trans_per_state)
    src = np.repeat(np.arange(states_num, dtype=np.uint64), trans_per_state)
    return HMM(
        states_num=states_num, emit_num=emit_range, trans_num=states_num * trans_per_state,
        trans_rows=src, trans_cols=dst.reshape(-1).astype(np.uint64),
        trans_probs=to_modified_prob(trans_p.reshape(-1)),
        emissions=np.ascontiguousarray(to_modified_prob(emis_p).T),
        non_zero_start_probs=num_non_zero_start_probs,
        start_probabilities_cols=np.arange(num_non_zero_start_probs, dtype=np.uint64),
        start_probabilities=to_modified_prob(start_p),
    )


def random_sequences(how_much_to_gen: int = 3, seq_size: int = 7000, emit_range: int = 20,
                     seed: int = 7000, dtype=np.uint8):
    """-> (symbols[dtype], offsets uint64[n+1]); equal lengths, uniform symbols."""
    rng = np.random.default_rng(seed)
    symbols = rng.integers(0, emit_range, size=how_much_to_gen * seq_size, dtype=np.int64).astype(dtype)
    offsets = (np.arange(how_much_to_gen + 1, dtype=np.uint64) * np.uint64(seq_size))
    return symbols, offsets


def write_chmm(hmm_probs: dict, path: str) -> None:
    """Writes a .chmm text file from plain probabilities (keys: states, start=[(state,p)],
    emissions[state][symbol], transitions=[(src,dst,p)]) in the reference's format
    (data_reader.cpp:7-15)."""
    with open(path, "w") as f:
        f.write(f"{hmm_probs['states']}\n{len(hmm_probs['start'])}\n")
        for s, p in hmm_probs["start"]:
            f.write(f"{s} {p}\n")
        em = hmm_probs["emissions"]
        f.write(f"{len(em[0])}\n")
        for row in em:
            f.write(" ".join(str(x) for x in row) + "\n")
        f.write(f"{len(hmm_probs['transitions'])}\n")
        for s, d, p in hmm_probs["transitions"]:
            f.write(f"{s} {d} {p}\n")


def write_ess(seqs, path: str, len_per_line: int = 100) -> None:
    """Writes sequences in the reference's .ess format (data_reader.cpp:81-91)."""
    with open(path, "w") as f:
        f.write(f"{len(seqs)}\n")
        for i, s in enumerate(seqs):
            s = [int(x) for x in s]
            f.write(f"{i} {len(s)}\n")
            for k in range(0, len(s), len_per_line):
                f.write(" ".join(str(x) for x in s[k:k + len_per_line]) + "\n")

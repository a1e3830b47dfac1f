"""Seeded restatements of chmm_gen.py / ess_gen.py."""
import numpy as np


def test_random_hmm_shape_and_distribution():
    from spec_viterbi_b200 import generators
    h = generators.random_hmm(states_num=2405, seed=2405)
    assert (h.states_num, h.trans_num, h.emit_num, h.non_zero_start_probs) == (2405, 7215, 20, 2)
    src = h.trans_rows.reshape(2405, 3)
    dst = h.trans_cols.reshape(2405, 3)
    assert np.all(src == np.arange(2405)[:, None])
    assert all(len(set(r)) == 3 for r in dst.tolist())  # distinct destinations (chmm_gen.py:50-57)
    assert list(h.start_probabilities_cols) == [0, 1]
    p = np.exp2(-h.trans_probs.astype(np.float64)).reshape(2405, 3)
    assert np.all(np.abs(p.sum(1) - 1) < 1e-3)  # truncated to 4 decimals
    # truncation creates exact zeros -> +inf weights, about 1%
    frac_inf = np.isposinf(h.trans_probs).mean()
    assert 0.001 < frac_inf < 0.03
    h2 = generators.random_hmm(states_num=2405, seed=2405)
    assert np.array_equal(h.trans_cols, h2.trans_cols) and np.array_equal(h.emissions, h2.emissions)


def test_random_sequences():
    from spec_viterbi_b200 import generators
    s, o = generators.random_sequences(5, 700, seed=1)
    assert s.dtype == np.uint8 and s.size == 3500 and s.max() < 20 and list(o) == [0, 700, 1400, 2100, 2800, 3500]
    c = np.bincount(s, minlength=20)
    assert c.min() > 100

"""K0 (host-side repack for the lockstep kernel) checked without a GPU: the schedule's register
images, replayed on the host exactly as a CTA consumes them, reproduce the oracle bit for bit;
the bank-colouring search leaves few conflicts."""
import numpy as np
import pytest

from conftest import same_bits


@pytest.mark.parametrize("name", ["100", "1200", "2405"])
def test_profile_models(name, api, oracle, load_model, load_seqs):
    hmm = load_model(name)
    seq = load_seqs("covid-19")[2]
    stats, out = api.k0_lockstep_check(hmm, seq)
    assert stats["valid"] == 1 and stats["heavy_rows"] == 2 and stats["split_states"] == 0
    assert same_bits(out, oracle.plain(hmm, seq))
    # band + dense column: conflict free by construction
    assert stats["wavefronts"] <= 1.02 * stats["ideal_wavefronts"]


@pytest.mark.parametrize("n,seed", [(3, 0), (150, 7), (2405, 2405)])
def test_random_models(n, seed, api, oracle):
    from spec_viterbi_b200 import generators
    hmm = generators.random_hmm(states_num=n, trans_per_state=min(3, n), num_non_zero_start_probs=min(2, n), seed=seed)
    seq = np.random.default_rng(seed).integers(0, 20, 120).astype(np.uint64)
    stats, out = api.k0_lockstep_check(hmm, seq)
    assert stats["valid"] == 1
    assert same_bits(out, oracle.plain(hmm, seq))
    assert stats["colour_cost_after"] <= stats["colour_cost_before"]
    assert stats["wavefronts"] <= 1.25 * stats["ideal_wavefronts"]


def test_high_in_degree_is_split(api, oracle):
    rng = np.random.default_rng(5)
    n = 300
    src = np.concatenate([np.arange(40), rng.integers(0, n, 900)]).astype(np.uint64)
    dst = np.concatenate([np.full(40, 7), rng.integers(0, n, 900)]).astype(np.uint64)
    w = rng.uniform(0.01, 9, src.size).astype(np.float32)
    hmm = api.HMM(n, 5, src.size, src, dst, w, rng.uniform(0.1, 6, (5, n)).astype(np.float32), 2,
                  np.array([4, 10], np.uint64), np.array([0.5, 2.0], np.float32))
    seq = rng.integers(0, 5, 80).astype(np.uint64)
    stats, out = api.k0_lockstep_check(hmm, seq)
    assert stats["valid"] == 1 and stats["split_states"] > 0
    assert same_bits(out, oracle.plain(hmm, seq))

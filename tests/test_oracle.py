"""Pins the CPU oracle (oracle/sv_oracle.c): reference golden vectors, the reference's own
sources (oracle/_ref, when built) and the committed fixtures generated from them."""
import numpy as np
import pytest

from conftest import PROFILE_MODELS, same_bits

# tests/test_helper.h:17-22 of the reference (probabilities; compared in -log2 space, abs tol 1.0)
REF_GOLDEN = [
    [25.6574, 24.4874, np.inf],
    [-np.log2(0.04608), -np.log2(0.10752)],
    [-np.log2(0.00882), -np.log2(0.02646)],
    [np.inf, -np.log2(2.82e-6), -np.log2(1.81e-5), -np.log2(6.05e-6)],
]


def ref_almost_equal(x, y):
    return (np.isposinf(x) and np.isposinf(y)) or abs(float(x) - float(y)) <= 1.0


@pytest.mark.parametrize("i", range(4))
def test_reference_known_answers_plain(i, oracle, load_model, load_seqs):
    v = oracle.plain(load_model(f"{i}_test_chmm"), load_seqs(f"{i}_test_seq")[0])
    assert len(v) == len(REF_GOLDEN[i])
    for a, b in zip(v, REF_GOLDEN[i]):
        assert ref_almost_equal(a, b)
        if np.isfinite(b):
            assert abs(a - b) < 5e-3  # goldens are rounded decimals


@pytest.mark.parametrize("i", range(4))
@pytest.mark.parametrize("level", [1, 2, 3])
def test_reference_known_answers_spec(i, level, oracle, load_model, load_seqs, golden):
    sp = oracle.spec(load_model(f"{i}_test_chmm"), level)
    v = sp.run(load_seqs(f"{i}_test_seq")[0])
    for a, b in zip(v, REF_GOLDEN[i]):
        assert ref_almost_equal(a, b)
    assert same_bits(v, golden[f"tiny{i}_spec{level}"])


@pytest.mark.parametrize("name", PROFILE_MODELS)
def test_plain_matches_reference_fixture(name, oracle, load_model, load_seqs, golden):
    v = oracle.plain(load_model(name), load_seqs("emit_3_3500_20")[0])
    assert same_bits(v, golden[f"{name}_emit3_3500_seq0_plain"])


def test_survey_anchors(oracle, load_model, load_seqs):
    v = oracle.plain(load_model("2405"), load_seqs("emit_3_3500_20")[0])
    assert np.all(np.isfinite(v)) and int(np.argmin(v)) == 0
    assert abs(v.min() - 15772.576) < 2e-3 and abs(v.max() - 15806.021) < 2e-3
    v = oracle.plain(load_model("100"), load_seqs("emit_3_7000_20")[0])
    assert abs(v.min() - 31834.846) < 4e-3


@pytest.mark.parametrize("k", [0, 4, 9])
def test_ragged_covid_fixture(k, oracle, load_model, load_seqs, golden):
    v = oracle.plain(load_model("2405"), load_seqs("covid-19")[k])
    assert same_bits(v, golden[f"2405_covid_seq{k}_plain"])


@pytest.mark.parametrize("name", ["100", "200"])
def test_spec2_matches_reference_fixture(name, oracle, load_model, load_seqs, golden):
    sp = oracle.spec(load_model(name), 2)
    for k in range(3):
        assert same_bits(sp.run(load_seqs("emit_3_3500_20")[k]), golden[f"{name}_emit3_3500_seq{k}_spec2"])


def test_config3_matches_reference_fixture(oracle, load_model, load_seqs, golden):
    """BASELINE.json configs[2] (2405.chmm x emit_50_3500_20, level 2): the fixture was produced by the
    reference's GraphBLAS_spec_impl itself (tools/make_fixtures.py --config3); handlers are built lazily
    here, ~35 s."""
    sp = oracle.spec(load_model("2405"), 2)
    assert same_bits(sp.run(load_seqs("emit_50_3500_20")[49]), golden["2405_emit50_3500_seq49_spec2"])


def test_spec1_is_bitwise_plain(oracle, load_model, load_seqs):
    hmm = load_model("300")
    seq = load_seqs("emit_3_3500_20")[1]
    assert same_bits(oracle.spec(hmm, 1).run(seq), oracle.plain(hmm, seq))


def test_batch_equals_single(oracle, load_model, load_seqs, api):
    hmm = load_model("100")
    seqs = load_seqs("covid-19")
    syms, offs = api.pack_sequences(seqs)
    out = oracle.plain_batch(hmm, syms, offs, threads=2)
    for k, s in enumerate(seqs):
        assert same_bits(out[k], oracle.plain(hmm, s))


def test_oracle_vs_reference_sources(oracle, load_model, load_seqs):
    """The reference's own GraphBLAS_* sources (compiled unmodified on the shim) agree bit for bit."""
    from oracle import pyoracle
    if not pyoracle.reference_available():
        pytest.skip("oracle/_ref not built (reference tree absent)")
    ref = pyoracle.Reference()
    seqs = load_seqs("covid-19")
    for name in ("100", "600"):
        hmm = load_model(name)
        for k in (0, 2, 5):
            assert same_bits(ref.plain(hmm, seqs[k]), oracle.plain(hmm, seqs[k]))
    hmm = load_model("100")
    for level in (1, 2):
        sp = ref.spec_create(hmm, level)
        osp = oracle.spec(hmm, level)
        for k in (0, 2, 4):
            assert same_bits(ref.spec_run(sp, seqs[k]), osp.run(seqs[k]))
        ref.spec_destroy(sp)


def test_oracle_random_model_vs_reference(oracle):
    """chmm_gen-style model: explicit +inf weights, two start states, unreachable handling."""
    from oracle import pyoracle
    from spec_viterbi_b200 import generators
    hmm = generators.random_hmm(states_num=150, seed=7)
    syms, offs = generators.random_sequences(3, 60, seed=3, dtype=np.uint64)
    assert np.isposinf(hmm.trans_probs).any() or np.isposinf(hmm.emissions).any()
    if not pyoracle.reference_available():
        pytest.skip("oracle/_ref not built")
    ref = pyoracle.Reference()
    for k in range(3):
        s = syms[int(offs[k]):int(offs[k + 1])]
        assert same_bits(ref.plain(hmm, s), oracle.plain(hmm, s))
    sp = ref.spec_create(hmm, 2)
    osp = oracle.spec(hmm, 2)
    s = syms[:60]
    assert same_bits(ref.spec_run(sp, s), osp.run(s))
    ref.spec_destroy(sp)

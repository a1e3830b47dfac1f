"""GPU parity, plain path (K1/K2 kernels through the C ABI) -- bit-exact against the oracle and
the committed reference outputs.  Reference: GraphBLAS_impl.cpp:4-93; harness protocol of
tests/test_semantic_equality.cpp:19-98 and tests/test_helper.h:44-57."""
import numpy as np
import pytest

from conftest import PROFILE_MODELS, best_state, same_bits

pytestmark = pytest.mark.gpu

KERNELS = ["generic", "es", "ls"]


def _model(api, hmm, kernel):
    from spec_viterbi_b200 import _lib
    m = api.Model(hmm)
    if kernel == "generic":
        m.set_kernel(_lib.KERNEL_GENERIC)
    elif kernel == "es":
        if m.info().es_threads == 0:
            pytest.skip("model does not fit the edge-stationary kernel")
        m.set_kernel(_lib.KERNEL_ES)
    elif kernel == "ls":
        if m.info().ls_threads == 0:
            pytest.skip("model does not fit the lockstep kernel")
        m.set_kernel(_lib.KERNEL_LS)
    return m


@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("i", range(4))
def test_reference_known_answers(i, kernel, api, load_model, load_seqs, golden):
    m = _model(api, load_model(f"{i}_test_chmm"), kernel)
    v = m.viterbi(load_seqs(f"{i}_test_seq")[0])
    assert same_bits(v, golden[f"tiny{i}_plain"])
    for a, b in zip(v, golden[f"tiny{i}_plain"]):
        assert api.almost_equal(a, b)


@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("name", PROFILE_MODELS)
def test_semantic_equality_all_models(name, kernel, api, oracle, load_model, load_seqs, golden):
    """24 shipped models x the 3 sequences of emit_3_3500_20, one batched call."""
    hmm = load_model(name)
    seqs = load_seqs("emit_3_3500_20")
    m = _model(api, hmm, kernel)
    syms, offs = api.pack_sequences(seqs)
    out, best, arg = m.viterbi_batch(syms, offs)
    assert same_bits(out[0], golden[f"{name}_emit3_3500_seq0_plain"])
    for k in range(3):
        exp = oracle.plain(hmm, seqs[k])
        assert same_bits(out[k], exp), (name, k)
        bv, bi = best_state(exp)
        assert best[k] == bv and arg[k] == bi


@pytest.mark.parametrize("kernel", KERNELS)
def test_ragged_covid(kernel, api, oracle, load_model, load_seqs, golden):
    hmm = load_model("2405")
    seqs = load_seqs("covid-19")
    m = _model(api, hmm, kernel)
    syms, offs = api.pack_sequences(seqs)
    out, best, arg = m.viterbi_batch(syms, offs)
    for k in (0, 4, 9):
        assert same_bits(out[k], golden[f"2405_covid_seq{k}_plain"])
    exp = oracle.plain_batch(hmm, syms, offs, threads=4)
    assert same_bits(out, exp)
    # single-sequence entry point == run_Viterbi(hmm, seq)
    assert same_bits(m.viterbi(seqs[6]), exp[6])


@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("n,seed", [(1, 0), (2, 1), (33, 2), (150, 7), (700, 3), (2405, 2405)])
def test_random_models(n, seed, kernel, api, oracle):
    """chmm_gen-style models: uniform out-degree, +inf weights, two start states, unreachable
    states; ragged batch with length-1 sequences."""
    from spec_viterbi_b200 import generators
    hmm = generators.random_hmm(states_num=n, trans_per_state=min(3, n), num_non_zero_start_probs=min(2, n), seed=seed)
    rng = np.random.default_rng(seed)
    lens = [1, 2, 3, 17, 64, 200, 5, 1, 129] + [int(x) for x in rng.integers(1, 300, 40)]
    seqs = [rng.integers(0, 20, L).astype(np.uint64) for L in lens]
    m = _model(api, hmm, kernel)
    syms, offs = api.pack_sequences(seqs)
    out, best, arg = m.viterbi_batch(syms, offs)
    exp = oracle.plain_batch(hmm, syms, offs, threads=4)
    assert same_bits(out, exp)
    for k in range(len(seqs)):
        bv, bi = best_state(exp[k])
        assert (best[k] == bv or (np.isposinf(bv) and np.isposinf(best[k]))) and arg[k] == bi


@pytest.mark.parametrize("n,seed", [(150, 7), (2405, 2405)])
def test_two_halves_lockstep_variant(n, seed, api, oracle, monkeypatch):
    """Opt-in variant of K2 (sv_plain_ls2.cu): two independent 256-thread halves per CTA, 10 rows per
    thread; same bits as the oracle, including the refill path (more sequences than slots)."""
    from spec_viterbi_b200 import generators, _lib
    monkeypatch.setenv("SV_LS_HALVES", "2")
    hmm = generators.random_hmm(states_num=n, trans_per_state=3, num_non_zero_start_probs=2, seed=seed)
    rng = np.random.default_rng(seed)
    lens = [1, 2, 3, 17, 64, 200, 5, 1, 129] + [int(x) for x in rng.integers(1, 120, 1400)]
    seqs = [rng.integers(0, 20, L).astype(np.uint64) for L in lens]
    m = _model(api, hmm, "ls")
    info = m.info()
    assert info.ls_threads == 256 and info.ls_smem_bytes == 164352  # 2 halves x 2 buffers x 2568 x 16 B
    syms, offs = api.pack_sequences(seqs)
    out, best, arg = m.viterbi_batch(syms, offs)
    exp = oracle.plain_batch(hmm, syms, offs, threads=8)
    assert same_bits(out, exp)
    for k in range(0, len(seqs), 37):
        bv, bi = best_state(exp[k])
        assert (best[k] == bv or (np.isposinf(bv) and np.isposinf(best[k]))) and arg[k] == bi


def test_high_in_degree_and_duplicates(api, oracle):
    """Dense destination rows, duplicate COO entries (FIRST wins) and duplicate start entries."""
    rng = np.random.default_rng(5)
    n = 300
    src = np.concatenate([np.arange(n), np.arange(n), rng.integers(0, n, 900), [3, 3]]).astype(np.uint64)
    dst = np.concatenate([np.zeros(n), np.full(n, 7), rng.integers(0, n, 900), [9, 9]]).astype(np.uint64)
    w = rng.uniform(0.01, 9, src.size).astype(np.float32)
    hmm = api.HMM(n, 5, src.size, src, dst, w, rng.uniform(0.1, 6, (5, n)).astype(np.float32), 3,
                  np.array([4, 4, 10], np.uint64), np.array([0.5, 0.1, 2.0], np.float32))
    seqs = [rng.integers(0, 5, L).astype(np.uint64) for L in (50, 1, 9, 300)]
    syms, offs = api.pack_sequences(seqs)
    exp = oracle.plain_batch(hmm, syms, offs, threads=2)
    for kernel in KERNELS:
        m = _model(api, hmm, kernel)
        out, _, _ = m.viterbi_batch(syms, offs)
        assert same_bits(out, exp), kernel


def test_many_sequences_queue(api, oracle):
    """More sequences than resident CTAs: the persistent work queue hands out the rest."""
    from spec_viterbi_b200 import generators
    hmm = generators.random_hmm(states_num=500, seed=11)
    syms, offs = generators.random_sequences(1500, 40, seed=4, dtype=np.uint64)
    exp = oracle.plain_batch(hmm, syms, offs, threads=8)
    for kernel in KERNELS:
        m = _model(api, hmm, kernel)
        out, best, arg = m.viterbi_batch(syms, offs)
        assert same_bits(out, exp), kernel


def test_adaptor_mirrors_reference_interface(api, oracle, load_model, load_seqs):
    impl = api.B200_impl()
    hmm = load_model("100")
    seq = load_seqs("emit_3_3500_20")[2]
    assert same_bits(impl.run_Viterbi(hmm, seq), oracle.plain(hmm, seq))
    hmm2 = load_model("200")  # a different HMM on the same object re-uploads
    assert same_bits(impl.run_Viterbi(hmm2, seq), oracle.plain(hmm2, seq))


def test_argument_errors(api, load_model):
    from spec_viterbi_b200._lib import SvError
    m = api.Model(load_model("0_test_chmm"))
    with pytest.raises(SvError):
        m.viterbi(np.zeros(0, np.uint64))          # empty sequence
    with pytest.raises(SvError):
        m.viterbi(np.array([0, 99], np.uint64))    # symbol >= emit_num
    # packed u8 input is validated on the device (no host scan): bad symbol in the word body and in the tail
    for pos in (5, 40):
        syms = np.zeros(41, np.uint8)
        syms[pos] = 200
        with pytest.raises(SvError):
            m.viterbi_batch(syms, np.array([0, 20, 41], np.uint64))
    out, _, _ = m.viterbi_batch(np.zeros(41, np.uint8), np.array([0, 20, 41], np.uint64))  # still usable
    assert out.shape[0] == 2

"""GPU parity, plain path (K1/K2 kernels through the C ABI) -- bit-exact against the oracle and
the committed reference outputs.  Reference: GraphBLAS_impl.cpp:4-93; harness protocol of
tests/test_semantic_equality.cpp:19-98 and tests/test_helper.h:44-57."""
import numpy as np
import pytest

from conftest import PROFILE_MODELS, best_state, same_bits

pytestmark = pytest.mark.gpu

KERNELS = ["generic", "es", "ls", "chain", "auto"]


def _model(api, hmm, kernel, must_fit=False):
    """Model with the kernel forced (K0 plans are lazy: set_kernel builds the plan and fails with
    SV_ERR_UNSUPPORTED when the model does not fit that kernel -> skip, unless the case says it must)."""
    from spec_viterbi_b200 import _lib
    m = api.Model(hmm)
    want = {"generic": _lib.KERNEL_GENERIC, "es": _lib.KERNEL_ES, "ls": _lib.KERNEL_LS, "chain": _lib.KERNEL_CHAIN,
            "auto": _lib.KERNEL_AUTO}[kernel]
    try:
        m.set_kernel(want)
    except _lib.SvError as e:
        if must_fit or e.code != 5:
            raise
        pytest.skip(f"model does not fit the {kernel} kernel")
    return m


def _ran(m, kernel):
    """The kernel that was asked for is the one the last launch used."""
    from spec_viterbi_b200 import _lib
    want = {"generic": _lib.KERNEL_GENERIC, "es": _lib.KERNEL_ES, "ls": _lib.KERNEL_LS, "chain": _lib.KERNEL_CHAIN}
    return kernel == "auto" or m.info().kernel == want[kernel]


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
    m = _model(api, hmm, kernel, must_fit=True)  # every shipped profile HMM fits all four kernels
    syms, offs = api.pack_sequences(seqs)
    out, best, arg = m.viterbi_batch(syms, offs)
    assert _ran(m, kernel)
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


@pytest.mark.parametrize("nseq", [300, 590, 1400])
def test_lockstep_sequences_per_cta(nseq, api, oracle):
    """K2 runs 4 sequences per CTA when 8 per CTA would leave SMs idle (batch <= 4 x SMs: the strong-scaling
    shards of config 5) and 8 otherwise; same bits either way, including the refill path."""
    from spec_viterbi_b200 import generators
    hmm = generators.random_hmm(states_num=700, trans_per_state=3, num_non_zero_start_probs=2, seed=3)
    rng = np.random.default_rng(nseq)
    lens = [1, 2, 3, 17, 64, 200, 5, 1, 129] + [int(x) for x in rng.integers(1, 120, nseq - 9)]
    seqs = [rng.integers(0, 20, L).astype(np.uint64) for L in lens]
    m = _model(api, hmm, "ls", must_fit=True)
    syms, offs = api.pack_sequences(seqs)
    out, best, arg = m.viterbi_batch(syms, offs)
    info = m.info()
    assert _ran(m, "ls") and info.ls_seqs_per_cta == (4 if nseq <= 4 * 148 else 8)
    exp = oracle.plain_batch(hmm, syms, offs, threads=8)
    assert same_bits(out, exp)
    for k in range(0, len(seqs), 37):
        bv, bi = best_state(exp[k])
        assert (best[k] == bv or (np.isposinf(bv) and np.isposinf(best[k]))) and arg[k] == bi


@pytest.mark.parametrize("name,nseq", [("100", 5), ("100", 700), ("100", 6000), ("300", 150), ("300", 300), ("1200", 40),
                                       ("2405", 2), ("2405", 149), ("2405", 600), ("2405", 1300)])
def test_chain_kernel_batches(name, nseq, api, oracle, load_model):
    """Chain kernel on profile HMMs at batch sizes that select 1, 2, 4 and 8 sequences per CTA and all three CTA
    sizes, ragged lengths with length-1 and length-2 sequences, more sequences than resident slots (refill)."""
    hmm = load_model(name)
    rng = np.random.default_rng(nseq)
    lens = ([1, 2, 3, 17, 64, 200, 5, 1, 129] + [int(x) for x in rng.integers(1, 90, max(0, nseq - 9))])[:nseq]
    seqs = [rng.integers(0, 20, L).astype(np.uint64) for L in lens]
    m = _model(api, hmm, "chain", must_fit=True)
    syms, offs = api.pack_sequences(seqs)
    out, best, arg = m.viterbi_batch(syms, offs)
    assert _ran(m, "chain") and m.info().chain_seqs_per_cta in (1, 2, 4, 8)
    exp = oracle.plain_batch(hmm, syms, offs, threads=8)
    assert same_bits(out, exp)
    for k in range(0, len(seqs), 7):
        bv, bi = best_state(exp[k])
        assert best[k] == bv and arg[k] == bi
    # vectors only / best only
    out2, _, _ = m.viterbi_batch(syms, offs, want_best=False)
    _, best2, arg2 = m.viterbi_batch(syms, offs, want_vectors=False)
    assert same_bits(out2, exp) and np.array_equal(best2, best) and np.array_equal(arg2, arg)


@pytest.mark.parametrize("kernel", ["ls", "chain"])
@pytest.mark.parametrize("ragged", [False, True])
def test_tail_balance(kernel, ragged, api, oracle, load_model):
    """The last, partial wave of a batch of 8-sequence CTAs is handed out 4 per CTA: as a second launch of the
    4-sequence kernel when all sequences have one length, inside the kernel (second float4 group idle and skipped) when
    they are ragged.  Same bits; every sequence processed exactly once."""
    from spec_viterbi_b200 import generators, _lib
    sms = 148
    hmm = load_model("2405") if kernel == "chain" else generators.random_hmm(states_num=500, seed=21)  # one CTA per SM
    n_seq = 8 * sms + 4 * sms - 37           # one full wave + a tail of at most 4 per CTA
    rng = np.random.default_rng(5)
    lens = rng.integers(20, 60, n_seq) if ragged else np.full(n_seq, 41)
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    syms = rng.integers(0, 20, int(offs[-1])).astype(np.uint8)
    m = _model(api, hmm, kernel, must_fit=True)
    out, best, arg = m.viterbi_batch(syms, offs)
    assert _ran(m, kernel) and m.last_launches() == (1 if ragged else 2)
    exp = oracle.plain_batch(hmm, syms.astype(np.uint64), offs, threads=16)
    assert same_bits(out, exp)
    assert np.array_equal(best, exp.min(1)) and np.array_equal(arg, exp.argmin(1))


def test_chain_kernel_variants(api, oracle):
    """Chain structure with shuffled state numbering, a broken chain, +inf band weights, non-uniform hub weights,
    a second start state in the middle of the chain; and a model with no reachable state (argmin = -1)."""
    rng = np.random.default_rng(3)
    L = 1500
    n = L + 2
    perm = rng.permutation(n)
    N, C = 0, L + 1
    src, dst, w = [N, C], [N, C], [0.3, 0.2]
    for k in range(1, L + 1):
        src += [N, k, k]
        dst += [k, N, C]
        w += [float(rng.uniform(1, 9)), 1.0, 1.25]
        if k < L and k % 211 != 0:
            src += [k]
            dst += [k + 1]
            w += [float(rng.uniform(0.1, 3)) if k % 97 else np.inf]
    src = perm[np.array(src)].astype(np.uint64)
    dst = perm[np.array(dst)].astype(np.uint64)
    em = rng.uniform(0.5, 7, (6, n)).astype(np.float32)
    hmm = api.HMM(n, 6, src.size, src, dst, np.array(w, np.float32), em, 2,
                  np.array([perm[N], perm[4]], np.uint64), np.array([0.0, 2.0], np.float32))
    seqs = [rng.integers(0, 6, T).astype(np.uint64) for T in (1, 2, 3, 50, 400, 7, 1, 33, 120, 5, 64, 2)]
    syms, offs = api.pack_sequences(seqs)
    m = _model(api, hmm, "chain", must_fit=True)
    out, best, arg = m.viterbi_batch(syms, offs)
    exp = oracle.plain_batch(hmm, syms, offs, threads=4)
    assert _ran(m, "chain") and same_bits(out, exp)
    for k in range(len(seqs)):
        bv, bi = best_state(exp[k])
        assert best[k] == bv and arg[k] == bi
    dead = api.HMM(n, 6, src.size, src, dst, np.array(w, np.float32), em, 0, np.zeros(0, np.uint64), np.zeros(0, np.float32))
    md = _model(api, dead, "chain", must_fit=True)
    out, best, arg = md.viterbi_batch(syms, offs)
    assert np.all(np.isposinf(out)) and np.all(np.isposinf(best)) and np.all(arg == -1)


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


def test_multi_device_entry_point(api, oracle, load_model):
    """sv_multi_*: the C-level multi-GPU batch call (replicas, contiguous symbol-balanced ranges, one host thread
    per device).  On a one-GPU box two replicas on device 0 exercise the same code; all devices when there are more."""
    from spec_viterbi_b200 import generators, _lib
    ndev = _lib.lib().sv_device_count()
    devices = list(range(ndev)) if ndev > 1 else [0, 0]
    rng = np.random.default_rng(8)
    for hmm in (load_model("300"), generators.random_hmm(states_num=400, seed=6)):
        lens = [int(x) for x in rng.integers(1, 200, 333)]
        seqs = [rng.integers(0, 20, L).astype(np.uint64) for L in lens]
        syms, offs = api.pack_sequences(seqs)
        mm = api.MultiModel(hmm, devices)
        assert mm.ndev == len(devices)
        exp = oracle.plain_batch(hmm, syms, offs, threads=8)
        for sy in (syms, syms.astype(np.uint8)):
            out, best, arg = mm.viterbi_batch(sy, offs)
            assert same_bits(out, exp)
            for k in range(0, len(seqs), 11):
                bv, bi = best_state(exp[k])
                assert (best[k] == bv or (np.isposinf(bv) and np.isposinf(best[k]))) and arg[k] == bi
        # fewer sequences than devices, and a symbol out of range on the LAST device's share
        out, _, _ = mm.viterbi_batch(syms[: int(offs[1])], offs[:2])
        assert same_bits(out[0], exp[0])
        bad = syms.copy()
        bad[-1] = 77
        with pytest.raises(_lib.SvError):
            mm.viterbi_batch(bad, offs)
        mm.close()


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


@pytest.mark.parametrize("kernel", ["generic", "es", "ls", "chain"])
def test_results_written_straight_to_pinned_host_memory(kernel, api, oracle, load_model):
    """A page-locked result array is device-accessible: sv_viterbi_batch* hands it to the step kernel, which stores every
    final vector when its sequence finishes (no staging buffer, no device-to-host copy afterwards).  Same bits as the
    staged path (pageable array) and as the oracle; one full wave + a tail (two launches for equal lengths).  (The chain
    kernel keeps the staging buffer -- its stores are strided -- and must give the same result through it.)"""
    import torch
    from spec_viterbi_b200 import generators
    hmm = load_model("300") if kernel == "chain" else generators.random_hmm(states_num=300, seed=31)
    rng = np.random.default_rng(9)
    for lens in (rng.integers(1, 70, 333), np.full(8 * 148 + 300, 23)):
        offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
        syms = rng.integers(0, 20, int(offs[-1])).astype(np.uint8)
        m = _model(api, hmm, kernel, must_fit=True)
        nseq, n = len(lens), hmm.states_num
        pinned = torch.full((nseq, n), float("nan"), dtype=torch.float32).pin_memory()
        out_p, best_p, arg_p = m.viterbi_batch(syms, offs, out=pinned.numpy())
        assert _ran(m, kernel)
        out_s, best_s, arg_s = m.viterbi_batch(syms, offs)           # pageable: staged through the device buffer
        exp = oracle.plain_batch(hmm, syms.astype(np.uint64), offs, threads=8)
        assert same_bits(out_p, exp) and same_bits(out_s, exp)
        assert np.array_equal(best_p, best_s) and np.array_equal(arg_p, arg_s)
        assert out_p.ctypes.data == pinned.numpy().ctypes.data      # the caller's array, filled in place

"""GPU parity, specialised path: handler precompute (K3 dense (min,+) GEMM, K3' sparse-LHS) and
the level-L run (K4), bit-exact against the same-level oracle.  Reference:
GraphBLAS_spec_impl.cpp:15-36,50-97,103-182; tests/test_helper.h:59-73 (levels 1..3)."""
import numpy as np
import pytest

from conftest import same_bits

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("i", range(4))
@pytest.mark.parametrize("level", [1, 2, 3])
@pytest.mark.parametrize("mode", ["auto", "table"])
def test_reference_known_answers(i, level, mode, api, load_model, load_seqs, golden):
    from spec_viterbi_b200 import _lib
    impl = api.B200_spec_impl(level, run_mode=_lib.SPEC_RUN_AUTO if mode == "auto" else _lib.SPEC_RUN_TABLE)
    impl.spec_with(load_model(f"{i}_test_chmm"))
    v = impl.run_Viterbi_spec(load_seqs(f"{i}_test_seq")[0])
    assert impl.get_level() == level
    assert same_bits(v, golden[f"tiny{i}_spec{level}"])


@pytest.mark.parametrize("method", ["sparse", "dense", "pair", "auto"])
@pytest.mark.parametrize("name", ["100", "200"])
def test_level2_reference_fixture(name, method, api, load_model, load_seqs, golden):
    from spec_viterbi_b200 import _lib
    impl = api.B200_spec_impl(2, method={"dense": _lib.SPEC_DENSE_GEMM, "sparse": _lib.SPEC_SPARSE_LHS,
                                         "pair": _lib.SPEC_PAIR_ROWS, "auto": _lib.SPEC_AUTO}[method],
                              run_mode=_lib.SPEC_RUN_TABLE)  # the precomputed handlers are what is under test
    impl.spec_with(load_model(name))
    out = impl.run_Viterbi_spec_batch(load_seqs("emit_3_3500_20"))
    for k in range(3):
        assert same_bits(out[k], golden[f"{name}_emit3_3500_seq{k}_spec2"])


@pytest.mark.parametrize("name", ["100", "300"])
def test_handlers_bit_exact(name, api, oracle, load_model):
    """Every compared pair handler equals the oracle's M_b (x) M_a, for all three precompute kernels."""
    from spec_viterbi_b200 import _lib
    hmm = load_model(name)
    m = api.Model(hmm)
    osp = oracle.spec(hmm, 2)
    sparse = api.Spec(m, 2, _lib.SPEC_SPARSE_LHS)
    dense = api.Spec(m, 2, _lib.SPEC_DENSE_GEMM)
    pair = api.Spec(m, 2, _lib.SPEC_PAIR_ROWS)
    for t in (0, 1, 19, 20, 57, 211, 399):
        exp = osp.handler(t)
        assert same_bits(sparse.handler(t), exp), t
        assert same_bits(dense.handler(t), exp), t
        assert same_bits(pair.handler(t), exp), t


def test_config3_reference_fixture(api, load_model, load_seqs, golden):
    """BASELINE.json configs[2] for real: 2405.chmm x emit_50_3500_20 at level 2 (400 handlers, 9.27 GB),
    through the precomputed table + K4 AND through the implicit run (K4i), against vectors produced by the
    reference's own GraphBLAS_spec_impl (GraphBLAS_spec_impl.cpp:50-97; tools/make_fixtures.py --config3)."""
    from spec_viterbi_b200 import _lib
    hmm = load_model("2405")
    seqs = load_seqs("emit_50_3500_20")
    syms, offs = api.pack_sequences([seqs[0], seqs[49]])
    m = api.Model(hmm)
    sp = api.Spec(m, 2)
    for mode in (_lib.SPEC_RUN_TABLE, _lib.SPEC_RUN_IMPLICIT):
        sp.set_run_mode(mode)
        out, best, arg = sp.run_batch(syms, offs)
        for j, k in enumerate((0, 49)):
            exp = golden[f"2405_emit50_3500_seq{k}_spec2"]
            assert same_bits(out[j], exp), (mode, k)
            assert best[j] == exp.min() and arg[j] == exp.argmin()
    sp.close()
    m.close()


def test_config3_pair_handlers_vs_oracle(api, oracle, load_model):
    """K3'' (the default level-2 precompute) on 2405.chmm against the ORACLE's M_b (x) M_a, element-wise."""
    from spec_viterbi_b200 import _lib
    hmm = load_model("2405")
    m = api.Model(hmm)
    osp = oracle.spec(hmm, 2)
    for t in (0, 211, 399):
        sp = api.Spec(m, 2, _lib.SPEC_PAIR_ROWS, t, t + 1)
        assert same_bits(sp.handler(t), osp.handler(t)), t
        sp.close()
    m.close()


def test_pair_rows_equals_sparse_lhs_on_large_models(api, load_model):
    """K3'' against K3' (itself pinned to the oracle above) on the largest profile HMM (dense rows,
    one dense column) and on a random sparse model with colliding (k, value) pairs."""
    from spec_viterbi_b200 import generators, _lib
    for hmm in (load_model("2405"), generators.random_hmm(states_num=700, seed=5),
                generators.random_hmm(states_num=300, trans_per_state=9, seed=6),     # > 32 pairs per row: barrier path
                generators.random_hmm(states_num=60, trans_per_state=40, seed=7)):   # every row heavy: CSC walk
        m = api.Model(hmm)
        E = hmm.emit_num
        for t0 in (0, 7 * E + 3, E * E - 2):
            a = api.Spec(m, 2, _lib.SPEC_SPARSE_LHS, t0, t0 + 2)
            b = api.Spec(m, 2, _lib.SPEC_PAIR_ROWS, t0, t0 + 2)
            for t in (t0, t0 + 1):
                assert same_bits(a.handler(t), b.handler(t)), t
            a.close()
            b.close()


def test_random_model_levels(api, oracle):
    from spec_viterbi_b200 import generators, _lib
    hmm = generators.random_hmm(states_num=130, seed=9)
    rng = np.random.default_rng(1)
    seqs = [rng.integers(0, 20, L).astype(np.uint64) for L in (1, 2, 3, 4, 5, 6, 7, 50, 301)]
    for level, method in ((1, _lib.SPEC_AUTO), (2, _lib.SPEC_SPARSE_LHS), (2, _lib.SPEC_DENSE_GEMM),
                          (2, _lib.SPEC_PAIR_ROWS)):
        impl = api.B200_spec_impl(level, method=method, run_mode=_lib.SPEC_RUN_TABLE)
        impl.spec_with(hmm)
        out = impl.run_Viterbi_spec_batch(seqs)
        osp = oracle.spec(hmm, level)
        for k, s in enumerate(seqs):
            assert same_bits(out[k], osp.run(s)), (level, k)


def test_level3_small_alphabet(api, oracle):
    rng = np.random.default_rng(3)
    n, e = 40, 4
    src = rng.integers(0, n, 160).astype(np.uint64)
    dst = rng.integers(0, n, 160).astype(np.uint64)
    hmm = api.HMM(n, e, 160, src, dst, rng.uniform(0.1, 5, 160).astype(np.float32),
                  rng.uniform(0.1, 5, (e, n)).astype(np.float32), 2, np.array([0, 5], np.uint64),
                  np.array([1.0, 2.0], np.float32))
    impl = api.B200_spec_impl(3)
    impl.spec_with(hmm)
    osp = oracle.spec(hmm, 3)
    for L in (1, 2, 3, 4, 5, 6, 7, 8, 100):
        s = rng.integers(0, e, L).astype(np.uint64)
        assert same_bits(impl.run_Viterbi_spec(s), osp.run(s)), L


def test_level3_on_demand(api, oracle, load_model):
    """level 3 WITHOUT the level-3 table.  SV_SPEC_ONDEMAND keeps only the level-2 table and builds,
    per chunk, the level-3 handler each sequence observes (H(k1,k2,k3) = M_k3 (x) H(k1,k2),
    GraphBLAS_spec_impl.cpp:15-36) before applying it: the bits of the full-table run and of the oracle."""
    from spec_viterbi_b200 import generators, _lib
    rng = np.random.default_rng(13)
    # (a) against the full level-3 table and the oracle on a model whose 8000 handlers fit easily
    hmm = generators.random_hmm(states_num=90, seed=4)
    seqs = [rng.integers(0, 20, L).astype(np.uint64) for L in (1, 2, 3, 4, 5, 6, 7, 8, 40, 121)]
    syms, offs = api.pack_sequences(seqs)
    m = api.Model(hmm)
    full = api.Spec(m, 3, _lib.SPEC_SPARSE_LHS)
    exp, ebest, earg = full.run_batch(syms, offs)
    full.close()
    od = api.Spec(m, 3, _lib.SPEC_ONDEMAND)
    got, best, arg = od.run_batch(syms, offs)
    assert same_bits(got, exp) and np.array_equal(best, ebest) and np.array_equal(arg, earg)
    osp = oracle.spec(hmm, 3)
    for k in (0, 3, 6, 8, 9):
        assert same_bits(got[k], osp.run(seqs[k])), k
    with pytest.raises(_lib.SvError):   # there is no level-3 table to read a handler from
        od.handler(5)
    od.close()
    # (b) the case the mode exists for: 2405.chmm at level 3 (full table: 185 GB) on one GPU, against the oracle
    hmm = load_model("2405")
    seqs = [rng.integers(0, 20, L).astype(np.uint64) for L in (20, 13, 1, 31)]
    syms, offs = api.pack_sequences(seqs)
    m = api.Model(hmm)
    od = api.Spec(m, 3)                 # AUTO: the full table cannot fit -> on demand
    got, best, arg = od.run_batch(syms, offs)
    osp = oracle.spec(hmm, 3)
    for k in range(len(seqs)):
        exp = osp.run(seqs[k])
        assert same_bits(got[k], exp), k
        assert best[k] == exp.min() and arg[k] == exp.argmin()
    # the reference interface takes it too
    impl = api.B200_spec_impl(3)
    impl.spec_with(hmm)
    assert same_bits(impl.run_Viterbi_spec(seqs[1]), osp.run(seqs[1]))


def test_spec_vs_plain_cross_path(api, load_model, load_seqs):
    """Across paths (spec level 2 vs plain) fp32 re-association differs in every element.  north_star asks for <= 1e-5
    relative on final scores; SURVEY A.2 measured 9.6e-6 on this input with the reference itself and up to 1.6e-5 on
    others, so the bound asserted here is 2e-5 (documented in DESIGN.md section 2) plus the reference's own criterion
    (abs <= 1.0 or both +inf, HMM.h:43-49) and equal argmin.  Positions that are +inf on both sides are masked."""
    hmm = load_model("300")
    seqs = load_seqs("emit_3_3500_20")
    plain = api.B200_impl().run_Viterbi_batch(hmm, seqs)
    impl = api.B200_spec_impl(2)
    impl.spec_with(hmm)
    spec = impl.run_Viterbi_spec_batch(seqs)
    both_inf = np.isposinf(plain) & np.isposinf(spec)
    assert np.array_equal(np.isposinf(plain), np.isposinf(spec))
    fin = ~both_inf
    assert np.all(np.abs(plain[fin] - spec[fin]) <= 1.0)
    assert np.max(np.abs(plain[fin] - spec[fin]) / np.abs(plain[fin])) < 2e-5
    assert np.array_equal(plain.argmin(1), spec.argmin(1))


def test_handler_table_save_load(api, oracle, tmp_path):
    """a saved level-2 table attaches to the same model without recomputation and
    gives the same bits; it is refused for a different model; tuple shards round-trip too."""
    from spec_viterbi_b200 import generators, _lib
    hmm = generators.random_hmm(states_num=150, trans_per_state=3, num_non_zero_start_probs=2, seed=9)
    seqs = [np.random.default_rng(k).integers(0, 20, L).astype(np.uint64) for k, L in enumerate((1, 2, 7, 64, 201))]
    syms, offs = api.pack_sequences(seqs)
    m = api.Model(hmm)
    built = api.Spec(m, 2)
    built.set_run_mode(_lib.SPEC_RUN_TABLE)  # the table is what is under test
    exp, _, _ = built.run_batch(syms, offs)
    osp = oracle.spec(hmm, 2)
    assert same_bits(exp[4], osp.run(seqs[4]))
    path = tmp_path / "level2.svht"
    built.save(path)
    assert path.stat().st_size == 64 + 400 * 150 * 150 * 4
    loaded = api.Spec.load(m, path)
    assert loaded.level == 2 and loaded.table()[1:] == (0, 400)
    assert same_bits(loaded.handler(123), built.handler(123))
    loaded.set_run_mode(_lib.SPEC_RUN_TABLE)
    got, _, _ = loaded.run_batch(syms, offs)
    assert same_bits(got, exp)
    # a shard of the tuple range
    shard = api.Spec(m, 2, tuple_begin=100, tuple_end=160)
    shard.save(tmp_path / "shard.svht")
    back = api.Spec.load(m, tmp_path / "shard.svht")
    assert back.table()[1:] == (100, 160) and same_bits(back.handler(159), built.handler(159))
    # another model (one weight differs) must not accept the table
    other = generators.random_hmm(states_num=150, trans_per_state=3, num_non_zero_start_probs=2, seed=10)
    with pytest.raises(_lib.SvError):
        api.Spec.load(api.Model(other), path)
    # not a table file / level 1 has nothing to save
    (tmp_path / "junk").write_bytes(b"x" * 100)
    with pytest.raises(_lib.SvError):
        api.Spec.load(m, tmp_path / "junk")
    with pytest.raises(_lib.SvError):
        api.Spec(m, 1).save(tmp_path / "l1")


IMPLICIT_MODES = ["step", "auto"]


def _implicit(mode):
    from spec_viterbi_b200 import _lib
    return {"step": _lib.SPEC_RUN_IMPLICIT, "auto": _lib.SPEC_RUN_AUTO}[mode]


@pytest.mark.parametrize("mode", IMPLICIT_MODES)
def test_implicit_run_reference_fixture(mode, api, load_model, load_seqs, golden):
    """K4i (level-2 chunks recomputed from the sparse factors, no table) gives the reference's bits:
    profile HMMs (two dense rows, one dense column) against the fixtures generated from
    GraphBLAS_spec_impl (level 2), with an EMPTY handler table."""
    from spec_viterbi_b200 import _lib
    for name in ("100", "200"):
        hmm = load_model(name)
        m = api.Model(hmm)
        sp = api.Spec(m, 2, tuple_begin=400, tuple_end=400)  # nothing precomputed
        sp.set_run_mode(_implicit(mode))
        syms, offs = api.pack_sequences(load_seqs("emit_3_3500_20"))
        out, best, arg = sp.run_batch(syms, offs)
        for k in range(3):
            assert same_bits(out[k], golden[f"{name}_emit3_3500_seq{k}_spec2"])
            assert best[k] == out[k].min() and arg[k] == out[k].argmin()
        with pytest.raises(_lib.SvError):  # the table path refuses an incomplete table
            sp.set_run_mode(_lib.SPEC_RUN_TABLE)
            sp.run_batch(syms, offs)


@pytest.mark.parametrize("mode", IMPLICIT_MODES)
def test_implicit_run_equals_table_run(mode, api, oracle, load_model, load_seqs):
    """K4i against table + K4 (pinned to the oracle above) and the oracle itself: ragged lengths
    (1..7 cover init-only, tail-only and odd/even chunking), random sparse models with +inf weights,
    a model with 3 dense rows, the largest profile HMM on the ragged covid set."""
    from spec_viterbi_b200 import generators, _lib
    rng = np.random.default_rng(11)
    dense3 = generators.random_hmm(states_num=90, trans_per_state=3, seed=12)
    extra_dst = np.repeat(np.array([4, 17, 60], np.uint64), 90)       # three states reachable from everywhere
    extra_src = np.tile(np.arange(90, dtype=np.uint64), 3)
    keep = ~np.isin(dense3.trans_rows * 1000 + dense3.trans_cols, extra_src * 1000 + extra_dst)
    cases = [(generators.random_hmm(states_num=130, seed=9), True),
             (generators.random_hmm(states_num=700, trans_per_state=5, seed=5), True),
             (api.HMM(90, 20, int(keep.sum()) + 270,
                      np.concatenate([dense3.trans_rows[keep], extra_src]),
                      np.concatenate([dense3.trans_cols[keep], extra_dst]),
                      np.concatenate([dense3.trans_probs[keep], rng.uniform(0.1, 6, 270).astype(np.float32)]),
                      dense3.emissions, dense3.non_zero_start_probs, dense3.start_probabilities_cols,
                      dense3.start_probabilities), True)]
    for hmm, use_oracle in cases:
        m = api.Model(hmm)
        seqs = [rng.integers(0, 20, L).astype(np.uint64) for L in (1, 2, 3, 4, 5, 6, 7, 50, 301, 1000)]
        syms, offs = api.pack_sequences(seqs)
        table = api.Spec(m, 2)
        table.set_run_mode(_lib.SPEC_RUN_TABLE)
        exp, ebest, earg = table.run_batch(syms, offs)
        table.set_run_mode(_implicit(mode))
        got, best, arg = table.run_batch(syms, offs)
        assert same_bits(got, exp)
        assert np.array_equal(best, ebest) and np.array_equal(arg, earg)
        if use_oracle:
            osp = oracle.spec(hmm, 2)
            for k in (0, 1, 2, 6, 8):
                assert same_bits(got[k], osp.run(seqs[k])), k
    hmm = load_model("2405")
    m = api.Model(hmm)
    syms, offs = api.pack_sequences(load_seqs("covid-19"))
    table = api.Spec(m, 2)
    table.set_run_mode(_lib.SPEC_RUN_TABLE)
    exp, _, _ = table.run_batch(syms, offs)
    table.set_run_mode(_implicit(mode))
    got, _, _ = table.run_batch(syms, offs)
    assert same_bits(got, exp)


def test_tuple_shards_gathered_into_one_table(api, oracle):
    """The multi-GPU precompute flow on one device (two 'ranks'): each builds its tuple shard into its slice
    of ONE caller-owned full-size table, then sv_spec_set_range(full base, 0, 0, ntuples) publishes the whole
    table (what follows the all-gather).  The run must give the bits of a locally built table; ranges that
    start below the table's first tuple are rejected."""
    import torch
    from spec_viterbi_b200 import generators, _lib
    hmm = generators.random_hmm(states_num=120, trans_per_state=3, seed=21)
    n, E = 120, 20
    seqs = [np.random.default_rng(k).integers(0, E, L).astype(np.uint64) for k, L in enumerate((2, 3, 9, 64, 301))]
    syms, offs = api.pack_sequences(seqs)
    m = api.Model(hmm)
    ref = api.Spec(m, 2)
    ref.set_run_mode(_lib.SPEC_RUN_TABLE)
    exp, _, _ = ref.run_batch(syms, offs)
    table = torch.full((E * E, n, n), float("nan"), dtype=torch.float32, device="cuda")
    half = 137  # a ragged cut
    a = api.Spec(m, 2, tuple_begin=0, tuple_end=half, d_table=table[0:half].data_ptr())
    b = api.Spec(m, 2, tuple_begin=half, tuple_end=E * E, d_table=table[half:].data_ptr())
    torch.cuda.synchronize()
    assert not torch.isnan(table).any()
    for sp in (a, b):
        sp.set_run_mode(_lib.SPEC_RUN_TABLE)
        with pytest.raises(_lib.SvError):  # a shard alone cannot run
            sp.run_batch(syms, offs)
    with pytest.raises(_lib.SvError):      # b's table starts at tuple `half`: tuple 0 is not in it
        b.set_range(table[half:].data_ptr(), half, 0, E * E)
    b.set_range(table.data_ptr(), 0, 0, E * E)
    got, _, _ = b.run_batch(syms, offs)
    assert same_bits(got, exp)
    assert same_bits(b.handler(3), ref.handler(3)) and same_bits(b.handler(399), ref.handler(399))
    assert same_bits(got[4], oracle.spec(hmm, 2).run(seqs[4]))


def test_concurrent_streams_on_one_model(api, oracle):
    """Two launches of ONE model overlapping on two streams (each takes its own work-queue counter) and two
    back-to-back runs of one spec on two streams (ordered on the device): all results are the oracle's."""
    import torch
    from spec_viterbi_b200 import generators
    hmm = generators.random_hmm(states_num=300, seed=4)
    m = api.Model(hmm)
    batches = [generators.random_sequences(600, 400, seed=70 + k, dtype=np.uint8) for k in range(2)]
    seqs = [m.upload(sy, of) for sy, of in batches]
    reps = 3
    outs = [[torch.full((600, 300), float("nan"), dtype=torch.float32, device="cuda") for _ in range(2)] for _ in range(reps)]
    streams = [torch.cuda.Stream() for _ in range(2)]
    torch.cuda.synchronize()
    for rep in range(reps):
        for k in range(2):
            m.viterbi_device(seqs[k], outs[rep][k].data_ptr(), 0, 0, streams[k].cuda_stream)
    torch.cuda.synchronize()
    for k in range(2):
        sy, of = batches[k]
        exp = oracle.plain_batch(hmm, sy[: int(of[8])].astype(np.uint64), of[:9], threads=4)
        for rep in range(reps):
            got = outs[rep][k].cpu().numpy()
            assert same_bits(got[:8], exp)
            assert not np.isnan(got).any()  # every sequence was processed in every launch: no row left unwritten
            assert same_bits(got, outs[0][k].cpu().numpy())
    outs = outs[0]
    sp = api.Spec(m, 2, tuple_begin=400, tuple_end=400)
    for k in range(2):
        sp.run_device(seqs[k], outs[k].data_ptr(), 0, 0, streams[k].cuda_stream)
    torch.cuda.synchronize()
    osp = oracle.spec(hmm, 2)
    for k in range(2):
        sy, of = batches[k]
        assert same_bits(outs[k][5].cpu().numpy(), osp.run(sy[int(of[5]):int(of[6])].astype(np.uint64)))


def test_implicit_run_unsupported(api):
    """More dense rows than K4i stages (or level != 2): SV_ERR_UNSUPPORTED, mode unchanged."""
    from spec_viterbi_b200 import generators, _lib
    m = api.Model(generators.random_hmm(states_num=60, trans_per_state=40, seed=7))  # every row heavy
    sp = api.Spec(m, 2)
    with pytest.raises(_lib.SvError):
        sp.set_run_mode(_lib.SPEC_RUN_IMPLICIT)
    s = np.arange(9, dtype=np.uint64) % 20
    assert sp.run(s).shape == (60,)


# ------------------------------------------------------------------ K4c: chain run ----
CHAIN_MODELS = ["100", "200", "300", "400", "700", "1200", "2405"]


@pytest.mark.parametrize("name", CHAIN_MODELS)
def test_chain_run_equals_implicit_and_oracle(name, api, oracle, load_model):
    """K4c (level 2 of a chain + specials model in one persistent launch, O(N) per chunk) against K4i (pinned to the
    reference fixtures and the oracle above) and, on the smaller models, against the oracle itself: ragged lengths
    1..7 (init only, tail only, odd / even chunking), all three CTA sizes (100/200: 128 threads, 300/400: 256,
    700+: 512), 200.chmm and 400.chmm carry a negative weight (shuffle reduction instead of REDUX)."""
    from spec_viterbi_b200 import _lib
    hmm = load_model(name)
    m = api.Model(hmm)
    rng = np.random.default_rng(int(name))
    seqs = [rng.integers(0, 20, L).astype(np.uint64) for L in (1, 2, 3, 4, 5, 6, 7, 50, 301, 1000, 8, 33)]
    syms, offs = api.pack_sequences(seqs)
    sp = api.Spec(m, 2, tuple_begin=400, tuple_end=400)  # nothing precomputed
    sp.set_run_mode(_lib.SPEC_RUN_IMPLICIT)
    exp, ebest, earg = sp.run_batch(syms, offs)
    sp.set_run_mode(_lib.SPEC_RUN_CHAIN)
    got, best, arg = sp.run_batch(syms, offs)
    assert m.last_launches() == 1
    assert same_bits(got, exp)
    assert np.array_equal(best, ebest) and np.array_equal(arg, earg)
    # vectors only / best only; u8 symbols; AUTO picks the chain run
    got2, _, _ = sp.run_batch(syms.astype(np.uint8), offs, want_best=False)
    _, best2, arg2 = sp.run_batch(syms, offs, want_vectors=False)
    assert same_bits(got2, exp) and np.array_equal(best2, ebest) and np.array_equal(arg2, earg)
    sp.set_run_mode(_lib.SPEC_RUN_AUTO)
    got3, _, _ = sp.run_batch(syms, offs)
    assert m.last_launches() == 1 and same_bits(got3, exp)
    if name in ("100", "200", "300"):
        osp = oracle.spec(hmm, 2)
        for k in (0, 1, 2, 3, 6, 8, 10):
            assert same_bits(got[k], osp.run(seqs[k])), k


def test_chain_run_reference_fixtures(api, load_model, load_seqs, golden):
    """K4c against vectors produced by the reference's own GraphBLAS_spec_impl at level 2 (GraphBLAS_spec_impl.cpp:
    50-97): 100.chmm / 200.chmm x emit_3_3500_20 and BASELINE.json configs[2] for real (2405.chmm x sequences 0 and 49
    of emit_50_3500_20)."""
    from spec_viterbi_b200 import _lib
    for name in ("100", "200"):
        m = api.Model(load_model(name))
        sp = api.Spec(m, 2, tuple_begin=400, tuple_end=400)
        sp.set_run_mode(_lib.SPEC_RUN_CHAIN)
        out, best, arg = sp.run_batch(*api.pack_sequences(load_seqs("emit_3_3500_20")))
        for k in range(3):
            exp = golden[f"{name}_emit3_3500_seq{k}_spec2"]
            assert same_bits(out[k], exp)
            assert best[k] == exp.min() and arg[k] == exp.argmin()
    m = api.Model(load_model("2405"))
    sp = api.Spec(m, 2, tuple_begin=400, tuple_end=400)
    sp.set_run_mode(_lib.SPEC_RUN_CHAIN)
    seqs = load_seqs("emit_50_3500_20")
    out, best, arg = sp.run_batch(*api.pack_sequences([seqs[0], seqs[49]]))
    for j, k in enumerate((0, 49)):
        exp = golden[f"2405_emit50_3500_seq{k}_spec2"]
        assert same_bits(out[j], exp), k
        assert best[j] == exp.min() and arg[j] == exp.argmin()


def test_chain_run_queue_and_covid(api, load_model, load_seqs):
    """More sequences than resident CTAs (the persistent queue hands out the rest) and the ragged covid set on the
    largest profile HMM, against K4i."""
    from spec_viterbi_b200 import _lib
    rng = np.random.default_rng(17)
    m = api.Model(load_model("300"))
    seqs = [rng.integers(0, 20, int(L)).astype(np.uint64) for L in rng.integers(1, 120, 1500)]
    syms, offs = api.pack_sequences(seqs)
    sp = api.Spec(m, 2, tuple_begin=400, tuple_end=400)
    sp.set_run_mode(_lib.SPEC_RUN_IMPLICIT)
    exp, ebest, earg = sp.run_batch(syms, offs)
    sp.set_run_mode(_lib.SPEC_RUN_CHAIN)
    got, best, arg = sp.run_batch(syms, offs)
    assert same_bits(got, exp) and np.array_equal(best, ebest) and np.array_equal(arg, earg)
    m = api.Model(load_model("2405"))
    syms, offs = api.pack_sequences(load_seqs("covid-19"))
    sp = api.Spec(m, 2, tuple_begin=400, tuple_end=400)
    sp.set_run_mode(_lib.SPEC_RUN_IMPLICIT)
    exp, ebest, earg = sp.run_batch(syms, offs)
    sp.set_run_mode(_lib.SPEC_RUN_CHAIN)
    got, best, arg = sp.run_batch(syms, offs)
    assert same_bits(got, exp) and np.array_equal(best, ebest) and np.array_equal(arg, earg)


def test_chain_run_variants(api, oracle):
    """Chain structure with shuffled state numbering, broken chains, +inf band weights, non-uniform hub weights, a
    hub -> C edge and a second start state, against the level-2 oracle; a model with no reachable state; and models
    without the structure: CHAIN is refused, AUTO falls back to K4i / the table."""
    from spec_viterbi_b200 import generators, _lib
    rng = np.random.default_rng(3)
    L = 700
    n = L + 2
    perm = rng.permutation(n)
    N, C = 0, L + 1
    src, dst, w = [N, C, N], [N, C, C], [0.3, 0.2, 4.5]
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
    seqs = [rng.integers(0, 6, T).astype(np.uint64) for T in (1, 2, 3, 50, 401, 7, 1, 33, 120, 5, 64, 2)]
    syms, offs = api.pack_sequences(seqs)
    m = api.Model(hmm)
    sp = api.Spec(m, 2, tuple_begin=36, tuple_end=36)
    sp.set_run_mode(_lib.SPEC_RUN_CHAIN)
    out, best, arg = sp.run_batch(syms, offs)
    osp = oracle.spec(hmm, 2)
    for k in range(len(seqs)):
        exp = osp.run(seqs[k])
        assert same_bits(out[k], exp), k
        assert best[k] == exp.min() and arg[k] == exp.argmin()
    dead = api.HMM(n, 6, src.size, src, dst, np.array(w, np.float32), em, 0, np.zeros(0, np.uint64), np.zeros(0, np.float32))
    sd = api.Spec(api.Model(dead), 2, tuple_begin=36, tuple_end=36)
    sd.set_run_mode(_lib.SPEC_RUN_CHAIN)
    out, best, arg = sd.run_batch(syms, offs)
    assert np.all(np.isposinf(out)) and np.all(np.isposinf(best)) and np.all(arg == -1)
    # no chain structure
    rnd = generators.random_hmm(states_num=130, seed=9)
    mr = api.Model(rnd)
    sr = api.Spec(mr, 2)
    with pytest.raises(_lib.SvError):
        sr.set_run_mode(_lib.SPEC_RUN_CHAIN)
    syms, offs = api.pack_sequences([rng.integers(0, 20, T).astype(np.uint64) for T in (1, 2, 9, 40)])
    sr.set_run_mode(_lib.SPEC_RUN_TABLE)
    exp, _, _ = sr.run_batch(syms, offs)
    sr.set_run_mode(_lib.SPEC_RUN_AUTO)
    got, _, _ = sr.run_batch(syms, offs)
    assert same_bits(got, exp) and mr.last_launches() > 1
    # level 3 of a chain model: CHAIN is level 2 only
    s3 = api.Spec(api.Model(hmm), 3, tuple_begin=216, tuple_end=216)
    with pytest.raises(_lib.SvError):
        s3.set_run_mode(_lib.SPEC_RUN_CHAIN)


def test_deferred_level2_table(api, oracle, load_model):
    """spec_with at level 2 of a chain + specials model (AUTO) defers the handler table -- the run (K4c) does not read it:
    no device memory is taken until something needs the handlers (the table run, sv_spec_get_handler, sv_spec_table,
    sv_spec_save), and what is then built is the same table."""
    import torch
    from spec_viterbi_b200 import _lib
    hmm = load_model("1200")
    m = api.Model(hmm)
    n = hmm.states_num
    table_bytes = 400 * n * n * 4
    rng = np.random.default_rng(2)
    syms, offs = api.pack_sequences([rng.integers(0, 20, T).astype(np.uint64) for T in (1, 2, 7, 40, 301)])
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    sp = api.Spec(m, 2)                                   # AUTO: deferred
    got, best, arg = sp.run_batch(syms, offs)             # K4c
    assert m.last_launches() == 1
    assert free0 - torch.cuda.mem_get_info()[0] < table_bytes // 4
    sp.set_run_mode(_lib.SPEC_RUN_TABLE)                  # the table run builds it
    exp, ebest, earg = sp.run_batch(syms, offs)
    assert free0 - torch.cuda.mem_get_info()[0] >= table_bytes
    assert same_bits(got, exp) and np.array_equal(best, ebest) and np.array_equal(arg, earg)
    ptr, tb, te = sp.table()
    assert ptr and (tb, te) == (0, 400)
    sp.close()
    # get_handler on a deferred spec; an explicit method is never deferred
    hmm = load_model("100")
    m = api.Model(hmm)
    lazy = api.Spec(m, 2)
    eager = api.Spec(m, 2, _lib.SPEC_PAIR_ROWS)
    osp = oracle.spec(hmm, 2)
    for t in (0, 57, 399):
        assert same_bits(lazy.handler(t), osp.handler(t)) and same_bits(eager.handler(t), osp.handler(t))


@pytest.mark.parametrize("seed", range(6))
def test_random_chain_models_plain_and_level2(seed, api, oracle):
    """The randomized chain + specials models of tests/conftest.py::random_chain_hmm on the GPU: the chain kernel
    (tabulated inner adds) against the plain oracle and K4c against the level-2 oracle, ragged lengths."""
    from conftest import random_chain_hmm
    from spec_viterbi_b200 import _lib
    hmm, E, rng = random_chain_hmm(api, seed)
    seqs = [rng.integers(0, E, T).astype(np.uint64) for T in (1, 2, 3, 4, 5, 8, 33, 2, 17, 64)]
    syms, offs = api.pack_sequences(seqs)
    m = api.Model(hmm)
    m.set_kernel(_lib.KERNEL_CHAIN)
    out, best, arg = m.viterbi_batch(syms, offs)
    assert m.info().kernel == _lib.KERNEL_CHAIN
    assert same_bits(out, oracle.plain_batch(hmm, syms, offs))
    sp = api.Spec(m, 2)
    sp.set_run_mode(_lib.SPEC_RUN_CHAIN)
    out2, best2, arg2 = sp.run_batch(syms, offs)
    osp = oracle.spec(hmm, 2)
    for k, seq in enumerate(seqs):
        exp = osp.run(seq)
        assert same_bits(out2[k], exp), (seed, k)
        assert best2[k] == exp.min() and arg2[k] == exp.argmin()

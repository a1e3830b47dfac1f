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


# ---------------------------------------------------------------- chain kernel K0 ----
@pytest.mark.parametrize("name", ["100", "300", "1200", "2405"])
def test_chain_plan_profile_models(name, api, oracle, load_model, load_seqs):
    """Every shipped profile HMM has the chain + specials structure; the kernel's algebra (band edge from
    the neighbouring position, hub broadcast, dense rows through the chain minimum) replayed on the host
    gives the oracle's bits."""
    hmm = load_model(name)
    n = hmm.states_num
    for seq in (load_seqs("covid-19")[2], load_seqs("covid-19")[0][:1], load_seqs("emit_3_3500_20")[1][:500]):
        stats, out = api.k0_chain_check(hmm, seq)
        assert stats["valid"] == 1 and stats["specials"] == 2 and stats["chain_states"] == n - 2
        assert {stats["special0"], stats["special1"]} == {0, n - 1} and stats["special0"] == 0  # hub = N state
        assert stats["threads"] * 5 >= n - 2 and stats["nonneg"] == 1
        assert same_bits(out, oracle.plain(hmm, seq))


@pytest.mark.parametrize("name", ["100", "200", "300"])
def test_chain_pair_replay_vs_oracle(name, api, oracle, load_model):
    """K4c's algebra (level-2 chunks through the collapsed two-step paths, sv_k0_chain_pair_check) replayed on the host
    against the level-2 oracle: lengths 1..8 cover init only, tail only and odd / even chunking; 200.chmm carries a
    negative weight."""
    hmm = load_model(name)
    osp = oracle.spec(hmm, 2)
    rng = np.random.default_rng(int(name))
    for T in (1, 2, 3, 4, 5, 6, 7, 8, 61, 200):
        seq = rng.integers(0, 20, T).astype(np.uint64)
        stats, out = api.k0_chain_check(hmm, seq, level=2)
        assert stats["valid"] == 1
        assert same_bits(out, osp.run(seq)), T


def test_chain_pair_replay_config3_fixture(api, load_model, load_seqs, golden):
    """BASELINE.json configs[2] without a device: the K4c algebra replayed on the host reproduces, bit for bit, the
    vectors the reference's own GraphBLAS_spec_impl produced for 2405.chmm x emit_50_3500_20 at level 2
    (tools/make_fixtures.py --config3), and the level-2 fixtures of 100.chmm / 200.chmm."""
    hmm = load_model("2405")
    seqs = load_seqs("emit_50_3500_20")
    for k in (0, 49):
        stats, out = api.k0_chain_check(hmm, seqs[k], level=2)
        assert stats["valid"] == 1 and same_bits(out, golden[f"2405_emit50_3500_seq{k}_spec2"]), k
    for name in ("100", "200"):
        hmm = load_model(name)
        for k, seq in enumerate(load_seqs("emit_3_3500_20")):
            _, out = api.k0_chain_check(hmm, seq, level=2)
            assert same_bits(out, golden[f"{name}_emit3_3500_seq{k}_spec2"]), (name, k)


def test_chain_plan_rejects_other_structures(api):
    from spec_viterbi_b200 import generators
    # random sparse graph: no specials
    st, _ = api.k0_chain_check(generators.random_hmm(states_num=150, seed=7))
    assert st["valid"] == 0
    # profile HMM with one dense-row weight changed: the chain-minimum identity no longer applies
    import copy
    from conftest import GOLDEN
    hmm = api.read_HMM(str(GOLDEN / "models" / "100.chmmb"))
    bad = copy.deepcopy(hmm)
    k = int(np.flatnonzero((bad.trans_cols == 0) & (bad.trans_rows == 5))[0])  # M_5 -> N
    bad.trans_probs = bad.trans_probs.copy()
    bad.trans_probs[k] = np.float32(2.5)
    assert api.k0_chain_check(bad)[0]["valid"] == 0
    # an extra skip edge M_3 -> M_7 breaks the simple-path property
    skip = api.HMM(hmm.states_num, hmm.emit_num, hmm.trans_num + 1,
                   np.append(hmm.trans_rows, np.uint64(3)), np.append(hmm.trans_cols, np.uint64(7)),
                   np.append(hmm.trans_probs, np.float32(1.5)), hmm.emissions, hmm.non_zero_start_probs,
                   hmm.start_probabilities_cols, hmm.start_probabilities)
    assert api.k0_chain_check(skip)[0]["valid"] == 0


def test_chain_plan_variants(api, oracle):
    """Shuffled state numbering, a broken chain (two paths), +inf weights, one special only, unreachable states."""
    rng = np.random.default_rng(3)
    L = 37
    n = L + 2
    perm = rng.permutation(n)          # state k of the canonical model is called perm[k]
    N, C = 0, L + 1
    src, dst, w = [], [], []
    src += [N]; dst += [N]; w += [0.3]
    src += [C]; dst += [C]; w += [0.2]
    for k in range(1, L + 1):
        src += [N]; dst += [k]; w += [float(rng.uniform(1, 9))]      # hub -> chain, different weights
        src += [k]; dst += [N]; w += [1.0]                            # chain -> N, uniform
        src += [k]; dst += [C]; w += [1.25]                           # chain -> C, uniform
        if k < L and k != 17:                                          # band, broken after state 17
            src += [k]; dst += [k + 1]; w += [float(rng.uniform(0.1, 3)) if k != 9 else np.inf]
    src = perm[np.array(src)].astype(np.uint64)
    dst = perm[np.array(dst)].astype(np.uint64)
    em = rng.uniform(0.5, 7, (6, n)).astype(np.float32)
    hmm = api.HMM(n, 6, src.size, src, dst, np.array(w, np.float32), em, 2,
                  np.array([perm[N], perm[4]], np.uint64), np.array([0.0, 2.0], np.float32))
    for T in (1, 2, 3, 50):
        seq = rng.integers(0, 6, T).astype(np.uint64)
        st, out = api.k0_chain_check(hmm, seq)
        assert st["valid"] == 1 and st["chain_states"] == L and st["special0"] == perm[N]
        assert same_bits(out, oracle.plain(hmm, seq))
    # no start state at all: everything stays +inf
    dead = api.HMM(n, 6, src.size, src, dst, np.array(w, np.float32), em, 0, np.zeros(0, np.uint64), np.zeros(0, np.float32))
    st, out = api.k0_chain_check(dead, np.array([1, 2, 3], np.uint64))
    assert st["valid"] == 1 and np.all(np.isposinf(out))


# ---------------------------------------------------------------- lazy / cached plans ----
def test_quick_plan_is_fast_and_exact(api, oracle, monkeypatch):
    """The first batch runs on a short colouring search (the full one continues in the background on a GPU
    box): the quick plan must still replay to the oracle's bits and stay within 1.35x of the ideal wavefronts."""
    import time
    from spec_viterbi_b200 import generators
    hmm = generators.random_hmm(states_num=2405, seed=2405)
    seq = np.random.default_rng(1).integers(0, 20, 60).astype(np.uint64)
    monkeypatch.setenv("SV_LS_SA_ITERS", str(20 * 2405))  # the quick search of sv_model_ensure_ls
    t0 = time.perf_counter()
    stats, out = api.k0_lockstep_check(hmm, seq)
    dt = time.perf_counter() - t0
    assert stats["valid"] == 1 and same_bits(out, oracle.plain(hmm, seq))
    assert stats["wavefronts"] <= 1.35 * stats["ideal_wavefronts"]
    assert dt < 2.0  # ~0.15 s on 8 cores; the bound only catches a return to multi-second planning


def test_multi_split(api):
    """sv_multi_viterbi_batch* deals contiguous ranges of about equal symbol counts."""
    offs = np.arange(0, 4097, dtype=np.uint64) * np.uint64(7000)
    for parts in (1, 2, 4, 8):
        cut = api.multi_split(offs, parts)
        assert cut[0] == 0 and cut[-1] == 4096 and np.all(np.diff(cut.astype(np.int64)) == 4096 // parts)
    lens = np.random.default_rng(0).integers(1, 8000, 300)
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64) + np.uint64(123)  # non-zero base
    for parts in (2, 3, 8):
        cut = api.multi_split(offs, parts).astype(np.int64)
        assert cut[0] == 0 and cut[-1] == 300 and np.all(np.diff(cut) >= 0)
        loads = np.array([int(offs[cut[k + 1]] - offs[cut[k]]) for k in range(parts)])
        assert loads.max() - loads.min() <= 2 * lens.max()
    assert list(api.multi_split(np.array([0, 5, 9], np.uint64), 8)[[0, -1]]) == [0, 2]  # more devices than sequences


def test_plan_cache_round_trip(api, oracle, monkeypatch, tmp_path):
    """The on-disk plan cache: a stored lockstep plan is loaded by the next process-level use (same statistics, same
    replayed bits, no search), another model never matches it, and a damaged file is ignored and rebuilt."""
    import time
    from spec_viterbi_b200 import generators
    monkeypatch.setenv("SV_PLAN_CACHE_DIR", str(tmp_path))
    monkeypatch.setenv("SV_K0_CHECK_USE_CACHE", "1")
    monkeypatch.setenv("SV_LS_SA_ITERS", "400000")
    hmm = generators.random_hmm(states_num=900, seed=17)
    seq = np.random.default_rng(2).integers(0, 20, 80).astype(np.uint64)
    t0 = time.perf_counter()
    st1, out1 = api.k0_lockstep_check(hmm, seq)
    t_build = time.perf_counter() - t0
    files = list(tmp_path.glob("ls-*.svplan"))
    assert len(files) == 1 and st1["valid"] == 1
    t0 = time.perf_counter()
    st2, out2 = api.k0_lockstep_check(hmm, seq)
    t_load = time.perf_counter() - t0
    assert st2 == st1 and same_bits(out2, out1) and same_bits(out1, oracle.plain(hmm, seq))
    assert t_load < t_build  # no colouring search the second time
    other = generators.random_hmm(states_num=900, seed=18)
    api.k0_lockstep_check(other, seq)
    assert len(list(tmp_path.glob("ls-*.svplan"))) == 2  # keyed by content: a second file, not a false hit
    # damage the first file: truncated, then garbage of the right length -> ignored, rebuilt, replaced
    blob = files[0].read_bytes()
    files[0].write_bytes(blob[: len(blob) // 2])
    st3, out3 = api.k0_lockstep_check(hmm, seq)
    assert st3["valid"] == 1 and same_bits(out3, out1)
    files[0].write_bytes(blob[:200] + bytes(len(blob) - 200))
    st4, out4 = api.k0_lockstep_check(hmm, seq)
    assert st4["valid"] == 1 and same_bits(out4, out1)
    # one flipped payload byte with every size field intact: caught by the payload checksum (a wrong weight in an
    # accepted plan would be a silently wrong result)
    good = files[0].read_bytes()
    bad = bytearray(good)
    bad[len(bad) // 3] ^= 0x40
    files[0].write_bytes(bytes(bad))
    st5, out5 = api.k0_lockstep_check(hmm, seq)
    assert st5["valid"] == 1 and same_bits(out5, out1)
    assert files[0].read_bytes() == good  # rebuilt (deterministic search) and stored again
    monkeypatch.setenv("SV_PLAN_CACHE", "0")   # disabled: nothing is read or written
    for f in tmp_path.glob("ls-*.svplan"):
        f.unlink()
    api.k0_lockstep_check(hmm, seq)
    assert not list(tmp_path.glob("ls-*.svplan"))


@pytest.mark.parametrize("seed", range(6))
def test_chain_pair_replay_random_chain_models(seed, api, oracle):
    """Random chain + specials models -- shuffled numbering, several paths, +inf band / hub weights, non-uniform hub
    weights, optional C -> N and N -> C edges, one or two specials, a negative weight, several start states -- through
    the K4c algebra (host replay) against the level-2 oracle, and through the plain chain algebra against the plain
    oracle."""
    from conftest import random_chain_hmm
    hmm, E, rng = random_chain_hmm(api, seed)
    osp = oracle.spec(hmm, 2)
    for T in (1, 2, 3, 4, 5, 8, 33):
        seq = rng.integers(0, E, T).astype(np.uint64)
        st2, out2 = api.k0_chain_check(hmm, seq, level=2)
        st1, out1 = api.k0_chain_check(hmm, seq, level=1)
        assert st2["valid"] == 1 and st1["valid"] == 1, (seed, st1)
        assert same_bits(out2, osp.run(seq)), (seed, T)
        assert same_bits(out1, oracle.plain(hmm, seq)), (seed, T)

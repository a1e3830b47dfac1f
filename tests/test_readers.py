"""Loaders either side of the hot path: .chmm/.ess text (reference data_reader.cpp formats),
the packed caches, and the reference's reader unit tests (tests/test_chmm_reader.cpp:3-31,
tests/test_ess_reader.cpp:3-10)."""
from pathlib import Path

import numpy as np
import pytest

from conftest import GOLDEN, same_bits


def test_chmm_reader_reference_case(load_model, api):
    h = load_model("0_test_chmm")
    assert h.states_num == 3 and h.non_zero_start_probs == 2 and h.emit_num == 4 and h.trans_num == 4
    assert list(h.start_probabilities_cols) == [0, 1]
    assert same_bits(h.start_probabilities, api.to_modified_prob([0.5, 0.5]))
    assert list(h.trans_rows) == [0, 0, 1, 1] and list(h.trans_cols) == [0, 1, 0, 1]
    assert h.emissions.shape == (4, 3)  # [emit][state]


def test_ess_reader_reference_case(load_seqs):
    s = load_seqs("0_test_seq")
    assert [list(map(int, x)) for x in s] == [[2, 2, 1, 0, 1, 3, 2, 0, 0], [3, 2, 1, 0]]


def test_shipped_shapes(load_model, load_seqs):
    h = load_model("2405")
    assert (h.states_num, h.trans_num, h.emit_num) == (2407, 9621, 20)
    assert [len(s) for s in load_seqs("covid-19")] == [38, 805, 275, 419, 61, 121, 43, 121, 97, 7096,
                                                      4405, 1273, 492, 75, 222, 73]
    assert [len(s) for s in load_seqs("emit_50_3500_20")] == [3500] * 50


def test_text_round_trip(tmp_path, api):
    from spec_viterbi_b200 import generators
    probs = dict(states=3, start=[(0, 1.0), (2, 0.25)],
                 emissions=[[0.5, 0.5], [0.1, 0.9], [0.0, 1.0]],
                 transitions=[(0, 1, 0.3), (0, 2, 0.7), (1, 1, 1.0), (2, 0, 0.0)])
    p = tmp_path / "m.chmm"
    generators.write_chmm(probs, str(p))
    h = api.read_HMM(str(p))
    assert h.states_num == 3 and h.emit_num == 2 and h.trans_num == 4
    assert np.signbit(h.start_probabilities[0]) and h.start_probabilities[0] == 0  # -0.0 for p = 1
    assert np.isposinf(h.emissions[0, 2]) and np.isposinf(h.trans_probs[3])
    assert same_bits(h.trans_probs[:3], api.to_modified_prob([0.3, 0.7, 1.0]))
    assert same_bits(h.emissions[:, 1], api.to_modified_prob([0.1, 0.9]))
    # packed cache round trip is bit exact
    pb = tmp_path / "m.chmmb"
    api.pack_HMM_file(str(p), str(pb))
    h2 = api.read_HMM(str(pb))
    for f in ("trans_rows", "trans_cols", "trans_probs", "emissions", "start_probabilities",
              "start_probabilities_cols"):
        assert np.array_equal(getattr(h, f), getattr(h2, f)), f
    seqs = [[0, 1, 1, 0] * 60, [1], [0] * 7]
    e = tmp_path / "s.ess"
    generators.write_ess(seqs, str(e))
    got = api.read_emit_seq(str(e))
    assert [list(map(int, x)) for x in got] == seqs
    api.pack_ess_file(str(e), str(tmp_path / "s.essb"))
    got = api.read_emit_seq(str(tmp_path / "s.essb"))
    assert [list(map(int, x)) for x in got] == seqs


def test_bad_files(tmp_path, api):
    from spec_viterbi_b200._lib import SvError
    with pytest.raises(SvError):
        api.read_HMM(str(tmp_path / "missing.chmm"))
    bad = tmp_path / "bad.ess"
    bad.write_text("2\n0 2\n1 1\n5 1\n0\n")  # running index broken (data_reader.cpp:113-119)
    with pytest.raises(SvError):
        api.read_emit_seq(str(bad))


def test_readers_match_reference_reader(api):
    """Bit-for-bit against the reference's own read_HMM where the reference tree is mounted."""
    from oracle import pyoracle
    ref_root = Path("/root/reference")
    if not (ref_root.exists() and pyoracle.reference_available()):
        pytest.skip("reference tree absent")
    ref = pyoracle.Reference()
    for name in ("100", "2405"):
        r = ref.read_hmm(str(ref_root / "chmm_files" / f"{name}.chmm"))
        for src in (str(ref_root / "chmm_files" / f"{name}.chmm"), str(GOLDEN / "models" / f"{name}.chmmb")):
            h = api.read_HMM(src)
            assert h.states_num == r["states_num"] and h.trans_num == r["trans_num"]
            for f in ("trans_rows", "trans_cols", "trans_probs", "emissions", "start_probabilities",
                      "start_probabilities_cols"):
                assert np.array_equal(getattr(h, f), r[f]), (name, f)


def test_corrupt_packed_files_are_rejected(api, tmp_path):
    """Packed caches are validated against the file size before anything is allocated or copied: a bad symbol
    width, non-monotonic offsets, oversized counts and out-of-range indices give an error, never a crash."""
    import struct
    from spec_viterbi_b200 import _lib
    good = (GOLDEN / "seqs" / "0_test_seq.essb").read_bytes()
    nseq, total, width = struct.unpack_from("<3Q", good, 8)
    assert width == 1

    def essb(nseq_, total_, width_, offs, payload):
        return good[:8] + struct.pack("<3Q", nseq_, total_, width_) + struct.pack(f"<{len(offs)}Q", *offs) + payload

    offs = list(struct.unpack_from(f"<{nseq + 1}Q", good, 32))
    payload = good[32 + 8 * (nseq + 1):]
    cases = {
        "width8": essb(nseq, total, 8, offs, payload * 8),
        "width3": essb(nseq, total, 3, offs, payload * 3),
        "huge": essb(1 << 60, total, 1, offs, payload),
        "nonmono": essb(nseq, total, 1, [0] + [total + 7] * (nseq - 1) + [total], payload),
        "short": good[:-1],
        "offs_beyond": essb(nseq, total, 1, offs[:-1] + [total + 7], payload),
    }
    for name, blob in cases.items():
        p = tmp_path / f"{name}.essb"
        p.write_bytes(blob)
        with pytest.raises(_lib.SvError):
            api.read_emit_seq(str(p))
    hm = (GOLDEN / "models" / "0_test_chmm.chmmb").read_bytes()
    states, emits, trans, nstart = struct.unpack_from("<4Q", hm, 8)
    for name, hdr in {"states": (1 << 50, emits, trans, nstart), "trans": (states, emits, 1 << 61, nstart),
                      "zero": (0, emits, trans, nstart)}.items():
        p = tmp_path / f"{name}.chmmb"
        p.write_bytes(hm[:8] + struct.pack("<4Q", *hdr) + hm[40:])
        with pytest.raises(_lib.SvError):
            api.read_HMM(str(p))
    # a transition index >= states
    bad = bytearray(hm)
    at = 40 + nstart * 8 + states * emits * 4
    struct.pack_into("<I", bad, at, states + 3)
    (tmp_path / "idx.chmmb").write_bytes(bytes(bad))
    with pytest.raises(_lib.SvError):
        api.read_HMM(str(tmp_path / "idx.chmmb"))

"""The C-ABI library loads and exports every symbol include/spec_viterbi_b200.h declares;
on a host without a GPU every compute call fails loudly (no CPU fallback)."""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent


def declared_functions():
    text = (ROOT / "include" / "spec_viterbi_b200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(sv_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_expected_surface():
    names = declared_functions()
    for must in ("sv_model_create", "sv_viterbi", "sv_viterbi_batch", "sv_spec_create", "sv_spec_run",
                 "sv_spec_run_batch", "sv_last_error"):
        assert must in names


def test_library_exports_every_declared_symbol():
    from spec_viterbi_b200 import _lib
    lib = _lib.lib()
    for name in declared_functions():
        assert hasattr(lib, name), f"{name} declared in the header but not exported"
    # and the ctypes table covers the header
    assert set(declared_functions()) <= set(_lib.SIGNATURES)


def test_product_never_references_oracle():
    pkg = ROOT / "spec_viterbi_b200"
    for p in list(pkg.rglob("*.py")) + list(pkg.rglob("*.cu")) + list(pkg.rglob("*.cpp")) + list(pkg.rglob("*.h")):
        s = p.read_text()
        assert "pyoracle" not in s and "sv_oracle" not in s and "libsv_ref" not in s, p


def test_no_gpu_fails_loudly(api):
    from spec_viterbi_b200 import _lib
    lib = _lib.lib()
    if lib.sv_device_count() > 0:
        pytest.skip("a GPU is present")
    hmm = api.read_HMM(str(ROOT / "tests/golden/models/0_test_chmm.chmmb"))
    with pytest.raises(_lib.SvError) as ei:
        api.Model(hmm)
    assert ei.value.code == 2 and "no CPU fallback" in str(ei.value)
    with pytest.raises(_lib.SvError):
        api.B200_impl().run_Viterbi(hmm, np.array([0, 1], np.uint64))

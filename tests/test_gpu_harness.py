"""The C++ drop-in: B200_impl / B200_spec_impl (host/B200_impl.h) behind the reference's abstract
classes, driven by the clone of tests/test_semantic_equality.cpp over the shipped models."""
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
pytestmark = pytest.mark.gpu


def test_semantic_equality_cpp(tmp_path):
    exe = ROOT / "harness" / "test_semantic_equality"
    if not exe.exists():
        subprocess.run(["make", "-C", str(ROOT / "harness")], check=True, capture_output=True)
    r = subprocess.run([str(exe), str(ROOT / "tests/golden/models"), str(ROOT / "tests/golden/seqs/emit_3_3500_20.essb"), "250"],
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "PASSED: 24 models x 3 sequences" in r.stdout

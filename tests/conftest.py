"""Shared fixtures.  `-m "not gpu"` covers the oracle, host logic and the ABI surface; `-m gpu`
are the parity tests proper (CUDA path through the C ABI vs the oracle / golden vectors)."""
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
GOLDEN = ROOT / "tests" / "golden"

PROFILE_MODELS = ["100", "200", "300", "400", "500", "600", "700", "800", "900", "1001", "1100",
                  "1200", "1301", "1400", "1509", "1600", "1705", "1799", "1901", "2050", "2138",
                  "2207", "2365", "2405"]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def api():
    from spec_viterbi_b200 import api as _api
    return _api


@pytest.fixture(scope="session")
def oracle():
    from oracle import pyoracle
    return pyoracle.Oracle()


@pytest.fixture(scope="session")
def golden():
    return dict(np.load(GOLDEN / "expected" / "reference_outputs.npz"))


@pytest.fixture(scope="session")
def load_model(api):
    cache = {}

    def _load(name):
        if name not in cache:
            cache[name] = api.read_HMM(str(GOLDEN / "models" / f"{name}.chmmb"))
        return cache[name]
    return _load


@pytest.fixture(scope="session")
def load_seqs(api):
    cache = {}

    def _load(name):
        if name not in cache:
            cache[name] = api.read_emit_seq(str(GOLDEN / "seqs" / f"{name}.essb"))
        return cache[name]
    return _load


def same_bits(a, b):
    """Bit-exact comparison with +0 == -0 (start score of p=1.0 is -0.0)."""
    a = np.asarray(a, np.float32)
    b = np.asarray(b, np.float32)
    return a.shape == b.shape and bool(np.all((a == b) | (np.isnan(a) & np.isnan(b))))


def best_state(v):
    v = np.asarray(v)
    if np.all(np.isposinf(v)):
        return np.float32(np.inf), -1
    i = int(np.argmin(v))  # lowest index attaining the minimum
    return v[i], i

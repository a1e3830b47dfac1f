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


def random_chain_hmm(api, seed):
    """A random chain + specials model (the structure of the reference's profile HMMs): shuffled numbering,
    several paths, +inf band / hub weights, non-uniform hub weights, optional C -> N and N -> C edges, one or two
    specials, a negative weight (seed 4), two start states.  -> (hmm, emits, rng)"""
    rng = np.random.default_rng(100 + seed)
    L = int(rng.integers(7, 60))
    two = seed % 3 != 0                       # seeds 0, 3: a single special state
    n = L + (2 if two else 1)
    E = int(rng.integers(2, 6))
    N, C = 0, L + 1
    src, dst, w = [N], [N], [float(rng.uniform(0.1, 2))]
    if two:
        src += [C]; dst += [C]; w += [float(rng.uniform(0.1, 2))]
        if seed % 2:
            src += [C, N]; dst += [N, C]; w += [float(rng.uniform(1, 5)), float(rng.uniform(1, 5))]
    uN, uC = float(rng.uniform(0.5, 2)), float(rng.uniform(0.5, 2))
    for k in range(1, L + 1):
        src += [N, k]; dst += [k, N]
        w += [float(rng.uniform(1, 9)) if rng.random() > 0.15 else np.inf, uN]
        if two:
            src += [k]; dst += [C]; w += [uC]
        if k < L and rng.random() > 0.2:
            src += [k]; dst += [k + 1]
            w += [float(rng.uniform(0.1, 3)) if rng.random() > 0.1 else np.inf]
    if seed == 4:
        w[2] = -0.25                           # a "probability" > 1 on the first hub -> chain edge (index 2: N -> 1)
    perm = rng.permutation(n)
    hmm = api.HMM(n, E, len(src), perm[np.array(src)].astype(np.uint64), perm[np.array(dst)].astype(np.uint64),
                  np.array(w, np.float32), rng.uniform(0.2, 7, (E, n)).astype(np.float32), 2,
                  np.array([perm[N], perm[min(3, L)]], np.uint64), np.array([0.0, 1.5], np.float32))
    return hmm, E, rng

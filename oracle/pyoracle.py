"""ctypes bindings of the CPU checkers (TEST INFRASTRUCTURE -- see oracle/sv_oracle.h).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  `Oracle` = this repo's C restatement (libsv_oracle.so, always buildable);
`Reference` = the reference's own GraphBLAS_* sources compiled unmodified against the
GraphBLAS shim (oracle/_ref/libsv_ref.so, built where /root/reference exists, prebuilt
elsewhere).
"""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
ORACLE_SO = HERE / "libsv_oracle.so"
REF_SO = HERE / "_ref" / "libsv_ref.so"

u64p = C.POINTER(C.c_uint64)
f32p = C.POINTER(C.c_float)


class SvoModel(C.Structure):
    _fields_ = [("states", C.c_uint64), ("emits", C.c_uint64), ("trans", C.c_uint64),
                ("nstart", C.c_uint64), ("trans_src", u64p), ("trans_dst", u64p),
                ("trans_w", f32p), ("emis", f32p), ("start_cols", u64p), ("start_w", f32p)]


def build(force: bool = False) -> None:
    """make -C oracle (C restatement always; _ref only where the reference tree exists)."""
    if force or not ORACLE_SO.exists() or (Path("/root/reference").exists() and not REF_SO.exists()):
        subprocess.run(["make", "-C", str(HERE)], check=True, capture_output=True)


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


class _ModelView:
    """Keeps the numpy arrays alive next to the C struct that points into them."""

    def __init__(self, hmm):
        h = hmm.normalized() if hasattr(hmm, "normalized") else hmm
        self.h = h
        self.n = int(h.states_num)
        self.c = SvoModel(h.states_num, h.emit_num, h.trans_num, h.non_zero_start_probs,
                          _p(h.trans_rows, C.c_uint64), _p(h.trans_cols, C.c_uint64),
                          _p(h.trans_probs, C.c_float), _p(h.emissions, C.c_float),
                          _p(h.start_probabilities_cols, C.c_uint64),
                          _p(h.start_probabilities, C.c_float))


class Oracle:
    def __init__(self):
        build()
        lib = C.CDLL(str(ORACLE_SO))
        lib.svo_plain.restype = C.c_int
        lib.svo_plain.argtypes = [C.POINTER(SvoModel), u64p, C.c_uint64, f32p]
        lib.svo_plain_batch.restype = C.c_int
        lib.svo_plain_batch.argtypes = [C.POINTER(SvoModel), u64p, u64p, C.c_uint64, f32p, C.c_int]
        lib.svo_spec_create.restype = C.c_void_p
        lib.svo_spec_create.argtypes = [C.POINTER(SvoModel), C.c_uint64, C.c_int]
        lib.svo_spec_destroy.argtypes = [C.c_void_p]
        lib.svo_spec_run.restype = C.c_int
        lib.svo_spec_run.argtypes = [C.c_void_p, u64p, C.c_uint64, f32p]
        lib.svo_spec_handler.restype = f32p
        lib.svo_spec_handler.argtypes = [C.c_void_p, C.c_uint64]
        self.lib = lib

    def plain(self, hmm, seq) -> np.ndarray:
        mv = _ModelView(hmm)
        s = np.ascontiguousarray(seq, np.uint64)
        out = np.empty(mv.n, np.float32)
        rc = self.lib.svo_plain(C.byref(mv.c), _p(s, C.c_uint64), s.size, _p(out, C.c_float))
        if rc != 0:
            raise RuntimeError(f"svo_plain failed: {rc}")
        return out

    def plain_batch(self, hmm, symbols, offsets, threads: int = 1) -> np.ndarray:
        mv = _ModelView(hmm)
        symbols = np.ascontiguousarray(symbols, np.uint64)
        offsets = np.ascontiguousarray(offsets, np.uint64)
        nseq = offsets.size - 1
        out = np.empty((nseq, mv.n), np.float32)
        rc = self.lib.svo_plain_batch(C.byref(mv.c), _p(symbols, C.c_uint64), _p(offsets, C.c_uint64),
                                      nseq, _p(out, C.c_float), threads)
        if rc != 0:
            raise RuntimeError(f"svo_plain_batch failed: {rc}")
        return out

    def spec(self, hmm, level: int, eager: bool = False) -> "OracleSpec":
        return OracleSpec(self, hmm, level, eager)


class OracleSpec:
    def __init__(self, oracle: Oracle, hmm, level: int, eager: bool):
        self.o = oracle
        self.mv = _ModelView(hmm)
        self.h = oracle.lib.svo_spec_create(C.byref(self.mv.c), level, 1 if eager else 0)
        if not self.h:
            raise RuntimeError("svo_spec_create failed")

    def run(self, seq) -> np.ndarray:
        s = np.ascontiguousarray(seq, np.uint64)
        out = np.empty(self.mv.n, np.float32)
        rc = self.o.lib.svo_spec_run(self.h, _p(s, C.c_uint64), s.size, _p(out, C.c_float))
        if rc != 0:
            raise RuntimeError(f"svo_spec_run failed: {rc}")
        return out

    def handler(self, tuple_index: int) -> np.ndarray:
        ptr = self.o.lib.svo_spec_handler(self.h, tuple_index)
        if not ptr:
            raise RuntimeError("svo_spec_handler failed")
        n = self.mv.n
        return np.ctypeslib.as_array(ptr, shape=(n, n)).copy()

    def close(self):
        if self.h:
            self.o.lib.svo_spec_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def reference_available() -> bool:
    build()
    return REF_SO.exists()


class Reference:
    """The reference's own CPU backends (GraphBLAS_impl / GraphBLAS_spec_impl on the shim)."""

    def __init__(self):
        build()
        if not REF_SO.exists():
            raise RuntimeError("oracle/_ref/libsv_ref.so is not built (reference tree absent)")
        lib = C.CDLL(str(REF_SO))
        lib.svref_plain.restype = C.c_int
        lib.svref_plain.argtypes = [C.POINTER(SvoModel), u64p, C.c_uint64, f32p]
        lib.svref_plain_batch.restype = C.c_int
        lib.svref_plain_batch.argtypes = [C.POINTER(SvoModel), u64p, u64p, C.c_uint64, f32p, C.c_int]
        lib.svref_spec_create.restype = C.c_void_p
        lib.svref_spec_create.argtypes = [C.POINTER(SvoModel), C.c_uint64]
        lib.svref_spec_respec.restype = C.c_int
        lib.svref_spec_respec.argtypes = [C.c_void_p, C.POINTER(SvoModel)]
        lib.svref_spec_run.restype = C.c_int
        lib.svref_spec_run.argtypes = [C.c_void_p, u64p, C.c_uint64, C.c_uint64, f32p]
        lib.svref_spec_run_batch.restype = C.c_int
        lib.svref_spec_run_batch.argtypes = [C.c_void_p, u64p, u64p, C.c_uint64, C.c_uint64, f32p, C.c_int]
        lib.svref_spec_destroy.argtypes = [C.c_void_p]
        lib.svref_read_hmm.restype = C.c_void_p
        lib.svref_read_hmm.argtypes = [C.c_char_p]
        lib.svref_hmm_dims.argtypes = [C.c_void_p, u64p]
        lib.svref_hmm_copy.argtypes = [C.c_void_p, u64p, u64p, f32p, f32p, u64p, f32p]
        lib.svref_hmm_free.argtypes = [C.c_void_p]
        self.lib = lib

    def plain(self, hmm, seq) -> np.ndarray:
        mv = _ModelView(hmm)
        s = np.ascontiguousarray(seq, np.uint64)
        out = np.empty(mv.n, np.float32)
        rc = self.lib.svref_plain(C.byref(mv.c), _p(s, C.c_uint64), s.size, _p(out, C.c_float))
        if rc != 0:
            raise RuntimeError("svref_plain failed")
        return out

    def plain_batch(self, hmm, symbols, offsets, threads: int = 1) -> np.ndarray:
        mv = _ModelView(hmm)
        symbols = np.ascontiguousarray(symbols, np.uint64)
        offsets = np.ascontiguousarray(offsets, np.uint64)
        nseq = offsets.size - 1
        out = np.empty((nseq, mv.n), np.float32)
        rc = self.lib.svref_plain_batch(C.byref(mv.c), _p(symbols, C.c_uint64), _p(offsets, C.c_uint64),
                                        nseq, _p(out, C.c_float), threads)
        if rc != 0:
            raise RuntimeError("svref_plain_batch failed")
        return out

    def spec_create(self, hmm, level: int):
        mv = _ModelView(hmm)
        h = self.lib.svref_spec_create(C.byref(mv.c), level)
        return (h, mv)

    def spec_run(self, spec, seq) -> np.ndarray:
        h, mv = spec
        s = np.ascontiguousarray(seq, np.uint64)
        out = np.empty(mv.n, np.float32)
        rc = self.lib.svref_spec_run(h, _p(s, C.c_uint64), s.size, mv.n, _p(out, C.c_float))
        if rc != 0:
            raise RuntimeError("svref_spec_run failed")
        return out

    def spec_run_batch(self, spec, symbols, offsets, threads: int = 1) -> np.ndarray:
        h, mv = spec
        symbols = np.ascontiguousarray(symbols, np.uint64)
        offsets = np.ascontiguousarray(offsets, np.uint64)
        nseq = offsets.size - 1
        out = np.empty((nseq, mv.n), np.float32)
        rc = self.lib.svref_spec_run_batch(h, _p(symbols, C.c_uint64), _p(offsets, C.c_uint64), nseq,
                                           mv.n, _p(out, C.c_float), threads)
        if rc != 0:
            raise RuntimeError("svref_spec_run_batch failed")
        return out

    def spec_destroy(self, spec) -> None:
        self.lib.svref_spec_destroy(spec[0])

    def read_hmm(self, path: str):
        """The reference's read_HMM -> dict of numpy arrays."""
        h = self.lib.svref_read_hmm(str(path).encode())
        dims = (C.c_uint64 * 4)()
        self.lib.svref_hmm_dims(h, dims)
        n, e, t, s = (int(x) for x in dims)
        out = dict(states_num=n, emit_num=e, trans_num=t, non_zero_start_probs=s,
                   trans_rows=np.zeros(t, np.uint64), trans_cols=np.zeros(t, np.uint64),
                   trans_probs=np.zeros(t, np.float32), emissions=np.zeros((e, n), np.float32),
                   start_probabilities_cols=np.zeros(s, np.uint64),
                   start_probabilities=np.zeros(s, np.float32))
        self.lib.svref_hmm_copy(h, _p(out["trans_rows"], C.c_uint64), _p(out["trans_cols"], C.c_uint64),
                                _p(out["trans_probs"], C.c_float), _p(out["emissions"], C.c_float),
                                _p(out["start_probabilities_cols"], C.c_uint64),
                                _p(out["start_probabilities"], C.c_float))
        self.lib.svref_hmm_free(h)
        return out

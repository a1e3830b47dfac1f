"""ctypes binding of libspecviterbi_b200.so (the C ABI in include/spec_viterbi_b200.h).

The library is the product: if it is missing it is built with nvcc (spec_viterbi_b200.build);
if it cannot be loaded every call fails loudly -- there is no Python or CPU fallback.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

_PKG = Path(__file__).resolve().parent
LIB_PATH = _PKG / "libspecviterbi_b200.so"

SV_OK = 0
KERNEL_AUTO, KERNEL_GENERIC, KERNEL_ES, KERNEL_LS, KERNEL_CHAIN = 0, 1, 2, 3, 4
SPEC_AUTO, SPEC_DENSE_GEMM, SPEC_SPARSE_LHS, SPEC_PAIR_ROWS, SPEC_ONDEMAND = 0, 1, 2, 3, 4
SPEC_RUN_TABLE, SPEC_RUN_IMPLICIT, SPEC_RUN_CHAIN, SPEC_RUN_AUTO = 0, 1, 2, 3

u64p = C.POINTER(C.c_uint64)
u8p = C.POINTER(C.c_uint8)
f32p = C.POINTER(C.c_float)
i32p = C.POINTER(C.c_int32)
vp = C.c_void_p


class ModelInfo(C.Structure):
    _fields_ = [
        ("states", C.c_uint64), ("emits", C.c_uint64), ("nnz", C.c_uint64),
        ("max_in_degree", C.c_uint64),
        ("device", C.c_int32), ("kernel", C.c_int32),
        ("es_threads", C.c_int32), ("es_seqs_per_cta", C.c_int32),
        ("es_ranks", C.c_int32), ("es_heavy_rows", C.c_int32),
        ("es_slots_per_thread", C.c_uint64),
        ("es_gather_wavefronts", C.c_uint64), ("es_gather_ideal", C.c_uint64),
        ("es_smem_bytes", C.c_uint64), ("es_split_states", C.c_uint64),
        ("ls_threads", C.c_int32), ("ls_seqs_per_cta", C.c_int32),
        ("ls_ranks", C.c_int32), ("ls_heavy_rows", C.c_int32),
        ("ls_tmem_cols", C.c_int32), ("ls_reserved", C.c_int32),
        ("ls_gather_wavefronts", C.c_uint64), ("ls_gather_ideal", C.c_uint64),
        ("ls_smem_bytes", C.c_uint64), ("ls_split_states", C.c_uint64),
        ("ls_colour_cost_before", C.c_uint64), ("ls_colour_cost_after", C.c_uint64),
        ("ls_sa_iters", C.c_uint64),
        ("chain_threads", C.c_int32), ("chain_states", C.c_int32), ("chain_specials", C.c_int32),
        ("chain_seqs_per_cta", C.c_int32), ("chain_tmem_cols", C.c_int32), ("chain_reserved", C.c_int32),
        ("plan_ms", C.c_double),
    ]


# name -> (restype, argtypes); this table is also what tests check against the header
SIGNATURES = {
    "sv_last_error": (C.c_char_p, []),
    "sv_version": (C.c_char_p, []),
    "sv_device_count": (C.c_int, []),
    "sv_model_create": (C.c_int, [C.c_uint64, C.c_uint64, C.c_uint64, u64p, u64p, f32p, f32p,
                                  C.c_uint64, u64p, f32p, C.c_int, C.POINTER(vp)]),
    "sv_model_destroy": (None, [vp]),
    "sv_model_set_kernel": (C.c_int, [vp, C.c_int]),
    "sv_model_prepare": (C.c_int, [vp]),
    "sv_model_get_info": (C.c_int, [vp, C.POINTER(ModelInfo)]),
    "sv_model_last_launches": (C.c_int, [vp]),
    "sv_seqs_create": (C.c_int, [vp, u64p, u64p, C.c_uint64, C.POINTER(vp)]),
    "sv_seqs_create_u8": (C.c_int, [vp, u8p, u64p, C.c_uint64, C.POINTER(vp)]),
    "sv_seqs_destroy": (None, [vp]),
    "sv_seqs_count": (C.c_uint64, [vp]),
    "sv_seqs_total_symbols": (C.c_uint64, [vp]),
    "sv_viterbi": (C.c_int, [vp, u64p, C.c_uint64, f32p]),
    "sv_viterbi_batch": (C.c_int, [vp, u64p, u64p, C.c_uint64, f32p, f32p, i32p]),
    "sv_viterbi_batch_u8": (C.c_int, [vp, u8p, u64p, C.c_uint64, f32p, f32p, i32p]),
    "sv_viterbi_device": (C.c_int, [vp, vp, vp, vp, vp, vp]),
    "sv_viterbi_device_static": (C.c_int, [vp, vp, vp, vp, vp, vp]),
    "sv_spec_create": (C.c_int, [vp, C.c_uint64, C.c_int, C.c_uint64, C.c_uint64, vp, vp,
                                 C.POINTER(vp)]),
    "sv_spec_destroy": (None, [vp]),
    "sv_spec_level": (C.c_uint64, [vp]),
    "sv_spec_set_range": (C.c_int, [vp, vp, C.c_uint64, C.c_uint64, C.c_uint64]),
    "sv_spec_set_run_mode": (C.c_int, [vp, C.c_int]),
    "sv_spec_table": (vp, [vp, u64p, u64p]),
    "sv_spec_get_handler": (C.c_int, [vp, C.c_uint64, f32p]),
    "sv_spec_save": (C.c_int, [vp, C.c_char_p]),
    "sv_spec_load": (C.c_int, [vp, C.c_char_p, vp, C.POINTER(vp)]),
    "sv_spec_run": (C.c_int, [vp, u64p, C.c_uint64, f32p]),
    "sv_spec_run_batch": (C.c_int, [vp, u64p, u64p, C.c_uint64, f32p, f32p, i32p]),
    "sv_spec_run_batch_u8": (C.c_int, [vp, u8p, u64p, C.c_uint64, f32p, f32p, i32p]),
    "sv_spec_run_device": (C.c_int, [vp, vp, vp, vp, vp, vp]),
    "sv_io_hmm_open": (C.c_int, [C.c_char_p, C.POINTER(vp), u64p]),
    "sv_io_hmm_copy": (C.c_int, [vp, u64p, u64p, f32p, f32p, u64p, f32p]),
    "sv_io_hmm_write_packed": (C.c_int, [vp, C.c_char_p]),
    "sv_io_hmm_close": (None, [vp]),
    "sv_io_ess_open": (C.c_int, [C.c_char_p, C.POINTER(vp), u64p, u64p]),
    "sv_io_ess_copy": (C.c_int, [vp, u64p, u64p]),
    "sv_io_ess_write_packed": (C.c_int, [vp, C.c_char_p]),
    "sv_io_ess_close": (None, [vp]),
    "sv_k0_lockstep_check": (C.c_int, [C.c_uint64, C.c_uint64, C.c_uint64, u64p, u64p, f32p, f32p,
                                       C.c_uint64, u64p, f32p, u64p, C.c_uint64, f32p, u64p]),
    "sv_k0_chain_check": (C.c_int, [C.c_uint64, C.c_uint64, C.c_uint64, u64p, u64p, f32p, f32p,
                                    C.c_uint64, u64p, f32p, u64p, C.c_uint64, f32p, u64p]),
    "sv_k0_chain_pair_check": (C.c_int, [C.c_uint64, C.c_uint64, C.c_uint64, u64p, u64p, f32p, f32p,
                                         C.c_uint64, u64p, f32p, u64p, C.c_uint64, f32p, u64p]),
    "sv_multi_create": (C.c_int, [C.c_uint64, C.c_uint64, C.c_uint64, u64p, u64p, f32p, f32p,
                                  C.c_uint64, u64p, f32p, C.POINTER(C.c_int), C.c_int, C.POINTER(vp)]),
    "sv_multi_destroy": (None, [vp]),
    "sv_multi_device_count": (C.c_int, [vp]),
    "sv_multi_model": (vp, [vp, C.c_int]),
    "sv_multi_prepare": (C.c_int, [vp]),
    "sv_multi_viterbi_batch": (C.c_int, [vp, u64p, u64p, C.c_uint64, f32p, f32p, i32p]),
    "sv_multi_viterbi_batch_u8": (C.c_int, [vp, u8p, u64p, C.c_uint64, f32p, f32p, i32p]),
    "sv_multi_split": (C.c_int, [u64p, C.c_uint64, C.c_int, u64p]),
}

# the micro-benchmark library (include/sv_probe_b200.h) -- separate from the product
PROBE_LIB_PATH = _PKG / "libsvprobe_b200.so"
PROBE_SIGNATURES = {
    "sv_probe_last_error": (C.c_char_p, []),
    "sv_probe_tmem": (C.c_int, [C.c_int, C.c_int, C.POINTER(C.c_uint), C.POINTER(C.c_double)]),
    "sv_probe_fp32": (C.c_int, [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "sv_probe_smem": (C.c_int, [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
}
_probe = None


def probe_lib() -> C.CDLL:
    global _probe
    if _probe is None:
        if not PROBE_LIB_PATH.exists():
            from . import build as _build
            _build.build()
        handle = C.CDLL(str(PROBE_LIB_PATH))
        for name, (res, args) in PROBE_SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _probe = handle
    return _probe

_lib = None


class SvError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"spec_viterbi_b200 error {code}: {msg}")
        self.code = code


def lib() -> C.CDLL:
    """Loads (building first if necessary) the native library; raises if that fails."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            from . import build as _build
            _build.build()
        handle = C.CDLL(str(LIB_PATH))
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)  # AttributeError if the symbol is not exported
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(rc: int) -> None:
    if rc != SV_OK:
        raise SvError(rc, lib().sv_last_error().decode("utf-8", "replace"))

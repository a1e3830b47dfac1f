"""spec_viterbi_b200 -- B200-native (min,+) Viterbi backend behind the Spec_Viterbi interface.

The product is libspecviterbi_b200.so (hand-written sm_100a CUDA behind the C ABI in
include/spec_viterbi_b200.h); this package is the Python mirror of the reference's
operator interface plus seeded restatements of its data generators.
"""
from .api import (HMM, B200_impl, B200_spec_impl, Model, Seqs, Spec, almost_equal, pack_sequences,
                  read_emit_seq, read_HMM, to_modified_prob, zero_prob)
from ._lib import (KERNEL_AUTO, KERNEL_ES, KERNEL_GENERIC, KERNEL_LS, SPEC_AUTO, SPEC_DENSE_GEMM,
                   SPEC_SPARSE_LHS, SvError)

__all__ = [
    "HMM", "B200_impl", "B200_spec_impl", "Model", "Seqs", "Spec", "almost_equal",
    "pack_sequences", "read_emit_seq", "read_HMM", "to_modified_prob", "zero_prob",
    "KERNEL_AUTO", "KERNEL_ES", "KERNEL_GENERIC", "KERNEL_LS", "SPEC_AUTO", "SPEC_DENSE_GEMM",
    "SPEC_SPARSE_LHS", "SvError",
]

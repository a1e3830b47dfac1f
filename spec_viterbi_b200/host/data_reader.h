// Loaders for the reference's text formats (.chmm / .ess) and for this repo's packed
// binary caches (.chmmb / .essb).
//
// Text formats follow reference `Viterbi_impl/data_reader.cpp:7-15` (.chmm) and
// `:81-91` (.ess); semantics mirrored: every probability is converted with
// HMM::to_modified_prob, emissions are transposed to [symbol][state], an unopenable
// file or an out-of-order sequence index yields an empty object.
#pragma once

#include "HMM.h"

#include <string>

// Text .chmm -> HMM (probabilities stored as -log2).  Empty HMM on failure.
HMM read_HMM(const std::string& HMM_file_name);

// Text .ess -> sequences.  Empty vector on failure.
HMM::Emit_seq_vec_t read_emit_seq(const std::string& emit_seq_file_name);

// Packed binary caches (bit-exact round trip of the already converted weights).
bool write_HMM_packed(const HMM& hmm, const std::string& file_name);
HMM read_HMM_packed(const std::string& file_name);
bool write_emit_seq_packed(const HMM::Emit_seq_vec_t& seqs, const std::string& file_name);
HMM::Emit_seq_vec_t read_emit_seq_packed(const std::string& file_name);

// Dispatch on extension: ".chmmb"/".essb" -> packed, anything else -> text.
HMM read_HMM_any(const std::string& file_name);
HMM::Emit_seq_vec_t read_emit_seq_any(const std::string& file_name);

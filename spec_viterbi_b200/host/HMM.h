// Host-side HMM model type for the B200 Viterbi backend.
//
// Clean-room mirror of the reference boundary type so that code written against
// the reference (`Viterbi_impl/HMM.h:7-60`) compiles unchanged against this header:
// same member names, same typedef names, same static helpers, same semantics
//   * probabilities are stored as -log2(p) in fp32 ("modified probability"),
//   * probability 0 is +inf (`zero_prob`),
//   * transitions are COO triplets (src = trans_rows, dst = trans_cols),
//   * emissions are indexed [symbol][state],
//   * the start vector is sparse (cols + values).
// Nothing else from the reference is needed by the hot path.
#pragma once

#include <cmath>
#include <cstddef>
#include <limits>
#include <vector>

class HMM {
  public:
    // ---- scalar and container types (reference: HMM.h:9-16) ----
    using Probability_t = float; // plain probability as read from a .chmm file
    using Mod_prob_t = float;    // -log2(probability)
    using Index_t = std::size_t;
    using Emit_t = std::size_t;
    using Mod_prob_vec_t = std::vector<Mod_prob_t>;
    using Index_vec_t = std::vector<Index_t>;
    using Emit_seq_t = std::vector<Emit_t>;
    using Emit_seq_vec_t = std::vector<Emit_seq_t>;

    // Hash of a symbol tuple; the reference keys its level-L handler map with it
    // (HMM.h:18-26).  The B200 backend indexes handlers directly (base-E number)
    // and only keeps this for source compatibility.
    struct Emit_seq_hasher {
        std::size_t operator()(const Emit_seq_t& key) const {
            std::size_t h = key.size();
            for (const auto sym : key) {
                h ^= sym + 0x9e3779b9 + (h << 6) + (h >> 2);
            }
            return h;
        }
    };

    // ---- model data (reference: HMM.h:28-38) ----
    Index_t states_num = 0;
    Index_t emit_num = 0;
    Index_t trans_num = 0;
    Index_vec_t trans_rows;                // source state of each transition
    Index_vec_t trans_cols;                // destination state of each transition
    Mod_prob_vec_t trans_probs;            // -log2 P(src -> dst)
    std::vector<Mod_prob_vec_t> emissions; // [symbol][state] -> -log2 P(symbol | state)
    Index_t non_zero_start_probs = 0;
    Index_vec_t start_probabilities_cols; // states with a non-zero start probability
    Mod_prob_vec_t start_probabilities;   // their -log2 start probabilities

    // ---- helpers on modified probabilities (reference: HMM.h:41-59) ----
    static constexpr Mod_prob_t zero_prob = std::numeric_limits<Mod_prob_t>::infinity();

    // The reference's own cross-backend tolerance: both +inf, or |x-y| <= 1.0.
    static bool almost_equal(Mod_prob_t x, Mod_prob_t y) {
        if (x == zero_prob && y == zero_prob) {
            return true;
        }
        return std::fabs(x - y) <= 1.0;
    }

    // fp32 -log2; non-positive probabilities map to +inf.  p == 1 gives -0.0f, exactly
    // as the reference's `-1 * std::log2(x)` does; the device code treats +0 and -0 alike.
    static Mod_prob_t to_modified_prob(Probability_t p) {
        return (p > 0.0) ? static_cast<Mod_prob_t>(-1 * std::log2(p)) : zero_prob;
    }

    static bool is_not_zero_mod_prob(Mod_prob_t x) { return !almost_equal(x, zero_prob); }
};

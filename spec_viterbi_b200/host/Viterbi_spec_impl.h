// Abstract specialised-Viterbi backend interface (mirror of reference
// `Viterbi_impl/Viterbi_spec_impl.h:6-24`).  The HMM is the static argument:
// `spec_with` precomputes, for every tuple of `level` symbols, the (min,+) product
// of the per-symbol step matrices; `run_Viterbi_spec` then consumes `level`
// symbols per product.
#pragma once

#include "HMM.h"

class Viterbi_spec_impl {
  public:
    Viterbi_spec_impl() = default;
    explicit Viterbi_spec_impl(std::size_t level_) : level(level_) {}
    virtual ~Viterbi_spec_impl() = default;

    // Drops any previous tables and rebuilds them for `hmm`.
    virtual void spec_with(const HMM& hmm) = 0;

    [[nodiscard]] virtual HMM::Mod_prob_vec_t
    run_Viterbi_spec(const HMM::Emit_seq_t& seq) const = 0;

    [[nodiscard]] std::size_t get_level() const { return level; }

  protected:
    std::size_t level = 1; // number of observation handlers folded into one product
};

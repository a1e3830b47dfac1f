// Abstract plain-Viterbi backend interface (mirror of reference
// `Viterbi_impl/Viterbi_impl.h:6-11`): one call = one observation sequence,
// result = the final score vector (-log2 space), one entry per state.
#pragma once

#include "HMM.h"

class Viterbi_impl {
  public:
    virtual ~Viterbi_impl() = default;

    [[nodiscard]] virtual HMM::Mod_prob_vec_t run_Viterbi(const HMM& hmm,
                                                          const HMM::Emit_seq_t& seq) const = 0;
};

// Shared bits of the C++ harness clones (test + bench infrastructure, not product code).
// Protocol follows the reference's harness: benchmark/benchmark_helper.h:14,37-66 (10 timed
// runs, median of the sorted integer-millisecond samples), tests/test_helper.h:25-42
// (compare_two_answers: element order, HMM::almost_equal, then sizes).
#pragma once

#include "../oracle/sv_oracle.h"
#include "../spec_viterbi_b200/host/B200_impl.h"
#include "../spec_viterbi_b200/host/data_reader.h"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <functional>
#include <string>
#include <vector>

namespace harness {

constexpr int TIMES_TO_RUN = 10; // benchmark_helper.h:14

inline svo_model oracle_view(const HMM& h, std::vector<float>& emis_flat) {
    emis_flat.resize(h.emit_num * h.states_num);
    for (size_t e = 0; e < h.emit_num; ++e)
        std::memcpy(emis_flat.data() + e * h.states_num, h.emissions[e].data(), h.states_num * sizeof(float));
    svo_model m;
    m.states = h.states_num;
    m.emits = h.emit_num;
    m.trans = h.trans_num;
    m.nstart = h.non_zero_start_probs;
    m.trans_src = reinterpret_cast<const uint64_t*>(h.trans_rows.data());
    m.trans_dst = reinterpret_cast<const uint64_t*>(h.trans_cols.data());
    m.trans_w = h.trans_probs.data();
    m.emis = emis_flat.data();
    m.start_cols = reinterpret_cast<const uint64_t*>(h.start_probabilities_cols.data());
    m.start_w = h.start_probabilities.data();
    return m;
}

// CPU column of the harness: the oracle's restatement of GraphBLAS_impl::run_Viterbi.
inline HMM::Mod_prob_vec_t oracle_run(const HMM& h, const HMM::Emit_seq_t& seq) {
    std::vector<float> emis;
    const svo_model m = oracle_view(h, emis);
    HMM::Mod_prob_vec_t out(h.states_num);
    svo_plain(&m, reinterpret_cast<const uint64_t*>(seq.data()), seq.size(), out.data());
    return out;
}

// tests/test_helper.h:25-42
inline bool compare_two_answers(const HMM::Mod_prob_vec_t& lhs, const HMM::Mod_prob_vec_t& rhs) {
    for (size_t i = 0; i < lhs.size(); ++i) {
        if (i >= rhs.size() || !HMM::almost_equal(lhs[i], rhs[i])) return false;
    }
    return lhs.size() == rhs.size();
}

inline bool bit_equal(const HMM::Mod_prob_vec_t& lhs, const HMM::Mod_prob_vec_t& rhs) {
    if (lhs.size() != rhs.size()) return false;
    for (size_t i = 0; i < lhs.size(); ++i)
        if (!(lhs[i] == rhs[i])) return false; // +0 == -0, inf == inf; no NaN occurs
    return true;
}

// median of TIMES_TO_RUN integer-millisecond samples (benchmark_helper.h:45-66)
inline long median_ms(const std::function<void()>& fn) {
    std::vector<long> t;
    for (int i = 0; i < TIMES_TO_RUN; ++i) {
        const auto a = std::chrono::steady_clock::now();
        fn();
        const auto b = std::chrono::steady_clock::now();
        t.push_back(static_cast<long>(std::chrono::duration_cast<std::chrono::milliseconds>(b - a).count()));
    }
    std::sort(t.begin(), t.end());
    return (t[TIMES_TO_RUN / 2 - 1] + t[TIMES_TO_RUN / 2]) / 2;
}

inline double median_us(const std::function<void()>& fn) {
    std::vector<double> t;
    for (int i = 0; i < TIMES_TO_RUN; ++i) {
        const auto a = std::chrono::steady_clock::now();
        fn();
        const auto b = std::chrono::steady_clock::now();
        t.push_back(std::chrono::duration<double, std::micro>(b - a).count());
    }
    std::sort(t.begin(), t.end());
    return (t[TIMES_TO_RUN / 2 - 1] + t[TIMES_TO_RUN / 2]) / 2;
}

} // namespace harness

// extern "C" driver around the REFERENCE's own CPU backends (test infrastructure).
//
// This translation unit is compiled together with the reference sources taken where they
// lie under /root/reference (GraphBLAS_impl.cpp, GraphBLAS_spec_impl.cpp,
// GraphBLAS_helper.cpp, data_reader.cpp; nothing is copied into this repo) and with
// oracle/grb_shim, into oracle/_ref/libsv_ref.so (see oracle/Makefile).  It lets tests and
// bench.py --impl reference call the unmodified reference through plain pointers.
#include "GraphBLAS_helper.h"
#include "GraphBLAS_impl.h"
#include "GraphBLAS_spec_impl.h"
#include "data_reader.h"

#include "sv_oracle.h" // svo_model (plain C view of an HMM)

#include <atomic>
#include <cstring>
#include <thread>
#include <vector>

namespace {

HMM to_reference_hmm(const svo_model* m) {
    HMM h;
    h.states_num = m->states;
    h.emit_num = m->emits;
    h.trans_num = m->trans;
    h.trans_rows.assign(m->trans_src, m->trans_src + m->trans);
    h.trans_cols.assign(m->trans_dst, m->trans_dst + m->trans);
    h.trans_probs.assign(m->trans_w, m->trans_w + m->trans);
    h.emissions.resize(m->emits);
    for (uint64_t e = 0; e < m->emits; ++e) {
        h.emissions[e].assign(m->emis + e * m->states, m->emis + (e + 1) * m->states);
    }
    h.non_zero_start_probs = m->nstart;
    h.start_probabilities_cols.assign(m->start_cols, m->start_cols + m->nstart);
    h.start_probabilities.assign(m->start_w, m->start_w + m->nstart);
    return h;
}

void launch() { GraphBLAS_helper::get_instance().launch_GraphBLAS(); }

} // namespace

extern "C" {

// GraphBLAS_impl::run_Viterbi, one sequence.
int svref_plain(const svo_model* m, const uint64_t* seq, uint64_t T, float* out) {
    launch();
    const HMM h = to_reference_hmm(m);
    const HMM::Emit_seq_t s(seq, seq + T);
    const auto res = GraphBLAS_impl().run_Viterbi(h, s);
    if (res.size() != m->states) return -1;
    std::memcpy(out, res.data(), res.size() * sizeof(float));
    return 0;
}

// One run_Viterbi call per sequence (the reference harness protocol, bench_Viterbi.h:51-57),
// spread over `threads` host threads; out is [nseq][N].
int svref_plain_batch(const svo_model* m, const uint64_t* symbols, const uint64_t* offsets,
                      uint64_t nseq, float* out, int threads) {
    launch();
    const HMM h = to_reference_hmm(m);
    std::atomic<uint64_t> next{0};
    std::atomic<int> rc{0};
    auto work = [&]() {
        const GraphBLAS_impl impl;
        for (;;) {
            const uint64_t s = next.fetch_add(1);
            if (s >= nseq) break;
            const HMM::Emit_seq_t seq(symbols + offsets[s], symbols + offsets[s + 1]);
            const auto res = impl.run_Viterbi(h, seq);
            if (res.size() != m->states) {
                rc = -1;
                continue;
            }
            std::memcpy(out + s * m->states, res.data(), res.size() * sizeof(float));
        }
    };
    if (threads < 1) threads = 1;
    std::vector<std::thread> pool;
    for (int t = 1; t < threads; ++t) pool.emplace_back(work);
    work();
    for (auto& t : pool) t.join();
    return rc;
}

// GraphBLAS_spec_impl: spec_with (= constructor with HMM) and run_Viterbi_spec.
void* svref_spec_create(const svo_model* m, uint64_t level) {
    launch();
    const HMM h = to_reference_hmm(m);
    auto* impl = new GraphBLAS_spec_impl(level);
    impl->spec_with(h);
    return impl;
}

int svref_spec_respec(void* spec, const svo_model* m) {
    const HMM h = to_reference_hmm(m);
    static_cast<GraphBLAS_spec_impl*>(spec)->spec_with(h);
    return 0;
}

int svref_spec_run(void* spec, const uint64_t* seq, uint64_t T, uint64_t states, float* out) {
    const HMM::Emit_seq_t s(seq, seq + T);
    const auto res = static_cast<const GraphBLAS_spec_impl*>(spec)->run_Viterbi_spec(s);
    if (res.size() != states) return -1;
    std::memcpy(out, res.data(), res.size() * sizeof(float));
    return 0;
}

// run_Viterbi_spec over many sequences on `threads` host threads (the method is const and
// only reads the precomputed tables).
int svref_spec_run_batch(void* spec, const uint64_t* symbols, const uint64_t* offsets,
                         uint64_t nseq, uint64_t states, float* out, int threads) {
    const auto* impl = static_cast<const GraphBLAS_spec_impl*>(spec);
    std::atomic<uint64_t> next{0};
    std::atomic<int> rc{0};
    auto work = [&]() {
        for (;;) {
            const uint64_t s = next.fetch_add(1);
            if (s >= nseq) break;
            const HMM::Emit_seq_t seq(symbols + offsets[s], symbols + offsets[s + 1]);
            const auto res = impl->run_Viterbi_spec(seq);
            if (res.size() != states) {
                rc = -1;
                continue;
            }
            std::memcpy(out + s * states, res.data(), res.size() * sizeof(float));
        }
    };
    if (threads < 1) threads = 1;
    std::vector<std::thread> pool;
    for (int t = 1; t < threads; ++t) pool.emplace_back(work);
    work();
    for (auto& t : pool) t.join();
    return rc;
}

void svref_spec_destroy(void* spec) { delete static_cast<GraphBLAS_spec_impl*>(spec); }

// The reference's text loaders, for cross-checking this repo's data_reader.
void* svref_read_hmm(const char* path) { return new HMM(read_HMM(path)); }
void svref_hmm_dims(const void* h, uint64_t dims[4]) {
    const auto* m = static_cast<const HMM*>(h);
    dims[0] = m->states_num;
    dims[1] = m->emit_num;
    dims[2] = m->trans_num;
    dims[3] = m->non_zero_start_probs;
}
void svref_hmm_copy(const void* h, uint64_t* src, uint64_t* dst, float* w, float* emis,
                    uint64_t* start_cols, float* start_w) {
    const auto* m = static_cast<const HMM*>(h);
    std::memcpy(src, m->trans_rows.data(), m->trans_rows.size() * sizeof(uint64_t));
    std::memcpy(dst, m->trans_cols.data(), m->trans_cols.size() * sizeof(uint64_t));
    std::memcpy(w, m->trans_probs.data(), m->trans_probs.size() * sizeof(float));
    for (size_t e = 0; e < m->emissions.size(); ++e) {
        std::memcpy(emis + e * m->states_num, m->emissions[e].data(),
                    m->emissions[e].size() * sizeof(float));
    }
    std::memcpy(start_cols, m->start_probabilities_cols.data(),
                m->start_probabilities_cols.size() * sizeof(uint64_t));
    std::memcpy(start_w, m->start_probabilities.data(),
                m->start_probabilities.size() * sizeof(float));
}
void svref_hmm_free(void* h) { delete static_cast<HMM*>(h); }

} // extern "C"

// B200_impl / B200_spec_impl -- the drop-in third GPU backend for Spec_Viterbi.
//
// Header-only C++ adaptors from the reference's abstract interfaces
//   Viterbi_impl       (reference Viterbi_impl/Viterbi_impl.h:6-11)
//   Viterbi_spec_impl  (reference Viterbi_impl/Viterbi_spec_impl.h:6-24)
// to the extern "C" ABI of libspecviterbi_b200.so (include/spec_viterbi_b200.h).  They take the
// place of CUSP_impl / cuASR_impl and CUSP_spec_impl / cuASR_spec_impl in the harnesses
// (benchmark/bench_Viterbi.h:26-31, bench_Viterbi_spec.h:31-47, tests/test_semantic_equality.cpp
// :36-56): same constructor shapes, same method names, same return-by-value vectors.
//
// Differences that a caller can observe:
//   * the reference backends rebuild / re-upload the model on every run_Viterbi call
//     (GraphBLAS_impl.cpp:24-45, CUSP_impl.cu:10, cuASR_impl.cu:11); B200_impl keeps the device
//     model of the last HMM it saw (content fingerprint), so only the first call pays the upload and
//     the K0 schedule (built on first use; harness/bench_viterbi.cpp reports the rebuild-per-run figure
//     beside the cached one);
//   * errors: the reference prints in debug builds and carries on; these adaptors throw
//     std::runtime_error with sv_last_error() (there is no CPU fallback to carry on with);
//   * additions: run_Viterbi_batch / run_Viterbi_spec_batch take all sequences of a set in one
//     call (sequences as SpMM columns) -- the reason this backend exists.
#pragma once

#include "HMM.h"
#include "Viterbi_impl.h"
#include "Viterbi_spec_impl.h"
#include "spec_viterbi_b200.h"

#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

namespace b200_detail {

inline void check(int rc, const char* what) {
    if (rc != SV_OK) throw std::runtime_error(std::string(what) + ": " + sv_last_error());
}

// Content fingerprint: decides whether the cached device model is still valid.  The HMM may be mutated in
// place between calls, so the content is hashed on every call -- 8 bytes per multiply (about 50 us for the
// 0.4 MB of a 2405-state model; round 1 hashed byte by byte, ~0.4 ms, comparable to a whole short sequence).
inline std::uint64_t fingerprint(const HMM& h) {
    std::uint64_t x = 0x9E3779B97F4A7C15ull;
    auto mix = [&](const void* p, std::size_t n) {
        const auto* b = static_cast<const unsigned char*>(p);
        std::size_t i = 0;
        for (; i + 8 <= n; i += 8) {
            std::uint64_t w;
            std::memcpy(&w, b + i, 8);
            x = (x ^ w) * 0x100000001B3ull;
            x ^= x >> 32;
        }
        std::uint64_t tail = n;
        if (i < n) std::memcpy(&tail, b + i, n - i);
        x = (x ^ tail) * 0x100000001B3ull;
        x ^= x >> 29;
    };
    const std::size_t dims[4] = {h.states_num, h.emit_num, h.trans_num, h.non_zero_start_probs};
    mix(dims, sizeof(dims));
    mix(h.trans_rows.data(), h.trans_rows.size() * sizeof(HMM::Index_t));
    mix(h.trans_cols.data(), h.trans_cols.size() * sizeof(HMM::Index_t));
    mix(h.trans_probs.data(), h.trans_probs.size() * sizeof(float));
    for (const auto& row : h.emissions) mix(row.data(), row.size() * sizeof(float));
    mix(h.start_probabilities_cols.data(), h.start_probabilities_cols.size() * sizeof(HMM::Index_t));
    mix(h.start_probabilities.data(), h.start_probabilities.size() * sizeof(float));
    return x;
}

inline sv_model* make_model(const HMM& h, int device) {
    static_assert(sizeof(HMM::Index_t) == sizeof(std::uint64_t), "size_t must be 64 bit");
    std::vector<float> emis(h.emit_num * h.states_num);
    for (std::size_t e = 0; e < h.emit_num; ++e)
        std::memcpy(emis.data() + e * h.states_num, h.emissions[e].data(), h.states_num * sizeof(float));
    sv_model* m = nullptr;
    check(sv_model_create(h.states_num, h.emit_num, h.trans_num,
                          reinterpret_cast<const std::uint64_t*>(h.trans_rows.data()),
                          reinterpret_cast<const std::uint64_t*>(h.trans_cols.data()), h.trans_probs.data(),
                          emis.data(), h.non_zero_start_probs,
                          reinterpret_cast<const std::uint64_t*>(h.start_probabilities_cols.data()),
                          h.start_probabilities.data(), device, &m),
          "sv_model_create");
    return m;
}

inline sv_multi* make_multi(const HMM& h, const std::vector<int>& devices) {
    std::vector<float> emis(h.emit_num * h.states_num);
    for (std::size_t e = 0; e < h.emit_num; ++e)
        std::memcpy(emis.data() + e * h.states_num, h.emissions[e].data(), h.states_num * sizeof(float));
    sv_multi* m = nullptr;
    check(sv_multi_create(h.states_num, h.emit_num, h.trans_num,
                          reinterpret_cast<const std::uint64_t*>(h.trans_rows.data()),
                          reinterpret_cast<const std::uint64_t*>(h.trans_cols.data()), h.trans_probs.data(),
                          emis.data(), h.non_zero_start_probs,
                          reinterpret_cast<const std::uint64_t*>(h.start_probabilities_cols.data()),
                          h.start_probabilities.data(), devices.data(), static_cast<int>(devices.size()), &m),
          "sv_multi_create");
    return m;
}

inline void pack(const HMM::Emit_seq_vec_t& seqs, std::vector<std::uint64_t>& symbols,
                 std::vector<std::uint64_t>& offsets) {
    offsets.assign(seqs.size() + 1, 0);
    for (std::size_t s = 0; s < seqs.size(); ++s) offsets[s + 1] = offsets[s] + seqs[s].size();
    symbols.resize(offsets.back());
    for (std::size_t s = 0; s < seqs.size(); ++s)
        std::memcpy(symbols.data() + offsets[s], seqs[s].data(), seqs[s].size() * sizeof(std::uint64_t));
}

inline std::vector<HMM::Mod_prob_vec_t> unpack(const std::vector<float>& flat, std::size_t nseq, std::size_t n) {
    std::vector<HMM::Mod_prob_vec_t> res(nseq);
    for (std::size_t s = 0; s < nseq; ++s) res[s].assign(flat.begin() + s * n, flat.begin() + (s + 1) * n);
    return res;
}

} // namespace b200_detail

class B200_impl : public Viterbi_impl {
  public:
    explicit B200_impl(int device_ = 0) : devices{device_} {}
    // Several GPUs of one box: the HMM is replicated, run_Viterbi_batch deals the sequences to the devices
    // (sv_multi_*); run_Viterbi (one sequence) uses the first device.
    explicit B200_impl(std::vector<int> devices_) : devices(std::move(devices_)) {
        if (devices.empty()) throw std::runtime_error("B200_impl: empty device list");
    }
    B200_impl(const B200_impl&) = delete;
    B200_impl& operator=(const B200_impl&) = delete;
    ~B200_impl() override { drop(); }

    // One sequence -> final score vector (states_num entries, +inf = unreachable).
    [[nodiscard]] HMM::Mod_prob_vec_t run_Viterbi(const HMM& hmm, const HMM::Emit_seq_t& seq) const override {
        sv_model* m = model_for(hmm);
        HMM::Mod_prob_vec_t res(hmm.states_num);
        b200_detail::check(sv_viterbi(m, reinterpret_cast<const std::uint64_t*>(seq.data()), seq.size(), res.data()),
                           "sv_viterbi");
        return res;
    }

    // All sequences of a set in one launch per device; result[s] == run_Viterbi(hmm, seqs[s]) bit for bit.
    [[nodiscard]] std::vector<HMM::Mod_prob_vec_t> run_Viterbi_batch(const HMM& hmm,
                                                                    const HMM::Emit_seq_vec_t& seqs) const {
        sv_model* m = model_for(hmm);
        std::vector<std::uint64_t> symbols, offsets;
        b200_detail::pack(seqs, symbols, offsets);
        std::vector<float> flat(seqs.size() * hmm.states_num);
        if (multi != nullptr)
            b200_detail::check(sv_multi_viterbi_batch(multi, symbols.data(), offsets.data(), seqs.size(), flat.data(),
                                                      nullptr, nullptr),
                               "sv_multi_viterbi_batch");
        else
            b200_detail::check(sv_viterbi_batch(m, symbols.data(), offsets.data(), seqs.size(), flat.data(), nullptr, nullptr),
                               "sv_viterbi_batch");
        return b200_detail::unpack(flat, seqs.size(), hmm.states_num);
    }

  private:
    void drop() const {
        if (multi != nullptr) sv_multi_destroy(multi); // owns its replicas, `model` is one of them
        else sv_model_destroy(model);
        multi = nullptr;
        model = nullptr;
    }
    sv_model* model_for(const HMM& hmm) const {
        const std::uint64_t key = b200_detail::fingerprint(hmm);
        if (model == nullptr || key != model_key) {
            drop();
            if (devices.size() > 1) {
                multi = b200_detail::make_multi(hmm, devices);
                model = sv_multi_model(multi, 0);
            } else {
                model = b200_detail::make_model(hmm, devices[0]);
            }
            model_key = key;
        }
        return model;
    }
    std::vector<int> devices;
    mutable sv_model* model = nullptr;
    mutable sv_multi* multi = nullptr;
    mutable std::uint64_t model_key = 0;
};

class B200_spec_impl : public Viterbi_spec_impl {
  public:
    explicit B200_spec_impl(std::size_t level_, int device_ = 0, int method_ = SV_SPEC_AUTO)
        : Viterbi_spec_impl(level_), device(device_), method(method_) {}
    B200_spec_impl(const HMM& hmm, std::size_t level_, int device_ = 0, int method_ = SV_SPEC_AUTO)
        : Viterbi_spec_impl(level_), device(device_), method(method_) {
        spec_with(hmm);
    }
    B200_spec_impl(const B200_spec_impl&) = delete;
    B200_spec_impl& operator=(const B200_spec_impl&) = delete;
    ~B200_spec_impl() override { drop(); }

    // Drops the previous tables, then rebuilds them (GraphBLAS_spec_impl.cpp:45-48).
    void spec_with(const HMM& hmm) override {
        drop();
        model = b200_detail::make_model(hmm, device);
        states = hmm.states_num;
        b200_detail::check(sv_spec_create(model, level, method, 0, 0, nullptr, nullptr, &spec), "sv_spec_create");
    }

    [[nodiscard]] HMM::Mod_prob_vec_t run_Viterbi_spec(const HMM::Emit_seq_t& seq) const override {
        if (spec == nullptr) throw std::runtime_error("B200_spec_impl: spec_with(hmm) must be called first");
        HMM::Mod_prob_vec_t res(states);
        b200_detail::check(sv_spec_run(spec, reinterpret_cast<const std::uint64_t*>(seq.data()), seq.size(), res.data()),
                           "sv_spec_run");
        return res;
    }

    // Not in the reference: how a level-2 chunk is applied (sv_spec_set_run_mode) -- SV_SPEC_RUN_TABLE reads the
    // precomputed handler (the reference's GrB_mxm per chunk), SV_SPEC_RUN_IMPLICIT recomputes the entries from
    // the sparse factors (bit-identical, no table traffic), SV_SPEC_RUN_CHAIN runs level 2 of a chain + specials
    // model (every profile HMM the reference ships) in one persistent launch at O(N) per chunk (bit-identical);
    // SV_SPEC_RUN_AUTO (the default) picks the fastest eligible one.  Throws if the model / level is not eligible.
    void set_run_mode(int mode) {
        if (spec == nullptr) throw std::runtime_error("B200_spec_impl: spec_with(hmm) must be called first");
        b200_detail::check(sv_spec_set_run_mode(spec, mode), "sv_spec_set_run_mode");
    }

    // Not in the reference (it recomputes in every spec_with): keep the handler tables of level >= 2
    // across processes.  spec_with_saved attaches a table written by save_tables; it throws if the file
    // was built from a different model.
    void save_tables(const std::string& path) const {
        if (spec == nullptr) throw std::runtime_error("B200_spec_impl: spec_with(hmm) must be called first");
        b200_detail::check(sv_spec_save(spec, path.c_str()), "sv_spec_save");
    }
    void spec_with_saved(const HMM& hmm, const std::string& path) {
        drop();
        model = b200_detail::make_model(hmm, device);
        states = hmm.states_num;
        b200_detail::check(sv_spec_load(model, path.c_str(), nullptr, &spec), "sv_spec_load");
        if (sv_spec_level(spec) != level) {
            drop();
            throw std::runtime_error("B200_spec_impl: saved tables are of another level");
        }
    }

    [[nodiscard]] std::vector<HMM::Mod_prob_vec_t> run_Viterbi_spec_batch(const HMM::Emit_seq_vec_t& seqs) const {
        if (spec == nullptr) throw std::runtime_error("B200_spec_impl: spec_with(hmm) must be called first");
        std::vector<std::uint64_t> symbols, offsets;
        b200_detail::pack(seqs, symbols, offsets);
        std::vector<float> flat(seqs.size() * states);
        b200_detail::check(sv_spec_run_batch(spec, symbols.data(), offsets.data(), seqs.size(), flat.data(), nullptr, nullptr),
                           "sv_spec_run_batch");
        return b200_detail::unpack(flat, seqs.size(), states);
    }

  private:
    void drop() {
        sv_spec_destroy(spec);
        spec = nullptr;
        sv_model_destroy(model);
        model = nullptr;
    }
    int device, method;
    sv_model* model = nullptr;
    sv_spec* spec = nullptr;
    std::size_t states = 0;
};

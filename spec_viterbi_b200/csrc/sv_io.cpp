// C ABI over the host-side loaders (spec_viterbi_b200/host/data_reader.*): the file formats
// on either side of the hot path (.chmm / .ess text as in the reference's
// Viterbi_impl/data_reader.cpp, plus this repo's packed binary caches).
#include "sv_internal.h"

#include "../host/data_reader.h"

#include <cstring>
#include <new>
#include <string>

struct sv_hmm_file {
    HMM hmm;
};
struct sv_ess_file {
    HMM::Emit_seq_vec_t seqs;
    uint64_t total = 0;
};

extern "C" int sv_io_hmm_open(const char* path, sv_hmm_file** out, uint64_t dims[4]) {
    SV_API_BEGIN
    if (path == nullptr || out == nullptr || dims == nullptr)
        return sv_fail(SV_ERR_ARG, "sv_io_hmm_open: NULL");
    *out = nullptr;
    auto* f = new (std::nothrow) sv_hmm_file;
    if (f == nullptr) return sv_fail(SV_ERR_MEMORY, "sv_io_hmm_open: out of memory");
    f->hmm = read_HMM_any(path);
    if (f->hmm.states_num == 0) {
        delete f;
        return sv_fail(SV_ERR_IO, std::string("cannot read HMM file: ") + path);
    }
    dims[0] = f->hmm.states_num;
    dims[1] = f->hmm.emit_num;
    dims[2] = f->hmm.trans_num;
    dims[3] = f->hmm.non_zero_start_probs;
    *out = f;
    return SV_OK;
    SV_API_END
}

extern "C" int sv_io_hmm_copy(const sv_hmm_file* f, uint64_t* trans_src, uint64_t* trans_dst,
                              float* trans_w, float* emissions, uint64_t* start_cols,
                              float* start_w) {
    SV_API_BEGIN
    if (f == nullptr) return sv_fail(SV_ERR_ARG, "sv_io_hmm_copy: NULL");
    const HMM& h = f->hmm;
    for (size_t i = 0; i < h.trans_num; ++i) {
        if (trans_src) trans_src[i] = h.trans_rows[i];
        if (trans_dst) trans_dst[i] = h.trans_cols[i];
        if (trans_w) trans_w[i] = h.trans_probs[i];
    }
    if (emissions) {
        for (size_t e = 0; e < h.emit_num; ++e)
            std::memcpy(emissions + e * h.states_num, h.emissions[e].data(),
                        h.states_num * sizeof(float));
    }
    for (size_t i = 0; i < h.non_zero_start_probs; ++i) {
        if (start_cols) start_cols[i] = h.start_probabilities_cols[i];
        if (start_w) start_w[i] = h.start_probabilities[i];
    }
    return SV_OK;
    SV_API_END
}

extern "C" int sv_io_hmm_write_packed(const sv_hmm_file* f, const char* path) {
    SV_API_BEGIN
    if (f == nullptr || path == nullptr) return sv_fail(SV_ERR_ARG, "sv_io_hmm_write_packed: NULL");
    return write_HMM_packed(f->hmm, path) ? SV_OK : sv_fail(SV_ERR_IO, "cannot write packed HMM");
    SV_API_END
}

extern "C" void sv_io_hmm_close(sv_hmm_file* f) { delete f; }

extern "C" int sv_io_ess_open(const char* path, sv_ess_file** out, uint64_t* nseq,
                              uint64_t* total_symbols) {
    SV_API_BEGIN
    if (path == nullptr || out == nullptr || nseq == nullptr || total_symbols == nullptr)
        return sv_fail(SV_ERR_ARG, "sv_io_ess_open: NULL");
    *out = nullptr;
    auto* f = new (std::nothrow) sv_ess_file;
    if (f == nullptr) return sv_fail(SV_ERR_MEMORY, "sv_io_ess_open: out of memory");
    f->seqs = read_emit_seq_any(path);
    if (f->seqs.empty()) {
        delete f;
        return sv_fail(SV_ERR_IO, std::string("cannot read sequence file: ") + path);
    }
    for (const auto& s : f->seqs) f->total += s.size();
    *nseq = f->seqs.size();
    *total_symbols = f->total;
    *out = f;
    return SV_OK;
    SV_API_END
}

extern "C" int sv_io_ess_copy(const sv_ess_file* f, uint64_t* offsets, uint64_t* symbols) {
    SV_API_BEGIN
    if (f == nullptr || offsets == nullptr) return sv_fail(SV_ERR_ARG, "sv_io_ess_copy: NULL");
    uint64_t at = 0;
    for (size_t s = 0; s < f->seqs.size(); ++s) {
        offsets[s] = at;
        if (symbols) {
            for (const auto v : f->seqs[s]) symbols[at++] = v;
        } else {
            at += f->seqs[s].size();
        }
    }
    offsets[f->seqs.size()] = at;
    return SV_OK;
    SV_API_END
}

extern "C" int sv_io_ess_write_packed(const sv_ess_file* f, const char* path) {
    SV_API_BEGIN
    if (f == nullptr || path == nullptr) return sv_fail(SV_ERR_ARG, "sv_io_ess_write_packed: NULL");
    return write_emit_seq_packed(f->seqs, path) ? SV_OK
                                                : sv_fail(SV_ERR_IO, "cannot write packed sequences");
    SV_API_END
}

extern "C" void sv_io_ess_close(sv_ess_file* f) { delete f; }

// C-level multi-GPU entry points: the HMM replicated on several devices of one box, the sequences of a
// batch dealt to them (sequences are independent, no exchange inside the time loop, so there
// is no data-path collective -- every device delivers its share of the results straight into the
// caller's host arrays).  One host thread per device drives that device's replica through the ordinary
// single-device entry points, so everything the single-GPU path does (validation, kernel choice, ragged
// scheduling) is shared.
//
// Reference interface replaced: the harness loop "one run_Viterbi call per sequence"
// (benchmark/bench_Viterbi.h:51-57), batched and spread over the box's GPUs.
#include "sv_internal.h"

#include <algorithm>
#include <new>
#include <string>
#include <thread>

struct sv_multi {
    std::vector<sv_model*> models;
    std::vector<int> devices;
};

namespace {

template <typename F> int guarded(F&& f) {
    try {
        return f();
    } catch (const std::bad_alloc&) {
        return sv_fail(SV_ERR_MEMORY, "out of host memory");
    } catch (...) {
        return sv_fail(SV_ERR_MEMORY, "unexpected C++ exception");
    }
}

// Contiguous ranges of sequences with about equal symbol counts: device k gets [cut[k], cut[k+1]).
// Contiguity means the caller's symbol array is used in place (no host repacking); raggedness inside a
// range is the persistent kernels' business (longest-first work queue).
std::vector<uint64_t> split_by_symbols(const uint64_t* offsets, uint64_t nseq, int parts) {
    std::vector<uint64_t> cut(parts + 1, nseq);
    cut[0] = 0;
    const uint64_t first = offsets[0], total = offsets[nseq] - first;
    for (int k = 1; k < parts; ++k) {
        const uint64_t target = first + static_cast<uint64_t>(static_cast<unsigned __int128>(total) * k / parts);
        // the sequence boundary nearest to the target, never before the previous cut
        uint64_t q = static_cast<uint64_t>(std::lower_bound(offsets, offsets + nseq + 1, target) - offsets);
        if (q > 0 && target - offsets[q - 1] < offsets[q] - target) --q;
        cut[k] = std::min(nseq, std::max(q, cut[k - 1]));
    }
    return cut;
}

template <typename SrcT>
int multi_batch(sv_multi* mm, const SrcT* symbols, const uint64_t* offsets, uint64_t nseq, float* out, float* best,
                int32_t* best_state) {
    if (mm == nullptr || mm->models.empty()) return sv_fail(SV_ERR_ARG, "sv_multi_viterbi_batch: NULL handle");
    if (offsets == nullptr || nseq == 0) return sv_fail(SV_ERR_ARG, "sv_multi_viterbi_batch: empty batch");
    const int G = static_cast<int>(std::min<uint64_t>(mm->models.size(), nseq));
    const std::vector<uint64_t> cut = split_by_symbols(offsets, nseq, G);
    const uint64_t n = mm->models[0]->n;
    std::vector<int> rcs(G, SV_OK);
    std::vector<std::string> msgs(G);
    auto work = [&](int k) {
        const uint64_t q0 = cut[k], q1 = cut[k + 1];
        if (q1 <= q0) return;
        sv_model* m = mm->models[k];
        int rc;
        // offsets are passed as they are: the single-device entry points subtract offsets[q0] themselves
        if (sizeof(SrcT) == 1)
            rc = sv_viterbi_batch_u8(m, reinterpret_cast<const uint8_t*>(symbols), offsets + q0, q1 - q0,
                                     out ? out + q0 * n : nullptr, best ? best + q0 : nullptr,
                                     best_state ? best_state + q0 : nullptr);
        else
            rc = sv_viterbi_batch(m, reinterpret_cast<const uint64_t*>(symbols), offsets + q0, q1 - q0,
                                  out ? out + q0 * n : nullptr, best ? best + q0 : nullptr,
                                  best_state ? best_state + q0 : nullptr);
        rcs[k] = rc;
        if (rc != SV_OK) msgs[k] = sv_last_error(); // thread-local: carry it to the calling thread
    };
    std::vector<std::thread> pool;
    for (int k = 1; k < G; ++k) pool.emplace_back(work, k);
    work(0);
    for (auto& th : pool) th.join();
    for (int k = 0; k < G; ++k)
        if (rcs[k] != SV_OK) return sv_fail(rcs[k], "device " + std::to_string(mm->devices[k]) + ": " + msgs[k]);
    return SV_OK;
}

} // namespace

extern "C" int sv_multi_create(uint64_t states, uint64_t emits, uint64_t trans, const uint64_t* trans_src,
                               const uint64_t* trans_dst, const float* trans_w, const float* emissions,
                               uint64_t nstart, const uint64_t* start_cols, const float* start_w,
                               const int* devices, int ndev, sv_multi** out) {
    return guarded([&]() -> int {
        if (out == nullptr) return sv_fail(SV_ERR_ARG, "sv_multi_create: out is NULL");
        *out = nullptr;
        const int visible = sv_device_count();
        if (visible == 0) return sv_fail(SV_ERR_CUDA, "no CUDA device: this backend has no CPU fallback");
        if (ndev < 0 || (ndev > 0 && devices == nullptr)) return sv_fail(SV_ERR_ARG, "sv_multi_create: bad device list");
        sv_multi* mm = new sv_multi;
        if (ndev == 0) { // all visible devices
            for (int d = 0; d < visible; ++d) mm->devices.push_back(d);
        } else {
            mm->devices.assign(devices, devices + ndev);
        }
        for (const int d : mm->devices) {
            sv_model* m = nullptr;
            const int rc = sv_model_create(states, emits, trans, trans_src, trans_dst, trans_w, emissions, nstart,
                                           start_cols, start_w, d, &m);
            if (rc != SV_OK) {
                const std::string msg = sv_last_error();
                sv_multi_destroy(mm);
                return sv_fail(rc, msg);
            }
            mm->models.push_back(m);
        }
        *out = mm;
        return SV_OK;
    });
}

extern "C" void sv_multi_destroy(sv_multi* mm) {
    if (mm == nullptr) return;
    for (sv_model* m : mm->models) sv_model_destroy(m);
    delete mm;
}

extern "C" int sv_multi_device_count(const sv_multi* mm) { return mm ? static_cast<int>(mm->models.size()) : 0; }

extern "C" sv_model* sv_multi_model(sv_multi* mm, int k) {
    return (mm != nullptr && k >= 0 && k < static_cast<int>(mm->models.size())) ? mm->models[k] : nullptr;
}

extern "C" int sv_multi_prepare(sv_multi* mm) {
    return guarded([&]() -> int {
        if (mm == nullptr) return sv_fail(SV_ERR_ARG, "sv_multi_prepare: NULL handle");
        std::vector<int> rcs(mm->models.size(), SV_OK);
        std::vector<std::thread> pool;
        for (size_t k = 0; k < mm->models.size(); ++k)
            pool.emplace_back([&, k]() { rcs[k] = sv_model_prepare(mm->models[k]); });
        for (auto& th : pool) th.join();
        for (const int rc : rcs)
            if (rc != SV_OK) return sv_fail(rc, "sv_multi_prepare failed on a device");
        return SV_OK;
    });
}

extern "C" int sv_multi_viterbi_batch(sv_multi* mm, const uint64_t* symbols, const uint64_t* offsets, uint64_t nseq,
                                      float* out, float* best, int32_t* best_state) {
    return guarded([&]() { return multi_batch(mm, symbols, offsets, nseq, out, best, best_state); });
}

extern "C" int sv_multi_viterbi_batch_u8(sv_multi* mm, const uint8_t* symbols, const uint64_t* offsets, uint64_t nseq,
                                         float* out, float* best, int32_t* best_state) {
    return guarded([&]() { return multi_batch(mm, symbols, offsets, nseq, out, best, best_state); });
}

// Host-only: the split sv_multi_viterbi_batch* would use (cut[parts + 1]); lets tests check the partition
// without devices.
extern "C" int sv_multi_split(const uint64_t* offsets, uint64_t nseq, int parts, uint64_t* cut) {
    return guarded([&]() -> int {
        if (offsets == nullptr || cut == nullptr || parts < 1 || nseq == 0) return sv_fail(SV_ERR_ARG, "sv_multi_split: bad argument");
        const std::vector<uint64_t> c = split_by_symbols(offsets, nseq, parts);
        std::copy(c.begin(), c.end(), cut);
        return SV_OK;
    });
}

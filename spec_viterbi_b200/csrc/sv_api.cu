// C ABI: errors, model handle (K0 repack + upload), sequence batches, plain-path entry points.
#include "sv_internal.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <new>
#include <numeric>

// ------------------------------------------------------------------ errors ----
namespace {
thread_local std::string g_last_error;
}

void sv_set_error(const std::string& msg) { g_last_error = msg; }
int sv_fail(int code, const std::string& msg) {
    g_last_error = msg;
    return code;
}

extern "C" const char* sv_last_error(void) { return g_last_error.c_str(); }
extern "C" const char* sv_version(void) { return "spec_viterbi_b200 0.1 (sm_100a)"; }
extern "C" int sv_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

// ------------------------------------------------------------------- model ----
namespace {

// CSR of T' (row = destination) with FIRST-wins duplicate removal, matching
// GrB_Matrix_build(..., GrB_FIRST_FP32) in the reference (GraphBLAS_impl.cpp:42-45).
int build_csr(uint64_t n, uint64_t trans, const uint64_t* src, const uint64_t* dst, const float* w,
              SvHostCsr* out) {
    std::vector<uint64_t> order;
    order.reserve(trans);
    for (uint64_t e = 0; e < trans; ++e) {
        if (src[e] >= n || dst[e] >= n) return sv_fail(SV_ERR_ARG, "transition index >= states");
        order.push_back(e);
    }
    std::stable_sort(order.begin(), order.end(), [&](uint64_t a, uint64_t b) {
        return dst[a] != dst[b] ? dst[a] < dst[b] : src[a] < src[b];
    });
    out->n = n;
    out->rowptr.assign(n + 1, 0);
    out->col.clear();
    out->w.clear();
    for (size_t k = 0; k < order.size(); ++k) {
        const uint64_t e = order[k];
        if (k > 0 && dst[e] == dst[order[k - 1]] && src[e] == src[order[k - 1]]) continue;
        out->col.push_back(static_cast<uint32_t>(src[e]));
        out->w.push_back(w[e]);
        out->rowptr[dst[e] + 1] += 1;
    }
    for (uint64_t i = 0; i < n; ++i) out->rowptr[i + 1] += out->rowptr[i];
    return SV_OK;
}

template <typename T> size_t blob_reserve(size_t* cursor, size_t count) {
    const size_t at = (*cursor + 255) & ~size_t(255);
    *cursor = at + count * sizeof(T);
    return at;
}

int upload_generic(sv_model* m) {
    const uint64_t n = m->n;
    // light rows sorted by in-degree (descending) so neighbouring threads loop alike;
    // rows above the threshold are block-reduced
    const uint32_t heavy_threshold = 256;
    std::vector<uint32_t> light, heavy;
    for (uint32_t i = 0; i < n; ++i) {
        const uint32_t d = m->csr.rowptr[i + 1] - m->csr.rowptr[i];
        (d > heavy_threshold ? heavy : light).push_back(i);
    }
    std::stable_sort(light.begin(), light.end(), [&](uint32_t a, uint32_t b) {
        return (m->csr.rowptr[a + 1] - m->csr.rowptr[a]) > (m->csr.rowptr[b + 1] - m->csr.rowptr[b]);
    });
    size_t bytes = 0;
    const size_t o_rowptr = blob_reserve<uint32_t>(&bytes, n + 1);
    const size_t o_col = blob_reserve<uint32_t>(&bytes, m->csr.col.size());
    const size_t o_w = blob_reserve<float>(&bytes, m->csr.w.size());
    const size_t o_emis = blob_reserve<float>(&bytes, m->emis.size());
    const size_t o_start = blob_reserve<float>(&bytes, n);
    const size_t o_light = blob_reserve<uint32_t>(&bytes, light.size());
    const size_t o_heavy = blob_reserve<uint32_t>(&bytes, heavy.size());
    bytes += 256;
    std::vector<unsigned char> host(bytes, 0);
    auto put = [&](size_t at, const void* p, size_t nbytes) {
        if (nbytes) std::memcpy(host.data() + at, p, nbytes);
    };
    put(o_rowptr, m->csr.rowptr.data(), (n + 1) * 4);
    put(o_col, m->csr.col.data(), m->csr.col.size() * 4);
    put(o_w, m->csr.w.data(), m->csr.w.size() * 4);
    put(o_emis, m->emis.data(), m->emis.size() * 4);
    put(o_start, m->start.data(), n * 4);
    put(o_light, light.data(), light.size() * 4);
    put(o_heavy, heavy.data(), heavy.size() * 4);
    SV_CUDA(cudaMalloc(&m->d_blob, bytes));
    SV_CUDA(cudaMemcpy(m->d_blob, host.data(), bytes, cudaMemcpyHostToDevice));
    auto* base = static_cast<unsigned char*>(m->d_blob);
    m->gen.n = static_cast<int>(n);
    m->gen.emits = static_cast<int>(m->emits);
    m->gen.n_light = static_cast<int>(light.size());
    m->gen.n_heavy = static_cast<int>(heavy.size());
    m->gen.rowptr = reinterpret_cast<const uint32_t*>(base + o_rowptr);
    m->gen.col = reinterpret_cast<const uint32_t*>(base + o_col);
    m->gen.w = reinterpret_cast<const float*>(base + o_w);
    m->gen.emis = reinterpret_cast<const float*>(base + o_emis);
    m->gen.start = reinterpret_cast<const float*>(base + o_start);
    m->gen.light_rows = reinterpret_cast<const uint32_t*>(base + o_light);
    m->gen.heavy_rows = reinterpret_cast<const uint32_t*>(base + o_heavy);
    return SV_OK;
}

int upload_es(sv_model* m) {
    const SvEsPlan& p = m->es;
    size_t bytes = 0;
    const size_t o_w = blob_reserve<float>(&bytes, p.edge_w.size());
    const size_t o_off = blob_reserve<uint32_t>(&bytes, p.edge_off.size());
    const size_t o_deg = blob_reserve<uint8_t>(&bytes, p.deg.size());
    const size_t o_emis = blob_reserve<float>(&bytes, p.emis_p.size());
    const size_t o_start = blob_reserve<float>(&bytes, p.start_p.size());
    const size_t o_cptr = blob_reserve<uint32_t>(&bytes, p.copy_ptr.size());
    const size_t o_cp = blob_reserve<uint32_t>(&bytes, p.copy_p.size());
    bytes += 256;
    std::vector<unsigned char> host(bytes, 0);
    std::memcpy(host.data() + o_w, p.edge_w.data(), p.edge_w.size() * 4);
    std::memcpy(host.data() + o_off, p.edge_off.data(), p.edge_off.size() * 4);
    std::memcpy(host.data() + o_deg, p.deg.data(), p.deg.size());
    std::memcpy(host.data() + o_emis, p.emis_p.data(), p.emis_p.size() * 4);
    std::memcpy(host.data() + o_start, p.start_p.data(), p.start_p.size() * 4);
    std::memcpy(host.data() + o_cptr, p.copy_ptr.data(), p.copy_ptr.size() * 4);
    std::memcpy(host.data() + o_cp, p.copy_p.data(), p.copy_p.size() * 4);
    SV_CUDA(cudaMalloc(&m->d_es_blob, bytes));
    SV_CUDA(cudaMemcpy(m->d_es_blob, host.data(), bytes, cudaMemcpyHostToDevice));
    auto* base = static_cast<unsigned char*>(m->d_es_blob);
    SvEsDev& d = m->esdev;
    d.n = static_cast<int>(m->n);
    d.n_int = p.n_int;
    d.n_light = p.n_light;
    d.n_heavy = p.n_heavy;
    d.emits = static_cast<int>(m->emits);
    d.nslots = p.nslots;
    d.edge_w = reinterpret_cast<const float*>(base + o_w);
    d.edge_off = reinterpret_cast<const uint32_t*>(base + o_off);
    d.deg = reinterpret_cast<const uint8_t*>(base + o_deg);
    for (int h = 0; h < SV_ES_NHMAX; ++h) d.heavy_slots[h] = p.heavy_slots[h];
    d.emis_p = reinterpret_cast<const float*>(base + o_emis);
    d.start_p = reinterpret_cast<const float*>(base + o_start);
    d.copy_ptr = reinterpret_cast<const uint32_t*>(base + o_cptr);
    d.copy_p = reinterpret_cast<const uint32_t*>(base + o_cp);
    return SV_OK;
}

int chosen_kernel(const sv_model* m) {
    if (m->kernel_pref == SV_KERNEL_GENERIC) return SV_KERNEL_GENERIC;
    if (m->es.valid) return SV_KERNEL_ES;
    return SV_KERNEL_GENERIC;
}

} // namespace

extern "C" int sv_model_create(uint64_t states, uint64_t emits, uint64_t trans,
                               const uint64_t* trans_src, const uint64_t* trans_dst,
                               const float* trans_w, const float* emissions, uint64_t nstart,
                               const uint64_t* start_cols, const float* start_w, int device,
                               sv_model** out) {
    if (out == nullptr) return sv_fail(SV_ERR_ARG, "sv_model_create: out is NULL");
    *out = nullptr;
    if (states == 0 || emits == 0) return sv_fail(SV_ERR_ARG, "sv_model_create: empty model");
    if (states > (1u << 30) || emits > 65536)
        return sv_fail(SV_ERR_UNSUPPORTED, "sv_model_create: model too large");
    if ((trans > 0 && (!trans_src || !trans_dst || !trans_w)) || !emissions ||
        (nstart > 0 && (!start_cols || !start_w)))
        return sv_fail(SV_ERR_ARG, "sv_model_create: NULL array");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return sv_fail(SV_ERR_CUDA, "no CUDA device: this backend has no CPU fallback");
    }
    if (device < 0 || device >= ndev) return sv_fail(SV_ERR_ARG, "sv_model_create: bad device");

    sv_model* m = new (std::nothrow) sv_model;
    if (m == nullptr) return sv_fail(SV_ERR_MEMORY, "sv_model_create: out of host memory");
    m->device = device;
    m->n = states;
    m->emits = emits;
    int rc = build_csr(states, trans, trans_src, trans_dst, trans_w, &m->csr);
    if (rc != SV_OK) {
        delete m;
        return rc;
    }
    m->nnz = m->csr.col.size();
    for (uint64_t i = 0; i < states; ++i)
        m->max_deg = std::max<uint64_t>(m->max_deg, m->csr.rowptr[i + 1] - m->csr.rowptr[i]);
    m->emis.assign(emissions, emissions + emits * states);
    // dense start column: +inf where absent, FIRST wins (GraphBLAS_impl.cpp:18-21)
    m->start.assign(states, std::numeric_limits<float>::infinity());
    {
        std::vector<unsigned char> seen(states, 0);
        for (uint64_t k = 0; k < nstart; ++k) {
            if (start_cols[k] >= states) {
                delete m;
                return sv_fail(SV_ERR_ARG, "start state index >= states");
            }
            if (!seen[start_cols[k]]) {
                seen[start_cols[k]] = 1;
                m->start[start_cols[k]] = start_w[k];
            }
        }
    }

    auto fail = [&](int code) {
        sv_model_destroy(m);
        return code;
    };
    if (cudaSetDevice(device) != cudaSuccess)
        return fail(sv_fail(SV_ERR_CUDA, "cudaSetDevice failed"));
    int v = 0;
    cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, device);
    m->sm_count = v;
    cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, device);
    m->smem_optin = static_cast<size_t>(v);
    if (sv_generic_smem(SvGenericDev{static_cast<int>(states)}) + 4096 > m->smem_optin)
        return fail(sv_fail(SV_ERR_UNSUPPORTED, "model too large for the shared-memory kernels"));
    if ((rc = upload_generic(m)) != SV_OK) return fail(rc);
    if (cudaMalloc(&m->d_counter, 64) != cudaSuccess)
        return fail(sv_fail(SV_ERR_CUDA, "cudaMalloc(counter) failed"));
    if (cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking) != cudaSuccess)
        return fail(sv_fail(SV_ERR_CUDA, "cudaStreamCreate failed"));

    sv_es_build_plan(m->csr, m->emis, m->start, emits, m->smem_optin, &m->es);
    if (m->es.valid) {
        if ((rc = upload_es(m)) != SV_OK) return fail(rc);
        if (sv_es_prepare(m->es.variant_index, m->es.smem_bytes) != 0) {
            cudaGetLastError();
            m->es.valid = false; // kernel attribute rejected: keep the generic kernel
        }
    }
    *out = m;
    return SV_OK;
}

extern "C" void sv_model_destroy(sv_model* m) {
    if (m == nullptr) return;
    cudaSetDevice(m->device);
    if (m->d_blob) cudaFree(m->d_blob);
    if (m->d_es_blob) cudaFree(m->d_es_blob);
    if (m->d_counter) cudaFree(m->d_counter);
    if (m->stream) cudaStreamDestroy(m->stream);
    delete m;
}

extern "C" int sv_model_set_kernel(sv_model* m, int kernel) {
    if (m == nullptr) return sv_fail(SV_ERR_ARG, "sv_model_set_kernel: NULL model");
    if (kernel == SV_KERNEL_ES && !m->es.valid)
        return sv_fail(SV_ERR_UNSUPPORTED, "model does not fit the edge-stationary kernel");
    if (kernel < SV_KERNEL_AUTO || kernel > SV_KERNEL_ES)
        return sv_fail(SV_ERR_ARG, "sv_model_set_kernel: unknown kernel");
    m->kernel_pref = kernel;
    return SV_OK;
}

extern "C" int sv_model_get_info(const sv_model* m, sv_model_info* info) {
    if (m == nullptr || info == nullptr) return sv_fail(SV_ERR_ARG, "sv_model_get_info: NULL");
    std::memset(info, 0, sizeof(*info));
    info->states = m->n;
    info->emits = m->emits;
    info->nnz = m->nnz;
    info->max_in_degree = m->max_deg;
    info->device = m->device;
    info->kernel = chosen_kernel(m);
    if (m->es.valid) {
        info->es_threads = m->es.var.threads;
        info->es_seqs_per_cta = m->es.var.q;
        info->es_ranks = m->es.ranks;
        info->es_heavy_rows = m->es.n_heavy;
        info->es_slots_per_thread = m->es.nslots;
        info->es_gather_wavefronts = m->es.gather_wavefronts;
        info->es_gather_ideal = m->es.gather_ideal;
        info->es_smem_bytes = m->es.smem_bytes;
        info->es_split_states = m->es.split_states;
    }
    return SV_OK;
}

extern "C" int sv_model_last_launches(const sv_model* m) { return m ? m->last_launches : 0; }

// ----------------------------------------------------------------- sequences ----
namespace {

template <typename SrcT>
int seqs_create_impl(const sv_model* m, const SrcT* symbols, const uint64_t* offsets,
                     uint64_t nseq, sv_seqs** out) {
    if (out == nullptr) return sv_fail(SV_ERR_ARG, "sv_seqs_create: out is NULL");
    *out = nullptr;
    if (m == nullptr || offsets == nullptr) return sv_fail(SV_ERR_ARG, "sv_seqs_create: NULL");
    if (nseq == 0 || nseq > 0x7fffffffu) return sv_fail(SV_ERR_ARG, "sv_seqs_create: bad nseq");
    const uint64_t total = offsets[nseq];
    if (total > 0 && symbols == nullptr) return sv_fail(SV_ERR_ARG, "sv_seqs_create: NULL symbols");
    uint64_t max_len = 0;
    for (uint64_t s = 0; s < nseq; ++s) {
        if (offsets[s + 1] <= offsets[s])
            return sv_fail(SV_ERR_ARG, "sv_seqs_create: empty sequence (the reference requires "
                                       "at least one symbol, GraphBLAS_impl.cpp:59)");
        max_len = std::max(max_len, offsets[s + 1] - offsets[s]);
    }
    const int sym_bytes = m->emits <= 256 ? 1 : 2;
    std::vector<uint32_t> order(nseq);
    std::iota(order.begin(), order.end(), 0u);
    std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) {
        return (offsets[a + 1] - offsets[a]) > (offsets[b + 1] - offsets[b]);
    });
    size_t bytes = 0;
    const size_t o_sym = blob_reserve<uint8_t>(&bytes, total * sym_bytes + 16);
    const size_t o_off = blob_reserve<uint64_t>(&bytes, nseq + 1);
    const size_t o_ord = blob_reserve<uint32_t>(&bytes, nseq);
    bytes += 256;
    unsigned char* host = nullptr;
    if (cudaMallocHost(&host, bytes) != cudaSuccess) {
        cudaGetLastError();
        return sv_fail(SV_ERR_MEMORY, "sv_seqs_create: pinned staging allocation failed");
    }
    bool bad = false;
    if (sym_bytes == 1) {
        uint8_t* dst = host + o_sym;
        for (uint64_t k = 0; k < total; ++k) {
            const uint64_t v = symbols[k];
            bad |= v >= m->emits;
            dst[k] = static_cast<uint8_t>(v);
        }
    } else {
        uint16_t* dst = reinterpret_cast<uint16_t*>(host + o_sym);
        for (uint64_t k = 0; k < total; ++k) {
            const uint64_t v = symbols[k];
            bad |= v >= m->emits;
            dst[k] = static_cast<uint16_t>(v);
        }
    }
    if (bad) {
        cudaFreeHost(host);
        return sv_fail(SV_ERR_ARG, "sv_seqs_create: symbol >= emit_num");
    }
    std::memset(host + o_sym + total * sym_bytes, 0, 16);
    std::memcpy(host + o_off, offsets, (nseq + 1) * sizeof(uint64_t));
    std::memcpy(host + o_ord, order.data(), nseq * sizeof(uint32_t));

    sv_seqs* s = new (std::nothrow) sv_seqs;
    if (s == nullptr) {
        cudaFreeHost(host);
        return sv_fail(SV_ERR_MEMORY, "sv_seqs_create: out of host memory");
    }
    s->device = m->device;
    s->nseq = nseq;
    s->total = total;
    s->max_len = max_len;
    cudaError_t e = cudaSetDevice(m->device);
    if (e == cudaSuccess) e = cudaMalloc(&s->d_blob, bytes);
    if (e == cudaSuccess) e = cudaMemcpy(s->d_blob, host, bytes, cudaMemcpyHostToDevice);
    cudaFreeHost(host);
    if (e != cudaSuccess) {
        if (s->d_blob) cudaFree(s->d_blob);
        delete s;
        return sv_fail(SV_ERR_CUDA, std::string("sv_seqs_create: ") + cudaGetErrorString(e));
    }
    auto* base = static_cast<unsigned char*>(s->d_blob);
    s->dev.nseq = static_cast<int>(nseq);
    s->dev.sym_bytes = sym_bytes;
    s->dev.sym = base + o_sym;
    s->dev.off = reinterpret_cast<const uint64_t*>(base + o_off);
    s->dev.order = reinterpret_cast<const uint32_t*>(base + o_ord);
    *out = s;
    return SV_OK;
}

} // namespace

extern "C" int sv_seqs_create(const sv_model* m, const uint64_t* symbols, const uint64_t* offsets,
                              uint64_t nseq, sv_seqs** out) {
    return seqs_create_impl(m, symbols, offsets, nseq, out);
}
extern "C" int sv_seqs_create_u8(const sv_model* m, const uint8_t* symbols,
                                 const uint64_t* offsets, uint64_t nseq, sv_seqs** out) {
    return seqs_create_impl(m, symbols, offsets, nseq, out);
}
extern "C" void sv_seqs_destroy(sv_seqs* s) {
    if (s == nullptr) return;
    cudaSetDevice(s->device);
    if (s->d_blob) cudaFree(s->d_blob);
    delete s;
}
extern "C" uint64_t sv_seqs_count(const sv_seqs* s) { return s ? s->nseq : 0; }
extern "C" uint64_t sv_seqs_total_symbols(const sv_seqs* s) { return s ? s->total : 0; }

// ---------------------------------------------------------------- plain path ----
extern "C" int sv_viterbi_device(sv_model* m, const sv_seqs* seqs, float* d_out, float* d_best,
                                 int32_t* d_best_state, void* stream) {
    if (m == nullptr || seqs == nullptr) return sv_fail(SV_ERR_ARG, "sv_viterbi_device: NULL");
    if (seqs->device != m->device)
        return sv_fail(SV_ERR_ARG, "sv_viterbi_device: sequences live on another device");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    SV_CUDA(cudaSetDevice(m->device));
    SV_CUDA(cudaMemsetAsync(m->d_counter, 0, sizeof(unsigned), st));
    int rc;
    if (chosen_kernel(m) == SV_KERNEL_ES) {
        const int per_sm = std::max(1, sv_es_max_ctas_per_sm(m->es.variant_index, m->es.smem_bytes));
        const int64_t want = (static_cast<int64_t>(seqs->nseq) + m->es.var.q - 1) / m->es.var.q;
        const int grid = static_cast<int>(std::min<int64_t>(want, int64_t(m->sm_count) * per_sm));
        rc = sv_es_launch(m->es.variant_index, m->esdev, seqs->dev, d_out, d_best, d_best_state,
                          m->d_counter, grid, m->es.smem_bytes, st);
    } else {
        const int grid = static_cast<int>(std::min<uint64_t>(seqs->nseq, m->sm_count));
        rc = sv_generic_launch(m->gen, seqs->dev, d_out, d_best, d_best_state, m->d_counter, grid,
                               st);
    }
    m->last_launches = 1;
    if (rc != 0) return sv_fail(SV_ERR_CUDA, cudaGetErrorString(static_cast<cudaError_t>(rc)));
    return SV_OK;
}

extern "C" int sv_viterbi_batch(sv_model* m, const uint64_t* symbols, const uint64_t* offsets,
                                uint64_t nseq, float* out, float* best, int32_t* best_state) {
    if (m == nullptr) return sv_fail(SV_ERR_ARG, "sv_viterbi_batch: NULL model");
    sv_seqs* seqs = nullptr;
    int rc = sv_seqs_create(m, symbols, offsets, nseq, &seqs);
    if (rc != SV_OK) return rc;
    float *d_out = nullptr, *d_best = nullptr;
    int32_t* d_arg = nullptr;
    auto cleanup = [&]() {
        if (d_out) cudaFree(d_out);
        if (d_best) cudaFree(d_best);
        if (d_arg) cudaFree(d_arg);
        sv_seqs_destroy(seqs);
    };
    cudaError_t e = cudaSuccess;
    if (out && e == cudaSuccess) e = cudaMalloc(&d_out, nseq * m->n * sizeof(float));
    if (best && e == cudaSuccess) e = cudaMalloc(&d_best, nseq * sizeof(float));
    if (best_state && e == cudaSuccess) e = cudaMalloc(&d_arg, nseq * sizeof(int32_t));
    if (e != cudaSuccess) {
        cleanup();
        return sv_fail(SV_ERR_MEMORY, std::string("sv_viterbi_batch: ") + cudaGetErrorString(e));
    }
    rc = sv_viterbi_device(m, seqs, d_out, d_best, d_arg, m->stream);
    if (rc == SV_OK) {
        if (out)
            e = cudaMemcpyAsync(out, d_out, nseq * m->n * sizeof(float), cudaMemcpyDeviceToHost,
                                m->stream);
        if (best && e == cudaSuccess)
            e = cudaMemcpyAsync(best, d_best, nseq * sizeof(float), cudaMemcpyDeviceToHost,
                                m->stream);
        if (best_state && e == cudaSuccess)
            e = cudaMemcpyAsync(best_state, d_arg, nseq * sizeof(int32_t), cudaMemcpyDeviceToHost,
                                m->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(m->stream);
        if (e != cudaSuccess)
            rc = sv_fail(SV_ERR_CUDA, std::string("sv_viterbi_batch: ") + cudaGetErrorString(e));
    }
    cleanup();
    return rc;
}

extern "C" int sv_viterbi(sv_model* m, const uint64_t* seq, uint64_t len, float* out) {
    if (seq == nullptr || out == nullptr) return sv_fail(SV_ERR_ARG, "sv_viterbi: NULL");
    const uint64_t offsets[2] = {0, len};
    return sv_viterbi_batch(m, seq, offsets, 1, out, nullptr, nullptr);
}

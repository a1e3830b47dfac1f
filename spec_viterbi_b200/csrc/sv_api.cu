// C ABI: errors, model handle (K0 repack + upload), sequence batches, plain-path entry points.
#include "sv_internal.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <mutex>
#include <new>
#include <numeric>
#include <thread>
#include <type_traits>

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
    SV_API_BEGIN
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
    SV_API_END
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

int upload_es(sv_model* m, const SvEsPlan& p, SvEsDev& d, void** blob) {
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
    SV_CUDA(cudaMalloc(blob, bytes));
    SV_CUDA(cudaMemcpy(*blob, host.data(), bytes, cudaMemcpyHostToDevice));
    auto* base = static_cast<unsigned char*>(*blob);
    d.n = static_cast<int>(m->n);
    d.n_int = p.n_int;
    d.n_light = p.n_light;
    d.n_heavy = p.n_heavy;
    {
        auto ok = [](float x) { return !(x < 0.0f) && x == x; };
        bool nn = true;
        for (const float x : m->csr.w) nn &= ok(x);
        for (const float x : m->emis) nn &= ok(x);
        for (const float x : m->start) nn &= ok(x);
        d.nonneg = nn ? 1 : 0;
    }
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

int upload_ls(sv_model* m, const SvLsPlan& p, SvLsDev& d, void** blob) {
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
    SV_CUDA(cudaMalloc(blob, bytes));
    SV_CUDA(cudaMemcpy(*blob, host.data(), bytes, cudaMemcpyHostToDevice));
    auto* base = static_cast<unsigned char*>(*blob);
    d.n = static_cast<int>(m->n);
    d.n_int = p.n_int;
    d.n_lpad = p.n_lpad;
    d.n_pad = p.n_pad;
    d.n_heavy = p.n_heavy;
    d.emits = static_cast<int>(m->emits);
    d.ranks = p.ranks;
    d.tmem_cols = p.tmem_cols;
    d.nonneg = p.nonneg ? 1 : 0;
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

int upload_chain(sv_model* m) {
    const SvChainPlan& p = m->chain;
    size_t bytes = 0;
    const size_t o_wp = blob_reserve<float>(&bytes, p.w_pred.size());
    const size_t o_wh = blob_reserve<float>(&bytes, p.w_hub.size());
    const size_t o_emis = blob_reserve<float>(&bytes, p.emis_p.size());
    const size_t o_start = blob_reserve<float>(&bytes, p.start_p.size());
    const size_t o_es = blob_reserve<float>(&bytes, p.emis_s.size());
    const size_t o_orig = blob_reserve<int32_t>(&bytes, p.orig_of_pos.size());
    bytes += 256;
    std::vector<unsigned char> host(bytes, 0);
    std::memcpy(host.data() + o_wp, p.w_pred.data(), p.w_pred.size() * 4);
    std::memcpy(host.data() + o_wh, p.w_hub.data(), p.w_hub.size() * 4);
    std::memcpy(host.data() + o_emis, p.emis_p.data(), p.emis_p.size() * 4);
    std::memcpy(host.data() + o_start, p.start_p.data(), p.start_p.size() * 4);
    std::memcpy(host.data() + o_es, p.emis_s.data(), p.emis_s.size() * 4);
    std::memcpy(host.data() + o_orig, p.orig_of_pos.data(), p.orig_of_pos.size() * 4);
    SV_CUDA(cudaMalloc(&m->d_chain_blob, bytes));
    SV_CUDA(cudaMemcpy(m->d_chain_blob, host.data(), bytes, cudaMemcpyHostToDevice));
    auto* base = static_cast<unsigned char*>(m->d_chain_blob);
    SvChainDev& d = m->chaindev;
    d.n = static_cast<int>(m->n);
    d.n_pos = p.n_pos;
    d.emits = static_cast<int>(m->emits);
    d.nsp = p.nsp;
    d.tmem_cols = p.tmem_cols;
    d.nonneg = p.nonneg ? 1 : 0;
    for (int a = 0; a < SV_CHAIN_NSP; ++a) {
        d.special[a] = a < p.nsp ? static_cast<int>(p.special[a]) : -1;
        d.u[a] = p.u[a];
        d.start_s[a] = p.start_s[a];
        for (int b = 0; b < SV_CHAIN_NSP; ++b) d.ss[a][b] = p.ss[a][b];
    }
    d.w_pred = reinterpret_cast<const float*>(base + o_wp);
    d.w_hub = reinterpret_cast<const float*>(base + o_wh);
    d.emis_p = reinterpret_cast<const float*>(base + o_emis);
    d.start_p = reinterpret_cast<const float*>(base + o_start);
    d.emis_s = reinterpret_cast<const float*>(base + o_es);
    d.orig_of_pos = reinterpret_cast<const int32_t*>(base + o_orig);
    return SV_OK;
}

struct PlanTimer {
    sv_model* m;
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    explicit PlanTimer(sv_model* mm) : m(mm) {}
    ~PlanTimer() { m->plan_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count(); }
};

int host_threads() {
    const unsigned hc = std::thread::hardware_concurrency();
    return static_cast<int>(std::max(1u, std::min(16u, hc)));
}

} // namespace

// ------------------------------------------------------------ lazy K0 plans ----
// Plans are built the first time a launch needs them (a level >= 2 spec_with never does): ES ~5 ms,
// chain < 1 ms, lockstep: short colouring search now (~0.1 s for 2405 random states), full search on a
// background thread, finished plans kept in the on-disk plan cache (sv_plan_cache.cpp).
struct SvLsRefine {
    std::thread th;
    std::atomic<int> state{0}; // 0 = idle, 1 = running, 2 = plan ready, 3 = done (taken or failed)
    volatile int cancel = 0;
    SvLsPlan plan;
};

bool sv_model_ensure_es(sv_model* m) {
    if (m->es_state != 0) return m->es_state > 0;
    PlanTimer timer(m);
    m->es_state = -1;
    sv_es_build_plan(m->csr, m->emis, m->start, m->emits, m->smem_optin, &m->es);
    if (!m->es.valid) return false;
    if (upload_es(m, m->es, m->esdev, &m->d_es_blob) != SV_OK ||
        sv_es_prepare(m->es.variant_index, m->es.smem_bytes) != 0) {
        cudaGetLastError();
        m->es.valid = false; // kernel attribute rejected: keep the generic kernel
        return false;
    }
    m->es_state = 1;
    return true;
}

bool sv_model_ensure_es1(sv_model* m) {
    if (m->es1_state != 0) return m->es1_state > 0;
    m->es1_state = -1;
    if (!sv_model_ensure_es(m) || m->es.var.q != 2) return false; // one-sequence-per-CTA twin of a 2-sequence plan
    PlanTimer timer(m);
    sv_es_build_plan(m->csr, m->emis, m->start, m->emits, m->smem_optin, &m->es1, 1);
    if (!m->es1.valid) return false;
    if (upload_es(m, m->es1, m->esdev1, &m->d_es1_blob) != SV_OK ||
        sv_es_prepare(m->es1.variant_index, m->es1.smem_bytes) != 0) {
        cudaGetLastError();
        m->es1.valid = false;
        return false;
    }
    m->es1_state = 1;
    return true;
}

namespace {
bool install_ls(sv_model* m, SvLsPlan&& plan) {
    void* blob = nullptr;
    SvLsDev dev;
    if (upload_ls(m, plan, dev, &blob) != SV_OK || sv_ls_prepare(plan.variant_index, 2, plan.smem_bytes) != 0 ||
        sv_ls_prepare(plan.variant_index, 1, plan.smem_bytes4) != 0) {
        cudaGetLastError();
        if (blob) cudaFree(blob);
        return false;
    }
    if (m->d_ls_blob) m->retired_blobs.push_back(m->d_ls_blob); // a kernel in flight may still read it
    m->d_ls_blob = blob;
    m->lsdev = dev;
    m->ls = std::move(plan);
    return true;
}

// Takes over the refined plan if the background search has finished (wait = block until it has).
void adopt_refined_ls(sv_model* m, bool wait) {
    SvLsRefine* r = m->refine;
    if (r == nullptr) return;
    if (wait && r->th.joinable()) r->th.join();
    if (r->state.load(std::memory_order_acquire) != 2) return;
    if (r->th.joinable()) r->th.join();
    if (r->plan.valid && r->plan.gather_wavefronts < m->ls.gather_wavefronts) install_ls(m, std::move(r->plan));
    r->state.store(3, std::memory_order_release);
}
} // namespace

bool sv_model_ensure_ls(sv_model* m) {
    if (m->ls_state != 0) {
        if (m->ls_state > 0) adopt_refined_ls(m, false);
        return m->ls_state > 0;
    }
    PlanTimer timer(m);
    m->ls_state = -1;
    // refine policy: "bg" (default) = short search now, full search on a host thread; "sync" = full search now;
    // "off" = short search only
    const char* pol = std::getenv("SV_PLAN_REFINE");
    const std::string policy = pol ? pol : "bg";
    const uint64_t full_iters = sv_ls_default_sa_iters(m->n);
    // knee of (K0 time + first-batch kernel time), measured for 2405 random states: 20 N iterations leave ~1160 gather
    // wavefronts per step and group (full search: 919) = +1.5 ms on a 38 ms batch, for ~15 ms less planning than 60 N
    uint64_t quick_iters = std::min<uint64_t>(full_iters, 20ull * m->n);
    if (const char* q = std::getenv("SV_LS_SA_QUICK_ITERS")) quick_iters = std::strtoull(q, nullptr, 10);
    SvLsPlan plan;
    bool refined = sv_plan_cache_load(m->fingerprint, &plan);
    if (!refined) {
        SvLsPlanOpts opts;
        opts.threads = host_threads();
        opts.sa_iters = policy == "sync" ? full_iters : std::max<uint64_t>(1, quick_iters);
        sv_ls_build_plan(m->csr, m->emis, m->start, m->emits, m->smem_optin, &plan, opts);
        refined = opts.sa_iters >= full_iters || plan.colour_cost_after == 0;
        if (plan.valid && refined) sv_plan_cache_store(m->fingerprint, plan);
    }
    if (!plan.valid || !install_ls(m, std::move(plan))) return false;
    m->ls_state = 1;
    if (!refined && policy == "bg") {
        SvLsRefine* r = new (std::nothrow) SvLsRefine;
        if (r != nullptr) {
            m->refine = r;
            r->state.store(1);
            const uint64_t fp = m->fingerprint;
            r->th = std::thread([m, r, full_iters, fp]() {
                try {
                    SvLsPlanOpts opts;
                    opts.sa_iters = full_iters;
                    opts.cancel = &r->cancel;
                    opts.threads = 2;
                    // reads only data that is immutable after sv_model_create
                    sv_ls_build_plan(m->csr, m->emis, m->start, m->emits, m->smem_optin, &r->plan, opts);
                    if (r->plan.valid) sv_plan_cache_store(fp, r->plan);
                } catch (...) {
                    r->plan.valid = false;
                }
                r->state.store(2, std::memory_order_release);
            });
        }
    }
    return true;
}

bool sv_model_ensure_chain(sv_model* m) {
    if (m->chain_state != 0) return m->chain_state > 0;
    PlanTimer timer(m);
    m->chain_state = -1;
    sv_chain_build_plan(m->csr, m->emis, m->start, m->emits, &m->chain);
    if (!m->chain.valid) return false;
    if (upload_chain(m) != SV_OK) {
        cudaGetLastError();
        m->chain.valid = false;
        return false;
    }
    for (const int q : {8, 4, 2, 1}) {
        size_t smem = 0;
        int per_sm = 0;
        if (sv_chain_prepare(m->chain.threads, q, m->chain.nonneg, static_cast<int>(m->emits), &smem, &per_sm) != 0) {
            cudaGetLastError();
            m->chain.valid = false;
            return false;
        }
        m->chain_smem[q == 8 ? 3 : q == 4 ? 2 : q == 2 ? 1 : 0] = smem;
        m->chain_ctas_per_sm = per_sm;
    }
    m->chain_state = 1;
    return true;
}

namespace {

// Sequences per CTA of the chain kernel for a batch: the step of a CTA costs about (2 + Q) units (fixed part:
// barrier, managers, special rows; Q sequences of relaxations), CTAs run in waves over the resident slots.
int chain_q_for(const sv_model* m, uint64_t nseq) {
    const uint64_t slots = static_cast<uint64_t>(std::max(1, m->sm_count)) * std::max(1, m->chain_ctas_per_sm);
    int best_q = 1;
    uint64_t best_cost = ~0ull;
    for (const int q : {8, 4, 2, 1}) {
        const uint64_t grid = std::min<uint64_t>((nseq + q - 1) / q, slots);
        const uint64_t wave = grid * static_cast<uint64_t>(q);
        const uint64_t full = nseq / wave, rest = nseq - full * wave;
        // the partial last wave costs a whole step time -- or, for 8-sequence CTAs, a 4-sequence one when it is
        // handed out 4 per CTA (the tail balance of sv_viterbi_device)
        const uint64_t tail = rest == 0 ? 0 : (q == 8 && full > 0 && rest <= 4 * grid) ? 2 + 4 : 2 + q;
        const uint64_t cost = full * static_cast<uint64_t>(2 + q) + tail;
        if (cost < best_cost) {
            best_cost = cost;
            best_q = q;
        }
    }
    return best_q;
}

// Kernel a launch over `nseq` sequences will use.  AUTO: the chain kernel whenever the model has the chain +
// specials structure; else the lockstep kernel when the batch can fill its sequence lanes on every SM; else one
// or two sequences per CTA (ES); else generic.  Builds the plans it asks about.
int chosen_kernel(sv_model* m, uint64_t nseq) {
    switch (m->kernel_pref) {
    case SV_KERNEL_GENERIC: return SV_KERNEL_GENERIC;
    case SV_KERNEL_ES: return sv_model_ensure_es(m) ? SV_KERNEL_ES : SV_KERNEL_GENERIC;
    case SV_KERNEL_LS: return sv_model_ensure_ls(m) ? SV_KERNEL_LS : SV_KERNEL_GENERIC;
    case SV_KERNEL_CHAIN: return sv_model_ensure_chain(m) ? SV_KERNEL_CHAIN : SV_KERNEL_GENERIC;
    default: break;
    }
    if (sv_model_ensure_chain(m)) return SV_KERNEL_CHAIN;
    if (nseq >= static_cast<uint64_t>(2) * m->sm_count && sv_model_ensure_ls(m)) return SV_KERNEL_LS;
    if (sv_model_ensure_es(m)) return SV_KERNEL_ES;
    if (sv_model_ensure_ls(m)) return SV_KERNEL_LS;
    return SV_KERNEL_GENERIC;
}

} // namespace

extern "C" int sv_model_create(uint64_t states, uint64_t emits, uint64_t trans,
                               const uint64_t* trans_src, const uint64_t* trans_dst,
                               const float* trans_w, const float* emissions, uint64_t nstart,
                               const uint64_t* start_cols, const float* start_w, int device,
                               sv_model** out) {
    SV_API_BEGIN
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
    if (cudaMalloc(&m->d_counters, SV_COUNTER_RING * 64) != cudaSuccess)
        return fail(sv_fail(SV_ERR_CUDA, "cudaMalloc(counters) failed"));
    if (cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking) != cudaSuccess)
        return fail(sv_fail(SV_ERR_CUDA, "cudaStreamCreate failed"));
    m->ws_mutex = new (std::nothrow) std::mutex;
    if (m->ws_mutex == nullptr) return fail(sv_fail(SV_ERR_MEMORY, "sv_model_create: out of host memory"));
    m->fingerprint = sv_model_fingerprint(m->csr, m->emis, m->start, emits);
    // K0 plans (ES / lockstep / chain) are built by the first launch that needs them: sv_model_ensure_*
    *out = m;
    return SV_OK;
    SV_API_END
}

extern "C" void sv_model_destroy(sv_model* m) {
    if (m == nullptr) return;
    if (m->refine) { // stop the background search before its inputs go away
        m->refine->cancel = 1;
        if (m->refine->th.joinable()) m->refine->th.join();
        delete m->refine;
    }
    cudaSetDevice(m->device);
    if (m->d_blob) cudaFree(m->d_blob);
    if (m->d_es_blob) cudaFree(m->d_es_blob);
    if (m->d_es1_blob) cudaFree(m->d_es1_blob);
    if (m->d_ls_blob) cudaFree(m->d_ls_blob);
    for (void* b : m->retired_blobs) cudaFree(b);
    if (m->d_chain_blob) cudaFree(m->d_chain_blob);
    if (m->d_pair_blob) cudaFree(m->d_pair_blob);
    if (m->d_path_blob) cudaFree(m->d_path_blob);
    if (m->d_counters) cudaFree(m->d_counters);
    if (m->ws.seqs) sv_seqs_destroy(m->ws.seqs);
    if (m->ws.d_out) cudaFree(m->ws.d_out);
    if (m->ws.d_best) cudaFree(m->ws.d_best);
    if (m->ws.d_arg) cudaFree(m->ws.d_arg);
    if (m->stream) cudaStreamDestroy(m->stream);
    delete static_cast<std::mutex*>(m->ws_mutex);
    delete m;
}

extern "C" int sv_model_set_kernel(sv_model* m, int kernel) {
    SV_API_BEGIN
    if (m == nullptr) return sv_fail(SV_ERR_ARG, "sv_model_set_kernel: NULL model");
    if (kernel < SV_KERNEL_AUTO || kernel > SV_KERNEL_CHAIN)
        return sv_fail(SV_ERR_ARG, "sv_model_set_kernel: unknown kernel");
    SV_CUDA(cudaSetDevice(m->device));
    if (kernel == SV_KERNEL_ES && !sv_model_ensure_es(m))
        return sv_fail(SV_ERR_UNSUPPORTED, "model does not fit the edge-stationary kernel");
    if (kernel == SV_KERNEL_LS && !sv_model_ensure_ls(m))
        return sv_fail(SV_ERR_UNSUPPORTED, "model does not fit the lockstep kernel");
    if (kernel == SV_KERNEL_CHAIN && !sv_model_ensure_chain(m))
        return sv_fail(SV_ERR_UNSUPPORTED, "model does not have the chain + specials structure");
    m->kernel_pref = kernel;
    return SV_OK;
    SV_API_END
}

// Builds every plan now (what sv_model_create did eagerly in version 0.1) and waits for the background
// refinement of the lockstep plan, so that timings taken afterwards see the final schedule.
extern "C" int sv_model_prepare(sv_model* m) {
    SV_API_BEGIN
    if (m == nullptr) return sv_fail(SV_ERR_ARG, "sv_model_prepare: NULL model");
    SV_CUDA(cudaSetDevice(m->device));
    sv_model_ensure_chain(m);
    sv_model_ensure_es(m);
    sv_model_ensure_es1(m);
    if (sv_model_ensure_ls(m)) adopt_refined_ls(m, true);
    return SV_OK;
    SV_API_END
}

extern "C" int sv_model_get_info(const sv_model* m, sv_model_info* info) {
    SV_API_BEGIN
    if (m == nullptr || info == nullptr) return sv_fail(SV_ERR_ARG, "sv_model_get_info: NULL");
    std::memset(info, 0, sizeof(*info));
    info->states = m->n;
    info->emits = m->emits;
    info->nnz = m->nnz;
    info->max_in_degree = m->max_deg;
    info->device = m->device;
    info->kernel = m->last_kernel; // kernel of the last launch (0 before the first one)
    info->plan_ms = m->plan_ms;
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
    if (m->ls.valid) {
        info->ls_threads = m->ls.var.threads;
        info->ls_seqs_per_cta = m->last_seqs_per_cta && m->last_kernel == SV_KERNEL_LS ? m->last_seqs_per_cta : 8;
        info->ls_ranks = m->ls.ranks;
        info->ls_heavy_rows = m->ls.n_heavy;
        info->ls_tmem_cols = m->ls.tmem_cols;
        info->ls_gather_wavefronts = m->ls.gather_wavefronts;
        info->ls_gather_ideal = m->ls.gather_ideal;
        info->ls_smem_bytes = m->ls.smem_bytes;
        info->ls_split_states = m->ls.split_states;
        info->ls_colour_cost_before = m->ls.colour_cost_before;
        info->ls_colour_cost_after = m->ls.colour_cost_after;
        info->ls_sa_iters = m->ls.sa_iters;
    }
    if (m->chain.valid) {
        info->chain_threads = m->chain.threads;
        info->chain_states = m->chain.n_chain;
        info->chain_specials = m->chain.nsp;
        info->chain_seqs_per_cta = m->last_kernel == SV_KERNEL_CHAIN ? m->last_seqs_per_cta : 0;
        info->chain_tmem_cols = m->chain.tmem_cols;
    }
    return SV_OK;
    SV_API_END
}

extern "C" int sv_model_last_launches(const sv_model* m) { return m ? m->last_launches : 0; }

// ----------------------------------------------------------------- sequences ----
namespace {

int ensure_pinned(sv_seqs* s, size_t bytes) {
    if (s->h_cap >= bytes) return SV_OK;
    if (s->h_pinned) cudaFreeHost(s->h_pinned);
    s->h_pinned = nullptr;
    s->h_cap = 0;
    const size_t want = bytes + bytes / 4 + 4096;
    if (cudaMallocHost(&s->h_pinned, want) != cudaSuccess) {
        cudaGetLastError();
        return sv_fail(SV_ERR_MEMORY, "sequence upload: pinned staging allocation failed");
    }
    s->h_cap = want;
    return SV_OK;
}

int ensure_device(sv_seqs* s, size_t bytes) {
    if (s->d_cap >= bytes) return SV_OK;
    if (s->d_blob) cudaFree(s->d_blob);
    s->d_blob = nullptr;
    s->d_cap = 0;
    const size_t want = bytes + bytes / 4 + 4096;
    if (cudaMalloc(&s->d_blob, want) != cudaSuccess) {
        cudaGetLastError();
        return sv_fail(SV_ERR_MEMORY, "sequence upload: device allocation failed");
    }
    s->d_cap = want;
    return SV_OK;
}

// Device-side validation of a directly uploaded u8 symbol stream (symbol < emit_num): the host
// never touches the caller's buffer (a scalar scan of 29 MB costs more than its PCIe copy).
__global__ void sv_validate_u8_kernel(const uint8_t* __restrict__ sym, uint64_t total, unsigned emits,
                                      unsigned* __restrict__ flag) {
    const uint64_t words = total / 16;
    const uint4* __restrict__ w = reinterpret_cast<const uint4*>(sym);
    const unsigned lim = emits * 0x01010101u;
    unsigned bad = 0;
    for (uint64_t k = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x; k < words;
         k += static_cast<uint64_t>(gridDim.x) * blockDim.x) {
        const uint4 v = w[k];
        bad |= __vcmpgeu4(v.x, lim) | __vcmpgeu4(v.y, lim) | __vcmpgeu4(v.z, lim) | __vcmpgeu4(v.w, lim);
    }
    if (blockIdx.x == 0 && threadIdx.x < total - words * 16) bad |= sym[words * 16 + threadIdx.x] >= emits;
    if (bad) atomicOr(flag, 1u);
}

// (Re)fills a sequence batch: validates, packs symbols to u8/u16, builds the longest-first
// work order and uploads on `st` (asynchronous; the pinned staging buffer is owned by `s`, so
// the caller must not refill `s` before the stream has consumed it).  u8 input that needs no
// repacking is copied straight from the caller's buffer.
template <typename SrcT>
int seqs_assign(sv_seqs* s, const sv_model* m, const SrcT* symbols, const uint64_t* offsets,
                uint64_t nseq, cudaStream_t st) {
    if (m == nullptr || offsets == nullptr) return sv_fail(SV_ERR_ARG, "sequence upload: NULL");
    if (nseq == 0 || nseq > 0x7fffffffu) return sv_fail(SV_ERR_ARG, "sequence upload: bad nseq");
    const uint64_t first = offsets[0];
    const uint64_t total = offsets[nseq] - first;
    if (total > 0 && symbols == nullptr) return sv_fail(SV_ERR_ARG, "sequence upload: NULL symbols");
    uint64_t max_len = 0, min_len = ~0ull;
    for (uint64_t q = 0; q < nseq; ++q) {
        if (offsets[q + 1] <= offsets[q])
            return sv_fail(SV_ERR_ARG, "sequence upload: empty sequence (the reference requires "
                                       "at least one symbol, GraphBLAS_impl.cpp:59)");
        const uint64_t len = offsets[q + 1] - offsets[q];
        max_len = std::max(max_len, len);
        min_len = std::min(min_len, len);
    }
    const int sym_bytes = m->emits <= 256 ? 1 : 2;
    const bool direct = sizeof(SrcT) == 1 && sym_bytes == 1; // no repacking needed
    SV_CUDA(cudaSetDevice(m->device));

    size_t dbytes = 0;
    const size_t o_sym = blob_reserve<uint8_t>(&dbytes, total * sym_bytes + 16);
    const size_t o_off = blob_reserve<uint64_t>(&dbytes, nseq + 1);
    const size_t o_ord = blob_reserve<uint32_t>(&dbytes, nseq);
    const size_t o_flag = blob_reserve<uint32_t>(&dbytes, 4);
    dbytes += 256;
    // host staging: [offsets][order][symbols unless direct]
    size_t hbytes = 0;
    const size_t h_off = blob_reserve<uint64_t>(&hbytes, nseq + 1);
    const size_t h_ord = blob_reserve<uint32_t>(&hbytes, nseq);
    const size_t h_sym = blob_reserve<uint8_t>(&hbytes, direct ? 16 : total * sym_bytes + 16);
    const size_t h_flag = blob_reserve<uint32_t>(&hbytes, 4);
    int rc = ensure_pinned(s, hbytes);
    if (rc != SV_OK) return rc;
    if ((rc = ensure_device(s, dbytes)) != SV_OK) return rc;
    unsigned char* host = s->h_pinned;

    // validation (+ packing)
    const SrcT* src = symbols + first;
    unsigned long long worst = 0;
    if (direct) {
        // validated on the device after the copy (below)
    } else {
        // narrowing copy + range check, by several host threads for large batches (the u64 entry point of the
        // reference's interface moves 8 host bytes per symbol: one core caps it near 1 GB/s of symbols)
        auto pack_range = [&](auto* dst, uint64_t k0, uint64_t k1, unsigned long long* worst_out) {
            unsigned long long w = 0;
            for (uint64_t k = k0; k < k1; ++k) {
                const unsigned long long v = src[k];
                w = std::max(w, v);
                dst[k] = static_cast<std::remove_pointer_t<decltype(dst)>>(v);
            }
            *worst_out = w;
        };
        auto pack_all = [&](auto* dst) {
            const int nthr = total >= (1ull << 21) ? host_threads() : 1; // up to 16: the narrowing copy is bound by host memory bandwidth
            std::vector<unsigned long long> w(nthr, 0);
            std::vector<std::thread> pool;
            for (int t = 1; t < nthr; ++t)
                pool.emplace_back(pack_range, dst, total * t / nthr, total * (t + 1) / nthr, &w[t]);
            pack_range(dst, 0, total / nthr, &w[0]);
            for (auto& th : pool) th.join();
            for (const unsigned long long x : w) worst = std::max(worst, x);
            std::memset(dst + total, 0, 16);
        };
        if (sym_bytes == 1) pack_all(host + h_sym);
        else pack_all(reinterpret_cast<uint16_t*>(host + h_sym));
    }
    if (total > 0 && worst >= m->emits) return sv_fail(SV_ERR_ARG, "sequence upload: symbol >= emit_num");

    uint64_t* hoff = reinterpret_cast<uint64_t*>(host + h_off);
    for (uint64_t q = 0; q <= nseq; ++q) hoff[q] = offsets[q] - first;
    uint32_t* hord = reinterpret_cast<uint32_t*>(host + h_ord);
    std::iota(hord, hord + nseq, 0u);
    if (min_len != max_len) {
        std::stable_sort(hord, hord + nseq, [&](uint32_t a, uint32_t b) {
            return (offsets[a + 1] - offsets[a]) > (offsets[b + 1] - offsets[b]);
        });
    }

    auto* base = static_cast<unsigned char*>(s->d_blob);
    if (direct) {
        SV_CUDA(cudaMemcpyAsync(base + o_sym, src, total, cudaMemcpyHostToDevice, st));
        SV_CUDA(cudaMemsetAsync(base + o_sym + total, 0, 16, st));
        if (m->emits < 256 && total > 0) {
            unsigned* d_flag = reinterpret_cast<unsigned*>(base + o_flag);
            unsigned* h_flagp = reinterpret_cast<unsigned*>(host + h_flag);
            SV_CUDA(cudaMemsetAsync(d_flag, 0, sizeof(unsigned), st));
            sv_validate_u8_kernel<<<296, 256, 0, st>>>(base + o_sym, total, static_cast<unsigned>(m->emits), d_flag);
            SV_CUDA(cudaMemcpyAsync(h_flagp, d_flag, sizeof(unsigned), cudaMemcpyDeviceToHost, st));
            SV_CUDA(cudaStreamSynchronize(st)); // an invalid symbol must never reach a step kernel
            if (*h_flagp != 0) return sv_fail(SV_ERR_ARG, "sequence upload: symbol >= emit_num");
        }
    } else {
        SV_CUDA(cudaMemcpyAsync(base + o_sym, host + h_sym, total * sym_bytes + 16,
                                cudaMemcpyHostToDevice, st));
    }
    SV_CUDA(cudaMemcpyAsync(base + o_off, hoff, (nseq + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, st));
    SV_CUDA(cudaMemcpyAsync(base + o_ord, hord, nseq * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
    s->device = m->device;
    s->nseq = nseq;
    s->total = total;
    s->max_len = max_len;
    s->min_len = min_len;
    s->h2d_bytes = total * sym_bytes + (nseq + 1) * sizeof(uint64_t) + nseq * sizeof(uint32_t);
    s->dev.nseq = static_cast<int>(nseq);
    s->dev.sym_bytes = sym_bytes;
    s->dev.sym = base + o_sym;
    s->dev.off = reinterpret_cast<const uint64_t*>(base + o_off);
    s->dev.order = reinterpret_cast<const uint32_t*>(base + o_ord);
    return SV_OK;
}

template <typename SrcT>
int seqs_create_impl(const sv_model* m, const SrcT* symbols, const uint64_t* offsets,
                     uint64_t nseq, sv_seqs** out) {
    if (out == nullptr) return sv_fail(SV_ERR_ARG, "sv_seqs_create: out is NULL");
    *out = nullptr;
    if (m == nullptr) return sv_fail(SV_ERR_ARG, "sv_seqs_create: NULL model");
    sv_seqs* s = new (std::nothrow) sv_seqs;
    if (s == nullptr) return sv_fail(SV_ERR_MEMORY, "sv_seqs_create: out of host memory");
    s->device = m->device;
    int rc = seqs_assign(s, m, symbols, offsets, nseq, nullptr);
    if (rc == SV_OK && cudaStreamSynchronize(nullptr) != cudaSuccess)
        rc = sv_fail(SV_ERR_CUDA, "sv_seqs_create: upload failed");
    if (rc != SV_OK) {
        sv_seqs_destroy(s);
        return rc;
    }
    // the staging buffer is only needed for refills
    cudaFreeHost(s->h_pinned);
    s->h_pinned = nullptr;
    s->h_cap = 0;
    *out = s;
    return SV_OK;
}

} // namespace

extern "C" int sv_seqs_create(const sv_model* m, const uint64_t* symbols, const uint64_t* offsets,
                              uint64_t nseq, sv_seqs** out) {
    SV_API_BEGIN
    return seqs_create_impl(m, symbols, offsets, nseq, out);
    SV_API_END
}
extern "C" int sv_seqs_create_u8(const sv_model* m, const uint8_t* symbols,
                                 const uint64_t* offsets, uint64_t nseq, sv_seqs** out) {
    SV_API_BEGIN
    return seqs_create_impl(m, symbols, offsets, nseq, out);
    SV_API_END
}
extern "C" void sv_seqs_destroy(sv_seqs* s) {
    if (s == nullptr) return;
    cudaSetDevice(s->device);
    if (s->d_blob) cudaFree(s->d_blob);
    if (s->h_pinned) cudaFreeHost(s->h_pinned);
    delete s;
}
extern "C" uint64_t sv_seqs_count(const sv_seqs* s) { return s ? s->nseq : 0; }
extern "C" uint64_t sv_seqs_total_symbols(const sv_seqs* s) { return s ? s->total : 0; }

// Host-buffer workspace of a model: sequence batch + result buffers, grown on demand and
// reused by every sv_viterbi / sv_viterbi_batch* / sv_spec_run* call on that model.
int sv_ws_results(sv_model* m, uint64_t nseq, bool want_out, float** d_out, float** d_best,
                  int32_t** d_arg) {
    SvWorkspace& w = m->ws;
    if (want_out && w.out_cap < nseq * m->n) {
        if (w.d_out) cudaFree(w.d_out);
        w.d_out = nullptr;
        w.out_cap = 0;
        if (cudaMalloc(&w.d_out, nseq * m->n * sizeof(float)) != cudaSuccess) {
            cudaGetLastError();
            return sv_fail(SV_ERR_MEMORY, "result buffer does not fit device memory");
        }
        w.out_cap = nseq * m->n;
    }
    if (w.best_cap < nseq) {
        if (w.d_best) cudaFree(w.d_best);
        if (w.d_arg) cudaFree(w.d_arg);
        w.d_best = nullptr;
        w.d_arg = nullptr;
        w.best_cap = 0;
        if (cudaMalloc(&w.d_best, nseq * sizeof(float)) != cudaSuccess ||
            cudaMalloc(&w.d_arg, nseq * sizeof(int32_t)) != cudaSuccess) {
            cudaGetLastError();
            return sv_fail(SV_ERR_MEMORY, "result buffer does not fit device memory");
        }
        w.best_cap = nseq;
    }
    *d_out = want_out ? w.d_out : nullptr;
    *d_best = w.d_best;
    *d_arg = w.d_arg;
    return SV_OK;
}

template <typename SrcT>
int sv_ws_upload(sv_model* m, const SrcT* symbols, const uint64_t* offsets, uint64_t nseq) {
    if (m->ws.seqs == nullptr) {
        m->ws.seqs = new (std::nothrow) sv_seqs;
        if (m->ws.seqs == nullptr) return sv_fail(SV_ERR_MEMORY, "out of host memory");
        m->ws.seqs->device = m->device;
    }
    return seqs_assign(m->ws.seqs, m, symbols, offsets, nseq, m->stream);
}
template int sv_ws_upload<uint64_t>(sv_model*, const uint64_t*, const uint64_t*, uint64_t);
template int sv_ws_upload<uint8_t>(sv_model*, const uint8_t*, const uint64_t*, uint64_t);

int sv_ws_download(sv_model* m, uint64_t nseq, const float* d_out, const float* d_best,
                   const int32_t* d_arg, float* out, float* best, int32_t* best_state) {
    cudaError_t e = cudaSuccess;
    if (out) e = cudaMemcpyAsync(out, d_out, nseq * m->n * sizeof(float), cudaMemcpyDeviceToHost, m->stream);
    if (best && e == cudaSuccess)
        e = cudaMemcpyAsync(best, d_best, nseq * sizeof(float), cudaMemcpyDeviceToHost, m->stream);
    if (best_state && e == cudaSuccess)
        e = cudaMemcpyAsync(best_state, d_arg, nseq * sizeof(int32_t), cudaMemcpyDeviceToHost, m->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(m->stream);
    if (e != cudaSuccess) return sv_fail(SV_ERR_CUDA, std::string("result download: ") + cudaGetErrorString(e));
    return SV_OK;
}

// ---------------------------------------------------------------- plain path ----
extern "C" int sv_viterbi_device(sv_model* m, const sv_seqs* seqs, float* d_out, float* d_best,
                                 int32_t* d_best_state, void* stream) {
    SV_API_BEGIN
    if (m == nullptr || seqs == nullptr) return sv_fail(SV_ERR_ARG, "sv_viterbi_device: NULL");
    if (seqs->device != m->device)
        return sv_fail(SV_ERR_ARG, "sv_viterbi_device: sequences live on another device");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    SV_CUDA(cudaSetDevice(m->device));
    // a private work-queue counter per launch: launches of one model may overlap on different streams
    unsigned* counter = m->d_counters + static_cast<size_t>(m->launch_seq++ % SV_COUNTER_RING) * 16;
    SV_CUDA(cudaMemsetAsync(counter, 0, sizeof(unsigned), st));
    int rc;
    const int kernel = chosen_kernel(m, seqs->nseq);
    m->last_nseq = seqs->nseq;
    m->last_kernel = kernel;
    // 8-sequence CTAs and the last, partial wave of a batch (see SvSeqsDev::tail_begin): when what is left after the full
    // waves is at most 4 sequences per CTA, it is handed out 4 per CTA.  Sequences of ONE length finish together, so
    // the tail can even be a second launch of the 4-sequence instantiation behind the first (its step is shorter than
    // an 8-sequence CTA's with one group skipped); ragged batches keep one launch and the in-kernel tail.
    const bool tail_on = [] {
        const char* env = std::getenv("SV_TAIL_BALANCE");
        return !(env != nullptr && env[0] == '0');
    }();
    auto tail_of = [&](int grid, uint64_t* full, uint64_t* rest) {
        const uint64_t wave = 8ull * static_cast<uint64_t>(grid);
        *full = seqs->nseq / wave * wave;
        *rest = seqs->nseq - *full;
        return tail_on && *full > 0 && *rest > 0 && *rest <= 4ull * static_cast<uint64_t>(grid);
    };
    const bool one_length = seqs->min_len == seqs->max_len;
    auto next_counter = [&](unsigned** c) -> cudaError_t {
        *c = m->d_counters + static_cast<size_t>(m->launch_seq++ % SV_COUNTER_RING) * 16;
        return cudaMemsetAsync(*c, 0, sizeof(unsigned), st);
    };
    m->last_launches = 1;
    if (kernel == SV_KERNEL_CHAIN) {
        const int q = chain_q_for(m, seqs->nseq);
        const int64_t want = (static_cast<int64_t>(seqs->nseq) + q - 1) / q;
        const int grid = static_cast<int>(std::min<int64_t>(want, int64_t(m->sm_count) * m->chain_ctas_per_sm));
        m->last_seqs_per_cta = q;
        auto smem_of = [&](int qq) { return m->chain_smem[qq == 8 ? 3 : qq == 4 ? 2 : qq == 2 ? 1 : 0]; };
        uint64_t full = 0, rest = 0;
        SvSeqsDev d = seqs->dev;
        if (q == 8 && tail_of(grid, &full, &rest)) {
            if (one_length) {
                d.nseq = static_cast<int>(full);
                SvSeqsDev dt = seqs->dev;
                dt.order += full;
                dt.nseq = static_cast<int>(rest);
                rc = sv_chain_launch(m->chain.threads, 8, m->chaindev, d, d_out, d_best, d_best_state, counter, grid, smem_of(8), st);
                unsigned* c2 = nullptr;
                if (rc == 0) rc = next_counter(&c2);
                if (rc == 0)
                    rc = sv_chain_launch(m->chain.threads, 4, m->chaindev, dt, d_out, d_best, d_best_state, c2,
                                         static_cast<int>((rest + 3) / 4), smem_of(4), st);
                m->last_launches = 2;
            } else {
                d.tail_begin = static_cast<unsigned>(full);
                rc = sv_chain_launch(m->chain.threads, 8, m->chaindev, d, d_out, d_best, d_best_state, counter, grid, smem_of(8), st);
            }
        } else {
            rc = sv_chain_launch(m->chain.threads, q, m->chaindev, d, d_out, d_best, d_best_state, counter, grid, smem_of(q), st);
        }
    } else if (kernel == SV_KERNEL_LS) {
        // 8 sequences per CTA when that fills every SM, else 4 (twice the CTAs, about half the step time each)
        int nq4 = seqs->nseq <= static_cast<uint64_t>(4) * m->sm_count ? 1 : 2;
        if (const char* env = std::getenv("SV_LS_NQ4")) nq4 = std::atoi(env) == 1 ? 1 : 2;
        const int q = 4 * nq4;
        const int64_t want = (static_cast<int64_t>(seqs->nseq) + q - 1) / q;
        const int grid = static_cast<int>(std::min<int64_t>(want, m->sm_count));
        m->last_seqs_per_cta = q;
        uint64_t full = 0, rest = 0;
        SvSeqsDev d = seqs->dev;
        if (nq4 == 2 && tail_of(grid, &full, &rest)) {
            if (one_length) {
                d.nseq = static_cast<int>(full);
                SvSeqsDev dt = seqs->dev;
                dt.order += full;
                dt.nseq = static_cast<int>(rest);
                rc = sv_ls_launch(m->ls.variant_index, 2, m->lsdev, d, d_out, d_best, d_best_state, counter, grid, m->ls.smem_bytes, st);
                unsigned* c2 = nullptr;
                if (rc == 0) rc = next_counter(&c2);
                if (rc == 0)
                    rc = sv_ls_launch(m->ls.variant_index, 1, m->lsdev, dt, d_out, d_best, d_best_state, c2,
                                      static_cast<int>(std::min<uint64_t>((rest + 3) / 4, m->sm_count)), m->ls.smem_bytes4, st);
                m->last_launches = 2;
            } else {
                d.tail_begin = static_cast<unsigned>(full);
                rc = sv_ls_launch(m->ls.variant_index, 2, m->lsdev, d, d_out, d_best, d_best_state, counter, grid, m->ls.smem_bytes, st);
            }
        } else {
            rc = sv_ls_launch(m->ls.variant_index, nq4, m->lsdev, d, d_out, d_best, d_best_state, counter, grid,
                              nq4 == 2 ? m->ls.smem_bytes : m->ls.smem_bytes4, st);
        }
    } else if (kernel == SV_KERNEL_ES) {
        // a batch small enough that every sequence gets its own resident CTA runs one sequence per CTA
        // (shorter steps); larger batches advance two sequences per CTA in lockstep
        bool solo = false;
        if (m->es.var.q == 2 && static_cast<int64_t>(seqs->nseq) <= int64_t(m->sm_count) * 2 && sv_model_ensure_es1(m))
            solo = static_cast<int64_t>(seqs->nseq) <=
                   int64_t(m->sm_count) * std::max(1, sv_es_max_ctas_per_sm(m->es1.variant_index, m->es1.smem_bytes));
        const SvEsPlan& pl = solo ? m->es1 : m->es;
        const int per_sm = std::max(1, sv_es_max_ctas_per_sm(pl.variant_index, pl.smem_bytes));
        const int64_t want = (static_cast<int64_t>(seqs->nseq) + pl.var.q - 1) / pl.var.q;
        const int grid = static_cast<int>(std::min<int64_t>(want, int64_t(m->sm_count) * per_sm));
        m->last_seqs_per_cta = pl.var.q;
        rc = sv_es_launch(pl.variant_index, solo ? m->esdev1 : m->esdev, seqs->dev, d_out, d_best, d_best_state,
                          counter, grid, pl.smem_bytes, st);
    } else {
        const int grid = static_cast<int>(std::min<uint64_t>(seqs->nseq, m->sm_count));
        m->last_seqs_per_cta = 1;
        rc = sv_generic_launch(m->gen, seqs->dev, d_out, d_best, d_best_state, counter, grid, st);
    }
    if (rc != 0) return sv_fail(SV_ERR_CUDA, cudaGetErrorString(static_cast<cudaError_t>(rc)));
    return SV_OK;
    SV_API_END
}

namespace {
// The caller's result array, as a pointer the kernels can store to: page-locked host memory (cudaMallocHost /
// cudaHostAlloc / cudaHostRegister) is device-accessible under unified addressing, so the step kernels write every
// final vector straight into it when its sequence finishes -- 4 N bytes per sequence trickle over PCIe behind the
// computation instead of one S x N x 4 byte device-to-host copy after it (39 MB = 0.7 ms of a 40 ms batch for BASELINE
// config 5).  Pageable memory (or SV_DIRECT_HOST_OUT=0) keeps the device staging buffer and the copy.
float* direct_host_out(float* out) {
    static const bool on = [] {
        const char* e = std::getenv("SV_DIRECT_HOST_OUT");
        return !(e != nullptr && e[0] == '0');
    }();
    if (!on || out == nullptr) return nullptr;
    cudaPointerAttributes at{};
    if (cudaPointerGetAttributes(&at, out) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    if (at.type != cudaMemoryTypeHost || at.devicePointer == nullptr) return nullptr;
    return static_cast<float*>(at.devicePointer);
}

template <typename SrcT>
int viterbi_batch_impl(sv_model* m, const SrcT* symbols, const uint64_t* offsets, uint64_t nseq,
                       float* out, float* best, int32_t* best_state) {
    if (m == nullptr) return sv_fail(SV_ERR_ARG, "sv_viterbi_batch: NULL model");
    std::lock_guard<std::mutex> lock(*static_cast<std::mutex*>(m->ws_mutex)); // one workspace per model
    int rc = sv_ws_upload(m, symbols, offsets, nseq);
    if (rc != SV_OK) return rc;
    // not for the chain kernel: a thread owns five CONSECUTIVE positions there, so a warp's store covers 640 bytes with
    // 20-byte strides -- fine for L2, but over PCIe it becomes many small writes (measured: e2e 3.0e12 -> 2.35e12)
    float* const direct = chosen_kernel(m, nseq) == SV_KERNEL_CHAIN ? nullptr : direct_host_out(out);
    float *d_out, *d_best;
    int32_t* d_arg;
    if ((rc = sv_ws_results(m, nseq, out != nullptr && direct == nullptr, &d_out, &d_best, &d_arg)) != SV_OK) return rc;
    rc = sv_viterbi_device(m, m->ws.seqs, direct != nullptr ? direct : d_out, best ? d_best : nullptr,
                           best_state ? d_arg : nullptr, m->stream);
    if (rc != SV_OK) return rc;
    return sv_ws_download(m, nseq, d_out, d_best, d_arg, direct != nullptr ? nullptr : out, best, best_state);
}
} // namespace

extern "C" int sv_viterbi_batch(sv_model* m, const uint64_t* symbols, const uint64_t* offsets,
                                uint64_t nseq, float* out, float* best, int32_t* best_state) {
    SV_API_BEGIN
    return viterbi_batch_impl(m, symbols, offsets, nseq, out, best, best_state);
    SV_API_END
}

extern "C" int sv_viterbi_batch_u8(sv_model* m, const uint8_t* symbols, const uint64_t* offsets,
                                   uint64_t nseq, float* out, float* best, int32_t* best_state) {
    SV_API_BEGIN
    return viterbi_batch_impl(m, symbols, offsets, nseq, out, best, best_state);
    SV_API_END
}

extern "C" int sv_viterbi(sv_model* m, const uint64_t* seq, uint64_t len, float* out) {
    SV_API_BEGIN
    if (seq == nullptr || out == nullptr) return sv_fail(SV_ERR_ARG, "sv_viterbi: NULL");
    const uint64_t offsets[2] = {0, len};
    return sv_viterbi_batch(m, seq, offsets, 1, out, nullptr, nullptr);
    SV_API_END
}

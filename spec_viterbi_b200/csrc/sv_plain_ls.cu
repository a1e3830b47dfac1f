// Lockstep persistent (min,+) Viterbi step kernel for sm_100a ("LS kernel") -- the batched
// "sequences as SpMM columns" step  v_t[i] = min_j fl(fl(E[o_t][i] + T'[i][j]) + v_{t-1}[j]).
//
// One persistent CTA per SM walks ALL time steps of Q = 4 or 8 sequences in lockstep:
//   * transitions (weight, source offset) are REGISTER resident for the whole kernel: a thread
//     owns up to 5 destination rows ("ranks") and keeps their in-edges in a statically indexed
//     register slice -- the model is read from memory exactly once per CTA;
//   * the score vectors live in shared memory, interleaved [state][4 sequences] and double
//     buffered, so one LDS.128 feeds four relaxations and the only per-relaxation memory
//     traffic is 4 bytes over the shared-memory crossbar.  K0 (sv_ls_plan.cpp) colours the
//     states so that the 8 lanes of every LDS.128 phase hit 8 distinct bank groups;
//   * the emission table E[symbol][state] lives in TENSOR MEMORY: TMEM lane = the thread's lane
//     in its warp quadrant, TMEM column = (warp group, symbol, rank).  The symbol of a sequence
//     is CTA-uniform, so E[o][own row] is one tcgen05.ld with a run-time column -- a datapath
//     that is separate from the shared-memory crossbar and costs no shared-memory capacity
//     (192 KB at N = 2405, which is what limited the ES kernel to one sequence per SM);
//   * one __syncthreads per time step (two with block-reduced heavy rows); lanes 0..Q-1 of
//     warp 0 are the sequence managers: they stream the symbols (8-byte loads, one step ahead)
//     and publish a small control block per step; finished sequences are written out and
//     replaced from a global longest-first queue, so ragged batches need no padding.
//
// Reference semantics: GraphBLAS_impl.cpp:59,64-73 (add order), GraphBLAS_helper.cpp:65
// (unreachable -> +inf).  Tensor cores are not used ((min,+) is not an MMA contraction); only
// their memory is.
#include "sv_internal.h"

#include <type_traits>

namespace {

__device__ __forceinline__ float pos_inf() { return __int_as_float(0x7f800000); }

template <int... S> struct LShape {
    static constexpr int R = sizeof...(S);
    __host__ __device__ static constexpr int at(int r) {
        constexpr int v[] = {S...};
        return v[r];
    }
    __host__ __device__ static constexpr int base(int r) {
        int b = 0;
        for (int k = 0; k < r; ++k) b += at(k);
        return b;
    }
    __host__ __device__ static constexpr int total() { return base(R); }
};

// ---- tensor memory ----
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_slot, uint32_t ncols) {
    const uint32_t dst = static_cast<uint32_t>(__cvta_generic_to_shared(smem_slot));
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst), "r"(ncols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_free(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tmem_st1(uint32_t taddr, float v) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(taddr), "r"(__float_as_uint(v))
                 : "memory");
}
__device__ __forceinline__ uint32_t tmem_ld1(uint32_t taddr) {
    uint32_t r;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(r) : "r"(taddr) : "memory");
    return r;
}
// Waits for the outstanding tcgen05.ld's; the loaded registers are operands so that no use of
// them can be scheduled above the wait.
__device__ __forceinline__ void tmem_wait4(uint32_t (&r)[4]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;" : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3])::"memory");
}

// Per-step control block published by the manager lanes.
struct alignas(16) LsCtl {
    uint32_t col[8];   // TMEM column offset of each sequence's current symbol: symbol * RSTRIDE
    uint32_t init_mask; // sequences whose step is v_0 = E[o_0] + start
    uint32_t fin_mask;  // sequences for which this step is the last one
    uint32_t any_act;   // sequences that take part in this step (0 = CTA is done)
    uint32_t pad;
};

template <int Q>
__device__ __forceinline__ void relax4(float w, const float4 v, const float* e, float* acc) {
    // fl(fl(E + T') + v), exact min -- never re-associated
    acc[0] = fminf(acc[0], (e[0] + w) + v.x);
    acc[1] = fminf(acc[1], (e[1] + w) + v.y);
    acc[2] = fminf(acc[2], (e[2] + w) + v.z);
    acc[3] = fminf(acc[3], (e[3] + w) + v.w);
}

// Runs slots [0, d) of the register slice [BASE, BASE + SMAX) (d warp-uniform).  Fall-through
// switch: every register index is a compile-time constant.
template <int NQ4, int BASE, int SMAX, int NSLOTS>
__device__ __forceinline__ void ls_run_slots(int d, const float (&ew)[NSLOTS], const uint32_t (&eo)[NSLOTS],
                                             const unsigned char* vcur, uint32_t gstride,
                                             const float (&e)[4 * NQ4], float (&acc)[4 * NQ4]) {
#define SV_LS_SLOT(K)                                                                           \
    if constexpr ((K) < SMAX) {                                                                 \
        const unsigned char* src = vcur + eo[BASE + (K)];                                       \
        _Pragma("unroll") for (int g = 0; g < NQ4; ++g)                                         \
            relax4<4>(ew[BASE + (K)], *reinterpret_cast<const float4*>(src + g * gstride),      \
                      &e[4 * g], &acc[4 * g]);                                                  \
    }
    switch (d) {
    case 8: SV_LS_SLOT(7) [[fallthrough]];
    case 7: SV_LS_SLOT(6) [[fallthrough]];
    case 6: SV_LS_SLOT(5) [[fallthrough]];
    case 5: SV_LS_SLOT(4) [[fallthrough]];
    case 4: SV_LS_SLOT(3) [[fallthrough]];
    case 3: SV_LS_SLOT(2) [[fallthrough]];
    case 2: SV_LS_SLOT(1) [[fallthrough]];
    case 1: SV_LS_SLOT(0) [[fallthrough]];
    default: break;
    }
#undef SV_LS_SLOT
}

} // namespace

namespace {

template <int TPG, int NQ4, typename SH, int NH, int HS>
__global__ void __launch_bounds__(TPG, 1)
sv_ls_kernel(SvLsDev m, SvSeqsDev s, float* __restrict__ out, float* __restrict__ best,
             int32_t* __restrict__ arg, unsigned* __restrict__ counter) {
    constexpr int Q = 4 * NQ4;
    constexpr int NW = TPG / 32;
    constexpr int NL = SH::total();
    constexpr int NSL = NL + NH * HS;
    constexpr int RS = SV_LS_RSTRIDE;
    static_assert(SH::R == SV_ES_RMAX, "shape must list SV_ES_RMAX ranks");
    static_assert(NH <= SV_ES_NHMAX, "too many heavy rows");

    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int n_pad = m.n_pad;
    const uint32_t gstride = static_cast<uint32_t>(n_pad) * 16u;  // one float4 group
    const uint32_t bstride = gstride * NQ4;                       // one buffer
    float* red = reinterpret_cast<float*>(smem_raw + 2u * bstride); // [NH][NW][Q]
    __shared__ LsCtl ctl[2];
    __shared__ uint32_t s_sid[8];
    __shared__ uint32_t s_taddr;
    __shared__ float s_bv[NW];
    __shared__ int s_bi[NW];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned long long* __restrict__ sym64 = static_cast<const unsigned long long*>(s.sym);
    const int nranks = m.ranks;

    // ---- one-time: tensor memory <- emission table, registers <- this thread's edges ----
    if (warp == 0) tmem_alloc(&s_taddr, static_cast<uint32_t>(m.tmem_cols));
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tbase = s_taddr;
    // TMEM lanes of this warp's quadrant; columns of this warp group
    const uint32_t tE = tbase + (static_cast<uint32_t>(warp & 3) << 21) +
                        static_cast<uint32_t>(warp >> 2) * static_cast<uint32_t>(m.emits * RS);
    for (int o = 0; o < m.emits; ++o) {
        for (int r = 0; r < nranks; ++r)
            tmem_st1(tE + static_cast<uint32_t>(o * RS + r),
                     m.emis_p[static_cast<size_t>(o) * n_pad + r * TPG + tid]);
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");

    float ew[NSL];
    uint32_t eo[NSL];
#pragma unroll
    for (int k = 0; k < NSL; ++k) {
        ew[k] = m.edge_w[static_cast<size_t>(k) * TPG + tid];
        eo[k] = m.edge_off[static_cast<size_t>(k) * TPG + tid];
    }
    int deg[SV_ES_RMAX];
#pragma unroll
    for (int r = 0; r < SV_ES_RMAX; ++r) deg[r] = m.deg[warp * SV_ES_RMAX + r];

    // ---- sequence managers: lane q of warp 0 owns sequence slot q ----
    bool m_act = false;
    uint32_t m_t = 0, m_len = 0, m_sid = 0;
    uint64_t m_i = 0;                    // index (into the packed symbols) of the NEXT step's symbol
    unsigned long long m_cur = 0, m_nxt = 0;
    auto next_symbol = [&]() -> uint32_t {
        const uint32_t sy = static_cast<uint32_t>(m_cur >> ((m_i & 7u) * 8u)) & 0xffu;
        ++m_i;
        if ((m_i & 7u) == 0) {
            m_cur = m_nxt;
            m_nxt = sym64[(m_i >> 3) + 1];
        }
        return sy;
    };
    // Hands new sequences to the manager lanes in `mask` and patches the control block of the
    // coming step (fresh: overwrite it).  Called by all lanes of warp 0.
    auto refill = [&](uint32_t mask, LsCtl& dst, bool fresh) {
        bool started = false, one = false;
        uint32_t colv = 0;
        if (lane < Q && ((mask >> lane) & 1u)) {
            m_act = false;
            const unsigned w = atomicAdd(counter, 1u);
            if (w < static_cast<unsigned>(s.nseq)) {
                m_sid = s.order[w];
                const uint64_t b = s.off[m_sid];
                m_len = static_cast<uint32_t>(s.off[m_sid + 1] - b);
                m_t = 0;
                m_i = b;
                m_cur = sym64[m_i >> 3];
                m_nxt = sym64[(m_i >> 3) + 1];
                colv = next_symbol() * RS;
                m_act = true;
                started = true;
                one = m_len == 1;
                s_sid[lane] = m_sid;
            }
        }
        const uint32_t startm = __ballot_sync(0xffffffffu, started);
        const uint32_t onem = __ballot_sync(0xffffffffu, one);
        if (lane < Q && (fresh || ((mask >> lane) & 1u))) dst.col[lane] = colv;
        if (lane == 0) {
            if (fresh) {
                dst.init_mask = startm;
                dst.fin_mask = onem;
                dst.any_act = startm;
            } else {
                dst.init_mask |= startm;
                dst.fin_mask |= onem;
                dst.any_act |= startm;
            }
        }
        __syncwarp();
    };
    if (warp == 0) refill((1u << Q) - 1u, ctl[0], true);
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

    unsigned char* vcur = smem_raw;
    unsigned char* vnxt = smem_raw + bstride;

    for (uint32_t step = 0;; ++step) {
        const LsCtl& c = ctl[step & 1u];
        const uint32_t any_act = c.any_act;
        if (any_act == 0) break;
        const uint32_t init_mask = c.init_mask, fin_mask = c.fin_mask;
        uint32_t col[Q];
        {
            const uint4 c0 = *reinterpret_cast<const uint4*>(&c.col[0]);
            col[0] = c0.x; col[1] = c0.y; col[2] = c0.z; col[3] = c0.w;
            if constexpr (NQ4 == 2) {
                const uint4 c1 = *reinterpret_cast<const uint4*>(&c.col[4]);
                col[4] = c1.x; col[5] = c1.y; col[6] = c1.z; col[7] = c1.w;
            }
        }

        // ---- light rows ----
        auto rank_step = [&](auto rc) {
            constexpr int r = decltype(rc)::value;
            if (r < nranks) {
                const uint32_t p = static_cast<uint32_t>(tid + r * TPG);
                float e[Q], acc[Q];
#pragma unroll
                for (int g = 0; g < NQ4; ++g) {
                    uint32_t t4[4];
#pragma unroll
                    for (int k = 0; k < 4; ++k) t4[k] = tmem_ld1(tE + col[4 * g + k] + r);
                    tmem_wait4(t4);
#pragma unroll
                    for (int k = 0; k < 4; ++k) e[4 * g + k] = __uint_as_float(t4[k]);
                }
#pragma unroll
                for (int q = 0; q < Q; ++q) acc[q] = pos_inf();
                ls_run_slots<NQ4, SH::base(r), SH::at(r), NSL>(deg[r], ew, eo, vcur, gstride, e, acc);
                if (init_mask != 0) {
                    const float st = m.start_p[p];
#pragma unroll
                    for (int q = 0; q < Q; ++q)
                        if ((init_mask >> q) & 1u) acc[q] = e[q] + st; // v_0 = E[o_0] + start
                }
#pragma unroll
                for (int g = 0; g < NQ4; ++g)
                    *reinterpret_cast<float4*>(vnxt + g * gstride + p * 16u) =
                        make_float4(acc[4 * g], acc[4 * g + 1], acc[4 * g + 2], acc[4 * g + 3]);
            }
        };
        rank_step(std::integral_constant<int, 0>{});
        rank_step(std::integral_constant<int, 1>{});
        rank_step(std::integral_constant<int, 2>{});
        rank_step(std::integral_constant<int, 3>{});
        rank_step(std::integral_constant<int, 4>{});
        static_assert(SV_ES_RMAX == 5, "rank_step calls above assume 5 ranks");

        // ---- heavy rows: every thread relaxes a slice, block-wide min ----
        if constexpr (NH > 0) {
            if (m.n_heavy > 0) {
                auto heavy_step = [&](auto hc) {
                    constexpr int h = decltype(hc)::value;
                    if (h < m.n_heavy) {
                        const int ph = m.n_lpad + h;
                        float he[Q], acc[Q];
#pragma unroll
                        for (int q = 0; q < Q; ++q) {
                            he[q] = m.emis_p[static_cast<size_t>(col[q] / RS) * n_pad + ph];
                            acc[q] = pos_inf();
                        }
                        ls_run_slots<NQ4, NL + h * HS, HS, NSL>(m.heavy_slots[h], ew, eo, vcur, gstride, he, acc);
#pragma unroll
                        for (int q = 0; q < Q; ++q) {
#pragma unroll
                            for (int d = 16; d > 0; d >>= 1)
                                acc[q] = fminf(acc[q], __shfl_xor_sync(0xffffffffu, acc[q], d));
                        }
                        if (lane == 0) {
#pragma unroll
                            for (int q = 0; q < Q; ++q) red[(h * NW + warp) * Q + q] = acc[q];
                        }
                    }
                };
                heavy_step(std::integral_constant<int, 0>{});
                if constexpr (NH > 1) heavy_step(std::integral_constant<int, 1>{});
                static_assert(NH <= 2, "heavy_step calls above assume <= 2 heavy rows");
                __syncthreads();
                if (warp < m.n_heavy) {
                    // lane l: sequence q = l / LPQ, partials of warps (l % LPQ) + j * LPQ
                    constexpr int LPQ = 32 / Q;
                    const int h = warp, q = lane / LPQ, ph = m.n_lpad + h;
                    float x = pos_inf();
                    for (int w = lane % LPQ; w < NW; w += LPQ) x = fminf(x, red[(h * NW + w) * Q + q]);
#pragma unroll
                    for (int d = LPQ / 2; d > 0; d >>= 1) x = fminf(x, __shfl_xor_sync(0xffffffffu, x, d));
                    if ((init_mask >> q) & 1u)
                        x = m.emis_p[static_cast<size_t>(c.col[q] / RS) * n_pad + ph] + m.start_p[ph];
                    if (lane % LPQ == 0)
                        reinterpret_cast<float*>(vnxt + (q >> 2) * gstride + static_cast<uint32_t>(ph) * 16u)[q & 3] = x;
                }
            }
        }

        // ---- managers: publish the control block of the next step ----
        if (warp == 0) {
            LsCtl& nx = ctl[(step + 1u) & 1u];
            bool cont = false, last = false;
            uint32_t colv = 0;
            if (lane < Q && m_act) {
                if (m_t + 1u < m_len) {
                    cont = true;
                    colv = next_symbol() * RS;
                    last = m_t + 2u == m_len;
                    m_t += 1u;
                }
            }
            const uint32_t contm = __ballot_sync(0xffffffffu, cont);
            const uint32_t lastm = __ballot_sync(0xffffffffu, last);
            if (lane < Q) nx.col[lane] = colv;
            if (lane == 0) {
                nx.init_mask = 0;
                nx.fin_mask = lastm;
                nx.any_act = contm;
            }
        }
        __syncthreads();
        {
            unsigned char* sw = vcur;
            vcur = vnxt;
            vnxt = sw;
        }

        // ---- rare path: write out finished sequences, hand out new ones ----
        if (fin_mask != 0) {
            for (int q = 0; q < Q; ++q) {
                if (!((fin_mask >> q) & 1u)) continue;
                const uint32_t sid = s_sid[q];
                const float* vq = reinterpret_cast<const float*>(vcur + (q >> 2) * gstride) + (q & 3);
                float bv = pos_inf();
                int bi = 0x7fffffff;
                for (int i = tid; i < m.n; i += TPG) {
                    float x = pos_inf();
                    const uint32_t c1 = m.copy_ptr[i + 1];
                    for (uint32_t cc = m.copy_ptr[i]; cc < c1; ++cc) x = fminf(x, vq[m.copy_p[cc] * 4u]);
                    if (out != nullptr) out[static_cast<size_t>(sid) * m.n + i] = x;
                    if (x < bv) {
                        bv = x;
                        bi = i;
                    }
                }
                if (best != nullptr || arg != nullptr) {
#pragma unroll
                    for (int d = 16; d > 0; d >>= 1) {
                        const float ov = __shfl_xor_sync(0xffffffffu, bv, d);
                        const int oi = __shfl_xor_sync(0xffffffffu, bi, d);
                        if (ov < bv || (ov == bv && oi < bi)) {
                            bv = ov;
                            bi = oi;
                        }
                    }
                    if (lane == 0) {
                        s_bv[warp] = bv;
                        s_bi[warp] = bi;
                    }
                    __syncthreads();
                    if (warp == 0) {
                        bv = lane < NW ? s_bv[lane] : pos_inf();
                        bi = lane < NW ? s_bi[lane] : 0x7fffffff;
#pragma unroll
                        for (int d = 16; d > 0; d >>= 1) {
                            const float ov = __shfl_xor_sync(0xffffffffu, bv, d);
                            const int oi = __shfl_xor_sync(0xffffffffu, bi, d);
                            if (ov < bv || (ov == bv && oi < bi)) {
                                bv = ov;
                                bi = oi;
                            }
                        }
                        if (lane == 0) {
                            if (best != nullptr) best[sid] = bv;
                            if (arg != nullptr) arg[sid] = bi == 0x7fffffff ? -1 : bi;
                        }
                    }
                    __syncthreads();
                }
            }
            if (warp == 0) refill(fin_mask, ctl[(step + 1u) & 1u], false);
            __syncthreads();
        }
    }

    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) tmem_free(tbase, static_cast<uint32_t>(m.tmem_cols));
}

} // namespace

// ------------------------------------------------------------- variant table ----
namespace {

using LsSparse = LShape<8, 6, 5, 4, 3>;  // random sparse models (in-degree capped at 8)
using LsProfile = LShape<4, 3, 3, 2, 2>; // profile HMMs: <= 2 in-edges per light row + 2 dense rows

// order = order of preference in sv_ls_build_plan
const SvLsVariant kLsVariants[] = {
    // threads, nq4, shape, nh, hs
    {512, 2, {8, 6, 5, 4, 3}, 0, 0},
    {512, 2, {4, 3, 3, 2, 2}, 2, 5},
    {512, 1, {8, 6, 5, 4, 3}, 0, 0},
    {512, 1, {4, 3, 3, 2, 2}, 2, 5},
};
constexpr int kNumLsVariants = sizeof(kLsVariants) / sizeof(kLsVariants[0]);

const void* ls_variant_fn(int index) {
    switch (index) {
    case 0: return (const void*)sv_ls_kernel<512, 2, LsSparse, 0, 0>;
    case 1: return (const void*)sv_ls_kernel<512, 2, LsProfile, 2, 5>;
    case 2: return (const void*)sv_ls_kernel<512, 1, LsSparse, 0, 0>;
    case 3: return (const void*)sv_ls_kernel<512, 1, LsProfile, 2, 5>;
    default: return nullptr;
    }
}

} // namespace

int sv_ls_num_variants() { return kNumLsVariants; }
const SvLsVariant& sv_ls_variant(int index) { return kLsVariants[index]; }

int sv_ls_prepare(int index, size_t smem) {
    const void* fn = ls_variant_fn(index);
    if (fn == nullptr) return cudaErrorInvalidValue;
    return cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
}

int sv_ls_launch(int index, const SvLsDev& m, const SvSeqsDev& s, float* d_out, float* d_best,
                 int32_t* d_arg, unsigned* d_counter, int grid, size_t smem, cudaStream_t st) {
    const void* fn = ls_variant_fn(index);
    if (fn == nullptr || s.sym_bytes != 1) return cudaErrorInvalidValue;
    void* args[] = {(void*)&m, (void*)&s, (void*)&d_out, (void*)&d_best, (void*)&d_arg, (void*)&d_counter};
    cudaError_t e = cudaLaunchKernel(fn, dim3(grid), dim3(kLsVariants[index].threads), args, smem, st);
    return e != cudaSuccess ? e : cudaGetLastError();
}

// Edge-stationary persistent (min,+) Viterbi step kernel for sm_100a ("ES kernel").
//
// The batched step  v_t[i] = min_j fl(fl(E[o_t][i] + T'[i][j]) + v_{t-1}[j])  is bound by the
// gather of v[j] (4 B per relaxation) -- on B200 that is the shared-memory crossbar
// (128 B/clk/SM), not HBM and not the FP32 pipes.  Everything else is therefore moved
// OFF the crossbar:
//   * the transitions (weight, source offset) live in REGISTERS for the whole kernel:
//     each thread owns up to RMAX destination rows ("ranks") and holds their in-edges in a
//     statically indexed register file slice, so the model is read from memory exactly once;
//   * the emission table E[symbol][state] (192 KB at N=2407) is resident in shared memory
//     and read once per (row, sequence, step);
//   * the score vectors of Q sequences (interleaved [state][Q], double buffered) are the
//     only gathered data; K0 (sv_es_plan.cpp) orders every thread's edges so that the 16/32
//     lanes of one LDS hit distinct banks;
//   * one CTA walks all time steps of its sequences with one __syncthreads per step
//     (two when the model has block-reduced heavy rows); finished lanes pull the next
//     sequence from a global longest-first queue, so ragged batches need no padding.
// Per-rank trip counts are warp-uniform (rows are dealt to warps in in-degree order) and
// dispatched through a fall-through switch, which keeps every edge register statically
// indexed.  Rows whose in-degree exceeds the register budget are either split into
// virtual states at K0 (exact, because fl(u + min) = min fl(u + .)) or, for the dense rows
// of profile HMMs, spread over all threads and reduced with warp-shuffle mins.
//
// Reference semantics: GraphBLAS_impl.cpp:59,64-73 (add order), GraphBLAS_helper.cpp:65
// (unreachable -> +inf).  Tensor cores are not used: (min,+) is not an MMA contraction.
#include "sv_internal.h"

#include <type_traits>

namespace {

__device__ __forceinline__ float pos_inf() { return __int_as_float(0x7f800000); }

template <int... S> struct Shape {
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

template <int Q> struct VecOf;
template <> struct VecOf<1> {
    using type = float;
};
template <> struct VecOf<2> {
    using type = float2;
};

template <int Q>
__device__ __forceinline__ void load_v(const unsigned char* base, uint32_t off, float (&v)[Q]) {
    if constexpr (Q == 1) {
        v[0] = *reinterpret_cast<const float*>(base + off);
    } else {
        const float2 t = *reinterpret_cast<const float2*>(base + off);
        v[0] = t.x;
        v[1] = t.y;
    }
}

template <int Q>
__device__ __forceinline__ void store_v(unsigned char* base, uint32_t off, const float (&v)[Q]) {
    if constexpr (Q == 1) {
        *reinterpret_cast<float*>(base + off) = v[0];
    } else {
        *reinterpret_cast<float2*>(base + off) = make_float2(v[0], v[1]);
    }
}

// One relaxation per sequence lane: fl(fl(E + T') + v), then exact min.
template <int Q>
__device__ __forceinline__ void relax(float w, uint32_t off, const unsigned char* vcur,
                                      const float (&e)[Q], float (&acc)[Q]) {
    float v[Q];
    load_v<Q>(vcur, off, v);
#pragma unroll
    for (int q = 0; q < Q; ++q) {
        const float u = e[q] + w;
        acc[q] = fminf(acc[q], u + v[q]);
    }
}

// Runs slots [0, d) of a register slice of SMAX slots starting at BASE (d <= SMAX, warp-uniform):
// a binary tree of uniform branches into straight-line code for exactly d slots (every register
// index is a compile-time constant; a switch costs a constant-bank load and an indirect branch
// per slice, and this kernel is latency-bound).
template <int Q, int BASE, int D, int NSLOTS>
__device__ __forceinline__ void run_fixed(const float (&ew)[NSLOTS], const uint32_t (&eo)[NSLOTS],
                                          const unsigned char* vcur, const float (&e)[Q], float (&acc)[Q]) {
#pragma unroll
    for (int k = 0; k < D; ++k) relax<Q>(ew[BASE + k], eo[BASE + k], vcur, e, acc);
}
template <int Q, int BASE, int LO, int HI, int NSLOTS>
__device__ __forceinline__ void run_tree(int d, const float (&ew)[NSLOTS], const uint32_t (&eo)[NSLOTS],
                                         const unsigned char* vcur, const float (&e)[Q], float (&acc)[Q]) {
    if constexpr (LO == HI) {
        run_fixed<Q, BASE, LO, NSLOTS>(ew, eo, vcur, e, acc);
    } else {
        constexpr int MID = (LO + HI) / 2;
        if (d <= MID)
            run_tree<Q, BASE, LO, MID, NSLOTS>(d, ew, eo, vcur, e, acc);
        else
            run_tree<Q, BASE, MID + 1, HI, NSLOTS>(d, ew, eo, vcur, e, acc);
    }
}
template <int Q, int BASE, int SMAX, int NSLOTS>
__device__ __forceinline__ void run_slots(int d, const float (&ew)[NSLOTS],
                                          const uint32_t (&eo)[NSLOTS],
                                          const unsigned char* vcur, const float (&e)[Q],
                                          float (&acc)[Q]) {
    run_tree<Q, BASE, 0, SMAX, NSLOTS>(d, ew, eo, vcur, e, acc);
}

template <int TPG, int Q, typename SH, int NH, int HS, typename SymT>
__global__ void __launch_bounds__(TPG, (TPG >= 512 ? 1 : (TPG >= 256 ? 2 : 4)))
sv_es_kernel(SvEsDev m, SvSeqsDev s, float* __restrict__ out, float* __restrict__ best,
             int32_t* __restrict__ arg, unsigned* __restrict__ counter) {
    constexpr int NW = TPG / 32;
    constexpr int NL = SH::total();        // light slots
    constexpr int NSL = NL + NH * HS;      // all register slots
    static_assert(SH::R == SV_ES_RMAX, "shape must list SV_ES_RMAX ranks");
    static_assert(NH <= SV_ES_NHMAX, "too many heavy rows");

    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int n_int = m.n_int;
    const uint32_t vbytes = static_cast<uint32_t>(n_int) * Q * 4u; // one v buffer
    unsigned char* vbuf0 = smem_raw;
    unsigned char* vbuf1 = smem_raw + ((vbytes + 15u) & ~15u);
    float* red = reinterpret_cast<float*>(vbuf1 + ((vbytes + 15u) & ~15u)); // [2][NH][NW][Q], by step parity
    const float* Es = red + (NH > 0 ? 2 * NH * NW * Q : 0);                // [emits][n_int]
    __shared__ unsigned s_work[2];
    __shared__ float s_bv[NW];
    __shared__ int s_bi[NW];

    const int tid = threadIdx.x, lane = tid & 31;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0); // provably warp-uniform: uniform branches on deg[]
    const SymT* __restrict__ sym = static_cast<const SymT*>(s.sym);

    // ---- one-time loads: emission table -> smem, this thread's edges -> registers ----
    {
        float* Ew = const_cast<float*>(Es);
        const int total = m.emits * n_int;
        for (int k = tid; k < total; k += TPG) Ew[k] = m.emis_p[k];
    }
    float ew[NSL];
    uint32_t eo[NSL];
#pragma unroll
    for (int k = 0; k < NSL; ++k) {
        ew[k] = m.edge_w[static_cast<size_t>(k) * TPG + tid]; // K0 fills every slot (pads = +inf)
        eo[k] = m.edge_off[static_cast<size_t>(k) * TPG + tid];
    }
    int deg[SV_ES_RMAX];
#pragma unroll
    for (int r = 0; r < SV_ES_RMAX; ++r) deg[r] = m.deg[warp * SV_ES_RMAX + r];
    const int n_light = m.n_light;
    const int nranks = (n_light + TPG - 1) / TPG;
    __syncthreads();

    // ---- per-lane sequence state (CTA-uniform values) ----
    bool act[Q], init[Q];
    uint32_t sid[Q], rem[Q];      // sequence id, steps left
    const SymT* sp[Q];            // symbol of the current step
    uint32_t o_cur[Q], o_n1[Q];
#pragma unroll
    for (int q = 0; q < Q; ++q) {
        act[q] = false;
        init[q] = false;
        sid[q] = 0;
        rem[q] = 0;
        sp[q] = sym;
        o_cur[q] = 0;
        o_n1[q] = 0;
    }
    bool drained = false;
    [[maybe_unused]] int rpar = 0; // parity of the heavy-row partial buffer
    unsigned char* vcur = vbuf0;
    unsigned char* vnxt = vbuf1;

    for (;;) {
        // ---- (A) hand new sequences to idle lanes ----
        bool want = false;
#pragma unroll
        for (int q = 0; q < Q; ++q) want |= !act[q];
        if (want && !drained) {
            if (tid == 0) {
#pragma unroll
                for (int q = 0; q < Q; ++q)
                    if (!act[q]) s_work[q] = atomicAdd(counter, 1u);
            }
            __syncthreads();
#pragma unroll
            for (int q = 0; q < Q; ++q) {
                if (!act[q]) {
                    const unsigned w = s_work[q];
                    if (w >= static_cast<unsigned>(s.nseq)) {
                        drained = true;
                    } else {
                        sid[q] = s.order[w];
                        const uint64_t b0 = s.off[sid[q]];
                        rem[q] = static_cast<uint32_t>(s.off[sid[q] + 1] - b0);
                        sp[q] = sym + b0;
                        act[q] = true;
                        init[q] = true;
                        o_cur[q] = sp[q][0];
                        o_n1[q] = sp[q][1]; // the symbol buffer is padded past the end
                    }
                }
            }
        }
        bool any = false, any_init = false;
#pragma unroll
        for (int q = 0; q < Q; ++q) {
            any |= act[q];
            any_init |= init[q];
        }
        if (!any) break;

        // ---- (B) one time step for all lanes ----
        uint32_t o_n2[Q];
#pragma unroll
        for (int q = 0; q < Q; ++q) o_n2[q] = sp[q][2]; // consumed two steps later
        const float* erow[Q];
#pragma unroll
        for (int q = 0; q < Q; ++q) erow[q] = Es + static_cast<size_t>(o_cur[q]) * n_int;

        // warp minimum of NV values under ONE uniform branch (the integer REDUX is valid when all scores
        // are >= 0: float order == signed-int order)
        auto warp_min_all = [&](auto& x) {
            constexpr int NV = sizeof(x) / sizeof(float);
            if (m.nonneg) {
#pragma unroll
                for (int k = 0; k < NV; ++k)
                    (&x[0][0])[k] = __int_as_float(__reduce_min_sync(0xffffffffu, __float_as_int((&x[0][0])[k])));
            } else {
#pragma unroll
                for (int k = 0; k < NV; ++k) {
#pragma unroll
                    for (int d = 16; d > 0; d >>= 1)
                        (&x[0][0])[k] = fminf((&x[0][0])[k], __shfl_xor_sync(0xffffffffu, (&x[0][0])[k], d));
                }
            }
        };
        auto rank_step = [&](auto rc) {
            constexpr int r = decltype(rc)::value;
            if (r < nranks) {
                const int p = tid + r * TPG;
                const int pc = p < n_int ? p : n_int - 1;
                float e[Q], acc[Q];
#pragma unroll
                for (int q = 0; q < Q; ++q) {
                    e[q] = erow[q][pc];
                    acc[q] = pos_inf();
                }
                run_slots<Q, SH::base(r), SH::at(r), NSL>(deg[r], ew, eo, vcur, e, acc);
                if (p < n_light) store_v<Q>(vnxt, static_cast<uint32_t>(p) * Q * 4u, acc);
            }
        };
        if constexpr (NH > 0) {
            if (m.n_heavy > 0) {
                float hacc[NH][Q];
                auto heavy_step = [&](auto hc) {
                    constexpr int h = decltype(hc)::value;
                    float he[Q];
#pragma unroll
                    for (int q = 0; q < Q; ++q) {
                        he[q] = erow[q][n_light + (h < m.n_heavy ? h : 0)];
                        hacc[h][q] = pos_inf();
                    }
                    run_slots<Q, NL + h * HS, HS, NSL>(h < m.n_heavy ? m.heavy_slots[h] : 0, ew, eo, vcur, he, hacc[h]);
                    if (any_init) { // first step of a sequence: v_0 = E[o_0] + start enters as ONE lane's partial,
                                    // everybody else contributes +inf (min(x, +inf) = x: same bits)
#pragma unroll
                        for (int q = 0; q < Q; ++q)
                            if (init[q])
                                hacc[h][q] = (warp == 0 && lane == 0 && h < m.n_heavy) ? he[q] + m.start_p[n_light + h] : pos_inf();
                    }
                };
                heavy_step(std::integral_constant<int, 0>{});
                if constexpr (NH > 1) heavy_step(std::integral_constant<int, 1>{});
                static_assert(NH <= 2 && SV_ES_NHMAX == 2, "heavy_step calls above assume 2 heavy rows");
                warp_min_all(hacc);
                if (lane == 0) {
#pragma unroll
                    for (int h = 0; h < NH; ++h)
#pragma unroll
                        for (int q = 0; q < Q; ++q) red[((rpar * NH + h) * NW + warp) * Q + q] = hacc[h][q];
                }
            }
        }
        // light rows after the heavy partials: the REDUX / shared-memory write of the partials overlaps them
        rank_step(std::integral_constant<int, 0>{});
        rank_step(std::integral_constant<int, 1>{});
        rank_step(std::integral_constant<int, 2>{});
        rank_step(std::integral_constant<int, 3>{});
        rank_step(std::integral_constant<int, 4>{});
        static_assert(SV_ES_RMAX == 5, "rank_step calls above assume 5 ranks");
        if (any_init) { // rare: v_0 = E[o_0] + start overwrites the relaxed scores (same thread, program order)
#pragma unroll 1
            for (int r = 0; r < nranks; ++r) {
                const int p = tid + r * TPG;
                if (p < n_light) {
                    const float st = m.start_p[p];
#pragma unroll
                    for (int q = 0; q < Q; ++q)
                        if (init[q]) reinterpret_cast<float*>(vnxt)[p * Q + q] = erow[q][p] + st;
                }
            }
        }
        __syncthreads(); // the ONE barrier of a step
        {
            unsigned char* sw = vcur;
            vcur = vnxt;
            vnxt = sw;
        }
        // heavy rows: EVERY warp reduces the NW partials itself and writes the (identical) result into
        // the new current buffer -- its own lanes read it after the __syncwarp, so no second CTA barrier
        // is needed (the partials are double buffered by step parity)
        if constexpr (NH > 0) {
            if (m.n_heavy > 0) {
                static_assert(NW <= 32, "one lane per partial");
                float x[NH][Q];
#pragma unroll
                for (int h = 0; h < NH; ++h)
#pragma unroll
                    for (int q = 0; q < Q; ++q)
                        x[h][q] = lane < NW ? red[((rpar * NH + h) * NW + lane) * Q + q] : pos_inf();
                warp_min_all(x);
                if (lane == 0) {
#pragma unroll
                    for (int h = 0; h < NH; ++h)
                        if (h < m.n_heavy) store_v<Q>(vcur, static_cast<uint32_t>(n_light + h) * Q * 4u, x[h]);
                }
                __syncwarp();
            }
            rpar ^= 1;
        }

        // ---- (C) advance lanes; emit results of finished sequences ----
#pragma unroll
        for (int q = 0; q < Q; ++q) {
            init[q] = false;
            o_cur[q] = o_n1[q];
            o_n1[q] = o_n2[q];
            if (act[q]) {
                ++sp[q]; // only while active: an idle slot must not walk past the padded end of the symbols
                if (--rem[q] == 0) {
                    act[q] = false;
                    const float* vq = reinterpret_cast<const float*>(vcur) + q;
                    float bv = pos_inf();
                    int bi = 0x7fffffff;
                    for (int i = tid; i < m.n; i += TPG) {
                        float x = pos_inf();
                        const uint32_t c1 = m.copy_ptr[i + 1];
                        for (uint32_t c = m.copy_ptr[i]; c < c1; ++c)
                            x = fminf(x, vq[m.copy_p[c] * Q]);
                        if (out != nullptr) out[static_cast<size_t>(sid[q]) * m.n + i] = x;
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
                                if (best != nullptr) best[sid[q]] = bv;
                                if (arg != nullptr) arg[sid[q]] = bi == 0x7fffffff ? -1 : bi;
                            }
                        }
                        __syncthreads();
                    }
                }
            }
        }
    }
}

// ------------------------------------------------------------- variant table ----
using ShapeSparse = Shape<8, 6, 5, 4, 3>;  // random sparse models (in-degree capped at 8)
using ShapeProfile = Shape<4, 3, 3, 2, 2>; // profile HMMs: <=2 in-edges per light row

const SvEsVariant kVariants[] = {
    // threads, q, shape, nh, hs
    {512, 2, {8, 6, 5, 4, 3}, 0, 0},
    {512, 1, {8, 6, 5, 4, 3}, 0, 0},
    {256, 2, {8, 6, 5, 4, 3}, 0, 0},
    {256, 1, {8, 6, 5, 4, 3}, 0, 0},
    {128, 2, {8, 6, 5, 4, 3}, 0, 0},
    {128, 1, {8, 6, 5, 4, 3}, 0, 0},
    {512, 2, {4, 3, 3, 2, 2}, 2, 5},
    {512, 1, {4, 3, 3, 2, 2}, 2, 5},
    {256, 2, {4, 3, 3, 2, 2}, 2, 5},
    {256, 1, {4, 3, 3, 2, 2}, 2, 5},
    {128, 2, {4, 3, 3, 2, 2}, 2, 5},
    {128, 1, {4, 3, 3, 2, 2}, 2, 5},
};
constexpr int kNumVariants = sizeof(kVariants) / sizeof(kVariants[0]);

template <typename SymT> const void* variant_fn(int index) {
    switch (index) {
    case 0: return (const void*)sv_es_kernel<512, 2, ShapeSparse, 0, 0, SymT>;
    case 1: return (const void*)sv_es_kernel<512, 1, ShapeSparse, 0, 0, SymT>;
    case 2: return (const void*)sv_es_kernel<256, 2, ShapeSparse, 0, 0, SymT>;
    case 3: return (const void*)sv_es_kernel<256, 1, ShapeSparse, 0, 0, SymT>;
    case 4: return (const void*)sv_es_kernel<128, 2, ShapeSparse, 0, 0, SymT>;
    case 5: return (const void*)sv_es_kernel<128, 1, ShapeSparse, 0, 0, SymT>;
    case 6: return (const void*)sv_es_kernel<512, 2, ShapeProfile, 2, 5, SymT>;
    case 7: return (const void*)sv_es_kernel<512, 1, ShapeProfile, 2, 5, SymT>;
    case 8: return (const void*)sv_es_kernel<256, 2, ShapeProfile, 2, 5, SymT>;
    case 9: return (const void*)sv_es_kernel<256, 1, ShapeProfile, 2, 5, SymT>;
    case 10: return (const void*)sv_es_kernel<128, 2, ShapeProfile, 2, 5, SymT>;
    case 11: return (const void*)sv_es_kernel<128, 1, ShapeProfile, 2, 5, SymT>;
    default: return nullptr;
    }
}

} // namespace

int sv_es_num_variants() { return kNumVariants; }
const SvEsVariant& sv_es_variant(int index) { return kVariants[index]; }

int sv_es_prepare(int index, size_t smem) {
    for (int b = 0; b < 2; ++b) {
        const void* fn = b == 0 ? variant_fn<uint8_t>(index) : variant_fn<uint16_t>(index);
        if (fn == nullptr) return cudaErrorInvalidValue;
        cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             static_cast<int>(smem));
        if (e != cudaSuccess) return e;
    }
    return 0;
}

int sv_es_max_ctas_per_sm(int index, size_t smem) {
    int n = 0;
    const void* fn = variant_fn<uint8_t>(index);
    if (fn == nullptr) return 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, fn, kVariants[index].threads, smem) !=
        cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

int sv_es_launch(int index, const SvEsDev& m, const SvSeqsDev& s, float* d_out, float* d_best,
                 int32_t* d_arg, unsigned* d_counter, int grid, size_t smem, cudaStream_t st) {
    const void* fn = s.sym_bytes == 1 ? variant_fn<uint8_t>(index) : variant_fn<uint16_t>(index);
    if (fn == nullptr) return cudaErrorInvalidValue;
    void* args[] = {(void*)&m, (void*)&s, (void*)&d_out, (void*)&d_best, (void*)&d_arg,
                    (void*)&d_counter};
    cudaError_t e = cudaLaunchKernel(fn, dim3(grid), dim3(kVariants[index].threads), args, smem, st);
    return e != cudaSuccess ? e : cudaGetLastError();
}

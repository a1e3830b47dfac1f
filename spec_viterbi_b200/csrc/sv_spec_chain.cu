// K4c -- the level-2 specialised run for "chain + specials" models (every profile HMM the reference ships) in ONE
// persistent launch, O(N) work per chunk instead of O(N^2).
//
// Reference semantics (GraphBLAS_spec_impl.cpp:50-97 with the level-2 handlers of :15-36): after the first symbol
// (v = E[o_0] (+) start), symbols are consumed two at a time, v <- H(a,b) (x) v with the precomputed dense handler
// H(a,b) = M_b (x) M_a, M_o[i][j] = fl(E_o[i] + T'[i][j]); an odd remainder is one level-1 step.  Written out,
//
//     v'[i] = min_k fl( H[i][k] + v[k] ),   H[i][k] = min_j fl( M_b[i][j] + M_a[j][k] ).
//
// x -> fl(x + v[k]) is monotone, so it commutes with the min over j: v'[i] is the min over all two-step paths
// i <- j <- k of the singly rounded sums fl(fl(M_b[i][j] + M_a[j][k]) + v[k]) -- the same sums in the same
// association as table + K4 and K4i evaluate, hence the same bits.  K4i enumerates those paths (5.8 M per chunk
// for 2405.chmm, because every chain row reaches every chain state through the hub).  For a model with the chain
// structure detected by K0 (sv_chain_plan.cpp: chain rows have <= 2 in-edges, predecessor and hub; a special row's
// chain in-edges all carry ONE weight u_s) almost all of them collapse, because x -> fl(c + x) is monotone too:
//
//     min_k fl( fl(M_b[i][hub] + fl(E_a[hub] + u_hub)) + v[k] ) = fl( fl(M_b[i][hub] + fl(E_a[hub] + u_hub)) + min_k v[k] )
//
// bit for bit.  The same argument folds "special row <- chain state p <- hub" through min_p M_a[p][hub], which depends on
// the symbol only and is tabulated by K0.  What is left per chunk: four or five terms per chain row, two per chain
// position for the two special rows, a handful of special <-> special terms, one block-wide reduction of three values
// (chain minimum of v', the two special rows' minima over the chain) -- about 7 N terms, 17 k for 2405.chmm.
//
// One CTA walks ALL chunks of a sequence (and then takes the next sequence from a longest-first queue): a thread
// owns SV_CHAIN_RPT = 5 consecutive chain positions with v in registers; the emission table lives in shared memory
// ([symbol][position], 205 KB for 2405.chmm -- conflict free, the position stride between lanes is 5 words); the new
// scores are published to a double-buffered shared vector for the two neighbours a thread needs (p-1, p-2); one
// __syncthreads per chunk; the symbol stream is read as 8-byte words one word ahead.  No tensor cores ((min,+) is
// not an MMA contraction), no TMEM (the table is read by position, not by sequence).
#include "sv_internal.h"

namespace {

constexpr int R = SV_CHAIN_RPT;
constexpr int NSP = SV_CHAIN_NSP;
static_assert(R == 5 && NSP == 2, "layout");

__device__ __forceinline__ float pinf() { return __int_as_float(0x7f800000); }

// Warp-wide minimum, result in every lane.  NN (every score >= 0, -0 or +inf; checked by K0): the signed-int order
// of the bit patterns is the float order, one REDUX; otherwise five shuffle + min stages.
template <bool NN> __device__ __forceinline__ float warp_min_all(float x) {
    if constexpr (NN) {
        return __int_as_float(__reduce_min_sync(0xffffffffu, __float_as_int(x)));
    } else {
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) x = fminf(x, __shfl_xor_sync(0xffffffffu, x, d));
        return x;
    }
}

template <int TPB, bool NN>
__global__ void __launch_bounds__(TPB, 1)
sv_chain_pair_kernel(SvChainDev m, SvSeqsDev s, float* __restrict__ out, float* __restrict__ best,
                     int32_t* __restrict__ arg, unsigned* __restrict__ counter) {
    constexpr int NW = TPB / 32;
    constexpr int NPOS = TPB * R;
    constexpr int VS = NPOS + 4; // one score buffer: two +inf entries in front (positions -2, -1), two behind
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* sE = reinterpret_cast<float*>(smem_raw);        // [emits][NPOS]
    float* sV = sE + static_cast<size_t>(m.emits) * NPOS;  // [2][VS]
    float* sP = sV + 2 * VS;                               // [2][3][32] per-warp partial minima
    float* sEs = sP + 2 * 3 * 32;                          // [NSP + 1][emits] emissions of the special states, min_p M_o[p][hub]
    __shared__ unsigned s_w;
    __shared__ float s_bv[NW];
    __shared__ int s_bi[NW];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int p0 = tid * R;
    const int emits = m.emits;

    // ---- one-time: emission table -> shared memory, this thread's weights -> registers ----
    {
        const float4* src = reinterpret_cast<const float4*>(m.emis_p);
        float4* dst = reinterpret_cast<float4*>(sE);
        const int n4 = emits * NPOS / 4;
        for (int i = tid; i < n4; i += TPB) dst[i] = src[i];
        for (int i = tid; i < (NSP + 1) * emits; i += TPB) sEs[i] = m.emis_s[i];
        for (int i = tid; i < 2 * 3 * 32; i += TPB) sP[i] = pinf();
        if (tid < 2) {
            sV[tid] = pinf();
            sV[VS + tid] = pinf();
            sV[2 + NPOS + tid] = pinf();
            sV[VS + 2 + NPOS + tid] = pinf();
        }
    }
    float wp[R], wh[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        wp[r] = m.w_pred[p0 + r];
        wh[r] = m.w_hub[p0 + r];
    }
    const float wpm1 = p0 > 0 ? m.w_pred[p0 - 1] : pinf();
    const float whm1 = p0 > 0 ? m.w_hub[p0 - 1] : pinf();
    float u[NSP], ss[NSP][NSP];
#pragma unroll
    for (int a = 0; a < NSP; ++a) {
        u[a] = m.u[a];
#pragma unroll
        for (int b = 0; b < NSP; ++b) ss[a][b] = m.ss[a][b];
    }
    const bool has_s1_to_hub = ss[0][1] < pinf();
    const unsigned long long* __restrict__ sym64 = static_cast<const unsigned long long*>(s.sym);
    __syncthreads();

    for (;;) {
        if (tid == 0) s_w = atomicAdd(counter, 1u);
        __syncthreads();
        const unsigned w = s_w;
        if (w >= static_cast<unsigned>(s.nseq)) break;
        const uint32_t sid = s.order[w];
        const uint64_t b0 = s.off[sid];
        const uint32_t len = static_cast<uint32_t>(s.off[sid + 1] - b0);
        // symbol stream: 8-byte words, one word ahead (the buffer is padded, sv_internal.h)
        const unsigned long long* ptr = sym64 + (b0 >> 3);
        unsigned long long cur = *ptr++, nxt = *ptr++;
        uint32_t pos = static_cast<uint32_t>(b0 & 7u);
        auto next_symbol = [&]() -> int {
            const int sy = static_cast<int>(cur >> ((pos & 7u) * 8u)) & 0xff;
            if ((++pos & 7u) == 0) {
                cur = nxt;
                nxt = *ptr++;
            }
            return sy;
        };

        float v[R];          // scores of the own positions
        float hp[NSP] = {pinf(), pinf()}; // special rows: the terms every thread can form itself (pending the block minimum)
        float h[NSP], cm;    // scores of the special states, chain minimum (valid after phase_begin)
        uint32_t par = 0;    // buffer the LAST phase wrote

        // publishes the new scores and the per-warp partial minima {chain minimum, special row 0, special row 1}
        auto publish = [&](const float (&vn)[R], float part0, float part1) {
            par ^= 1u;
            float* vb = sV + par * VS + 2 + p0;
            float pm = vn[0];
#pragma unroll
            for (int r = 0; r < R; ++r) {
                vb[r] = vn[r];
                v[r] = vn[r];
                pm = fminf(pm, vn[r]);
            }
            pm = warp_min_all<NN>(pm);
            part0 = warp_min_all<NN>(part0);
            part1 = warp_min_all<NN>(part1);
            if (lane == 0) {
                float* pp = sP + par * 96 + warp;
                pp[0] = pm;
                pp[32] = part0;
                pp[64] = part1;
            }
            __syncthreads();
        };
        // after the barrier: every thread finishes the chain minimum and the two special scores of the phase
        auto phase_begin = [&]() {
            const float* pp = sP + par * 96 + lane; // entries >= NW stay +inf
            cm = warp_min_all<NN>(pp[0]);
            h[0] = fminf(hp[0], warp_min_all<NN>(pp[32]));
            h[1] = fminf(hp[1], warp_min_all<NN>(pp[64]));
        };

        // ---- first symbol: v = E[o_0] (+) start (GraphBLAS_spec_impl.cpp:54) ----
        {
            const int o = next_symbol();
            float vn[R];
#pragma unroll
            for (int r = 0; r < R; ++r) vn[r] = sE[o * NPOS + p0 + r] + m.start_p[p0 + r];
#pragma unroll
            for (int a = 0; a < NSP; ++a) hp[a] = sEs[a * emits + o] + m.start_s[a];
            publish(vn, pinf(), pinf());
        }
        uint32_t t = 1;
        // ---- chunks of two symbols (GraphBLAS_spec_impl.cpp:66-81) ----
        while (len - t >= 2u) {
            const int a = next_symbol(), b = next_symbol();
            t += 2u;
            phase_begin();
            const float* vb = sV + par * VS + 2 + p0;
            const float vm1 = vb[-1], vm2 = vb[-2];
            const float* ea = sE + a * NPOS + p0;
            const float* eb = sE + b * NPOS + p0;
            float Pa[R], Ha[R], Pb[R], Hb[R];
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const float xa = ea[r], xb = eb[r];
                Pa[r] = xa + wp[r]; // M_a[p][p-1]
                Ha[r] = xa + wh[r]; // M_a[p][hub]
                Pb[r] = xb + wp[r];
                Hb[r] = xb + wh[r];
            }
            const float xam1 = p0 > 0 ? ea[-1] : 0.f;
            const float Pam1 = xam1 + wpm1, Ham1 = xam1 + whm1;
            float Ua[NSP], Ub[NSP], Sa[NSP][NSP], Sb[NSP][NSP];
#pragma unroll
            for (int x = 0; x < NSP; ++x) {
                const float xa = sEs[x * emits + a], xb = sEs[x * emits + b];
                Ua[x] = xa + u[x]; // M_a[special x][any chain state]
                Ub[x] = xb + u[x];
#pragma unroll
                for (int y = 0; y < NSP; ++y) {
                    Sa[x][y] = xa + ss[x][y]; // M_a[special x][special y]
                    Sb[x][y] = xb + ss[x][y];
                }
            }
            float vn[R], part[NSP] = {pinf(), pinf()};
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const float v1 = r == 0 ? vm1 : v[r - 1];                      // v[p-1]
                const float v2 = r == 0 ? vm2 : r == 1 ? vm1 : v[r - 2];       // v[p-2]
                const float pP = r == 0 ? Pam1 : Pa[r - 1], pH = r == 0 ? Ham1 : Ha[r - 1];
                // chain row p: through its predecessor, through the hub
                float x = (Pb[r] + pP) + v2;
                x = fminf(x, (Pb[r] + pH) + h[0]);
                x = fminf(x, (Hb[r] + Ua[0]) + cm);
                x = fminf(x, (Hb[r] + Sa[0][0]) + h[0]);
                vn[r] = x;
                // special rows through chain state p and its predecessor (through p and the hub: below, once per thread)
#pragma unroll
                for (int y = 0; y < NSP; ++y) part[y] = fminf(part[y], (Ub[y] + Pa[r]) + v1);
            }
            if (has_s1_to_hub) { // chain row <- hub <- special 1: absent in the reference's models (no C -> N edge)
#pragma unroll
                for (int r = 0; r < R; ++r) vn[r] = fminf(vn[r], (Hb[r] + Sa[0][1]) + h[1]);
            }
            // special rows through the special states, and through (chain state p, hub): x -> fl(U + x) and x -> fl(x + h)
            // are monotone, so min_p fl(fl(U_b + M_a[p][hub]) + h_0) = fl(fl(U_b + min_p M_a[p][hub]) + h_0), bit for bit
            const float min_ha = sEs[NSP * emits + a];
#pragma unroll
            for (int x = 0; x < NSP; ++x) {
                float acc = (Ub[x] + min_ha) + h[0];
#pragma unroll
                for (int y = 0; y < NSP; ++y) {
                    acc = fminf(acc, (Sb[x][y] + Ua[y]) + cm);
#pragma unroll
                    for (int z = 0; z < NSP; ++z) acc = fminf(acc, (Sb[x][y] + Sa[y][z]) + h[z]);
                }
                hp[x] = acc;
            }
            publish(vn, part[0], part[1]);
        }
        // ---- odd remainder: one level-1 step (GraphBLAS_spec_impl.cpp:84-89) ----
        if (t < len) {
            const int o = next_symbol();
            phase_begin();
            const float vm1 = sV[par * VS + 2 + p0 - 1];
            const float* eo = sE + o * NPOS + p0;
            float vn[R];
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const float xo = eo[r];
                const float v1 = r == 0 ? vm1 : v[r - 1];
                vn[r] = fminf((xo + wp[r]) + v1, (xo + wh[r]) + h[0]);
            }
#pragma unroll
            for (int x = 0; x < NSP; ++x) {
                const float xo = sEs[x * emits + o];
                float acc = (xo + u[x]) + cm;
#pragma unroll
                for (int y = 0; y < NSP; ++y) acc = fminf(acc, (xo + ss[x][y]) + h[y]);
                hp[x] = acc;
            }
            publish(vn, pinf(), pinf());
        }
        phase_begin();

        // ---- write-out: final vector in original state order, (best, argmin) with the lowest index on ties ----
        float bv = pinf();
        int bi = 0x7fffffff;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int o = m.orig_of_pos[p0 + r];
            if (o < 0) continue;
            if (out != nullptr) out[static_cast<size_t>(sid) * m.n + o] = v[r];
            if (v[r] < bv || (v[r] == bv && o < bi)) {
                bv = v[r];
                bi = o;
            }
        }
        if (tid == 0) {
#pragma unroll
            for (int x = 0; x < NSP; ++x) {
                const int o = m.special[x];
                if (x >= m.nsp || o < 0) continue;
                if (out != nullptr) out[static_cast<size_t>(sid) * m.n + o] = h[x];
                if (h[x] < bv || (h[x] == bv && o < bi)) {
                    bv = h[x];
                    bi = o;
                }
            }
        }
        if (best != nullptr || arg != nullptr) {
            if (!(bv < pinf())) bi = 0x7fffffff; // +inf is "unreachable", never the best state
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
                bv = lane < NW ? s_bv[lane] : pinf();
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
        }
        __syncthreads(); // s_w, s_bv and the partials are reused by the next sequence
    }
}

template <int TPB, bool NN> const void* pair_fn() { return reinterpret_cast<const void*>(sv_chain_pair_kernel<TPB, NN>); }

const void* pair_variant(int threads, bool nonneg) {
    switch (threads) {
    case 128: return nonneg ? pair_fn<128, true>() : pair_fn<128, false>();
    case 256: return nonneg ? pair_fn<256, true>() : pair_fn<256, false>();
    case 512: return nonneg ? pair_fn<512, true>() : pair_fn<512, false>();
    default: return nullptr;
    }
}

} // namespace

size_t sv_chain_pair_smem(int threads, int emits) {
    const size_t npos = static_cast<size_t>(threads) * R;
    return (static_cast<size_t>(emits) * npos + 2 * (npos + 4) + 2 * 3 * 32 + static_cast<size_t>(NSP + 1) * emits) * sizeof(float);
}

int sv_chain_pair_prepare(int threads, bool nonneg, size_t smem, int* ctas_per_sm_out) {
    const void* fn = pair_variant(threads, nonneg);
    if (fn == nullptr) return cudaErrorInvalidValue;
    cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return e;
    int per_sm = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, threads, smem);
    if (e != cudaSuccess) return e;
    if (ctas_per_sm_out != nullptr) *ctas_per_sm_out = per_sm;
    return per_sm >= 1 ? 0 : cudaErrorInvalidConfiguration;
}

int sv_chain_pair_launch(int threads, bool nonneg, const SvChainDev& m, const SvSeqsDev& s, float* d_out, float* d_best,
                         int32_t* d_arg, unsigned* d_counter, int grid, size_t smem, cudaStream_t st) {
    const void* fn = pair_variant(threads, nonneg);
    if (fn == nullptr || s.sym_bytes != 1) return cudaErrorInvalidValue;
    void* args[] = {(void*)&m, (void*)&s, (void*)&d_out, (void*)&d_best, (void*)&d_arg, (void*)&d_counter};
    cudaError_t e = cudaLaunchKernel(fn, dim3(grid), dim3(threads), args, smem, st);
    return e != cudaSuccess ? e : cudaGetLastError();
}

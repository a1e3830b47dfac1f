// Generic persistent (min,+) Viterbi step kernel (any model size, any structure).
//
// One CTA walks ALL time steps of one observation sequence with the score vector resident
// in shared memory (double buffered), then pulls the next sequence from a global work
// queue (longest first).  The model stays in global memory: CSR of T' (row = destination),
// emission table [symbol][state]; on B200 both sit in L1/L2 after the first step.
//
//   v_0[i] = fl(E[o_0][i] + start[i])
//   v_t[i] = min_j fl( fl(E[o_t][i] + T'[i][j]) + v_{t-1}[j] )      (reference:
//            GraphBLAS_impl.cpp:59,64-73 -- the two adds are never re-associated)
//
// Rows are processed one per thread in descending in-degree order; rows with a very large
// in-degree ("heavy": the N and C states of the profile HMMs) are reduced by the whole CTA
// with warp-shuffle mins.  min is exact, so any split of a row is bit-exact.
#include "sv_internal.h"

#include <cfloat>

namespace {

constexpr int GEN_THREADS = 1024;
constexpr int GEN_WARPS = GEN_THREADS / 32;

__device__ __forceinline__ float warp_min(float x) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) x = fminf(x, __shfl_xor_sync(0xffffffffu, x, d));
    return x;
}

template <typename SymT>
__global__ void __launch_bounds__(GEN_THREADS, 1)
sv_plain_generic_kernel(SvGenericDev m, SvSeqsDev s, float* __restrict__ out,
                        float* __restrict__ best, int32_t* __restrict__ arg,
                        unsigned* __restrict__ counter) {
    extern __shared__ float smem[];
    float* cur = smem;
    float* nxt = smem + m.n;
    float* red = smem + 2 * m.n;               // [n_heavy][GEN_WARPS]
    __shared__ unsigned s_work;
    __shared__ float s_bv[GEN_WARPS];
    __shared__ int s_bi[GEN_WARPS];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const SymT* __restrict__ sym = static_cast<const SymT*>(s.sym);
    const float inf = __int_as_float(0x7f800000);

    for (;;) {
        if (tid == 0) s_work = atomicAdd(counter, 1u);
        __syncthreads();
        const unsigned work = s_work;
        __syncthreads();
        if (work >= static_cast<unsigned>(s.nseq)) break;
        const uint32_t sid = s.order[work];
        const uint64_t o0 = s.off[sid];
        const uint64_t len = s.off[sid + 1] - o0;

        // first symbol folded with the start vector
        {
            const float* __restrict__ em = m.emis + static_cast<size_t>(sym[o0]) * m.n;
            for (int i = tid; i < m.n; i += GEN_THREADS) cur[i] = em[i] + m.start[i];
        }
        __syncthreads();

        unsigned o_next = len > 1 ? sym[o0 + 1] : 0;
        for (uint64_t t = 1; t < len; ++t) {
            const unsigned o = o_next;
            if (t + 1 < len) o_next = sym[o0 + t + 1]; // prefetch: off the critical path
            const float* __restrict__ em = m.emis + static_cast<size_t>(o) * m.n;

            for (int r = tid; r < m.n_light; r += GEN_THREADS) {
                const uint32_t i = m.light_rows[r];
                const float e = em[i];
                float acc = inf;
                const uint32_t p1 = m.rowptr[i + 1];
                for (uint32_t p = m.rowptr[i]; p < p1; ++p) {
                    const float u = e + m.w[p];          // fl(E + T')
                    acc = fminf(acc, u + cur[m.col[p]]); // fl(. + v), exact min
                }
                nxt[i] = acc;
            }
            for (int h = 0; h < m.n_heavy; ++h) {
                const uint32_t i = m.heavy_rows[h];
                const float e = em[i];
                float acc = inf;
                const uint32_t p1 = m.rowptr[i + 1];
                for (uint32_t p = m.rowptr[i] + tid; p < p1; p += GEN_THREADS) {
                    const float u = e + m.w[p];
                    acc = fminf(acc, u + cur[m.col[p]]);
                }
                acc = warp_min(acc);
                if (lane == 0) red[h * GEN_WARPS + warp] = acc;
            }
            if (m.n_heavy > 0) {
                __syncthreads();
                for (int h = warp; h < m.n_heavy; h += GEN_WARPS) {
                    const float x = warp_min(red[h * GEN_WARPS + lane]);
                    if (lane == 0) nxt[m.heavy_rows[h]] = x;
                }
            }
            __syncthreads();
            float* sw = cur;
            cur = nxt;
            nxt = sw;
        }

        // results: full vector and/or (min score, lowest argmin)
        if (out != nullptr) {
            float* __restrict__ o = out + static_cast<size_t>(sid) * m.n;
            for (int i = tid; i < m.n; i += GEN_THREADS) o[i] = cur[i];
        }
        if (best != nullptr || arg != nullptr) {
            float bv = inf;
            int bi = 0x7fffffff;
            for (int i = tid; i < m.n; i += GEN_THREADS) {
                const float x = cur[i];
                if (x < bv) {
                    bv = x;
                    bi = i;
                }
            }
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
                bv = s_bv[lane];
                bi = s_bi[lane];
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
                    if (arg != nullptr) arg[sid] = (bi == 0x7fffffff) ? -1 : bi;
                }
            }
        }
        __syncthreads();
    }
}

} // namespace

size_t sv_generic_smem(const SvGenericDev& m) {
    return (2 * static_cast<size_t>(m.n) + static_cast<size_t>(m.n_heavy) * GEN_WARPS) *
           sizeof(float);
}

int sv_generic_launch(const SvGenericDev& m, const SvSeqsDev& s, float* d_out, float* d_best,
                      int32_t* d_arg, unsigned* d_counter, int grid, cudaStream_t st) {
    const size_t smem = sv_generic_smem(m);
    cudaError_t e;
    if (s.sym_bytes == 1) {
        e = cudaFuncSetAttribute(sv_plain_generic_kernel<uint8_t>,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        sv_plain_generic_kernel<uint8_t>
            <<<grid, GEN_THREADS, smem, st>>>(m, s, d_out, d_best, d_arg, d_counter);
    } else {
        e = cudaFuncSetAttribute(sv_plain_generic_kernel<uint16_t>,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        sv_plain_generic_kernel<uint16_t>
            <<<grid, GEN_THREADS, smem, st>>>(m, s, d_out, d_best, d_arg, d_counter);
    }
    return cudaGetLastError();
}

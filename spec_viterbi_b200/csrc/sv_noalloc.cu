// sv_viterbi_device_static: the plain (min,+) Viterbi run for callers that must not allocate device memory
// (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
//
// No counterpart in the reference.
//
// It is a variant of the generic persistent kernel (sv_plain_generic.cu): same arithmetic, same association
//   v_t[i] = min_j fl( fl(E[o_t][i] + T'[i][j]) + v_{t-1}[j] )       (GraphBLAS_impl.cpp:59,64-73)
// so its results are the bits of every other plain kernel here.  What differs is where the scratch lives:
// the two score vectors of every CTA and the work-queue counter sit in ONE file-scope device array of a
// fixed size instead of in memory obtained per model / per launch.  Consequences, all part of the contract
// (include/spec_viterbi_b200.h):
//   * limits: states <= SV_STATIC_MAX_STATES; larger models are refused with SV_ERR_ARG;
//   * every GPU that loads the library holds the array (SV_STATIC_SCRATCH_BYTES, about 4.6 MiB) until exit;
//   * all calls on one device share the array, which keeps the previous call's contents: the launch clears
//     the queue counter itself (a memset node on the caller's stream) and every score a step reads was
//     written earlier in the same call, so nothing else has to start at zero; two calls that overlap on
//     the GPU (different streams, no ordering between them) race on it -- order them.
//
// In host code the array's name is only a host-side stand-in for the device symbol, not a device address:
// the address comes from cudaGetSymbolAddress, once per device, with that device current, and is what
// the kernel is given.
#include "sv_internal.h"

#include <mutex>

namespace {

constexpr int SA_THREADS = 1024;
constexpr int SA_WARPS = SA_THREADS / 32;
constexpr int SA_MAX_STATES = 4096; // == SV_STATIC_MAX_STATES (checked below)
constexpr int SA_MAX_CTAS = 148;    // one persistent CTA per SM of a B200
constexpr int SA_MAX_HEAVY = 32;    // CTA-wide rows whose warp partials fit static shared memory
constexpr size_t SA_HEADER_FLOATS = 64; // 256 bytes: [0] = work-queue counter
constexpr size_t SA_SCRATCH_FLOATS = SA_HEADER_FLOATS + size_t(SA_MAX_CTAS) * 2 * SA_MAX_STATES;

static_assert(SA_MAX_STATES == SV_STATIC_MAX_STATES, "header and kernel disagree on the state limit");
static_assert(SA_SCRATCH_FLOATS * sizeof(float) == SV_STATIC_SCRATCH_BYTES, "header and kernel disagree on the size");

// The scratch itself.  Device code may name it; host code must not use the name as a pointer.
static __device__ __align__(256) float g_static_scratch[SA_SCRATCH_FLOATS];

__device__ __forceinline__ float warp_min_f(float x) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) x = fminf(x, __shfl_xor_sync(0xffffffffu, x, d));
    return x;
}

// `scratch` is the device address of g_static_scratch obtained by the host (cudaGetSymbolAddress).
// The score vectors are read and written by the same CTA only, with a barrier between a step's writes and
// the next step's reads, so plain (coherent) global accesses are enough; they must not go through the
// read-only path, hence no const / __restrict__ on them.
template <typename SymT>
__global__ void __launch_bounds__(SA_THREADS, 1)
sv_plain_static_kernel(SvGenericDev m, SvSeqsDev s, float* __restrict__ out, float* __restrict__ best,
                       int32_t* __restrict__ arg, float* scratch) {
    __shared__ unsigned s_work;
    __shared__ float s_red[SA_MAX_HEAVY * SA_WARPS];
    __shared__ float s_bv[SA_WARPS];
    __shared__ int s_bi[SA_WARPS];

    unsigned* counter = reinterpret_cast<unsigned*>(scratch);
    float* cur = scratch + SA_HEADER_FLOATS + static_cast<size_t>(blockIdx.x) * 2 * SA_MAX_STATES;
    float* nxt = cur + SA_MAX_STATES;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const SymT* __restrict__ sym = static_cast<const SymT*>(s.sym);
    const float inf = __int_as_float(0x7f800000);
    // rows beyond the CTA-wide budget are done one per thread like the light rows (min is exact: same bits)
    const int n_heavy = min(m.n_heavy, SA_MAX_HEAVY);

    for (;;) {
        if (tid == 0) s_work = atomicAdd(counter, 1u);
        __syncthreads();
        const unsigned work = s_work;
        __syncthreads();
        if (work >= static_cast<unsigned>(s.nseq)) break;
        const uint32_t sid = s.order[work];
        const uint64_t o0 = s.off[sid];
        const uint64_t len = s.off[sid + 1] - o0;

        {
            const float* __restrict__ em = m.emis + static_cast<size_t>(sym[o0]) * m.n;
            for (int i = tid; i < m.n; i += SA_THREADS) cur[i] = em[i] + m.start[i];
        }
        __syncthreads();

        unsigned o_next = len > 1 ? sym[o0 + 1] : 0;
        for (uint64_t t = 1; t < len; ++t) {
            const unsigned o = o_next;
            if (t + 1 < len) o_next = sym[o0 + t + 1];
            const float* __restrict__ em = m.emis + static_cast<size_t>(o) * m.n;

            for (int r = tid; r < m.n_light + (m.n_heavy - n_heavy); r += SA_THREADS) {
                const uint32_t i = r < m.n_light ? m.light_rows[r] : m.heavy_rows[n_heavy + (r - m.n_light)];
                const float e = em[i];
                float acc = inf;
                const uint32_t p1 = m.rowptr[i + 1];
                for (uint32_t p = m.rowptr[i]; p < p1; ++p) {
                    const float u = e + m.w[p];          // fl(E + T')
                    acc = fminf(acc, u + cur[m.col[p]]); // fl(. + v), exact min
                }
                nxt[i] = acc;
            }
            for (int h = 0; h < n_heavy; ++h) {
                const uint32_t i = m.heavy_rows[h];
                const float e = em[i];
                float acc = inf;
                const uint32_t p1 = m.rowptr[i + 1];
                for (uint32_t p = m.rowptr[i] + tid; p < p1; p += SA_THREADS) {
                    const float u = e + m.w[p];
                    acc = fminf(acc, u + cur[m.col[p]]);
                }
                acc = warp_min_f(acc);
                if (lane == 0) s_red[h * SA_WARPS + warp] = acc;
            }
            if (n_heavy > 0) {
                __syncthreads();
                for (int h = warp; h < n_heavy; h += SA_WARPS) {
                    const float x = warp_min_f(s_red[h * SA_WARPS + lane]);
                    if (lane == 0) nxt[m.heavy_rows[h]] = x;
                }
            }
            __syncthreads(); // orders this step's global writes before the next step's reads (same CTA)
            float* sw = cur;
            cur = nxt;
            nxt = sw;
        }

        if (out != nullptr) {
            float* __restrict__ o = out + static_cast<size_t>(sid) * m.n;
            for (int i = tid; i < m.n; i += SA_THREADS) o[i] = cur[i];
        }
        if (best != nullptr || arg != nullptr) {
            float bv = inf;
            int bi = 0x7fffffff;
            for (int i = tid; i < m.n; i += SA_THREADS) {
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

// Device address of the scratch, looked up once per device with that device current.
constexpr int SA_MAX_DEVICES = 64;
std::mutex g_addr_mutex;
float* g_scratch_addr[SA_MAX_DEVICES] = {};

int scratch_address(int device, float** out) {
    if (device < 0 || device >= SA_MAX_DEVICES) return sv_fail(SV_ERR_ARG, "sv_viterbi_device_static: device ordinal out of range");
    std::lock_guard<std::mutex> lock(g_addr_mutex);
    if (g_scratch_addr[device] == nullptr) {
        SV_CUDA(cudaSetDevice(device));
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess || p == nullptr)
            return sv_fail(SV_ERR_CUDA, std::string("sv_viterbi_device_static: cudaGetSymbolAddress: ") +
                                            cudaGetErrorString(e));
        g_scratch_addr[device] = static_cast<float*>(p);
    }
    *out = g_scratch_addr[device];
    return SV_OK;
}

} // namespace

extern "C" int sv_viterbi_device_static(const sv_model* m, const sv_seqs* seqs, float* d_out, float* d_best,
                                        int32_t* d_best_state, void* stream) {
    SV_API_BEGIN
    if (m == nullptr || seqs == nullptr) return sv_fail(SV_ERR_ARG, "sv_viterbi_device_static: NULL handle");
    if (seqs->device != m->device) return sv_fail(SV_ERR_ARG, "sv_viterbi_device_static: sequences live on another device");
    if (m->n > static_cast<uint64_t>(SA_MAX_STATES))
        return sv_fail(SV_ERR_ARG, "sv_viterbi_device_static: " + std::to_string(m->n) + " states exceed the fixed scratch (SV_STATIC_MAX_STATES = " +
                                       std::to_string(SA_MAX_STATES) + "); use sv_viterbi_device");
    if (seqs->nseq == 0) return sv_fail(SV_ERR_ARG, "sv_viterbi_device_static: empty batch");
    float* scratch = nullptr;
    const int rc = scratch_address(m->device, &scratch);
    if (rc != SV_OK) return rc;
    SV_CUDA(cudaSetDevice(m->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    // the queue counter is the only word that must start at zero (the rest of the array is written before it is read)
    SV_CUDA(cudaMemsetAsync(scratch, 0, SA_HEADER_FLOATS * sizeof(float), st));
    const int grid = static_cast<int>(std::min<uint64_t>(seqs->nseq, std::min(m->sm_count, SA_MAX_CTAS)));
    if (seqs->dev.sym_bytes == 1)
        sv_plain_static_kernel<uint8_t><<<grid, SA_THREADS, 0, st>>>(m->gen, seqs->dev, d_out, d_best, d_best_state, scratch);
    else
        sv_plain_static_kernel<uint16_t><<<grid, SA_THREADS, 0, st>>>(m->gen, seqs->dev, d_out, d_best, d_best_state, scratch);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return sv_fail(SV_ERR_CUDA, cudaGetErrorString(e));
    return SV_OK;
    SV_API_END
}

// FP32 pipe micro-benchmark: fixes the non-tensor roofline used for the (min,+) kernels.
// Measures lane-instructions/s of FADD, FMNMX, FMNMX3 and the step kernel's instruction
// mixes on every SM, plus the SM clock seen during the run (clock64 vs globaltimer).
#include "sv_internal.h"

namespace {

constexpr int PROBE_THREADS = 512;
constexpr int PROBE_ILP = 8;

__device__ __forceinline__ float min3(float a, float b, float c) {
    float d;
    asm("min.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}

template <int MIX>
__global__ void __launch_bounds__(PROBE_THREADS)
sv_probe_kernel(int iters, float seed, float* sink, unsigned long long* clk) {
    __shared__ float sh[PROBE_THREADS * 2];
    float a[PROBE_ILP], b[PROBE_ILP];
#pragma unroll
    for (int k = 0; k < PROBE_ILP; ++k) {
        a[k] = seed + k + threadIdx.x;
        b[k] = seed * 0.5f + k;
    }
    sh[threadIdx.x] = seed;
    sh[threadIdx.x + PROBE_THREADS] = seed + 1.0f;
    __syncthreads();
    unsigned long long c0 = 0, g0 = 0;
    if (threadIdx.x == 0) {
        c0 = clock64();
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g0));
    }
    const float c1 = seed * 1e-3f, c2 = seed * 2e-3f;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < PROBE_ILP; ++k) {
            if (MIX == 0) { // 2 FADD
                a[k] = a[k] + c1;
                b[k] = b[k] + c2;
            } else if (MIX == 1) { // 2 FMNMX
                a[k] = fminf(a[k], b[k] );
                b[k] = fmaxf(b[k], a[k]);
            } else if (MIX == 2) { // 2 FMNMX3
                a[k] = min3(a[k], b[k], c1);
                b[k] = min3(b[k], a[k], c2);
            } else if (MIX == 3) { // 2 FADD + 1 FMNMX3  (tabulated (min,+) MAC pair)
                const float x = a[k] + c1;
                const float y = a[k] + c2;
                b[k] = min3(b[k], x, y);
                a[k] = x;
            } else if (MIX == 4) { // 4 FADD + 1 FMNMX3  (oracle-order relaxation pair)
                const float u0 = b[k] + c1;
                const float u1 = b[k] + c2;
                const float x = u0 + a[k];
                const float y = u1 + a[k];
                a[k] = min3(a[k], x, y);
            } else { // LDS + 2 FADD + FMNMX (the gather-bound inner loop)
                const float v = sh[(threadIdx.x + k * 32 + it) & (2 * PROBE_THREADS - 1)];
                const float u = b[k] + c1;
                a[k] = fminf(a[k], u + v);
            }
        }
    }
    float r = 0.f;
#pragma unroll
    for (int k = 0; k < PROBE_ILP; ++k) r += a[k] + b[k];
    if (r == 12345.678f) sink[0] = r;
    if (threadIdx.x == 0) {
        unsigned long long g1;
        const unsigned long long c1k = clock64();
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g1));
        if (blockIdx.x == 0) {
            clk[0] = c1k - c0;
            clk[1] = g1 - g0;
        }
    }
}

template <int MIX>
int run_probe(int grid, int iters, float* d_sink, unsigned long long* d_clk, float* ms) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    sv_probe_kernel<MIX><<<grid, PROBE_THREADS>>>(iters / 8 + 1, 1.0f, d_sink, d_clk); // warm-up
    cudaEventRecord(e0);
    sv_probe_kernel<MIX><<<grid, PROBE_THREADS>>>(iters, 1.0f, d_sink, d_clk);
    cudaEventRecord(e1);
    cudaError_t e = cudaEventSynchronize(e1);
    cudaEventElapsedTime(ms, e0, e1);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    return e == cudaSuccess ? cudaGetLastError() : e;
}

} // namespace

extern "C" int sv_probe_fp32(int device, int mix, int iters, double* lane_instr_per_s,
                             double* sm_clock_mhz) {
    if (mix < 0 || mix > 5 || iters < 1 || lane_instr_per_s == nullptr)
        return sv_fail(SV_ERR_ARG, "sv_probe_fp32: bad argument");
    SV_CUDA(cudaSetDevice(device));
    int sms = 0;
    SV_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    float* d_sink = nullptr;
    unsigned long long* d_clk = nullptr;
    SV_CUDA(cudaMalloc(&d_sink, sizeof(float)));
    SV_CUDA(cudaMalloc(&d_clk, 2 * sizeof(unsigned long long)));
    const int grid = sms * 4;
    float ms = 0.f;
    int rc = 0;
    static const int per_iter[6] = {2, 2, 2, 3, 5, 4}; // instructions per ILP lane per iteration
    switch (mix) {
    case 0: rc = run_probe<0>(grid, iters, d_sink, d_clk, &ms); break;
    case 1: rc = run_probe<1>(grid, iters, d_sink, d_clk, &ms); break;
    case 2: rc = run_probe<2>(grid, iters, d_sink, d_clk, &ms); break;
    case 3: rc = run_probe<3>(grid, iters, d_sink, d_clk, &ms); break;
    case 4: rc = run_probe<4>(grid, iters, d_sink, d_clk, &ms); break;
    default: rc = run_probe<5>(grid, iters, d_sink, d_clk, &ms); break;
    }
    unsigned long long clk[2] = {0, 1};
    cudaMemcpy(clk, d_clk, sizeof(clk), cudaMemcpyDeviceToHost);
    cudaFree(d_sink);
    cudaFree(d_clk);
    if (rc != 0) return sv_fail(SV_ERR_CUDA, cudaGetErrorString((cudaError_t)rc));
    const double lanes = double(grid) * PROBE_THREADS * double(iters) * PROBE_ILP * per_iter[mix];
    *lane_instr_per_s = lanes / (double(ms) * 1e-3);
    if (sm_clock_mhz != nullptr) *sm_clock_mhz = clk[1] ? double(clk[0]) / double(clk[1]) * 1e3 : 0;
    return SV_OK;
}

// ------------------------------------------------------------------ TMEM probe ----
// Checks that tensor memory can serve as a per-SM lookup table for the emission scores
// (tcgen05.alloc / st / ld with a warp-uniform, run-time column) and measures the read rate.
namespace {

__device__ __forceinline__ void tmem_alloc_cols(uint32_t* smem_slot, uint32_t ncols) {
    const uint32_t dst = static_cast<uint32_t>(__cvta_generic_to_shared(smem_slot));
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}

__global__ void __launch_bounds__(512, 1)
sv_probe_tmem_kernel(int iters, unsigned* errors, unsigned long long* clk, float* sink) {
    __shared__ uint32_t s_taddr;
    const int tid = threadIdx.x, warp = tid >> 5;
    if (warp == 0) tmem_alloc_cols(&s_taddr, 512);
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tbase = s_taddr;
    const uint32_t quad = static_cast<uint32_t>(warp & 3) * 32u;   // TMEM lanes this warp may touch
    const uint32_t grp = static_cast<uint32_t>(warp >> 2);          // 4 warps share a quadrant
    const uint32_t my = tbase + (quad << 16) + grp * 128u;          // 128 columns per warp
    // fill: value encodes (tid, column)
    for (uint32_t c = 0; c < 128; ++c) {
        const uint32_t v = (static_cast<uint32_t>(tid) << 8) | c;
        asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(my + c), "r"(v) : "memory");
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    __syncthreads();
    // check with run-time columns, x1 and x4 forms
    unsigned bad = 0;
    for (uint32_t c = 0; c < 128; ++c) {
        uint32_t r;
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(r) : "r"(my + c) : "memory");
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        bad += r != ((static_cast<uint32_t>(tid) << 8) | c);
    }
    for (uint32_t c = 0; c + 4 <= 128; c += 3) {
        uint32_t r0, r1, r2, r3;
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];"
                     : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(my + c) : "memory");
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        const uint32_t e = (static_cast<uint32_t>(tid) << 8) | c;
        bad += (r0 != e) + (r1 != e + 1) + (r2 != e + 2) + (r3 != e + 3);
    }
    if (bad) atomicAdd(errors, bad);
    __syncthreads();
    // throughput: every warp streams x4 loads with a run-time column (like a symbol lookup)
    unsigned long long c0 = clock64();
    uint32_t acc = 0, col = 0;
    for (int it = 0; it < iters; ++it) {
        uint32_t r0, r1, r2, r3;
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];"
                     : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(my + col) : "memory");
        col = (col + 4) & 124u;
        uint32_t q0, q1, q2, q3;
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];"
                     : "=r"(q0), "=r"(q1), "=r"(q2), "=r"(q3) : "r"(my + col) : "memory");
        col = (col + 4) & 124u;
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        acc += r0 ^ r1 ^ r2 ^ r3 ^ q0 ^ q1 ^ q2 ^ q3;
    }
    unsigned long long c1 = clock64();
    if (acc == 0x12345u) sink[0] = 1.f;
    if (blockIdx.x == 0 && tid == 0) clk[0] = c1 - c0;
    __syncthreads();
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tbase), "r"(512u) : "memory");
}

} // namespace

extern "C" int sv_probe_tmem(int device, int iters, unsigned* errors, double* bytes_per_clk_per_sm) {
    if (errors == nullptr || bytes_per_clk_per_sm == nullptr || iters < 1)
        return sv_fail(SV_ERR_ARG, "sv_probe_tmem: bad argument");
    SV_CUDA(cudaSetDevice(device));
    int sms = 0;
    SV_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    unsigned* d_err = nullptr;
    unsigned long long* d_clk = nullptr;
    float* d_sink = nullptr;
    SV_CUDA(cudaMalloc(&d_err, sizeof(unsigned)));
    SV_CUDA(cudaMalloc(&d_clk, sizeof(unsigned long long)));
    SV_CUDA(cudaMalloc(&d_sink, sizeof(float)));
    SV_CUDA(cudaMemset(d_err, 0, sizeof(unsigned)));
    sv_probe_tmem_kernel<<<sms, 512>>>(iters, d_err, d_clk, d_sink);
    cudaError_t e = cudaDeviceSynchronize();
    unsigned long long clk = 1;
    cudaMemcpy(errors, d_err, sizeof(unsigned), cudaMemcpyDeviceToHost);
    cudaMemcpy(&clk, d_clk, sizeof(clk), cudaMemcpyDeviceToHost);
    cudaFree(d_err);
    cudaFree(d_clk);
    cudaFree(d_sink);
    if (e != cudaSuccess) return sv_fail(SV_ERR_CUDA, std::string("sv_probe_tmem: ") + cudaGetErrorString(e));
    // per iteration each of the 16 warps reads 2 x (32 lanes x 4 columns x 4 B)
    *bytes_per_clk_per_sm = double(iters) * 16.0 * 2.0 * 512.0 / double(clk ? clk : 1);
    return SV_OK;
}

// Micro-benchmarks that fix the rooflines of the (min,+) kernels -- built as a SEPARATE library
// (libsvprobe_b200.so, include/sv_probe_b200.h); nothing of the product links or loads it.
//   sv_probe_fp32: lane-instructions/s of FADD, FMNMX, FMNMX3 and the step kernels' instruction mixes
//                  on every SM, plus the SM clock seen during the run (clock64 vs globaltimer);
//   sv_probe_tmem: tensor memory as a per-SM lookup table (tcgen05.ld with a run-time column);
//   sv_probe_smem: the shared-memory pipe under the lockstep kernel's access pattern: conflict-free
//                  LDS.128 gathers at per-lane register addresses, alone and mixed with STS.128.
#include "sv_probe_b200.h"

#include <cuda_runtime.h>

#include <cstdint>
#include <string>

namespace {
thread_local std::string g_probe_error;
int sv_fail(int code, const std::string& msg) {
    g_probe_error = msg;
    return code;
}
enum { SV_OK = 0, SV_ERR_ARG = 1, SV_ERR_CUDA = 2 };
#define SV_CUDA(call)                                                                            \
    do {                                                                                         \
        cudaError_t sv_e_ = (call);                                                              \
        if (sv_e_ != cudaSuccess) {                                                              \
            return sv_fail(SV_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(sv_e_)); \
        }                                                                                        \
    } while (0)
} // namespace

extern "C" const char* sv_probe_last_error(void) { return g_probe_error.c_str(); }

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
            } else if (MIX == 4) { // 4 FADD + 1 FMNMX3  (oracle-order relaxation pair); b[k] is loop carried
                const float u0 = b[k] + c1;  // (round 1 had b[k] loop-invariant: u0, u1 were hoisted and the
                const float u1 = b[k] + c2;  //  row over-counted two adds per iteration)
                const float x = u0 + a[k];
                const float y = u1 + a[k];
                a[k] = min3(a[k], x, y);
                b[k] = u1;
            } else if (MIX == 6) { // 2 FADD2 (packed add.rn.f32x2: two IEEE adds per instruction, sm_100+)
                unsigned long long pa, pb, pc;
                asm("mov.b64 %0, {%1, %2};" : "=l"(pa) : "f"(a[k]), "f"(b[k]));
                asm("mov.b64 %0, {%1, %2};" : "=l"(pc) : "f"(c1), "f"(c2));
                asm("add.rn.f32x2 %0, %1, %2;" : "=l"(pb) : "l"(pa), "l"(pc));
                asm("add.rn.f32x2 %0, %1, %2;" : "=l"(pa) : "l"(pb), "l"(pc));
                asm("mov.b64 {%0, %1}, %2;" : "=f"(a[k]), "=f"(b[k]) : "l"(pa));
            } else if (MIX == 7) { // 2 FADD2 + 1 FMNMX3: the oracle-order relaxation pair with packed adds
                unsigned long long pa, pb, pc;
                asm("mov.b64 %0, {%1, %2};" : "=l"(pa) : "f"(b[k]), "f"(b[k]));
                asm("mov.b64 %0, {%1, %2};" : "=l"(pc) : "f"(c1), "f"(c2));
                asm("add.rn.f32x2 %0, %1, %2;" : "=l"(pb) : "l"(pa), "l"(pc));
                asm("mov.b64 %0, {%1, %2};" : "=l"(pc) : "f"(a[k]), "f"(a[k]));
                asm("add.rn.f32x2 %0, %1, %2;" : "=l"(pa) : "l"(pb), "l"(pc));
                float x, y;
                asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(pa));
                a[k] = min3(a[k], x, y);
                b[k] = y;
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
    if (mix < 0 || mix > 7 || iters < 1 || lane_instr_per_s == nullptr)
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
    // arithmetic operations per ILP lane per iteration (a packed FADD2 counts as its two adds)
    static const int per_iter[8] = {2, 2, 2, 3, 5, 4, 4, 5};
    switch (mix) {
    case 0: rc = run_probe<0>(grid, iters, d_sink, d_clk, &ms); break;
    case 1: rc = run_probe<1>(grid, iters, d_sink, d_clk, &ms); break;
    case 2: rc = run_probe<2>(grid, iters, d_sink, d_clk, &ms); break;
    case 3: rc = run_probe<3>(grid, iters, d_sink, d_clk, &ms); break;
    case 4: rc = run_probe<4>(grid, iters, d_sink, d_clk, &ms); break;
    case 6: rc = run_probe<6>(grid, iters, d_sink, d_clk, &ms); break;
    case 7: rc = run_probe<7>(grid, iters, d_sink, d_clk, &ms); break;
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

// ------------------------------------------------------------------ shared-memory probe ----
// The lockstep kernel's binding resource is the shared-memory pipe: per relaxation one lane reads 16 B
// (v[src] of 4 sequences) at a register address that K0 made bank-conflict free (the 8 lanes of an
// LDS.128 phase hit 8 distinct 16-byte bank groups), and per state update one lane stores 16 B.  This
// probe issues exactly those instructions, with 16 warps per SM and 8 loads in flight per thread, and
// reports the bytes per clock and SM the pipe sustains -- the MEASURED ceiling bench.py uses.
//   mode 0: LDS.128 gathers only            mode 1: 3 LDS.128 : 1 STS.128 (the kernel's ratio, nnz/N = 3)
//   mode 2: mode 1 plus the relaxations (8 FADD + 2 FMNMX3 per load), i.e. the whole inner loop
//   mode 3: LDS.32 conflict-free (128 B per warp instruction) for reference
namespace {

constexpr int SM_THREADS = 512;
constexpr int SM_POS = 2560;  // float4 positions (40 KB), as in the kernel: 5 ranks x 512 threads
constexpr int SM_SLOTS = 8;

template <int MODE>
__global__ void __launch_bounds__(SM_THREADS, 1)
sv_probe_smem_kernel(int iters, unsigned long long* clk, float* sink) {
    extern __shared__ __align__(16) unsigned char sm_raw[];
    float4* vbuf = reinterpret_cast<float4*>(sm_raw);
    const int tid = threadIdx.x, lane = tid & 31;
    for (int i = tid; i < 2 * SM_POS; i += SM_THREADS) vbuf[i] = make_float4(1.f + i, 2.f, 3.f, 4.f);
    // per-thread gather addresses: pseudo-random position whose bank group (position mod 8) is
    // (lane + slot) mod 8 -> the 8 lanes of every LDS.128 phase hit 8 distinct bank groups
    const uint32_t sbase = static_cast<uint32_t>(__cvta_generic_to_shared(sm_raw));
    uint32_t addr[SM_SLOTS];
#pragma unroll
    for (int k = 0; k < SM_SLOTS; ++k) {
        uint32_t h = static_cast<uint32_t>(tid) * 2654435761u + static_cast<uint32_t>(k) * 40503u + 12345u;
        h ^= h >> 15;
        h *= 2246822519u;
        h ^= h >> 13;
        const uint32_t pos = ((h % (SM_POS / 8)) * 8u) + ((static_cast<uint32_t>(lane) + k) & 7u);
        addr[k] = sbase + (MODE == 3 ? ((h % (SM_POS * 4 / 32)) * 32u + ((lane + k) & 31u)) * 4u : pos * 16u);
    }
    const uint32_t own = sbase + static_cast<uint32_t>(tid) * 16u;
    __syncthreads();
    float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f, acc3 = 0.f;
    const float w0 = 0.25f + lane, w1 = 0.5f + lane;
    const unsigned long long c0 = clock64();
    for (int it = 0; it < iters; ++it) {
        const uint32_t pofs = (it & 1) ? SM_POS * 16u : 0u; // double buffer parity, as in the kernel
        if (MODE == 3) {
#pragma unroll
            for (int k = 0; k < SM_SLOTS; ++k) {
                float x;
                asm volatile("ld.shared.f32 %0, [%1];" : "=f"(x) : "r"(addr[k] + pofs) : "memory");
                acc0 += x;
            }
            continue;
        }
        float4 g[SM_SLOTS];
#pragma unroll
        for (int k = 0; k < SM_SLOTS; ++k)
            asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];"
                         : "=f"(g[k].x), "=f"(g[k].y), "=f"(g[k].z), "=f"(g[k].w)
                         : "r"(addr[k] + pofs)
                         : "memory");
        if (MODE == 2) {
            // the "emission" operands are last iteration's results: loop carried across iterations (nothing
            // can be hoisted) but independent inside one, like the rows of a step
            const float p0 = acc0, p1 = acc1, p2 = acc2, p3 = acc3;
#pragma unroll
            for (int k = 0; k < SM_SLOTS; k += 2) { // a slot pair: 16 FADD + 4 FMNMX3, as pair4<false>
                acc0 = fminf(acc0, fminf((p0 + w0) + g[k].x, (p0 + w1) + g[k + 1].x));
                acc1 = fminf(acc1, fminf((p1 + w0) + g[k].y, (p1 + w1) + g[k + 1].y));
                acc2 = fminf(acc2, fminf((p2 + w0) + g[k].z, (p2 + w1) + g[k + 1].z));
                acc3 = fminf(acc3, fminf((p3 + w0) + g[k].w, (p3 + w1) + g[k + 1].w));
            }
        } else {
#pragma unroll
            for (int k = 0; k < SM_SLOTS; ++k) {
                acc0 += g[k].x;
                acc1 += g[k].y;
                acc2 += g[k].z;
                acc3 += g[k].w;
            }
        }
        if (MODE == 1 || MODE == 2) { // 8 loads : 8/3 stores -> 3 stores on two iterations of three, 2 on the third
            const int ns = (it % 3 == 2) ? 2 : 3;
            for (int k = 0; k < ns; ++k)
                asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(own + (pofs ^ (SM_POS * 16u)) + k * SM_THREADS * 16u),
                             "f"(acc0), "f"(acc1), "f"(acc2), "f"(acc3)
                             : "memory");
        }
    }
    __syncthreads(); // the CTA's time, not warp 0's: without a barrier the warps drift apart (oldest-first issue)
    const unsigned long long c1 = clock64();
    if (acc0 + acc1 + acc2 + acc3 == 12345.678f) sink[0] = acc0;
    if (tid == 0) clk[blockIdx.x] = c1 - c0;
}

} // namespace

extern "C" int sv_probe_smem(int device, int mode, int iters, double* bytes_per_clk_per_sm, double* loads_per_clk_per_sm) {
    // modes 1 and 2 only: their load / store counts were checked against ncu's wavefront counters
    // (profiles/r02_smem_probe_ncu.csv: 0.98 wavefronts per clock and SM); the loads-only loops (0, 3) are
    // restructured by ptxas and are not reported
    if ((mode != 1 && mode != 2) || iters < 3 || bytes_per_clk_per_sm == nullptr)
        return sv_fail(SV_ERR_ARG, "sv_probe_smem: bad argument (mode 1 or 2)");
    SV_CUDA(cudaSetDevice(device));
    int sms = 0;
    SV_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    unsigned long long* d_clk = nullptr;
    float* d_sink = nullptr;
    SV_CUDA(cudaMalloc(&d_clk, sms * sizeof(unsigned long long)));
    SV_CUDA(cudaMalloc(&d_sink, sizeof(float)));
    const size_t smem = 2 * SM_POS * 16;
    const void* fn = mode == 0 ? (const void*)sv_probe_smem_kernel<0> : mode == 1 ? (const void*)sv_probe_smem_kernel<1>
                     : mode == 2 ? (const void*)sv_probe_smem_kernel<2> : (const void*)sv_probe_smem_kernel<3>;
    SV_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    iters = iters / 3 * 3;
    cudaError_t e = cudaSuccess;
    for (int rep = 0; rep < 2 && e == cudaSuccess; ++rep) { // first launch warms up
        void* args[] = {(void*)&iters, (void*)&d_clk, (void*)&d_sink};
        e = cudaLaunchKernel(fn, dim3(sms), dim3(SM_THREADS), args, smem, nullptr);
        if (e == cudaSuccess) e = cudaDeviceSynchronize();
    }
    std::string err = e == cudaSuccess ? "" : cudaGetErrorString(e);
    unsigned long long clk = 1;
    {
        unsigned long long* h = new unsigned long long[sms];
        cudaMemcpy(h, d_clk, sms * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
        clk = 0;
        for (int i = 0; i < sms; ++i) clk = h[i] > clk ? h[i] : clk; // slowest SM
        delete[] h;
    }
    cudaFree(d_clk);
    cudaFree(d_sink);
    if (e != cudaSuccess) return sv_fail(SV_ERR_CUDA, "sv_probe_smem: " + err);
    const double loads = double(iters) * SM_SLOTS * SM_THREADS;              // lane loads per SM
    const double stores = (mode == 1 || mode == 2) ? double(iters) * (8.0 / 3.0) * SM_THREADS : 0.0;
    const double unit = mode == 3 ? 4.0 : 16.0;
    *bytes_per_clk_per_sm = (loads + stores) * unit / double(clk ? clk : 1);
    if (loads_per_clk_per_sm != nullptr) *loads_per_clk_per_sm = loads / double(clk ? clk : 1);
    return SV_OK;
}

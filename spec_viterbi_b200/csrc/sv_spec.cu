// Specialised path: precompute of the per-tuple handlers (K3 dense (min,+) GEMM, K3' sparse
// left factor) and the level-L run phase (K4 handler x vector, level-1 tail steps).
//
// Reference: GraphBLAS_spec_impl.cpp:103-182 (initializer), :15-36 (add_level:
// handler[key ++ i] = M_i (x) handler[key], sparse level-1 matrix on the LEFT), :50-97
// (run: first symbol folded with the start vector, chunks of `level` symbols from index 1,
// tail with level-1 matrices).  Handlers are stored dense (+inf = no entry): for the
// profile HMMs they are ~99.7% dense.  Every value is a min of singly rounded sums in the
// reference's association, so each level is bit-exact against the same-level oracle.
//
// Level 1 needs no tables: the plain kernels form fl(E + T') on the fly with identical bits.
#include "sv_internal.h"

#include <cooperative_groups.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <mutex>
#include <new>
#include <string>
#include <type_traits>
#include <vector>

struct sv_spec {
    sv_model* m = nullptr;
    uint64_t level = 1;
    uint64_t ntuples = 1;     // emits^level
    uint64_t t_begin = 0, t_end = 0; // valid tuple range of the table
    float* d_table = nullptr; // [(t_end-t_begin)][n][n]
    uint64_t table_first = 0; // tuple index stored at d_table[0]
    bool owns_table = false;
    // Level 2 of a chain + specials model in AUTO mode runs without the table (K4c), so the table is built by the first
    // call that needs it (SV_SPEC_RUN_TABLE, sv_spec_table, sv_spec_get_handler, sv_spec_save) instead of by
    // sv_spec_create: spec_with costs a model upload, and 4 N^2 emits^2 bytes (9.27 GB for 2405.chmm) stay free.
    bool deferred = false;
    void* stream_at_create = nullptr;
    int method = SV_SPEC_SPARSE_LHS;
    int run_mode = SV_SPEC_RUN_AUTO;  // AUTO: K4i (level-2 chunks recomputed from the sparse factors) when eligible, else K4
    cudaEvent_t ev_done = nullptr;    // end of the last run on this spec: runs share d_v and serialise on it
    // on-demand mode (level >= 3): the complete level-(L-1) table and a scratch of per-sequence level-L handlers
    bool ondemand = false;
    float* d_prev = nullptr;       // [emits^(L-1)][n][n]
    float* d_scratch = nullptr;    // [scratch_cap][n][n]
    size_t scratch_cap = 0;        // handlers
    // run-phase scratch
    float* d_v[2] = {nullptr, nullptr};
    uint64_t v_capacity = 0;
    // K4 runs groups of sequences on side streams so that the partial last wave of one group's
    // launch is filled by the next group's (sequences are independent)
    static constexpr int kRunStreams = 4;
    cudaStream_t aux[kRunStreams - 1] = {nullptr, nullptr, nullptr};
    cudaEvent_t ev_fork = nullptr, ev_join[kRunStreams - 1] = {nullptr, nullptr, nullptr};
};

namespace cg = cooperative_groups;

namespace {

__device__ __forceinline__ float pos_inf() { return __int_as_float(0x7f800000); }

__device__ __forceinline__ float warp_min(float x) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) x = fminf(x, __shfl_xor_sync(0xffffffffu, x, d));
    return x;
}

// ---------------------------------------------------------------- level 1 dense ----
// dst[o][i][j] = +inf, then fl(E[o][i] + T'[i][j]) on the pattern (ld = leading dimension).
__global__ void fill_inf_kernel(float* __restrict__ p, size_t count) {
    const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    for (size_t k = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; k < count; k += stride)
        p[k] = pos_inf();
}

__global__ void scatter_level1_kernel(SvGenericDev m, float* __restrict__ dst,
                                      float* __restrict__ dst_t, size_t ld, size_t mat_stride,
                                      int sym_begin, int sym_count) {
    // one thread per (row, symbol): walks the row's edges
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int o = sym_begin + blockIdx.y;
    if (i >= m.n || static_cast<int>(blockIdx.y) >= sym_count) return;
    const float e = m.emis[static_cast<size_t>(o) * m.n + i];
    float* row = dst + static_cast<size_t>(blockIdx.y) * mat_stride + static_cast<size_t>(i) * ld;
    float* colT = dst_t ? dst_t + static_cast<size_t>(blockIdx.y) * mat_stride + i : nullptr;
    for (uint32_t p = m.rowptr[i]; p < m.rowptr[i + 1]; ++p) {
        const float v = e + m.w[p];
        row[m.col[p]] = v;
        if (colT) colT[static_cast<size_t>(m.col[p]) * ld] = v; // transposed copy [j][i]
    }
}

// ------------------------------------------------------------ K3': sparse left factor ----
// out[t][i][k] = min_{j in row i of T'} fl( fl(E[sym(t)][i] + T'[i][j]) + prev[prefix(t)][j][k] )
// grid: x = column tile (blockDim.x * 4 columns), y = row i, z = tuple (relative)
__global__ void __launch_bounds__(256)
sparse_lhs_kernel(SvGenericDev m, const float* __restrict__ prev, uint64_t prev_first,
                  float* __restrict__ out, uint64_t tuple_first, int emits) {
    const int n = m.n;
    const int i = blockIdx.y;
    const uint64_t tuple = tuple_first + blockIdx.z;
    const int symb = static_cast<int>(tuple % emits);
    const uint64_t prefix = tuple / emits;
    const float* __restrict__ P = prev + (prefix - prev_first) * static_cast<size_t>(n) * n;
    float* __restrict__ O = out + static_cast<size_t>(blockIdx.z) * n * n + static_cast<size_t>(i) * n;
    const float e = m.emis[static_cast<size_t>(symb) * n + i];
    const uint32_t p0 = m.rowptr[i], p1 = m.rowptr[i + 1];
    for (int k0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4; k0 < n; k0 += gridDim.x * blockDim.x * 4) {
        float acc[4] = {pos_inf(), pos_inf(), pos_inf(), pos_inf()};
        for (uint32_t p = p0; p < p1; ++p) {
            const float a = e + m.w[p]; // level-1 value M_sym[i][j]
            const float* __restrict__ prow = P + static_cast<size_t>(m.col[p]) * n + k0;
#pragma unroll
            for (int c = 0; c < 4; ++c)
                if (k0 + c < n) acc[c] = fminf(acc[c], a + prow[c]);
        }
#pragma unroll
        for (int c = 0; c < 4; ++c)
            if (k0 + c < n) O[k0 + c] = acc[c];
    }
}

// The same product for ONE tuple per sequence -- the tuple sequence q observes in chunk `chunk` -- into a scratch
// handler per sequence (on-demand mode of levels >= 3: "recompute on demand per observed tuple").
// prev = the COMPLETE level-(L-1) table.  grid: x = column tile, y = row i, z = sequence (relative to seq_base).
struct OnDemandSeqs {
    int nseq;
    const void* sym;
    int sym_bytes;
    const uint64_t* off;
};
__global__ void __launch_bounds__(256)
ondemand_handler_kernel(SvGenericDev m, const float* __restrict__ prev, float* __restrict__ scratch, OnDemandSeqs s,
                        uint64_t chunk, int level, int emits, int seq_base) {
    const int n = m.n;
    const int i = blockIdx.y;
    const int q = seq_base + blockIdx.z;
    const uint64_t o0 = s.off[q];
    const uint64_t len = s.off[q + 1] - o0;
    if (chunk >= (len - 1) / level) return; // no chunk left: K4 copies the vector through
    uint64_t prefix = 0;
    unsigned symb = 0;
    for (int j = 0; j < level; ++j) {
        const uint64_t idx = o0 + 1 + chunk * level + j;
        const unsigned o = s.sym_bytes == 1 ? static_cast<const uint8_t*>(s.sym)[idx] : static_cast<const uint16_t*>(s.sym)[idx];
        if (j + 1 < level) prefix = prefix * emits + o;
        else symb = o; // the last consumed symbol is the left (sparse) factor: H(k1..kL) = M_kL (x) H(k1..k(L-1))
    }
    const float* __restrict__ P = prev + prefix * static_cast<size_t>(n) * n;
    float* __restrict__ O = scratch + static_cast<size_t>(blockIdx.z) * n * n + static_cast<size_t>(i) * n;
    const float e = m.emis[static_cast<size_t>(symb) * n + i];
    const uint32_t p0 = m.rowptr[i], p1 = m.rowptr[i + 1];
    for (int k0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4; k0 < n; k0 += gridDim.x * blockDim.x * 4) {
        float acc[4] = {pos_inf(), pos_inf(), pos_inf(), pos_inf()};
        for (uint32_t p = p0; p < p1; ++p) {
            const float a = e + m.w[p]; // level-1 value M_sym[i][j]
            const float* __restrict__ prow = P + static_cast<size_t>(m.col[p]) * n + k0;
#pragma unroll
            for (int c = 0; c < 4; ++c)
                if (k0 + c < n) acc[c] = fminf(acc[c], a + prow[c]);
        }
#pragma unroll
        for (int c = 0; c < 4; ++c)
            if (k0 + c < n) O[k0 + c] = acc[c];
    }
}

// K3' organised by PREFIX: one CTA owns (row i, prefix) and produces that row for all `emits` tuples
// that extend the prefix, so every P[prefix][j][k] is loaded once and used `emits` times
// (the kernel above re-reads it per tuple with a 16-byte lane stride: 4 L1 wavefronts per load and
// four partial stores per line).  One column per thread per iteration: loads and stores are 128
// contiguous bytes per warp.  Rows with more than 256 in-edges (the generic kernel's heavy list) are
// split into 256-column tiles on a separate launch.  Same association => same bits.
// EB = tuples of one prefix kept in registers at a time, KU = columns per thread (256 apart, so that a
// CTA writes KU KB of every tuple's row back to back)
#ifndef SV_LHS_EB
#define SV_LHS_EB 10
#define SV_LHS_KU 4
#endif
template <int EB, int KU>
__global__ void __launch_bounds__(256, 3)
sparse_lhs_prefix_kernel(SvGenericDev m, const float* __restrict__ prev, uint64_t prev_first,
                         float* __restrict__ out, uint64_t out_first, uint64_t tb, uint64_t te,
                         uint64_t prefix_first, int emits, int heavy_pass) {
    const int n = m.n, tid = threadIdx.x;
    int i, kbeg, kend;
    uint64_t prefix;
    if (heavy_pass) {
        i = static_cast<int>(m.heavy_rows[blockIdx.y]);
        kbeg = blockIdx.x * 256 * KU;
        kend = min(n, kbeg + 256 * KU);
        prefix = prefix_first + blockIdx.z;
    } else {
        i = blockIdx.x;
        kbeg = 0;
        kend = n;
        prefix = prefix_first + blockIdx.y;
    }
    const uint32_t p0 = m.rowptr[i], p1 = m.rowptr[i + 1];
    if (!heavy_pass && p1 - p0 > 256u) return; // same rule as the heavy list (sv_api.cu upload_generic)
    const uint64_t t_lo = max(tb, prefix * emits), t_hi = min(te, (prefix + 1) * emits);
    if (t_lo >= t_hi) return;
    const size_t nn = static_cast<size_t>(n) * n;
    const float* __restrict__ P = prev + (prefix - prev_first) * nn;
    for (uint64_t t0 = t_lo; t0 < t_hi; t0 += EB) {
        const int nb = static_cast<int>(min(static_cast<uint64_t>(EB), t_hi - t0));
        float e[EB];
#pragma unroll
        for (int b = 0; b < EB; ++b) e[b] = m.emis[static_cast<size_t>((t0 + min(b, nb - 1)) % emits) * n + i];
        float* __restrict__ O = out + (t0 - out_first) * nn + static_cast<size_t>(i) * n;
        for (int k0 = kbeg + tid; k0 < kend; k0 += 256 * KU) {
            float acc[EB][KU];
#pragma unroll
            for (int b = 0; b < EB; ++b)
#pragma unroll
                for (int u = 0; u < KU; ++u) acc[b][u] = pos_inf();
#pragma unroll 2
            for (uint32_t p = p0; p < p1; ++p) {
                const float w = m.w[p];
                const float* __restrict__ prow = P + static_cast<size_t>(m.col[p]) * n + k0;
                float pv[KU];
#pragma unroll
                for (int u = 0; u < KU; ++u) pv[u] = k0 + 256 * u < kend ? prow[256 * u] : pos_inf();
#pragma unroll
                for (int b = 0; b < EB; ++b) {
                    const float a = e[b] + w; // level-1 value M_sym[i][j]
#pragma unroll
                    for (int u = 0; u < KU; ++u) acc[b][u] = fminf(acc[b][u], a + pv[u]);
                }
            }
#pragma unroll
            for (int b = 0; b < EB; ++b)
                if (b < nb) {
#pragma unroll
                    for (int u = 0; u < KU; ++u)
                        if (k0 + 256 * u < kend) O[b * nn + k0 + 256 * u] = acc[b][u];
                }
        }
    }
}

// ------------------------------------------------------ K3'': level-2 pair rows, sparse x sparse ----
// Level 2 only: BOTH factors of H(a,b) = M_b (x) M_a have the sparsity of T', so one output row
//   out[t][i][k] = min_{j in in(i)} fl( fl(E[b][i] + T'[i][j]) + fl(E[a][j] + T'[j][k]) )
// needs no dense operand at all: the kernel is a pure streaming WRITE of the dense table
// (E^2 * N^2 * 4 bytes) with a handful of terms per row.  Rows of T' with more than kPairHeavy
// entries ("heavy": the dense rows of profile HMMs) are kept as dense vectors Wd[h][.]:
//   * light row i: dense terms (heavy j) are evaluated per column; the few (k, value) pairs that the
//     light j contribute are merged into the streamed 16-byte stores through a column bitmap with
//     prefix popcounts (fast path: <= 64 pairs, <= 2 dense terms; the tuple-dependent scalars of a whole
//     chunk of tuples are tabulated in shared memory first, so the stream has no dependent load), or
//     patched in after a CTA barrier (general path);
//   * heavy row i: column k walks its CSC list {j : k in in(j)} and looks T'[i][j] up in Wd.
// Same association as K3' => same bits.  grid: x = row i, y = tuple (relative); 256 threads.
constexpr int kPairHeavy = 32;    // rows with more in-edges are "heavy"
constexpr int kPairMaxHeavy = 64; // more heavy rows than this: use K3'
constexpr int kPairLongCol = 64;  // CSC columns longer than this are reduced by the whole CTA
constexpr int kPairMaxPairs = kPairHeavy * kPairHeavy;

struct SvPairDev {
    int n = 0, emits = 0, n_longcol = 0;
    const uint32_t* rowptr = nullptr; // CSR of T' (row = destination)
    const uint32_t* col = nullptr;
    const float* w = nullptr;
    const float* emis = nullptr;      // [emits][n]
    const int32_t* hidx = nullptr;    // [n] index of the row's dense copy, -1 = light row
    const float* wd = nullptr;        // [heavy][n] dense copies, +inf where absent
    const uint32_t* cscptr = nullptr; // CSC of T' (column = source)
    const uint32_t* cscrow = nullptr;
    const float* cscw = nullptr;
    const uint32_t* longcol = nullptr; // columns with more than kPairLongCol entries
    const uint32_t* hrow = nullptr;    // [heavy] row of each dense copy
    int n_heavy = 0;
    // K4i/K4p only (path_prepare): two-step paths i <- j <- k through LIGHT j, flattened so that a thread
    // meets no chain of dependent loads: {T'[i][j], T'[j][k], j, k}
    const uint4* lpath = nullptr;     // [path_slots][n] paths of light row i (slot-major: coalesced)
    const uint32_t* lcount = nullptr; // [n] paths of row i (0 for dense rows)
    const float* wcol = nullptr;      // [heavy][n] T'[i][hrow[h]], +inf where absent (dense column copies)
    const uint4* hpath = nullptr;     // paths of the dense rows, concatenated
    const uint32_t* hpath_ptr = nullptr; // [heavy + 1]
    int path_slots = 0;
    // r_h[k] = fl(E[a][dense row h] + T'[h][k]) tabulated per symbol: rh[(a * n_stage + s) * npad + k], +inf padded to
    // npad (a multiple of 16 floats), so that a CTA stages a row with ONE 1-D bulk copy (TMA) instead of recomputing it
    const float* rh = nullptr;
    int npad = 0;
    int n_stage = 0;                 // dense rows that are a SOURCE of some row: only these are staged as r_h
    int stage_h[4] = {0, 0, 0, 0};   // their dense-copy indices
    int stage_row = 0; // light pass: (n + 8) floats of dynamic shared memory hold the row's first dense source
};

#ifndef SV_PAIR_CHUNK
#define SV_PAIR_CHUNK 40
#endif
constexpr int kPairChunk = SV_PAIR_CHUNK; // consecutive tuples per CTA: the row's structure is set up once
constexpr int kPairFast = 64;  // pairs per row whose per-tuple values are tabulated in shared memory
constexpr int kPairBitWords = 1024; // column bitmap of the fast path: up to 32 * 1024 - 64 states

__global__ void __launch_bounds__(256)
pair_rows_kernel(SvPairDev d, float* __restrict__ out, uint64_t tuple_first, uint64_t tuple_count, int heavy_pass) {
    // light pass: x = row, y = chunk of kPairChunk tuples (heavy rows exit); heavy pass: x = heavy row
    // index, y = one tuple (their per-tuple work is ~100x a light row's: spread it over the grid)
    const int n = d.n, tid = threadIdx.x;
    const int i = heavy_pass ? static_cast<int>(d.hrow[blockIdx.x]) : static_cast<int>(blockIdx.x);
    const int chunk = heavy_pass ? 1 : kPairChunk;
    const uint64_t rel0 = static_cast<uint64_t>(blockIdx.y) * chunk;
    const int ntup = static_cast<int>(min(static_cast<uint64_t>(chunk), tuple_count - rel0));
    const int hi = d.hidx[i];
    __shared__ float s_red[8];
    if ((hi >= 0) != (heavy_pass != 0)) return;
    if (hi >= 0) { // ---- heavy row: per column, over the column's CSC list ----
        const float* __restrict__ Wi = d.wd + static_cast<size_t>(hi) * n;
        for (int s = 0; s < ntup; ++s) {
            const uint64_t tuple = tuple_first + rel0 + s;
            const int b = static_cast<int>(tuple % d.emits), a = static_cast<int>(tuple / d.emits);
            float* __restrict__ O = out + (rel0 + s) * static_cast<size_t>(n) * n + static_cast<size_t>(i) * n;
            const float eb = d.emis[static_cast<size_t>(b) * n + i];
            const float* __restrict__ Ea = d.emis + static_cast<size_t>(a) * n;
            for (int k = tid; k < n; k += 256) {
                const uint32_t p0 = d.cscptr[k], p1 = d.cscptr[k + 1];
                float acc = pos_inf();
                if (p1 - p0 <= static_cast<uint32_t>(kPairLongCol)) {
                    for (uint32_t p = p0; p < p1; ++p) {
                        const uint32_t j = d.cscrow[p];
                        acc = fminf(acc, (eb + Wi[j]) + (Ea[j] + d.cscw[p]));
                    }
                }
                O[k] = acc;
            }
            for (int c = 0; c < d.n_longcol; ++c) { // long columns: CTA-wide reduction, thread 0 overwrites
                const uint32_t k = d.longcol[c];
                float acc = pos_inf();
                for (uint32_t p = d.cscptr[k] + tid; p < d.cscptr[k + 1]; p += 256) {
                    const uint32_t j = d.cscrow[p];
                    acc = fminf(acc, (eb + Wi[j]) + (Ea[j] + d.cscw[p]));
                }
                acc = warp_min(acc);
                __syncthreads(); // orders the O[k] stores above / the previous round's s_red reads
                if ((tid & 31) == 0) s_red[tid >> 5] = acc;
                __syncthreads();
                if (tid == 0) {
                    float x = s_red[0];
#pragma unroll
                    for (int q = 1; q < 8; ++q) x = fminf(x, s_red[q]);
                    O[k] = x;
                }
            }
        }
        return;
    }
    // ---- light row: structure once, values per tuple ----
    __shared__ float s_w[kPairHeavy];       // T'[i][j_q]
    __shared__ uint32_t s_j[kPairHeavy];    // j_q
    __shared__ int s_h[kPairHeavy];         // dense copy of row j_q, -1 = light
    __shared__ uint32_t s_off[kPairHeavy + 1];
    __shared__ uint32_t s_pk[kPairMaxPairs]; // pairs contributed by the light j: column,
    __shared__ float s_pw[kPairMaxPairs];    //   T'[j][k],
    __shared__ uint8_t s_pq[kPairMaxPairs];  //   and which q they belong to
    __shared__ int s_nheavy;
    __shared__ uint8_t s_hq[kPairHeavy];     // the q with heavy j
    const uint32_t r0 = d.rowptr[i];
    const int di = static_cast<int>(d.rowptr[i + 1] - r0);
    if (tid < di) {
        const uint32_t j = d.col[r0 + tid];
        s_w[tid] = d.w[r0 + tid];
        s_j[tid] = j;
        s_h[tid] = d.hidx[j];
    }
    __syncthreads();
    if (tid == 0) {
        uint32_t o = 0;
        int nhv = 0;
        for (int q = 0; q < di; ++q) {
            s_off[q] = o;
            if (s_h[q] < 0)
                o += d.rowptr[s_j[q] + 1] - d.rowptr[s_j[q]];
            else
                s_hq[nhv++] = static_cast<uint8_t>(q);
        }
        s_off[di] = o;
        s_nheavy = nhv;
    }
    __syncthreads();
    const int npairs = static_cast<int>(s_off[di]);
    const int nhv = s_nheavy;
    if (tid < di && s_h[tid] < 0) {
        const uint32_t q0 = d.rowptr[s_j[tid]], q1 = d.rowptr[s_j[tid] + 1];
        for (uint32_t p = q0; p < q1; ++p) {
            s_pk[s_off[tid] + (p - q0)] = d.col[p];
            s_pw[s_off[tid] + (p - q0)] = d.w[p];
            s_pq[s_off[tid] + (p - q0)] = static_cast<uint8_t>(tid);
        }
    }
    __syncthreads();
    // ---- fast path (few pairs, at most two dense terms): everything a tuple needs from global memory --
    //      E_b[i], E_a[j_q] -- is tabulated for the whole chunk up front, so that the tuple loop is a pure
    //      store stream with no dependent load at the head of every tuple (that latency, paid kPairChunk
    //      times per CTA and again inside every patch, held the first version at a third of the HBM
    //      write rate).  Patches are merged into the streamed stores through a column bitmap. ----
    if (npairs <= kPairFast && nhv <= 2 && n <= 32 * kPairBitWords - 64) {
        __shared__ float s_val[kPairChunk][kPairFast]; // per tuple: min over the pairs of the u-th patched column
        __shared__ float s_hu[kPairChunk][2], s_hea[kPairChunk][2];
        __shared__ uint32_t s_bits[kPairBitWords];     // columns that receive a patch
        __shared__ uint8_t s_pref[kPairBitWords];      // patched columns below 32 * w
        __shared__ uint8_t s_first[kPairFast], s_u[kPairFast];
        const int nwords = (n >> 5) + 2;
        for (int w = tid; w < nwords; w += 256) s_bits[w] = 0u;
        if (tid < npairs) { // first occurrence of its column?
            bool first = true;
            for (int t2 = 0; t2 < tid; ++t2) first = first && s_pk[t2] != s_pk[tid];
            s_first[tid] = first;
        }
        __syncthreads();
        if (tid < npairs) {
            atomicOr(&s_bits[s_pk[tid] >> 5], 1u << (s_pk[tid] & 31u));
            int u = 0; // rank of this pair's column among the patched columns
            for (int t2 = 0; t2 < npairs; ++t2) u += (s_first[t2] && s_pk[t2] < s_pk[tid]) ? 1 : 0;
            s_u[tid] = static_cast<uint8_t>(u);
        }
        for (int w = tid; w < nwords; w += 256) {
            int c = 0;
            for (int t2 = 0; t2 < npairs; ++t2) c += (s_first[t2] && s_pk[t2] < 32u * w) ? 1 : 0;
            s_pref[w] = static_cast<uint8_t>(c);
        }
        __syncthreads();
        for (int idx = tid; idx < ntup * npairs; idx += 256) {
            const int s = idx / npairs, t = idx - s * npairs;
            if (!s_first[t]) continue;
            const uint64_t tuple = tuple_first + rel0 + s;
            const size_t b = tuple % d.emits, a = tuple / d.emits;
            const float eb = d.emis[b * n + i];
            float v = pos_inf();
            for (int t2 = t; t2 < npairs; ++t2)
                if (s_pk[t2] == s_pk[t]) {
                    const int q = s_pq[t2];
                    v = fminf(v, (eb + s_w[q]) + (d.emis[a * n + s_j[q]] + s_pw[t2]));
                }
            s_val[s][s_u[t]] = v;
        }
        if (tid < 2 * ntup && (tid & 1) < nhv) {
            const int s = tid >> 1, x = tid & 1, q = s_hq[x];
            const uint64_t tuple = tuple_first + rel0 + s;
            const size_t b = tuple % d.emits, a = tuple / d.emits;
            s_hu[s][x] = d.emis[b * n + i] + s_w[q];
            s_hea[s][x] = d.emis[a * n + s_j[q]];
        }
        const float* __restrict__ hp0 = nhv > 0 ? d.wd + static_cast<size_t>(s_h[s_hq[0]]) * n : d.wd;
        const float* __restrict__ hp1 = nhv > 1 ? d.wd + static_cast<size_t>(s_h[s_hq[1]]) * n : d.wd;
        // one dense source (every light row of a profile HMM): its dense copy is staged in shared memory
        // once per chunk, so that the stream reads it with two aligned LDS.128 per 16-byte store instead
        // of four scalar global loads with 64-bit addresses (the stream is issue-bound)
        extern __shared__ __align__(16) unsigned char s_dyn[];
        float* s_row = reinterpret_cast<float*>(s_dyn);
        const bool staged = d.stage_row != 0 && nhv == 1;
        if (staged)
            for (int x = tid; x < n + 8; x += 256) s_row[x] = x < n ? hp0[x] : pos_inf();
        __syncthreads();
        // (the per-tuple preamble is kept short -- the row pointer advances by n * n, nothing is re-derived
        //  from the tuple index: at ~20 instructions per 512-byte store the preamble was 40 % of all issue slots)
        float* __restrict__ O = out + rel0 * static_cast<size_t>(n) * n + static_cast<size_t>(i) * n;
        const size_t nn = static_cast<size_t>(n) * n;
        for (int s = 0; s < ntup; ++s, O += nn) {
            const float hu0 = s_hu[s][0], hea0 = s_hea[s][0], hu1 = s_hu[s][1], hea1 = s_hea[s][1];
            const float* __restrict__ val = s_val[s];
            const int kb = min(n, static_cast<int>((32u - (static_cast<uint32_t>(reinterpret_cast<uintptr_t>(O) >> 2) & 31u)) & 31u));
            const int nb = static_cast<int>(static_cast<uint32_t>(n - kb) >> 2);
            auto rank_of = [&](int k) { return s_pref[k >> 5] + __popc(s_bits[k >> 5] & ((1u << (k & 31)) - 1u)); };
            auto fix1 = [&](int k, float v) {
                if ((s_bits[k >> 5] >> (k & 31)) & 1u) v = fminf(v, val[rank_of(k)]);
                return v;
            };
            auto stream = [&](auto f) {
                if (tid < kb) O[tid] = fix1(tid, f(tid));
                // column k = kb + 4 * (tid + 256 * it): the bit offset inside a bitmap word is the same in
                // every iteration, the word index advances by 32
                const int k0 = kb + 4 * tid, sh = k0 & 31;
                const uint32_t* __restrict__ bw = s_bits + (k0 >> 5);
                const uint32_t* const bw_end = s_bits + ((kb + 4 * nb) >> 5) + ((((kb + 4 * nb) & 31) > sh) ? 1 : 0);
                float4* __restrict__ dst = reinterpret_cast<float4*>(O + k0);
                for (; bw < bw_end; bw += 32, dst += 256) { // the word pointer is the only loop counter
                    const int k = (static_cast<int>(bw - s_bits) << 5) + sh;
                    float v[4] = {f(k), f(k + 1), f(k + 2), f(k + 3)};
                    const uint32_t bits = __funnelshift_r(bw[0], bw[1], sh) & 15u;
                    if (bits) { // the patched columns of a row are ranked left to right
                        int u = rank_of(k);
                        if (bits & 1u) v[0] = fminf(v[0], val[u++]);
                        if (bits & 2u) v[1] = fminf(v[1], val[u++]);
                        if (bits & 4u) v[2] = fminf(v[2], val[u++]);
                        if (bits & 8u) v[3] = fminf(v[3], val[u]);
                    }
                    *dst = make_float4(v[0], v[1], v[2], v[3]);
                }
                for (int kk = kb + 4 * nb + tid; kk < n; kk += 256) O[kk] = fix1(kk, f(kk));
            };
            // same, the dense term read from the staged row: columns k..k+3 straddle two aligned float4;
            // the shift r = k & 3 is the same for every group of the tuple (compile-time in the body)
            auto stream_staged = [&](auto rc) {
                constexpr int r = decltype(rc)::value;
                auto f = [&](int k) { return hu0 + (hea0 + s_row[k]); };
                if (tid < kb) O[tid] = fix1(tid, f(tid));
                const int k0 = kb + 4 * tid, sh = k0 & 31;
                const uint32_t* __restrict__ bw = s_bits + (k0 >> 5);
                const uint32_t* const bw_end = s_bits + ((kb + 4 * nb) >> 5) + ((((kb + 4 * nb) & 31) > sh) ? 1 : 0);
                const float4* __restrict__ rp = reinterpret_cast<const float4*>(s_row) + (k0 >> 2);
                float4* __restrict__ dst = reinterpret_cast<float4*>(O + k0);
                for (; bw < bw_end; bw += 32, dst += 256, rp += 256) {
                    const int k = (static_cast<int>(bw - s_bits) << 5) + sh;
                    const float4 a = rp[0];
                    float x[4];
                    if constexpr (r == 0) {
                        x[0] = a.x; x[1] = a.y; x[2] = a.z; x[3] = a.w;
                    } else {
                        const float4 b = rp[1];
                        if constexpr (r == 1) { x[0] = a.y; x[1] = a.z; x[2] = a.w; x[3] = b.x; }
                        if constexpr (r == 2) { x[0] = a.z; x[1] = a.w; x[2] = b.x; x[3] = b.y; }
                        if constexpr (r == 3) { x[0] = a.w; x[1] = b.x; x[2] = b.y; x[3] = b.z; }
                    }
                    float v[4];
#pragma unroll
                    for (int c = 0; c < 4; ++c) v[c] = hu0 + (hea0 + x[c]);
                    const uint32_t bits = __funnelshift_r(bw[0], bw[1], sh) & 15u;
                    if (bits) {
                        int u = rank_of(k);
                        if (bits & 1u) v[0] = fminf(v[0], val[u++]);
                        if (bits & 2u) v[1] = fminf(v[1], val[u++]);
                        if (bits & 4u) v[2] = fminf(v[2], val[u++]);
                        if (bits & 8u) v[3] = fminf(v[3], val[u]);
                    }
                    *dst = make_float4(v[0], v[1], v[2], v[3]);
                }
                for (int kk = kb + 4 * nb + tid; kk < n; kk += 256) O[kk] = fix1(kk, f(kk));
            };
            if (nhv == 0)
                stream([&](int) { return pos_inf(); });
            else if (staged) {
                switch (kb & 3) {
                case 0: stream_staged(std::integral_constant<int, 0>{}); break;
                case 1: stream_staged(std::integral_constant<int, 1>{}); break;
                case 2: stream_staged(std::integral_constant<int, 2>{}); break;
                default: stream_staged(std::integral_constant<int, 3>{}); break;
                }
            } else if (nhv == 1)
                stream([&](int k) { return hu0 + (hea0 + hp0[k]); });
            else
                stream([&](int k) { return fminf(hu0 + (hea0 + hp0[k]), hu1 + (hea1 + hp1[k])); });
        }
        return;
    }
    for (int s = 0; s < ntup; ++s) {
        const uint64_t tuple = tuple_first + rel0 + s;
        const int b = static_cast<int>(tuple % d.emits), a = static_cast<int>(tuple / d.emits);
        float* __restrict__ O = out + (rel0 + s) * static_cast<size_t>(n) * n + static_cast<size_t>(i) * n;
        const float eb = d.emis[static_cast<size_t>(b) * n + i];
        const float* __restrict__ Ea = d.emis + static_cast<size_t>(a) * n;
        // dense terms (heavy j): min_q fl( fl(E_b[i] + T'[i][j_q]) + fl(E_a[j_q] + T'[j_q][k]) ); the per-tuple
        // invariants of the first two heavy j live in registers
        float hu0 = pos_inf(), hu1 = pos_inf(), hea0 = 0.f, hea1 = 0.f;
        const float* __restrict__ hp0 = d.wd;
        const float* __restrict__ hp1 = d.wd;
        if (nhv > 0) {
            const int q = s_hq[0];
            hu0 = eb + s_w[q];
            hea0 = Ea[s_j[q]];
            hp0 = d.wd + static_cast<size_t>(s_h[q]) * n;
        }
        if (nhv > 1) {
            const int q = s_hq[1];
            hu1 = eb + s_w[q];
            hea1 = Ea[s_j[q]];
            hp1 = d.wd + static_cast<size_t>(s_h[q]) * n;
        }
        auto dense_at = [&](int k) {
            float acc = pos_inf();
            if (nhv > 0) acc = hu0 + (hea0 + hp0[k]);
            if (nhv > 1) acc = fminf(acc, hu1 + (hea1 + hp1[k]));
            for (int x = 2; x < nhv; ++x) {
                const int q = s_hq[x];
                acc = fminf(acc, (eb + s_w[q]) + (Ea[s_j[q]] + d.wd[static_cast<size_t>(s_h[q]) * n + k]));
            }
            return acc;
        };
        // the row as 16-byte stores where its (odd-length) placement allows, scalar head and tail; the
        // streaming loop is specialised on the number of dense terms (0, 1, 2, more)
        // (the vector body starts on a 128-byte line, so that every warp store writes whole lines: partially
        //  written lines cost a read-modify-write in the memory system)
        const int kb = min(n, static_cast<int>((32u - ((reinterpret_cast<uintptr_t>(O) >> 2) & 31u)) & 31u));
        const int nb = (n - kb) / 4;
        auto stream = [&](auto f) {
            if (tid < kb) O[tid] = f(tid);
            for (int g = tid; g < nb; g += 256) {
                const int k = kb + 4 * g;
                *reinterpret_cast<float4*>(O + k) = make_float4(f(k), f(k + 1), f(k + 2), f(k + 3));
            }
            for (int k = kb + 4 * nb + tid; k < n; k += 256) O[k] = f(k);
        };
        if (nhv == 0)
            stream([&](int) { return pos_inf(); });
        else if (nhv == 1)
            stream([&](int k) { return hu0 + (hea0 + hp0[k]); });
        else if (nhv == 2)
            stream([&](int k) { return fminf(hu0 + (hea0 + hp0[k]), hu1 + (hea1 + hp1[k])); });
        else
            stream(dense_at);
        // patches: the (k, value) pairs of the light j.  A patch must land after the streamed store of
        // its column: with few pairs every thread applies the patches of the columns IT streamed
        // (program order, no barrier); with many, a barrier and one thread per pair.
        auto patched = [&](uint32_t k) {
            float v = dense_at(static_cast<int>(k));
            for (int t2 = 0; t2 < npairs; ++t2)
                if (s_pk[t2] == k) {
                    const int q = s_pq[t2];
                    v = fminf(v, (eb + s_w[q]) + (Ea[s_j[q]] + s_pw[t2]));
                }
            return v;
        };
        if (npairs > 0 && npairs <= 32) {
            for (int t = 0; t < npairs; ++t) {
                const int k = static_cast<int>(s_pk[t]);
                int owner;
                if (k < kb)
                    owner = k;
                else if (k < kb + 4 * nb)
                    owner = ((k - kb) >> 2) & 255;
                else
                    owner = (k - kb - 4 * nb) & 255;
                if (owner == tid) O[k] = patched(static_cast<uint32_t>(k));
            }
        } else if (npairs > 32) {
            __syncthreads(); // the row's stores are ordered before the patches
            for (int t = tid; t < npairs; t += 256) O[s_pk[t]] = patched(s_pk[t]); // equal k: equal values
        }
    }
}

// ------------------------------------------------------------ K3: dense (min,+) GEMM ----
// C[t] = A[b] (x) B[a] for pair tuple t = a*emits + b (a consumed first).  Operands are the
// dense level-1 matrices padded to ld = npad (multiple of 128) with +inf: B = L1[a]
// (row-major, k x n) and A^T = L1T[b] (the transposed copy, k x m), so both tiles are plain
// row copies.  128x128x16 CTA tile, 256 threads, 8x8 register tile, 3-stage cp.async ring.
// Inner loop: per two k, 8 LDS.128 feed 128 FADD + 64 FMNMX3 (3-input min, sm_100).
constexpr int GM = 128, GN = 128, GK = 16, GSTAGES = 3;
constexpr size_t GEMM_SMEM = static_cast<size_t>(GSTAGES) * GK * (GM + GN) * sizeof(float);

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
    const unsigned d = static_cast<unsigned>(__cvta_generic_to_shared(smem_dst));
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gmem_src));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
template <int N> __device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N));
}

__global__ void __launch_bounds__(256, 2)
dense_minplus_gemm_kernel(const float* __restrict__ L1, const float* __restrict__ L1T, size_t npad,
                          float* __restrict__ out, int n, uint64_t tuple_first, int emits) {
    extern __shared__ __align__(16) float gsm[];
    float(*As)[GK][GM] = reinterpret_cast<float(*)[GK][GM]>(gsm);
    float(*Bs)[GK][GN] = reinterpret_cast<float(*)[GK][GN]>(gsm + GSTAGES * GK * GM);
    const uint64_t tuple = tuple_first + blockIdx.z;
    const int sa = static_cast<int>(tuple / emits); // first consumed symbol -> right factor
    const int sb = static_cast<int>(tuple % emits); // second symbol -> left factor
    const float* __restrict__ AT = L1T + static_cast<size_t>(sb) * npad * npad;
    const float* __restrict__ B = L1 + static_cast<size_t>(sa) * npad * npad;
    const int m0 = blockIdx.y * GM, n0 = blockIdx.x * GN;
    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;

    // each thread copies 2 x 16 B of the A^T tile and 2 x 16 B of the B tile per stage
    const int c_k = tid >> 5, c_col = (tid & 31) * 4; // rows c_k and c_k + 8
    auto issue = [&](int kt, int stage) {
        const size_t k0 = static_cast<size_t>(kt) * GK;
        cp_async16(&As[stage][c_k][c_col], AT + (k0 + c_k) * npad + m0 + c_col);
        cp_async16(&As[stage][c_k + 8][c_col], AT + (k0 + c_k + 8) * npad + m0 + c_col);
        cp_async16(&Bs[stage][c_k][c_col], B + (k0 + c_k) * npad + n0 + c_col);
        cp_async16(&Bs[stage][c_k + 8][c_col], B + (k0 + c_k + 8) * npad + n0 + c_col);
    };

    float c[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) c[i][j] = pos_inf();

    const int ktiles = static_cast<int>(npad) / GK;
#pragma unroll
    for (int s = 0; s < GSTAGES - 1; ++s) {
        if (s < ktiles) issue(s, s);
        cp_async_commit();
    }
    for (int kt = 0; kt < ktiles; ++kt) {
        cp_async_wait<GSTAGES - 2>();
        __syncthreads(); // tile kt landed for everyone; stage (kt-1)%S is free again
        if (kt + GSTAGES - 1 < ktiles) issue(kt + GSTAGES - 1, (kt + GSTAGES - 1) % GSTAGES);
        cp_async_commit();
        const int buf = kt % GSTAGES;
#pragma unroll
        for (int kk = 0; kk < GK; kk += 2) {
            float a0[8], a1[8], b0[8], b1[8];
            *reinterpret_cast<float4*>(&a0[0]) = *reinterpret_cast<const float4*>(&As[buf][kk][ty * 4]);
            *reinterpret_cast<float4*>(&a0[4]) = *reinterpret_cast<const float4*>(&As[buf][kk][64 + ty * 4]);
            *reinterpret_cast<float4*>(&a1[0]) = *reinterpret_cast<const float4*>(&As[buf][kk + 1][ty * 4]);
            *reinterpret_cast<float4*>(&a1[4]) = *reinterpret_cast<const float4*>(&As[buf][kk + 1][64 + ty * 4]);
            *reinterpret_cast<float4*>(&b0[0]) = *reinterpret_cast<const float4*>(&Bs[buf][kk][tx * 4]);
            *reinterpret_cast<float4*>(&b0[4]) = *reinterpret_cast<const float4*>(&Bs[buf][kk][64 + tx * 4]);
            *reinterpret_cast<float4*>(&b1[0]) = *reinterpret_cast<const float4*>(&Bs[buf][kk + 1][tx * 4]);
            *reinterpret_cast<float4*>(&b1[4]) = *reinterpret_cast<const float4*>(&Bs[buf][kk + 1][64 + tx * 4]);
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    c[i][j] = fminf(c[i][j], fminf(a0[i] + b0[j], a1[i] + b1[j]));
        }
    }
    float* __restrict__ O = out + static_cast<size_t>(blockIdx.z) * n * n;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int row = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
        if (row >= n) continue;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int col = n0 + (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
            if (col < n) O[static_cast<size_t>(row) * n + col] = c[i][j];
        }
    }
}

// ---------------------------------------------------------------------- run phase ----
struct RunSeqs {
    int nseq;
    const void* sym;
    int sym_bytes;
    const uint64_t* off;
};

__device__ __forceinline__ unsigned load_sym(const RunSeqs& s, uint64_t idx) {
    return s.sym_bytes == 1 ? static_cast<const uint8_t*>(s.sym)[idx]
                            : static_cast<const uint16_t*>(s.sym)[idx];
}

// v0[s][i] = fl(E[o_0][i] + start[i])      (GraphBLAS_spec_impl.cpp:54,153-154)
// grid: x = sequence (up to 2^31-1 of them), y = tile of states
// (rows of v are ld floats apart; columns n .. ld-1 are padding and hold +inf)
__global__ void spec_init_kernel(SvGenericDev m, RunSeqs s, float* __restrict__ v, int ld) {
    const int q = blockIdx.x;
    const unsigned o = load_sym(s, s.off[q]);
    for (int i = blockIdx.y * blockDim.x + threadIdx.x; i < ld; i += gridDim.y * blockDim.x)
        v[static_cast<size_t>(q) * ld + i] = i < m.n ? m.emis[static_cast<size_t>(o) * m.n + i] + m.start[i] : pos_inf();
}

// K4: one chunk of `level` symbols: vout[s] = H[key(s, chunk)] (x) vin[s]; sequences whose
// chunks are exhausted copy vin -> vout.  grid: x = sequence, y = row tile; one warp per row.
__global__ void __launch_bounds__(256)
handler_gemv_kernel(int n, int emits, int level, RunSeqs s, uint64_t chunk,
                    const float* __restrict__ table, uint64_t table_first,
                    const float* __restrict__ vin, float* __restrict__ vout, int ld, int rows_per_cta, int seq_base) {
    extern __shared__ float sv[];
    const int q = seq_base + blockIdx.x;
    const uint64_t o0 = s.off[q];
    const uint64_t len = s.off[q + 1] - o0;
    const uint64_t nchunks = (len - 1) / level;
    const float* __restrict__ vi = vin + static_cast<size_t>(q) * ld;
    float* __restrict__ vo = vout + static_cast<size_t>(q) * ld;
    const int r0 = blockIdx.y * rows_per_cta;
    const int r1 = min(n, r0 + rows_per_cta);
    if (chunk >= nchunks) {
        for (int i = r0 + threadIdx.x; i < r1; i += blockDim.x) vo[i] = vi[i];
        return;
    }
    for (int k = threadIdx.x; k < n; k += blockDim.x) sv[k] = vi[k];
    uint64_t tuple = 0;
    for (int j = 0; j < level; ++j) tuple = tuple * emits + load_sym(s, o0 + 1 + chunk * level + j);
    // table_first == ~0: on-demand mode, `table` holds one handler per sequence of this launch
    const float* __restrict__ H = table_first == ~0ull ? table + static_cast<size_t>(blockIdx.x) * n * n
                                                       : table + (tuple - table_first) * static_cast<size_t>(n) * n;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int i = r0 + warp; i < r1; i += nw) {
        const float* __restrict__ hrow = H + static_cast<size_t>(i) * n;
        float acc = pos_inf();
        for (int k = lane; k < n; k += 32) acc = fminf(acc, hrow[k] + sv[k]);
        acc = warp_min(acc);
        if (lane == 0) vo[i] = acc;
    }
}

// K4i: the level-2 chunk WITHOUT reading the table ("recompute on demand").  fl(. + v[k]) is monotone,
// so it commutes with the min over j inside a handler entry:
//   vout[i] = min_k fl( H(a,b)[i][k] + v[k] )
//           = min_{j in in(i)} min_{k in in(j)} fl( fl( fl(E[b][i]+T'[i][j]) + fl(E[a][j]+T'[j][k]) ) + v[k] )
// -- the same singly rounded sums in the same association as table + K4, hence the same bits, but the
// work is the number of two-step paths instead of N^2 table reads.  One thread per light row i: light j
// are walked through the CSR; for a heavy j (dense row of T', staged per CTA as r_h[k] = fl(E[a][j] +
// T'[j][k]) beside v) the thread streams all k with broadcast 16-byte shared loads.  Heavy rows i get a
// CTA each (blockIdx.x >= ntiles).  Up to kImplicitMaxHeavy dense rows.
constexpr int kImplicitMaxHeavy = 4;
constexpr int kImplicitThreads = 128;

__device__ __forceinline__ float4 lds128(uint32_t addr) {
    float4 v; // volatile: never hoisted over the barriers that publish v and r_h
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}

#define SV_HEAVY_TERMS(r, v)               \
    b0 = fminf(b0, (u + (r).x) + (v).x);   \
    b1 = fminf(b1, (u + (r).y) + (v).y);   \
    b2 = fminf(b2, (u + (r).z) + (v).z);   \
    b3 = fminf(b3, (u + (r).w) + (v).w);

// Packed IEEE adds (sm_100+: add.rn.f32x2 -> FADD2): two independent round-to-nearest fp32 adds per instruction
// on a 64-bit register pair; each half is an ordinary IEEE add, so the association and the bits are unchanged.
typedef unsigned long long f32x2_t;
__device__ __forceinline__ f32x2_t pack2f(float lo, float hi) {
    f32x2_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2f(f32x2_t p, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(p));
}
__device__ __forceinline__ f32x2_t add2f(f32x2_t a, f32x2_t b) {
    f32x2_t r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
// 16 bytes of shared memory as two register pairs (volatile: never hoisted over the barriers that publish v and r_h)
__device__ __forceinline__ void lds128_pairs(uint32_t addr, f32x2_t& lo, f32x2_t& hi) {
    asm volatile("ld.shared.v2.b64 {%0, %1}, [%2];" : "=l"(lo), "=l"(hi) : "r"(addr));
}

// best[x] = min(best[x], min_k fl(fl(u[x] + r[k]) + v[k])) over all npad columns (npad is a multiple of 16, +inf
// padded) for R rows at once: a broadcast LDS costs shared-memory wavefronts like any other, so every loaded
// (r, v) quad is used for R rows.  16 columns per trip = 8 LDS.128 with immediate offsets, 16 R FADD2 (two adds each),
// 8 R FMNMX3: 2 instructions per term where scalar adds need 3.
template <int R>
__device__ __forceinline__ void heavy_stream_all(const float (&u)[R], uint32_t r_addr, uint32_t v_addr, int npad,
                                                 float (&best)[R]) {
    float b0[R], b1[R];
    f32x2_t uu[R];
#pragma unroll
    for (int x = 0; x < R; ++x) {
        b0[x] = b1[x] = pos_inf();
        uu[x] = pack2f(u[x], u[x]);
    }
    const uint32_t r_end = r_addr + static_cast<uint32_t>(npad) * 4u;
#pragma unroll 1
    for (; r_addr != r_end; r_addr += 64, v_addr += 64) {
        f32x2_t rl[4], rh[4], vl[4], vh[4];
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            lds128_pairs(r_addr + 16 * g, rl[g], rh[g]);
            lds128_pairs(v_addr + 16 * g, vl[g], vh[g]);
        }
#pragma unroll
        for (int g = 0; g < 4; ++g) {
#pragma unroll
            for (int x = 0; x < R; ++x) {
                float t0, t1, t2, t3;
                unpack2f(add2f(add2f(uu[x], rl[g]), vl[g]), t0, t1); // fl(fl(u + r) + v), columns 4g and 4g+1
                unpack2f(add2f(add2f(uu[x], rh[g]), vh[g]), t2, t3); // ... 4g+2 and 4g+3
                b0[x] = fminf(fminf(b0[x], t0), t1);
                b1[x] = fminf(fminf(b1[x], t2), t3);
            }
        }
    }
#pragma unroll
    for (int x = 0; x < R; ++x) best[x] = fminf(best[x], fminf(b0[x], b1[x]));
}

// the same columns dealt to the threads of a CTA (heavy rows): thread k0 of kstep takes every kstep-th float4
__device__ __forceinline__ float heavy_stream(float u, const float4* __restrict__ r4,
                                              const float4* __restrict__ v4, int n4, int k0, int kstep) {
    float b0 = pos_inf(), b1 = pos_inf(), b2 = pos_inf(), b3 = pos_inf();
    for (int k = k0; k < n4; k += kstep) {
        const float4 r = r4[k];
        const float4 v = v4[k];
        SV_HEAVY_TERMS(r, v)
    }
    return fminf(fminf(b0, b1), fminf(b2, b3));
}

// One light row i: returns min over its two-step paths (see above).  Must be called by whole warps
// (`live` = this lane owns a light row): the heavy streams are entered on a warp vote.
__device__ __forceinline__ float path_term(const uint4 t, float e, const float* __restrict__ ea,
                                           const float* __restrict__ sv) {
    return ((e + __uint_as_float(t.x)) + (ea[t.z] + __uint_as_float(t.y))) + sv[t.w];
}

// R light rows per thread (rows i0 + x * stride, x < R; `live` = bit x set when the lane owns that row).
// Must be called by whole warps: the heavy streams are entered on a warp vote.
template <int R>
__device__ __forceinline__ void implicit_light_rows(const SvPairDev& pd, const float* __restrict__ ea,
                                                    const float* __restrict__ eb, int i0, int stride, unsigned live,
                                                    const float* __restrict__ sv, const float* __restrict__ sr,
                                                    int npad, float (&best)[R]) {
    float uh[kImplicitMaxHeavy][R];
#pragma unroll
    for (int h = 0; h < kImplicitMaxHeavy; ++h)
#pragma unroll
        for (int x = 0; x < R; ++x) uh[h][x] = pos_inf();
#pragma unroll
    for (int x = 0; x < R; ++x) {
        best[x] = pos_inf();
        if (!((live >> x) & 1u)) continue;
        const int i = i0 + x * stride;
        const float e = eb[i];
#pragma unroll
        for (int h = 0; h < kImplicitMaxHeavy; ++h)
            if (h < pd.n_stage) uh[h][x] = e + pd.wcol[static_cast<size_t>(pd.stage_h[h]) * pd.n + i];
        const int cnt = static_cast<int>(pd.lcount[i]);
        const uint4* __restrict__ lp = pd.lpath + i;
        float bx = pos_inf();
        int c = 0;
        for (; c + 4 <= cnt; c += 4) { // independent loads, four in flight
            const uint4 t0 = lp[static_cast<size_t>(c) * pd.n], t1 = lp[static_cast<size_t>(c + 1) * pd.n];
            const uint4 t2 = lp[static_cast<size_t>(c + 2) * pd.n], t3 = lp[static_cast<size_t>(c + 3) * pd.n];
            bx = fminf(fminf(bx, path_term(t0, e, ea, sv)), path_term(t1, e, ea, sv));
            bx = fminf(fminf(bx, path_term(t2, e, ea, sv)), path_term(t3, e, ea, sv));
        }
        for (; c < cnt; ++c) bx = fminf(bx, path_term(lp[static_cast<size_t>(c) * pd.n], e, ea, sv));
        best[x] = bx;
    }
    const uint32_t sv_addr = static_cast<uint32_t>(__cvta_generic_to_shared(sv));
    const uint32_t sr_addr = static_cast<uint32_t>(__cvta_generic_to_shared(sr));
#pragma unroll
    for (int h = 0; h < kImplicitMaxHeavy; ++h) {
        if (h >= pd.n_stage) continue;
        bool any = false;
#pragma unroll
        for (int x = 0; x < R; ++x) any = any || uh[h][x] < pos_inf();
        if (__any_sync(0xffffffffu, any))
            heavy_stream_all<R>(uh[h], sr_addr + static_cast<uint32_t>(h * npad) * 4u, sv_addr, npad, best);
    }
}

// Heavy row number hme, spread over `nthr` threads (tid = this thread's place): returns its partial minimum.
__device__ __forceinline__ float implicit_heavy_row(const SvPairDev& pd, const float* __restrict__ ea,
                                                    const float* __restrict__ eb, int hme,
                                                    const float* __restrict__ sv, const float* __restrict__ sr,
                                                    int npad, int tid, int nthr) {
    const int n = pd.n;
    const float e = eb[pd.hrow[hme]];
    float best = pos_inf();
    const uint32_t p1 = pd.hpath_ptr[hme + 1];
    for (uint32_t p = pd.hpath_ptr[hme] + tid; p < p1; p += nthr) best = fminf(best, path_term(pd.hpath[p], e, ea, sv));
    const float4* __restrict__ v4 = reinterpret_cast<const float4*>(sv);
    for (int h = 0; h < pd.n_stage; ++h) {
        const float w = pd.wd[static_cast<size_t>(hme) * n + pd.hrow[pd.stage_h[h]]]; // T'[i][source row], +inf = absent
        if (w < pos_inf())
            best = fminf(best, heavy_stream(e + w, reinterpret_cast<const float4*>(sr + h * npad), v4, npad >> 2, tid, nthr));
    }
    return best;
}

// ---- TMA staging (1-D bulk copies global -> shared, completion on an mbarrier) ----
// The per-CTA inputs of a chunk -- the sequence's score vector (rows of npad floats in global memory, +inf padded)
// and the r_h rows of the chunk's first symbol (tabulated once per model, SvPairDev::rh) -- arrive as bulk copies
// issued by one thread instead of being loaded, recomputed and stored by all of them.
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "SV_MBAR_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra SV_MBAR_DONE;\n"
        "bra SV_MBAR_WAIT;\n"
        "SV_MBAR_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// 1-D bulk copy global -> shared (TMA), bytes a multiple of 16, both addresses 16-byte aligned
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
// per-chunk launch: grid x = sequence, y = row tile (then one CTA per heavy row)
template <int R>
__global__ void __launch_bounds__(kImplicitThreads)
implicit_pair_step_kernel(SvPairDev pd, RunSeqs s, uint64_t chunk, const float* __restrict__ vin,
                          float* __restrict__ vout, int seq_base, int ntiles) {
    extern __shared__ float4 sm4[];
    const int n = pd.n, npad = pd.npad; // rows of vin / vout are npad floats apart, columns n.. hold +inf
    float* sv = reinterpret_cast<float*>(sm4);  // [npad] v, +inf padded
    float* sr = sv + npad;                      // [n_stage][npad] r_h
    __shared__ float s_red[kImplicitThreads / 32 /*max*/];
    __shared__ __align__(8) uint64_t s_bar;
    const int nthr = blockDim.x; // 64 or 128
    const int q = seq_base + blockIdx.x;
    const uint64_t o0 = s.off[q];
    const uint64_t len = s.off[q + 1] - o0;
    const float* __restrict__ vi = vin + static_cast<size_t>(q) * npad;
    float* __restrict__ vo = vout + static_cast<size_t>(q) * npad;
    const bool heavy_cta = static_cast<int>(blockIdx.y) >= ntiles;
    const int hme = static_cast<int>(blockIdx.y) - ntiles;
    if (chunk >= (len - 1) / 2) {
        if (heavy_cta) {
            if (threadIdx.x == 0) vo[pd.hrow[hme]] = vi[pd.hrow[hme]];
        } else {
#pragma unroll
            for (int x = 0; x < R; ++x) {
                const int i = (blockIdx.y * R + x) * nthr + threadIdx.x;
                if (i < npad && (i >= n || pd.hidx[i] < 0)) vo[i] = vi[i]; // padding included
            }
        }
        return;
    }
    const unsigned a = load_sym(s, o0 + 1 + chunk * 2), b = load_sym(s, o0 + 2 + chunk * 2);
    const float* __restrict__ ea = pd.emis + static_cast<size_t>(a) * n;
    const float* __restrict__ eb = pd.emis + static_cast<size_t>(b) * n;
    if (threadIdx.x == 0) {
        mbar_init(&s_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        const unsigned row_bytes = static_cast<unsigned>(npad) * 4u;
        mbar_expect_tx(&s_bar, row_bytes * static_cast<unsigned>(1 + pd.n_stage));
        bulk_g2s(sv, vi, row_bytes, &s_bar);
        for (int h = 0; h < pd.n_stage; ++h)
            bulk_g2s(sr + h * npad, pd.rh + (static_cast<size_t>(a) * pd.n_stage + h) * npad, row_bytes, &s_bar);
    }
    __syncthreads(); // the initialised barrier is visible to every waiter
    mbar_wait(&s_bar, 0);
    if (!heavy_cta) {
        const int i0 = blockIdx.y * R * nthr + threadIdx.x;
        unsigned live = 0;
#pragma unroll
        for (int x = 0; x < R; ++x) {
            const int i = i0 + x * nthr;
            if (i < n && pd.hidx[i] < 0) live |= 1u << x;
        }
        float best[R];
        implicit_light_rows<R>(pd, ea, eb, i0, nthr, live, sv, sr, npad, best);
#pragma unroll
        for (int x = 0; x < R; ++x) {
            const int i = i0 + x * nthr;
            if ((live >> x) & 1u) vo[i] = best[x];
            else if (i >= n && i < npad) vo[i] = pos_inf(); // the padding the next chunk's bulk copy fetches
        }
        return;
    }
    float best = warp_min(implicit_heavy_row(pd, ea, eb, hme, sv, sr, npad, threadIdx.x, nthr));
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = best;
    __syncthreads();
    if (threadIdx.x == 0) {
        float r = s_red[0];
#pragma unroll
        for (int wv = 1; wv < nthr / 32; ++wv) r = fminf(r, s_red[wv]);
        vo[pd.hrow[hme]] = r;
    }
}

// Level-1 tail step number `tail` (0-based) of every sequence that still has one
// (GraphBLAS_spec_impl.cpp:84-89); others copy.  One CTA per sequence.
__global__ void __launch_bounds__(1024)
tail_step_kernel(SvGenericDev m, int level, RunSeqs s, int tail, const float* __restrict__ vin,
                 float* __restrict__ vout, int ld) {
    extern __shared__ float sv[];
    const int q = blockIdx.x;
    const uint64_t o0 = s.off[q];
    const uint64_t len = s.off[q + 1] - o0;
    const uint64_t nchunks = (len - 1) / level;
    const uint64_t ntail = (len - 1) % level;
    const float* __restrict__ vi = vin + static_cast<size_t>(q) * ld;
    float* __restrict__ vo = vout + static_cast<size_t>(q) * ld;
    if (static_cast<uint64_t>(tail) >= ntail) {
        for (int i = threadIdx.x; i < m.n; i += blockDim.x) vo[i] = vi[i];
        return;
    }
    for (int k = threadIdx.x; k < m.n; k += blockDim.x) sv[k] = vi[k];
    const unsigned o = load_sym(s, o0 + 1 + nchunks * level + tail);
    const float* __restrict__ em = m.emis + static_cast<size_t>(o) * m.n;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    // warp per row keeps the dense rows of profile HMMs cheap; light rows cost one pass
    for (int i = warp; i < m.n; i += nw) {
        const float e = em[i];
        float acc = pos_inf();
        for (uint32_t p = m.rowptr[i] + lane; p < m.rowptr[i + 1]; p += 32) {
            const float u = e + m.w[p];
            acc = fminf(acc, u + sv[m.col[p]]);
        }
        acc = warp_min(acc);
        if (lane == 0) vo[i] = acc;
    }
}

// (min, lowest argmin) of each final vector; optional copy-out is done by the caller.
__global__ void __launch_bounds__(256)
best_state_kernel(int n, int ld, const float* __restrict__ v, float* __restrict__ best,
                  int32_t* __restrict__ arg) {
    __shared__ float s_bv[8];
    __shared__ int s_bi[8];
    const int q = blockIdx.x;
    const float* __restrict__ x = v + static_cast<size_t>(q) * ld;
    float bv = pos_inf();
    int bi = 0x7fffffff;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        if (x[i] < bv) {
            bv = x[i];
            bi = i;
        }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int round = 0; round < 2; ++round) {
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, bv, d);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, d);
            if (ov < bv || (ov == bv && oi < bi)) {
                bv = ov;
                bi = oi;
            }
        }
        if (round == 0) {
            if (lane == 0) {
                s_bv[warp] = bv;
                s_bi[warp] = bi;
            }
            __syncthreads();
            bv = lane < 8 ? s_bv[lane] : pos_inf();
            bi = lane < 8 ? s_bi[lane] : 0x7fffffff;
        }
    }
    if (threadIdx.x == 0) {
        if (best) best[q] = bv;
        if (arg) arg[q] = bi == 0x7fffffff ? -1 : bi;
    }
}

uint64_t ipow(uint64_t b, uint64_t e) {
    uint64_t r = 1;
    for (uint64_t k = 0; k < e; ++k) {
        if (r > (~0ull) / b) return ~0ull;
        r *= b;
    }
    return r;
}

// Builds handlers [tb, te) of level `lvl` into dst (dense, ld = n) with the sparse-left
// kernel, recursing over the prefixes.  Level 1 = dense scatter of M_o.
bool pair_prepare(sv_model* m, SvPairDev* out);
int build_pair_level2(sv_model* m, const SvPairDev& pd, uint64_t tb, uint64_t te, float* dst, cudaStream_t st);

int build_sparse(sv_model* m, uint64_t lvl, uint64_t tb, uint64_t te, float* dst, cudaStream_t st) {
    const size_t n = m->n;
    const size_t mat = n * n;
    if (lvl == 1) {
        const size_t count = (te - tb) * mat;
        fill_inf_kernel<<<std::min<size_t>((count + 255) / 256, 148 * 16), 256, 0, st>>>(dst, count);
        dim3 grid(static_cast<unsigned>((n + 127) / 128), static_cast<unsigned>(te - tb));
        scatter_level1_kernel<<<grid, 128, 0, st>>>(m->gen, dst, nullptr, n, mat,
                                                    static_cast<int>(tb), static_cast<int>(te - tb));
        SV_CUDA(cudaGetLastError());
        return SV_OK;
    }
    const uint64_t pb = tb / m->emits, pe = (te - 1) / m->emits + 1;
    float* prev = nullptr;
    if (cudaMalloc(&prev, (pe - pb) * mat * sizeof(float)) != cudaSuccess) {
        cudaGetLastError();
        return sv_fail(SV_ERR_MEMORY, "spec precompute: prefix table does not fit device memory");
    }
    // the level-2 prefix table of a level-3 build comes from the streaming pair kernel when the model allows
    SvPairDev pd;
    int rc = (lvl == 3 && std::getenv("SV_SPEC_LHS_PER_TUPLE") == nullptr && pair_prepare(m, &pd))
                 ? build_pair_level2(m, pd, pb, pe, prev, st)
                 : build_sparse(m, lvl - 1, pb, pe, prev, st);
    if (rc == SV_OK) {
        const uint64_t ymax = 32768;
        const bool per_tuple = std::getenv("SV_SPEC_LHS_PER_TUPLE") != nullptr; // first version, kept for A/B runs
        for (uint64_t t0 = tb; per_tuple && t0 < te && rc == SV_OK; t0 += ymax) {
            const uint64_t cnt = std::min(ymax, te - t0);
            dim3 grid(static_cast<unsigned>((n + 1023) / 1024), static_cast<unsigned>(n),
                      static_cast<unsigned>(cnt));
            sparse_lhs_kernel<<<grid, 256, 0, st>>>(m->gen, prev, pb, dst + (t0 - tb) * mat, t0,
                                                    static_cast<int>(m->emits));
            if (cudaGetLastError() != cudaSuccess) rc = sv_fail(SV_ERR_CUDA, "sparse_lhs launch failed");
        }
        // the heavy rows' long-running CTAs go to a side stream: they write other rows than the light pass
        cudaStream_t hst = nullptr;
        cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
        if (!per_tuple && m->gen.n_heavy > 0) {
            if (cudaStreamCreateWithFlags(&hst, cudaStreamNonBlocking) != cudaSuccess ||
                cudaEventCreateWithFlags(&ev_fork, cudaEventDisableTiming) != cudaSuccess ||
                cudaEventCreateWithFlags(&ev_join, cudaEventDisableTiming) != cudaSuccess)
                rc = sv_fail(SV_ERR_CUDA, "spec precompute: side stream");
            if (rc == SV_OK) {
                cudaEventRecord(ev_fork, st); // prev is complete
                cudaStreamWaitEvent(hst, ev_fork, 0);
            }
        }
        for (uint64_t f0 = pb; !per_tuple && f0 < pe && rc == SV_OK; f0 += ymax) {
            const uint64_t cnt = std::min(ymax, pe - f0);
            if (m->gen.n_heavy > 0) {
                dim3 hgrid(static_cast<unsigned>((n + 255) / 256), static_cast<unsigned>(m->gen.n_heavy),
                           static_cast<unsigned>(cnt));
                sparse_lhs_prefix_kernel<20, 1><<<hgrid, 256, 0, hst>>>(m->gen, prev, pb, dst, tb, tb, te, f0,
                                                                        static_cast<int>(m->emits), 1);
            }
            dim3 grid(static_cast<unsigned>(n), static_cast<unsigned>(cnt));
            // (<20, 2> for short rows was tried on 500.chmm level 3: 11.1 ms vs 6.4 ms -- fewer concurrent
            //  output streams per CTA matter more than idle lanes)
            sparse_lhs_prefix_kernel<SV_LHS_EB, SV_LHS_KU><<<grid, 256, 0, st>>>(m->gen, prev, pb, dst, tb, tb, te, f0,
                                                                               static_cast<int>(m->emits), 0);
            if (cudaGetLastError() != cudaSuccess) rc = sv_fail(SV_ERR_CUDA, "sparse_lhs launch failed");
        }
        if (hst != nullptr) {
            if (ev_join != nullptr) {
                cudaEventRecord(ev_join, hst);
                cudaStreamWaitEvent(st, ev_join, 0);
            }
            cudaStreamSynchronize(hst);
            cudaStreamDestroy(hst);
        }
        if (ev_fork) cudaEventDestroy(ev_fork);
        if (ev_join) cudaEventDestroy(ev_join);
    }
    cudaError_t e = cudaStreamSynchronize(st);
    cudaFree(prev);
    if (rc == SV_OK && e != cudaSuccess) rc = sv_fail(SV_ERR_CUDA, cudaGetErrorString(e));
    return rc;
}

// Device images for K3'' (built on first use, cached in the model).  Returns false if the model
// has too many heavy rows for the dense copies (then K3' is used).
bool pair_prepare(sv_model* m, SvPairDev* out) {
    const size_t n = m->n, nnz = m->csr.col.size();
    if (m->pair_state < 0) return false;
    std::vector<int32_t> hidx(n, -1);
    std::vector<uint32_t> heavy;
    for (size_t i = 0; i < n; ++i)
        if (m->csr.rowptr[i + 1] - m->csr.rowptr[i] > static_cast<uint32_t>(kPairHeavy)) {
            hidx[i] = static_cast<int32_t>(heavy.size());
            heavy.push_back(static_cast<uint32_t>(i));
        }
    if (heavy.size() > static_cast<size_t>(kPairMaxHeavy)) {
        m->pair_state = -1;
        return false;
    }
    const size_t nh = std::max<size_t>(heavy.size(), 1);
    auto align = [](size_t x) { return (x + 255) / 256 * 256; };
    const size_t o_hidx = 0;
    const size_t o_wd = align(o_hidx + n * 4);
    const size_t o_cptr = align(o_wd + nh * n * 4);
    const size_t o_crow = align(o_cptr + (n + 1) * 4);
    const size_t o_cw = align(o_crow + std::max<size_t>(nnz, 1) * 4);
    const size_t o_long = align(o_cw + std::max<size_t>(nnz, 1) * 4);
    const size_t o_hrow = align(o_long + n * 4); // (at most n long columns)
    if (m->pair_state == 0) {
        std::vector<float> wd(nh * n, std::numeric_limits<float>::infinity());
        for (size_t h = 0; h < heavy.size(); ++h)
            for (uint32_t p = m->csr.rowptr[heavy[h]]; p < m->csr.rowptr[heavy[h] + 1]; ++p)
                wd[h * n + m->csr.col[p]] = m->csr.w[p];
        std::vector<uint32_t> cptr(n + 1, 0), crow(nnz), fillp;
        std::vector<float> cw(nnz);
        for (size_t p = 0; p < nnz; ++p) cptr[m->csr.col[p] + 1] += 1;
        for (size_t k = 0; k < n; ++k) cptr[k + 1] += cptr[k];
        fillp.assign(cptr.begin(), cptr.end() - 1);
        for (size_t i = 0; i < n; ++i)
            for (uint32_t p = m->csr.rowptr[i]; p < m->csr.rowptr[i + 1]; ++p) {
                const uint32_t q = fillp[m->csr.col[p]]++;
                crow[q] = static_cast<uint32_t>(i);
                cw[q] = m->csr.w[p];
            }
        std::vector<uint32_t> longcol;
        for (size_t k = 0; k < n; ++k)
            if (cptr[k + 1] - cptr[k] > static_cast<uint32_t>(kPairLongCol)) longcol.push_back(static_cast<uint32_t>(k));
        const size_t bytes = o_hrow + nh * 4 + 256;
        std::vector<unsigned char> host(bytes, 0);
        std::memcpy(host.data() + o_hidx, hidx.data(), n * 4);
        std::memcpy(host.data() + o_wd, wd.data(), wd.size() * 4);
        std::memcpy(host.data() + o_cptr, cptr.data(), cptr.size() * 4);
        if (nnz) std::memcpy(host.data() + o_crow, crow.data(), nnz * 4);
        if (nnz) std::memcpy(host.data() + o_cw, cw.data(), nnz * 4);
        if (!longcol.empty()) std::memcpy(host.data() + o_long, longcol.data(), longcol.size() * 4);
        if (!heavy.empty()) std::memcpy(host.data() + o_hrow, heavy.data(), heavy.size() * 4);
        if (cudaMalloc(&m->d_pair_blob, bytes) != cudaSuccess ||
            cudaMemcpy(m->d_pair_blob, host.data(), bytes, cudaMemcpyHostToDevice) != cudaSuccess) {
            cudaGetLastError();
            if (m->d_pair_blob) cudaFree(m->d_pair_blob);
            m->d_pair_blob = nullptr;
            m->pair_state = -1;
            return false;
        }
        m->pair_state = 1 + static_cast<int>(longcol.size());
    }
    auto* base = static_cast<unsigned char*>(m->d_pair_blob);
    out->n = static_cast<int>(n);
    out->emits = static_cast<int>(m->emits);
    out->n_longcol = m->pair_state - 1;
    out->rowptr = m->gen.rowptr;
    out->col = m->gen.col;
    out->w = m->gen.w;
    out->emis = m->gen.emis;
    out->hidx = reinterpret_cast<const int32_t*>(base + o_hidx);
    out->wd = reinterpret_cast<const float*>(base + o_wd);
    out->cscptr = reinterpret_cast<const uint32_t*>(base + o_cptr);
    out->cscrow = reinterpret_cast<const uint32_t*>(base + o_crow);
    out->cscw = reinterpret_cast<const float*>(base + o_cw);
    out->longcol = reinterpret_cast<const uint32_t*>(base + o_long);
    out->hrow = reinterpret_cast<const uint32_t*>(base + o_hrow);
    out->n_heavy = static_cast<int>(heavy.size());
    return true;
}

// K4i/K4p: flattens the two-step paths through light intermediate states (see SvPairDev).  Rows with more
// than kPathMaxSlots such paths make the model ineligible (the table run is used).
constexpr int kPathMaxSlots = 1024;

bool path_prepare(sv_model* m, SvPairDev* pd) {
    if (m->path_state < 0 || pd->n_heavy > kImplicitMaxHeavy) return false;
    const size_t n = m->n;
    const size_t nh = std::max<size_t>(static_cast<size_t>(pd->n_heavy), 1);
    auto align = [](size_t x) { return (x + 255) / 256 * 256; };
    if (m->path_state == 0) {
        const auto& rp = m->csr.rowptr;
        const auto& cl = m->csr.col;
        const auto& cw = m->csr.w;
        auto heavy = [&](size_t r) { return rp[r + 1] - rp[r] > static_cast<uint32_t>(kPairHeavy); };
        std::vector<uint32_t> hrow;
        for (size_t i = 0; i < n; ++i) if (heavy(i)) hrow.push_back(static_cast<uint32_t>(i));
        std::vector<uint32_t> lcount(n, 0), hptr(nh + 1, 0);
        size_t slots = 0;
        for (size_t i = 0; i < n; ++i) {
            size_t c = 0;
            for (uint32_t p = rp[i]; p < rp[i + 1]; ++p)
                if (!heavy(cl[p])) c += rp[cl[p] + 1] - rp[cl[p]];
            if (heavy(i)) {
                const size_t h = std::find(hrow.begin(), hrow.end(), static_cast<uint32_t>(i)) - hrow.begin();
                hptr[h + 1] = static_cast<uint32_t>(c);
            } else {
                lcount[i] = static_cast<uint32_t>(c);
                slots = std::max(slots, c);
            }
        }
        if (slots > static_cast<size_t>(kPathMaxSlots)) {
            m->path_state = -1;
            return false;
        }
        for (size_t h = 0; h < nh; ++h) hptr[h + 1] += hptr[h];
        const size_t htotal = hptr[nh];
        struct P4 { float w1, w2; uint32_t j, k; };
        std::vector<P4> lpath(std::max<size_t>(slots * n, 1), P4{0, 0, 0, 0}), hpath(std::max<size_t>(htotal, 1), P4{0, 0, 0, 0});
        std::vector<float> wcol(nh * n, std::numeric_limits<float>::infinity());
        unsigned used = 0;
        for (size_t i = 0; i < n; ++i) {
            size_t c = 0;
            const size_t hi = heavy(i) ? std::find(hrow.begin(), hrow.end(), static_cast<uint32_t>(i)) - hrow.begin() : 0;
            for (uint32_t p = rp[i]; p < rp[i + 1]; ++p) {
                const uint32_t j = cl[p];
                if (heavy(j)) {
                    const size_t hj = std::find(hrow.begin(), hrow.end(), j) - hrow.begin();
                    wcol[hj * n + i] = cw[p];
                    if (cw[p] < std::numeric_limits<float>::infinity()) used |= 1u << hj;
                    continue;
                }
                for (uint32_t t = rp[j]; t < rp[j + 1]; ++t, ++c) {
                    const P4 e{cw[p], cw[t], j, cl[t]};
                    if (heavy(i)) hpath[hptr[hi] + c] = e; else lpath[c * n + i] = e;
                }
            }
        }
        // r_h tables of the staged dense rows (those that are a source of some row), one per symbol
        const size_t npad = (n + 15) / 16 * 16;
        std::vector<uint32_t> stage;
        for (size_t h = 0; h < hrow.size(); ++h)
            if ((used >> h) & 1u) stage.push_back(static_cast<uint32_t>(h));
        std::vector<float> rh(std::max<size_t>(m->emits * stage.size() * npad, 1), std::numeric_limits<float>::infinity());
        for (size_t a = 0; a < m->emits; ++a)
            for (size_t sidx = 0; sidx < stage.size(); ++sidx) {
                const uint32_t row = hrow[stage[sidx]];
                const float e = m->emis[a * n + row];
                float* dst = &rh[(a * stage.size() + sidx) * npad];
                for (uint32_t p2 = rp[row]; p2 < rp[row + 1]; ++p2) dst[cl[p2]] = e + cw[p2]; // fl(E + T'), as the kernels form it
            }
        const size_t o_lp = 0;
        const size_t o_hp = align(o_lp + lpath.size() * 16);
        const size_t o_lc = align(o_hp + hpath.size() * 16);
        const size_t o_wc = align(o_lc + n * 4);
        const size_t o_ptr = align(o_wc + wcol.size() * 4);
        const size_t o_rh = align(o_ptr + (nh + 1) * 4);
        const size_t bytes = o_rh + rh.size() * 4 + 256;
        m->path_rh_offset = o_rh;
        std::vector<unsigned char> host(bytes, 0);
        std::memcpy(host.data() + o_rh, rh.data(), rh.size() * 4);
        std::memcpy(host.data() + o_lp, lpath.data(), lpath.size() * 16);
        std::memcpy(host.data() + o_hp, hpath.data(), hpath.size() * 16);
        std::memcpy(host.data() + o_lc, lcount.data(), n * 4);
        std::memcpy(host.data() + o_wc, wcol.data(), wcol.size() * 4);
        std::memcpy(host.data() + o_ptr, hptr.data(), (nh + 1) * 4);
        if (cudaMalloc(&m->d_path_blob, bytes) != cudaSuccess ||
            cudaMemcpy(m->d_path_blob, host.data(), bytes, cudaMemcpyHostToDevice) != cudaSuccess) {
            cudaGetLastError();
            if (m->d_path_blob) cudaFree(m->d_path_blob);
            m->d_path_blob = nullptr;
            m->path_state = -1;
            return false;
        }
        m->path_state = 1;
        m->path_slots = static_cast<int>(slots);
        m->path_heavy_total = static_cast<int>(htotal);
        m->path_used_mask = used;
    }
    const size_t lp_count = std::max<size_t>(static_cast<size_t>(m->path_slots) * n, 1);
    const size_t hp_count = std::max<size_t>(static_cast<size_t>(m->path_heavy_total), 1);
    const size_t o_lp = 0;
    const size_t o_hp = align(o_lp + lp_count * 16);
    const size_t o_lc = align(o_hp + hp_count * 16);
    const size_t o_wc = align(o_lc + n * 4);
    const size_t o_ptr = align(o_wc + nh * n * 4);
    auto* base = static_cast<unsigned char*>(m->d_path_blob);
    pd->lpath = reinterpret_cast<const uint4*>(base + o_lp);
    pd->hpath = reinterpret_cast<const uint4*>(base + o_hp);
    pd->lcount = reinterpret_cast<const uint32_t*>(base + o_lc);
    pd->wcol = reinterpret_cast<const float*>(base + o_wc);
    pd->hpath_ptr = reinterpret_cast<const uint32_t*>(base + o_ptr);
    pd->rh = reinterpret_cast<const float*>(base + m->path_rh_offset);
    pd->npad = static_cast<int>((n + 15) / 16 * 16);
    pd->path_slots = m->path_slots;
    pd->n_stage = 0;
    for (int h = 0; h < pd->n_heavy; ++h)
        if ((m->path_used_mask >> h) & 1u) pd->stage_h[pd->n_stage++] = h;
    return true;
}

// Level 2 with K3'': no dense prefix table, one streaming pass over the output.
int build_pair_level2(sv_model* m, const SvPairDev& pd, uint64_t tb, uint64_t te, float* dst, cudaStream_t st) {
    const size_t n = m->n;
    int rc = SV_OK;
    const uint64_t ymax = 32768ull;
    for (uint64_t t0 = tb; t0 < te && rc == SV_OK; t0 += ymax) {
        const uint64_t cnt = std::min(ymax, te - t0);
        dim3 grid(static_cast<unsigned>(n), static_cast<unsigned>((cnt + kPairChunk - 1) / kPairChunk));
        if (pd.n_heavy > 0) { // first: the few long-running CTAs
            dim3 hgrid(static_cast<unsigned>(pd.n_heavy), static_cast<unsigned>(cnt));
            pair_rows_kernel<<<hgrid, 256, 0, st>>>(pd, dst + (t0 - tb) * n * n, t0, cnt, 1);
        }
        SvPairDev pl = pd; // light pass: stage the dense source of a row when it fits beside 4 resident CTAs
        const size_t stage_bytes = (n + 8) * sizeof(float);
        pl.stage_row = pd.n_heavy > 0 && stage_bytes <= 28 * 1024;
        pair_rows_kernel<<<grid, 256, pl.stage_row ? stage_bytes : 0, st>>>(pl, dst + (t0 - tb) * n * n, t0, cnt, 0);
        if (cudaGetLastError() != cudaSuccess) rc = sv_fail(SV_ERR_CUDA, "pair_rows launch failed");
    }
    cudaError_t e = cudaStreamSynchronize(st);
    if (rc == SV_OK && e != cudaSuccess) rc = sv_fail(SV_ERR_CUDA, cudaGetErrorString(e));
    return rc;
}

int build_dense_level2(sv_model* m, uint64_t tb, uint64_t te, float* dst, cudaStream_t st) {
    const size_t n = m->n;
    const size_t npad = (n + GM - 1) / GM * GM;
    float* l1 = nullptr;
    const size_t count = 2 * m->emits * npad * npad; // row-major copy, then transposed copy
    if (cudaMalloc(&l1, count * sizeof(float)) != cudaSuccess) {
        cudaGetLastError();
        return sv_fail(SV_ERR_MEMORY, "spec precompute: padded level-1 operands do not fit");
    }
    fill_inf_kernel<<<148 * 16, 256, 0, st>>>(l1, count);
    dim3 sgrid(static_cast<unsigned>((n + 127) / 128), static_cast<unsigned>(m->emits));
    float* l1t = l1 + m->emits * npad * npad;
    scatter_level1_kernel<<<sgrid, 128, 0, st>>>(m->gen, l1, l1t, npad, npad * npad, 0,
                                                 static_cast<int>(m->emits));
    if (cudaFuncSetAttribute(dense_minplus_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             static_cast<int>(GEMM_SMEM)) != cudaSuccess) {
        cudaFree(l1);
        return sv_fail(SV_ERR_CUDA, "dense gemm: cannot reserve shared memory");
    }
    int rc = SV_OK;
    const uint64_t zmax = 32768;
    for (uint64_t t0 = tb; t0 < te && rc == SV_OK; t0 += zmax) {
        const uint64_t cnt = std::min(zmax, te - t0);
        dim3 grid(static_cast<unsigned>(npad / GN), static_cast<unsigned>(npad / GM),
                  static_cast<unsigned>(cnt));
        dense_minplus_gemm_kernel<<<grid, 256, GEMM_SMEM, st>>>(l1, l1t, npad, dst + (t0 - tb) * n * n,
                                                        static_cast<int>(n), t0,
                                                        static_cast<int>(m->emits));
        if (cudaGetLastError() != cudaSuccess) rc = sv_fail(SV_ERR_CUDA, "dense gemm launch failed");
    }
    cudaError_t e = cudaStreamSynchronize(st);
    cudaFree(l1);
    if (rc == SV_OK && e != cudaSuccess) rc = sv_fail(SV_ERR_CUDA, cudaGetErrorString(e));
    return rc;
}

} // namespace

namespace {
// K4c eligibility of a MODEL at level 2: chain + specials structure (K0, lazily), emission table fits one CTA's shared memory.
bool chain_pair_model_ok(sv_model* m) {
    if (m->chain_pair_state != 0) return m->chain_pair_state > 0;
    m->chain_pair_state = -1;
    if (!sv_model_ensure_chain(m)) return false;
    const size_t smem = sv_chain_pair_smem(m->chain.threads, static_cast<int>(m->emits));
    if (smem + 2048 > m->smem_optin) return false;
    int per_sm = 0;
    if (sv_chain_pair_prepare(m->chain.threads, m->chain.nonneg, smem, &per_sm) != 0) {
        cudaGetLastError();
        return false;
    }
    m->chain_pair_smem = smem;
    m->chain_pair_ctas_per_sm = per_sm;
    m->chain_pair_state = 1;
    return true;
}

// Builds a deferred level-2 table now (see sv_spec::deferred).
int ensure_table(sv_spec* s) {
    if (!s->deferred) return SV_OK;
    sv_model* m = s->m;
    SV_CUDA(cudaSetDevice(m->device));
    const size_t mat = static_cast<size_t>(m->n) * m->n;
    if (cudaMalloc(&s->d_table, s->ntuples * mat * sizeof(float)) != cudaSuccess) {
        cudaGetLastError();
        s->d_table = nullptr;
        return sv_fail(SV_ERR_MEMORY, "sv_spec: handler table does not fit device memory");
    }
    s->owns_table = true;
    cudaStream_t st = static_cast<cudaStream_t>(s->stream_at_create);
    SvPairDev pairdev;
    int rc = s->method == SV_SPEC_PAIR_ROWS && pair_prepare(m, &pairdev) ? build_pair_level2(m, pairdev, 0, s->ntuples, s->d_table, st)
                                                                        : build_sparse(m, s->level, 0, s->ntuples, s->d_table, st);
    if (rc == SV_OK && cudaStreamSynchronize(st) != cudaSuccess) rc = sv_fail(SV_ERR_CUDA, "sv_spec: precompute failed");
    if (rc != SV_OK) {
        cudaFree(s->d_table);
        s->d_table = nullptr;
        s->owns_table = false;
        return rc;
    }
    s->deferred = false;
    return SV_OK;
}
} // namespace

extern "C" int sv_spec_create(sv_model* m, uint64_t level, int method, uint64_t tuple_begin,
                              uint64_t tuple_end, float* d_table, void* stream, sv_spec** out) {
    SV_API_BEGIN
    if (out == nullptr) return sv_fail(SV_ERR_ARG, "sv_spec_create: out is NULL");
    *out = nullptr;
    if (m == nullptr || level < 1) return sv_fail(SV_ERR_ARG, "sv_spec_create: bad argument");
    const uint64_t ntuples = ipow(m->emits, level);
    if (ntuples == ~0ull) return sv_fail(SV_ERR_UNSUPPORTED, "sv_spec_create: emits^level overflows");
    if (tuple_begin == 0 && tuple_end == 0) tuple_end = ntuples;
    if (tuple_begin > tuple_end || tuple_end > ntuples)
        return sv_fail(SV_ERR_ARG, "sv_spec_create: bad tuple range");
    if (method == SV_SPEC_DENSE_GEMM && level != 2 && level != 1)
        return sv_fail(SV_ERR_UNSUPPORTED, "dense (min,+) GEMM precompute is implemented for level 2");
    SV_CUDA(cudaSetDevice(m->device));
    SvPairDev pairdev;
    const size_t mat_elems = static_cast<size_t>(m->n) * m->n;
    if (method == SV_SPEC_AUTO && level >= 3 && d_table == nullptr && tuple_begin == 0 && tuple_end == ntuples) {
        // the whole level-L table is asked for: on-demand when it cannot fit beside the level-(L-1) table
        size_t free_b = 0, total_b = 0;
        if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess) {
            const double need = static_cast<double>(ntuples) * mat_elems * sizeof(float) * (1.0 + 1.0 / m->emits);
            if (need > 0.9 * static_cast<double>(free_b)) method = SV_SPEC_ONDEMAND;
        }
    }
    if (method == SV_SPEC_ONDEMAND) {
        if (level < 3) {
            method = SV_SPEC_AUTO; // levels 1 and 2 have nothing to defer (level 2 runs implicitly, K4i)
        } else if (d_table != nullptr || !((tuple_begin == 0 && tuple_end == ntuples) || tuple_begin == tuple_end)) {
            return sv_fail(SV_ERR_ARG, "sv_spec_create: the on-demand mode has no level-L table (no tuple shards, no caller table)");
        }
    }
    // AUTO, the whole level-2 table, library-owned, chain + specials model: the run does not read it (K4c) -- defer it
    static const bool lazy_on = [] {
        const char* e = std::getenv("SV_SPEC_LAZY_TABLE");
        return !(e != nullptr && e[0] == '0');
    }();
    const bool defer = lazy_on && method == SV_SPEC_AUTO && level == 2 && d_table == nullptr && tuple_begin == 0 &&
                       tuple_end == ntuples && chain_pair_model_ok(m);
    if (method == SV_SPEC_AUTO) method = (level == 2 && pair_prepare(m, &pairdev)) ? SV_SPEC_PAIR_ROWS : SV_SPEC_SPARSE_LHS;
    if (method != SV_SPEC_SPARSE_LHS && method != SV_SPEC_DENSE_GEMM && method != SV_SPEC_PAIR_ROWS && method != SV_SPEC_ONDEMAND)
        return sv_fail(SV_ERR_ARG, "sv_spec_create: unknown method");
    if (method == SV_SPEC_PAIR_ROWS && (level != 2 || !pair_prepare(m, &pairdev)))
        return sv_fail(SV_ERR_UNSUPPORTED, "sv_spec_create: the pair-rows kernel builds level 2 of models with few dense rows");
    sv_spec* s = new (std::nothrow) sv_spec;
    if (s == nullptr) return sv_fail(SV_ERR_MEMORY, "sv_spec_create: out of host memory");
    s->m = m;
    s->level = level;
    s->ntuples = ntuples;
    s->method = method;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (method == SV_SPEC_ONDEMAND) {
        // the complete level-(L-1) table; level-L handlers are built per chunk by sv_spec_run*
        const uint64_t nprev = ipow(m->emits, level - 1);
        s->ondemand = true;
        s->t_begin = s->t_end = s->table_first = ntuples; // no level-L table
        if (cudaMalloc(&s->d_prev, nprev * mat_elems * sizeof(float)) != cudaSuccess) {
            cudaGetLastError();
            delete s;
            return sv_fail(SV_ERR_MEMORY, "sv_spec_create: the level-(L-1) table does not fit device memory");
        }
        int rc = (level == 3 && pair_prepare(m, &pairdev)) ? build_pair_level2(m, pairdev, 0, nprev, s->d_prev, st)
                                                          : build_sparse(m, level - 1, 0, nprev, s->d_prev, st);
        if (rc == SV_OK && cudaStreamSynchronize(st) != cudaSuccess) rc = sv_fail(SV_ERR_CUDA, "sv_spec_create: precompute failed");
        if (rc != SV_OK) {
            sv_spec_destroy(s);
            return rc;
        }
        *out = s;
        return SV_OK;
    }
    s->stream_at_create = stream;
    if (defer) {
        s->table_first = s->t_begin = 0;
        s->t_end = ntuples; // the logical range; the handlers appear on first use (ensure_table)
        s->deferred = true;
        *out = s;
        return SV_OK;
    }
    if (level > 1) {
        const size_t mat = static_cast<size_t>(m->n) * m->n;
        s->table_first = tuple_begin;
        s->t_begin = tuple_begin;
        s->t_end = tuple_end;
        if (d_table != nullptr) {
            s->d_table = d_table;
        } else if (tuple_end > tuple_begin) {
            if (cudaMalloc(&s->d_table, (tuple_end - tuple_begin) * mat * sizeof(float)) != cudaSuccess) {
                cudaGetLastError();
                delete s;
                return sv_fail(SV_ERR_MEMORY, "sv_spec_create: handler table does not fit device memory");
            }
            s->owns_table = true;
        }
        int rc = SV_OK;
        if (tuple_end > tuple_begin) {
            rc = method == SV_SPEC_DENSE_GEMM  ? build_dense_level2(m, tuple_begin, tuple_end, s->d_table, st)
                 : method == SV_SPEC_PAIR_ROWS ? build_pair_level2(m, pairdev, tuple_begin, tuple_end, s->d_table, st)
                                               : build_sparse(m, level, tuple_begin, tuple_end, s->d_table, st);
        }
        if (rc != SV_OK) {
            sv_spec_destroy(s);
            return rc;
        }
    }
    *out = s;
    return SV_OK;
    SV_API_END
}

extern "C" void sv_spec_destroy(sv_spec* s) {
    if (s == nullptr) return;
    cudaSetDevice(s->m->device);
    if (s->owns_table && s->d_table) cudaFree(s->d_table);
    for (int g = 0; g < sv_spec::kRunStreams - 1; ++g) {
        if (s->aux[g]) cudaStreamDestroy(s->aux[g]);
        if (s->ev_join[g]) cudaEventDestroy(s->ev_join[g]);
    }
    if (s->ev_fork) cudaEventDestroy(s->ev_fork);
    if (s->ev_done) cudaEventDestroy(s->ev_done);
    if (s->d_prev) cudaFree(s->d_prev);
    if (s->d_scratch) cudaFree(s->d_scratch);
    if (s->d_v[0]) cudaFree(s->d_v[0]);
    if (s->d_v[1]) cudaFree(s->d_v[1]);
    delete s;
}

extern "C" uint64_t sv_spec_level(const sv_spec* s) { return s ? s->level : 0; }

extern "C" int sv_spec_set_range(sv_spec* s, float* d_table, uint64_t table_first, uint64_t tuple_begin,
                                 uint64_t tuple_end) {
    SV_API_BEGIN
    if (s == nullptr || tuple_begin > tuple_end || tuple_end > s->ntuples || tuple_begin < table_first)
        return sv_fail(SV_ERR_ARG, "sv_spec_set_range: bad argument (need table_first <= tuple_begin <= tuple_end <= emits^level)");
    if (s->level < 2) return sv_fail(SV_ERR_ARG, "sv_spec_set_range: level 1 has no handler table");
    if (s->owns_table || s->deferred) return sv_fail(SV_ERR_ARG, "sv_spec_set_range: table is library-owned");
    if (d_table == nullptr && tuple_end > tuple_begin) return sv_fail(SV_ERR_ARG, "sv_spec_set_range: NULL table");
    s->d_table = d_table;
    s->table_first = table_first;
    s->t_begin = tuple_begin;
    s->t_end = tuple_end;
    return SV_OK;
    SV_API_END
}

extern "C" float* sv_spec_table(const sv_spec* s, uint64_t* tuple_begin, uint64_t* tuple_end) {
    if (s == nullptr) return nullptr;
    try {
        if (ensure_table(const_cast<sv_spec*>(s)) != SV_OK) return nullptr; // deferred level-2 table: built on first use
    } catch (...) {
        return nullptr;
    }
    if (tuple_begin) *tuple_begin = s->t_begin;
    if (tuple_end) *tuple_end = s->t_end;
    return s->d_table;
}

extern "C" int sv_spec_get_handler(const sv_spec* s, uint64_t tuple_index, float* out) {
    SV_API_BEGIN
    if (s == nullptr || out == nullptr) return sv_fail(SV_ERR_ARG, "sv_spec_get_handler: NULL");
    if (s->level < 2 || tuple_index < s->t_begin || tuple_index >= s->t_end)
        return sv_fail(SV_ERR_ARG, "sv_spec_get_handler: tuple not in the table");
    if (const int rc = ensure_table(const_cast<sv_spec*>(s)); rc != SV_OK) return rc;
    const size_t mat = static_cast<size_t>(s->m->n) * s->m->n;
    SV_CUDA(cudaSetDevice(s->m->device));
    SV_CUDA(cudaMemcpy(out, s->d_table + (tuple_index - s->table_first) * mat, mat * sizeof(float),
                       cudaMemcpyDeviceToHost));
    return SV_OK;
    SV_API_END
}

// ---------------------------------------------------------------- persistence ----
// Handler tables on disk (the reference rebuilds every table in each spec_with call,
// GraphBLAS_spec_impl.cpp:45-48).  File = 64-byte header + the dense handlers of [t_begin, t_end) in tuple
// order, row = destination state.  The header carries a fingerprint of the model the table was built
// from, so a table is never attached to a different model.
namespace {

constexpr char kTableMagic[8] = {'S', 'V', 'H', 'T', '0', '0', '0', '1'};
constexpr size_t kIoChunk = 64ull << 20; // bytes per pinned staging buffer (two of them, double buffered)

struct TableHeader {
    char magic[8];
    uint64_t states, emits, level, t_begin, t_end, fingerprint, reserved;
};
static_assert(sizeof(TableHeader) == 64, "table header layout");

uint64_t fnv1a(uint64_t h, const void* p, size_t bytes) {
    const unsigned char* b = static_cast<const unsigned char*>(p);
    for (size_t k = 0; k < bytes; ++k) h = (h ^ b[k]) * 1099511628211ull;
    return h;
}

// Every bit that enters a handler: T' (CSR, FIRST-wins duplicates already applied), emissions, dims.
uint64_t model_fingerprint(const sv_model* m) {
    uint64_t h = 1469598103934665603ull;
    const uint64_t dims[3] = {m->n, m->emits, m->nnz};
    h = fnv1a(h, dims, sizeof(dims));
    h = fnv1a(h, m->csr.rowptr.data(), m->csr.rowptr.size() * sizeof(uint32_t));
    h = fnv1a(h, m->csr.col.data(), m->csr.col.size() * sizeof(uint32_t));
    h = fnv1a(h, m->csr.w.data(), m->csr.w.size() * sizeof(float));
    h = fnv1a(h, m->emis.data(), m->emis.size() * sizeof(float));
    return h;
}

struct PinnedPair {
    void* p[2] = {nullptr, nullptr};
    cudaEvent_t ev[2] = {nullptr, nullptr};
    cudaStream_t st = nullptr;
    ~PinnedPair() {
        for (int k = 0; k < 2; ++k) {
            if (p[k]) cudaFreeHost(p[k]);
            if (ev[k]) cudaEventDestroy(ev[k]);
        }
        if (st) cudaStreamDestroy(st);
    }
    bool init() {
        if (cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking) != cudaSuccess) return false;
        for (int k = 0; k < 2; ++k)
            if (cudaMallocHost(&p[k], kIoChunk) != cudaSuccess ||
                cudaEventCreateWithFlags(&ev[k], cudaEventDisableTiming) != cudaSuccess)
                return false;
        return true;
    }
};

} // namespace

extern "C" int sv_spec_save(const sv_spec* s, const char* path) {
    SV_API_BEGIN
    if (s == nullptr || path == nullptr) return sv_fail(SV_ERR_ARG, "sv_spec_save: NULL");
    if (s->level < 2) return sv_fail(SV_ERR_ARG, "sv_spec_save: level 1 has no handler table (K2 forms E+T' on the fly)");
    if (const int rc = ensure_table(const_cast<sv_spec*>(s)); rc != SV_OK) return rc; // deferred level-2 table
    const sv_model* m = s->m;
    SV_CUDA(cudaSetDevice(m->device));
    SV_CUDA(cudaDeviceSynchronize()); // the table may still be under construction on a caller's stream
    PinnedPair io;
    if (!io.init()) {
        cudaGetLastError();
        return sv_fail(SV_ERR_MEMORY, "sv_spec_save: no pinned staging memory");
    }
    FILE* f = std::fopen(path, "wb");
    if (f == nullptr) return sv_fail(SV_ERR_IO, std::string("sv_spec_save: cannot open ") + path);
    TableHeader h{};
    std::memcpy(h.magic, kTableMagic, 8);
    h.states = m->n; h.emits = m->emits; h.level = s->level;
    h.t_begin = s->t_begin; h.t_end = s->t_end; h.fingerprint = model_fingerprint(m);
    bool ok = std::fwrite(&h, sizeof(h), 1, f) == 1;
    const size_t mat = static_cast<size_t>(m->n) * m->n * sizeof(float);
    const char* src = reinterpret_cast<const char*>(s->d_table) + (s->t_begin - s->table_first) * mat;
    const size_t total = (s->t_end - s->t_begin) * mat;
    // D2H of chunk k+1 overlaps the fwrite of chunk k
    size_t issued = 0, written = 0;
    int k = 0;
    auto issue = [&](int b) {
        const size_t len = std::min(kIoChunk, total - issued);
        cudaMemcpyAsync(io.p[b], src + issued, len, cudaMemcpyDeviceToHost, io.st);
        cudaEventRecord(io.ev[b], io.st);
        issued += len;
    };
    if (total > 0) issue(0);
    while (ok && written < total) {
        const int b = k & 1;
        if (issued < total) issue(b ^ 1);
        if (cudaEventSynchronize(io.ev[b]) != cudaSuccess) { ok = false; break; }
        const size_t len = std::min(kIoChunk, total - written);
        ok = std::fwrite(io.p[b], 1, len, f) == len;
        written += len;
        ++k;
    }
    cudaStreamSynchronize(io.st);
    ok = (std::fclose(f) == 0) && ok;
    if (cudaGetLastError() != cudaSuccess) ok = false;
    if (!ok) {
        std::remove(path);
        return sv_fail(SV_ERR_IO, std::string("sv_spec_save: write failed: ") + path);
    }
    return SV_OK;
    SV_API_END
}

extern "C" int sv_spec_load(sv_model* m, const char* path, float* d_table, sv_spec** out) {
    SV_API_BEGIN
    if (out == nullptr) return sv_fail(SV_ERR_ARG, "sv_spec_load: out is NULL");
    *out = nullptr;
    if (m == nullptr || path == nullptr) return sv_fail(SV_ERR_ARG, "sv_spec_load: NULL");
    FILE* f = std::fopen(path, "rb");
    if (f == nullptr) return sv_fail(SV_ERR_IO, std::string("sv_spec_load: cannot open ") + path);
    TableHeader h{};
    auto fail = [&](int code, const std::string& msg) {
        std::fclose(f);
        return sv_fail(code, "sv_spec_load: " + msg + ": " + path);
    };
    if (std::fread(&h, sizeof(h), 1, f) != 1 || std::memcmp(h.magic, kTableMagic, 8) != 0)
        return fail(SV_ERR_IO, "not a handler-table file");
    if (h.states != m->n || h.emits != m->emits || h.fingerprint != model_fingerprint(m))
        return fail(SV_ERR_ARG, "table was built from a different model");
    const uint64_t ntuples = h.level >= 2 ? ipow(m->emits, h.level) : ~0ull;
    if (ntuples == ~0ull || h.t_begin > h.t_end || h.t_end > ntuples) return fail(SV_ERR_IO, "corrupt header");
    const size_t mat = static_cast<size_t>(m->n) * m->n * sizeof(float);
    const size_t total = (h.t_end - h.t_begin) * mat;
    if (cudaSetDevice(m->device) != cudaSuccess) return fail(SV_ERR_CUDA, "cudaSetDevice");
    PinnedPair io;
    if (!io.init()) {
        cudaGetLastError();
        return fail(SV_ERR_MEMORY, "no pinned staging memory");
    }
    sv_spec* s = new (std::nothrow) sv_spec;
    if (s == nullptr) return fail(SV_ERR_MEMORY, "out of host memory");
    s->m = m;
    s->level = h.level;
    s->ntuples = ntuples;
    s->method = SV_SPEC_SPARSE_LHS;
    s->table_first = s->t_begin = h.t_begin;
    s->t_end = h.t_end;
    if (d_table != nullptr) {
        s->d_table = d_table;
    } else if (total > 0) {
        if (cudaMalloc(&s->d_table, total) != cudaSuccess) {
            cudaGetLastError();
            delete s;
            return fail(SV_ERR_MEMORY, "handler table does not fit device memory");
        }
        s->owns_table = true;
    }
    // fread of chunk k+1 overlaps the H2D of chunk k
    char* dst = reinterpret_cast<char*>(s->d_table);
    size_t done = 0;
    bool ok = true;
    for (int k = 0; ok && done < total; ++k) {
        const int b = k & 1;
        if (k >= 2 && cudaEventSynchronize(io.ev[b]) != cudaSuccess) { ok = false; break; }
        const size_t len = std::min(kIoChunk, total - done);
        if (std::fread(io.p[b], 1, len, f) != len) { ok = false; break; }
        cudaMemcpyAsync(dst + done, io.p[b], len, cudaMemcpyHostToDevice, io.st);
        cudaEventRecord(io.ev[b], io.st);
        done += len;
    }
    if (cudaStreamSynchronize(io.st) != cudaSuccess || cudaGetLastError() != cudaSuccess) ok = false;
    if (!ok) {
        sv_spec_destroy(s);
        return fail(SV_ERR_IO, "short or unreadable file");
    }
    std::fclose(f);
    *out = s;
    return SV_OK;
    SV_API_END
}

namespace {
// light rows per thread of K4i (register blocking of the heavy streams); SV_IMPLICIT_ROWS = 1 | 2 | 4
int implicit_rows_per_thread() {
    const char* e = std::getenv("SV_IMPLICIT_ROWS");
    const int r = e ? std::atoi(e) : 2;
    return r == 1 || r == 4 ? r : 2;
}

// threads per CTA of K4i; SV_IMPLICIT_THREADS = 64 | 128
int implicit_threads() {
    const char* e = std::getenv("SV_IMPLICIT_THREADS");
    return e && std::atoi(e) == 64 ? 64 : kImplicitThreads;
}

// sequence groups (side streams) of K4i; SV_IMPLICIT_GROUPS = 1..4
int implicit_groups() {
    const char* e = std::getenv("SV_IMPLICIT_GROUPS");
    // one grid of 21 x 50 CTAs is 14 CTAs more than one resident wave (7 per SM): 93 ms; four groups on
    // four streams fill each other's tails: 76 ms
    const int g = e ? std::atoi(e) : sv_spec::kRunStreams;
    return g >= 1 && g <= sv_spec::kRunStreams ? g : sv_spec::kRunStreams;
}

size_t implicit_smem(const SvPairDev& pd) {
    return static_cast<size_t>(1 + pd.n_stage) * ((pd.n + 15) / 16 * 16) * sizeof(float);
}

// K4i applies to level 2 of models with at most kImplicitMaxHeavy dense rows whose staged copies fit
// the shared memory of one CTA.
bool implicit_supported(sv_spec* s, SvPairDev* pd) {
    if (s->level != 2 || !pair_prepare(s->m, pd) || pd->n_heavy > kImplicitMaxHeavy || !path_prepare(s->m, pd)) return false;
    const size_t bytes = implicit_smem(*pd);
    if (bytes > 200 * 1024) return false;
    return cudaFuncSetAttribute(implicit_pair_step_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(bytes)) == cudaSuccess &&
           cudaFuncSetAttribute(implicit_pair_step_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(bytes)) == cudaSuccess &&
           cudaFuncSetAttribute(implicit_pair_step_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(bytes)) == cudaSuccess;
}
} // namespace

namespace {
// K4c eligibility: level 2, chain + specials structure (K0, lazily), emission table fits one CTA's shared memory.
bool chain_pair_supported(sv_spec* s) {
    if (s->level != 2 || s->ondemand) return false;
    return chain_pair_model_ok(s->m);
}
} // namespace

extern "C" int sv_spec_set_run_mode(sv_spec* s, int mode) {
    SV_API_BEGIN
    if (s == nullptr || (mode != SV_SPEC_RUN_AUTO && mode != SV_SPEC_RUN_TABLE && mode != SV_SPEC_RUN_IMPLICIT &&
                         mode != SV_SPEC_RUN_CHAIN))
        return sv_fail(SV_ERR_ARG, "sv_spec_set_run_mode: bad argument");
    if (mode == SV_SPEC_RUN_CHAIN && s->level > 1) {
        SV_CUDA(cudaSetDevice(s->m->device));
        if (!chain_pair_supported(s))
            return sv_fail(SV_ERR_UNSUPPORTED,
                           "sv_spec_set_run_mode: chain run needs level 2, the chain + specials structure and an emission "
                           "table that fits shared memory");
    }
    if (mode == SV_SPEC_RUN_IMPLICIT && s->level > 1) {
        SV_CUDA(cudaSetDevice(s->m->device));
        SvPairDev pd;
        if (!implicit_supported(s, &pd)) {
            cudaGetLastError();
            return sv_fail(SV_ERR_UNSUPPORTED, "sv_spec_set_run_mode: implicit run needs level 2 and at most 4 dense rows");
        }
    }
    s->run_mode = mode;
    return SV_OK;
    SV_API_END
}

extern "C" int sv_spec_run_device(sv_spec* s, const sv_seqs* seqs, float* d_out, float* d_best,
                                  int32_t* d_best_state, void* stream) {
    SV_API_BEGIN
    if (s == nullptr || seqs == nullptr) return sv_fail(SV_ERR_ARG, "sv_spec_run_device: NULL");
    sv_model* m = s->m;
    if (s->level == 1) return sv_viterbi_device(m, seqs, d_out, d_best, d_best_state, stream);
    if (seqs->device != m->device)
        return sv_fail(SV_ERR_ARG, "sv_spec_run_device: sequences live on another device");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    SV_CUDA(cudaSetDevice(m->device));
    // AUTO: chain + specials models run all their chunks in ONE persistent launch (K4c: identical bits, O(N) per chunk)
    if (s->run_mode == SV_SPEC_RUN_CHAIN || s->run_mode == SV_SPEC_RUN_AUTO) {
        const bool chain = seqs->dev.sym_bytes == 1 && chain_pair_supported(s);
        if (!chain && s->run_mode == SV_SPEC_RUN_CHAIN)
            return sv_fail(SV_ERR_UNSUPPORTED, "sv_spec_run: chain run needs level 2, the chain + specials structure, u8 symbols");
        if (chain) {
            // a private work-queue counter per launch, as in sv_viterbi_device: nothing else is shared between runs
            unsigned* counter = m->d_counters + static_cast<size_t>(m->launch_seq++ % SV_COUNTER_RING) * 16;
            SV_CUDA(cudaMemsetAsync(counter, 0, sizeof(unsigned), st));
            const int64_t slots = static_cast<int64_t>(std::max(1, m->sm_count)) * std::max(1, m->chain_pair_ctas_per_sm);
            const int grid = static_cast<int>(std::max<int64_t>(1, std::min<int64_t>(slots, static_cast<int64_t>(seqs->nseq))));
            const int rc = sv_chain_pair_launch(m->chain.threads, m->chain.nonneg, m->chaindev, seqs->dev, d_out, d_best,
                                                d_best_state, counter, grid, m->chain_pair_smem, st);
            if (rc != 0) return sv_fail(SV_ERR_CUDA, std::string("chain pair kernel launch: ") + cudaGetErrorString(static_cast<cudaError_t>(rc)));
            m->last_launches = 1;
            return SV_OK;
        }
    }
    SvPairDev pd;
    // AUTO: the implicit run whenever the model is eligible (identical bits, ~5x faster, no table traffic)
    bool implicit = !s->ondemand && s->run_mode != SV_SPEC_RUN_TABLE && implicit_supported(s, &pd);
    if (s->run_mode == SV_SPEC_RUN_IMPLICIT && !implicit)
        return sv_fail(SV_ERR_UNSUPPORTED, "sv_spec_run: implicit run needs level 2 and at most 4 dense rows");
    if (!implicit) cudaGetLastError();
    if (!implicit && !s->ondemand)
        if (const int rc = ensure_table(s); rc != SV_OK) return rc; // deferred level-2 table: the table run needs it
    if (!implicit && !s->ondemand && (s->t_begin != 0 || s->t_end != s->ntuples))
        return sv_fail(SV_ERR_ARG, "sv_spec_run: the handler table is not complete on this device");
    // runs on one spec share its state vectors: order this run after the previous one, whatever its stream
    if (s->ev_done == nullptr) SV_CUDA(cudaEventCreateWithFlags(&s->ev_done, cudaEventDisableTiming));
    else SV_CUDA(cudaStreamWaitEvent(st, s->ev_done, 0));
    const size_t n = m->n;
    // row stride of the state vectors: the implicit run fetches rows with bulk copies (64-byte rows, +inf padded)
    const size_t npad_rows = (n + 15) / 16 * 16;
    const size_t ld = npad_rows;
    const uint64_t nseq = seqs->nseq;
    if (s->v_capacity < nseq) {
        for (int b = 0; b < 2; ++b) {
            if (s->d_v[b]) cudaFree(s->d_v[b]);
            s->d_v[b] = nullptr;
            SV_CUDA(cudaMalloc(&s->d_v[b], nseq * npad_rows * sizeof(float)));
        }
        s->v_capacity = nseq;
    }
    RunSeqs rs{static_cast<int>(nseq), seqs->dev.sym, seqs->dev.sym_bytes, seqs->dev.off};
    int launches = 0;
    int cur = 0;
    {
        dim3 grid(static_cast<unsigned>(nseq), static_cast<unsigned>(std::min<size_t>((ld + 255) / 256, 65535)));
        spec_init_kernel<<<grid, 256, 0, st>>>(m->gen, rs, s->d_v[cur], static_cast<int>(ld));
        SV_CUDA(cudaGetLastError()); // a failed first launch must not let the chunk kernels run on garbage
        ++launches;
    }
    uint64_t max_chunks = (seqs->max_len - 1) / s->level;
    // rows per CTA: at least ~6 waves of resident CTAs (8 x 256 threads per SM), so that the last,
    // partial wave costs a few percent instead of the ~20 % it cost with 64 rows at 50 sequences
    int rows_per_cta = 64;
    {
        const uint64_t slots = static_cast<uint64_t>(std::max(1, m->sm_count)) * 8ull;
        while (rows_per_cta > 8 && ((n + rows_per_cta - 1) / rows_per_cta) * nseq < 6ull * slots) rows_per_cta /= 2;
        while ((n + rows_per_cta - 1) / rows_per_cta > 65535) rows_per_cta *= 2;
    }
    // v is staged in dynamic shared memory (n floats): opt in above 48 KB
    if (n * sizeof(float) > 48 * 1024) {
        SV_CUDA(cudaFuncSetAttribute(handler_gemv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     static_cast<int>(n * sizeof(float))));
        SV_CUDA(cudaFuncSetAttribute(tail_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     static_cast<int>(n * sizeof(float))));
    }
    // sequence groups on up to kRunStreams streams (group 0 stays on the caller's stream)
    int ngroups = static_cast<int>(std::min<uint64_t>(sv_spec::kRunStreams, nseq));
    if (implicit) ngroups = std::min(ngroups, implicit_groups());
    if (ngroups > 1 && s->ev_fork == nullptr) {
        SV_CUDA(cudaEventCreateWithFlags(&s->ev_fork, cudaEventDisableTiming));
        for (int g = 0; g < sv_spec::kRunStreams - 1; ++g) {
            SV_CUDA(cudaStreamCreateWithFlags(&s->aux[g], cudaStreamNonBlocking));
            SV_CUDA(cudaEventCreateWithFlags(&s->ev_join[g], cudaEventDisableTiming));
        }
    }
    if (ngroups > 1 && max_chunks > 0) {
        SV_CUDA(cudaEventRecord(s->ev_fork, st));
        for (int g = 1; g < ngroups; ++g) SV_CUDA(cudaStreamWaitEvent(s->aux[g - 1], s->ev_fork, 0));
    }
    if (s->ondemand && max_chunks > 0) {
        // On-demand levels >= 3: per chunk, K3' builds the handler each sequence needs from the level-(L-1) table into
        // a scratch (one handler per sequence of the group), then K4 applies it.  Groups bound the scratch to ~8 GB.
        const size_t mat = n * n;
        const uint64_t gmax = std::max<uint64_t>(1, std::min<uint64_t>(nseq, (8ull << 30) / (mat * sizeof(float))));
        if (s->scratch_cap < gmax) {
            if (s->d_scratch) cudaFree(s->d_scratch);
            s->d_scratch = nullptr;
            s->scratch_cap = 0;
            SV_CUDA(cudaMalloc(&s->d_scratch, gmax * mat * sizeof(float)));
            s->scratch_cap = gmax;
        }
        if (n > 65535) return sv_fail(SV_ERR_UNSUPPORTED, "sv_spec_run: on-demand mode supports up to 65535 states");
        OnDemandSeqs os{static_cast<int>(nseq), seqs->dev.sym, seqs->dev.sym_bytes, seqs->dev.off};
        for (uint64_t c = 0; c < max_chunks; ++c) {
            for (uint64_t q0 = 0; q0 < nseq; q0 += gmax) {
                const uint64_t q1 = std::min(nseq, q0 + gmax);
                dim3 bgrid(static_cast<unsigned>((n + 1023) / 1024), static_cast<unsigned>(n), static_cast<unsigned>(q1 - q0));
                ondemand_handler_kernel<<<bgrid, 256, 0, st>>>(m->gen, s->d_prev, s->d_scratch, os, c, static_cast<int>(s->level),
                                                               static_cast<int>(m->emits), static_cast<int>(q0));
                dim3 grid(static_cast<unsigned>(q1 - q0), static_cast<unsigned>((n + rows_per_cta - 1) / rows_per_cta));
                handler_gemv_kernel<<<grid, 256, n * sizeof(float), st>>>(
                    static_cast<int>(n), static_cast<int>(m->emits), static_cast<int>(s->level), rs, c, s->d_scratch, ~0ull,
                    s->d_v[cur], s->d_v[cur ^ 1], static_cast<int>(ld), rows_per_cta, static_cast<int>(q0));
                launches += 2;
            }
            cur ^= 1;
        }
        SV_CUDA(cudaGetLastError());
        max_chunks = 0;
    }
    for (uint64_t c = 0; c < max_chunks; ++c) {
        for (int g = 0; g < ngroups; ++g) {
            const uint64_t q0 = nseq * g / ngroups, q1 = nseq * (g + 1) / ngroups;
            if (implicit) {
                static const int rows = implicit_rows_per_thread();
                static const int threads = implicit_threads();
                const int per_cta = threads * rows;
                const int ntiles = static_cast<int>((n + per_cta - 1) / per_cta);
                dim3 igrid(static_cast<unsigned>(q1 - q0), static_cast<unsigned>(ntiles + pd.n_heavy));
                auto kern = rows == 1 ? implicit_pair_step_kernel<1> : rows == 2 ? implicit_pair_step_kernel<2> : implicit_pair_step_kernel<4>;
                kern<<<igrid, threads, implicit_smem(pd), g == 0 ? st : s->aux[g - 1]>>>(
                    pd, rs, c, s->d_v[cur], s->d_v[cur ^ 1], static_cast<int>(q0), ntiles);
                ++launches;
                continue;
            }
            dim3 grid(static_cast<unsigned>(q1 - q0), static_cast<unsigned>((n + rows_per_cta - 1) / rows_per_cta));
            handler_gemv_kernel<<<grid, 256, n * sizeof(float), g == 0 ? st : s->aux[g - 1]>>>(
                static_cast<int>(n), static_cast<int>(m->emits), static_cast<int>(s->level), rs, c,
                s->d_table, s->table_first, s->d_v[cur], s->d_v[cur ^ 1], static_cast<int>(ld), rows_per_cta, static_cast<int>(q0));
            ++launches;
        }
        cur ^= 1;
    }
    if (ngroups > 1 && max_chunks > 0) {
        for (int g = 1; g < ngroups; ++g) {
            SV_CUDA(cudaEventRecord(s->ev_join[g - 1], s->aux[g - 1]));
            SV_CUDA(cudaStreamWaitEvent(st, s->ev_join[g - 1], 0));
        }
    }
    for (uint64_t tl = 0; tl + 1 < s->level; ++tl) {
        tail_step_kernel<<<static_cast<unsigned>(nseq), 1024, n * sizeof(float), st>>>(
            m->gen, static_cast<int>(s->level), rs, static_cast<int>(tl), s->d_v[cur], s->d_v[cur ^ 1], static_cast<int>(ld));
        cur ^= 1;
        ++launches;
    }
    if (d_out != nullptr)
        SV_CUDA(cudaMemcpy2DAsync(d_out, n * sizeof(float), s->d_v[cur], ld * sizeof(float), n * sizeof(float), nseq,
                                  cudaMemcpyDeviceToDevice, st));
    if (d_best != nullptr || d_best_state != nullptr) {
        best_state_kernel<<<static_cast<unsigned>(nseq), 256, 0, st>>>(static_cast<int>(n), static_cast<int>(ld), s->d_v[cur],
                                                                      d_best, d_best_state);
        ++launches;
    }
    SV_CUDA(cudaGetLastError());
    SV_CUDA(cudaEventRecord(s->ev_done, st));
    m->last_launches = launches;
    return SV_OK;
    SV_API_END
}

namespace {
template <typename SrcT>
int spec_run_batch_impl(sv_spec* s, const SrcT* symbols, const uint64_t* offsets, uint64_t nseq,
                        float* out, float* best, int32_t* best_state) {
    if (s == nullptr) return sv_fail(SV_ERR_ARG, "sv_spec_run_batch: NULL spec");
    sv_model* m = s->m;
    std::lock_guard<std::mutex> lock(*static_cast<std::mutex*>(m->ws_mutex)); // one workspace per model
    int rc = sv_ws_upload(m, symbols, offsets, nseq);
    if (rc != SV_OK) return rc;
    float *d_out, *d_best;
    int32_t* d_arg;
    if ((rc = sv_ws_results(m, nseq, out != nullptr, &d_out, &d_best, &d_arg)) != SV_OK) return rc;
    rc = sv_spec_run_device(s, m->ws.seqs, d_out, best ? d_best : nullptr,
                            best_state ? d_arg : nullptr, m->stream);
    if (rc != SV_OK) return rc;
    return sv_ws_download(m, nseq, d_out, d_best, d_arg, out, best, best_state);
}
} // namespace

extern "C" int sv_spec_run_batch(sv_spec* s, const uint64_t* symbols, const uint64_t* offsets,
                                 uint64_t nseq, float* out, float* best, int32_t* best_state) {
    SV_API_BEGIN
    return spec_run_batch_impl(s, symbols, offsets, nseq, out, best, best_state);
    SV_API_END
}

extern "C" int sv_spec_run_batch_u8(sv_spec* s, const uint8_t* symbols, const uint64_t* offsets,
                                    uint64_t nseq, float* out, float* best, int32_t* best_state) {
    SV_API_BEGIN
    return spec_run_batch_impl(s, symbols, offsets, nseq, out, best, best_state);
    SV_API_END
}

extern "C" int sv_spec_run(sv_spec* s, const uint64_t* seq, uint64_t len, float* out) {
    SV_API_BEGIN
    if (seq == nullptr || out == nullptr) return sv_fail(SV_ERR_ARG, "sv_spec_run: NULL");
    const uint64_t offsets[2] = {0, len};
    return sv_spec_run_batch(s, seq, offsets, 1, out, nullptr, nullptr);
    SV_API_END
}

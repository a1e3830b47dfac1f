// Internal declarations shared by the translation units of libspecviterbi_b200.so.
// Nothing here is part of the public ABI (include/spec_viterbi_b200.h).
#pragma once

#include "spec_viterbi_b200.h"

#include <cuda_runtime.h>

#include <cstdint>
#include <string>
#include <vector>

// ---------------------------------------------------------------- errors ----
void sv_set_error(const std::string& msg);
int sv_fail(int code, const std::string& msg);

#define SV_CUDA(call)                                                                            \
    do {                                                                                         \
        cudaError_t sv_e_ = (call);                                                              \
        if (sv_e_ != cudaSuccess) {                                                              \
            return sv_fail(SV_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(sv_e_)); \
        }                                                                                        \
    } while (0)

// ------------------------------------------------------- generic kernel ----
// Model view for the generic persistent kernel: CSR of the transposed transition
// matrix T' (row = destination state), dense emission table and dense start column.
struct SvGenericDev {
    int n = 0;                  // states
    int emits = 0;
    int n_light = 0;            // rows processed one per thread (sorted by in-degree, descending)
    int n_heavy = 0;            // rows processed by the whole block
    const uint32_t* rowptr = nullptr; // [n+1]
    const uint32_t* col = nullptr;    // [nnz] source state
    const float* w = nullptr;         // [nnz] -log2 P(src->dst)
    const float* emis = nullptr;      // [emits][n]
    const float* start = nullptr;     // [n], +inf where no start entry
    const uint32_t* light_rows = nullptr;
    const uint32_t* heavy_rows = nullptr;
};

// Device-resident batch of observation sequences.
struct SvSeqsDev {
    int nseq = 0;
    int sym_bytes = 1;               // 1 = u8 symbols, 2 = u16
    const void* sym = nullptr;       // packed symbols, sequences concatenated in input order
    const uint64_t* off = nullptr;   // [nseq+1]
    const uint32_t* order = nullptr; // sequence ids, longest first (work-queue order)
};

// ------------------------------------------------ edge-stationary kernel ----
constexpr int SV_ES_RMAX = 5;      // light rows ("ranks") per thread
constexpr int SV_ES_NHMAX = 2;     // heavy rows (block-reduced)

// Compile-time variant of the edge-stationary kernel.
struct SvEsVariant {
    int threads;                 // threads per CTA (TPG)
    int q;                       // sequences advanced in lockstep per CTA (1 or 2)
    int shape[SV_ES_RMAX];       // max register slots per rank (staircase)
    int nh;                      // heavy rows supported (0 or SV_ES_NHMAX)
    int hs;                      // register slots per heavy row per thread
};

// Host-side schedule produced by K0 for one variant.
struct SvEsPlan {
    bool valid = false;
    SvEsVariant var{};
    int variant_index = -1;
    int n_int = 0;       // internal states (after in-degree capping by state splitting)
    int n_light = 0;     // internal states owned as light rows: p in [0, n_light)
    int n_heavy = 0;     // heavy rows: p in [n_light, n_int)
    int nslots = 0;      // sum(shape) + nh*hs
    int ranks = 0;       // ranks actually used
    uint64_t gather_wavefronts = 0, gather_ideal = 0;
    uint64_t smem_bytes = 0;
    uint64_t split_states = 0;
    // per-thread edge registers, slot-major: [nslots][threads]
    std::vector<float> edge_w;
    std::vector<uint32_t> edge_off; // byte offset of v[src] inside one v buffer: p_src * q * 4
    std::vector<uint8_t> deg;       // [warps][RMAX] slots to run per rank (warp-uniform)
    std::vector<uint8_t> heavy_slots; // [NHMAX] slots to run per heavy row (CTA-uniform)
    std::vector<float> emis_p;      // [emits][n_int] emission table in internal order
    std::vector<float> start_p;     // [n_int]
    std::vector<uint32_t> orig;     // [n_int] original state of internal state p
    std::vector<uint32_t> copy_ptr; // [n+1] internal copies of each original state
    std::vector<uint32_t> copy_p;   // [n_int]
};

struct SvEsDev {
    int n = 0, n_int = 0, n_light = 0, n_heavy = 0, emits = 0, nslots = 0;
    int nonneg = 0; // scores are >= 0: warp minima may use the integer REDUX unit
    const float* edge_w = nullptr;
    const uint32_t* edge_off = nullptr;
    const uint8_t* deg = nullptr;
    uint8_t heavy_slots[SV_ES_NHMAX] = {0, 0};
    const float* emis_p = nullptr;
    const float* start_p = nullptr;
    const uint32_t* copy_ptr = nullptr;
    const uint32_t* copy_p = nullptr;
};

int sv_es_num_variants();
const SvEsVariant& sv_es_variant(int index);
// Launch of variant `index`; returns cudaError_t as int.
int sv_es_launch(int index, const SvEsDev& m, const SvSeqsDev& s, float* d_out, float* d_best,
                 int32_t* d_arg, unsigned* d_counter, int grid, size_t smem, cudaStream_t st);
int sv_es_max_ctas_per_sm(int index, size_t smem);
int sv_es_prepare(int index, size_t smem); // sets the dynamic smem attribute

// ------------------------------------------------------ lockstep (LS) kernel ----
// Edge-stationary like the ES kernel, but 4 or 8 sequences advance in lockstep per CTA
// (score vectors interleaved [state][4] so one LDS.128 feeds four relaxations) and the
// emission table lives in TENSOR MEMORY (tcgen05.ld lookup), which frees the shared memory
// for the score vectors.  K0 colours the states so that the 8 lanes of every LDS.128 phase hit
// 8 distinct 16-byte bank groups.
constexpr int SV_LS_RSTRIDE = SV_ES_RMAX; // TMEM columns per symbol (one per rank) of the 5-rank variants
constexpr int SV_LS_RMAX = 10;            // rank capacity of any lockstep variant

struct SvLsVariant {
    int threads;                 // threads that share one set of rows (a CTA, or one HALF of a CTA)
    int nq4;                     // float4 groups of sequences per set: Q = 4 * nq4
    int shape[SV_LS_RMAX];       // max register slots per rank
    int nh;                      // heavy rows supported
    int hs;                      // register slots per heavy row per thread
    int ranks;                   // rows per thread (5 or 10)
    int halves;                  // independent thread sets per CTA (1, or 2 with their own barriers)
};

struct SvLsPlan {
    bool valid = false;
    SvLsVariant var{};
    int variant_index = -1;
    int n_int = 0;    // internal states (after splitting), light + heavy
    int n_lpad = 0;   // light positions = ranks * threads (holes included)
    int n_pad = 0;    // positions per score buffer (n_lpad + heavy, rounded up to 8)
    int n_heavy = 0, nslots = 0, ranks = 0, tmem_cols = 0;
    bool nonneg = false; // every finite score is >= 0 (or -0): float order == signed-int order
    uint64_t gather_wavefronts = 0, gather_ideal = 0, smem_bytes = 0, split_states = 0;
    uint64_t colour_cost_before = 0, colour_cost_after = 0; // excess bank-group loads (K0 search)
    std::vector<float> edge_w;       // [nslots][threads]
    std::vector<uint32_t> edge_off;  // [nslots][threads] byte offset of v[src] in a buffer: p * 16
    std::vector<uint8_t> deg;        // [warps][RMAX]
    std::vector<uint8_t> heavy_slots;
    std::vector<float> emis_p;       // [emits][n_pad]
    std::vector<float> start_p;      // [n_pad]
    std::vector<uint32_t> copy_ptr;  // [n+1]
    std::vector<uint32_t> copy_p;    // [n_int]
};

struct SvLsDev {
    int n = 0, n_int = 0, n_lpad = 0, n_pad = 0, n_heavy = 0, emits = 0, ranks = 0, tmem_cols = 0;
    int nonneg = 0; // scores are >= 0: warp minima may use the integer REDUX unit
    const float* edge_w = nullptr;
    const uint32_t* edge_off = nullptr;
    const uint8_t* deg = nullptr;    // [warps][variant ranks]
    uint8_t heavy_slots[SV_ES_NHMAX] = {0, 0};
    const float* emis_p = nullptr;
    const float* start_p = nullptr;
    const uint32_t* copy_ptr = nullptr;
    const uint32_t* copy_p = nullptr;
};

int sv_ls_num_variants();
const SvLsVariant& sv_ls_variant(int index);
int sv_ls_prepare(int index, size_t smem);
int sv_ls_launch(int index, const SvLsDev& m, const SvSeqsDev& s, float* d_out, float* d_best,
                 int32_t* d_arg, unsigned* d_counter, int grid, size_t smem, cudaStream_t st);

// ----------------------------------------------------------- host model ----
struct SvHostCsr {
    uint64_t n = 0;
    std::vector<uint32_t> rowptr, col;
    std::vector<float> w;
};

struct sv_seqs;
// Reusable buffers behind the host-pointer entry points (sv_viterbi*, sv_spec_run*).
struct SvWorkspace {
    sv_seqs* seqs = nullptr;
    float* d_out = nullptr;
    size_t out_cap = 0; // floats
    float* d_best = nullptr;
    int32_t* d_arg = nullptr;
    size_t best_cap = 0; // sequences
};

struct sv_model {
    int device = 0;
    int sm_count = 0;
    size_t smem_optin = 0;
    uint64_t n = 0, emits = 0, nnz = 0, max_deg = 0;
    SvHostCsr csr;
    std::vector<float> emis;  // [emits][n]
    std::vector<float> start; // [n]
    // generic kernel device data
    SvGenericDev gen;
    void* d_blob = nullptr; // single allocation backing all device arrays
    unsigned* d_counter = nullptr;
    // edge-stationary plan
    SvEsPlan es;
    SvEsDev esdev;
    void* d_es_blob = nullptr;
    // the same with ONE sequence per CTA: lower step latency, used when every sequence of a batch
    // gets its own resident CTA (the latency cases: a handful of long sequences)
    SvEsPlan es1;
    SvEsDev esdev1;
    void* d_es1_blob = nullptr;
    // lockstep plan
    SvLsPlan ls;
    SvLsDev lsdev;
    void* d_ls_blob = nullptr;
    // level-2 pair precompute (sv_spec.cu): dense copies of the heavy rows + CSC, built on first use
    void* d_pair_blob = nullptr;
    void* d_path_blob = nullptr; // K4i/K4p: flattened two-step paths (sv_spec.cu: path_prepare)
    int path_state = 0;          // 0 = not built, 1 = ready, -1 = not eligible
    int path_slots = 0, path_heavy_total = 0;
    unsigned path_used_mask = 0; // dense rows that some row has a finite edge from (only those are staged)
    int pair_state = 0; // 0 = not built, 1 = ready, -1 = model not eligible (too many heavy rows)
    int kernel_pref = SV_KERNEL_AUTO;
    int last_launches = 0;
    int last_kernel = 0;
    uint64_t last_nseq = 0;
    // scratch for host-buffer entry points
    cudaStream_t stream = nullptr;
    SvWorkspace ws;
};

struct sv_seqs {
    int device = 0;
    uint64_t nseq = 0, total = 0, max_len = 0;
    uint64_t h2d_bytes = 0; // bytes the last (re)fill copied to the device
    SvSeqsDev dev;
    void* d_blob = nullptr;
    size_t d_cap = 0;
    unsigned char* h_pinned = nullptr; // staging for offsets / order / repacked symbols
    size_t h_cap = 0;
};

// host-buffer workspace helpers (sv_api.cu)
template <typename SrcT>
int sv_ws_upload(sv_model* m, const SrcT* symbols, const uint64_t* offsets, uint64_t nseq);
int sv_ws_results(sv_model* m, uint64_t nseq, bool want_out, float** d_out, float** d_best,
                  int32_t** d_arg);
int sv_ws_download(sv_model* m, uint64_t nseq, const float* d_out, const float* d_best,
                   const int32_t* d_arg, float* out, float* best, int32_t* best_state);

// K0: build the edge-stationary schedule (sv_es_plan.cpp); plan.valid=false if no variant fits.
// force_q: 0 = prefer two sequences per CTA, 1 / 2 = only variants with that many.
void sv_es_build_plan(const SvHostCsr& csr, const std::vector<float>& emis,
                      const std::vector<float>& start, uint64_t emits, size_t smem_limit,
                      SvEsPlan* plan, int force_q = 0);

// K0 for the lockstep kernel (sv_ls_plan.cpp); plan.valid=false if the model does not fit.
void sv_ls_build_plan(const SvHostCsr& csr, const std::vector<float>& emis,
                      const std::vector<float>& start, uint64_t emits, size_t smem_limit,
                      SvLsPlan* plan);

// generic kernel launcher (sv_plain_generic.cu)
int sv_generic_launch(const SvGenericDev& m, const SvSeqsDev& s, float* d_out, float* d_best,
                      int32_t* d_arg, unsigned* d_counter, int grid, cudaStream_t st);
size_t sv_generic_smem(const SvGenericDev& m);

// Internal declarations shared by the translation units of libspecviterbi_b200.so.
// Nothing here is part of the public ABI (include/spec_viterbi_b200.h).
#pragma once

#include "spec_viterbi_b200.h"

#include <cuda_runtime.h>

#include <cstdint>
#include <exception>
#include <new>
#include <string>
#include <vector>

// ---------------------------------------------------------------- errors ----
void sv_set_error(const std::string& msg);
int sv_fail(int code, const std::string& msg);

// Every int-returning extern "C" entry point wraps its body in these: no C++ exception (std::bad_alloc /
// std::length_error from the std::vector-based planners and readers) may cross the C ABI.
#define SV_API_BEGIN try {
#define SV_API_END                                                                               \
    }                                                                                            \
    catch (const std::bad_alloc&) { return sv_fail(SV_ERR_MEMORY, "out of host memory"); }      \
    catch (const std::exception& sv_ex_) {                                                       \
        return sv_fail(SV_ERR_MEMORY, std::string("internal error: ") + sv_ex_.what());         \
    }                                                                                            \
    catch (...) { return sv_fail(SV_ERR_MEMORY, "internal error: unknown C++ exception"); }

#define SV_CUDA(call)                                                                            \
    do {                                                                                         \
        cudaError_t sv_e_ = (call);                                                              \
        if (sv_e_ != cudaSuccess) {                                                              \
            return sv_fail(SV_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(sv_e_)); \
        }                                                                                        \
    } while (0)

struct SvHostCsr {
    uint64_t n = 0;
    std::vector<uint32_t> rowptr, col;
    std::vector<float> w;
};

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
    // Set per launch by sv_viterbi_device (8-sequence CTAs only): queue positions >= tail_begin are handed out 4 per
    // CTA.  Sequences of equal length finish together, and a queue that gives 8 to every CTA that asks leaves most SMs
    // idle in the last wave (4096 sequences on 148 SMs: 68 CTAs would take 8, 80 none); when the remainder after the
    // full waves is at most 4 per CTA, every CTA takes 4 and skips its idle float4 group (about half the step time).
    unsigned tail_begin = 0xffffffffu;
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

// A variant = a register staircase; each is compiled for 8 (nq4 = 2) and 4 (nq4 = 1) sequences per
// CTA, and the launch picks nq4 from the batch size (the register images do not depend on it).
struct SvLsVariant {
    int threads;                 // threads per CTA
    int shape[SV_LS_RMAX];       // max register slots per rank
    int nh;                      // heavy rows supported
    int hs;                      // register slots per heavy row per thread
    int ranks;                   // rows per thread
};

// Knobs of one K0 run for the lockstep kernel.
struct SvLsPlanOpts {
    uint64_t sa_iters = 0;               // bank-colouring search length (0 = the full default)
    const volatile int* cancel = nullptr; // polled by the search: non-zero = give up (plan stays invalid)
    int threads = 1;                     // host threads for the per-octet slot assignment
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
    uint64_t gather_wavefronts = 0, gather_ideal = 0, split_states = 0;
    uint64_t smem_bytes = 0;  // dynamic shared memory of the 8-sequence launch (nq4 = 2)
    uint64_t smem_bytes4 = 0; // ... of the 4-sequence launch (nq4 = 1)
    uint64_t sa_iters = 0;    // length of the colouring search that produced this plan
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
int sv_ls_prepare(int index, int nq4, size_t smem);
int sv_ls_launch(int index, int nq4, const SvLsDev& m, const SvSeqsDev& s, float* d_out, float* d_best,
                 int32_t* d_arg, unsigned* d_counter, int grid, size_t smem, cudaStream_t st);

// ------------------------------------------------------------- chain kernel ----
// Register-resident step for "chain + specials" models -- the structure of every profile HMM the
// reference ships (chmm_files/silent_hmm_to_chmm.py:118-144): after removing at most two SPECIAL
// states (N and C: fed by / feeding every match state) the transition graph is a union of simple
// paths.  A thread owns SV_CHAIN_RPT consecutive chain positions of Q sequences IN REGISTERS; the
// band edge comes from the neighbouring register (or one shuffle), the hub state (special 0) is a
// broadcast, and a special row whose chain in-edges all carry one weight u needs only the minimum
// over the chain, because fl(c + x) is monotone:  min_k fl(fl(E+u) + v[k]) = fl(fl(E+u) + min_k v[k]).
// No per-relaxation shared-memory gather is left, so the kernel is FP32-issue bound.
constexpr int SV_CHAIN_RPT = 5;  // chain positions per thread (= TMEM columns per symbol)
constexpr int SV_CHAIN_NSP = 2;  // special states supported
constexpr int SV_CHAIN_TMEM_CS = SV_CHAIN_RPT + 1; // TMEM columns per symbol: fl(E + w_pred) of 5 positions, fl(E + w_hub) of one

struct SvChainPlan {
    bool valid = false;
    int threads = 0;   // CTA size: 128, 256 or 512
    int n_chain = 0;   // chain states (positions 0 .. n_chain-1; the rest up to n_pos are padding)
    int n_pos = 0;     // threads * SV_CHAIN_RPT
    int nsp = 0;       // special states
    int tmem_cols = 0;
    bool nonneg = false;
    uint32_t special[SV_CHAIN_NSP] = {0, 0};       // original ids
    float u[SV_CHAIN_NSP];                          // weight of every chain -> special s edge (+inf: none)
    float ss[SV_CHAIN_NSP][SV_CHAIN_NSP];           // ss[s][s2] = weight of special s2 -> special s
    float start_s[SV_CHAIN_NSP];
    std::vector<float> w_pred, w_hub;               // [n_pos] weight of (p-1) -> p and of special 0 -> p
    std::vector<float> emis_p, start_p;             // [emits][n_pos], [n_pos]
    std::vector<float> emis_s;                      // [NSP + 1][emits]: special states, then min_p fl(E[o][p] + w_hub[p])
    std::vector<int32_t> orig_of_pos;               // [n_pos] original state, -1 = padding
};

struct SvChainDev {
    int n = 0, n_pos = 0, emits = 0, nsp = 0, tmem_cols = 0, nonneg = 0;
    int special[SV_CHAIN_NSP] = {-1, -1};
    float u[SV_CHAIN_NSP], ss[SV_CHAIN_NSP][SV_CHAIN_NSP], start_s[SV_CHAIN_NSP];
    const float* w_pred = nullptr;
    const float* w_hub = nullptr;
    const float* emis_p = nullptr;
    const float* start_p = nullptr;
    const float* emis_s = nullptr;
    const int32_t* orig_of_pos = nullptr;
};

// K0 (sv_chain_plan.cpp): detects the structure; plan.valid = false if the model does not have it.
void sv_chain_build_plan(const SvHostCsr& csr, const std::vector<float>& emis, const std::vector<float>& start,
                         uint64_t emits, SvChainPlan* plan);
// sequences per CTA that are compiled: 8, 4, 2, 1
int sv_chain_prepare(int threads, int q, bool nonneg, int emits, size_t* smem_out, int* ctas_per_sm_out);
int sv_chain_launch(int threads, int q, const SvChainDev& m, const SvSeqsDev& s, float* d_out, float* d_best,
                    int32_t* d_arg, unsigned* d_counter, int grid, size_t smem, cudaStream_t st);

// K4c (sv_spec_chain.cu): the level-2 specialised run of a chain + specials model in one persistent launch
size_t sv_chain_pair_smem(int threads, int emits);
int sv_chain_pair_prepare(int threads, bool nonneg, size_t smem, int* ctas_per_sm_out);
int sv_chain_pair_launch(int threads, bool nonneg, const SvChainDev& m, const SvSeqsDev& s, float* d_out, float* d_best,
                         int32_t* d_arg, unsigned* d_counter, int grid, size_t smem, cudaStream_t st);

// ----------------------------------------------------------- host model ----
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

// Background refinement of the lockstep plan (sv_api.cu): the first batch runs on a short colouring
// search; the full-length search continues on a host thread and its plan is swapped in (and written
// to the plan cache) when it is ready.
struct SvLsRefine;

constexpr int SV_COUNTER_RING = 64; // work-queue counters, one per launch in flight

struct sv_model {
    int device = 0;
    int sm_count = 0;
    size_t smem_optin = 0;
    uint64_t n = 0, emits = 0, nnz = 0, max_deg = 0;
    uint64_t fingerprint = 0;
    SvHostCsr csr;
    std::vector<float> emis;  // [emits][n]
    std::vector<float> start; // [n]
    // generic kernel device data
    SvGenericDev gen;
    void* d_blob = nullptr; // single allocation backing all device arrays
    // Persistent kernels pull sequences from a global counter.  Every launch takes the next counter of
    // a ring, so launches of one model that overlap on different streams never share a queue.
    unsigned* d_counters = nullptr; // [SV_COUNTER_RING], 64 bytes apart
    unsigned launch_seq = 0;
    // K0 plans are built on first use (a level >= 2 spec_with never needs them)
    int es_state = 0, es1_state = 0, ls_state = 0, chain_state = 0; // 0 = not tried, 1 = ready, -1 = does not fit
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
    std::vector<void*> retired_blobs; // plans replaced by the background refinement (kernels may still read them)
    SvLsRefine* refine = nullptr;
    // chain plan (profile HMMs: sv_chain_plan.cpp / sv_plain_chain.cu)
    SvChainPlan chain;
    SvChainDev chaindev;
    void* d_chain_blob = nullptr;
    size_t chain_smem[4] = {0, 0, 0, 0}; // dynamic shared memory of the Q = 1, 2, 4, 8 instantiations
    int chain_ctas_per_sm = 1;
    int chain_pair_state = 0;    // K4c: 0 = not tried, 1 = ready, -1 = not eligible (no chain structure / table does not fit)
    size_t chain_pair_smem = 0;
    int chain_pair_ctas_per_sm = 1;
    // level-2 pair precompute (sv_spec.cu): dense copies of the heavy rows + CSC, built on first use
    void* d_pair_blob = nullptr;
    void* d_path_blob = nullptr; // K4i: flattened two-step paths (sv_spec.cu: path_prepare)
    int path_state = 0;          // 0 = not built, 1 = ready, -1 = not eligible
    int path_slots = 0, path_heavy_total = 0;
    size_t path_rh_offset = 0;   // byte offset of the per-symbol r_h tables inside d_path_blob
    unsigned path_used_mask = 0; // dense rows that some row has a finite edge from (only those are staged)
    int pair_state = 0; // 0 = not built, 1 = ready, -1 = model not eligible (too many heavy rows)
    int kernel_pref = SV_KERNEL_AUTO;
    int last_launches = 0;
    int last_kernel = 0;
    int last_seqs_per_cta = 0;
    uint64_t last_nseq = 0;
    double plan_ms = 0; // host time spent in K0 planners so far (reported by sv_model_get_info)
    // scratch for host-buffer entry points (single flight: the host-buffer calls of one model serialise)
    cudaStream_t stream = nullptr;
    SvWorkspace ws;
    void* ws_mutex = nullptr; // std::mutex
};

// Plans are lazy: these build (once) and upload the plan a kernel needs; false = the model does not fit it.
bool sv_model_ensure_es(sv_model* m);
bool sv_model_ensure_es1(sv_model* m);
bool sv_model_ensure_ls(sv_model* m);
bool sv_model_ensure_chain(sv_model* m);

struct sv_seqs {
    int device = 0;
    uint64_t nseq = 0, total = 0, max_len = 0, min_len = 0;
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
                      SvLsPlan* plan, const SvLsPlanOpts& opts = SvLsPlanOpts());
uint64_t sv_ls_default_sa_iters(uint64_t states);

// Plan cache (sv_plan_cache.cpp): the finished lockstep plan of a model, keyed by a fingerprint of
// T', the emissions and the start column, kept on disk across processes.
uint64_t sv_model_fingerprint(const SvHostCsr& csr, const std::vector<float>& emis,
                              const std::vector<float>& start, uint64_t emits);
bool sv_plan_cache_load(uint64_t fingerprint, SvLsPlan* plan);
void sv_plan_cache_store(uint64_t fingerprint, const SvLsPlan& plan);

// generic kernel launcher (sv_plain_generic.cu)
int sv_generic_launch(const SvGenericDev& m, const SvSeqsDev& s, float* d_out, float* d_best,
                      int32_t* d_arg, unsigned* d_counter, int grid, cudaStream_t st);
size_t sv_generic_smem(const SvGenericDev& m);

/* spec_viterbi_b200 -- C ABI of the B200-native (min,+) Viterbi backend.
 *
 * This is the drop-in boundary: a third GPU backend for IvanTyulyandin/Spec_Viterbi that
 * replaces the CUSP and cuASR backends behind the reference's `Viterbi_impl` /
 * `Viterbi_spec_impl` interface.  Every entry point is `extern "C"`, takes plain pointers
 * and sizes, returns an int status (0 = SV_OK) and never throws.  The thin C++ adaptors
 * `B200_impl` / `B200_spec_impl` (spec_viterbi_b200/host/) and the Python mirror
 * (spec_viterbi_b200/api.py, ctypes) sit directly on these functions.
 *
 * Reference interfaces replaced (file:line under the reference tree):
 *   sv_model_create      <- the per-call model build/upload of every backend
 *                           (GraphBLAS_impl.cpp:24-45, CUSP_helper.cu:32-72 init_matrices_from_HMM,
 *                            cuASR_helper.cu:167-225); HMM layout = Viterbi_impl/HMM.h:28-38
 *   sv_viterbi           <- Viterbi_impl::run_Viterbi (Viterbi_impl.h:8-9; GraphBLAS_impl.cpp:4-93,
 *                           CUSP_impl.cu:5-23, cuASR_impl.cu:6-24)
 *   sv_viterbi_batch     <- the harness loop "one run_Viterbi call per sequence"
 *                           (benchmark/bench_Viterbi.h:51-57), batched
 *   sv_spec_create       <- Viterbi_spec_impl::spec_with (Viterbi_spec_impl.h:11;
 *                           GraphBLAS_spec_impl.cpp:45-48,103-182, add_level :15-36;
 *                           CUSP_spec_impl.cu:81-96, cuASR_spec_impl.cu:81-107)
 *   sv_spec_run[_batch]  <- Viterbi_spec_impl::run_Viterbi_spec (Viterbi_spec_impl.h:13-14;
 *                           GraphBLAS_spec_impl.cpp:50-97) and bench_Viterbi_spec.h:74-80
 *   sv_io_*              <- read_HMM / read_emit_seq (Viterbi_impl/data_reader.h:8,11)
 *
 * Numerics contract: scores are fp32 -log2 probabilities, +inf = impossible.
 * Plain path and every spec level are bit-exact against the same-level GraphBLAS
 * formulation:  v'[i] = min_j fl(fl(E[o][i] + T'[i][j]) + v[j]).
 *
 * There is no CPU fallback: without a CUDA device every compute call fails with
 * SV_ERR_CUDA and a message in sv_last_error().
 */
#ifndef SPEC_VITERBI_B200_H
#define SPEC_VITERBI_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum {
    SV_OK = 0,
    SV_ERR_ARG = 1,     /* bad argument (null pointer, empty sequence, symbol >= E, ...) */
    SV_ERR_CUDA = 2,    /* CUDA runtime error, or no device */
    SV_ERR_MEMORY = 3,  /* host or device allocation failed / table does not fit */
    SV_ERR_IO = 4,      /* file could not be read / parsed */
    SV_ERR_UNSUPPORTED = 5
};

/* Thread-local message of the last failing call on this thread ("" if none). */
const char* sv_last_error(void);
/* Library version string, e.g. "spec_viterbi_b200 0.1 (sm_100a)". */
const char* sv_version(void);
/* Number of visible CUDA devices (0 on a CPU-only host; never fails). */
int sv_device_count(void);

/* ------------------------------------------------------------------ model ---- */
typedef struct sv_model sv_model;

/* Kernel selection for the plain / level-1 step (sv_model_set_kernel). */
enum {
    SV_KERNEL_AUTO = 0,
    SV_KERNEL_GENERIC = 1, /* CSR in L1/L2, any model size */
    SV_KERNEL_ES = 2,      /* edge-stationary: transitions in registers, E and v in smem */
    SV_KERNEL_LS = 3,      /* lockstep: 4/8 sequences per CTA, transitions in registers, v in
                              smem as float4, E in tensor memory (the batched throughput kernel) */
    SV_KERNEL_CHAIN = 4    /* chain + specials models (every profile HMM the reference ships): v of 1..8
                              sequences in REGISTERS, band edge from the neighbouring register, dense rows
                              through the chain minimum, E in tensor memory; FP32-issue bound */
};

/* Builds the device-resident model from the reference's HMM layout:
 * COO transitions (src = HMM::trans_rows, dst = HMM::trans_cols, w = -log2 p),
 * emissions [emits][states], sparse start vector.  Duplicate COO / start entries keep
 * the FIRST occurrence (GrB_FIRST_FP32).  `device` is a CUDA ordinal.
 * Cost: CSR build + one upload (about a millisecond at 2405 states).  The kernel-specific K0
 * schedules are built by the FIRST launch that needs them (chain < 1 ms, edge-stationary ~5 ms,
 * lockstep ~0.1 s with the full bank-colouring search continuing on a background host thread and its
 * result kept in an on-disk plan cache: $SV_PLAN_CACHE_DIR, default ~/.cache/spec_viterbi_b200,
 * SV_PLAN_CACHE=0 disables; SV_PLAN_REFINE=bg|sync|off).  A level >= 2 spec never builds them.
 *
 * Concurrency: a model may be used from several host threads / streams.  sv_viterbi_device launches
 * are independent (each takes its own work-queue counter); the host-buffer entry points
 * (sv_viterbi, sv_viterbi_batch*, sv_spec_run, sv_spec_run_batch*) share one workspace per model and
 * serialise on it; sv_spec_run_device calls on ONE spec are ordered on the device (they share the
 * spec's state vectors).  Creating / destroying handles is not thread-safe against their use. */
int sv_model_create(uint64_t states, uint64_t emits, uint64_t trans, const uint64_t* trans_src,
                    const uint64_t* trans_dst, const float* trans_w, const float* emissions,
                    uint64_t nstart, const uint64_t* start_cols, const float* start_w, int device,
                    sv_model** out);
void sv_model_destroy(sv_model* m);
int sv_model_set_kernel(sv_model* m, int kernel);
/* Builds every K0 schedule now and waits for the background refinement of the lockstep plan, so
 * that measurements taken afterwards see the final schedules (optional). */
int sv_model_prepare(sv_model* m);

typedef struct sv_model_info {
    uint64_t states, emits, nnz;   /* nnz after duplicate removal */
    uint64_t max_in_degree;
    int32_t device;
    int32_t kernel;                /* SV_KERNEL_* of the last sv_viterbi* launch (0 before the first) */
    /* edge-stationary schedule (zeros when the model does not fit that kernel) */
    int32_t es_threads, es_seqs_per_cta, es_ranks, es_heavy_rows;
    uint64_t es_slots_per_thread;  /* register-resident edge slots per thread (padded) */
    uint64_t es_gather_wavefronts; /* smem wavefronts of the v gather per step (scheduled) */
    uint64_t es_gather_ideal;      /* same with zero bank conflicts */
    uint64_t es_smem_bytes;
    uint64_t es_split_states;      /* virtual states added by in-degree capping */
    /* lockstep schedule (zeros when the model does not fit that kernel, or not asked for yet) */
    int32_t ls_threads, ls_seqs_per_cta, ls_ranks, ls_heavy_rows;
    int32_t ls_tmem_cols, ls_reserved;
    uint64_t ls_gather_wavefronts; /* LDS.128 wavefronts of the v gather per step and float4 group */
    uint64_t ls_gather_ideal;      /* same with zero bank conflicts */
    uint64_t ls_smem_bytes;
    uint64_t ls_split_states;
    uint64_t ls_colour_cost_before, ls_colour_cost_after; /* K0 bank-colouring search */
    uint64_t ls_sa_iters;          /* length of the search behind the current lockstep plan */
    /* chain schedule (zeros when the model is not chain + specials, or not asked for yet) */
    int32_t chain_threads, chain_states, chain_specials, chain_seqs_per_cta, chain_tmem_cols, chain_reserved;
    double plan_ms;                /* host milliseconds spent in K0 planners so far */
} sv_model_info;
int sv_model_get_info(const sv_model* m, sv_model_info* info);

/* ------------------------------------------------ observation sequences ---- */
/* A device-resident batch of sequences: symbols validated (< emits), packed to u8/u16,
 * ordered longest-first for the persistent scheduler. */
typedef struct sv_seqs sv_seqs;
int sv_seqs_create(const sv_model* m, const uint64_t* symbols, const uint64_t* offsets,
                   uint64_t nseq, sv_seqs** out);
/* Same, from symbols already packed one byte each (alphabet <= 256). */
int sv_seqs_create_u8(const sv_model* m, const uint8_t* symbols, const uint64_t* offsets,
                      uint64_t nseq, sv_seqs** out);
void sv_seqs_destroy(sv_seqs* s);
uint64_t sv_seqs_count(const sv_seqs* s);
uint64_t sv_seqs_total_symbols(const sv_seqs* s);

/* ------------------------------------------------------------ plain path ---- */
/* One sequence, host buffers: out[states].  Equivalent of run_Viterbi(hmm, seq). */
int sv_viterbi(sv_model* m, const uint64_t* seq, uint64_t len, float* out);

/* Many sequences, host buffers.  symbols = concatenated sequences, offsets[nseq+1].
 * Any of out ([nseq][states]), best ([nseq], min final score) and best_state ([nseq],
 * lowest state index attaining it, -1 if all +inf) may be NULL. */
int sv_viterbi_batch(sv_model* m, const uint64_t* symbols, const uint64_t* offsets, uint64_t nseq,
                     float* out, float* best, int32_t* best_state);
/* Same with one byte per symbol (alphabet <= 256; the .essb cache layout): the symbols are
 * copied to the device straight from the caller's buffer (pin it for full PCIe speed). */
int sv_viterbi_batch_u8(sv_model* m, const uint8_t* symbols, const uint64_t* offsets, uint64_t nseq,
                        float* out, float* best, int32_t* best_state);

/* Device-resident variant: `seqs` from sv_seqs_create, d_* are device pointers on the
 * model's device (may be NULL), `stream` is a cudaStream_t (NULL = default stream).
 * Asynchronous with respect to the host. */
int sv_viterbi_device(sv_model* m, const sv_seqs* seqs, float* d_out, float* d_best,
                      int32_t* d_best_state, void* stream);
/* No counterpart in the reference: added for callers that cannot
 * call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same inputs, outputs and results (bit for bit) as sv_viterbi_device, but the call allocates NOTHING: no
 * cudaMalloc / cudaFree, no lazy K0 plan, no library routine that allocates.  It runs a variant of the generic
 * kernel whose scratch (the score vectors of every CTA and the work-queue counter) is a fixed file-scope device
 * array of SV_STATIC_SCRATCH_BYTES that every GPU loading the library holds until exit.
 *   - limit: models with more than SV_STATIC_MAX_STATES states are refused (SV_ERR_ARG);
 *   - all calls on one device share that array and it keeps the previous call's contents; the call clears the
 *     one word that must start at zero (the queue counter) on `stream` itself;
 *   - calls that overlap on the GPU (different streams without an ordering between them) RACE on the array:
 *     order them (same stream, or an event).  Calls on different devices are independent.
 * `m` and `seqs` are the ordinary handles (created earlier, where allocation is allowed); d_* as above. */
#define SV_STATIC_MAX_STATES 4096
#define SV_STATIC_SCRATCH_BYTES 4849920 /* (64 + 148 CTAs * 2 vectors * 4096 states) floats */
int sv_viterbi_device_static(const sv_model* m, const sv_seqs* seqs, float* d_out, float* d_best,
                             int32_t* d_best_state, void* stream);
/* Kernel launches issued by the last sv_viterbi*_ / sv_spec_run* call on this model. */
int sv_model_last_launches(const sv_model* m);

/* ------------------------------------------------------------- multi-GPU ---- */
/* The HMM replicated on several devices of one box, the sequences of a batch dealt to them in contiguous
 * ranges of about equal symbol counts (independent sequences, no exchange inside the time loop,
 * hence no collective: each device delivers its share of the results into the caller's host arrays).
 * One host thread per device drives that device's replica.  devices = NULL, ndev = 0: every visible device.
 * Replaces the harness loop of benchmark/bench_Viterbi.h:51-57 at box scale. */
typedef struct sv_multi sv_multi;
int sv_multi_create(uint64_t states, uint64_t emits, uint64_t trans, const uint64_t* trans_src,
                    const uint64_t* trans_dst, const float* trans_w, const float* emissions, uint64_t nstart,
                    const uint64_t* start_cols, const float* start_w, const int* devices, int ndev,
                    sv_multi** out);
void sv_multi_destroy(sv_multi* mm);
int sv_multi_device_count(const sv_multi* mm);
sv_model* sv_multi_model(sv_multi* mm, int k);    /* replica k (owned by mm) */
int sv_multi_prepare(sv_multi* mm);               /* sv_model_prepare on every replica, in parallel */
int sv_multi_viterbi_batch(sv_multi* mm, const uint64_t* symbols, const uint64_t* offsets, uint64_t nseq,
                           float* out, float* best, int32_t* best_state);
int sv_multi_viterbi_batch_u8(sv_multi* mm, const uint8_t* symbols, const uint64_t* offsets, uint64_t nseq,
                              float* out, float* best, int32_t* best_state);
/* Host only: the partition the two calls above use; cut[parts + 1], device k gets sequences [cut[k], cut[k+1]). */
int sv_multi_split(const uint64_t* offsets, uint64_t nseq, int parts, uint64_t* cut);

/* ------------------------------------------------------ specialised path ---- */
typedef struct sv_spec sv_spec;

enum {
    SV_SPEC_AUTO = 0,
    SV_SPEC_DENSE_GEMM = 1, /* K3: register-blocked dense (min,+) GEMM, 2*N^3 ops / handler */
    SV_SPEC_SPARSE_LHS = 2, /* K3': sparse level-1 matrix times dense handler, 2*nnz*N ops */
    SV_SPEC_PAIR_ROWS = 3,  /* K3'': level 2 only, sparse x sparse rows streamed into the dense table;
                               AUTO picks it for level 2 when the model has few dense rows */
    SV_SPEC_ONDEMAND = 4    /* level >= 3: precompute only the level-(L-1) table; every chunk of a run builds the
                               level-L handlers of the tuples the batch actually observes (K3' into a scratch, one per
                               sequence) and applies them (K4) -- same sums, same association, same bits as the full
                               table.  Level 3 of a 2405-state model needs 9.3 GB instead of 185 GB.  AUTO picks it
                               when the full level-L table would not fit the device (tuple range must be the whole
                               range or empty) */
};

/* spec_with(hmm): level-1 tables, and for level > 1 the dense handler of every symbol
 * tuple in [tuple_begin, tuple_end) (pass 0, 0 for all emits^level tuples).  Tuple index =
 * sum_j k_j * emits^(level-j), k_1 = first consumed symbol.  If d_table is not NULL it is a device
 * buffer the caller owns and handler `tuple_begin` is written at d_table[0] (so it needs
 * (tuple_end-tuple_begin)*states*states floats); otherwise the table is allocated internally.
 * Deferred table: with method AUTO, level 2, all tuples and no caller buffer, a model with the chain + specials
 * structure gets its table on FIRST USE (SV_SPEC_RUN_TABLE, sv_spec_table, sv_spec_get_handler, sv_spec_save) instead
 * of here, because its default run (SV_SPEC_RUN_CHAIN) never reads it: spec_with then costs a model upload and
 * 4*states^2*emits^2 bytes of device memory stay free.  SV_SPEC_LAZY_TABLE=0 in the environment, or any explicit
 * method, builds it here as the reference's spec_with does (GraphBLAS_spec_impl.cpp:168-181). */
int sv_spec_create(sv_model* m, uint64_t level, int method, uint64_t tuple_begin,
                   uint64_t tuple_end, float* d_table, void* stream, sv_spec** out);
void sv_spec_destroy(sv_spec* s);
uint64_t sv_spec_level(const sv_spec* s);
/* Re-points a spec with a CALLER-OWNED table at `d_table`, whose first handler is tuple `table_first`,
 * and declares tuples [tuple_begin, tuple_end) of it valid (table_first <= tuple_begin).  This is the
 * step after an all-gather of tuple shards: every rank builds its shard with sv_spec_create into its
 * slice of one full-size buffer, the shards are exchanged, then each rank calls
 * sv_spec_set_range(s, full_buffer, 0, 0, emits^level).  Ranges below table_first are rejected. */
int sv_spec_set_range(sv_spec* s, float* d_table, uint64_t table_first, uint64_t tuple_begin,
                      uint64_t tuple_end);
/* Device pointer of the handler table (its first handler is tuple *table_first) and the valid range. */
float* sv_spec_table(const sv_spec* s, uint64_t* tuple_begin, uint64_t* tuple_end);
/* Copies one dense handler (states*states floats, row = destination state) to the host. */
int sv_spec_get_handler(const sv_spec* s, uint64_t tuple_index, float* out);

/* Handler-table persistence (level >= 2).  The reference rebuilds every table in each spec_with
 * call (GraphBLAS_spec_impl.cpp:45-48); these two keep a built table across processes.
 * sv_spec_save writes the valid tuple range of the table (64-byte header + dense handlers);
 * sv_spec_load attaches a saved table to `m` without recomputing it -- it fails with SV_ERR_ARG
 * if the file was built from a different model (dimensions and a fingerprint of T' and the
 * emissions are checked).  d_table: optional caller-owned device buffer for the loaded range
 * (as in sv_spec_create), NULL = allocate. */
int sv_spec_save(const sv_spec* s, const char* path);
int sv_spec_load(sv_model* m, const char* path, float* d_table, sv_spec** out);

/* How sv_spec_run* applies a level-2 chunk (GraphBLAS_spec_impl.cpp:69-79, one GrB_mxm per chunk):
 * SV_SPEC_RUN_TABLE reads the precomputed dense handler (K4, 4*N^2 bytes per chunk);
 * SV_SPEC_RUN_IMPLICIT recomputes the handler entries it needs from the two sparse level-1 factors
 * (K4i) -- same sums, same association, same bits, no table traffic; the table may then be empty
 * (sv_spec_create with tuple_begin == tuple_end == emits^level).  Level 2, models with at most 4
 * dense rows; otherwise SV_ERR_UNSUPPORTED and the mode is unchanged.
 * SV_SPEC_RUN_CHAIN (K4c): level 2 of a model with the chain + specials structure of the reference's profile HMMs
 * (detected, never assumed): the two-step paths through the dense rows collapse onto the chain minimum because
 * x -> fl(c + x) is monotone, so a chunk costs O(N) instead of O(N^2) and a whole batch is ONE persistent launch --
 * still the sums of GraphBLAS_spec_impl.cpp:15-36,69-79 in their association, hence the same bits.  u8 symbols, emission
 * table must fit shared memory; otherwise SV_ERR_UNSUPPORTED and the mode is unchanged.
 * SV_SPEC_RUN_AUTO (default) = CHAIN, else IMPLICIT, whenever the spec is eligible, else TABLE.  Level 1 ignores the mode. */
enum {
    SV_SPEC_RUN_TABLE = 0,
    SV_SPEC_RUN_IMPLICIT = 1, /* K4i */
    SV_SPEC_RUN_CHAIN = 2,    /* K4c */
    SV_SPEC_RUN_AUTO = 3
};
int sv_spec_set_run_mode(sv_spec* s, int mode);

int sv_spec_run(sv_spec* s, const uint64_t* seq, uint64_t len, float* out);
int sv_spec_run_batch(sv_spec* s, const uint64_t* symbols, const uint64_t* offsets, uint64_t nseq,
                      float* out, float* best, int32_t* best_state);
int sv_spec_run_batch_u8(sv_spec* s, const uint8_t* symbols, const uint64_t* offsets, uint64_t nseq,
                         float* out, float* best, int32_t* best_state);
int sv_spec_run_device(sv_spec* s, const sv_seqs* seqs, float* d_out, float* d_best,
                       int32_t* d_best_state, void* stream);

/* ---------------------------------------------------- files (host only) ---- */
/* .chmm text / .chmmb packed (by extension).  dims = {states, emits, trans, nstart}. */
typedef struct sv_hmm_file sv_hmm_file;
int sv_io_hmm_open(const char* path, sv_hmm_file** out, uint64_t dims[4]);
int sv_io_hmm_copy(const sv_hmm_file* f, uint64_t* trans_src, uint64_t* trans_dst, float* trans_w,
                   float* emissions, uint64_t* start_cols, float* start_w);
int sv_io_hmm_write_packed(const sv_hmm_file* f, const char* path);
void sv_io_hmm_close(sv_hmm_file* f);

/* .ess text / .essb packed (by extension). */
typedef struct sv_ess_file sv_ess_file;
int sv_io_ess_open(const char* path, sv_ess_file** out, uint64_t* nseq, uint64_t* total_symbols);
int sv_io_ess_copy(const sv_ess_file* f, uint64_t* offsets /*[nseq+1]*/, uint64_t* symbols);
int sv_io_ess_write_packed(const sv_ess_file* f, const char* path);
void sv_io_ess_close(sv_ess_file* f);

/* K0 self-check (host only, no GPU, NOT a compute path): builds the lockstep schedule of a model
 * and replays its register images for one sequence the way a CTA would, so the K0
 * transformation can be verified against the oracle without a device.  stats[12] = {valid,
 * threads, sequences per CTA, ranks, heavy rows, split states, gather wavefronts, ideal
 * wavefronts, colour cost before / after the search, smem bytes, tmem columns}.  seq/out may be
 * NULL to get the statistics only. */
int sv_k0_lockstep_check(uint64_t states, uint64_t emits, uint64_t trans, const uint64_t* trans_src,
                         const uint64_t* trans_dst, const float* trans_w, const float* emissions,
                         uint64_t nstart, const uint64_t* start_cols, const float* start_w,
                         const uint64_t* seq, uint64_t len, float* out, uint64_t stats[12]);

/* The same for the chain kernel: detects the chain + specials structure and replays the kernel's
 * algebra (band edge from position p-1, hub broadcast, special rows through the chain minimum) on the
 * host.  stats[8] = {valid, threads, chain states, specials, tmem columns, nonneg, special 0, special 1}. */
int sv_k0_chain_check(uint64_t states, uint64_t emits, uint64_t trans, const uint64_t* trans_src,
                      const uint64_t* trans_dst, const float* trans_w, const float* emissions,
                      uint64_t nstart, const uint64_t* start_cols, const float* start_w,
                      const uint64_t* seq, uint64_t len, float* out, uint64_t stats[8]);

/* The same for K4c, the level-2 run of a chain + specials model (SV_SPEC_RUN_CHAIN): first symbol, chunks of two
 * symbols through the collapsed two-step paths, odd tail as one level-1 step, replayed on the host; to be compared
 * with the level-2 semantics of GraphBLAS_spec_impl.cpp:15-36,50-97. */
int sv_k0_chain_pair_check(uint64_t states, uint64_t emits, uint64_t trans, const uint64_t* trans_src,
                           const uint64_t* trans_dst, const float* trans_w, const float* emissions,
                           uint64_t nstart, const uint64_t* start_cols, const float* start_w,
                           const uint64_t* seq, uint64_t len, float* out, uint64_t stats[8]);

#ifdef __cplusplus
}
#endif
#endif

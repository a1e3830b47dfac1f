/* CPU ORACLE for the (min,+) Viterbi hot path -- TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C restatement of the arithmetic of the reference's GraphBLAS backends
 * (reference files: Viterbi_impl/GraphBLAS_impl.cpp, GraphBLAS_spec_impl.cpp,
 * GraphBLAS_helper.cpp).  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may use it; the product path
 * (spec_viterbi_b200/csrc) never links or calls anything in oracle/.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks this file against
 *   (1) the reference's four golden vectors (tests/test_helper.h:17-22), levels 1..3,
 *   (2) the reference's own sources compiled unmodified against oracle/grb_shim
 *       (oracle/_ref/libsv_ref.so), bit for bit, on the shipped models,
 *   (3) committed fixtures generated from (2) under tests/golden/.
 *
 * All arithmetic is IEEE fp32 add and exact min; every output is a min over
 * singly rounded sums, so only the parenthesisation matters:
 *   plain / level 1:  v'[i] = min_j fl( fl(E[o][i] + T'[i][j]) + v[j] )
 *   level L handler:  H_(k1..kL) = M_kL (x) H_(k1..k(L-1)),  M_o[i][j] = fl(E[o][i] + T'[i][j])
 * where T'[dst][src] is the transposed transition matrix in -log2 space.
 */
#ifndef SV_ORACLE_H
#define SV_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Borrowed view of an HMM in the reference's layout (HMM.h:28-38). */
typedef struct svo_model {
    uint64_t states;            /* N */
    uint64_t emits;             /* E */
    uint64_t trans;             /* nnz (COO entries, duplicates allowed: FIRST wins) */
    uint64_t nstart;            /* entries of the sparse start vector */
    const uint64_t* trans_src;  /* HMM::trans_rows */
    const uint64_t* trans_dst;  /* HMM::trans_cols */
    const float* trans_w;       /* -log2 P(src->dst) */
    const float* emis;          /* [E][N], -log2 P(symbol | state) */
    const uint64_t* start_cols; /* states with a start entry */
    const float* start_w;       /* their -log2 start probability */
} svo_model;

/* GraphBLAS_impl::run_Viterbi (GraphBLAS_impl.cpp:4-93).  out[N].  T >= 1. */
int svo_plain(const svo_model* m, const uint64_t* seq, uint64_t T, float* out);

/* Many sequences (packed symbols + offsets[nseq+1]) on `threads` OpenMP threads;
 * out is [nseq][N].  Used as the timed "port" CPU baseline. */
int svo_plain_batch(const svo_model* m, const uint64_t* symbols, const uint64_t* offsets,
                    uint64_t nseq, float* out, int threads);

/* GraphBLAS_spec_impl (GraphBLAS_spec_impl.cpp:103-182 initializer, :15-36 add_level,
 * :50-97 run).  Handlers are dense N*N fp32 (+inf where the reference has no entry),
 * built lazily per symbol tuple unless eager != 0. */
typedef struct svo_spec svo_spec;
svo_spec* svo_spec_create(const svo_model* m, uint64_t level, int eager);
void svo_spec_destroy(svo_spec* s);
int svo_spec_run(svo_spec* s, const uint64_t* seq, uint64_t T, float* out);
/* Dense level-`level` handler of tuple index sum_j k_j * E^(level-j) (k_1 = first
 * consumed symbol); returns a pointer owned by the spec object (N*N floats,
 * row = destination state), or NULL. */
const float* svo_spec_handler(svo_spec* s, uint64_t tuple_index);
/* Level-1 tables: M_o values on the CSR pattern of T' and s_o = E[o] + start. */
uint64_t svo_spec_csr(const svo_spec* s, const uint64_t** rowptr, const uint32_t** col);
const float* svo_spec_level1_values(const svo_spec* s, uint64_t symbol); /* [nnz_csr] */
const float* svo_spec_start(const svo_spec* s, uint64_t symbol);         /* [N] */

/* (min score, lowest argmin) of a final vector; +inf/-1 if every entry is +inf. */
void svo_best_state(const float* v, uint64_t n, float* best, int64_t* arg);

#ifdef __cplusplus
}
#endif
#endif

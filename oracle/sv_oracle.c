/* CPU ORACLE -- test infrastructure only; see sv_oracle.h for the contract.
 * Build: gcc -O2 -fopenmp -ffp-contract=off -fPIC -shared sv_oracle.c -o libsv_oracle.so
 */
#include "sv_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ---- CSR of the transposed transition matrix, duplicates resolved FIRST-wins ----
 * Reference: GrB_Matrix_build_FP32(T', I=trans_cols, J=trans_rows, X, nnz, GrB_FIRST_FP32)
 * (GraphBLAS_impl.cpp:42-45, GraphBLAS_spec_impl.cpp:111-114). */
typedef struct csr_t {
    uint64_t n, nnz;
    uint64_t* rowptr; /* [n+1], row = destination state */
    uint32_t* col;    /* source state */
    float* w;
} csr_t;

typedef struct coo_key {
    uint64_t dst, src, pos;
} coo_key;

static int cmp_key(const void* a, const void* b) {
    const coo_key* x = (const coo_key*)a;
    const coo_key* y = (const coo_key*)b;
    if (x->dst != y->dst) return x->dst < y->dst ? -1 : 1;
    if (x->src != y->src) return x->src < y->src ? -1 : 1;
    if (x->pos != y->pos) return x->pos < y->pos ? -1 : 1;
    return 0;
}

static int csr_build(const svo_model* m, csr_t* c) {
    const uint64_t n = m->states, nz = m->trans;
    coo_key* k = (coo_key*)malloc((nz ? nz : 1) * sizeof(coo_key));
    c->n = n;
    c->rowptr = (uint64_t*)calloc(n + 1, sizeof(uint64_t));
    c->col = (uint32_t*)malloc((nz ? nz : 1) * sizeof(uint32_t));
    c->w = (float*)malloc((nz ? nz : 1) * sizeof(float));
    if (!k || !c->rowptr || !c->col || !c->w) return -1;
    uint64_t kept = 0;
    for (uint64_t e = 0; e < nz; ++e) {
        if (m->trans_src[e] >= n || m->trans_dst[e] >= n) continue; /* out of range: dropped */
        k[kept].dst = m->trans_dst[e];
        k[kept].src = m->trans_src[e];
        k[kept].pos = e;
        ++kept;
    }
    qsort(k, kept, sizeof(coo_key), cmp_key);
    uint64_t out = 0;
    for (uint64_t e = 0; e < kept; ++e) {
        if (e > 0 && k[e].dst == k[e - 1].dst && k[e].src == k[e - 1].src) continue; /* FIRST */
        c->col[out] = (uint32_t)k[e].src;
        c->w[out] = m->trans_w[k[e].pos];
        c->rowptr[k[e].dst + 1] += 1;
        ++out;
    }
    for (uint64_t i = 0; i < n; ++i) c->rowptr[i + 1] += c->rowptr[i];
    c->nnz = out;
    free(k);
    return 0;
}

static void csr_free(csr_t* c) {
    free(c->rowptr);
    free(c->col);
    free(c->w);
}

/* Dense start column: +inf where the sparse start vector has no entry, FIRST wins
 * (GraphBLAS_impl.cpp:18-21). */
static void dense_start(const svo_model* m, float* s) {
    const uint64_t n = m->states;
    unsigned char* seen = (unsigned char*)calloc(n ? n : 1, 1);
    for (uint64_t i = 0; i < n; ++i) s[i] = INFINITY;
    for (uint64_t k = 0; k < m->nstart; ++k) {
        const uint64_t c = m->start_cols[k];
        if (c < n && !seen[c]) {
            seen[c] = 1;
            s[c] = m->start_w[k];
        }
    }
    free(seen);
}

/* One step: GraphBLAS_impl.cpp:65 (prob_x_trans = Em[o] (x) T') then :69 (next = pxt (x) v). */
static void plain_step(const csr_t* c, const float* em_o, const float* v, float* next) {
    for (uint64_t i = 0; i < c->n; ++i) {
        float best = INFINITY;
        const float e = em_o[i];
        for (uint64_t p = c->rowptr[i]; p < c->rowptr[i + 1]; ++p) {
            const float pxt = e + c->w[p];         /* fl(E + T') */
            const float cand = pxt + v[c->col[p]]; /* fl(. + v)  */
            best = cand < best ? cand : best;
        }
        next[i] = best;
    }
}

static int plain_with_csr(const svo_model* m, const csr_t* c, const float* start,
                          const uint64_t* seq, uint64_t T, float* out, float* tmp) {
    const uint64_t n = m->states;
    if (T == 0) return -1;
    for (uint64_t t = 0; t < T; ++t)
        if (seq[t] >= m->emits) return -2;
    /* GraphBLAS_impl.cpp:59: v0 = Em[seq0] (x) start */
    const float* e0 = m->emis + seq[0] * n;
    float* cur = out;
    float* nxt = tmp;
    for (uint64_t i = 0; i < n; ++i) cur[i] = e0[i] + start[i];
    for (uint64_t t = 1; t < T; ++t) {
        plain_step(c, m->emis + seq[t] * n, cur, nxt);
        float* sw = cur;
        cur = nxt;
        nxt = sw;
    }
    if (cur != out) memcpy(out, cur, n * sizeof(float));
    return 0;
}

int svo_plain(const svo_model* m, const uint64_t* seq, uint64_t T, float* out) {
    csr_t c;
    if (csr_build(m, &c) != 0) return -1;
    float* start = (float*)malloc((m->states ? m->states : 1) * sizeof(float));
    float* tmp = (float*)malloc((m->states ? m->states : 1) * sizeof(float));
    dense_start(m, start);
    const int rc = plain_with_csr(m, &c, start, seq, T, out, tmp);
    free(start);
    free(tmp);
    csr_free(&c);
    return rc;
}

int svo_plain_batch(const svo_model* m, const uint64_t* symbols, const uint64_t* offsets,
                    uint64_t nseq, float* out, int threads) {
    csr_t c;
    if (csr_build(m, &c) != 0) return -1;
    const uint64_t n = m->states;
    float* start = (float*)malloc((n ? n : 1) * sizeof(float));
    dense_start(m, start);
    int rc = 0;
    if (threads < 1) threads = 1;
#pragma omp parallel num_threads(threads)
    {
        float* tmp = (float*)malloc((n ? n : 1) * sizeof(float));
#pragma omp for schedule(dynamic, 1)
        for (int64_t s = 0; s < (int64_t)nseq; ++s) {
            const int r = plain_with_csr(m, &c, start, symbols + offsets[s],
                                         offsets[s + 1] - offsets[s], out + (uint64_t)s * n, tmp);
            if (r != 0) {
#pragma omp atomic write
                rc = r;
            }
        }
        free(tmp);
    }
    free(start);
    csr_free(&c);
    return rc;
}

/* ---- specialised path ---- */
struct svo_spec {
    svo_model m; /* borrowed pointers; emis is only used during create */
    uint64_t level;
    uint64_t ntuples; /* E^level */
    csr_t c;
    float* lvl1;     /* [E][nnz]  M_o on the pattern of T' (GraphBLAS_spec_impl.cpp:156-157) */
    float* start;    /* [E][N]    s_o = Em[o] (x) start    (GraphBLAS_spec_impl.cpp:153-154) */
    float** handler; /* [E^level] dense N*N or NULL (lazy) */
    float* scratch;  /* [N] */
};

/* dst = M_sym (x) src, both dense N*N (src may be NULL: then dst = dense(M_sym)).
 * add_level: handler_new = updater[i] (x) prev  (GraphBLAS_spec_impl.cpp:28). */
static void left_multiply(const svo_spec* s, uint64_t sym, const float* src, float* dst) {
    const uint64_t n = s->c.n;
    const float* mv = s->lvl1 + sym * s->c.nnz;
    for (uint64_t i = 0; i < n; ++i) {
        float* row = dst + i * n;
        for (uint64_t k = 0; k < n; ++k) row[k] = INFINITY;
        for (uint64_t p = s->c.rowptr[i]; p < s->c.rowptr[i + 1]; ++p) {
            const float a = mv[p];
            const uint64_t j = s->c.col[p];
            if (src == NULL) {
                row[j] = a;
            } else {
                const float* srow = src + j * n;
                for (uint64_t k = 0; k < n; ++k) {
                    const float cand = a + srow[k];
                    row[k] = cand < row[k] ? cand : row[k];
                }
            }
        }
    }
}

/* Handler of the tuple (k1..kL) encoded base E, k1 most significant. */
static float* build_handler(svo_spec* s, uint64_t tuple) {
    const uint64_t n = s->c.n, E = s->m.emits, L = s->level;
    uint64_t key[64];
    for (uint64_t j = L; j-- > 0;) {
        key[j] = tuple % E;
        tuple /= E;
    }
    float* a = (float*)malloc(n * n * sizeof(float));
    float* b = L > 1 ? (float*)malloc(n * n * sizeof(float)) : NULL;
    if (!a || (L > 1 && !b)) return NULL;
    left_multiply(s, key[0], NULL, a);
    for (uint64_t j = 1; j < L; ++j) {
        left_multiply(s, key[j], a, b);
        float* sw = a;
        a = b;
        b = sw;
    }
    free(b);
    return a;
}

svo_spec* svo_spec_create(const svo_model* m, uint64_t level, int eager) {
    if (level < 1 || level > 63) return NULL;
    svo_spec* s = (svo_spec*)calloc(1, sizeof(svo_spec));
    if (!s) return NULL;
    s->m = *m;
    s->level = level;
    if (csr_build(m, &s->c) != 0) {
        free(s);
        return NULL;
    }
    const uint64_t n = m->states, E = m->emits, nz = s->c.nnz;
    s->ntuples = 1;
    for (uint64_t j = 0; j < level; ++j) s->ntuples *= E;
    s->lvl1 = (float*)malloc((E * nz ? E * nz : 1) * sizeof(float));
    s->start = (float*)malloc((E * n ? E * n : 1) * sizeof(float));
    s->scratch = (float*)malloc((n ? n : 1) * sizeof(float));
    s->handler = (float**)calloc(level > 1 ? s->ntuples : 1, sizeof(float*));
    float* st = (float*)malloc((n ? n : 1) * sizeof(float));
    dense_start(m, st);
    for (uint64_t o = 0; o < E; ++o) {
        const float* em = m->emis + o * n;
        for (uint64_t i = 0; i < n; ++i) {
            s->start[o * n + i] = em[i] + st[i];
            for (uint64_t p = s->c.rowptr[i]; p < s->c.rowptr[i + 1]; ++p)
                s->lvl1[o * nz + p] = em[i] + s->c.w[p];
        }
    }
    free(st);
    if (eager && level > 1) {
#pragma omp parallel for schedule(dynamic, 1)
        for (int64_t t = 0; t < (int64_t)s->ntuples; ++t) s->handler[t] = build_handler(s, (uint64_t)t);
    }
    return s;
}

void svo_spec_destroy(svo_spec* s) {
    if (!s) return;
    if (s->level > 1)
        for (uint64_t t = 0; t < s->ntuples; ++t) free(s->handler[t]);
    free(s->handler);
    free(s->lvl1);
    free(s->start);
    free(s->scratch);
    csr_free(&s->c);
    free(s);
}

const float* svo_spec_handler(svo_spec* s, uint64_t tuple_index) {
    if (!s || s->level < 2 || tuple_index >= s->ntuples) return NULL;
    if (!s->handler[tuple_index]) s->handler[tuple_index] = build_handler(s, tuple_index);
    return s->handler[tuple_index];
}

uint64_t svo_spec_csr(const svo_spec* s, const uint64_t** rowptr, const uint32_t** col) {
    *rowptr = s->c.rowptr;
    *col = s->c.col;
    return s->c.nnz;
}
const float* svo_spec_level1_values(const svo_spec* s, uint64_t symbol) {
    return s->lvl1 + symbol * s->c.nnz;
}
const float* svo_spec_start(const svo_spec* s, uint64_t symbol) {
    return s->start + symbol * s->c.n;
}

int svo_spec_run(svo_spec* s, const uint64_t* seq, uint64_t T, float* out) {
    const uint64_t n = s->c.n, E = s->m.emits, L = s->level, nz = s->c.nnz;
    if (T == 0) return -1;
    for (uint64_t t = 0; t < T; ++t)
        if (seq[t] >= E) return -2;
    float* cur = out;
    float* nxt = s->scratch;
    /* GraphBLAS_spec_impl.cpp:54: result = dup(emit_pr_x_start_pr[seq0]) */
    memcpy(cur, s->start + seq[0] * n, n * sizeof(float));
    uint64_t i = 1;
    if (L > 1) {
        /* GraphBLAS_spec_impl.cpp:66-81: chunks of `level` symbols starting at index 1 */
        while (T - i >= L) {
            uint64_t tuple = 0;
            for (uint64_t j = 0; j < L; ++j, ++i) tuple = tuple * E + seq[i];
            const float* H = svo_spec_handler(s, tuple);
            if (!H) return -3;
            for (uint64_t r = 0; r < n; ++r) {
                const float* hrow = H + r * n;
                float best = INFINITY;
                for (uint64_t k = 0; k < n; ++k) {
                    const float cand = hrow[k] + cur[k];
                    best = cand < best ? cand : best;
                }
                nxt[r] = best;
            }
            float* sw = cur;
            cur = nxt;
            nxt = sw;
        }
    }
    /* GraphBLAS_spec_impl.cpp:84-89: tail with level-1 matrices */
    for (; i < T; ++i) {
        const float* mv = s->lvl1 + seq[i] * nz;
        for (uint64_t r = 0; r < n; ++r) {
            float best = INFINITY;
            for (uint64_t p = s->c.rowptr[r]; p < s->c.rowptr[r + 1]; ++p) {
                const float cand = mv[p] + cur[s->c.col[p]];
                best = cand < best ? cand : best;
            }
            nxt[r] = best;
        }
        float* sw = cur;
        cur = nxt;
        nxt = sw;
    }
    if (cur != out) {
        memcpy(out, cur, n * sizeof(float));
    }
    return 0;
}

void svo_best_state(const float* v, uint64_t n, float* best, int64_t* arg) {
    float b = INFINITY;
    int64_t a = -1;
    for (uint64_t i = 0; i < n; ++i) {
        if (v[i] < b) {
            b = v[i];
            a = (int64_t)i;
        }
    }
    *best = b;
    *arg = a;
}

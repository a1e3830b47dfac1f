// Implementation of the GraphBLAS subset declared in GraphBLAS.h (test infrastructure).
// Storage: CSR by row, column indices sorted within a row.  mxm: Gustavson row-wise
// product with a dense accumulator; result pattern = structural intersection, exactly as
// the GraphBLAS semantics prescribe (an output entry exists iff some j has both operands).
#include "GraphBLAS.h"

#include <algorithm>
#include <cstring>
#include <limits>
#include <numeric>
#include <vector>

struct sv_grb_type_opaque {
    int id;
};
struct sv_grb_binop_opaque {
    int id;
};
struct sv_grb_semiring_opaque {
    int id;
};
struct sv_grb_desc_opaque {
    int id;
};
struct sv_grb_matrix_opaque {
    GrB_Index nrows = 0, ncols = 0;
    std::vector<GrB_Index> rowptr; // nrows + 1
    std::vector<GrB_Index> col;
    std::vector<float> val;
};

namespace {
sv_grb_type_opaque fp32_type{1};
sv_grb_binop_opaque first_fp32{1};
sv_grb_semiring_opaque min_plus_fp32{1};
} // namespace

extern "C" {

GrB_Type GrB_FP32 = &fp32_type;
GrB_BinaryOp GrB_FIRST_FP32 = &first_fp32;
GrB_Semiring GrB_MIN_PLUS_SEMIRING_FP32 = &min_plus_fp32;

GrB_Info GrB_init(GrB_Mode) { return GrB_SUCCESS; }
GrB_Info GrB_finalize(void) { return GrB_SUCCESS; }

GrB_Info GrB_Matrix_new(GrB_Matrix* A, GrB_Type type, GrB_Index nrows, GrB_Index ncols) {
    if (A == nullptr || type == nullptr) return GrB_NULL_POINTER;
    auto* m = new sv_grb_matrix_opaque;
    m->nrows = nrows;
    m->ncols = ncols;
    m->rowptr.assign(nrows + 1, 0);
    *A = m;
    return GrB_SUCCESS;
}

GrB_Info GrB_Matrix_dup(GrB_Matrix* C, const GrB_Matrix A) {
    if (C == nullptr || A == nullptr) return GrB_NULL_POINTER;
    *C = new sv_grb_matrix_opaque(*A);
    return GrB_SUCCESS;
}

GrB_Info GrB_Matrix_clear(GrB_Matrix A) {
    if (A == nullptr) return GrB_NULL_POINTER;
    A->rowptr.assign(A->nrows + 1, 0);
    A->col.clear();
    A->val.clear();
    return GrB_SUCCESS;
}

GrB_Info GrB_Matrix_free(GrB_Matrix* A) {
    if (A != nullptr && *A != nullptr) {
        delete *A;
        *A = nullptr;
    }
    return GrB_SUCCESS;
}

GrB_Info GrB_Matrix_wait(GrB_Matrix* A) { return A == nullptr ? GrB_NULL_POINTER : GrB_SUCCESS; }

GrB_Info GrB_Matrix_nrows(GrB_Index* n, const GrB_Matrix A) {
    if (n == nullptr || A == nullptr) return GrB_NULL_POINTER;
    *n = A->nrows;
    return GrB_SUCCESS;
}
GrB_Info GrB_Matrix_ncols(GrB_Index* n, const GrB_Matrix A) {
    if (n == nullptr || A == nullptr) return GrB_NULL_POINTER;
    *n = A->ncols;
    return GrB_SUCCESS;
}
GrB_Info GrB_Matrix_nvals(GrB_Index* n, const GrB_Matrix A) {
    if (n == nullptr || A == nullptr) return GrB_NULL_POINTER;
    *n = A->col.size();
    return GrB_SUCCESS;
}

GrB_Info GrB_Matrix_build_FP32(GrB_Matrix C, const GrB_Index* I, const GrB_Index* J,
                               const float* X, GrB_Index nvals, const GrB_BinaryOp dup) {
    if (C == nullptr || (nvals > 0 && (I == nullptr || J == nullptr || X == nullptr)))
        return GrB_NULL_POINTER;
    if (dup != GrB_FIRST_FP32) return GrB_DOMAIN_MISMATCH;
    if (!C->col.empty()) return GrB_OUTPUT_NOT_EMPTY;
    for (GrB_Index t = 0; t < nvals; ++t)
        if (I[t] >= C->nrows || J[t] >= C->ncols) return GrB_INVALID_INDEX;
    std::vector<GrB_Index> order(nvals);
    std::iota(order.begin(), order.end(), GrB_Index(0));
    std::stable_sort(order.begin(), order.end(), [&](GrB_Index a, GrB_Index b) {
        return I[a] != I[b] ? I[a] < I[b] : J[a] < J[b];
    });
    C->rowptr.assign(C->nrows + 1, 0);
    C->col.reserve(nvals);
    C->val.reserve(nvals);
    for (GrB_Index p = 0; p < nvals; ++p) {
        const GrB_Index t = order[p];
        if (p > 0 && I[t] == I[order[p - 1]] && J[t] == J[order[p - 1]]) {
            continue; // dup = FIRST: the earliest tuple (stable order) was kept
        }
        C->col.push_back(J[t]);
        C->val.push_back(X[t]);
        C->rowptr[I[t] + 1] += 1;
    }
    for (GrB_Index r = 0; r < C->nrows; ++r) C->rowptr[r + 1] += C->rowptr[r];
    return GrB_SUCCESS;
}

GrB_Info GrB_Matrix_extractTuples_FP32(GrB_Index* I, GrB_Index* J, float* X, GrB_Index* nvals,
                                       const GrB_Matrix A) {
    if (nvals == nullptr || A == nullptr) return GrB_NULL_POINTER;
    if (*nvals < A->col.size()) return GrB_INVALID_VALUE; // insufficient space
    GrB_Index p = 0;
    for (GrB_Index r = 0; r < A->nrows; ++r) {
        for (GrB_Index q = A->rowptr[r]; q < A->rowptr[r + 1]; ++q, ++p) {
            if (I != nullptr) I[p] = r;
            if (J != nullptr) J[p] = A->col[q];
            if (X != nullptr) X[p] = A->val[q];
        }
    }
    *nvals = p;
    return GrB_SUCCESS;
}

GrB_Info GrB_mxm(GrB_Matrix C, const GrB_Matrix Mask, const GrB_BinaryOp accum,
                 const GrB_Semiring semiring, const GrB_Matrix A, const GrB_Matrix B,
                 const GrB_Descriptor desc) {
    if (C == nullptr || A == nullptr || B == nullptr || semiring == nullptr)
        return GrB_NULL_POINTER;
    if (Mask != nullptr || accum != nullptr || desc != nullptr) return GrB_INVALID_VALUE;
    if (semiring != GrB_MIN_PLUS_SEMIRING_FP32) return GrB_DOMAIN_MISMATCH;
    if (A->ncols != B->nrows || C->nrows != A->nrows || C->ncols != B->ncols)
        return GrB_DIMENSION_MISMATCH;

    sv_grb_matrix_opaque out; // built aside: C may alias A or B
    out.nrows = A->nrows;
    out.ncols = B->ncols;
    out.rowptr.assign(out.nrows + 1, 0);
    std::vector<float> acc(out.ncols);
    std::vector<unsigned char> hit(out.ncols, 0);
    std::vector<GrB_Index> touched;
    touched.reserve(out.ncols);
    for (GrB_Index i = 0; i < A->nrows; ++i) {
        touched.clear();
        for (GrB_Index p = A->rowptr[i]; p < A->rowptr[i + 1]; ++p) {
            const float a = A->val[p];
            const GrB_Index j = A->col[p];
            for (GrB_Index q = B->rowptr[j]; q < B->rowptr[j + 1]; ++q) {
                const GrB_Index k = B->col[q];
                const float t = a + B->val[q]; // "plus" multiply, one fp32 rounding
                if (!hit[k]) {
                    hit[k] = 1;
                    acc[k] = t;
                    touched.push_back(k);
                } else if (t < acc[k]) { // "min" monoid, exact
                    acc[k] = t;
                }
            }
        }
        if (touched.size() > 1) std::sort(touched.begin(), touched.end());
        for (const GrB_Index k : touched) {
            out.col.push_back(k);
            out.val.push_back(acc[k]);
            hit[k] = 0;
        }
        out.rowptr[i + 1] = out.col.size();
    }
    C->rowptr.swap(out.rowptr);
    C->col.swap(out.col);
    C->val.swap(out.val);
    return GrB_SUCCESS;
}

} // extern "C"

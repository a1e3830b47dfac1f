/* Minimal stand-in for the GraphBLAS C API 1.3 subset that the reference's CPU backends
 * call (17 symbols; call sites listed in SURVEY.md section 8c).  TEST INFRASTRUCTURE:
 * it exists so that the reference's own GraphBLAS_impl.cpp / GraphBLAS_spec_impl.cpp /
 * GraphBLAS_helper.cpp compile UNMODIFIED from /root/reference into oracle/_ref/ and can
 * serve as the parity oracle and the timed CPU reference arm.  SuiteSparse:GraphBLAS
 * (un-vendored, un-pinned dependency of the reference, README.md:23) is not installed.
 *
 * Bit-faithfulness: the only arithmetic is GrB_mxm over GrB_MIN_PLUS_SEMIRING_FP32.
 * Every output entry is a min over singly rounded fp32 sums a(i,j)+b(j,k) taken over the
 * structural intersection; min is exact, so any conforming library produces the same bits.
 */
#ifndef SV_GRB_SHIM_GRAPHBLAS_H
#define SV_GRB_SHIM_GRAPHBLAS_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef uint64_t GrB_Index;

typedef enum {
    GrB_SUCCESS = 0,
    GrB_NO_VALUE = 1,
    GrB_UNINITIALIZED_OBJECT = 2,
    GrB_NULL_POINTER = 4,
    GrB_INVALID_VALUE = 5,
    GrB_INVALID_INDEX = 6,
    GrB_DOMAIN_MISMATCH = 7,
    GrB_DIMENSION_MISMATCH = 8,
    GrB_OUTPUT_NOT_EMPTY = 9,
    GrB_OUT_OF_MEMORY = 10,
    GrB_PANIC = 13
} GrB_Info;

typedef enum { GrB_NONBLOCKING = 0, GrB_BLOCKING = 1 } GrB_Mode;

typedef struct sv_grb_type_opaque* GrB_Type;
typedef struct sv_grb_binop_opaque* GrB_BinaryOp;
typedef struct sv_grb_semiring_opaque* GrB_Semiring;
typedef struct sv_grb_desc_opaque* GrB_Descriptor;
typedef struct sv_grb_matrix_opaque* GrB_Matrix;

extern GrB_Type GrB_FP32;
extern GrB_BinaryOp GrB_FIRST_FP32;
extern GrB_Semiring GrB_MIN_PLUS_SEMIRING_FP32;

#define GrB_NULL NULL

GrB_Info GrB_init(GrB_Mode mode);
GrB_Info GrB_finalize(void);

GrB_Info GrB_Matrix_new(GrB_Matrix* A, GrB_Type type, GrB_Index nrows, GrB_Index ncols);
GrB_Info GrB_Matrix_dup(GrB_Matrix* C, const GrB_Matrix A);
GrB_Info GrB_Matrix_clear(GrB_Matrix A);
GrB_Info GrB_Matrix_free(GrB_Matrix* A);
GrB_Info GrB_Matrix_wait(GrB_Matrix* A);

GrB_Info GrB_Matrix_nrows(GrB_Index* nrows, const GrB_Matrix A);
GrB_Info GrB_Matrix_ncols(GrB_Index* ncols, const GrB_Matrix A);
GrB_Info GrB_Matrix_nvals(GrB_Index* nvals, const GrB_Matrix A);

GrB_Info GrB_Matrix_build_FP32(GrB_Matrix C, const GrB_Index* I, const GrB_Index* J,
                               const float* X, GrB_Index nvals, const GrB_BinaryOp dup);
GrB_Info GrB_Matrix_extractTuples_FP32(GrB_Index* I, GrB_Index* J, float* X, GrB_Index* nvals,
                                       const GrB_Matrix A);

/* C = A (semiring) B; Mask, accum and desc must be GrB_NULL (all the reference uses). */
GrB_Info GrB_mxm(GrB_Matrix C, const GrB_Matrix Mask, const GrB_BinaryOp accum,
                 const GrB_Semiring semiring, const GrB_Matrix A, const GrB_Matrix B,
                 const GrB_Descriptor desc);

#ifdef __cplusplus
}
#endif
#endif

/* linalg.h -- the reference's dense helper API (reference src/linalg.h:53-206): the Cholesky bookkeeping struct and
 * the GEMM / Cholesky / triangular-solve entry points, same signatures. This build has exactly one arithmetic backend
 * (the sm_100a kernels): the compile-time backend switch of the reference collapses to it, and every helper call runs
 * one device kernel on host-resident blocks (rslqr_cuda_la_*, include/rslqr_cuda.h). The solve path does not use
 * these helpers; `CholeskyInfo.success` (0 = ok) is also where the solve publishes its per-factor status words. */
#ifndef RSLQR_LINALG_H
#define RSLQR_LINALG_H
#include "matrix.h"
#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
  char uplo;    /* 'L' once factored */
  int success;  /* 0 = factorization succeeded, non-zero otherwise */
  char lib;     /* 'C' = CUDA kernels (the reference uses 'B','E','I') */
  void* fact;   /* unused (no external factor object) */
  int is_freed;
} CholeskyInfo;

/* backend flags of the reference (src/linalg.h:17-39): none of its CPU backends exists here */
static const int kUseMKL = 0;
static const int kUseEigen = 0;
static const int kUseClap = 0;
static const int kUseBLAS = 0;

CholeskyInfo DefaultCholeskyInfo(void);
void FreeFactorization(CholeskyInfo* cholinfo);

int MatrixAddition(Matrix* A, Matrix* B, double alpha);                     /* B += alpha A */
int MatrixCholeskyFactorize(Matrix* mat);                                   /* in-place lower Cholesky, 0 = ok */
int MatrixCholeskyFactorizeWithInfo(Matrix* mat, CholeskyInfo* cholinfo);   /* + cholinfo->success = return value */
int MatrixCholeskySolve(Matrix* A, Matrix* b);                              /* b <- (L L^T)^-1 b, A holds L */
int MatrixCholeskySolveWithInfo(Matrix* A, Matrix* b, CholeskyInfo* cholinfo);
/* C = beta C + alpha op(A) op(B) */
void MatrixMultiply(Matrix* A, Matrix* B, Matrix* C, bool tA, bool tB, double alpha, double beta);
void MatrixSymmetricMultiply(Matrix* Asym, Matrix* B, Matrix* C, double alpha, double beta);
void MatrixCopyDiagonal(Matrix* dest, Matrix* src);                         /* dest = diag(src) */

enum MatrixLinearAlgebraLibrary { libBLAS = 0, libMKL = 1, libEigen = 2, libInternal = 3, libCUDA = 4 };
enum MatrixLinearAlgebraLibrary MatrixGetLinearAlgebraLibrary(void);
void MatrixPrintLinearAlgebraLibrary(void);

#ifdef __cplusplus
}
#endif
#endif

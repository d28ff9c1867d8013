/* linalg_custom.h -- the reference's "clap" dense primitives (reference src/linalg_custom.h:44-159) with identical
 * signatures and return codes. In this build each call runs one device kernel (rslqr_cuda_la_*, include/rslqr_cuda.h)
 * on host-resident column-major blocks: they are helpers for callers and for the reference's own unit tests, not the
 * solve path. -1 is returned (and a message printed) when no CUDA device is usable. */
#ifndef RSLQR_LINALG_CUSTOM_H
#define RSLQR_LINALG_CUSTOM_H
#include "matrix.h"
#ifdef __cplusplus
extern "C" {
#endif

enum { clap_kCholeskySuccess = 0, clap_kCholeskyFail = -1 };

int clap_MatrixAddition(Matrix* A, Matrix* B, double alpha);  /* B += alpha A */
int clap_MatrixScale(Matrix* A, double alpha);                /* A *= alpha */
/* C = beta C + alpha op(A) op(B) */
int clap_MatrixMultiply(Matrix* A, Matrix* B, Matrix* C, bool tA, bool tB, double alpha, double beta);
int clap_MatrixTransposeMultiply(Matrix* A, Matrix* B, Matrix* C);  /* C = A^T B */
/* C = beta C + alpha Asym B, only the lower triangle of Asym is read */
int clap_SymmetricMatrixMultiply(Matrix* Asym, Matrix* B, Matrix* C, double alpha, double beta);
int clap_AddDiagonal(Matrix* A, double alpha);                /* A += alpha I */
int clap_CholeskyFactorize(Matrix* A);                        /* in-place lower Cholesky; clap_kCholeskyFail on a pivot <= 0 */
int clap_CholeskySolve(Matrix* L, Matrix* b);                 /* b <- (L L^T)^-1 b */
int clap_LowerTriBackSub(Matrix* L, Matrix* b, bool istransposed); /* b <- L^-1 b or L^-T b */

#ifdef __cplusplus
}
#endif
#endif

/* linalg.h -- bookkeeping type for a Cholesky factorization (reference src/linalg.h:53-59).
 * This build has exactly one arithmetic backend (the sm_100a kernels), so the reference's
 * compile-time backend switch and host GEMM/Cholesky entry points are not reproduced; the struct
 * stays because NdLqrCholeskyFactors exposes an array of it and `success` (0 = ok) is where the
 * device's per-factor status words are published after a solve. */
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

CholeskyInfo DefaultCholeskyInfo(void);
void FreeFactorization(CholeskyInfo* cholinfo);

enum MatrixLinearAlgebraLibrary { libBLAS = 0, libMKL = 1, libEigen = 2, libInternal = 3, libCUDA = 4 };
enum MatrixLinearAlgebraLibrary MatrixGetLinearAlgebraLibrary(void);
void MatrixPrintLinearAlgebraLibrary(void);

#ifdef __cplusplus
}
#endif
#endif

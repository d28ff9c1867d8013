/* host_linalg.c -- the reference's dense helper API (reference src/linalg.h:103-206, src/linalg_custom.h:44-159,
 * src/linalg_utils.h:14-17) bound to the device primitives of include/rslqr_cuda.h (rslqr_cuda_la_*).
 *
 * These functions are not on the solve path (ndlqr_Solve never calls them here; its contractions are fused into the
 * solve kernels). They exist so that code written against the reference's public headers -- its own unit tests
 * test/linalg_test.c, test/linalg_custom_test.c, callers that assemble their own blocks -- links and runs against
 * this library. Every call moves its operands to the GPU, runs one kernel and moves the result back; without a CUDA
 * device it fails loudly (no CPU arithmetic in this library). */
#include <stdio.h>
#include <time.h>

#include "linalg.h"
#include "linalg_custom.h"
#include "linalg_utils.h"
#include "rslqr_cuda.h"

static int la_report(int rc, const char* fn) {
  if (rc != RSLQR_OK) fprintf(stderr, "ERROR: %s failed on the GPU: %s\n", fn, rslqr_cuda_la_last_error());
  return rc == RSLQR_OK ? 0 : -1;
}

/* ---- linalg_custom.h ---- */
int clap_MatrixAddition(Matrix* A, Matrix* B, double alpha) {
  if (!A || !B) return -1;
  return la_report(rslqr_cuda_la_axpy(A->data, B->data, (long long)MatrixNumElements(A), alpha), "clap_MatrixAddition");
}

int clap_MatrixScale(Matrix* A, double alpha) {
  if (!A) return -1;
  return la_report(rslqr_cuda_la_axpy(NULL, A->data, (long long)MatrixNumElements(A), alpha), "clap_MatrixScale");
}

int clap_MatrixMultiply(Matrix* A, Matrix* B, Matrix* C, bool tA, bool tB, double alpha, double beta) {
  if (!A || !B || !C) return -1;
  return la_report(rslqr_cuda_la_gemm(A->data, A->rows, A->cols, B->data, B->rows, B->cols, C->data, C->rows, C->cols,
                                      tA ? 1 : 0, tB ? 1 : 0, alpha, beta, 0), "clap_MatrixMultiply");
}

int clap_MatrixTransposeMultiply(Matrix* A, Matrix* B, Matrix* C) { return clap_MatrixMultiply(A, B, C, true, false, 1.0, 0.0); }

int clap_SymmetricMatrixMultiply(Matrix* Asym, Matrix* B, Matrix* C, double alpha, double beta) {
  if (!Asym || !B || !C) return -1;
  return la_report(rslqr_cuda_la_gemm(Asym->data, Asym->rows, Asym->cols, B->data, B->rows, B->cols, C->data, C->rows,
                                      C->cols, 0, 0, alpha, beta, 1), "clap_SymmetricMatrixMultiply");
}

int clap_AddDiagonal(Matrix* A, double alpha) {
  if (!A) return -1;
  return la_report(rslqr_cuda_la_add_diagonal(A->data, A->rows, alpha), "clap_AddDiagonal");
}

int clap_CholeskyFactorize(Matrix* A) {
  if (!A) return -1;
  int info = 0;
  if (la_report(rslqr_cuda_la_cholesky(A->data, A->rows, &info), "clap_CholeskyFactorize") != 0) return clap_kCholeskyFail;
  return info == 0 ? clap_kCholeskySuccess : clap_kCholeskyFail;
}

int clap_LowerTriBackSub(Matrix* L, Matrix* b, bool istransposed) {
  if (!L || !b) return -1;
  return la_report(rslqr_cuda_la_trisolve(L->data, b->rows, b->data, b->cols, istransposed ? 1 : 0), "clap_LowerTriBackSub");
}

int clap_CholeskySolve(Matrix* L, Matrix* b) {
  if (!L || !b) return -1;
  return la_report(rslqr_cuda_la_trisolve(L->data, b->rows, b->data, b->cols, 2), "clap_CholeskySolve");
}

/* ---- linalg.h (the backend switch of the reference collapses to the one backend there is) ---- */
int MatrixAddition(Matrix* A, Matrix* B, double alpha) {
  if (!A || !B) return -1;
  return clap_MatrixAddition(A, B, alpha);
}

int MatrixCholeskyFactorize(Matrix* mat) { return clap_CholeskyFactorize(mat); }

int MatrixCholeskyFactorizeWithInfo(Matrix* mat, CholeskyInfo* cholinfo) {
  const int out = clap_CholeskyFactorize(mat);
  if (cholinfo) {
    cholinfo->lib = 'C';
    cholinfo->is_freed = 1;
    cholinfo->success = out; /* 0 = success, as the reference ends up storing (src/linalg.c:84) */
    cholinfo->uplo = 'L';
  }
  return out;
}

int MatrixCholeskySolve(Matrix* A, Matrix* b) { return clap_CholeskySolve(A, b); }

int MatrixCholeskySolveWithInfo(Matrix* A, Matrix* b, CholeskyInfo* cholinfo) {
  (void)cholinfo;
  return clap_CholeskySolve(A, b);
}

void MatrixMultiply(Matrix* A, Matrix* B, Matrix* C, bool tA, bool tB, double alpha, double beta) {
  clap_MatrixMultiply(A, B, C, tA, tB, alpha, beta);
}

void MatrixSymmetricMultiply(Matrix* Asym, Matrix* B, Matrix* C, double alpha, double beta) {
  clap_SymmetricMatrixMultiply(Asym, B, C, alpha, beta);
}

/* dest = diag(src): a copy, no arithmetic */
void MatrixCopyDiagonal(Matrix* dest, Matrix* src) {
  if (!dest || !src) return;
  MatrixSetConst(dest, 0.0);
  const int len = MatrixNumElements(src);
  for (int i = 0; i < len; ++i) MatrixSetElement(dest, i, i, src->data[i]);
}

/* ---- linalg_utils.h: CPU-time stopwatch around helper calls (reference src/linalg_utils.c) ---- */
static clock_t la_clock_start;
static double la_clock_ms;
void MatrixLinAlgTimeStart(void) { la_clock_start = clock(); }
void MatrixLinAlgTimeStop(void) { la_clock_ms += (double)(clock() - la_clock_start) * 1000.0 / (double)CLOCKS_PER_SEC; }
void MatrixLinAlgTimeReset(void) { la_clock_ms = 0.0; }
double MatrixGetLinAlgTimeMilliseconds(void) { return la_clock_ms; }

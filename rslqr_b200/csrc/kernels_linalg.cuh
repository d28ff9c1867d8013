// kernels_linalg.cuh -- the reference's dense fp64 primitives (reference src/linalg_custom.c:10-139, dispatched by
// src/linalg.c:38-213) as device kernels behind the C-ABI (rslqr_cuda_la_*). They serve the PUBLIC helper API
// (MatrixMultiply, MatrixCholeskyFactorize, clap_* ... used by the reference's own unit tests and by callers that
// build their own blocks); the solve path does not go through them -- its contractions are fused into the kernels of
// kernels_scan.cuh / kernels_fused.cuh / kernels_staged*.cuh. Shapes are run-time values, matrices column-major with
// leading dimension = rows (the reference's Matrix, src/matrix.h:71-75). One launch per call; correctness over speed.
#pragma once
#include <cuda_runtime.h>

namespace rslqr {

__device__ __forceinline__ double la_at(const double* M, int rows, int i, int j, int transposed) {
  return transposed ? M[j + (size_t)rows * i] : M[i + (size_t)rows * j];
}

// C(n x p) = beta C + alpha op(A) op(B), inner dimension m; the k-innermost accumulation order of
// clap_MatrixMultiply (src/linalg_custom.c:20-46). sym != 0: A is symmetric, only its lower triangle is read
// (clap_SymmetricMatrixMultiply, :48-83). One thread per element of C.
__global__ void la_gemm_kernel(const double* A, int arows, const double* B, int brows, double* C, int crows, int n, int m,
                               int p, int tA, int tB, double alpha, double beta, int sym) {
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < (long long)n * p;
       e += (long long)gridDim.x * blockDim.x) {
    const int i = (int)(e % n), j = (int)(e / n);
    double c = C[i + (size_t)crows * j] * beta;
    for (int k = 0; k < m; ++k) {
      const double a = sym ? (i < k ? A[k + (size_t)arows * i] : A[i + (size_t)arows * k]) : la_at(A, arows, i, k, tA);
      c += alpha * a * la_at(B, brows, k, j, tB);
    }
    C[i + (size_t)crows * j] = c;
  }
}

// B += alpha A (clap_MatrixAddition, :10-15); A == nullptr: B *= alpha (clap_MatrixScale, :17-22)
__global__ void la_axpy_kernel(const double* A, double* B, long long count, double alpha) {
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < count; e += (long long)gridDim.x * blockDim.x)
    B[e] = A ? B[e] + alpha * A[e] : B[e] * alpha;
}

// A(i,i) += alpha (clap_AddDiagonal, :85-92)
__global__ void la_add_diag_kernel(double* A, int n, double alpha) {
  for (int i = threadIdx.x; i < n; i += blockDim.x) A[i + (size_t)n * i] += alpha;
}

// In-place lower Cholesky, left-looking, one CTA (clap_CholeskyFactorize, :94-118): column j minus the products of the
// columns before it, pivot check (<= 0 fails and leaves the rest untouched), scale by 1/sqrt. The strict upper
// triangle is not touched. info[0] = 0 ok, -1 failed.
__global__ void la_cholesky_kernel(double* A, int n, int* info) {
  __shared__ int failed;
  if (threadIdx.x == 0) failed = 0;
  __syncthreads();
  for (int j = 0; j < n; ++j) {
    for (int i = j + threadIdx.x; i < n; i += blockDim.x) {
      double v = A[i + (size_t)n * j];
      for (int k = 0; k < j; ++k) v -= A[i + (size_t)n * k] * A[j + (size_t)n * k];
      A[i + (size_t)n * j] = v;
    }
    __syncthreads();
    const double ajj = A[j + (size_t)n * j];
    if (!(ajj > 0.0)) {
      if (threadIdx.x == 0) failed = 1;
      __syncthreads();
      break;
    }
    __syncthreads();
    const double r = sqrt(ajj);
    for (int i = j + threadIdx.x; i < n; i += blockDim.x) A[i + (size_t)n * j] /= r;
    __syncthreads();
  }
  if (threadIdx.x == 0) info[0] = failed ? -1 : 0;
}

// b(n x m) <- L^-1 b (transposed == 0) or L^-T b (transposed != 0), L lower triangular n x n
// (clap_LowerTriBackSub, :120-139). One thread per right-hand-side column.
__global__ void la_trisolve_kernel(const double* L, int n, double* b, int m, int transposed) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= m) return;
  double* x = b + (size_t)n * k;
  if (!transposed) {
    for (int j = 0; j < n; ++j) {
      const double xj = x[j] / L[j + (size_t)n * j];
      x[j] = xj;
      for (int i = j + 1; i < n; ++i) x[i] -= L[i + (size_t)n * j] * xj;
    }
  } else {
    for (int j = n - 1; j >= 0; --j) {
      const double xj = x[j] / L[j + (size_t)n * j];
      x[j] = xj;
      for (int i = 0; i < j; ++i) x[i] -= L[j + (size_t)n * i] * xj;
    }
  }
}

}  // namespace rslqr

// rslqr_linalg.cu -- C-ABI of the dense helper primitives (include/rslqr_cuda.h, rslqr_cuda_la_*): host blocks in,
// one kernel of kernels_linalg.cuh, host blocks out. Backs the reference-named helper API of include/linalg.h and
// include/linalg_custom.h (host_linalg.c). No arithmetic happens on the host.
#include <cuda_runtime.h>

#include <mutex>
#include <string>

#include "kernels_linalg.cuh"
#include "rslqr_cuda.h"

using namespace rslqr;

namespace {
// one grow-only device scratch per device, shared by all calls (they are synchronous)
struct Scratch {
  double* buf = nullptr;
  size_t cap = 0;
  int* info = nullptr;
};
std::mutex g_mu;
Scratch g_scratch[64];
thread_local std::string g_la_err;

int la_fail(int code, const char* what, cudaError_t e = cudaSuccess) {
  g_la_err = std::string(what) + (e != cudaSuccess ? std::string(": ") + cudaGetErrorString(e) : std::string());
  return code;
}

// scratch of at least `doubles` doubles on the current device (nullptr on failure)
Scratch* scratch(size_t doubles, int* rc) {
  int dev = 0, ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) {
    cudaGetLastError();
    *rc = la_fail(RSLQR_E_NODEVICE, "no CUDA device available (the dense helpers have no CPU fallback)");
    return nullptr;
  }
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { *rc = la_fail(RSLQR_E_CUDA, "cudaGetDevice failed"); return nullptr; }
  Scratch& s = g_scratch[dev];
  if (!s.info && cudaMalloc((void**)&s.info, sizeof(int)) != cudaSuccess) { cudaGetLastError(); *rc = la_fail(RSLQR_E_NOMEM, "scratch allocation failed"); return nullptr; }
  if (doubles > s.cap) {
    cudaFree(s.buf);
    s.buf = nullptr; s.cap = 0;
    const size_t want = doubles < 4096 ? 4096 : doubles;
    if (cudaMalloc((void**)&s.buf, sizeof(double) * want) != cudaSuccess) { cudaGetLastError(); *rc = la_fail(RSLQR_E_NOMEM, "scratch allocation failed"); return nullptr; }
    s.cap = want;
  }
  return &s;
}
#define LA_CU(call)                                                        \
  do {                                                                     \
    cudaError_t e_ = (call);                                               \
    if (e_ != cudaSuccess) return la_fail(RSLQR_E_CUDA, #call, e_);        \
  } while (0)
}  // namespace

extern "C" const char* rslqr_cuda_la_last_error(void) { return g_la_err.c_str(); }

extern "C" int rslqr_cuda_la_gemm(const double* A, int arows, int acols, const double* B, int brows, int bcols, double* Cm,
                                  int crows, int ccols, int tA, int tB, double alpha, double beta, int symmetric) {
  if (!A || !B || !Cm || arows <= 0 || acols <= 0 || brows <= 0 || bcols <= 0 || crows <= 0 || ccols <= 0)
    return la_fail(RSLQR_E_ARG, "la_gemm: bad argument");
  const int n = tA ? acols : arows, m = tA ? arows : acols, p = tB ? brows : bcols;
  if ((tB ? bcols : brows) != m || crows < n || ccols < p) return la_fail(RSLQR_E_ARG, "la_gemm: dimension mismatch");
  std::lock_guard<std::mutex> lock(g_mu);
  const size_t na = (size_t)arows * acols, nb = (size_t)brows * bcols, nc = (size_t)crows * ccols;
  int rc = 0;
  Scratch* s = scratch(na + nb + nc, &rc);
  if (!s) return rc;
  double *dA = s->buf, *dB = dA + na, *dC = dB + nb;
  LA_CU(cudaMemcpy(dA, A, sizeof(double) * na, cudaMemcpyHostToDevice));
  LA_CU(cudaMemcpy(dB, B, sizeof(double) * nb, cudaMemcpyHostToDevice));
  LA_CU(cudaMemcpy(dC, Cm, sizeof(double) * nc, cudaMemcpyHostToDevice));
  const long long tot = (long long)n * p;
  la_gemm_kernel<<<(unsigned)((tot + 127) / 128 < 4096 ? (tot + 127) / 128 : 4096), 128>>>(dA, arows, dB, brows, dC, crows, n, m, p, tA, tB, alpha, beta, symmetric);
  LA_CU(cudaGetLastError());
  LA_CU(cudaMemcpy(Cm, dC, sizeof(double) * nc, cudaMemcpyDeviceToHost));
  return RSLQR_OK;
}

extern "C" int rslqr_cuda_la_axpy(const double* A, double* B, long long count, double alpha) {
  if (!B || count <= 0) return la_fail(RSLQR_E_ARG, "la_axpy: bad argument");
  std::lock_guard<std::mutex> lock(g_mu);
  int rc = 0;
  Scratch* s = scratch((size_t)count * 2, &rc);
  if (!s) return rc;
  double *dA = s->buf, *dB = dA + count;
  if (A) LA_CU(cudaMemcpy(dA, A, sizeof(double) * count, cudaMemcpyHostToDevice));
  LA_CU(cudaMemcpy(dB, B, sizeof(double) * count, cudaMemcpyHostToDevice));
  la_axpy_kernel<<<(unsigned)((count + 255) / 256 < 4096 ? (count + 255) / 256 : 4096), 256>>>(A ? dA : nullptr, dB, count, alpha);
  LA_CU(cudaGetLastError());
  LA_CU(cudaMemcpy(B, dB, sizeof(double) * count, cudaMemcpyDeviceToHost));
  return RSLQR_OK;
}

extern "C" int rslqr_cuda_la_add_diagonal(double* A, int n, double alpha) {
  if (!A || n <= 0) return la_fail(RSLQR_E_ARG, "la_add_diagonal: bad argument");
  std::lock_guard<std::mutex> lock(g_mu);
  int rc = 0;
  Scratch* s = scratch((size_t)n * n, &rc);
  if (!s) return rc;
  LA_CU(cudaMemcpy(s->buf, A, sizeof(double) * n * n, cudaMemcpyHostToDevice));
  la_add_diag_kernel<<<1, 128>>>(s->buf, n, alpha);
  LA_CU(cudaGetLastError());
  LA_CU(cudaMemcpy(A, s->buf, sizeof(double) * n * n, cudaMemcpyDeviceToHost));
  return RSLQR_OK;
}

extern "C" int rslqr_cuda_la_cholesky(double* A, int n, int* info) {
  if (!A || n <= 0 || !info) return la_fail(RSLQR_E_ARG, "la_cholesky: bad argument");
  std::lock_guard<std::mutex> lock(g_mu);
  int rc = 0;
  Scratch* s = scratch((size_t)n * n, &rc);
  if (!s) return rc;
  LA_CU(cudaMemcpy(s->buf, A, sizeof(double) * n * n, cudaMemcpyHostToDevice));
  la_cholesky_kernel<<<1, n >= 128 ? 256 : (n > 32 ? 128 : 32)>>>(s->buf, n, s->info);
  LA_CU(cudaGetLastError());
  LA_CU(cudaMemcpy(A, s->buf, sizeof(double) * n * n, cudaMemcpyDeviceToHost));
  LA_CU(cudaMemcpy(info, s->info, sizeof(int), cudaMemcpyDeviceToHost));
  return RSLQR_OK;
}

// which: 0 = L^-1 b, 1 = L^-T b, 2 = both in that order ((L L^T)^-1 b, clap_CholeskySolve)
extern "C" int rslqr_cuda_la_trisolve(const double* L, int n, double* b, int m, int which) {
  if (!L || !b || n <= 0 || m <= 0 || which < 0 || which > 2) return la_fail(RSLQR_E_ARG, "la_trisolve: bad argument");
  std::lock_guard<std::mutex> lock(g_mu);
  int rc = 0;
  Scratch* s = scratch((size_t)n * n + (size_t)n * m, &rc);
  if (!s) return rc;
  double *dL = s->buf, *db = dL + (size_t)n * n;
  LA_CU(cudaMemcpy(dL, L, sizeof(double) * n * n, cudaMemcpyHostToDevice));
  LA_CU(cudaMemcpy(db, b, sizeof(double) * n * m, cudaMemcpyHostToDevice));
  if (which != 1) la_trisolve_kernel<<<(m + 63) / 64, 64>>>(dL, n, db, m, 0);
  if (which != 0) la_trisolve_kernel<<<(m + 63) / 64, 64>>>(dL, n, db, m, 1);
  LA_CU(cudaGetLastError());
  LA_CU(cudaMemcpy(b, db, sizeof(double) * n * m, cudaMemcpyDeviceToHost));
  return RSLQR_OK;
}

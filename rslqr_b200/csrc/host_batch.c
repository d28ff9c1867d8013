/* host_batch.c -- batched front end over the C-ABI (include/rslqr_batch.h). Gathers LQRProblem
 * knot blocks into pinned staging memory, uploads, runs the device solve, copies solutions back. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "rslqr_batch.h"
#include "rslqr_cuda.h"

struct NdLqrBatchSolver {
  int n, m, N, batch, count;
  rslqr_cuda_solver* dev;
  double* stage_knots; /* pinned */
  double* stage_x0;    /* pinned */
  double* stage_soln;  /* pinned */
  double solve_ms;
};

static double now_ms(void) {
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return 1e3 * (double)ts.tv_sec + 1e-6 * (double)ts.tv_nsec;
}

NdLqrBatchSolver* ndlqr_NewBatchSolver(int n, int m, int N, int batch, int device) {
  NdLqrBatchSolver* bs = (NdLqrBatchSolver*)calloc(1, sizeof(NdLqrBatchSolver));
  if (!bs) return NULL;
  bs->n = n; bs->m = m; bs->N = N; bs->batch = batch;
  if (rslqr_cuda_create(n, m, N, batch, device, 0, &bs->dev) != RSLQR_OK) {
    /* shapes that do not fit the fused kernel need the dense factor array */
    if (rslqr_cuda_create(n, m, N, batch, device, RSLQR_KEEP_FACT, &bs->dev) != RSLQR_OK) {
      fprintf(stderr, "ERROR: ndlqr_NewBatchSolver: %s\n", rslqr_cuda_last_error());
      free(bs);
      return NULL;
    }
  } else if (rslqr_cuda_get_mode(bs->dev) == RSLQR_MODE_STAGED) {
    rslqr_cuda_destroy(bs->dev);
    if (rslqr_cuda_create(n, m, N, batch, device, RSLQR_KEEP_FACT, &bs->dev) != RSLQR_OK) {
      fprintf(stderr, "ERROR: ndlqr_NewBatchSolver: %s\n", rslqr_cuda_last_error());
      free(bs);
      return NULL;
    }
  }
  const size_t stride = (size_t)ndlqr_LQRDataBlockSize(n, m);
  bs->stage_knots = (double*)rslqr_cuda_host_alloc(sizeof(double) * (size_t)batch * N * stride);
  bs->stage_x0 = (double*)rslqr_cuda_host_alloc(sizeof(double) * (size_t)batch * n);
  bs->stage_soln = (double*)rslqr_cuda_host_alloc(sizeof(double) * (size_t)batch * N * (2 * n + m));
  if (!bs->stage_knots || !bs->stage_x0 || !bs->stage_soln) {
    fprintf(stderr, "ERROR: ndlqr_NewBatchSolver: pinned allocation failed\n");
    ndlqr_FreeBatchSolver(bs);
    return NULL;
  }
  return bs;
}

int ndlqr_FreeBatchSolver(NdLqrBatchSolver* bs) {
  if (!bs) return -1;
  if (bs->dev) rslqr_cuda_destroy(bs->dev);
  rslqr_cuda_host_free(bs->stage_knots);
  rslqr_cuda_host_free(bs->stage_x0);
  rslqr_cuda_host_free(bs->stage_soln);
  free(bs);
  return 0;
}

int ndlqr_BatchInitializeFlat(NdLqrBatchSolver* bs, const double* knots, const double* x0, int count) {
  if (!bs || !knots || !x0 || count < 0 || count > bs->batch) return -1;
  bs->count = count;
  if (rslqr_cuda_upload(bs->dev, knots, x0, 0, count) != RSLQR_OK) {
    fprintf(stderr, "ERROR: ndlqr_BatchInitializeFlat: %s\n", rslqr_cuda_last_error());
    return -1;
  }
  return 0;
}

int ndlqr_BatchInitialize(NdLqrBatchSolver* bs, const LQRProblem* const* probs, int count) {
  if (!bs || !probs || count < 0 || count > bs->batch) return -1;
  const size_t stride = (size_t)ndlqr_LQRDataBlockSize(bs->n, bs->m);
  for (int p = 0; p < count; ++p) {
    const LQRProblem* pr = probs[p];
    if (!pr || pr->nhorizon != bs->N) return -1;
    for (int k = 0; k < bs->N; ++k) {
      const LQRData* kd = pr->lqrdata[k];
      if (kd->nstates != bs->n || kd->ninputs != bs->m) return -1;
      memcpy(bs->stage_knots + ((size_t)p * bs->N + k) * stride, kd->Q, sizeof(double) * stride);
    }
    memcpy(bs->stage_x0 + (size_t)p * bs->n, pr->x0, sizeof(double) * (size_t)bs->n);
  }
  return ndlqr_BatchInitializeFlat(bs, bs->stage_knots, bs->stage_x0, count);
}

int ndlqr_BatchSolve(NdLqrBatchSolver* bs) {
  if (!bs) return -1;
  const double t0 = now_ms();
  if (rslqr_cuda_get_mode(bs->dev) == RSLQR_MODE_STAGED) rslqr_cuda_reset(bs->dev);
  if (rslqr_cuda_solve(bs->dev, 1) != RSLQR_OK) {
    fprintf(stderr, "ERROR: ndlqr_BatchSolve: %s\n", rslqr_cuda_last_error());
    return -1;
  }
  bs->solve_ms = now_ms() - t0;
  return 0;
}

int ndlqr_BatchGetNumVars(NdLqrBatchSolver* bs) { return bs ? (2 * bs->n + bs->m) * bs->N - bs->m : -1; }

int ndlqr_BatchCopySolutions(NdLqrBatchSolver* bs, double* soln, int count) {
  if (!bs || !soln || count < 0 || count > bs->batch) return -1;
  const size_t per = (size_t)bs->N * (size_t)(2 * bs->n + bs->m);
  if (rslqr_cuda_download_solution(bs->dev, bs->stage_soln, 0, count) != RSLQR_OK) return -1;
  memcpy(soln, bs->stage_soln, sizeof(double) * per * (size_t)count);
  return 0;
}

int ndlqr_BatchCopySolution(NdLqrBatchSolver* bs, int problem, double* soln) {
  if (!bs || !soln || problem < 0 || problem >= bs->batch) return -1;
  if (rslqr_cuda_download_solution(bs->dev, bs->stage_soln, problem, 1) != RSLQR_OK) return -1;
  const int nvars = ndlqr_BatchGetNumVars(bs);
  memcpy(soln, bs->stage_soln, sizeof(double) * (size_t)nvars);
  return nvars;
}

int ndlqr_BatchGetStatus(NdLqrBatchSolver* bs, int* flags, int count) {
  if (!bs || !flags || count < 0 || count > bs->batch) return -1;
  return rslqr_cuda_download_flags(bs->dev, flags, 0, count) == RSLQR_OK ? 0 : -1;
}

double ndlqr_BatchGetSolveTimeMs(NdLqrBatchSolver* bs) { return bs ? bs->solve_ms : -1.0; }
void* ndlqr_BatchDeviceHandle(NdLqrBatchSolver* bs) { return bs ? (void*)bs->dev : NULL; }

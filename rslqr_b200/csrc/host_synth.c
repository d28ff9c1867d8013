/* host_synth.c -- reproducible synthetic LQR problems for the benchmark configurations
 * (the reference ships only n=6 problems, reference lqr_prob*.json).
 * Inputs only: the same buffers are handed to the GPU path and to the oracle. */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "rslqr_batch.h"

typedef struct { uint64_t state; } SplitMix;

static uint64_t sm_next(SplitMix* g) {
  uint64_t z = (g->state += 0x9E3779B97F4A7C15ULL);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}
static double sm_uni(SplitMix* g, double lo, double hi) {
  const double u = (double)(sm_next(g) >> 11) * (1.0 / 9007199254740992.0);
  return lo + (hi - lo) * u;
}

static void fill_one(int n, int m, int N, uint64_t seed, int id, double* knots, double* x0) {
  const double dt = 0.02;
  const int stride = ndlqr_LQRDataBlockSize(n, m);
  const int quad = (n == 12 && m == 4);
  SplitMix g = {seed * 0x100000001B3ULL + (uint64_t)id};
  for (int i = 0; i < n; ++i) x0[i] = sm_uni(&g, -1, 1);
  double mass = 1, J[3] = {1, 1, 1};
  if (quad) {
    mass = 0.5 * (1 + 0.2 * sm_uni(&g, -1, 1));
    const double J0[3] = {0.23, 0.23, 0.40};
    for (int a = 0; a < 3; ++a) J[a] = J0[a] * (1 + 0.2 * sm_uni(&g, -1, 1));
  }
  for (int k = 0; k < N; ++k) {
    double* kn = knots + (size_t)k * stride;
    double *Q = kn, *R = Q + n, *q = R + m, *r = q + n, *c = r + m, *A = c + 1, *B = A + n * n, *d = B + n * m;
    const int last = (k == N - 1);
    memset(kn, 0, sizeof(double) * (size_t)stride);
    for (int i = 0; i < n; ++i) {
      Q[i] = sm_uni(&g, 0.5, 2.0) * (last ? 10.0 : 1.0);
      q[i] = sm_uni(&g, -1, 1);
      d[i] = dt * sm_uni(&g, -0.1, 0.1);
    }
    for (int i = 0; i < m; ++i) {
      R[i] = sm_uni(&g, 0.05, 0.5);
      r[i] = sm_uni(&g, -1, 1);
    }
    for (int i = 0; i < n; ++i) A[i + n * i] = 1.0;
    if (quad) {
      const double gk = 9.81 * (1 + 0.05 * sm_uni(&g, -1, 1));
      const double L = 0.175, ctau = 0.0245;
      for (int a = 0; a < 3; ++a) {
        A[a + n * (6 + a)] = dt;        /* p <- v      */
        A[(3 + a) + n * (9 + a)] = dt;  /* theta <- w  */
      }
      A[6 + n * 4] = dt * gk;           /* v_x <- theta_y */
      A[7 + n * 3] = -dt * gk;          /* v_y <- theta_x */
      for (int j = 0; j < 4; ++j) {
        B[8 + n * j] = dt / mass;                              /* v_z  */
        B[11 + n * j] = ((j & 1) ? -1.0 : 1.0) * dt * ctau / J[2]; /* w_z */
      }
      B[9 + n * 1] = dt * L / J[0];
      B[9 + n * 3] = -dt * L / J[0];
      B[10 + n * 2] = dt * L / J[1];
      B[10 + n * 0] = -dt * L / J[1];
    } else {
      const int h = n / 2;
      const double sc = dt * 0.1 / sqrt((double)n);
      for (int i = 0; i < h; ++i) A[i + n * (h + i)] = dt;
      for (int e = 0; e < n * n; ++e) A[e] += sc * sm_uni(&g, -1, 1);
      for (int e = 0; e < n * m; ++e) B[e] = dt * sm_uni(&g, -1, 1);
    }
    if (last) { /* the terminal knot only contributes Q and q */
      memset(R, 0, sizeof(double) * (size_t)m);
      memset(r, 0, sizeof(double) * (size_t)m);
      memset(d, 0, sizeof(double) * (size_t)n);
      memset(A, 0, sizeof(double) * (size_t)(n * n + n * m));
    }
  }
}

int rslqr_synth_fill(int n, int m, int N, uint64_t seed, int first_id, int count, double* knots, double* x0) {
  if (n <= 0 || m <= 0 || N <= 0 || count < 0 || !knots || !x0) return -1;
  const size_t per = (size_t)N * (size_t)ndlqr_LQRDataBlockSize(n, m);
  for (int p = 0; p < count; ++p) fill_one(n, m, N, seed, first_id + p, knots + per * p, x0 + (size_t)n * p);
  return 0;
}

LQRProblem* rslqr_synth_problem(int n, int m, int N, uint64_t seed, int id) {
  LQRProblem* prob = ndlqr_NewLQRProblem(n, m, N);
  if (!prob) return NULL;
  const int stride = ndlqr_LQRDataBlockSize(n, m);
  double* flat = (double*)malloc(sizeof(double) * (size_t)N * stride);
  if (!flat) { ndlqr_FreeLQRProblem(prob); return NULL; }
  fill_one(n, m, N, seed, id, flat, prob->x0);
  for (int k = 0; k < N; ++k) memcpy(prob->lqrdata[k]->Q, flat + (size_t)k * stride, sizeof(double) * (size_t)stride);
  free(flat);
  return prob;
}

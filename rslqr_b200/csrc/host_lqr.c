/* host_lqr.c -- LQRData / LQRProblem containers (include/lqr_data.h, include/lqr_problem.h).
 * Semantics follow the reference's src/lqr_data.c:10-140 and src/lqr_problem.c:7-52: every knot
 * owns ONE heap block ordered Q R q r c A B d, which is also the "flat knot block" the CUDA layer
 * uploads verbatim (include/rslqr_cuda.h). */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "lqr_problem.h"

int ndlqr_LQRDataBlockSize(int n, int m) { return 2 * n + 2 * m + 1 + n * n + n * m + n; }

LQRData* ndlqr_NewLQRData(int nstates, int ninputs) {
  const int n = nstates, m = ninputs;
  LQRData* kd = (LQRData*)malloc(sizeof(LQRData));
  double* blk = (double*)malloc(sizeof(double) * (size_t)ndlqr_LQRDataBlockSize(n, m));
  if (!kd || !blk) {
    free(kd);
    free(blk);
    return NULL;
  }
  kd->nstates = n;
  kd->ninputs = m;
  kd->Q = blk;
  kd->R = kd->Q + n;
  kd->q = kd->R + m;
  kd->r = kd->q + n;
  kd->c = kd->r + m;
  kd->A = kd->c + 1;
  kd->B = kd->A + n * n;
  kd->d = kd->B + n * m;
  return kd;
}

int ndlqr_FreeLQRData(LQRData* lqrdata) {
  if (!lqrdata) return -1;
  free(lqrdata->Q); /* head of the single block */
  free(lqrdata);
  return 0;
}

int ndlqr_InitializeLQRData(LQRData* lqrdata, double* Q, double* R, double* q, double* r,
                            double c, double* A, double* B, double* d) {
  if (!lqrdata) return -1;
  const size_t n = (size_t)lqrdata->nstates, m = (size_t)lqrdata->ninputs;
  memcpy(lqrdata->Q, Q, n * sizeof(double));
  memcpy(lqrdata->R, R, m * sizeof(double));
  memcpy(lqrdata->q, q, n * sizeof(double));
  memcpy(lqrdata->r, r, m * sizeof(double));
  lqrdata->c[0] = c;
  memcpy(lqrdata->A, A, n * n * sizeof(double));
  memcpy(lqrdata->B, B, n * m * sizeof(double));
  memcpy(lqrdata->d, d, n * sizeof(double));
  return 0;
}

int ndlqr_CopyLQRData(LQRData* dest, LQRData* src) {
  if (dest->nstates != src->nstates || dest->ninputs != src->ninputs) {
    fprintf(stderr, "Can't copy LQRData of different sizes: (%d,%d) and (%d,%d).\n",
            dest->nstates, dest->ninputs, src->nstates, src->ninputs);
    return -1;
  }
  memcpy(dest->Q, src->Q,
         sizeof(double) * (size_t)ndlqr_LQRDataBlockSize(src->nstates, src->ninputs));
  return 0;
}

static Matrix view(int rows, int cols, double* p) {
  Matrix v = {rows, cols, p};
  return v;
}
Matrix ndlqr_GetA(LQRData* k) { return view(k->nstates, k->nstates, k->A); }
Matrix ndlqr_GetB(LQRData* k) { return view(k->nstates, k->ninputs, k->B); }
Matrix ndlqr_Getd(LQRData* k) { return view(k->nstates, 1, k->d); }
Matrix ndlqr_GetQ(LQRData* k) { return view(k->nstates, 1, k->Q); }
Matrix ndlqr_GetR(LQRData* k) { return view(k->ninputs, 1, k->R); }
Matrix ndlqr_Getq(LQRData* k) { return view(k->nstates, 1, k->q); }
Matrix ndlqr_Getr(LQRData* k) { return view(k->ninputs, 1, k->r); }

static void print_flat(const char* label, Matrix v) {
  printf("%s = [", label);
  for (int e = 0; e < v.rows * v.cols; ++e) printf("%6.2f ", v.data[e]);
  printf("]\n");
}

void ndlqr_PrintLQRData(LQRData* k) {
  printf("LQR Data with n=%d, m=%d:\n", k->nstates, k->ninputs);
  print_flat("Q", ndlqr_GetQ(k));
  print_flat("R", ndlqr_GetR(k));
  print_flat("q", ndlqr_Getq(k));
  print_flat("r", ndlqr_Getr(k));
  printf("c = %f\n", k->c[0]);
  Matrix A = ndlqr_GetA(k), B = ndlqr_GetB(k);
  printf("A:\n");
  PrintMatrix(&A);
  printf("B:\n");
  PrintMatrix(&B);
  print_flat("d", ndlqr_Getd(k));
}

/* ---- LQRProblem ---- */

LQRProblem* ndlqr_NewLQRProblem(int nstates, int ninputs, int nhorizon) {
  if (nhorizon <= 0) {
    fprintf(stderr, "ERROR: Horizon must be positive.\n");
    return NULL;
  }
  LQRProblem* prob = (LQRProblem*)malloc(sizeof(LQRProblem));
  LQRData** knots = (LQRData**)malloc(sizeof(LQRData*) * (size_t)nhorizon);
  if (!prob || !knots) {
    fprintf(stderr, "ERROR: Couldn't allocate memory for LQRProblem.\n");
    free(prob);
    free(knots);
    return NULL;
  }
  for (int k = 0; k < nhorizon; ++k) knots[k] = ndlqr_NewLQRData(nstates, ninputs);
  prob->nhorizon = nhorizon;
  prob->x0 = (double*)malloc(sizeof(double) * (size_t)nstates);
  prob->lqrdata = knots;
  return prob;
}

int ndlqr_InitializeLQRProblem(LQRProblem* lqrproblem, double* x0, LQRData** lqrdata) {
  if (!lqrproblem) return -1;
  for (int k = 0; k < lqrproblem->nhorizon; ++k) ndlqr_CopyLQRData(lqrproblem->lqrdata[k], lqrdata[k]);
  memcpy(lqrproblem->x0, x0, sizeof(double) * (size_t)lqrproblem->lqrdata[0]->nstates);
  return 0;
}

int ndlqr_FreeLQRProblem(LQRProblem* lqrprob) {
  if (!lqrprob) return -1;
  if (lqrprob->lqrdata)
    for (int k = 0; k < lqrprob->nhorizon; ++k) ndlqr_FreeLQRData(lqrprob->lqrdata[k]);
  free(lqrprob->x0);
  free(lqrprob->lqrdata);
  free(lqrprob);
  return 0;
}

/* lqr_data.h -- one knot point of a time-varying LQR problem (reference src/lqr_data.h:64-76).
 * cost  0.5 x'Qx + q'x + 0.5 u'Ru + r'u + c,   dynamics  x+ = A x + B u + d.
 * Q and R are DIAGONALS (length n and m). All arrays live in ONE heap block that starts at `Q`,
 * in the order Q R q r c A B d (reference src/lqr_data.c:24-36); the batched CUDA upload relies on
 * that order because it copies the block verbatim (see rslqr_cuda.h, "flat knot block"). */
#ifndef RSLQR_LQR_DATA_H
#define RSLQR_LQR_DATA_H
#include "matrix.h"
#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
  int nstates;
  int ninputs;
  double* Q; /* (n)   diagonal of the state cost Hessian   */
  double* R; /* (m)   diagonal of the input cost Hessian   */
  double* q; /* (n)   */
  double* r; /* (m)   */
  double* c; /* (1)   */
  double* A; /* (n,n) column-major */
  double* B; /* (n,m) column-major */
  double* d; /* (n)   */
} LQRData;

LQRData* ndlqr_NewLQRData(int nstates, int ninputs);
int ndlqr_FreeLQRData(LQRData* lqrdata);
int ndlqr_InitializeLQRData(LQRData* lqrdata, double* Q, double* R, double* q, double* r,
                            double c, double* A, double* B, double* d); /* copies in */
int ndlqr_CopyLQRData(LQRData* dest, LQRData* src); /* -1 on dimension mismatch */
Matrix ndlqr_GetA(LQRData* lqrdata);
Matrix ndlqr_GetB(LQRData* lqrdata);
Matrix ndlqr_Getd(LQRData* lqrdata);
Matrix ndlqr_GetQ(LQRData* lqrdata);
Matrix ndlqr_GetR(LQRData* lqrdata);
Matrix ndlqr_Getq(LQRData* lqrdata);
Matrix ndlqr_Getr(LQRData* lqrdata);
void ndlqr_PrintLQRData(LQRData* lqrdata);
/* doubles in the single heap block of one knot: 2n + 2m + 1 + n*n + n*m + n (additive helper) */
int ndlqr_LQRDataBlockSize(int nstates, int ninputs);

#ifdef __cplusplus
}
#endif
#endif

/* lqr_problem.h -- a whole LQR problem: horizon, initial state, one LQRData per knot
 * (reference src/lqr_problem.h:31-35). The terminal knot only contributes Q and q. */
#ifndef RSLQR_LQR_PROBLEM_H
#define RSLQR_LQR_PROBLEM_H
#include "lqr_data.h"
#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
  int nhorizon;
  double* x0;
  LQRData** lqrdata;
} LQRProblem;

LQRProblem* ndlqr_NewLQRProblem(int nstates, int ninputs, int nhorizon); /* NULL if nhorizon<=0 */
int ndlqr_InitializeLQRProblem(LQRProblem* lqrproblem, double* x0, LQRData** lqrdata);
int ndlqr_FreeLQRProblem(LQRProblem* lqrprob);

#ifdef __cplusplus
}
#endif
#endif

/* solve.h -- the solve entry points (reference src/solve.h:37,59,72; src/solve.c:38,192,197).
 *
 * ndlqr_Solve keeps the reference contract: destructive (soln holds the rhs on entry and the
 * solution [lambda_0,x_0,u_0,lambda_1,...,x_{N-1}] on return), blocking, returns 0 (or -1 when the device is
 * unavailable). The reference additionally requires `fact` to have been reset since the previous solve
 * (src/solve.h:28-32); here every solve starts from a cleared device factor array by itself. The work is done
 * by the sm_100a kernels behind rslqr_cuda.h. */
#ifndef RSLQR_SOLVE_H
#define RSLQR_SOLVE_H
#include "solver.h"
#ifdef __cplusplus
extern "C" {
#endif

int ndlqr_Solve(NdLqrSolver* solver);
Matrix ndlqr_GetSolution(NdLqrSolver* solver);              /* (nvars,1) view, solver-owned */
int ndlqr_CopySolution(NdLqrSolver* solver, double* soln);  /* returns nvars, -1 on NULL */

/* Convenience named by the task statement (no reference counterpart): initialize + solve. */
int ndlqr_SolveLQRProblem(const LQRProblem* lqrprob, NdLqrSolver* solver);

/* Factor once, solve many (no reference counterpart: docs/rslqr_usage.dox:53-54 notes that the reference cannot
 * separate the two halves). After ndlqr_InitializeWithLQRProblem (and usually an ndlqr_Solve), write a NEW right-hand
 * side into solver->soln->data (same layout and sign convention as Initialize leaves there, src/solver.c:138-190) and
 * call this: only the rhs phase (src/solve.c:141-181) runs, on the staged factorization the device holds (computed on
 * the first call if the last solve took the on-chip path). solver->fact is not touched. Returns 0, or -1 when no
 * problem has been initialized or there is no device. */
int ndlqr_SolveRhsOnly(NdLqrSolver* solver);

/* Refresh solver->fact with the reference's dense factor array (F(k, level) in NdData order, src/nddata.c:82-93) of
 * the initialized problem. ndlqr_Solve does not do this by default (see solver->mirror_fact): the array is
 * N log2(N) (2n+m) n doubles and nothing on the solve path reads it back. The stage functions of
 * nested_dissection.h keep the mirror current themselves. Returns 0 or -1. */
int ndlqr_SyncFactMirror(NdLqrSolver* solver);

#ifdef __cplusplus
}
#endif
#endif

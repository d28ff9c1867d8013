/* solve.h -- the solve entry points (reference src/solve.h:37,59,72; src/solve.c:38,192,197).
 *
 * ndlqr_Solve keeps the reference contract: destructive (soln holds the rhs on entry and the
 * solution [lambda_0,x_0,u_0,lambda_1,...,x_{N-1}] on return), `fact` must have been reset since
 * the previous solve, blocking, returns 0 (or -1 when the device is unavailable). The work is done
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
 * separate the two halves). After an ndlqr_Solve, write a NEW right-hand side into solver->soln->data (same layout
 * and sign convention as ndlqr_InitializeWithLQRProblem leaves there, src/solver.c:138-190) and call this: only the
 * rhs phase (src/solve.c:141-181) runs, on the factorization the device still holds. solver->fact is not
 * touched. Returns 0, or -1 when there is no factorization (no previous solve, or a solver created with
 * mirror_fact == 0) or no device. */
int ndlqr_SolveRhsOnly(NdLqrSolver* solver);

#ifdef __cplusplus
}
#endif
#endif

/* solver.h -- the rsLQR solver object (reference src/solver.h:31-39,114-130, src/solver.c).
 *
 * Same public fields, in the same order, as the reference's NdLqrSolver; `device` is appended.
 * Host arrays (`data`, `fact`, `soln`, `diagonals`, `cholfacts`) are MIRRORS: the arithmetic of
 * ndlqr_Solve runs on the GPU (rslqr_cuda.h) and the mirrors are refreshed when it returns.
 * There is no CPU solve path: every entry point that computes fails with -1 (and a message on
 * stderr) when no CUDA device / context is available. */
#ifndef RSLQR_SOLVER_H
#define RSLQR_SOLVER_H
#include "binary_tree.h"
#include "cholesky_factors.h"
#include "linalg.h"
#include "lqr_problem.h"
#include "nddata.h"
#ifdef __cplusplus
extern "C" {
#endif

/* Per-phase times of the last solve in ms. The reference fills these from omp_get_wtime around
 * each OpenMP phase (src/solve.c:15-25); here they are CUDA-event times of the per-phase kernel
 * launches (staged mode) and t_total_ms is the host wall time of ndlqr_Solve. */
typedef struct {
  double t_total_ms;
  double t_leaves_ms;
  double t_products_ms;
  double t_cholesky_ms;
  double t_cholsolve_ms;
  double t_shur_ms;
  int num_threads;
} NdLqrProfile;

NdLqrProfile ndlqr_NewNdLqrProfile(void);
void ndlqr_ResetProfile(NdLqrProfile* prof);
void ndlqr_CopyProfile(NdLqrProfile* dest, NdLqrProfile* src);
void ndlqr_PrintProfile(NdLqrProfile* profile);
void ndlqr_CompareProfile(NdLqrProfile* base, NdLqrProfile* prof);

typedef struct {
  int nstates;
  int ninputs;
  int nhorizon;
  int depth;
  int nvars;
  OrderedBinaryTree tree;
  Matrix* diagonals;  /* [2k] = Q_k (n x n), [2k+1] = R_k (m x m); hold the Cholesky factors after a solve */
  NdData* data;       /* KKT column blocks (host mirror, filled by Initialize) */
  NdData* fact;       /* factorization (host mirror, refreshed after Solve) */
  NdData* soln;       /* rhs before Solve, solution after */
  NdLqrCholeskyFactors* cholfacts;
  double solve_time_ms;
  double linalg_time_ms;
  NdLqrProfile profile;
  int num_threads;    /* kept for API compatibility; the GPU path ignores it */
  /* ---- appended by this build ---- */
  void* device;       /* rslqr_cuda_solver* (NULL when no CUDA device could be opened) */
  double* flat_knots; /* staging copy of the problem in flat knot-block form */
  double* flat_x0;
  int mirror_fact;    /* 0 (default): ndlqr_Solve returns the solution (on-chip kernels), solver->fact is refreshed by
                         ndlqr_SyncFactMirror / the stage calls; 1 (or RSLQR_MIRROR_FACT=1 in the environment):
                         every ndlqr_Solve runs the staged kernels and copies the dense `fact` array back */
  int pinned_fact, pinned_soln, pinned_knots; /* the mirrors are page-locked (rslqr_cuda_host_register) */
  int initialized;    /* a problem has been uploaded (ndlqr_InitializeWithLQRProblem) */
  int fact_dirty;     /* the factor mirror has been written since the last reset */
  int* status_host;   /* pinned staging buffer for the 3N-1 status words (cholinfo[].success) */
} NdLqrSolver;

NdLqrSolver* ndlqr_NewNdLqrSolver(int nstates, int ninputs, int nhorizon);
int ndlqr_FreeNdLqrSolver(NdLqrSolver* solver);
int ndlqr_InitializeWithLQRProblem(const LQRProblem* lqrprob, NdLqrSolver* solver);
void ndlqr_ResetSolver(NdLqrSolver* solver);
void ndlqr_PrintSolveSummary(NdLqrSolver* solver);
int ndlqr_GetNumVars(NdLqrSolver* solver);
int ndlqr_SetNumThreads(NdLqrSolver* solver, int num_threads);
int ndlqr_GetNumThreads(NdLqrSolver* solver);
int ndlqr_PrintSolveProfile(NdLqrSolver* solver);
NdLqrProfile ndlqr_GetProfile(NdLqrSolver* solver);

#ifdef __cplusplus
}
#endif
#endif

/* host_solver.c -- the solver object and the solve entry points of the public C API
 * (include/solver.h, solve.h, nested_dissection.h), bound to the CUDA layer (include/rslqr_cuda.h).
 *
 * What stays on the host: allocation of the mirrors, copying the caller's LQRProblem into the
 * KKT mirror / rhs exactly where the reference puts it (reference src/solver.c:122-194), and moving
 * bytes to and from the device. All arithmetic of the solve happens in the kernels. */
#include <math.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>

#include "nested_dissection.h"
#include "rslqr_cuda.h"
#include "rslqr_host_internal.h"
#include "solve.h"
#include "tree_math.h"

/* ---- registry of live solvers (lets the reference-signature stage calls find their solver) ----
 * A growing array behind a mutex: any number of solvers, created and freed from any thread. */
static NdLqrSolver** g_live = NULL;
static int g_live_cap = 0;
static pthread_mutex_t g_live_mu = PTHREAD_MUTEX_INITIALIZER;

static int registry_add(NdLqrSolver* s) {
  int ok = 0;
  pthread_mutex_lock(&g_live_mu);
  for (int i = 0; i < g_live_cap && !ok; ++i)
    if (!g_live[i]) { g_live[i] = s; ok = 1; }
  if (!ok) {
    const int cap = g_live_cap ? 2 * g_live_cap : 64;
    NdLqrSolver** grown = (NdLqrSolver**)realloc(g_live, sizeof(NdLqrSolver*) * (size_t)cap);
    if (grown) {
      memset(grown + g_live_cap, 0, sizeof(NdLqrSolver*) * (size_t)(cap - g_live_cap));
      grown[g_live_cap] = s;
      g_live = grown;
      g_live_cap = cap;
      ok = 1;
    }
  }
  pthread_mutex_unlock(&g_live_mu);
  return ok ? 0 : -1;
}
static void registry_remove(NdLqrSolver* s) {
  pthread_mutex_lock(&g_live_mu);
  for (int i = 0; i < g_live_cap; ++i)
    if (g_live[i] == s) g_live[i] = NULL;
  pthread_mutex_unlock(&g_live_mu);
}
NdLqrSolver* rslqr_owner_of(const NdData* nd, int* which) {
  NdLqrSolver* found = NULL;
  pthread_mutex_lock(&g_live_mu);
  for (int i = 0; i < g_live_cap && !found; ++i) {
    NdLqrSolver* s = g_live[i];
    if (!s) continue;
    if (s->data == nd) { if (which) *which = 0; found = s; }
    else if (s->fact == nd) { if (which) *which = 1; found = s; }
    else if (s->soln == nd) { if (which) *which = 2; found = s; }
  }
  pthread_mutex_unlock(&g_live_mu);
  return found;
}

static double wall_ms(void) {
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return 1e3 * (double)ts.tv_sec + 1e-6 * (double)ts.tv_nsec;
}

/* ---- profile ---- */
NdLqrProfile ndlqr_NewNdLqrProfile(void) {
  NdLqrProfile p = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, -1};
  return p;
}
void ndlqr_ResetProfile(NdLqrProfile* p) {
  p->t_total_ms = p->t_leaves_ms = p->t_products_ms = 0.0;
  p->t_cholesky_ms = p->t_cholsolve_ms = p->t_shur_ms = 0.0;
}
void ndlqr_CopyProfile(NdLqrProfile* dest, NdLqrProfile* src) { *dest = *src; }
void ndlqr_PrintProfile(NdLqrProfile* p) {
  printf("Solved with %d threads\n", p->num_threads);
  printf("Solve Total:    %.3f ms\n", p->t_total_ms);
  printf("Solve Leaves:   %.3f ms\n", p->t_leaves_ms);
  printf("Solve Products: %.3f ms\n", p->t_products_ms);
  printf("Solve Cholesky: %.3f ms\n", p->t_cholesky_ms);
  printf("Solve Solve:    %.3f ms\n", p->t_cholsolve_ms);
  printf("Solve Shur:     %.3f ms\n", p->t_shur_ms);
}
static void ratio_line(const char* label, double base, double now) {
  printf("%s%.3f / %.3f (%.2f speedup)\n", label, base, now, base / now);
}
void ndlqr_CompareProfile(NdLqrProfile* base, NdLqrProfile* prof) {
  printf("Num Threads:     %d / %d\n", base->num_threads, prof->num_threads);
  ratio_line("Solve Total:     ", base->t_total_ms, prof->t_total_ms);
  ratio_line("Solve Leaves:    ", base->t_leaves_ms, prof->t_leaves_ms);
  ratio_line("Solve Products:  ", base->t_products_ms, prof->t_products_ms);
  ratio_line("Solve Cholesky:  ", base->t_cholesky_ms, prof->t_cholesky_ms);
  ratio_line("Solve CholSolve: ", base->t_cholsolve_ms, prof->t_cholsolve_ms);
  ratio_line("Solve Shur Comp: ", base->t_shur_ms, prof->t_shur_ms);
}

/* ---- lifecycle ---- */
NdLqrSolver* ndlqr_NewNdLqrSolver(int nstates, int ninputs, int nhorizon) {
  const int n = nstates, m = ninputs, N = nhorizon;
  if (n <= 0 || m <= 0 || !IsPowerOfTwo(N) || N < 2) {
    fprintf(stderr, "ERROR: rsLQR needs positive dimensions and a power-of-two horizon >= 2.\n");
    return NULL;
  }
  NdLqrSolver* s = (NdLqrSolver*)calloc(1, sizeof(NdLqrSolver));
  if (!s) return NULL;
  s->nstates = n;
  s->ninputs = m;
  s->nhorizon = N;
  s->tree = ndlqr_BuildTree(N);
  s->depth = s->tree.depth;
  s->nvars = (2 * n + m) * N - m;
  const int per_knot = n * n + m * m;
  double* dstore = (double*)calloc((size_t)per_knot * N, sizeof(double));
  s->diagonals = (Matrix*)malloc(sizeof(Matrix) * 2 * (size_t)N);
  for (int k = 0; k < N; ++k) {
    const Matrix Qk = {n, n, dstore + (size_t)k * per_knot};
    const Matrix Rk = {m, m, dstore + (size_t)k * per_knot + n * n};
    s->diagonals[2 * k] = Qk;
    s->diagonals[2 * k + 1] = Rk;
  }
  s->data = ndlqr_NewNdData(n, m, N, n);
  s->fact = ndlqr_NewNdData(n, m, N, n);
  s->soln = ndlqr_NewNdData(n, m, N, 1);
  s->cholfacts = ndlqr_NewCholeskyFactors(s->depth, N);
  s->profile = ndlqr_NewNdLqrProfile();
  const long procs = sysconf(_SC_NPROCESSORS_ONLN);
  s->num_threads = (int)(procs > 1 ? procs / 2 : 1);
  s->flat_knots = (double*)calloc((size_t)N * ndlqr_LQRDataBlockSize(n, m), sizeof(double));
  s->flat_x0 = (double*)calloc((size_t)n, sizeof(double));
  /* The dense factor mirror solver->fact is refreshed on request (ndlqr_SyncFactMirror, the stage calls), not by every
   * ndlqr_Solve: the solve then runs the on-chip kernels and moves only the solution. RSLQR_MIRROR_FACT=1 (or
   * solver->mirror_fact = 1) restores "every solve leaves the reference's fact array in solver->fact". */
  const char* mf = getenv("RSLQR_MIRROR_FACT");
  s->mirror_fact = mf ? (atoi(mf) != 0) : 0;
  rslqr_cuda_solver* dev = NULL;
  if (rslqr_cuda_create(n, m, N, 1, 0, RSLQR_PROFILE, &dev) != RSLQR_OK) {
    fprintf(stderr, "ERROR: rsLQR could not open a CUDA device (%s). ndlqr_Solve will fail; there is no CPU path.\n",
            rslqr_cuda_last_error());
    dev = NULL;
  }
  s->device = dev;
  if (dev) {
    /* page-lock the mirrors the solve copies to and from (best effort: a refused range just stays pageable); the
     * factor mirror (N log2(N) blocks: 617 MB at n=12, N=16384) stays untouched calloc memory until it is first
     * downloaded into (pin_fact_mirror) */
    s->pinned_soln = rslqr_cuda_host_register(s->soln->data, sizeof(double) * (size_t)(2 * n + m) * (size_t)N) == RSLQR_OK;
    s->pinned_knots = rslqr_cuda_host_register(s->flat_knots, sizeof(double) * (size_t)N * ndlqr_LQRDataBlockSize(n, m)) == RSLQR_OK;
    s->status_host = (int*)rslqr_cuda_host_alloc(sizeof(int) * 3 * (size_t)N);
  }
  if (registry_add(s) != 0) fprintf(stderr, "WARNING: rsLQR solver registry is out of memory; stage calls on this solver will fail.\n");
  return s;
}

static void pin_fact_mirror(NdLqrSolver* s) {
  if (s->pinned_fact || !s->device) return;
  const size_t fs = (size_t)(2 * s->nstates + s->ninputs) * (size_t)s->nstates;
  s->pinned_fact = rslqr_cuda_host_register(s->fact->data, sizeof(double) * fs * (size_t)s->nhorizon * (size_t)s->depth) == RSLQR_OK;
}

int ndlqr_FreeNdLqrSolver(NdLqrSolver* solver) {
  if (!solver) return -1;
  registry_remove(solver);
  if (solver->pinned_fact) rslqr_cuda_host_unregister(solver->fact->data);
  if (solver->pinned_soln) rslqr_cuda_host_unregister(solver->soln->data);
  if (solver->pinned_knots) rslqr_cuda_host_unregister(solver->flat_knots);
  if (solver->status_host) rslqr_cuda_host_free(solver->status_host);
  if (solver->device) rslqr_cuda_destroy((rslqr_cuda_solver*)solver->device);
  ndlqr_FreeTree(&solver->tree);
  ndlqr_FreeNdData(solver->data);
  ndlqr_FreeNdData(solver->fact);
  ndlqr_FreeNdData(solver->soln);
  ndlqr_FreeCholeskyFactors(solver->cholfacts);
  free(solver->diagonals[0].data);
  free(solver->diagonals);
  free(solver->flat_knots);
  free(solver->flat_x0);
  free(solver);
  return 0;
}

void ndlqr_ResetSolver(NdLqrSolver* solver) {
  ndlqr_ResetNdData(solver->data);
  /* the factor mirror is only ever written by a download (fact_dirty): an untouched one is still all zeros */
  if (solver->fact_dirty) ndlqr_ResetNdData(solver->fact);
  solver->fact_dirty = 0;
  ndlqr_ResetNdData(solver->soln);
  ndlqr_ResetProfile(&solver->profile);
  for (int i = 0; i < 2 * solver->nhorizon; ++i) MatrixSetConst(&solver->diagonals[i], 0.0);
  if (solver->device) rslqr_cuda_reset((rslqr_cuda_solver*)solver->device);
}

/* Copy the problem into the mirrors where the reference puts it (reference src/solver.c:122-194)
 * and ship the flat knot blocks to the device. */
int ndlqr_InitializeWithLQRProblem(const LQRProblem* lqrprob, NdLqrSolver* solver) {
  const int n = solver->nstates, m = solver->ninputs, N = solver->nhorizon;
  if (lqrprob->nhorizon != N) return -1;
  const int blk = ndlqr_LQRDataBlockSize(n, m);
  NdFactor* col;
  NdFactor* rhs;
  ndlqr_GetNdFactor(solver->soln, 0, 0, &rhs);
  memcpy(rhs->lambda.data, lqrprob->x0, sizeof(double) * (size_t)n);
  memcpy(solver->flat_x0, lqrprob->x0, sizeof(double) * (size_t)n);
  for (int k = 0; k < N; ++k) {
    LQRData* kd = lqrprob->lqrdata[k];
    const int terminal = (k == N - 1);
    if (!terminal && (kd->nstates != n || kd->ninputs != m)) return -1;
    memcpy(solver->flat_knots + (size_t)k * blk, kd->Q, sizeof(double) * (size_t)blk);
    ndlqr_GetNdFactor(solver->soln, k, 0, &rhs);
    memcpy(rhs->state.data, kd->q, sizeof(double) * (size_t)n);
    Matrix Qk = solver->diagonals[2 * k];
    MatrixSetConst(&Qk, 0.0);
    for (int i = 0; i < n; ++i) Qk.data[i + n * i] = kd->Q[i];
    if (terminal) break;
    memcpy(rhs->input.data, kd->r, sizeof(double) * (size_t)m);
    Matrix Rk = solver->diagonals[2 * k + 1];
    MatrixSetConst(&Rk, 0.0);
    for (int i = 0; i < m; ++i) Rk.data[i + m * i] = kd->R[i];
    /* column of the separator after knot k: (A_k^T, B_k^T) at knot k, (-I, 0) at knot k+1 */
    const int lv = rs_knot_level(k);
    ndlqr_GetNdFactor(solver->data, k, lv, &col);
    Matrix A = {n, n, kd->A}, B = {n, m, kd->B};
    MatrixCopyTranspose(&col->state, &A);
    MatrixCopyTranspose(&col->input, &B);
    ndlqr_GetNdFactor(solver->data, k + 1, lv, &col);
    MatrixSetConst(&col->state, 0.0);
    for (int i = 0; i < n; ++i) col->state.data[i + n * i] = -1.0;
    MatrixSetConst(&col->input, 0.0);
    ndlqr_GetNdFactor(solver->soln, k + 1, 0, &rhs);
    memcpy(rhs->lambda.data, kd->d, sizeof(double) * (size_t)n);
  }
  for (int i = 0; i < solver->nvars; ++i) solver->soln->data[i] *= -1; /* rhs = -[...] */
  if (solver->device &&
      rslqr_cuda_upload((rslqr_cuda_solver*)solver->device, solver->flat_knots, solver->flat_x0, 0, 1) != RSLQR_OK) {
    fprintf(stderr, "ERROR: upload to the GPU failed: %s\n", rslqr_cuda_last_error());
    return -1;
  }
  if (solver->device) rslqr_cuda_invalidate_factorization((rslqr_cuda_solver*)solver->device); /* new Q, R, A, B */
  solver->initialized = 1;
  return 0;
}

/* ---- mirrors <-> device ---- */
static int push_rhs(NdLqrSolver* s) {
  return rslqr_cuda_upload_rhs_async((rslqr_cuda_solver*)s->device, s->soln->data, 0, 1);
}

/* The leaf Cholesky factors the reference leaves in `diagonals` (sqrt of the diagonal Q_k, R_k, reference
 * src/nested_dissection.c:61-73) are mirror bookkeeping like `fact`: refreshed together with it (mirror_fact,
 * ndlqr_SyncFactMirror, stage calls), from the flat knot blocks, since only the staged kernels materialise them on
 * the device. */
static void refresh_leaf_factors(NdLqrSolver* s) {
  const int n = s->nstates, m = s->ninputs, N = s->nhorizon;
  const int blk = ndlqr_LQRDataBlockSize(n, m);
  for (int k = 0; k < N; ++k) {
    const double* kd = s->flat_knots + (size_t)k * blk; /* Q (n) then R (m) lead the block */
    for (int i = 0; i < n; ++i) s->diagonals[2 * k].data[i + n * i] = kd[i] > 0 ? sqrt(kd[i]) : kd[i];
    if (k < N - 1)
      for (int i = 0; i < m; ++i) s->diagonals[2 * k + 1].data[i + m * i] = kd[n + i] > 0 ? sqrt(kd[n + i]) : kd[n + i];
  }
}

static int pull_results(NdLqrSolver* s, int with_fact) {
  rslqr_cuda_solver* dev = (rslqr_cuda_solver*)s->device;
  const int N = s->nhorizon;
  /* the solution and the status words were queued behind the solve by ndlqr_Solve; one sync covers them */
  int* status = s->status_host; /* pinned, allocated with the solver */
  if (!status) return RSLQR_E_NOMEM;
  int rc = rslqr_cuda_download_solution_async(dev, s->soln->data, 0, 1);
  if (!rc) rc = rslqr_cuda_download_status_async(dev, status, 0);
  if (!rc) rc = rslqr_cuda_sync(dev);
  if (!rc && with_fact) {
    pin_fact_mirror(s);
    rc = rslqr_cuda_download_fact(dev, s->fact->data, 0);
    s->fact_dirty = 1;
  }
  if (rc) return rc;
  /* status words -> cholinfo[].success */
  for (int f = 0; f < s->cholfacts->numfacts; ++f) {
    s->cholfacts->cholinfo[f].success = status[f];
    s->cholfacts->cholinfo[f].uplo = 'L';
    s->cholfacts->cholinfo[f].lib = 'C';
  }
  if (with_fact) refresh_leaf_factors(s);
  (void)N;
  return rc;
}

/* ---- solve ---- */
int ndlqr_Solve(NdLqrSolver* solver) {
  if (!solver) return -1;
  if (!solver->device) {
    fprintf(stderr, "ERROR: ndlqr_Solve: no CUDA device is attached to this solver (no CPU fallback).\n");
    return -1;
  }
  rslqr_cuda_solver* dev = (rslqr_cuda_solver*)solver->device;
  const double t0 = wall_ms();
  int rc = push_rhs(solver);
  if (!rc) rc = rslqr_cuda_set_mode(dev, solver->mirror_fact ? RSLQR_MODE_STAGED : RSLQR_MODE_AUTO);
  if (!rc) rc = rslqr_cuda_solve_async(dev, 0);
  if (!rc) rc = pull_results(solver, solver->mirror_fact);
  if (rc) {
    fprintf(stderr, "ERROR: ndlqr_Solve failed on the GPU: %s\n", rslqr_cuda_last_error());
    return -1;
  }
  double ms[7] = {0};
  rslqr_cuda_get_profile(dev, ms);
  solver->profile.t_leaves_ms += ms[1];
  solver->profile.t_products_ms += ms[2];
  solver->profile.t_cholesky_ms += ms[3];
  solver->profile.t_cholsolve_ms += ms[4];
  solver->profile.t_shur_ms += ms[5];
  solver->solve_time_ms = wall_ms() - t0;
  solver->linalg_time_ms = ms[0];
  solver->profile.t_total_ms = solver->solve_time_ms;
  solver->profile.num_threads = solver->num_threads;
  return 0;
}

int ndlqr_SolveRhsOnly(NdLqrSolver* solver) {
  if (!solver) return -1;
  if (!solver->device) {
    fprintf(stderr, "ERROR: ndlqr_SolveRhsOnly: no CUDA device is attached to this solver (no CPU fallback).\n");
    return -1;
  }
  if (!solver->initialized) return -1; /* nothing to factor yet */
  rslqr_cuda_solver* dev = (rslqr_cuda_solver*)solver->device;
  const double t0 = wall_ms();
  /* the rhs phase runs on the staged factorization; a solver whose last solve took the on-chip path (which keeps
   * none) factors here, once */
  int rc = rslqr_cuda_set_mode(dev, RSLQR_MODE_STAGED);
  if (!rc && !rslqr_cuda_has_factorization(dev)) rc = rslqr_cuda_factor_async(dev);
  if (!rc) rc = push_rhs(solver);
  if (!rc) rc = rslqr_cuda_solve_rhs_async(dev, 0);
  if (!rc) rc = pull_results(solver, 0);
  if (rc) {
    fprintf(stderr, "ERROR: ndlqr_SolveRhsOnly failed: %s\n", rslqr_cuda_last_error());
    return -1;
  }
  double ms[7] = {0};
  rslqr_cuda_get_profile(dev, ms);
  solver->solve_time_ms = wall_ms() - t0;
  solver->linalg_time_ms = ms[0];
  return 0;
}

/* solver->fact <- the reference's factor array for the problem the solver was initialized with (what ndlqr_Solve of
 * the reference leaves there, src/solve.c:57-133): downloaded if the device holds a staged factorization, else
 * factored first by the staged kernels. */
int ndlqr_SyncFactMirror(NdLqrSolver* solver) {
  if (!solver || !solver->device || !solver->initialized) return -1;
  rslqr_cuda_solver* dev = (rslqr_cuda_solver*)solver->device;
  int rc = rslqr_cuda_set_mode(dev, RSLQR_MODE_STAGED);
  if (!rc && !rslqr_cuda_has_factorization(dev)) rc = rslqr_cuda_factor_async(dev);
  if (!rc) {
    pin_fact_mirror(solver);
    rc = rslqr_cuda_download_fact(dev, solver->fact->data, 0);
    solver->fact_dirty = 1;
    refresh_leaf_factors(solver);
  }
  if (rc) fprintf(stderr, "ERROR: ndlqr_SyncFactMirror failed: %s\n", rslqr_cuda_last_error());
  return rc ? -1 : 0;
}

int ndlqr_SolveLQRProblem(const LQRProblem* lqrprob, NdLqrSolver* solver) {
  if (!lqrprob || !solver) return -1;
  ndlqr_ResetSolver(solver);
  if (ndlqr_InitializeWithLQRProblem(lqrprob, solver) != 0) return -1;
  return ndlqr_Solve(solver);
}

Matrix ndlqr_GetSolution(NdLqrSolver* solver) {
  Matrix view = {solver->nvars, 1, solver->soln->data};
  return view;
}

int ndlqr_CopySolution(NdLqrSolver* solver, double* soln) {
  if (!solver) return -1;
  memcpy(soln, solver->soln->data, sizeof(double) * (size_t)solver->nvars);
  return solver->nvars;
}

/* ---- accessors ---- */
void ndlqr_PrintSolveSummary(NdLqrSolver* solver) {
  printf("rsLQR Solve Summary\n");
  printf("-------------------\n");
  printf("  rsLQR: nested-dissection LQR solver (B200 CUDA build of the RExLab algorithm).\n\n");
  printf("  Solve time:  %f ms\n", solver->solve_time_ms);
  printf("  Kernel time: %f ms\n", solver->linalg_time_ms);
  printf("  Solved with %d threads.\n", solver->num_threads);
  printf("  ");
  MatrixPrintLinearAlgebraLibrary();
}
int ndlqr_GetNumVars(NdLqrSolver* solver) { return solver->nvars; }
int ndlqr_SetNumThreads(NdLqrSolver* solver, int num_threads) {
  if (!solver) return -1;
  solver->num_threads = num_threads;
  return 0;
}
int ndlqr_GetNumThreads(NdLqrSolver* solver) { return solver ? solver->num_threads : -1; }
int ndlqr_PrintSolveProfile(NdLqrSolver* solver) {
  if (!solver) return -1;
  ndlqr_PrintProfile(&solver->profile);
  return 0;
}
NdLqrProfile ndlqr_GetProfile(NdLqrSolver* solver) { return solver->profile; }

/* ---- stage calls (reference src/nested_dissection.c) ---------------------------------------
 * The HOST mirrors are authoritative between calls: solver->soln and solver->fact are pushed before and pulled after
 * each call, so a caller can interleave stage calls with reads of, and helper calls on, views of solver->fact / soln
 * (the reference's own test does: MatrixCholeskyFactorize / MatrixCholeskySolve on F->lambda between
 * ndlqr_FactorInnerProduct and ndlqr_ComputeShurCompliment, test/nested_dissection_test.c:138-190). */
static int push_fact(NdLqrSolver* s) {
  pin_fact_mirror(s);
  s->fact_dirty = 1;
  return rslqr_cuda_upload_fact((rslqr_cuda_solver*)s->device, s->fact->data, 0);
}

static int run_stage(NdLqrSolver* s, int stage, int level, int item, int col) {
  if (!s) return -1;
  if (!s->device) {
    fprintf(stderr, "ERROR: stage call without a CUDA device (no CPU fallback).\n");
    return -1;
  }
  rslqr_cuda_solver* dev = (rslqr_cuda_solver*)s->device;
  int rc = push_rhs(s);
  if (!rc) rc = push_fact(s);
  if (!rc) rc = rslqr_cuda_stage(dev, stage, level, item, col, 0);
  if (!rc) rc = pull_results(s, 1);
  if (rc) fprintf(stderr, "ERROR: GPU stage %d failed: %s\n", stage, rslqr_cuda_last_error());
  return rc ? -1 : 0;
}

int ndlqr_SolveLeaf(NdLqrSolver* solver, int index) {
  if (!solver || index < 0 || index >= solver->nhorizon) return -1;
  return run_stage(solver, RSLQR_STAGE_LEAVES, 0, index, -1);
}

int ndlqr_SolveLeaves(NdLqrSolver* solver) { return run_stage(solver, RSLQR_STAGE_LEAVES, 0, -1, -1); }

bool ndlqr_ShouldCalcLambda(OrderedBinaryTree* tree, int index, int i) {
  const BinaryNode* nd = tree->node_list + index;
  const bool opens_half = (i == nd->left_inds.start) || (i == nd->right_inds.start);
  return !opens_half || i == 0;
}

static int leaf_of_index(int index, int level) { return index >> (level + 1); }

int ndlqr_SolverFactorInnerProduct(NdLqrSolver* solver, int index, int data_level, int fact_level, int use_rhs) {
  if (!solver || rs_sep_index(leaf_of_index(index, data_level), data_level) != index) return -1;
  return run_stage(solver, use_rhs ? RSLQR_STAGE_RHS_PRODUCTS : RSLQR_STAGE_PRODUCTS, data_level,
                   leaf_of_index(index, data_level), use_rhs ? -1 : fact_level);
}

int ndlqr_SolverFactorCholesky(NdLqrSolver* solver, int leaf, int level) {
  return run_stage(solver, RSLQR_STAGE_CHOLESKY, level, leaf, -1);
}

int ndlqr_SolverSolveCholeskyFactor(NdLqrSolver* solver, int index, int level, int upper_level, int use_rhs) {
  if (!solver || rs_sep_index(leaf_of_index(index, level), level) != index) return -1;
  if (!use_rhs && upper_level <= level) {
    fprintf(stderr, "ERROR: `upper_level` must be greater than `level`.");
    return -1;
  }
  return run_stage(solver, use_rhs ? RSLQR_STAGE_RHS_SOLVES : RSLQR_STAGE_SOLVES, level,
                   leaf_of_index(index, level), use_rhs ? -1 : upper_level);
}

int ndlqr_SolverUpdateShurFactor(NdLqrSolver* solver, int index, int i, int level, int upper_level,
                                 bool calc_lambda, int use_rhs) {
  if (!solver) return -1;
  if (!solver->device) {
    fprintf(stderr, "ERROR: stage call without a CUDA device (no CPU fallback).\n");
    return -1;
  }
  rslqr_cuda_solver* dev = (rslqr_cuda_solver*)solver->device;
  int rc = push_rhs(solver);
  if (!rc) rc = push_fact(solver);
  if (!rc) rc = rslqr_cuda_schur_task(dev, index, i, level, upper_level, calc_lambda ? 1 : 0, use_rhs, 0);
  if (!rc) rc = pull_results(solver, 1);
  if (rc) fprintf(stderr, "ERROR: GPU Schur update failed: %s\n", rslqr_cuda_last_error());
  return rc ? -1 : 0;
}

/* reference signatures: the NdData pointers identify the solver and whether the rhs is meant */
int ndlqr_FactorInnerProduct(NdData* data, NdData* fact, int index, int data_level, int fact_level) {
  int which = -1;
  NdLqrSolver* s = rslqr_owner_of(fact, &which);
  (void)data;
  if (!s || which == 0) return -1;
  return ndlqr_SolverFactorInnerProduct(s, index, data_level, fact_level, which == 2);
}

int ndlqr_SolveCholeskyFactor(NdData* fact, CholeskyInfo* cholinfo, int index, int level, int upper_level) {
  (void)cholinfo;
  if (!fact) return -1;
  int which = -1;
  NdLqrSolver* s = rslqr_owner_of(fact, &which);
  if (!s || which != 1) return -1;
  return ndlqr_SolverSolveCholeskyFactor(s, index, level, upper_level, 0);
}

int ndlqr_UpdateShurFactor(NdData* fact, NdData* soln, int index, int i, int level, int upper_level,
                           bool calc_lambda) {
  if (!fact || !soln) return -1;
  int which = -1;
  NdLqrSolver* s = rslqr_owner_of(soln, &which);
  if (!s || which == 0) return -1;
  return ndlqr_SolverUpdateShurFactor(s, index, i, level, upper_level, calc_lambda, which == 2);
}

int ndlqr_ComputeShurCompliment(NdLqrSolver* solver, int index, int level, int upper_level) {
  if (!solver) return -1;
  const BinaryNode* nd = solver->tree.node_list + index;
  const int use_rhs = (upper_level == 0);
  for (int k = nd->left_inds.start; k <= nd->right_inds.stop; ++k) {
    const bool calc = ndlqr_ShouldCalcLambda(&solver->tree, index, k);
    if (ndlqr_SolverUpdateShurFactor(solver, index, k, level, upper_level, calc, use_rhs) != 0) return -1;
  }
  return 0;
}

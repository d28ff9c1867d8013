/* rslqr_batch.h -- additive API (no reference counterpart): many independent LQR problems of one
 * shape solved in a single device pass, plus the synthetic problem generator used by the
 * benchmarks. Mirrors the single-solver flow of the reference
 * (New -> Initialize -> Solve -> CopySolution -> Free; reference src/solver.h:84-96). */
#ifndef RSLQR_BATCH_H
#define RSLQR_BATCH_H
#include <stdint.h>
#include "lqr_problem.h"
#ifdef __cplusplus
extern "C" {
#endif

typedef struct NdLqrBatchSolver NdLqrBatchSolver;

NdLqrBatchSolver* ndlqr_NewBatchSolver(int nstates, int ninputs, int nhorizon, int batch, int device);
int ndlqr_FreeBatchSolver(NdLqrBatchSolver* bs);
/* copy `count` problems (count <= batch) to the device; the caller keeps ownership */
int ndlqr_BatchInitialize(NdLqrBatchSolver* bs, const LQRProblem* const* probs, int count);
/* same from flat knot blocks: knots[count][N][ndlqr_LQRDataBlockSize], x0[count][n] */
int ndlqr_BatchInitializeFlat(NdLqrBatchSolver* bs, const double* knots, const double* x0, int count);
int ndlqr_BatchSolve(NdLqrBatchSolver* bs);              /* 0, or -1 (+stderr) on failure */
int ndlqr_BatchGetNumVars(NdLqrBatchSolver* bs);
/* solution of one problem, nvars doubles; returns nvars or -1 */
int ndlqr_BatchCopySolution(NdLqrBatchSolver* bs, int problem, double* soln);
/* all solutions, [count][N*(2n+m)] doubles (each padded like solver->soln->data); returns 0 / -1 */
int ndlqr_BatchCopySolutions(NdLqrBatchSolver* bs, double* soln, int count);
/* per problem 0 = all Cholesky factorizations succeeded */
int ndlqr_BatchGetStatus(NdLqrBatchSolver* bs, int* flags, int count);
double ndlqr_BatchGetSolveTimeMs(NdLqrBatchSolver* bs);
void* ndlqr_BatchDeviceHandle(NdLqrBatchSolver* bs);      /* rslqr_cuda_solver* */

/* Synthetic problems (SplitMix64 stream seeded by seed*0x100000001B3 + id; quadrotor hover
 * linearisation for n=12,m=4, random near-integrator otherwise). Fills flat knot blocks. */
int rslqr_synth_fill(int nstates, int ninputs, int nhorizon, uint64_t seed, int first_id, int count,
                     double* knots, double* x0);
LQRProblem* rslqr_synth_problem(int nstates, int ninputs, int nhorizon, uint64_t seed, int id);

/* Flat binary container for batches of problems (the reference only has JSON, src/json_utils.h:24-61,
 * which is impractical for 4096 x 256 knots). A 64-byte header (magic "RSLQRB01", n, m, N, count, byte-order tag),
 * then per problem x0[n] followed by N flat knot blocks in the LQRData order Q,R,q,r,c,A,B,d
 * (src/lqr_data.c:25-36) -- the layout rslqr_cuda_upload / ndlqr_BatchInitializeFlat take, so a range of problems
 * is read straight into the upload buffers. All return 0 / -1 (+stderr); the reader returns NULL on failure. */
int rslqr_WriteProblemsBinary(const char* path, int nstates, int ninputs, int nhorizon, int count, const double* knots,
                              const double* x0);
int rslqr_ReadProblemsBinaryHeader(const char* path, int* nstates, int* ninputs, int* nhorizon, int* count);
/* problems [first, first+count) -> knots[count][N][block], x0[count][n] */
int rslqr_ReadProblemsBinary(const char* path, int first, int count, double* knots, double* x0);
int rslqr_WriteLQRProblemsBinary(const char* path, const LQRProblem* const* probs, int count);
LQRProblem* rslqr_ReadLQRProblemBinary(const char* path, int index);  /* free with ndlqr_FreeLQRProblem */

#ifdef __cplusplus
}
#endif
#endif

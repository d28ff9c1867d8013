/* rslqr_cuda.h -- the C-ABI boundary between the C host layer and the sm_100a kernels.
 *
 * Plain C: opaque handle, ints, doubles and raw pointers only; no C++ or torch types. The
 * reference has no FFI layer (it is one C library, reference src/ndlqr.h); these entry points are
 * what its public functions bind to in this build:
 *
 *   ndlqr_NewNdLqrSolver            (reference src/solver.c:61)   -> rslqr_cuda_create
 *   ndlqr_InitializeWithLQRProblem  (reference src/solver.c:122)  -> rslqr_cuda_upload (+ _pack_rhs)
 *   ndlqr_ResetSolver / ResetNdData (reference src/solver.c:98)   -> rslqr_cuda_reset
 *   ndlqr_Solve                     (reference src/solve.c:38)    -> rslqr_cuda_solve
 *   ndlqr_SolveLeaf ... UpdateShurFactor (reference src/nested_dissection.c:10-171)
 *                                                                 -> rslqr_cuda_stage
 *   ndlqr_GetSolution / CopySolution (reference src/solve.c:192,197) -> rslqr_cuda_download_solution
 *   solver->fact mirror             (reference src/nddata.c:82)   -> rslqr_cuda_download_fact
 *   cholinfo->success               (reference src/linalg.c:84)   -> rslqr_cuda_download_status
 *   NdLqrProfile                    (reference src/solve.c:15-25) -> rslqr_cuda_get_profile
 *
 * A handle owns `batch` independent problems of identical shape (n, m, N); the single-solver
 * API uses batch = 1. Inputs are "flat knot blocks": per problem N blocks of
 * rslqr_cuda_knot_stride(n,m) doubles in the LQRData order  Q(n) R(m) q(n) r(m) c A(n*n) B(n*m)
 * d(n)  (reference src/lqr_data.c:25-36) plus x0(n). Outputs per problem: the solution vector of
 * N*(2n+m) doubles [lambda_k, x_k, u_k]_k (reference src/solve.h:47-53) and optionally the factor
 * array `fact` of log2(N)*N*(2n+m)*n doubles in the reference's NdData order.
 *
 * Every function returns 0 on success and a negative RSLQR_E_* code otherwise;
 * rslqr_cuda_last_error() gives the message. All calls are synchronous with respect to the host
 * unless stated. There is no CPU fallback: without a usable CUDA device create() fails. */
#ifndef RSLQR_CUDA_H
#define RSLQR_CUDA_H
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct rslqr_cuda_solver rslqr_cuda_solver;

enum {
  RSLQR_OK = 0,
  RSLQR_E_ARG = -1,      /* bad argument (NULL, non power-of-two horizon, range) */
  RSLQR_E_NODEVICE = -2, /* no CUDA device / driver */
  RSLQR_E_CUDA = -3,     /* a CUDA call failed */
  RSLQR_E_NOMEM = -4,    /* device or pinned allocation failed */
  RSLQR_E_UNSUPPORTED = -5
};

/* create() flags */
enum {
  RSLQR_KEEP_FACT = 1u << 0, /* allocate the dense `fact` array at creation; without it the array is
                                allocated (and zeroed) on first need: staged mode, a stage call,
                                factor-once, or rslqr_cuda_ensure_fact */
  RSLQR_PROFILE = 1u << 1    /* record CUDA events around each phase in staged mode */
};

/* solve modes */
enum {
  RSLQR_MODE_AUTO = 0,   /* scan kernels when the shape is covered, else tiered, else staged */
  RSLQR_MODE_STAGED = 1, /* one launch per phase per level (leaf, products, Cholesky, solves, Schur, rhs) */
  RSLQR_MODE_FUSED = 2,  /* tiered: a CTA keeps a group of knots in shared memory for the first levels */
  RSLQR_MODE_SCAN = 3    /* every level on boundary entries, one warp per subtree (kernels_scan.cuh) */
};

/* stages for rslqr_cuda_stage(); `level`, `item`, `col` select one task, -1 = all tasks of the phase */
enum {
  RSLQR_STAGE_PACK_RHS = 0,    /* build the negated rhs from the flat problem (solver.c:138-190) */
  RSLQR_STAGE_LEAVES = 1,      /* ndlqr_SolveLeaf, item = knot                                   */
  RSLQR_STAGE_PRODUCTS = 2,    /* ndlqr_FactorInnerProduct, item = leaf, col = fact level        */
  RSLQR_STAGE_CHOLESKY = 3,    /* separator Cholesky, item = leaf                                */
  RSLQR_STAGE_SOLVES = 4,      /* ndlqr_SolveCholeskyFactor, item = leaf, col = upper level      */
  RSLQR_STAGE_SCHUR = 5,       /* ndlqr_UpdateShurFactor, item = knot, col = upper level         */
  RSLQR_STAGE_RHS_PRODUCTS = 6,
  RSLQR_STAGE_RHS_SOLVES = 7,
  RSLQR_STAGE_RHS_SCHUR = 8,
  RSLQR_STAGE_LEAF_RHS = 9     /* the rhs half of ndlqr_SolveLeaf alone (stored factorization, new rhs) */
};

int rslqr_cuda_device_count(void);
const char* rslqr_cuda_last_error(void);
int rslqr_cuda_knot_stride(int nstates, int ninputs);

int rslqr_cuda_create(int nstates, int ninputs, int nhorizon, int batch, int device,
                      unsigned flags, rslqr_cuda_solver** out);
int rslqr_cuda_destroy(rslqr_cuda_solver* h);

/* host -> device: problems [first, first+count). `knots` = count*N*stride doubles, `x0` = count*n. */
int rslqr_cuda_upload(rslqr_cuda_solver* h, const double* knots, const double* x0, int first,
                      int count);
/* host -> device: overwrite the rhs / solution vector of problems [first, first+count) */
int rslqr_cuda_upload_rhs(rslqr_cuda_solver* h, const double* soln, int first, int count);

int rslqr_cuda_reset(rslqr_cuda_solver* h);            /* zero fact, soln and status           */
/* The dense factor array [batch][K][N][fs] in the reference's NdData order (src/nddata.c:93):
 *   ensure_fact        allocates + zeroes it if it does not exist yet (no-op otherwise);
 *   zero_fact          zeroes it (ndlqr_ResetNdData(solver->fact), reference src/nddata.c:71) and drops the factorization;
 *   has_factorization  1 when it holds a finished staged factorization of the uploaded problems. */
int rslqr_cuda_ensure_fact(rslqr_cuda_solver* h);
int rslqr_cuda_zero_fact(rslqr_cuda_solver* h);
int rslqr_cuda_has_factorization(rslqr_cuda_solver* h);
/* forget the stored factorization without touching memory (the uploaded Q, R, A, B changed) */
int rslqr_cuda_invalidate_factorization(rslqr_cuda_solver* h);
/* RSLQR_MODE_STAGED allocates the dense factor array on first use (rslqr_cuda_ensure_fact). */
int rslqr_cuda_set_mode(rslqr_cuda_solver* h, int mode);
int rslqr_cuda_get_mode(rslqr_cuda_solver* h);         /* the mode solve() will actually use   */

/* Whole solve for all problems of the handle: pack rhs (if `pack_rhs`), leaves, K factor levels,
 * K rhs levels. Blocks until the result is on the device. Every solve starts from cleared status words and
 * flags, and (staged mode) from a zeroed factor array, so a handle can be re-solved without rslqr_cuda_reset. */
int rslqr_cuda_solve(rslqr_cuda_solver* h, int pack_rhs);
/* Same, but only enqueues on the handle's stream (for overlap with copies); pair with _sync. */
int rslqr_cuda_solve_async(rslqr_cuda_solver* h, int pack_rhs);

/* Factor once, solve many (the reference cannot separate the two halves of ndlqr_Solve,
 * docs/rslqr_usage.dox:53-54). Staged mode only (it is the mode that keeps the reference's `fact` array):
 *   rslqr_cuda_factor_async    zeroes `fact`, then leaves + the K factor levels (src/solve.c:57-133).
 *   rslqr_cuda_solve_rhs_async applies the stored factorization to a new right-hand side: pack it from the
 *                              flat problems (`pack_rhs`, e.g. after re-uploading q, r, d, x0) or take the one
 *                              uploaded with rslqr_cuda_upload_rhs; then the rhs half of the leaf stage and the
 *                              K rhs levels (src/solve.c:141-181).
 * A staged rslqr_cuda_solve also leaves a valid factorization behind. As with getrf / getrs the caller keeps the
 * two consistent: re-uploaded knot blocks may change q, r, d, x0 but not Q, R, A, B. rslqr_cuda_reset drops the
 * factorization; solve_rhs then fails with RSLQR_E_UNSUPPORTED. Both calls only enqueue. */
int rslqr_cuda_factor_async(rslqr_cuda_solver* h);
/* Comparison solver: the Riccati recursion of the reference's own baseline, ndlqr_BackwardPass /
 * ndlqr_ForwardPass (src/riccati_solve.c:27-150), one warp per problem; writes the same [lambda_k, x_k, u_k] vector
 * the nested-dissection solve produces (rhs taken from the uploaded problems). n=12, m=4 only; needs batch*N*1136
 * bytes of extra device memory, allocated on the first call. Enqueues only. */
int rslqr_cuda_riccati_solve_async(rslqr_cuda_solver* h);
int rslqr_cuda_solve_rhs_async(rslqr_cuda_solver* h, int pack_rhs);
int rslqr_cuda_sync(rslqr_cuda_solver* h);
/* End to end from HOST buffers (pinned for full PCIe rate): upload `count` problems, solve them (rhs packed from
 * the problem data, as ndlqr_InitializeWithLQRProblem + ndlqr_Solve do, reference src/solver.c:122, src/solve.c:38)
 * and copy the solutions back, in chunks of `chunk` problems (<= 0: automatic) pipelined over three streams so
 * that copies and kernels overlap. Asynchronous: pair with rslqr_cuda_sync. */
int rslqr_cuda_solve_host(rslqr_cuda_solver* h, const double* knots, const double* x0, double* soln, int count,
                          int chunk);

/* One phase (or one task of it) of the staged path, on the dense factor array (created on first use). `problem` -1 = all. */
int rslqr_cuda_stage(rslqr_cuda_solver* h, int stage, int level, int item, int col, int problem);
/* ndlqr_UpdateShurFactor with an explicit calc_lambda (the reference lets the caller choose). */
int rslqr_cuda_schur_task(rslqr_cuda_solver* h, int index, int knot, int level, int upper_level,
                          int calc_lambda, int use_rhs, int problem);

/* device -> host */
int rslqr_cuda_download_solution(rslqr_cuda_solver* h, double* soln, int first, int count);
int rslqr_cuda_download_fact(rslqr_cuda_solver* h, double* fact, int problem);
/* host -> device: overwrite one problem's factor array (enqueues; the stage calls of the host API push the host mirror
 * first, so that helper calls made on views of solver->fact between two stage calls are seen by the next one) */
int rslqr_cuda_upload_fact(rslqr_cuda_solver* h, const double* fact, int problem);
/* per problem: 3N-1 status words in NdLqrCholeskyFactors slot order, 0 = ok, -1 = pivot <= 0 */
int rslqr_cuda_download_status(rslqr_cuda_solver* h, int* status, int problem);
/* per problem summary: 0 = every factorization succeeded */
int rslqr_cuda_download_flags(rslqr_cuda_solver* h, int* flags, int first, int count);
/* sqrt(diag Q_k), sqrt(diag R_k): the leaf Cholesky factors, [N][n+m] per problem */
int rslqr_cuda_download_leaf_factors(rslqr_cuda_solver* h, double* out, int problem);

/* ms: [0] total [1] leaves [2] products [3] cholesky [4] solves [5] schur [6] rhs; staged+PROFILE */
int rslqr_cuda_get_profile(rslqr_cuda_solver* h, double ms[7]);

/* raw device pointers for callers that keep data resident (benchmarks, other CUDA code).
 * Layouts as for upload / download. `fact` is NULL until the factor array exists. */
int rslqr_cuda_device_pointers(rslqr_cuda_solver* h, double** knots, double** x0, double** soln,
                               double** fact);
void* rslqr_cuda_stream(rslqr_cuda_solver* h);          /* cudaStream_t the handle launches on */
int rslqr_cuda_kernel_launches(rslqr_cuda_solver* h);   /* kernels launched by the last solve  */
/* average device time (ms) of the dominant kernel over the launches since the last call with
 * reset != 0; measured with CUDA events on the handle's stream when enabled by enable != 0 */
int rslqr_cuda_kernel_timing(rslqr_cuda_solver* h, int enable, int reset, double* avg_ms, int* launches);

/* CUDA-event stopwatch on the handle's stream: mark(slot) records event `slot` (0..7) after the work
 * enqueued so far; elapsed_ms(a,b) synchronises on b and returns the device time between them (<0 on error) */
int rslqr_cuda_mark(rslqr_cuda_solver* h, int slot);
double rslqr_cuda_elapsed_ms(rslqr_cuda_solver* h, int slot_a, int slot_b);
/* asynchronous copies on the handle's stream (host buffers should be pinned); pair with _sync */
int rslqr_cuda_upload_async(rslqr_cuda_solver* h, const double* knots, const double* x0, int first, int count);
int rslqr_cuda_download_solution_async(rslqr_cuda_solver* h, double* soln, int first, int count);
int rslqr_cuda_upload_rhs_async(rslqr_cuda_solver* h, const double* soln, int first, int count);
int rslqr_cuda_download_status_async(rslqr_cuda_solver* h, int* status, int problem);

/* ---- one long horizon split over GPUs (BASELINE configs[3]): shard r owns knots [first_knot, first_knot+shard_knots)
 * of a horizon of `nhorizon` knots. Upload as usual (knots = the shard's knot blocks, x0 = x0 for the first shard, else
 * d of the knot before the shard). Solve = shard_reduce (leaves + all local levels -> one boundary entry + [A|B] of the
 * last knot in d_payload) ; all-gather of the payloads over NVLink (NCCL by the caller, or rslqr_cuda_shard_exchange
 * within one process) ; shard_finish (the top log2(nranks) separator levels on the gathered entries, redundantly on
 * every rank, then the local top-down pass and rhs propagation). All asynchronous on the handle's stream.
 * Replaces the top levels of the reference's loop, src/solve.c:68-182, which has no distributed form. */
int rslqr_cuda_create_shard(int nstates, int ninputs, int nhorizon, int first_knot, int shard_knots, int device,
                            unsigned flags, rslqr_cuda_solver** out);
size_t rslqr_cuda_shard_payload_doubles(rslqr_cuda_solver* h);
int rslqr_cuda_shard_reduce(rslqr_cuda_solver* h, double* d_payload);
int rslqr_cuda_shard_finish(rslqr_cuda_solver* h, const double* d_gathered, int rank, int nranks);
int rslqr_cuda_shard_exchange(rslqr_cuda_solver** hs, int nranks, double** d_payload, double** d_gathered);
/* Exchange without a collective: shard_push stores the payload into every rank's gathered buffer over NVLink peer
 * mappings (peer_gathered[r], peer_flags[r]: rank r's buffers as mapped into this process) and raises flag `rank` there
 * to `epoch`; shard_finish_wait = shard_finish whose cross-GPU tier kernel first waits (bounded) until the local flags of
 * all ranks have reached `epoch`. Epochs must increase by one per solve and the gathered buffer alternate between two
 * halves, so that a rank one solve ahead never overwrites data still being read. Scan kernels only. */
int rslqr_cuda_shard_push(rslqr_cuda_solver* h, const double* d_payload, double* const* peer_gathered,
                          unsigned long long* const* peer_flags, int rank, int nranks, unsigned long long epoch);
int rslqr_cuda_shard_finish_wait(rslqr_cuda_solver* h, const double* d_gathered, const unsigned long long* d_flags,
                                 unsigned long long epoch, int rank, int nranks);
/* enqueue on a caller-owned cudaStream_t (e.g. the stream NCCL collectives are ordered on) */
int rslqr_cuda_set_stream(rslqr_cuda_solver* h, void* stream);

/* ---- dense helper primitives on the device (kernels_linalg.cuh): what the reference-named helper API of
 * include/linalg.h / include/linalg_custom.h (MatrixMultiply, MatrixCholeskyFactorize, clap_* ...; reference
 * src/linalg.h:103-191, src/linalg_custom.h:44-159) binds to. HOST pointers in and out, column-major, leading dimension =
 * rows, synchronous, on the current CUDA device; not on the solve path. rslqr_cuda_la_last_error() has the message. */
const char* rslqr_cuda_la_last_error(void);
/* C = beta C + alpha op(A) op(B); symmetric != 0: A symmetric, lower triangle read (tA, tB ignored) */
int rslqr_cuda_la_gemm(const double* A, int arows, int acols, const double* B, int brows, int bcols, double* C, int crows,
                       int ccols, int tA, int tB, double alpha, double beta, int symmetric);
int rslqr_cuda_la_axpy(const double* A, double* B, long long count, double alpha); /* B += alpha A; A == NULL: B *= alpha */
int rslqr_cuda_la_add_diagonal(double* A, int n, double alpha);
int rslqr_cuda_la_cholesky(double* A, int n, int* info);          /* in-place lower Cholesky; *info = 0 ok, -1 pivot <= 0 */
int rslqr_cuda_la_trisolve(const double* L, int n, double* b, int m, int which); /* 0: L^-1 b, 1: L^-T b, 2: both */

/* pinned host memory helpers so that callers can stage uploads at full PCIe rate */
/* Solve without any allocation. No counterpart in the reference: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same kernels and launch order as rslqr_cuda_solve_async in RSLQR_MODE_SCAN with the rhs packed from the problem
 * data, but without a handle: the caller passes DEVICE buffers on `device` (16-byte aligned)
 *     d_knots [batch][N][stride]  flat knot blocks (rslqr_cuda_knot_stride)     d_x0   [batch][n]
 *     d_soln  [batch][N][2n+m]    out, [lambda_k, x_k, u_k]                      d_flags [batch] out, 0 = solved
 * and everything the kernels keep between launches lives in one static device array of RSLQR_STATIC_SCRATCH_BYTES
 * (32 MiB; every GPU that loads the library holds it until exit). The call enqueues on `stream` (NULL = the legacy
 * default stream) and calls no cudaMalloc / cudaFree or any other allocating routine. Shapes: those of the scan
 * path (33 (n, m) pairs with n, m <= 16: the list RSLQR_DUAL_SHAPES in rslqr_b200/csrc/kernels_dual.cuh), power-of-two N >= 8; about 0.98 MB of scratch per problem at
 * n=12, m=4, N=256 (33 problems fit). A size beyond the array is rejected with RSLQR_E_ARG and a message that
 * states the bytes needed. ALL calls on a device share the array: calls that overlap on the GPU (different streams,
 * or graphs replayed concurrently) race; serialise them. */
#define RSLQR_STATIC_SCRATCH_BYTES 33554432u /* 32 MiB */
int rslqr_cuda_solve_static(int n, int m, int nhorizon, int batch, int device, const double* d_knots,
                            const double* d_x0, double* d_soln, int* d_flags, void* stream);

void* rslqr_cuda_host_alloc(size_t bytes);
void rslqr_cuda_host_free(void* p);
/* page-lock (and later release) memory the caller already owns, e.g. the malloc'ed mirrors behind solver->fact /
 * solver->soln of the host API, so that their copies run at the pinned rate and truly asynchronously. 0, or
 * RSLQR_E_CUDA when the range cannot be locked (the copies then simply stay pageable). */
int rslqr_cuda_host_register(void* p, size_t bytes);
int rslqr_cuda_host_unregister(void* p);

#ifdef __cplusplus
}
#endif
#endif

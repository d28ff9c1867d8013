// rslqr_cuda.cu -- implementation of the C-ABI in include/rslqr_cuda.h: device memory, copies,
// kernel launches. No arithmetic happens on the host in this file.
#include "rslqr_cuda.h"

#include <cuda_runtime.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "kernels_fused.cuh"
#include "kernels_scan.cuh"
#include "kernels_dual.cuh"
#include "kernels_big.cuh"
#include "kernels_riccati.cuh"
#include "kernels_staged.cuh"
#include "kernels_staged_mma.cuh"

using namespace rslqr;

static thread_local std::string g_err;
static int fail(int code, const std::string& msg) {
  g_err = msg;
  return code;
}
#define CU(call)                                                                         \
  do {                                                                                   \
    cudaError_t e_ = (call);                                                             \
    if (e_ != cudaSuccess)                                                               \
      return fail(e_ == cudaErrorMemoryAllocation ? RSLQR_E_NOMEM : RSLQR_E_CUDA,        \
                  std::string(#call) + ": " + cudaGetErrorString(e_));                   \
  } while (0)

// create() flag used by rslqr_cuda_create_shard only: plan the scan path on boundary entries at every level (the
// shard exchange and the cross-GPU tier work on them)
static const unsigned RSLQR_INTERNAL_SHARD = 1u << 31;

struct PhaseSpan { int phase; cudaEvent_t t0, t1; };

struct rslqr_cuda_solver {
  Shape s;
  Bufs b;
  int batch, device;
  unsigned flags;
  int mode;            // requested
  cudaStream_t stream;
  double* d_knots; double* d_x0;
  int launches;
  // staged-mode profile
  std::vector<PhaseSpan> spans;
  std::vector<cudaEvent_t> evpool; size_t evused;
  double prof[7];
  // dominant-kernel timing
  bool ktime_on; std::vector<std::pair<cudaEvent_t, cudaEvent_t>> kspans; double k_ms; int k_n;
  FusedPlan fused; int force_dfma;
  ScanPlan scan; int g0_pref;
  bool pdl;   // latency mode (batch < 64): kernels are chained by programmatic dependent launch (kernels_staged.cuh)
  BigPlan big{};  // dual-form path for large blocks (n <= 64, m <= 16 without an instance of the warp kernels): buffers allocated on the first solve
  int factored = 0;  // `fact` holds a finished staged factorization of the uploaded problems
  bool prof_keep = false;  // keep the profile spans across enqueues (the chunks of solve_host)
  double* ric_gain = nullptr;  // [batch][N][K | dv | P | p] of the Riccati comparison solver (allocated on first use)
  bool shard; bool own_stream;
  int r_first, r_count;            // problem range the next enqueue covers (whole batch by default)
  cudaStream_t s_h2d, s_d2h; bool pipe_made; std::vector<cudaEvent_t> ev_up, ev_done; cudaEvent_t ev_join;
  cudaEvent_t marks[8]; bool marks_made;
};

extern "C" const char* rslqr_cuda_last_error(void) { return g_err.c_str(); }

extern "C" int rslqr_cuda_device_count(void) {
  int c = 0;
  if (cudaGetDeviceCount(&c) != cudaSuccess) { cudaGetLastError(); return 0; }
  return c;
}

extern "C" int rslqr_cuda_knot_stride(int n, int m) { return 2 * n + 2 * m + 1 + n * n + n * m + n; }

static Shape make_shape(int n, int m, int N) {
  Shape s;
  s.n = n; s.m = m; s.N = N; s.K = rs_log2(N);
  s.fs = (2 * n + m) * n; s.zs = 2 * n + m;
  s.stride = rslqr_cuda_knot_stride(n, m);
  s.oQ = 0; s.oR = n; s.oq = n + m; s.orr = 2 * n + m; s.oA = 2 * n + 2 * m + 1;
  s.oB = s.oA + n * n; s.od = s.oB + n * m;
  s.kofs = 0; s.Ng = N;
  return s;
}

extern "C" int rslqr_cuda_create(int n, int m, int N, int batch, int device, unsigned flags,
                                 rslqr_cuda_solver** out) {
  if (!out) return fail(RSLQR_E_ARG, "out is NULL");
  *out = nullptr;
  if (n <= 0 || m <= 0 || N < 2 || (N & (N - 1)) || batch <= 0)
    return fail(RSLQR_E_ARG, "need n,m,batch > 0 and a power-of-two horizon >= 2");
  int ndev = rslqr_cuda_device_count();
  if (ndev <= 0) return fail(RSLQR_E_NODEVICE, "no CUDA device available (there is no CPU fallback)");
  if (device < 0 || device >= ndev) return fail(RSLQR_E_ARG, "device index out of range");
  CU(cudaSetDevice(device));
  rslqr_cuda_solver* h = new rslqr_cuda_solver();
  h->s = make_shape(n, m, N);
  h->batch = batch; h->device = device; h->flags = flags; h->mode = RSLQR_MODE_AUTO;
  h->shard = false; h->own_stream = true;
  h->marks_made = false; h->pipe_made = false; h->r_first = 0; h->r_count = batch;
  h->launches = 0; h->evused = 0; h->ktime_on = false; h->k_ms = 0; h->k_n = 0;
  memset(h->prof, 0, sizeof(h->prof));
  memset(&h->b, 0, sizeof(h->b));
  h->d_knots = h->d_x0 = nullptr;
  const Shape& s = h->s;
  cudaError_t e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
  if (e != cudaSuccess) { delete h; return fail(RSLQR_E_CUDA, cudaGetErrorString(e)); }
  auto alloc = [&](void** p, size_t bytes) { return cudaMalloc(p, bytes); };
  bool ok = true;
  ok &= alloc((void**)&h->d_knots, sizeof(double) * (size_t)batch * N * s.stride) == cudaSuccess;
  ok &= alloc((void**)&h->d_x0, sizeof(double) * (size_t)batch * n) == cudaSuccess;
  ok &= alloc((void**)&h->b.soln, sizeof(double) * (size_t)batch * N * s.zs) == cudaSuccess;
  // status words and per-problem flags share one allocation ([batch][3N] then [batch]): a full-batch solve clears
  // both with one memset node
  ok &= alloc((void**)&h->b.status, sizeof(int) * ((size_t)batch * 3 * N + (size_t)batch)) == cudaSuccess;
  if (ok) h->b.flags = h->b.status + (size_t)batch * 3 * N;
  ok &= alloc((void**)&h->b.leafd, sizeof(double) * (size_t)batch * N * (n + m)) == cudaSuccess;
  // RSLQR_KEEP_FACT allocates the dense factor array up front; without it the array appears on first need
  // (staged mode, stage calls, factor-once: ensure_fact)
  if (ok && (flags & RSLQR_KEEP_FACT))
    ok &= alloc((void**)&h->b.fact, sizeof(double) * (size_t)batch * s.K * N * s.fs) == cudaSuccess;
  if (!ok) {
    cudaGetLastError();
    rslqr_cuda_destroy(h);
    return fail(RSLQR_E_NOMEM, "device allocation failed");
  }
  h->b.knots = h->d_knots; h->b.x0 = h->d_x0;
  {
    const char* g0 = getenv("RSLQR_G0");
    const char* fd = getenv("RSLQR_FORCE_DFMA");
    h->force_dfma = fd ? atoi(fd) : 0;
    memset(&h->scan, 0, sizeof(h->scan));
    memset(&h->fused, 0, sizeof(h->fused));
    h->big.covered = big_shape_ok(n, m, N) && !(flags & RSLQR_INTERNAL_SHARD);
    // shards take the hybrid form too (their exchange payload is then the shard's dual element); RSLQR_SHARD_CLASSIC=1
    // keeps the boundary-entry shard path of round 1 (A/B runs)
    const char* sc = getenv("RSLQR_SHARD_CLASSIC");
    if (scan_plan(h->s, h->batch, &h->scan, g_err, !((flags & RSLQR_INTERNAL_SHARD) && sc && atoi(sc)))) {
      rslqr_cuda_destroy(h);
      return RSLQR_E_NOMEM;
    }
    // the tiered buffers (dense state of every knot) are allocated up front only for shapes the scan kernels
    // do not cover; otherwise on the first set_mode(RSLQR_MODE_FUSED)
    h->g0_pref = g0 ? atoi(g0) : 8;
    const char* pv = getenv("RSLQR_PDL");   // 0 switches programmatic dependent launch off (A/B runs)
    h->pdl = batch < 64 && !(pv && !atoi(pv));
    if (!h->scan.f.available && fused_plan(h->s, h->batch, h->g0_pref, &h->fused, g_err)) {
      rslqr_cuda_destroy(h);
      return RSLQR_E_NOMEM;
    }
  }
  *out = h;
  int rc = rslqr_cuda_reset(h);
  if (rc) { rslqr_cuda_destroy(h); *out = nullptr; }
  return rc;
}

extern "C" int rslqr_cuda_destroy(rslqr_cuda_solver* h) {
  if (!h) return fail(RSLQR_E_ARG, "handle is NULL");
  cudaSetDevice(h->device);
  cudaStreamSynchronize(h->stream);
  cudaFree(h->d_knots); cudaFree(h->d_x0); cudaFree(h->b.soln); cudaFree(h->b.fact);
  cudaFree(h->b.status); cudaFree(h->b.leafd); cudaFree(h->ric_gain);
  fused_free(&h->fused);
  scan_free(&h->scan);
  big_free(&h->big);
  for (auto ev : h->evpool) cudaEventDestroy(ev);
  for (auto& pr : h->kspans) { cudaEventDestroy(pr.first); cudaEventDestroy(pr.second); }
  if (h->marks_made) for (auto ev : h->marks) cudaEventDestroy(ev);
  if (h->pipe_made) {
    for (auto ev : h->ev_up) cudaEventDestroy(ev);
    for (auto ev : h->ev_done) cudaEventDestroy(ev);
    cudaEventDestroy(h->ev_join); cudaStreamDestroy(h->s_h2d); cudaStreamDestroy(h->s_d2h);
  }
  if (h->own_stream) cudaStreamDestroy(h->stream);
  delete h;
  return RSLQR_OK;
}

static bool range_ok(const rslqr_cuda_solver* h, int first, int count) {
  return first >= 0 && count >= 0 && first + count <= h->batch;
}

extern "C" int rslqr_cuda_upload(rslqr_cuda_solver* h, const double* knots, const double* x0,
                                 int first, int count) {
  if (!h || !knots || !x0 || !range_ok(h, first, count)) return fail(RSLQR_E_ARG, "upload: bad argument");
  CU(cudaSetDevice(h->device));
  const Shape& s = h->s;
  const size_t per = (size_t)s.N * s.stride;
  CU(cudaMemcpyAsync(h->d_knots + first * per, knots, sizeof(double) * per * count, cudaMemcpyHostToDevice, h->stream));
  CU(cudaMemcpyAsync(h->d_x0 + (size_t)first * s.n, x0, sizeof(double) * (size_t)s.n * count, cudaMemcpyHostToDevice, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  return RSLQR_OK;
}

extern "C" int rslqr_cuda_upload_rhs(rslqr_cuda_solver* h, const double* soln, int first, int count) {
  if (!h || !soln || !range_ok(h, first, count)) return fail(RSLQR_E_ARG, "upload_rhs: bad argument");
  CU(cudaSetDevice(h->device));
  const size_t per = (size_t)h->s.N * h->s.zs;
  CU(cudaMemcpyAsync(h->b.soln + first * per, soln, sizeof(double) * per * count, cudaMemcpyHostToDevice, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  return RSLQR_OK;
}

extern "C" int rslqr_cuda_reset(rslqr_cuda_solver* h) {
  if (!h) return fail(RSLQR_E_ARG, "handle is NULL");
  CU(cudaSetDevice(h->device));
  const Shape& s = h->s;
  if (h->b.fact) CU(cudaMemsetAsync(h->b.fact, 0, sizeof(double) * (size_t)h->batch * s.K * s.N * s.fs, h->stream));
  CU(cudaMemsetAsync(h->b.soln, 0, sizeof(double) * (size_t)h->batch * s.N * s.zs, h->stream));
  CU(cudaMemsetAsync(h->b.status, 0, sizeof(int) * ((size_t)h->batch * 3 * s.N + (size_t)h->batch), h->stream));
  CU(cudaStreamSynchronize(h->stream));
  h->factored = 0;
  return RSLQR_OK;
}

// The dense factor array [batch][K][N][fs] of the staged path (reference NdData order, src/nddata.c:93): allocated and
// zeroed on first need unless RSLQR_KEEP_FACT asked for it at creation.
static int ensure_fact(rslqr_cuda_solver* h) {
  if (h->b.fact) return RSLQR_OK;
  const Shape& s = h->s;
  const size_t bytes = sizeof(double) * (size_t)h->batch * s.K * s.N * s.fs;
  if (cudaMalloc((void**)&h->b.fact, bytes) != cudaSuccess) {
    cudaGetLastError();
    h->b.fact = nullptr;
    return fail(RSLQR_E_NOMEM, "device allocation of the dense factor array failed");
  }
  CU(cudaMemsetAsync(h->b.fact, 0, bytes, h->stream));
  h->factored = 0;
  return RSLQR_OK;
}

extern "C" int rslqr_cuda_ensure_fact(rslqr_cuda_solver* h) {
  if (!h) return fail(RSLQR_E_ARG, "handle is NULL");
  if (h->shard) return fail(RSLQR_E_UNSUPPORTED, "a shard keeps no factor array");
  CU(cudaSetDevice(h->device));
  return ensure_fact(h);
}

extern "C" int rslqr_cuda_zero_fact(rslqr_cuda_solver* h) {
  if (!h) return fail(RSLQR_E_ARG, "handle is NULL");
  CU(cudaSetDevice(h->device));
  h->factored = 0;
  if (!h->b.fact) return RSLQR_OK;
  const Shape& s = h->s;
  CU(cudaMemsetAsync(h->b.fact, 0, sizeof(double) * (size_t)h->batch * s.K * s.N * s.fs, h->stream));
  return RSLQR_OK;
}

extern "C" int rslqr_cuda_has_factorization(rslqr_cuda_solver* h) { return h ? h->factored : 0; }
extern "C" int rslqr_cuda_invalidate_factorization(rslqr_cuda_solver* h) {
  if (!h) return fail(RSLQR_E_ARG, "handle is NULL");
  h->factored = 0;
  return RSLQR_OK;
}

// status words and flags of the problem range the next enqueue covers: cleared at the start of EVERY solve, so a
// failure (or an exchange time-out) of one solve is not reported for the next one
static int clear_status(rslqr_cuda_solver* h) {
  const Shape& s = h->s;
  if (h->r_first == 0 && h->r_count == h->batch) {
    CU(cudaMemsetAsync(h->b.status, 0, sizeof(int) * ((size_t)h->batch * 3 * s.N + (size_t)h->batch), h->stream));
  } else {
    CU(cudaMemsetAsync(h->b.status + (size_t)h->r_first * 3 * s.N, 0, sizeof(int) * (size_t)h->r_count * 3 * s.N, h->stream));
    CU(cudaMemsetAsync(h->b.flags + h->r_first, 0, sizeof(int) * (size_t)h->r_count, h->stream));
  }
  return RSLQR_OK;
}

extern "C" int rslqr_cuda_set_mode(rslqr_cuda_solver* h, int mode) {
  if (!h || mode < RSLQR_MODE_AUTO || mode > RSLQR_MODE_SCAN) return fail(RSLQR_E_ARG, "set_mode: bad argument");
  if (mode == RSLQR_MODE_STAGED) {
    if (h->shard) return fail(RSLQR_E_UNSUPPORTED, "a shard has no staged mode");
    CU(cudaSetDevice(h->device));
    int rc = ensure_fact(h);
    if (rc) return rc;
  }
  if (mode == RSLQR_MODE_FUSED && !h->fused.available) {
    CU(cudaSetDevice(h->device));
    if (fused_plan(h->s, h->batch, h->g0_pref, &h->fused, g_err)) return RSLQR_E_NOMEM;
    if (!h->fused.available) return fail(RSLQR_E_UNSUPPORTED, "shape is not covered by the tiered (fused) kernels");
  }
  if (mode == RSLQR_MODE_SCAN && !h->scan.f.available && !h->big.covered)
    return fail(RSLQR_E_UNSUPPORTED, "shape is not covered by the scan kernels");
  h->mode = mode;
  return RSLQR_OK;
}

extern "C" int rslqr_cuda_get_mode(rslqr_cuda_solver* h) {
  if (!h) return fail(RSLQR_E_ARG, "handle is NULL");
  if (h->mode != RSLQR_MODE_AUTO) return h->mode;
  if (h->scan.f.available || h->big.covered) return RSLQR_MODE_SCAN;
  if (h->fused.available) return RSLQR_MODE_FUSED;
  return RSLQR_MODE_STAGED;
}

// ------------------------------------------------------------------ staged launches
static inline int round32(int x) { return ((x + 31) / 32) * 32; }
static inline int clampi(int x, int lo, int hi) { return x < lo ? lo : (x > hi ? hi : x); }

static int launch_stage(rslqr_cuda_solver* h, int stage, int level, Sel sel) {
  const Shape& s = h->s;
  const int n = s.n, nn = n * n;
  if (sel.np <= 0 || sel.nitems <= 0) return RSLQR_OK;
  const long long tasks = (long long)sel.np * sel.nitems;
  // grid.x of every stage kernel is a small multiple of `tasks` (times at most zs / 128 * 3): keep it inside 2^31 - 1
  if (tasks * 3 > 0x7fffffffLL || tasks * (long long)s.zs * s.N / 128 > 0x7fffffffLL * 64)
    return fail(RSLQR_E_ARG, "stage launch too large for one grid: split the batch");
  cudaStream_t st = h->stream;
  switch (stage) {
    case RSLQR_STAGE_PACK_RHS: {
      sel.item0 = 0; sel.nitems = s.N;
      const long long tot = (long long)sel.np * s.N * s.zs;
      pack_rhs_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(s, h->b, sel);
      break;
    }
    case RSLQR_STAGE_LEAVES:
      leaf_kernel<<<(unsigned)tasks, clampi(round32(nn), 32, 256), 0, st>>>(s, h->b, sel);
      break;
    case RSLQR_STAGE_PRODUCTS:
      if (n == 64 && s.m == 16 && !getenv("RSLQR_NO_MMA_STAGED")) {
        constexpr int smem = 64 * (80 + 4) * (int)sizeof(double);
        products_mma_kernel<64, 16><<<(unsigned)(tasks * (sel.col < 0 ? 3 : 1)), 128, smem, st>>>(s, h->b, sel, level);
      } else if (n == 16 && s.m == 16 && !getenv("RSLQR_NO_MMA_STAGED")) {
        products_mma_kernel<16, 16><<<(unsigned)(tasks * (sel.col < 0 ? 3 : 1)), 32, 16 * 36 * sizeof(double), st>>>(s, h->b, sel, level);
      } else
      products_kernel<<<(unsigned)(tasks * (sel.col < 0 ? 3 : 1)), clampi(round32(nn), 32, 256), 0, st>>>(s, h->b, sel, level);
      break;
    case RSLQR_STAGE_CHOLESKY:
      cholesky_kernel<<<(unsigned)tasks, clampi(round32(n), 32, 256), sizeof(double) * nn, st>>>(s, h->b, sel, level);
      break;
    case RSLQR_STAGE_SOLVES:
      if (n == 64 && s.m == 16 && !getenv("RSLQR_NO_MMA_STAGED")) {
        constexpr int smem = (64 * 66 + 64 * 65 + 64) * (int)sizeof(double);
        CU(smem_optin(solves_cols_kernel<64>, smem));
        solves_cols_kernel<64><<<(unsigned)(tasks * (sel.col < 0 ? 2 : 1)), 64, smem, st>>>(s, h->b, sel, level);
      } else if (n == 16 && s.m == 16 && !getenv("RSLQR_NO_MMA_STAGED")) {
        solves_cols_kernel<16><<<(unsigned)(tasks * (sel.col < 0 ? 2 : 1)), 16, (16 * 18 + 16 * 17 + 16) * sizeof(double), st>>>(s, h->b, sel, level);
      } else
      solves_kernel<<<(unsigned)(tasks * (sel.col < 0 ? 2 : 1)), clampi(round32(nn), 32, 256), sizeof(double) * 2 * nn, st>>>(s, h->b, sel, level);
      break;
    case RSLQR_STAGE_SCHUR:
      if (n == 64 && s.m == 16 && !getenv("RSLQR_NO_MMA_STAGED"))
        schur_mma_kernel<64, 16><<<(unsigned)(tasks * (sel.col < 0 ? 2 : 1)), 32 * 9, sizeof(double) * 68 * 64, st>>>(s, h->b, sel, level);
      else if (n == 16 && s.m == 16 && !getenv("RSLQR_NO_MMA_STAGED"))
        schur_mma_kernel<16, 16><<<(unsigned)(tasks * (sel.col < 0 ? 2 : 1)), 32 * 3, sizeof(double) * 20 * 16, st>>>(s, h->b, sel, level);
      else
      schur_kernel<<<(unsigned)(tasks * (sel.col < 0 ? 2 : 1)), clampi(round32(s.fs), 32, 256), sizeof(double) * nn, st>>>(s, h->b, sel, level);
      break;
    case RSLQR_STAGE_RHS_PRODUCTS:
      rhs_products_kernel<<<(unsigned)((tasks * n + 127) / 128), 128, 0, st>>>(s, h->b, sel, level);
      break;
    case RSLQR_STAGE_RHS_SOLVES:
      if (n <= 64 && !getenv("RSLQR_NO_WARP_RHS_SOLVE")) {  // one warp per separator, no CTA barrier
        const int wpc = n > 32 ? 2 : 4;
        const size_t smem = sizeof(double) * (size_t)wpc * n * (n + 1);
        CU(smem_optin(rhs_solves_warp_kernel, (int)smem));
        rhs_solves_warp_kernel<<<(unsigned)((tasks + wpc - 1) / wpc), 32 * wpc, smem, st>>>(s, h->b, sel, level);
      } else
      rhs_solves_kernel<<<(unsigned)tasks, clampi(round32(n), 32, 256), sizeof(double) * (nn + n), st>>>(s, h->b, sel, level);
      break;
    case RSLQR_STAGE_RHS_SCHUR:
      rhs_schur_kernel<<<(unsigned)((tasks * s.zs + 127) / 128), 128, 0, st>>>(s, h->b, sel, level);
      break;
    case RSLQR_STAGE_LEAF_RHS:
      leaf_rhs_kernel<<<(unsigned)((tasks * (n + s.m) + 127) / 128), 128, 0, st>>>(s, h->b, sel);
      break;
    default:
      return fail(RSLQR_E_ARG, "unknown stage");
  }
  CU(cudaGetLastError());
  ++h->launches;
  return RSLQR_OK;
}

static cudaEvent_t take_event(rslqr_cuda_solver* h) {
  if (h->evused == h->evpool.size()) {
    cudaEvent_t ev;
    cudaEventCreate(&ev);
    h->evpool.push_back(ev);
  }
  return h->evpool[h->evused++];
}

static int timed_stage(rslqr_cuda_solver* h, int phase, int stage, int level, const Sel& sel) {
  const bool prof = (h->flags & RSLQR_PROFILE) != 0;
  PhaseSpan sp;
  if (prof) {
    // a stage starts where the previous one of the same enqueue ended: one event record per stage, not two
    sp.phase = phase;
    if (h->spans.empty()) { sp.t0 = take_event(h); cudaEventRecord(sp.t0, h->stream); }
    else sp.t0 = h->spans.back().t1;
    sp.t1 = take_event(h);
  }
  int rc = launch_stage(h, stage, level, sel);
  if (prof) { cudaEventRecord(sp.t1, h->stream); h->spans.push_back(sp); }
  return rc;
}

static Sel all_items(const rslqr_cuda_solver* h, int nitems) {
  Sel sel;
  sel.p0 = h->r_first; sel.np = h->r_count; sel.item0 = 0; sel.nitems = nitems; sel.col = -1;
  sel.calc_lambda = -1; sel.index = -1;
  return sel;
}

// the two halves of the staged solve: factorization (reference src/solve.c:57-133) and rhs phase (:141-181)
static int staged_factor_levels(rslqr_cuda_solver* h) {
  const Shape& s = h->s;
  int rc = 0;
  CU(smem_optin(solves_kernel, (int)(s.n * s.n * 2 * sizeof(double))));
  for (int level = 0; level < s.K && !rc; ++level) {
    const int leaves = 1 << (s.K - level - 1);
    rc |= timed_stage(h, 2, RSLQR_STAGE_PRODUCTS, level, all_items(h, leaves));
    rc |= timed_stage(h, 3, RSLQR_STAGE_CHOLESKY, level, all_items(h, leaves));
    if (level < s.K - 1) {
      rc |= timed_stage(h, 4, RSLQR_STAGE_SOLVES, level, all_items(h, leaves));
      rc |= timed_stage(h, 5, RSLQR_STAGE_SCHUR, level, all_items(h, s.N));
    }
  }
  return rc;
}

static int staged_rhs_levels(rslqr_cuda_solver* h) {
  const Shape& s = h->s;
  int rc = 0;
  for (int level = 0; level < s.K && !rc; ++level) {
    const int leaves = 1 << (s.K - level - 1);
    rc |= timed_stage(h, 6, RSLQR_STAGE_RHS_PRODUCTS, level, all_items(h, leaves));
    rc |= timed_stage(h, 6, RSLQR_STAGE_RHS_SOLVES, level, all_items(h, leaves));
    rc |= timed_stage(h, 6, RSLQR_STAGE_RHS_SCHUR, level, all_items(h, s.N));
  }
  return rc;
}

static int solve_staged(rslqr_cuda_solver* h, int pack_rhs) {
  const Shape& s = h->s;
  int rc = ensure_fact(h);
  if (rc) return rc;
  // the staged kernels ACCUMULATE into the factor array (products: -F + ..., Schur: F - ...), exactly like the
  // reference, whose contract is a zeroed `fact` before every solve (src/solve.h:28-32); the array is cleared here
  // so that a second Initialize + Solve on the same handle cannot factorize on top of the previous factors
  h->factored = 0;
  CU(cudaMemsetAsync(h->b.fact + (size_t)h->r_first * s.K * s.N * s.fs, 0, sizeof(double) * (size_t)h->r_count * s.K * s.N * s.fs,
                     h->stream));
  if (pack_rhs) rc |= timed_stage(h, 6, RSLQR_STAGE_PACK_RHS, 0, all_items(h, s.N));
  rc |= timed_stage(h, 1, RSLQR_STAGE_LEAVES, 0, all_items(h, s.N));
  if (!rc) rc = staged_factor_levels(h);
  if (!rc) rc = staged_rhs_levels(h);
  if (!rc && h->r_first == 0 && h->r_count == h->batch) h->factored = 1;
  return rc;
}

static int staged_entry(rslqr_cuda_solver* h, const char* what) {
  if (!h) return fail(RSLQR_E_ARG, "handle is NULL");
  if (h->shard) return fail(RSLQR_E_UNSUPPORTED, "a shard has no stored factorization");
  if (rslqr_cuda_get_mode(h) != RSLQR_MODE_STAGED) return fail(RSLQR_E_UNSUPPORTED, what);
  CU(cudaSetDevice(h->device));
  int rc = ensure_fact(h);
  if (rc) return rc;
  h->launches = 0;
  if (!h->prof_keep) { h->spans.clear(); h->evused = 0; }
  return RSLQR_OK;
}

extern "C" int rslqr_cuda_factor_async(rslqr_cuda_solver* h) {
  int rc = staged_entry(h, "factor: only the staged mode keeps a factorization (rslqr_cuda_set_mode(RSLQR_MODE_STAGED))");
  if (rc) return rc;
  const Shape& s = h->s;
  h->factored = 0;
  rc = clear_status(h);
  if (rc) return rc;
  CU(cudaMemsetAsync(h->b.fact + (size_t)h->r_first * s.K * s.N * s.fs, 0, sizeof(double) * (size_t)h->r_count * s.K * s.N * s.fs,
                     h->stream));
  rc = timed_stage(h, 1, RSLQR_STAGE_LEAVES, 0, all_items(h, s.N));  // (also scales whatever rhs is in soln: harmless)
  if (!rc) rc = staged_factor_levels(h);
  if (!rc && h->r_first == 0 && h->r_count == h->batch) h->factored = 1;
  return rc;
}

extern "C" int rslqr_cuda_solve_rhs_async(rslqr_cuda_solver* h, int pack_rhs) {
  int rc = staged_entry(h, "solve_rhs: only the staged mode keeps a factorization (rslqr_cuda_set_mode(RSLQR_MODE_STAGED))");
  if (rc) return rc;
  if (!h->factored) return fail(RSLQR_E_UNSUPPORTED, "solve_rhs: no factorization (call rslqr_cuda_factor_async or a staged solve first)");
  const Shape& s = h->s;
  if (pack_rhs) rc |= timed_stage(h, 6, RSLQR_STAGE_PACK_RHS, 0, all_items(h, s.N));
  rc |= timed_stage(h, 6, RSLQR_STAGE_LEAF_RHS, 0, all_items(h, s.N));
  if (!rc) rc = staged_rhs_levels(h);
  return rc;
}

// The Riccati recursion of the reference's own baseline solver (src/riccati_solve.c:27-150) on the device: one warp
// per problem, result in the same [lambda, x, u] layout as the nested-dissection solve. Comparison solver only.
extern "C" int rslqr_cuda_riccati_solve_async(rslqr_cuda_solver* h) {
  if (!h) return fail(RSLQR_E_ARG, "handle is NULL");
  if (h->shard) return fail(RSLQR_E_UNSUPPORTED, "riccati: not for shards");
  const Shape& s = h->s;
  if (!(s.n == 12 && s.m == 4 && s.N >= 2)) return fail(RSLQR_E_UNSUPPORTED, "riccati: the device kernel covers n=12, m=4");
  CU(cudaSetDevice(h->device));
  if (!h->ric_gain)
    CU(cudaMalloc(&h->ric_gain, sizeof(double) * (size_t)h->batch * s.N * RicDim<12, 4>::GSZ));
  h->launches = 0;
  {
    int rc = clear_status(h);
    if (rc) return rc;
  }
  const size_t f = (size_t)h->r_first;
  Bufs b = h->b;
  b.knots += f * s.N * s.stride; b.x0 += f * s.n; b.soln += f * s.N * s.zs; b.flags += f;
  if (launch_riccati<12, 4>(s, b, h->ric_gain + f * s.N * RicDim<12, 4>::GSZ, h->r_count, h->stream, g_err)) return RSLQR_E_CUDA;
  h->launches = 1;
  return RSLQR_OK;
}

extern "C" int rslqr_cuda_solve_async(rslqr_cuda_solver* h, int pack_rhs) {
  if (!h) return fail(RSLQR_E_ARG, "handle is NULL");
  if (h->shard) return fail(RSLQR_E_UNSUPPORTED, "a shard is solved with shard_reduce / exchange / shard_finish");
  CU(cudaSetDevice(h->device));
  h->launches = 0;
  if (!h->prof_keep) {  // (solve_host accumulates the spans of all its chunks)
    h->spans.clear();
    h->evused = 0;
    memset(h->prof, 0, sizeof(h->prof));
  }
  {
    int rc = clear_status(h);
    if (rc) return rc;
  }
  const int mode = rslqr_cuda_get_mode(h);
  if (mode == RSLQR_MODE_SCAN && !h->scan.f.available) {  // large blocks: elements in HBM, a CTA per merge (kernels_big.cuh)
    if (big_alloc(h->s, h->batch, &h->big, g_err)) return RSLQR_E_NOMEM;
    const Shape& s = h->s;
    const size_t f = (size_t)h->r_first;
    Bufs b = h->b;
    b.knots += f * s.N * s.stride; b.x0 += f * s.n; b.soln += f * s.N * s.zs;
    b.status += f * 3 * s.N; b.flags += f;
    BigPlan q = h->big;
    q.bb.entA += f * s.N * big_ent_doubles(q.ns); q.bb.entB += f * (s.N / 2) * big_ent_doubles(q.ns);
    q.bb.rec += f * s.N * big_rec_doubles(q.ns); q.bb.lam += f * (s.N + 1) * q.ns;
    q.bb.ufac += f * s.N * big_fac_doubles(q.ns);
    int rc = big_launch(q, s, b, h->r_count, pack_rhs, h->stream, g_err);
    if (rc) return rc;
    h->launches = q.launches;
    return RSLQR_OK;
  }
  if (mode == RSLQR_MODE_SCAN) {
    PdlScope pdl(h->pdl && !h->ktime_on);
    std::pair<cudaEvent_t, cudaEvent_t> kt;
    if (h->ktime_on) { cudaEventCreate(&kt.first); cudaEventCreate(&kt.second); }
    const Shape& s = h->s;
    const size_t f = (size_t)h->r_first;
    Bufs b = h->b;
    b.knots += f * s.N * s.stride; b.x0 += f * s.n; b.soln += f * s.N * s.zs;
    b.status += f * 3 * s.N; b.flags += f; b.leafd += f * s.N * (s.n + s.m);
    ScanPlan q = h->scan;
    {
      const size_t ent = q.ent, wsz = (2 * (size_t)s.n + 1) * s.n;
      q.f.fb.sep += f * q.f.M0 * wsz; q.f.fb.lamhat += f * (q.f.M0 + 1) * s.n;
      q.sb.sep_in += f * q.f.M0 * 3 * wsz; q.sb.ent0 += f * q.f.M0 * ent;
      q.f.fb.ent[0] = q.sb.ent0;
      for (int t = 1; t < q.f.ntiers; ++t) q.f.fb.ent[t] += f * q.f.tiers[t].Min * ent;
    }
    int rc = scan_launch(q, s, b, h->r_count, pack_rhs, h->stream, g_err, h->ktime_on ? kt.first : nullptr,
                         h->ktime_on ? kt.second : nullptr);
    if (h->ktime_on) h->kspans.push_back(kt);
    if (rc) return rc;
    h->launches = q.f.launches;
    return RSLQR_OK;
  }
  if (mode == RSLQR_MODE_FUSED) {
    std::pair<cudaEvent_t, cudaEvent_t> kt;
    if (h->ktime_on) { cudaEventCreate(&kt.first); cudaEventCreate(&kt.second); }
    // the kernels index problems from 0: hand them views that start at the first problem of the range
    const Shape& s = h->s;
    const size_t f = (size_t)h->r_first;
    Bufs b = h->b;
    b.knots += f * s.N * s.stride; b.x0 += f * s.n; b.soln += f * s.N * s.zs;
    b.status += f * 3 * s.N; b.flags += f; b.leafd += f * s.N * (s.n + s.m);
    FusedPlan q = h->fused;
    if (q.M0 > 1) {
      const size_t ent = fused_ent_doubles(s.n, s.m), wsz = (2 * (size_t)s.n + 1) * s.n;
      q.fb.stL += f * s.N * s.zs * s.n; q.fb.stR += f * s.N * s.zs * s.n; q.fb.stz += f * s.N * s.zs;
      q.fb.sep += f * q.M0 * wsz; q.fb.lamhat += f * (q.M0 + 1) * s.n;
      for (int t = 0; t < q.ntiers; ++t) q.fb.ent[t] += f * q.tiers[t].Min * ent;
    }
    int rc = fused_launch(q, s, b, h->r_count, pack_rhs, h->force_dfma, h->stream, g_err,
                          h->ktime_on ? kt.first : nullptr, h->ktime_on ? kt.second : nullptr);
    if (h->ktime_on) h->kspans.push_back(kt);
    if (rc) return rc;
    h->launches = q.launches;
    return RSLQR_OK;
  }
  return solve_staged(h, pack_rhs);
}

// Host buffers in, host solutions out: uploads, solves and downloads are cut into chunks of problems and
// overlapped on three streams (copy-in, compute, copy-out), so the PCIe transfers hide the kernels.
extern "C" int rslqr_cuda_solve_host(rslqr_cuda_solver* h, const double* knots, const double* x0, double* soln,
                                     int count, int chunk) {
  if (!h || !knots || !x0 || !soln || count <= 0 || count > h->batch) return fail(RSLQR_E_ARG, "solve_host: bad argument");
  CU(cudaSetDevice(h->device));
  const Shape& s = h->s;
  if (chunk <= 0) {
    // ~64 MB of input per chunk keeps the first upload (which nothing overlaps) short
    const size_t per = sizeof(double) * (size_t)s.N * s.stride;
    chunk = (int)((64u << 20) / per);
    if (chunk < 1) chunk = 1;
  }
  if (chunk > count) chunk = count;
  const int nch = (count + chunk - 1) / chunk;
  if (!h->pipe_made) {
    CU(cudaStreamCreateWithFlags(&h->s_h2d, cudaStreamNonBlocking));
    CU(cudaStreamCreateWithFlags(&h->s_d2h, cudaStreamNonBlocking));
    CU(cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming));
    h->pipe_made = true;
  }
  while ((int)h->ev_up.size() < nch) {
    cudaEvent_t a, b2;
    CU(cudaEventCreateWithFlags(&a, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&b2, cudaEventDisableTiming));
    h->ev_up.push_back(a); h->ev_done.push_back(b2);
  }
  // the copy-in must not overtake work already queued on the compute stream that still reads the inputs
  CU(cudaEventRecord(h->ev_join, h->stream));
  CU(cudaStreamWaitEvent(h->s_h2d, h->ev_join, 0));
  CU(cudaStreamWaitEvent(h->s_d2h, h->ev_join, 0));
  const size_t kper = (size_t)s.N * s.stride, zper = (size_t)s.N * s.zs;
  int launches = 0, rc = 0;
  for (int c = 0; c < nch && !rc; ++c) {
    const int first = c * chunk, cnt = (first + chunk <= count) ? chunk : count - first;
    CU(cudaMemcpyAsync(h->d_knots + first * kper, knots + first * kper, sizeof(double) * kper * cnt, cudaMemcpyHostToDevice, h->s_h2d));
    CU(cudaMemcpyAsync(h->d_x0 + (size_t)first * s.n, x0 + (size_t)first * s.n, sizeof(double) * (size_t)s.n * cnt, cudaMemcpyHostToDevice, h->s_h2d));
    CU(cudaEventRecord(h->ev_up[c], h->s_h2d));
    CU(cudaStreamWaitEvent(h->stream, h->ev_up[c], 0));
    h->r_first = first; h->r_count = cnt;
    h->prof_keep = c > 0;  // RSLQR_PROFILE: the phase times cover all chunks, not only the last one
    rc = rslqr_cuda_solve_async(h, 1);
    h->prof_keep = false;
    launches += h->launches;
    h->r_first = 0; h->r_count = h->batch;
    if (rc) break;
    CU(cudaEventRecord(h->ev_done[c], h->stream));
    CU(cudaStreamWaitEvent(h->s_d2h, h->ev_done[c], 0));
    CU(cudaMemcpyAsync(soln + first * zper, h->b.soln + first * zper, sizeof(double) * zper * cnt, cudaMemcpyDeviceToHost, h->s_d2h));
  }
  h->launches = launches;
  if (rc) return rc;
  CU(cudaEventRecord(h->ev_join, h->s_d2h));
  CU(cudaStreamWaitEvent(h->stream, h->ev_join, 0));
  return RSLQR_OK;
}

extern "C" int rslqr_cuda_sync(rslqr_cuda_solver* h) {
  if (!h) return fail(RSLQR_E_ARG, "handle is NULL");
  CU(cudaSetDevice(h->device));
  CU(cudaStreamSynchronize(h->stream));
  if (!h->spans.empty()) {
    memset(h->prof, 0, sizeof(h->prof));
    for (auto& sp : h->spans) {
      float ms = 0;
      cudaEventElapsedTime(&ms, sp.t0, sp.t1);
      h->prof[sp.phase] += ms;
      h->prof[0] += ms;
    }
    h->spans.clear();
  }
  for (auto& pr : h->kspans) {
    float ms = 0;
    if (cudaEventElapsedTime(&ms, pr.first, pr.second) == cudaSuccess) { h->k_ms += ms; ++h->k_n; }
    cudaEventDestroy(pr.first); cudaEventDestroy(pr.second);
  }
  h->kspans.clear();
  return RSLQR_OK;
}

extern "C" int rslqr_cuda_solve(rslqr_cuda_solver* h, int pack_rhs) {
  int rc = rslqr_cuda_solve_async(h, pack_rhs);
  if (rc) return rc;
  return rslqr_cuda_sync(h);
}

extern "C" int rslqr_cuda_stage(rslqr_cuda_solver* h, int stage, int level, int item, int col, int problem) {
  if (!h) return fail(RSLQR_E_ARG, "handle is NULL");
  if (h->shard) return fail(RSLQR_E_UNSUPPORTED, "stage calls are not available on a shard");
  const Shape& s = h->s;
  if (level < 0 || level >= s.K || problem >= h->batch) return fail(RSLQR_E_ARG, "stage: level/problem out of range");
  CU(cudaSetDevice(h->device));
  {
    int rc = ensure_fact(h);
    if (rc) return rc;
  }
  const bool per_knot = stage == RSLQR_STAGE_LEAVES || stage == RSLQR_STAGE_SCHUR ||
                        stage == RSLQR_STAGE_RHS_SCHUR || stage == RSLQR_STAGE_PACK_RHS;
  const int nall = per_knot ? s.N : (1 << (s.K - level - 1));
  if (item >= nall || col >= s.K) return fail(RSLQR_E_ARG, "stage: item/col out of range");
  Sel sel = all_items(h, nall);
  if (problem >= 0) { sel.p0 = problem; sel.np = 1; }
  if (item >= 0) { sel.item0 = item; sel.nitems = 1; }
  sel.col = col;
  CU(smem_optin(solves_kernel, (int)(s.n * s.n * 2 * sizeof(double))));
  int rc = launch_stage(h, stage, level, sel);
  if (rc) return rc;
  CU(cudaStreamSynchronize(h->stream));
  return RSLQR_OK;
}

extern "C" int rslqr_cuda_schur_task(rslqr_cuda_solver* h, int index, int knot, int level, int upper_level,
                                     int calc_lambda, int use_rhs, int problem) {
  if (!h) return fail(RSLQR_E_ARG, "handle is NULL");
  if (h->shard) return fail(RSLQR_E_UNSUPPORTED, "stage calls are not available on a shard");
  {
    CU(cudaSetDevice(h->device));
    int rc = ensure_fact(h);
    if (rc) return rc;
  }
  const Shape& s = h->s;
  if (level < 0 || level >= s.K || upper_level < 0 || upper_level >= s.K || knot < 0 || knot >= s.N ||
      index < 0 || index >= s.N - 1 || problem >= h->batch)
    return fail(RSLQR_E_ARG, "schur_task: argument out of range");
  CU(cudaSetDevice(h->device));
  Sel sel = all_items(h, 1);
  if (problem >= 0) { sel.p0 = problem; sel.np = 1; }
  sel.item0 = knot; sel.col = upper_level; sel.calc_lambda = calc_lambda ? 1 : 0; sel.index = index;
  int rc = launch_stage(h, use_rhs ? RSLQR_STAGE_RHS_SCHUR : RSLQR_STAGE_SCHUR, level, sel);
  if (rc) return rc;
  CU(cudaStreamSynchronize(h->stream));
  return RSLQR_OK;
}

// ------------------------------------------------------------------ device -> host
extern "C" int rslqr_cuda_download_solution(rslqr_cuda_solver* h, double* soln, int first, int count) {
  if (!h || !soln || !range_ok(h, first, count)) return fail(RSLQR_E_ARG, "download_solution: bad argument");
  CU(cudaSetDevice(h->device));
  const size_t per = (size_t)h->s.N * h->s.zs;
  CU(cudaMemcpyAsync(soln, h->b.soln + first * per, sizeof(double) * per * count, cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  return RSLQR_OK;
}

extern "C" int rslqr_cuda_download_fact(rslqr_cuda_solver* h, double* fact, int problem) {
  if (!h || !fact || problem < 0 || problem >= h->batch) return fail(RSLQR_E_ARG, "download_fact: bad argument");
  if (!h->b.fact) return fail(RSLQR_E_UNSUPPORTED, "download_fact: no factor array yet (staged mode, a stage call or rslqr_cuda_ensure_fact creates it)");
  CU(cudaSetDevice(h->device));
  const size_t per = (size_t)h->s.K * h->s.N * h->s.fs;
  CU(cudaMemcpyAsync(fact, h->b.fact + problem * per, sizeof(double) * per, cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  return RSLQR_OK;
}

// host -> device copy of one problem's dense factor array (the stage calls of the host API treat the host mirror as
// authoritative between calls: the reference's tests run helper routines on views of solver->fact in between)
extern "C" int rslqr_cuda_upload_fact(rslqr_cuda_solver* h, const double* fact, int problem) {
  if (!h || !fact || problem < 0 || problem >= h->batch) return fail(RSLQR_E_ARG, "upload_fact: bad argument");
  if (h->shard) return fail(RSLQR_E_UNSUPPORTED, "a shard keeps no factor array");
  CU(cudaSetDevice(h->device));
  int rc = ensure_fact(h);
  if (rc) return rc;
  const size_t per = (size_t)h->s.K * h->s.N * h->s.fs;
  h->factored = 0;
  CU(cudaMemcpyAsync(h->b.fact + problem * per, fact, sizeof(double) * per, cudaMemcpyHostToDevice, h->stream));
  return RSLQR_OK;
}

extern "C" int rslqr_cuda_download_status(rslqr_cuda_solver* h, int* status, int problem) {
  if (!h || !status || problem < 0 || problem >= h->batch) return fail(RSLQR_E_ARG, "download_status: bad argument");
  CU(cudaSetDevice(h->device));
  const size_t per = (size_t)3 * h->s.N;
  CU(cudaMemcpyAsync(status, h->b.status + problem * per, sizeof(int) * (per - 1), cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  return RSLQR_OK;
}

extern "C" int rslqr_cuda_download_flags(rslqr_cuda_solver* h, int* flags, int first, int count) {
  if (!h || !flags || !range_ok(h, first, count)) return fail(RSLQR_E_ARG, "download_flags: bad argument");
  CU(cudaSetDevice(h->device));
  CU(cudaMemcpyAsync(flags, h->b.flags + first, sizeof(int) * count, cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  return RSLQR_OK;
}

extern "C" int rslqr_cuda_download_leaf_factors(rslqr_cuda_solver* h, double* out, int problem) {
  if (!h || !out || problem < 0 || problem >= h->batch) return fail(RSLQR_E_ARG, "download_leaf_factors: bad argument");
  CU(cudaSetDevice(h->device));
  const size_t per = (size_t)h->s.N * (h->s.n + h->s.m);
  CU(cudaMemcpyAsync(out, h->b.leafd + problem * per, sizeof(double) * per, cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  return RSLQR_OK;
}

extern "C" int rslqr_cuda_get_profile(rslqr_cuda_solver* h, double ms[7]) {
  if (!h || !ms) return fail(RSLQR_E_ARG, "get_profile: bad argument");
  memcpy(ms, h->prof, sizeof(h->prof));
  return RSLQR_OK;
}

extern "C" int rslqr_cuda_device_pointers(rslqr_cuda_solver* h, double** knots, double** x0, double** soln, double** fact) {
  if (!h) return fail(RSLQR_E_ARG, "handle is NULL");
  if (knots) *knots = h->d_knots;
  if (x0) *x0 = h->d_x0;
  if (soln) *soln = h->b.soln;
  if (fact) *fact = h->b.fact;
  return RSLQR_OK;
}

extern "C" void* rslqr_cuda_stream(rslqr_cuda_solver* h) { return h ? (void*)h->stream : nullptr; }
extern "C" int rslqr_cuda_kernel_launches(rslqr_cuda_solver* h) { return h ? h->launches : RSLQR_E_ARG; }

extern "C" int rslqr_cuda_kernel_timing(rslqr_cuda_solver* h, int enable, int reset, double* avg_ms, int* launches) {
  if (!h) return fail(RSLQR_E_ARG, "handle is NULL");
  if (avg_ms) *avg_ms = h->k_n ? h->k_ms / h->k_n : 0.0;
  if (launches) *launches = h->k_n;
  if (reset) { h->k_ms = 0; h->k_n = 0; }
  h->ktime_on = enable != 0;
  return RSLQR_OK;
}

extern "C" void* rslqr_cuda_host_alloc(size_t bytes) {
  void* p = nullptr;
  if (cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess) {
    cudaGetLastError();
    g_err = "cudaHostAlloc failed";
    return nullptr;
  }
  return p;
}
extern "C" void rslqr_cuda_host_free(void* p) {
  if (p) cudaFreeHost(p);
}
extern "C" int rslqr_cuda_host_register(void* p, size_t bytes) {
  if (!p || !bytes) return fail(RSLQR_E_ARG, "host_register: bad argument");
  if (cudaHostRegister(p, bytes, cudaHostRegisterDefault) != cudaSuccess) {
    cudaGetLastError();
    return fail(RSLQR_E_CUDA, "cudaHostRegister failed");
  }
  return RSLQR_OK;
}
extern "C" int rslqr_cuda_host_unregister(void* p) {
  if (!p) return fail(RSLQR_E_ARG, "host_unregister: bad argument");
  if (cudaHostUnregister(p) != cudaSuccess) {
    cudaGetLastError();
    return fail(RSLQR_E_CUDA, "cudaHostUnregister failed");
  }
  return RSLQR_OK;
}

extern "C" int rslqr_cuda_mark(rslqr_cuda_solver* h, int slot) {
  if (!h || slot < 0 || slot >= 8) return fail(RSLQR_E_ARG, "mark: bad argument");
  CU(cudaSetDevice(h->device));
  if (!h->marks_made) {
    for (auto& ev : h->marks) CU(cudaEventCreate(&ev));
    h->marks_made = true;
  }
  CU(cudaEventRecord(h->marks[slot], h->stream));
  return RSLQR_OK;
}

extern "C" double rslqr_cuda_elapsed_ms(rslqr_cuda_solver* h, int a, int b) {
  if (!h || !h->marks_made || a < 0 || a >= 8 || b < 0 || b >= 8) return -1.0;
  cudaSetDevice(h->device);
  if (cudaEventSynchronize(h->marks[b]) != cudaSuccess) return -1.0;
  float ms = 0;
  if (cudaEventElapsedTime(&ms, h->marks[a], h->marks[b]) != cudaSuccess) return -1.0;
  return ms;
}

extern "C" int rslqr_cuda_upload_async(rslqr_cuda_solver* h, const double* knots, const double* x0, int first, int count) {
  if (!h || !knots || !x0 || !range_ok(h, first, count)) return fail(RSLQR_E_ARG, "upload_async: bad argument");
  CU(cudaSetDevice(h->device));
  const Shape& s = h->s;
  const size_t per = (size_t)s.N * s.stride;
  CU(cudaMemcpyAsync(h->d_knots + first * per, knots, sizeof(double) * per * count, cudaMemcpyHostToDevice, h->stream));
  CU(cudaMemcpyAsync(h->d_x0 + (size_t)first * s.n, x0, sizeof(double) * (size_t)s.n * count, cudaMemcpyHostToDevice, h->stream));
  return RSLQR_OK;
}

extern "C" int rslqr_cuda_upload_rhs_async(rslqr_cuda_solver* h, const double* soln, int first, int count) {
  if (!h || !soln || !range_ok(h, first, count)) return fail(RSLQR_E_ARG, "upload_rhs_async: bad argument");
  CU(cudaSetDevice(h->device));
  const size_t per = (size_t)h->s.N * h->s.zs;
  CU(cudaMemcpyAsync(h->b.soln + first * per, soln, sizeof(double) * per * count, cudaMemcpyHostToDevice, h->stream));
  return RSLQR_OK;
}

extern "C" int rslqr_cuda_download_status_async(rslqr_cuda_solver* h, int* status, int problem) {
  if (!h || !status || problem < 0 || problem >= h->batch) return fail(RSLQR_E_ARG, "download_status_async: bad argument");
  CU(cudaSetDevice(h->device));
  const size_t per = (size_t)3 * h->s.N;
  CU(cudaMemcpyAsync(status, h->b.status + problem * per, sizeof(int) * (per - 1), cudaMemcpyDeviceToHost, h->stream));
  return RSLQR_OK;
}

extern "C" int rslqr_cuda_download_solution_async(rslqr_cuda_solver* h, double* soln, int first, int count) {
  if (!h || !soln || !range_ok(h, first, count)) return fail(RSLQR_E_ARG, "download_solution_async: bad argument");
  CU(cudaSetDevice(h->device));
  const size_t per = (size_t)h->s.N * h->s.zs;
  CU(cudaMemcpyAsync(soln, h->b.soln + first * per, sizeof(double) * per * count, cudaMemcpyDeviceToHost, h->stream));
  return RSLQR_OK;
}

// ------------------------------------------------------------------ shards of one long horizon
extern "C" int rslqr_cuda_create_shard(int n, int m, int nhorizon, int first_knot, int shard_knots, int device,
                                       unsigned flags, rslqr_cuda_solver** out) {
  if (!out) return fail(RSLQR_E_ARG, "out is NULL");
  *out = nullptr;
  if (nhorizon < 2 || (nhorizon & (nhorizon - 1)) || shard_knots < 16 || (shard_knots & (shard_knots - 1)) ||
      first_knot < 0 || first_knot % shard_knots || first_knot + shard_knots > nhorizon)
    return fail(RSLQR_E_ARG, "shard: need power-of-two horizon and shard size (>= 16), first_knot a multiple of the shard size");
  int rc = rslqr_cuda_create(n, m, shard_knots, 1, device, (flags & ~RSLQR_KEEP_FACT) | RSLQR_INTERNAL_SHARD, out);
  if (rc) return rc;
  rslqr_cuda_solver* h = *out;
  if (h->scan.dual) {
    // the shard kernels exist for the boundary-entry form only: a shape the scan path covers in dual form runs its
    // shards on the tiered kernels
    scan_free(&h->scan);
    if (!h->fused.available && fused_plan(h->s, h->batch, h->g0_pref, &h->fused, g_err)) {
      rslqr_cuda_destroy(h);
      *out = nullptr;
      return RSLQR_E_NOMEM;
    }
  }
  if (!(h->scan.f.available && h->scan.f.ntiers >= 1) && !(h->fused.available && h->fused.ntiers >= 1)) {
    rslqr_cuda_destroy(h);
    *out = nullptr;
    return fail(RSLQR_E_UNSUPPORTED, "shards need a shape covered by the scan or tiered kernels");
  }
  h->shard = true;
  h->s.kofs = first_knot;
  h->s.Ng = nhorizon;
  return RSLQR_OK;
}

static size_t shard_payload(const rslqr_cuda_solver* h) {
  return (h->scan.f.available && h->scan.hybrid) ? dual_ent_doubles(h->s.n) : fused_payload_doubles(h->s.n, h->s.m);
}
extern "C" size_t rslqr_cuda_shard_payload_doubles(rslqr_cuda_solver* h) { return h ? shard_payload(h) : 0; }

extern "C" int rslqr_cuda_shard_reduce(rslqr_cuda_solver* h, double* d_payload) {
  if (!h || !d_payload || !h->shard) return fail(RSLQR_E_ARG, "shard_reduce: bad argument");
  CU(cudaSetDevice(h->device));
  PdlScope pdl(h->pdl);
  {
    int rc = clear_status(h);  // (a time-out or failure of the previous solve must not stick)
    if (rc) return rc;
  }
  if (rslqr_cuda_get_mode(h) == RSLQR_MODE_SCAN) {
    int rc = h->scan.hybrid ? dual_shard_reduce(h->scan, h->s, h->b, 1, h->stream, d_payload, g_err)
                            : scan_shard_reduce(h->scan, h->s, h->b, 1, h->stream, d_payload, g_err);
    h->launches = h->scan.f.launches;
    return rc;
  }
  int rc = fused_shard_reduce(h->fused, h->s, h->b, 1, h->force_dfma, h->stream, d_payload, g_err);
  h->launches = h->fused.launches;
  return rc;
}

extern "C" int rslqr_cuda_shard_finish(rslqr_cuda_solver* h, const double* d_gathered, int rank, int nranks) {
  if (!h || !h->shard || nranks < 1 || nranks > 8 || (nranks & (nranks - 1)) || rank < 0 || rank >= nranks ||
      (nranks > 1 && !d_gathered) || h->s.kofs != rank * h->s.N || h->s.Ng != nranks * h->s.N)
    return fail(RSLQR_E_ARG, "shard_finish: bad argument (rank / nranks must match the shard's position; nranks <= 8)");
  CU(cudaSetDevice(h->device));
  PdlScope pdl(h->pdl);
  if (rslqr_cuda_get_mode(h) == RSLQR_MODE_SCAN) {
    int rc = h->scan.hybrid ? dual_shard_finish(h->scan, h->s, h->b, 1, h->stream, d_gathered, rank, nranks, g_err)
                            : scan_shard_finish(h->scan, h->s, h->b, 1, h->stream, d_gathered, rank, nranks, g_err);
    h->launches = h->scan.f.launches;
    return rc;
  }
  int rc = fused_shard_finish(h->fused, h->s, h->b, h->stream, d_gathered, rank, nranks, g_err);
  h->launches = h->fused.launches;
  return rc;
}

// Peer-to-peer exchange (NVLink): instead of an all-gather collective, shard_push stores this rank's payload straight
// into every rank's gathered buffer and raises a flag there; shard_finish_wait lets the cross-GPU tier kernel wait for
// the flags of all ranks. `peer_gathered[r]` / `peer_flags[r]` are rank r's buffers as mapped into THIS process
// (torch symmetric memory / cudaIpc; plain device pointers when the ranks are handles of one process).
extern "C" int rslqr_cuda_shard_push(rslqr_cuda_solver* h, const double* d_payload, double* const* peer_gathered,
                                     unsigned long long* const* peer_flags, int rank, int nranks, unsigned long long epoch) {
  if (!h || !h->shard || !d_payload || !peer_gathered || !peer_flags || nranks < 1 || nranks > 8 || rank < 0 || rank >= nranks)
    return fail(RSLQR_E_ARG, "shard_push: bad argument");
  CU(cudaSetDevice(h->device));
  PeerPtrs pp;
  memset(&pp, 0, sizeof(pp));
  for (int r = 0; r < nranks; ++r) { pp.gathered[r] = peer_gathered[r]; pp.flags[r] = peer_flags[r]; }
  const int pay = (int)shard_payload(h);
  PdlScope pdl(h->pdl);
  launch_k(shard_push_kernel, dim3(nranks), dim3(256), (size_t)0, h->stream, d_payload, pp, rank, pay, epoch);
  CU(cudaGetLastError());
  ++h->launches;
  return RSLQR_OK;
}

extern "C" int rslqr_cuda_shard_finish_wait(rslqr_cuda_solver* h, const double* d_gathered, const unsigned long long* d_flags,
                                            unsigned long long epoch, int rank, int nranks) {
  if (!h || !h->shard || nranks < 1 || nranks > 8 || (nranks & (nranks - 1)) || rank < 0 || rank >= nranks ||
      (nranks > 1 && (!d_gathered || !d_flags)) || h->s.kofs != rank * h->s.N || h->s.Ng != nranks * h->s.N)
    return fail(RSLQR_E_ARG, "shard_finish_wait: bad argument");
  if (rslqr_cuda_get_mode(h) != RSLQR_MODE_SCAN) return fail(RSLQR_E_UNSUPPORTED, "the peer-to-peer exchange needs the scan kernels");
  CU(cudaSetDevice(h->device));
  PdlScope pdl(h->pdl);
  int rc = h->scan.hybrid ? dual_shard_finish(h->scan, h->s, h->b, 1, h->stream, d_gathered, rank, nranks, g_err, d_flags, epoch)
                          : scan_shard_finish(h->scan, h->s, h->b, 1, h->stream, d_gathered, rank, nranks, g_err, d_flags, epoch);
  h->launches = h->scan.f.launches;
  return rc;
}

// Enqueue on a caller-owned stream (e.g. the one NCCL collectives are ordered on) instead of the handle's own.
extern "C" int rslqr_cuda_set_stream(rslqr_cuda_solver* h, void* stream) {
  if (!h) return fail(RSLQR_E_ARG, "handle is NULL");
  CU(cudaSetDevice(h->device));
  CU(cudaStreamSynchronize(h->stream));
  if (h->own_stream) cudaStreamDestroy(h->stream);
  h->stream = (cudaStream_t)stream;
  h->own_stream = false;
  return RSLQR_OK;
}

// Single-process exchange for `nranks` shard handles (any devices): the all-gather of the payloads as peer copies.
// d_payload[r] / d_gathered[r] are device buffers on shard r's device (payload / nranks * payload doubles).
extern "C" int rslqr_cuda_shard_exchange(rslqr_cuda_solver** hs, int nranks, double** d_payload, double** d_gathered) {
  if (!hs || !d_payload || !d_gathered || nranks < 1) return fail(RSLQR_E_ARG, "shard_exchange: bad argument");
  const size_t pay = shard_payload(hs[0]);
  for (int r = 0; r < nranks; ++r) {  // payloads complete
    CU(cudaSetDevice(hs[r]->device));
    CU(cudaStreamSynchronize(hs[r]->stream));
  }
  for (int dst = 0; dst < nranks; ++dst) {
    CU(cudaSetDevice(hs[dst]->device));
    for (int src = 0; src < nranks; ++src)
      CU(cudaMemcpyPeerAsync(d_gathered[dst] + src * pay, hs[dst]->device, d_payload[src], hs[src]->device,
                             sizeof(double) * pay, hs[dst]->stream));
  }
  return RSLQR_OK;
}

// ------------------------------------------------------------------ solve without any allocation
// rslqr_cuda_solve_static: the scan-path solve (the launches of rslqr_cuda_solve_async in RSLQR_MODE_SCAN, same
// kernels, same order) for callers that cannot allocate: code under CUDA graph capture, tight serving loops. No handle:
// the caller passes the device buffers of the problems, the solutions and the per-problem flags; everything the
// kernels keep between launches (elements of the subtrees, separator records, multipliers of the upper separators,
// status words) lives in ONE file-scope device array of RSLQR_STATIC_SCRATCH_BYTES that every GPU loading the library
// holds until exit. In host code `g_static_scratch` is only the host-side stand-in of that array: its device address
// comes from cudaGetSymbolAddress, once per device, with that device current (static_scratch_base).
// All calls on a device share the array: calls that overlap on the GPU (different streams) race.
static __device__ __align__(256) unsigned char g_static_scratch[RSLQR_STATIC_SCRATCH_BYTES];

static int static_scratch_base(int device, unsigned char** out) {
  constexpr int MAXDEV = 64;
  static std::mutex mu;
  static unsigned char* base[MAXDEV] = {};
  if (device < 0 || device >= MAXDEV) return fail(RSLQR_E_ARG, "solve_static: device index out of range");
  std::lock_guard<std::mutex> lock(mu);
  if (!base[device]) {  // (the caller made `device` current)
    unsigned char* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, g_static_scratch);
    if (e != cudaSuccess || !p) {
      cudaGetLastError();
      return fail(RSLQR_E_CUDA, std::string("solve_static: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
    }
    base[device] = p;
  }
  *out = base[device];
  return RSLQR_OK;
}

extern "C" int rslqr_cuda_solve_static(int n, int m, int nhorizon, int batch, int device, const double* d_knots,
                                       const double* d_x0, double* d_soln, int* d_flags, void* stream) {
  if (!d_knots || !d_x0 || !d_soln || !d_flags) return fail(RSLQR_E_ARG, "solve_static: a buffer is NULL");
  if (batch <= 0 || nhorizon < 2 * SCAN_LEAF || (nhorizon & (nhorizon - 1)))
    return fail(RSLQR_E_ARG, "solve_static: need batch > 0 and a power-of-two horizon >= 8");
  if (!dual_shape_ok(n, m)) return fail(RSLQR_E_UNSUPPORTED, "solve_static: shape not covered by the scan kernels");
  const int ndev = rslqr_cuda_device_count();
  if (ndev <= 0) return fail(RSLQR_E_NODEVICE, "no CUDA device available (there is no CPU fallback)");
  if (device < 0 || device >= ndev) return fail(RSLQR_E_ARG, "solve_static: device index out of range");
  CU(cudaSetDevice(device));
  const Shape s = make_shape(n, m, nhorizon);
  // the plan scan_plan would make, with its buffers carved out of the static array instead of allocated
  ScanPlan p;
  memset(&p, 0, sizeof(p));
  FusedPlan& f = p.f;
  p.hybrid = n == 12 && m == 4;
  p.dual = !p.hybrid;
  p.ent = dual_ent_doubles(n);
  f.ns = n; f.ni = m; f.G0 = SCAN_LEAF; f.M0 = nhorizon / SCAN_LEAF; f.use_dmma = 1;
  {
    int tier_e = batch >= 64 ? 4 : 8;  // (scan_plan's rule)
    for (;;) {
      int M = f.M0, span = f.G0;
      f.ntiers = 0;
      bool fits = true;
      while (M > 1) {
        if (f.ntiers == 8) { fits = false; break; }
        const int e = M < tier_e ? M : tier_e;
        f.tiers[f.ntiers++] = FusedTier{M, e, span};
        M /= e; span *= e;
      }
      if (fits) break;
      if (tier_e == 8) return fail(RSLQR_E_UNSUPPORTED, "solve_static: horizon too long for the tier slots");
      tier_e = 8;
    }
  }
  const size_t wsz = (2 * (size_t)n + 1) * n, B = (size_t)batch;
  size_t need = 0;  // bytes, every piece 256-byte aligned (the bulk copies need 16)
  auto piece = [&](size_t bytes) { const size_t at = need; need += (bytes + 255) & ~(size_t)255; return at; };
  const size_t o_sep = piece(sizeof(double) * B * f.M0 * wsz);
  const size_t o_lam = piece(sizeof(double) * B * (f.M0 + 1) * n);
  const size_t o_sin = piece(sizeof(double) * B * f.M0 * 3 * wsz);
  const size_t o_ent0 = piece(sizeof(double) * B * f.M0 * p.ent);
  size_t o_ent[8] = {};
  for (int t = 1; t < f.ntiers; ++t) o_ent[t] = piece(sizeof(double) * B * f.tiers[t].Min * p.ent);
  const size_t status_bytes = sizeof(int) * B * 3 * nhorizon;
  const size_t o_status = piece(status_bytes);
  if (need > (size_t)RSLQR_STATIC_SCRATCH_BYTES) {
    char msg[256];
    snprintf(msg, sizeof(msg),
             "solve_static: n=%d m=%d N=%d batch=%d needs %zu bytes of scratch, the static array holds %zu "
             "(RSLQR_STATIC_SCRATCH_BYTES); use a handle (rslqr_cuda_create) for larger problems",
             n, m, nhorizon, batch, need, (size_t)RSLQR_STATIC_SCRATCH_BYTES);
    return fail(RSLQR_E_ARG, msg);
  }
  unsigned char* base = nullptr;
  {
    int rc = static_scratch_base(device, &base);
    if (rc) return rc;
  }
  f.fb.sep = reinterpret_cast<double*>(base + o_sep);
  f.fb.lamhat = reinterpret_cast<double*>(base + o_lam);
  p.sb.sep_in = reinterpret_cast<double*>(base + o_sin);
  p.sb.ent0 = reinterpret_cast<double*>(base + o_ent0);
  f.fb.ent[0] = p.sb.ent0;
  for (int t = 1; t < f.ntiers; ++t) f.fb.ent[t] = reinterpret_cast<double*>(base + o_ent[t]);
  f.bytes = need;
  f.available = true;
  Bufs b;
  memset(&b, 0, sizeof(b));
  b.knots = d_knots; b.x0 = d_x0; b.soln = d_soln; b.flags = d_flags;
  b.status = reinterpret_cast<int*>(base + o_status);
  cudaStream_t st = (cudaStream_t)stream;
  // the array keeps the last call's contents: status words and flags must start at zero, the rest is written before
  // it is read
  CU(cudaMemsetAsync(b.status, 0, status_bytes, st));
  CU(cudaMemsetAsync(d_flags, 0, sizeof(int) * B, st));
  PdlScope pdl(false);
  return scan_launch(p, s, b, batch, 1, st, g_err);
}

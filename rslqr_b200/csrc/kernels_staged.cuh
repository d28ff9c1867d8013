// kernels_staged.cuh -- "staged" device path: one launch per phase per level, factors resident in HBM.
//
// Each kernel is the batched, structurally-sparse form of one phase of the reference's OpenMP loop
// (reference src/solve.c:59-182): a CTA owns one task (problem, item, column) and its threads own the
// output elements of that task. Shapes (n, m, N) are runtime values, so this path covers every
// configuration; the fused cluster kernel (kernels_fused.cuh) replaces it when a problem fits on chip.
//
// Storage per problem (reference src/nddata.c:42-53,93): fact[(k + N*level)*fs], fs = (2n+m)*n,
// block = [Lambda n*n | X n*n | U m*n] column-major; soln[k*(2n+m)] = [lambda_k | x_k | u_k].
// The KKT `data` array of the reference is never materialised: C(i,j) = (A_i^T, B_i^T) and
// C(i+1,j) = (-I, 0) are read straight from the flat knot blocks.
#pragma once
#include <cuda_runtime.h>
#include <map>
#include <mutex>
#include <utility>
#include "tree_math.h"

namespace rslqr {

// Opt-in to more than 48 KB of dynamic shared memory. cudaFuncSetAttribute applies to the CURRENT device only, so
// what has been granted is remembered per (kernel, device): a process that opens handles on several GPUs gets the
// attribute on each of them (a process-wide flag would leave every device but the first without it).
template <class Kernel>
inline cudaError_t smem_optin(Kernel kernel, int bytes) {
  if (bytes <= 48 * 1024) return cudaSuccess;
  static std::mutex mu;
  static std::map<std::pair<const void*, int>, int> granted;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  std::lock_guard<std::mutex> lock(mu);
  int& have = granted[std::make_pair(reinterpret_cast<const void*>(kernel), dev)];
  if (bytes <= have) return cudaSuccess;
  e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e == cudaSuccess) have = bytes;
  return e;
}

// ---------------------------------------------------------------------------------------------------------
// Programmatic dependent launch (latency mode: one long horizon, few CTAs per kernel). A kernel launched through
// launch_k while pdl_mode() is on carries cudaLaunchAttributeProgrammaticStreamSerialization: it may become resident
// while its predecessor in the stream still runs, and waits in pdl_sync() (griddepcontrol.wait: the predecessor grid
// complete, its writes visible) before it touches any global memory. Launch latency and the CTA prologue then overlap
// the predecessor instead of following it. Every kernel launched through launch_k MUST start with pdl_sync(); without
// the attribute both instructions are no-ops.
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void pdl_sync() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
}
inline bool& pdl_mode() {
  static thread_local bool on = false;
  return on;
}
struct PdlScope {
  bool prev;
  explicit PdlScope(bool on) : prev(pdl_mode()) { pdl_mode() = on; }
  ~PdlScope() { pdl_mode() = prev; }
  PdlScope(const PdlScope&) = delete;
  PdlScope& operator=(const PdlScope&) = delete;
};
template <typename... KA, typename... A>
inline cudaError_t launch_k(void (*kern)(KA...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, A&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at{};
  at.id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at.val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = &at;
  cfg.numAttrs = pdl_mode() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KA>(std::forward<A>(args))...);
}

struct Shape {
  int n, m, N, K;
  int fs;      // (2n+m)*n
  int zs;      // 2n+m
  int stride;  // doubles per flat knot block
  int oQ, oR, oq, orr, oA, oB, od;  // offsets inside a knot block
  int kofs, Ng;  // shard of a longer horizon (tiered path only): global index of local knot 0, global horizon
};

struct Bufs {
  const double* knots;  // [batch][N][stride]
  const double* x0;     // [batch][n]
  double* soln;         // [batch][N][zs]
  double* fact;         // [batch][K][N][fs]   (nullptr without KEEP_FACT)
  int* status;          // [batch][3N]
  int* flags;           // [batch]
  double* leafd;        // [batch][N][n+m]  sqrt of the diagonals (leaf Cholesky factors)
};

// which tasks a launch covers
struct Sel {
  int p0, np;        // problems [p0, p0+np)
  int item0, nitems; // items (knots or leaves) [item0, item0+nitems)
  int col;           // explicit column / upper level, or -1 = structurally active set
  int calc_lambda;   // -1 = rule of the reference (rs_calc_lambda), 0/1 = forced (single Schur task)
  int index;         // explicit separator index for a single Schur task, else -1
};

__device__ __forceinline__ const double* knot_ptr(const Shape& s, const Bufs& b, int prob, int k) {
  return b.knots + ((size_t)prob * s.N + k) * s.stride;
}
__device__ __forceinline__ double* fact_ptr(const Shape& s, const Bufs& b, int prob, int k, int level) {
  return b.fact + ((size_t)prob * s.K * s.N + (size_t)level * s.N + k) * s.fs;
}
__device__ __forceinline__ double* soln_ptr(const Shape& s, const Bufs& b, int prob, int k) {
  return b.soln + ((size_t)prob * s.N + k) * s.zs;
}

// ---------------------------------------------------------------------------------------------
// pack: rhs = -[x0 | q_0 | r_0 | d_0 | q_1 | r_1 | ... | d_{N-2} | q_{N-1} | 0]
// (reference src/solver.c:138-139,156-157,176,181,188-190). One thread per rhs entry.
// ---------------------------------------------------------------------------------------------
__global__ void pack_rhs_kernel(Shape s, Bufs b, Sel sel) {
  const size_t per = (size_t)s.N * s.zs;
  const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= per * sel.np) return;
  const int prob = sel.p0 + (int)(tid / per);
  const int e = (int)(tid % per);
  const int k = e / s.zs, r = e % s.zs;
  const double* kn = knot_ptr(s, b, prob, k);
  double v;
  if (r < s.n) {
    v = (k == 0) ? b.x0[(size_t)prob * s.n + r] : knot_ptr(s, b, prob, k - 1)[s.od + r];
  } else if (r < 2 * s.n) {
    v = kn[s.oq + r - s.n];
  } else {
    v = (k < s.N - 1) ? kn[s.orr + r - 2 * s.n] : 0.0;
  }
  b.soln[(size_t)prob * per + e] = (k == s.N - 1 && r >= 2 * s.n) ? 0.0 : -v;
}

// ---------------------------------------------------------------------------------------------
// leaves: ndlqr_SolveLeaf (reference src/nested_dissection.c:10-105). One CTA per (problem, knot).
// Q_k, R_k are diagonal, so their Cholesky factors are sqrt(diag) and D^{-1}C is a row scaling.
// ---------------------------------------------------------------------------------------------
__global__ void leaf_kernel(Shape s, Bufs b, Sel sel) {
  const int task = blockIdx.x;
  const int prob = sel.p0 + task / sel.nitems;
  const int k = sel.item0 + task % sel.nitems;
  const int n = s.n, m = s.m, nn = n * n;
  const double* kn = knot_ptr(s, b, prob, k);
  const double* Qd = kn + s.oQ;
  const double* Rd = kn + s.oR;
  const double* A = kn + s.oA;
  const double* B = kn + s.oB;
  double* z = soln_ptr(s, b, prob, k);
  int* st = b.status + (size_t)prob * 3 * s.N;
  const bool last = (k == s.N - 1);
  const int t = threadIdx.x, nt = blockDim.x;

  // status words + the leaf Cholesky factors sqrt(Q), sqrt(R)
  for (int i = t; i < n + m; i += nt) {
    const bool isR = i >= n;
    if (isR && last) { b.leafd[((size_t)prob * s.N + k) * (n + m) + i] = 0.0; continue; }
    const double d = isR ? Rd[i - n] : Qd[i];
    b.leafd[((size_t)prob * s.N + k) * (n + m) + i] = d > 0 ? sqrt(d) : d;
    if (!(d > 0)) { st[isR ? rs_slot_R(k) : rs_slot_Q(k)] = -1; b.flags[prob] = 1; }
  }

  if (k == 0) {  // nested_dissection.c:24-59
    double* F = fact_ptr(s, b, prob, 0, 0);
    for (int e = t; e < nn; e += nt) {
      const int i = e % n, c = e / n;
      F[e] = -A[c + n * i];  // Lambda = -A0^T
      F[nn + e] = 0.0;       // X = 0
    }
    for (int e = t; e < m * n; e += nt) {
      const int i = e % m, c = e / m;
      F[2 * nn + e] = B[c + n * i] / Rd[i];  // U = R \ B0^T
    }
    // rhs: zy <- -Q zy - zx ; zx <- -zy(old) ; zu <- R \ zu
    for (int i = t; i < n; i += nt) {
      const double zy = z[i], zx = z[n + i];
      z[i] = -Qd[i] * zy - zx;
      z[n + i] = -zy;
    }
    for (int i = t; i < m; i += nt) z[2 * n + i] = z[2 * n + i] / Rd[i];
    return;
  }
  if (!last) {  // nested_dissection.c:70-87
    double* F = fact_ptr(s, b, prob, k, rs_knot_level(k));
    for (int e = t; e < nn; e += nt) {
      const int i = e % n, c = e / n;
      F[nn + e] = A[c + n * i] / Qd[i];  // X = Q \ A^T
    }
    for (int e = t; e < m * n; e += nt) {
      const int i = e % m, c = e / m;
      F[2 * nn + e] = B[c + n * i] / Rd[i];  // U = R \ B^T
    }
    for (int i = t; i < m; i += nt) z[2 * n + i] = z[2 * n + i] / Rd[i];
  }
  for (int i = t; i < n; i += nt) z[n + i] = z[n + i] / Qd[i];  // :89-90
  double* Fp = fact_ptr(s, b, prob, k, rs_knot_level(k - 1));    // :96-102
  for (int e = t; e < nn; e += nt) {
    const int i = e % n, c = e / n;
    Fp[nn + e] = (i == c) ? -1.0 / Qd[i] : 0.0;  // X = Q \ (-I)
  }
  for (int e = t; e < m * n; e += nt) Fp[2 * nn + e] = 0.0;
}

// ---------------------------------------------------------------------------------------------
// The rhs half of ndlqr_SolveLeaf alone (reference src/nested_dissection.c:41-59, 84-90): used when a stored
// factorization is applied to a new right-hand side (factor once, solve many). Q, R are read from the knot blocks.
// ---------------------------------------------------------------------------------------------
__global__ void leaf_rhs_kernel(Shape s, Bufs b, Sel sel) {
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int per = s.n + s.m;
  if (gid >= (long long)sel.np * sel.nitems * per) return;
  const int i = (int)(gid % per);
  const long long task = gid / per;
  const int prob = sel.p0 + (int)(task / sel.nitems);
  const int k = sel.item0 + (int)(task % sel.nitems);
  const int n = s.n;
  const double* kn = knot_ptr(s, b, prob, k);
  double* z = soln_ptr(s, b, prob, k);
  if (i >= n) {  // z_u <- R \ z_u (no input at the last knot)
    if (k != s.N - 1) z[2 * n + (i - n)] = z[2 * n + (i - n)] / kn[s.oR + (i - n)];
    return;
  }
  const double Qd = kn[s.oQ + i];
  if (k == 0) {  // zy <- -Q zy - zx ; zx <- -zy(old)
    const double zy = z[i], zx = z[n + i];
    z[i] = -Qd * zy - zx;
    z[n + i] = -zy;
  } else {
    z[n + i] = z[n + i] / Qd;
  }
}

// decode (leaf, column-slot) for the product / solve kernels. Slot 0 = own column (products only),
// 1 = left boundary column, 2 = right boundary column. Returns -1 when the slot is absent.
__device__ __forceinline__ int active_col(const Shape& s, int leaf, int level, int slot) {
  if (slot == 0) return level;
  const int a = leaf << (level + 1);
  const int bnd = a + (2 << level) - 1;
  return slot == 1 ? rs_left_col(a) : rs_right_col(bnd, s.N);
}

// ---------------------------------------------------------------------------------------------
// products: ndlqr_FactorInnerProduct (reference src/nested_dissection.c:114-134) with
// C(i,j) = (A_i^T, B_i^T), C(i+1,j) = (-I, 0):
//   S = F(i+1,p).Lambda <- A_i F(i,p).X + B_i F(i,p).U - F(i+1,p).X - S_old
// grid.x = np * nitems * 3 (column slots) when sel.col < 0, else np * nitems.
// ---------------------------------------------------------------------------------------------
__global__ void products_kernel(Shape s, Bufs b, Sel sel, int level) {
  int task = blockIdx.x, slot = 0, p;
  if (sel.col < 0) { slot = task % 3; task /= 3; }
  const int prob = sel.p0 + task / sel.nitems;
  const int leaf = sel.item0 + task % sel.nitems;
  p = sel.col < 0 ? active_col(s, leaf, level, slot) : sel.col;
  if (p < 0) return;
  const int i = rs_sep_index(leaf, level);
  const int n = s.n, m = s.m, nn = n * n;
  const double* kn = knot_ptr(s, b, prob, i);
  const double* A = kn + s.oA;
  const double* B = kn + s.oB;
  const double* F1 = fact_ptr(s, b, prob, i, p);
  double* F2 = fact_ptr(s, b, prob, i + 1, p);
  for (int e = threadIdx.x; e < nn; e += blockDim.x) {
    const int r = e % n, c = e / n;
    double acc = -F2[e];
    for (int l = 0; l < n; ++l) acc += A[r + n * l] * F1[nn + l + n * c];
    for (int l = 0; l < m; ++l) acc += B[r + n * l] * F1[2 * nn + l + m * c];
    acc -= F2[nn + e];
    F2[e] = acc;  // only thread e touches element e of S; F2.X is not written
  }
}

// rhs variant (reference src/solve.c:142-148): z_lambda(i+1) <- A_i zx(i) + B_i zu(i) - zx(i+1) - z_lambda(i+1)
__global__ void rhs_products_kernel(Shape s, Bufs b, Sel sel, int level) {
  const int gw = (blockIdx.x * blockDim.x + threadIdx.x);  // one thread per (task, row)
  const int n = s.n, m = s.m;
  const int task = gw / n, r = gw % n;
  if (task >= sel.np * sel.nitems) return;
  const int prob = sel.p0 + task / sel.nitems;
  const int leaf = sel.item0 + task % sel.nitems;
  const int i = rs_sep_index(leaf, level);
  const double* kn = knot_ptr(s, b, prob, i);
  const double* A = kn + s.oA;
  const double* B = kn + s.oB;
  const double* z1 = soln_ptr(s, b, prob, i);
  double* z2 = soln_ptr(s, b, prob, i + 1);
  double acc = -z2[r];
  for (int l = 0; l < n; ++l) acc += A[r + n * l] * z1[n + l];
  for (int l = 0; l < m; ++l) acc += B[r + n * l] * z1[2 * n + l];
  acc -= z2[n + r];
  z2[r] = acc;
}

// ---------------------------------------------------------------------------------------------
// separator Cholesky (reference src/solve.c:87-98 -> src/linalg_custom.c:88-111): in-place lower
// factor of the n x n SPD block F(index+1, level).Lambda. One CTA per (problem, leaf); the matrix is
// staged in shared memory; column j is finished by threads i >= j (left-looking, like the reference).
// A pivot <= 0 stops the factorization where the reference stops and raises the status word.
// dynamic smem: n*n doubles.
// ---------------------------------------------------------------------------------------------
__global__ void cholesky_kernel(Shape s, Bufs b, Sel sel, int level) {
  extern __shared__ double sm[];
  __shared__ int failed;
  const int prob = sel.p0 + blockIdx.x / sel.nitems;
  const int leaf = sel.item0 + blockIdx.x % sel.nitems;
  const int n = s.n, nn = n * n;
  double* M = fact_ptr(s, b, prob, rs_sep_index(leaf, level) + 1, level);
  for (int e = threadIdx.x; e < nn; e += blockDim.x) sm[e] = M[e];
  if (threadIdx.x == 0) failed = 0;
  __syncthreads();
  for (int j = 0; j < n; ++j) {
    for (int i = j + threadIdx.x; i < n; i += blockDim.x) {
      double v = sm[i + n * j];
      int k = 0;
      for (; k + 8 <= j; k += 8) {  // operands of 8 steps in flight; the subtractions keep the reference's order
        double a[8], c[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) { a[u] = sm[i + n * (k + u)]; c[u] = sm[j + n * (k + u)]; }
#pragma unroll
        for (int u = 0; u < 8; ++u) v -= a[u] * c[u];
      }
      for (; k < j; ++k) v -= sm[i + n * k] * sm[j + n * k];
      sm[i + n * j] = v;
    }
    __syncthreads();
    const double d = sm[j + n * j];
    if (!(d > 0)) {
      if (threadIdx.x == 0) failed = 1;
      break;  // uniform: every thread reads the same d
    }
    const double rt = sqrt(d);
    __syncthreads();
    for (int i = j + threadIdx.x; i < n; i += blockDim.x) sm[i + n * j] = sm[i + n * j] / rt;
    __syncthreads();
  }
  __syncthreads();
  for (int e = threadIdx.x; e < nn; e += blockDim.x) M[e] = sm[e];
  if (threadIdx.x == 0 && failed) {
    b.status[(size_t)prob * 3 * s.N + rs_slot_S(leaf, level, s.N)] = -1;
    b.flags[prob] = 1;
  }
}

// G (n x w, in shared memory, leading dim n) <- (L L^T)^{-1} G with all threads of the CTA.
// Same operation order as clap_LowerTriBackSub (reference src/linalg_custom.c:113-132).
__device__ __forceinline__ void block_chol_solve(const double* L, double* G, int n, int w) {
  for (int sweep = 0; sweep < 2; ++sweep) {
    for (int step = 0; step < n; ++step) {
      const int j = sweep ? n - 1 - step : step;
      const double piv = L[j + n * j];
      for (int c = threadIdx.x; c < w; c += blockDim.x) G[j + n * c] = G[j + n * c] / piv;
      __syncthreads();
      const int rows = sweep ? j : n - 1 - j;  // rows still coupled to x_j
      for (int e = threadIdx.x; e < rows * w; e += blockDim.x) {
        const int c = e / rows;
        const int i = sweep ? e % rows : j + 1 + e % rows;
        const double lij = sweep ? L[j + n * i] : L[i + n * j];
        G[i + n * c] -= lij * G[j + n * c];
      }
      __syncthreads();
    }
  }
}

// ---------------------------------------------------------------------------------------------
// solves: ndlqr_SolveCholeskyFactor (reference src/nested_dissection.c:136-152):
//   F(index+1,p).Lambda <- S^{-1} F(index+1,p).Lambda, p in {left, right boundary column}.
// grid.x = np*nitems*2 when sel.col < 0, else np*nitems. dynamic smem: 2*n*n doubles.
// ---------------------------------------------------------------------------------------------
__global__ void solves_kernel(Shape s, Bufs b, Sel sel, int level) {
  extern __shared__ double sm[];
  int task = blockIdx.x, slot = 0;
  if (sel.col < 0) { slot = 1 + task % 2; task /= 2; }
  const int prob = sel.p0 + task / sel.nitems;
  const int leaf = sel.item0 + task % sel.nitems;
  const int p = sel.col < 0 ? active_col(s, leaf, level, slot) : sel.col;
  if (p < 0) return;
  const int n = s.n, nn = n * n;
  const int i1 = rs_sep_index(leaf, level) + 1;
  const double* L = fact_ptr(s, b, prob, i1, level);
  double* G = fact_ptr(s, b, prob, i1, p);
  double* sL = sm;
  double* sG = sm + nn;
  for (int e = threadIdx.x; e < nn; e += blockDim.x) { sL[e] = L[e]; sG[e] = G[e]; }
  __syncthreads();
  block_chol_solve(sL, sG, n, n);
  for (int e = threadIdx.x; e < nn; e += blockDim.x) G[e] = sG[e];
}

// rhs variant (reference src/solve.c:152-170): z_lambda(index+1) <- S^{-1} z_lambda(index+1).
// dynamic smem: n*n + n doubles.
__global__ void rhs_solves_kernel(Shape s, Bufs b, Sel sel, int level) {
  extern __shared__ double sm[];
  const int prob = sel.p0 + blockIdx.x / sel.nitems;
  const int leaf = sel.item0 + blockIdx.x % sel.nitems;
  const int n = s.n, nn = n * n;
  const int i1 = rs_sep_index(leaf, level) + 1;
  const double* L = fact_ptr(s, b, prob, i1, level);
  double* z = soln_ptr(s, b, prob, i1);
  double* sL = sm;
  double* sz = sm + nn;
  for (int e = threadIdx.x; e < nn; e += blockDim.x) sL[e] = L[e];
  for (int e = threadIdx.x; e < n; e += blockDim.x) sz[e] = z[e];
  __syncthreads();
  block_chol_solve(sL, sz, n, 1);
  for (int e = threadIdx.x; e < n; e += blockDim.x) z[e] = sz[e];
}

// rhs variant for n <= 64 without CTA barriers: ONE WARP per separator. L is staged into the warp's shared memory with
// an odd leading dimension (column reads of the forward sweep and row reads of the backward sweep are both
// conflict-free), the right-hand side lives in registers (rows lane and lane + 32), the pivot element travels by
// shuffle. Same operation order per row as clap_LowerTriBackSub / clap_CholeskySolve (src/linalg_custom.c:113-150):
// subtract the products in increasing (forward) / decreasing (backward) column order, then scale by the pivot
// (multiplication by its reciprocal instead of the reference's division: at most one ulp apart per step).
// dynamic smem: warps * n * (n + 1) doubles.
__global__ void rhs_solves_warp_kernel(Shape s, Bufs b, Sel sel, int level) {
  extern __shared__ double sm[];
  const int wpc = blockDim.x >> 5, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long task = (long long)blockIdx.x * wpc + warp;
  if (task >= (long long)sel.np * sel.nitems) return;
  const int prob = sel.p0 + (int)(task / sel.nitems);
  const int leaf = sel.item0 + (int)(task % sel.nitems);
  const int n = s.n, ld = n + 1;
  const int i1 = rs_sep_index(leaf, level) + 1;
  const double* L = fact_ptr(s, b, prob, i1, level);
  double* z = soln_ptr(s, b, prob, i1);
  double* sL = sm + (size_t)warp * n * ld;
  const int r0 = lane, r1 = lane + 32;
  // column by column (coalesced global reads, no integer division), 8 columns of loads in flight
  for (int c0 = 0; c0 < n; c0 += 8) {
    double v0[8], v1[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int c = c0 + u;
      v0[u] = (c < n && r0 < n) ? __ldg(L + r0 + n * c) : 0.0;
      v1[u] = (c < n && r1 < n) ? __ldg(L + r1 + n * c) : 0.0;
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int c = c0 + u;
      if (c < n && r0 < n) sL[r0 + ld * c] = v0[u];
      if (c < n && r1 < n) sL[r1 + ld * c] = v1[u];
    }
  }
  double y0 = r0 < n ? z[r0] : 0.0, y1 = r1 < n ? z[r1] : 0.0;
  __syncwarp();
  // reciprocal pivots once per row (the division would otherwise sit on the dependent chain of every step)
  const double inv0 = r0 < n ? 1.0 / sL[r0 + ld * r0] : 0.0, inv1 = r1 < n ? 1.0 / sL[r1 + ld * r1] : 0.0;
  for (int j = 0; j < n; ++j) {  // L y = z
    const double yj = __shfl_sync(0xffffffffu, j < 32 ? y0 : y1, j & 31) * __shfl_sync(0xffffffffu, j < 32 ? inv0 : inv1, j & 31);
    if (r0 == j) y0 = yj;
    if (r1 == j) y1 = yj;
    if (r0 > j && r0 < n) y0 -= sL[r0 + ld * j] * yj;
    if (r1 > j && r1 < n) y1 -= sL[r1 + ld * j] * yj;
  }
  for (int j = n - 1; j >= 0; --j) {  // L^T x = y
    const double xj = __shfl_sync(0xffffffffu, j < 32 ? y0 : y1, j & 31) * __shfl_sync(0xffffffffu, j < 32 ? inv0 : inv1, j & 31);
    if (r0 == j) y0 = xj;
    if (r1 == j) y1 = xj;
    if (r0 < j) y0 -= sL[j + ld * r0] * xj;
    if (r1 < j) y1 -= sL[j + ld * r1] * xj;
  }
  if (r0 < n) z[r0] = y0;
  if (r1 < n) z[r1] = y1;
}

// ---------------------------------------------------------------------------------------------
// Schur update: ndlqr_UpdateShurFactor (reference src/nested_dissection.c:154-171):
//   F(k,p) -= F(k,level) * f,  f = F(index+1,p).Lambda, Lambda rows only if calc_lambda.
// One CTA per (problem, knot, boundary column); f is staged in shared memory (n*n doubles).
// grid.x = np*nitems*2 when sel.col < 0, else np*nitems.
// ---------------------------------------------------------------------------------------------
__global__ void schur_kernel(Shape s, Bufs b, Sel sel, int level) {
  extern __shared__ double sf[];
  int task = blockIdx.x, slot = 0;
  if (sel.col < 0) { slot = 1 + task % 2; task /= 2; }
  const int prob = sel.p0 + task / sel.nitems;
  const int k = sel.item0 + task % sel.nitems;
  const int n = s.n, m = s.m, nn = n * n;
  int p = sel.col;
  if (p < 0) {
    const int a = rs_subtree_first(k, level);
    p = slot == 1 ? rs_left_col(a) : rs_right_col(a + (2 << level) - 1, s.N);
    if (p < 0) return;
  }
  const int index = sel.index >= 0 ? sel.index : rs_sep_of_knot(k, level);
  const int calc = sel.calc_lambda >= 0 ? sel.calc_lambda : rs_calc_lambda(k, level);
  const double* f = fact_ptr(s, b, prob, index + 1, p);
  const double* Fk = fact_ptr(s, b, prob, k, level);
  double* g = fact_ptr(s, b, prob, k, p);
  for (int e = threadIdx.x; e < nn; e += blockDim.x) sf[e] = f[e];
  __syncthreads();
  const int first = calc ? 0 : nn;
  for (int e = first + threadIdx.x; e < s.fs; e += blockDim.x) {
    // element e lives in block Lambda / X (leading dim n) or U (leading dim m)
    int base, ld, r, c;
    if (e < 2 * nn) { base = e < nn ? 0 : nn; ld = n; }
    else { base = 2 * nn; ld = m; }
    r = (e - base) % ld;
    c = (e - base) / ld;
    double acc = g[e];
    for (int l = 0; l < n; ++l) acc -= Fk[base + r + ld * l] * sf[l + n * c];
    g[e] = acc;
  }
}

// rhs variant (reference src/solve.c:175-180): soln(k) -= F(k,level) * z_lambda(index+1).
// One thread per (problem, knot, row).
__global__ void rhs_schur_kernel(Shape s, Bufs b, Sel sel, int level) {
  const int gw = blockIdx.x * blockDim.x + threadIdx.x;
  const int task = gw / s.zs, e = gw % s.zs;
  if (task >= sel.np * sel.nitems) return;
  const int prob = sel.p0 + task / sel.nitems;
  const int k = sel.item0 + task % sel.nitems;
  const int n = s.n, m = s.m, nn = n * n;
  const int index = sel.index >= 0 ? sel.index : rs_sep_of_knot(k, level);
  const int calc = sel.calc_lambda >= 0 ? sel.calc_lambda : rs_calc_lambda(k, level);
  if (e < n && !calc) return;
  const double* zf = soln_ptr(s, b, prob, index + 1);  // lambda part = solved separator
  const double* Fk = fact_ptr(s, b, prob, k, level);
  double* z = soln_ptr(s, b, prob, k);
  int base, ld, r;
  if (e < 2 * n) { base = e < n ? 0 : nn; ld = n; r = e < n ? e : e - n; }
  else { base = 2 * nn; ld = m; r = e - 2 * n; }
  double acc = z[e];
  for (int l = 0; l < n; ++l) acc -= Fk[base + r + ld * l] * zf[l];
  z[e] = acc;
}

}  // namespace rslqr

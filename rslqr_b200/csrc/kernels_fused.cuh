// kernels_fused.cuh -- "tiered" device path: the bottom levels of the nested dissection stay on chip.
//
// The reference walks the binary tree level by level over ALL knots of a problem (reference
// src/solve.c:68-182), streaming the whole factor array through memory once per level. Here a CTA owns
// a group of G = 2^g consecutive knots and runs the leaf stage and the first g levels for that group out of
// shared memory (tier 0); the few remaining levels only ever need the first and the last knot of each group,
// so they run on compact "boundary entries" (tiers >= 1); a top-down pass turns the separator solves of those
// upper levels into final multipliers, and one streaming kernel applies them to every knot.
//
// State kept per knot k (replaces the K columns F(k,0..K-1) of the reference's NdData, src/nddata.h:83-93):
//   L(k), R(k) : (2n+m) x n blocks = the two structurally non-zero "boundary" columns of the subtree that
//                currently contains k: L couples to the separator on the left of the
//                subtree (lambda_a), R to the one on its right (lambda_{b+1}); rows = [Lambda | X | U].
//   z(k)       : the right-hand side / solution column (the rhs rides along as a 25th column instead of the
//                reference's second pass over `fact`, src/solve.c:137-182; the reference notes this option
//                itself, src/nddata.h:72-75).
// After the leaves L(k) = F(k, lev(k-1)) and R(k) = F(k, lev(k)). Merging the sibling subtrees [a,i] and
// [i+1,b] at level j (separator variable lambda_{i+1}) is, in the reference's terms:
//   ndlqr_FactorInnerProduct (src/nested_dissection.c:114):
//        S   = A_i X_R(i) + B_i U_R(i) - X_L(i+1)             (own column j)
//        S_L = A_i X_L(i) + B_i U_L(i)                         (column p_L; F(i+1,p_L) is still zero)
//        S_R = -X_R(i+1)                                        (column p_R; F(i,p_R) is still zero)
//        s   = A_i z_x(i) + B_i z_u(i) - z_x(i+1) - z_lambda(i+1)
//   Cholesky + ndlqr_SolveCholeskyFactor (src/solve.c:87-98, src/nested_dissection.c:136):
//        [f_L f_R zs] = S^{-1} [S_L S_R s]
//   ndlqr_UpdateShurFactor (src/nested_dissection.c:154) for every knot k of [a,b], own = R(k) if k <= i else L(k):
//        L(k) <- (k <= i ? L(k) : 0) - own f_L ;  R(k) <- (k <= i ? 0 : R(k)) - own f_R ;  z(k) -= own zs
//        (the Lambda rows the reference skips via ndlqr_ShouldCalcLambda are exactly the rows where own is zero)
//   and F(i+1,.).Lambda <- f_. : Lambda rows of knot i+1 become (f_L, f_R, zs).
// All arithmetic is IEEE fp64. Shapes are compile-time (n, m) so blocks live in registers / fragments.
#pragma once
#include <cuda_runtime.h>
#include <cstring>
#include <string>
#include "kernels_staged.cuh"

namespace rslqr {

inline size_t fused_ent_doubles(int n, int m) { return ((size_t)2 * (2 * n + m) * n + (2 * n + m) + n + 1) & ~(size_t)1; }

struct FusedBufs {
  double* stL;     // [batch][N*ZS][NS]   state after tier 0, row-major rows
  double* stR;     // [batch][N*ZS][NS]
  double* stz;     // [batch][N*ZS]
  double* ent[8];  // per tier t >= 1 input entries: [batch][M_{t-1}][ENT]
  double* sep;     // [batch][M0][(2NS+1)*NS]  upper separators: [f_L | f_R | zs], slot q = (i+1)/G0
  double* lamhat;  // [batch][M0+1][NS]        final multipliers of the upper separators (slots 0 and M0 = 0)
};

template <int NS, int NI>
struct Dim {
  static constexpr int ZS = 2 * NS + NI;         // rows per knot
  static constexpr int XU = NS + NI;             // X and U rows
  static constexpr int W = 2 * NS + 1;           // columns [L | R | z]
  static constexpr int WSZ = W * NS;             // one separator record [f_L | f_R | zs], column-major ld NS
  static constexpr int SLOT = ZS * NS;           // one block, row-major [ZS][NS]
  // boundary entry: L rows, R rows, z, lambda-rhs of the first knot; padded to an even count so that entries stay
  // 16-byte aligned for the TMA bulk copies
  static constexpr int ENT = (2 * SLOT + ZS + NS + 1) & ~1;
};

// ---- layout of the dual-form entries (kernels_dual.cuh; also written by scan_leaf_kernel) ----
template <int NS>
struct DualDim {
  static constexpr int NP = (NS + 3) & ~3;          // padded block size (multiple of 4: DMMA k-steps, 16-byte rows)
  static constexpr int LD = NP + 2;                 // row stride of a block (rows 16-byte aligned, bank-staggered)
  static constexpr int BLK = NP * LD;
  static constexpr int OK = 0, OC = BLK, OS = 2 * BLK, OKS = 3 * BLK, OSS = 3 * BLK + NP;
  static constexpr int ENT = 3 * BLK + 2 * NP;      // even -> entries stay 16-byte aligned
  static constexpr int WSZ = (2 * NS + 1) * NS;     // separator record [f_L | f_R | zs], column-major ld NS
  static constexpr int WP = (2 * NP + 1) * NP;      // the same in shared memory, ld NP
  static constexpr int LR = NP * (NP + 2);          // factor scratch of spine_core2<NP>
};

inline size_t dual_ent_doubles(int n) { const int np = (n + 3) & ~3; return (size_t)3 * np * (np + 2) + 2 * np; }

// ---- TMA 1-D bulk copies (cp.async.bulk) and the mbarrier that tracks them ----
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, int bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, int parity) {
  unsigned ok;
  do {
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                 : "=r"(ok)
                 : "r"(smem_u32(bar)), "r"(parity)
                 : "memory");
  } while (!ok);
}
// global -> shared, completion counted in bytes on `bar`; 16-byte aligned addresses, size a multiple of 16
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, int bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// shared -> global as part of the thread's current bulk group
__device__ __forceinline__ void bulk_s2g(void* dst, const void* src, int bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
// make generic-proxy writes to shared memory visible to the async proxy (TMA) before a bulk store
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ double fast_rcp(double d) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
  // one cubically convergent step, r (1 + e + e^2) with e = 1 - d r: the seed is good to 2^-19.9 (measured on B200 over 4 M
  // inputs), e^3 is below the rounding error of a double (measured max relative error 1.112e-16 against 1.110e-16 for two
  // Newton steps); three dependent FMAs instead of four (this sits on every pivot chain)
  const double e = fma(-d, r, 1.0);
  const double t = fma(e, e, e);
  r = fma(r, t, r);
  return r;
}

__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(d0), "+d"(d1)
               : "d"(a), "d"(b));
}

// st.shared.f64 predicated on `on` (a per-lane predicate the compiler must not turn into a branch)
__device__ __forceinline__ void sts_pred(unsigned saddr, double v, int on) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.s32 p, %2, 0;\n\t@p st.shared.f64 [%0], %1;\n\t}" ::"r"(saddr), "d"(v), "r"(on) : "memory");
}
__device__ __forceinline__ double shfl_d(double v, int src) { return __shfl_sync(0xffffffffu, v, src); }
// warp index of the calling thread as a value the compiler KNOWS to be warp-uniform (broadcast from lane 0): branches on it
// are then uniform branches, and the shuffles inside them need no WARPSYNC / ENDCOLLECTIVE wrappers (which also fence the
// instruction scheduler around every shuffle of the elimination sweeps)
__device__ __forceinline__ int warp_index_sync() { return __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0); }

// factor scratch of spine_core2: Lt row-major (row j = Lt(j, .)) with leading dimension NS + 2: rows stay 16-byte
// aligned for the 128-bit broadcast loads, and the 128-bit row stores of 8 neighbouring lanes (28 words apart for
// NS = 12) fall into distinct banks
template <int NS>
__host__ __device__ constexpr int factor_ld() { return NS + 2; }
template <int NS>
__host__ __device__ constexpr int factor_scratch_doubles() { return NS * factor_ld<NS>(); }

// ---------------------------------------------------------------------------------------------------------
// LDL^T sweep with the right-hand sides riding along as extra ROWS, then the backward sweep.
//   lanes [0,NS)    : acc = row i of S                                   (acc2 unused)
//   lanes [NS,2NS)  : acc = S_L(:,c), acc2 = S_R(:,c), c = lane - NS     (two right-hand sides per lane)
//   lane  2NS       : acc = s
// Eliminating an appended row y with the pivot rows of S leaves the multipliers y_j <- (D^-1 Lt^-1 y)_j behind: the
// forward substitution and the diagonal scaling of EVERY right-hand side come out of the same pivot-row shuffles the
// factorization needs anyway (no factor loads, no parking of the right-hand sides in shared memory). Only the
// backward sweep Lt^-T reads the factor (row-major Lr, 128-bit broadcast loads shared by both columns of a lane).
// On return Wout = [f_L | f_R | zs] (column-major, ld NS) and acc / acc2 still hold the lane's solved columns.
// Same S = Lt D Lt^T as clap_CholeskyFactorize (src/linalg_custom.c:88) up to the diagonal scaling.
// ---------------------------------------------------------------------------------------------------------
//   S0 (dual path, kernels_dual.cuh): the right-hand side s rides in acc2 of lane 0 instead of in acc of lane 2NS, so
//   that 2 NS lanes suffice (NS = 16 fits in one warp); zs then comes out of lane 0's acc2.
template <int NS, bool S0 = false>
__device__ __forceinline__ int spine_core2(double (&acc)[NS], double (&acc2)[NS], double* Lr, double* Wout) {
  static_assert(NS % 2 == 0 && 2 * NS + (S0 ? 0 : 1) <= 32, "pairs; one warp per separator");
  constexpr int LDF = factor_ld<NS>();
  const int lane = threadIdx.x & 31;
  int bad = 0;
#pragma unroll
  for (int j = 0; j < NS; ++j) {
    const double d = shfl_d(acc[j], j);
    if (!(d > 0.0)) bad = 1;
    const double dinv = fast_rcp(d);
    const double lij = acc[j] * dinv, lij2 = acc2[j] * dinv;
#pragma unroll
    for (int k = j + 1; k < NS; ++k) {
      const double pk = shfl_d(acc[k], j);  // pivot row element S'(j,k)
      acc[k] = fma(-lij, pk, acc[k]);
      acc2[k] = fma(-lij2, pk, acc2[k]);
    }
    acc[j] = lane == j ? dinv : lij;
    acc2[j] = lij2;
  }
  if (lane < NS) {
#pragma unroll
    for (int k = 0; k < NS; k += 2) *reinterpret_cast<double2*>(Lr + lane * LDF + k) = make_double2(acc[k], acc[k + 1]);
  }
  __syncwarp();
#pragma unroll
  for (int j = NS - 1; j > 0; --j) {
#pragma unroll
    for (int p = 0; p < j; p += 2) {
      const double2 l = *reinterpret_cast<const double2*>(Lr + j * LDF + p);
      acc[p] = fma(-l.x, acc[j], acc[p]);
      acc2[p] = fma(-l.x, acc2[j], acc2[p]);
      if (p + 1 < j) {
        acc[p + 1] = fma(-l.y, acc[j], acc[p + 1]);
        acc2[p + 1] = fma(-l.y, acc2[j], acc2[p + 1]);
      }
    }
  }
  // a lane's solved column is contiguous in W (column-major, ld NS): 128-bit stores (Wout is 16-byte aligned)
  if constexpr (S0) {
    if (lane >= NS && lane < 2 * NS) {
      const int c = lane - NS;
#pragma unroll
      for (int r = 0; r < NS; r += 2) {
        *reinterpret_cast<double2*>(Wout + c * NS + r) = make_double2(acc[r], acc[r + 1]);
        *reinterpret_cast<double2*>(Wout + (NS + c) * NS + r) = make_double2(acc2[r], acc2[r + 1]);
      }
    } else if (lane == 0) {
#pragma unroll
      for (int r = 0; r < NS; r += 2) *reinterpret_cast<double2*>(Wout + 2 * NS * NS + r) = make_double2(acc2[r], acc2[r + 1]);
    }
    __syncwarp();
    return bad;
  }
  if (lane >= NS && lane <= 2 * NS) {
    const int c = lane - NS;  // lane 2NS: c = NS -> column 2NS of W (zs)
    double* w1 = lane < 2 * NS ? Wout + c * NS : Wout + 2 * NS * NS;
#pragma unroll
    for (int r = 0; r < NS; r += 2) *reinterpret_cast<double2*>(w1 + r) = make_double2(acc[r], acc[r + 1]);
    if (lane < 2 * NS) {
#pragma unroll
      for (int r = 0; r < NS; r += 2) *reinterpret_cast<double2*>(Wout + (NS + c) * NS + r) = make_double2(acc2[r], acc2[r + 1]);
    }
  }
  __syncwarp();
  return bad;
}

// ---------------------------------------------------------------------------------------------------------
// Same sweep as spine_core2 with the pivot rows broadcast through shared memory instead of by shuffles: at step j
// lane j scales its row, U(j, k) = S'(j, k) / d_j = Lt(k, j), and stores [1/d_j | U(j, j+1..)] as row j of Ur (one
// lane, 128-bit stores: one wavefront each); every lane reads the row back with 128-bit broadcast loads (one
// wavefront per pair of elements, against two per element for a 64-bit shuffle) and eliminates with its own
// unscaled element: acc[k] -= acc[j] * U(j, k). Row j of Ur is exactly what the backward sweep x_j = y_j -
// sum_{k>j} Lt(k, j) x_k needs, so the factor is never stored a second time. LSU wavefronts per merge: 42 stores
// + 42 + 36 broadcast loads = 120 (spine_core2: 178 shuffle + 12 store + 36 load = 226). Same arithmetic, same
// rounding as spine_core2 (lij * pk = (acc[j] / d) * pk there, acc[j] * (pk / d) here: one rounding apart).
// ---------------------------------------------------------------------------------------------------------
template <int NS>
__device__ __forceinline__ int spine_core3(double (&acc)[NS], double (&acc2)[NS], double* Ur, double* Wout) {
  static_assert(NS % 2 == 0 && 2 * NS + 1 <= 32, "pairs; one warp per separator");
  constexpr int LDF = factor_ld<NS>();
  const int lane = threadIdx.x & 31;
  int bad = 0;
#pragma unroll
  for (int j = 0; j < NS; ++j) {
    const int j0 = j & ~1;  // first element of the 128-bit pair that holds 1/d_j
    if (lane == j) {
      const double d = acc[j];
      if (!(d > 0.0)) bad = 1;
      const double dinv = fast_rcp(d);
      double row[NS];
#pragma unroll
      for (int k = j0; k < NS; ++k) row[k] = k < j ? 0.0 : (k == j ? dinv : acc[k] * dinv);
#pragma unroll
      for (int k = j0; k < NS; k += 2) *reinterpret_cast<double2*>(Ur + j * LDF + k) = make_double2(row[k], row[k + 1]);
    }
    __syncwarp();
    const double m1 = acc[j], m2 = acc2[j];
    double dinv = 0.0;
#pragma unroll
    for (int k = j0; k < NS; k += 2) {
      const double2 u = *reinterpret_cast<const double2*>(Ur + j * LDF + k);
      if (k == j) dinv = u.x;
      else if (k > j) { acc[k] = fma(-m1, u.x, acc[k]); acc2[k] = fma(-m2, u.x, acc2[k]); }
      if (k + 1 == j) dinv = u.y;
      else if (k + 1 > j) { acc[k + 1] = fma(-m1, u.y, acc[k + 1]); acc2[k + 1] = fma(-m2, u.y, acc2[k + 1]); }
    }
    acc[j] = m1 * dinv;
    acc2[j] = m2 * dinv;
  }
  bad = __any_sync(0xffffffffu, bad);
  // backward sweep: x_j = y_j - sum_{k > j} U(j, k) x_k, the latest x last so that the dependent chain is one FMA per j
#pragma unroll
  for (int j = NS - 2; j >= 0; --j) {
    const int k0 = (j + 1) & ~1;
    double2 u[NS / 2];
#pragma unroll
    for (int k = k0; k < NS; k += 2) u[k / 2] = *reinterpret_cast<const double2*>(Ur + j * LDF + k);
#pragma unroll
    for (int k = NS - 1; k > j; --k) {
      const double uk = (k & 1) ? u[k / 2].y : u[k / 2].x;
      acc[j] = fma(-uk, acc[k], acc[j]);
      acc2[j] = fma(-uk, acc2[k], acc2[j]);
    }
  }
  if (lane >= NS && lane <= 2 * NS) {
    const int c = lane - NS;  // lane 2NS: c = NS -> column 2NS of W (zs)
    double* w1 = lane < 2 * NS ? Wout + c * NS : Wout + 2 * NS * NS;
#pragma unroll
    for (int r = 0; r < NS; r += 2) *reinterpret_cast<double2*>(w1 + r) = make_double2(acc[r], acc[r + 1]);
    if (lane < 2 * NS) {
#pragma unroll
      for (int r = 0; r < NS; r += 2) *reinterpret_cast<double2*>(Wout + (NS + c) * NS + r) = make_double2(acc2[r], acc2[r + 1]);
    }
  }
  __syncwarp();
  return bad;
}

#ifndef RSLQR_SPINE_SMEM
#define RSLQR_SPINE_SMEM 0
#endif
template <int NS>
__device__ __forceinline__ int spine_core(double (&acc)[NS], double (&acc2)[NS], double* Lr, double* Wout) {
#if RSLQR_SPINE_SMEM
  return spine_core3<NS>(acc, acc2, Lr, Wout);
#else
  return spine_core2<NS>(acc, acc2, Lr, Wout);
#endif
}

// ---------------------------------------------------------------------------------------------------------
// One separator, one warp: inner products -> factorization -> solves. Lane roles:
//   lanes [0,NS)      : column c of S  (then, for the solves, column c of S_R -> f_R)
//   lanes [NS,2NS)    : column c of S_L -> f_L
//   lane  2NS         : s -> zs
// S is symmetric, so "lane c holds column c" doubles as "lane c holds row c" for the row-per-lane
// factorization. The factorization is S = Lt D Lt^T (unit lower Lt): same factor as the reference's
// Cholesky (src/linalg_custom.c:88) up to the diagonal scaling L = Lt sqrt(D), without a square root or a
// division on the dependent chain; the solves are Lt^{-T} D^{-1} Lt^{-1}.
//   AB    : global, A_i (NS x NS column-major) followed by B_i (NS x NI column-major)
//   LXi/RXi : X rows of knot i in its L / R block (row-major, ld NS), the U rows follow directly
//   zxi   : z_x(i) followed by z_u(i)
//   LXn/RXn : X rows of knot i+1; zxn = z_x(i+1); zln = z_lambda(i+1)
//   Lsm   : scratch for Lt (F2: factor_scratch_doubles, 16-byte aligned, directly behind Wout; else NS*NS); Wout: [f_L | f_R | zs] column-major ld NS
//   F2    : the right-hand sides ride through the factorization (spine_core2) instead of being parked and solved after it
// Returns non-zero (warp-uniform) when a pivot is <= 0 (reference: clap_CholeskyFactorize fails).
// ---------------------------------------------------------------------------------------------------------
//   before_write : hook called after the tensor-core products and before the first store into Wout / Lsm (a caller
//           whose copy engine may still be reading those regions waits there instead of before the merge)
struct NoHook { __device__ __forceinline__ void operator()() const {} };
template <int NS, int NI, bool LDG = true, bool F2 = false, class Hook = NoHook>
__device__ __forceinline__ int spine_merge(const double* __restrict__ AB, const double* LXi, const double* RXi,
                                           const double* zxi, const double* LXn, const double* RXn,
                                           const double* zxn, const double* zln, double* Lsm, double* Wout,
                                           Hook before_write = Hook()) {
  static_assert(2 * NS + 1 <= 32, "one warp per separator needs 2n+1 <= 32");
  const int lane = threadIdx.x & 31;
  const int c = lane < NS ? lane : (lane < 2 * NS ? lane - NS : 0);
  auto ldab = [](const double* p) -> double {  // [A|B] lives in global memory (read-only path) or in shared memory
    if constexpr (LDG) return __ldg(p); else return *p;
  };
  const double* src;
  int ss;
  if (lane < NS) { src = RXi + c; ss = NS; }
  else if (lane < 2 * NS) { src = LXi + c; ss = NS; }
  else { src = zxi; ss = 1; }  // lanes > 2NS compute a harmless copy of the rhs column
  double acc[NS];
  constexpr bool MMA_P1 = ((NS + NI) % 4 == 0) && (NS > 8) && (NS <= 16) && (2 * NS + 1 > 24);
  if constexpr (MMA_P1) {
    // [A_i | B_i] (NS x (NS+NI), padded to 16 rows) times [X_R;U_R | X_L;U_L | z_x;z_u](i) ((NS+NI) x (2NS+1),
    // padded to 32 columns) on the fp64 tensor cores: 2 x 4 tiles, (NS+NI)/4 k-steps; the 8x8 result fragments
    // are dropped into Wout (column-major, ld NS) and picked up again one column per lane.
    // F2: the transpose buffer has an odd leading dimension (column reads of 16 lanes hit 16 distinct bank pairs)
    // and runs (2NS+1) doubles over the end of W: Wout must have (2NS+1)(NS+1) doubles (the overflow may be the head
    // of the Lt scratch, which is written only after the buffer has been read back).
    constexpr int KS = (NS + NI) / 4;
    constexpr int TLD = F2 ? NS + 1 : NS;
    const int q = lane & 3, rr = lane >> 2;
    double d[2][4][2];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) d[mt][nt][0] = d[mt][nt][1] = 0.0;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      const int l = 4 * ks + q;
      const double a0 = ldab(AB + l * NS + rr);
      const double a1 = (8 + rr < NS) ? ldab(AB + l * NS + 8 + rr) : 0.0;
      double bf[4];
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
        const int n = 8 * nt + rr;
        bf[nt] = n < NS ? RXi[l * NS + n] : (n < 2 * NS ? LXi[l * NS + n - NS] : (n == 2 * NS ? zxi[l] : 0.0));
      }
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
        dmma884(d[0][nt][0], d[0][nt][1], a0, bf[nt]);
        dmma884(d[1][nt][0], d[1][nt][1], a1, bf[nt]);
      }
    }
    before_write();
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int nt = 0; nt < 4; ++nt)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int row = 8 * mt + rr, col = 8 * nt + 2 * q + h;
          if (row < NS && col <= 2 * NS) Wout[col * TLD + row] = d[mt][nt][h];
        }
    __syncwarp();
    const int mc = lane <= 2 * NS ? lane : 2 * NS;
#pragma unroll
    for (int r = 0; r < NS; ++r) acc[r] = Wout[mc * TLD + r];
    __syncwarp();
  } else {
    before_write();
#pragma unroll
    for (int r = 0; r < NS; ++r) acc[r] = 0.0;
#pragma unroll 4
    for (int l = 0; l < NS + NI; ++l) {
      const double x = src[l * ss];
      const double* col = AB + l * NS;  // column l of [A | B] (B follows A, both have NS rows)
#pragma unroll
      for (int r = 0; r < NS; ++r) acc[r] = fma(ldab(col + r), x, acc[r]);
    }
  }
  if constexpr (F2) {
    // right-hand sides stay in registers: lanes [NS,2NS) carry S_L (acc) and S_R = -X_R(i+1) (acc2), lane 2NS carries s
    double acc2[NS];
#pragma unroll
    for (int r = 0; r < NS; ++r) {
      const double ln = LXn[r * NS + c], rn = RXn[r * NS + c], zz = zxn[r] + zln[r];
      acc[r] = lane < NS ? acc[r] - ln : (lane >= 2 * NS ? acc[r] - zz : acc[r]);
      acc2[r] = (lane >= NS && lane < 2 * NS) ? -rn : 0.0;
    }
    return spine_core<NS>(acc, acc2, Lsm, Wout);
  }
  const int mycol = lane <= 2 * NS ? lane : 2 * NS;
  if (lane < NS) {
#pragma unroll
    for (int r = 0; r < NS; ++r) acc[r] -= LXn[r * NS + c];
  } else {
    if (lane >= 2 * NS) {
#pragma unroll
      for (int r = 0; r < NS; ++r) acc[r] = acc[r] - zxn[r] - zln[r];
    }
    // park the rhs columns in Wout so that only one register array is live during the factorization
#pragma unroll
    for (int r = 0; r < NS; ++r) Wout[mycol * NS + r] = acc[r];
  }
  // ---- factorization, lane i < NS holds row i of S in acc[] (the other lanes run along on junk) ----
  int bad = 0;
#pragma unroll
  for (int j = 0; j < NS; ++j) {
    const double d = shfl_d(acc[j], j);
    if (!(d > 0.0)) bad = 1;
    const double dinv = fast_rcp(d);
    const double lij = acc[j] * dinv;  // Lt(i,j) for lane i > j
#pragma unroll
    for (int k = j + 1; k < NS; ++k) {
      const double pk = shfl_d(acc[k], j);  // pivot row element S'(j,k) = S'(k,j)
      acc[k] = fma(-lij, pk, acc[k]);
    }
    acc[j] = lane == j ? dinv : lij;  // the diagonal slot of Lt carries 1/D(j)
  }
  if (lane < NS) {
#pragma unroll
    for (int k = 0; k < NS; ++k) Lsm[lane * NS + k] = acc[k];
  }
  __syncwarp();
  // ---- solves: lanes < NS take the S_R columns, the others pick their parked column up again ----
  if (lane < NS) {
#pragma unroll
    for (int r = 0; r < NS; ++r) acc[r] = -RXn[r * NS + c];
  } else {
#pragma unroll
    for (int r = 0; r < NS; ++r) acc[r] = Wout[mycol * NS + r];
  }
#pragma unroll
  for (int j = 0; j < NS; ++j) {
#pragma unroll
    for (int i = j + 1; i < NS; ++i) acc[i] = fma(-Lsm[i * NS + j], acc[j], acc[i]);
  }
#pragma unroll
  for (int j = 0; j < NS; ++j) acc[j] *= Lsm[j * NS + j];
#pragma unroll
  for (int j = NS - 1; j >= 0; --j) {
#pragma unroll
    for (int i = 0; i < j; ++i) acc[i] = fma(-Lsm[j * NS + i], acc[j], acc[i]);
  }
  __syncwarp();
  if (lane <= 2 * NS) {
    const int wc = lane < NS ? NS + lane : (lane < 2 * NS ? lane - NS : 2 * NS);
#pragma unroll
    for (int r = 0; r < NS; ++r) Wout[wc * NS + r] = acc[r];
  }
  __syncwarp();
  return bad;
}

// ---------------------------------------------------------------------------------------------------------
// Schur update of a whole group at level j (ndlqr_UpdateShurFactor over every knot of the group).
// Two implementations with identical results up to summation order:
//   schur_dmma : fp64 tensor cores, mma.sync m8n8k4. A knot pair is 2*ZS rows = an integer number of 8-row
//                tiles (needs 2*ZS % 8 == 0, NS % 4 == 0); warp w owns knot pair w, so all of its tiles see the
//                same separator and the B fragments [-f_L | -f_R] are loaded once per level.
//   schur_dfma : plain DFMA, one thread per row (any shape).
// Ls/Rs: [G*ZS][NS] row-major, zv: [G*ZS], W: per separator [f_L | f_R | zs] column-major ld NS.
// ---------------------------------------------------------------------------------------------------------

template <int NS, int NI, bool LEVEL0>
__device__ __forceinline__ void schur_dmma_level(double* Ls, double* Rs, double* zv, const double* W, int j) {
  using D = Dim<NS, NI>;
  constexpr int ZS = D::ZS, KS = NS / 4, NT = 2 * NS / 8, TILES = 2 * ZS / 8;
  static_assert((2 * ZS) % 8 == 0 && NS % 4 == 0 && NS % 2 == 0, "tile shapes");
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q = lane & 3, rr = lane >> 2;
  const int s = warp >> j;  // separator (local) of knot pair `warp` at level j
  const double* Wf = W + (s << j) * D::WSZ;  // records are indexed by the first warp of the subtree
  double bneg[KS][NT], zq[KS];
#pragma unroll
  for (int ks = 0; ks < KS; ++ks) {
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) bneg[ks][nt] = -Wf[(8 * nt + rr) * NS + 4 * ks + q];
    zq[ks] = Wf[2 * NS * NS + 4 * ks + q];
  }
  // per-lane bases: row (8t + rr) of the pair, A fragment column q, C fragment columns 2q, 2q+1
  const int rowbase = (warp * 2 * ZS + rr) * NS;
  // side of the pair above level 0: 0 = left child (own = R block), 1 = right child (own = L block);
  // at level 0 the first knot of the pair is a left child and the second a right child
  const int pside = LEVEL0 ? 0 : (warp >> (j - 1)) & 1;
  // which block column pair (8nt + 2q, +1) lives in: L for n < NS
  bool inL[NT];
  int coff[NT];
#pragma unroll
  for (int nt = 0; nt < NT; ++nt) {
    const int n = 8 * nt + 2 * q;
    inL[nt] = n < NS;
    coff[nt] = inL[nt] ? n : n - NS;
  }
#pragma unroll
  for (int t = 0; t < TILES; ++t) {
    int side;
    if (LEVEL0) side = (8 * t + 7 < ZS) ? 0 : ((8 * t >= ZS) ? 1 : (8 * t + rr >= ZS ? 1 : 0));
    else side = pside;
    const double* own = (side ? Ls : Rs) + rowbase + t * 8 * NS + q;
    double a[KS];
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) a[ks] = own[4 * ks];
    double c[NT][2];
    double* dst[NT];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      dst[nt] = (inL[nt] ? Ls : Rs) + rowbase + t * 8 * NS + coff[nt];
      const bool isown = inL[nt] == (side == 1);
      if (isown) { c[nt][0] = 0.0; c[nt][1] = 0.0; }
      else { const double2 v = *reinterpret_cast<const double2*>(dst[nt]); c[nt][0] = v.x; c[nt][1] = v.y; }
    }
#pragma unroll
    for (int ks = 0; ks < KS; ++ks)
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) dmma884(c[nt][0], c[nt][1], a[ks], bneg[ks][nt]);
    double part = 0.0;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) part = fma(a[ks], zq[ks], part);
    part += __shfl_xor_sync(0xffffffffu, part, 1);
    part += __shfl_xor_sync(0xffffffffu, part, 2);
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) *reinterpret_cast<double2*>(dst[nt]) = make_double2(c[nt][0], c[nt][1]);
    if (q == 0) zv[warp * 2 * ZS + 8 * t + rr] -= part;
  }
  // F(i+1, .).Lambda <- (f_L, f_R, zs): the Lambda rows of local knot (2s+1) << j, owned by this warp if it
  // lies in its pair (own was zero there, so the update above left keep - 0 behind)
  const int ksep = (2 * s + 1) << j;
  if ((ksep >> 1) == warp) {
    __syncwarp();
    for (int e = lane; e < D::W * NS; e += 32) {
      const int col = e / NS, row = e - col * NS;
      const double v = Wf[e];
      if (col < NS) Ls[(ksep * ZS + row) * NS + col] = v;
      else if (col < 2 * NS) Rs[(ksep * ZS + row) * NS + col - NS] = v;
      else zv[ksep * ZS + row] = v;
    }
  }
}

template <int NS, int NI>
__device__ __forceinline__ void schur_dmma(double* Ls, double* Rs, double* zv, const double* W, int j) {
  if (j == 0) schur_dmma_level<NS, NI, true>(Ls, Rs, zv, W, 0);
  else schur_dmma_level<NS, NI, false>(Ls, Rs, zv, W, j);
}

// one row: [L R z](R) <- keep - own * [f_L f_R zs]; `side` selects own (0: R block, 1: L block)
template <int NS>
__device__ __forceinline__ void schur_row(double* Lrow, double* Rrow, double* zr, const double* Wf, int side,
                                          bool special, int row) {
  double o[NS];
  const double* own = side ? Lrow : Rrow;
#pragma unroll
  for (int k = 0; k < NS; ++k) o[k] = own[k];
#pragma unroll 2
  for (int c = 0; c < NS; ++c) {
    double aL = side ? 0.0 : Lrow[c];
    double aR = side ? Rrow[c] : 0.0;
#pragma unroll
    for (int k = 0; k < NS; ++k) {
      aL = fma(-o[k], Wf[c * NS + k], aL);
      aR = fma(-o[k], Wf[(NS + c) * NS + k], aR);
    }
    if (special) { aL = Wf[c * NS + row]; aR = Wf[(NS + c) * NS + row]; }
    Lrow[c] = aL;
    Rrow[c] = aR;
  }
  double az = *zr;
#pragma unroll
  for (int k = 0; k < NS; ++k) az = fma(-o[k], Wf[2 * NS * NS + k], az);
  if (special) az = Wf[2 * NS * NS + row];
  *zr = az;
}

template <int NS, int NI, int G>
__device__ __forceinline__ void schur_dfma(double* Ls, double* Rs, double* zv, const double* W, int j) {
  using D = Dim<NS, NI>;
  constexpr int ZS = D::ZS;
  for (int R = threadIdx.x; R < G * ZS; R += blockDim.x) {
    const int kn = R / ZS, row = R - kn * ZS;
    const int side = (kn >> j) & 1;
    const int s = kn >> (j + 1);
    const bool special = (kn == ((2 * s + 1) << j)) && (row < NS);
    schur_row<NS>(Ls + R * NS, Rs + R * NS, zv + R, W + (s << j) * D::WSZ, side, special, row);
  }
}

// ---------------------------------------------------------------------------------------------------------
// Lean Schur update of a merged boundary entry on the fp64 tensor cores (ndlqr_UpdateShurFactor restricted to the
// rows that stay boundary rows). Entry rows 0..NS-1 = X(a) come from entry sA (own = its R block, keep L), rows
// NS..ZS-1 = X(b), U(b) from entry sB (own = its L block, keep R):
//     [L R z](row) <- keep - own(row) * [f_L f_R zs]
// written to dA (rows < NS) / dB (rows >= NS), which may alias the sources. One warp; ZS rows = 4 tiles of 8
// (the last one half empty), 2NS = 24 columns = 3 tiles, NS = 12 = 3 k-steps: 36 DMMA.
// ---------------------------------------------------------------------------------------------------------
template <int NS, int NI>
__device__ __forceinline__ void lean_schur_dmma(const double* sA, const double* sB, double* dA, double* dB, const double* Wf) {
  using D = Dim<NS, NI>;
  constexpr int ZS = D::ZS, SLOT = D::SLOT, KS = NS / 4, NT = 2 * NS / 8, TILES = (ZS + 7) / 8;
  static_assert(NS % 4 == 0 && (2 * NS) % 8 == 0, "tile shapes");
  const int lane = threadIdx.x & 31, q = lane & 3, rr = lane >> 2;
  double bneg[KS][NT], zq[KS];
#pragma unroll
  for (int ks = 0; ks < KS; ++ks) {
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) bneg[ks][nt] = -Wf[(8 * nt + rr) * NS + 4 * ks + q];
    zq[ks] = Wf[2 * NS * NS + 4 * ks + q];
  }
#pragma unroll
  for (int t = 0; t < TILES; ++t) {
    const int row = 8 * t + rr;
    const bool valid = row < ZS;
    const int side = row < NS ? 0 : 1;
    const double* src = side ? sB : sA;
    double* dst = side ? dB : dA;
    const int rb = (valid ? row : 0) * NS;
    const double* own = src + (side ? 0 : SLOT) + rb + q;
    double a[KS];
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) a[ks] = valid ? own[4 * ks] : 0.0;
    double c[NT][2];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const int n = 8 * nt + 2 * q;
      const bool inL = n < NS;
      const bool keep = inL != (side == 1);  // side 0 keeps its L columns, side 1 its R columns
      if (keep) {
        const double2 v = *reinterpret_cast<const double2*>(src + (inL ? 0 : SLOT) + rb + (inL ? n : n - NS));
        c[nt][0] = v.x; c[nt][1] = v.y;
      } else { c[nt][0] = 0.0; c[nt][1] = 0.0; }
    }
    const double z0 = src[2 * SLOT + (valid ? row : 0)];
#pragma unroll
    for (int ks = 0; ks < KS; ++ks)
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) dmma884(c[nt][0], c[nt][1], a[ks], bneg[ks][nt]);
    double part = 0.0;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) part = fma(a[ks], zq[ks], part);
    part += __shfl_xor_sync(0xffffffffu, part, 1);
    part += __shfl_xor_sync(0xffffffffu, part, 2);
    if (valid) {
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        const int n = 8 * nt + 2 * q;
        const bool inL = n < NS;
        *reinterpret_cast<double2*>(dst + (inL ? 0 : SLOT) + rb + (inL ? n : n - NS)) = make_double2(c[nt][0], c[nt][1]);
      }
      if (q == 0) dst[2 * SLOT + row] = z0 - part;
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// Tier 0: one CTA = (problem, group of G consecutive knots). Leaves (ndlqr_SolveLeaf, reference
// src/nested_dissection.c:10-105, with the rhs built on the fly like src/solver.c:138-190 when pack_rhs != 0)
// and levels 0 .. log2(G)-1, all in shared memory. G/2 warps: warp w is the spine warp of separator w and
// owns knot pair w in the Schur update.
// Output: the group's state (L, R, z) to fb.stL/stR/stz and its boundary entry to fb.ent[0]; when the group
// is the whole problem (N == G) the solution goes straight to b.soln.
// dynamic smem: (2*G*SLOT + G*ZS + (G/2)*WSZ) doubles.
// ---------------------------------------------------------------------------------------------------------
template <int NS, int NI, int G>
constexpr int tier0_smem_doubles() {
  using D = Dim<NS, NI>;
  return 2 * G * D::SLOT + G * D::ZS + (G / 2) * D::WSZ;
}

template <int NS, int NI, int G, bool DMMA>
__global__ void __launch_bounds__(16 * G, 512 / (16 * G)) tier0_kernel(Shape s, Bufs b, FusedBufs fb, int pack_rhs, int M0) {
  using D = Dim<NS, NI>;
  constexpr int ZS = D::ZS, SLOT = D::SLOT, WSZ = D::WSZ;
  extern __shared__ __align__(16) double smf[];
  double* Ls = smf;
  double* Rs = Ls + G * SLOT;
  double* zv = Rs + G * SLOT;
  double* W = zv + G * ZS;
  const int tid = threadIdx.x, nthr = blockDim.x;
  const int warp = warp_index_sync(), lane = tid & 31;
  const int prob = blockIdx.x / M0, grp = blockIdx.x % M0;
  const int k0 = grp * G;
  const int N = s.N, Ng = s.Ng;

  // ---- leaves: L(k) = F(k, lev(k-1)), R(k) = F(k, lev(k)) ----
  // pass A: one TMA bulk copy brings the group's G flat knot blocks (contiguous in HBM) into the L area,
  // which is only needed (zeroed) after the leaves. [A_k | B_k] column-major is exactly [X rows | U rows] of
  // (A_k^T, B_k^T) row-major, so R(k) is a row-scaled copy of it.
  constexpr int VEC = 2 * NS + 2 * NI, AB = NS * (NS + NI);
  constexpr int STRIDE = VEC + 1 + AB + NS;          // doubles per flat knot block (reference src/lqr_data.c:25-36)
  constexpr int OA = VEC + 1, OD = OA + AB;
  static_assert(STRIDE <= SLOT && (G * STRIDE) % 2 == 0, "raw knot blocks are staged in the L area");
  static_assert(G * (NS + NI) + NS <= (G / 2) * WSZ, "reciprocals are staged in W");
  __shared__ __align__(8) unsigned long long mbar;
  double* raw = Ls;                    // [G][STRIDE]
  double* inv = W;                     // [G][NS+NI]  1/Q, 1/R
  double* prevd = W + G * (NS + NI);   // d_{k0-1} (or x0 for the first group)
  const double* kn0 = knot_ptr(s, b, prob, k0);
  if (tid == 0) {
    mbar_init(&mbar, 1);
    fence_mbar_init();
  }
  __syncthreads();
  if (tid == 0) {
    mbar_expect_tx(&mbar, G * STRIDE * (int)sizeof(double));
    bulk_g2s(raw, kn0, G * STRIDE * (int)sizeof(double), &mbar);
  }
  // lambda-rhs of the group's first knot: x0 (or, for a shard, d of the knot before it) travels in b.x0
  if (tid < NS) prevd[tid] = (k0 == 0) ? b.x0[(size_t)prob * NS + tid] : (kn0 - STRIDE)[OD + tid];
  for (int idx = tid; idx < G * NS * NS; idx += nthr) {
    const int kl = idx / (NS * NS), e = idx - kl * NS * NS;
    Rs[kl * SLOT + e] = 0.0;
  }
  const double* zin = soln_ptr(s, b, prob, k0);
  mbar_wait(&mbar, 0);
  __syncthreads();  // prevd
  // pass B: rhs of the leaves (reference src/solver.c:138-190 for the packing, src/nested_dissection.c:45-59,
  // 87-95 for the leaf solves) and 1/Q, 1/R with the leaf "Cholesky" status (Q_k, R_k are diagonal: the
  // factorization fails on an entry <= 0, reference src/linalg_custom.c:88)
  for (int idx = tid; idx < G * ZS; idx += nthr) {
    const int kl = idx / ZS, row = idx - kl * ZS;
    const int k = k0 + kl;
    const int kg = s.kofs + k;  // global knot index (shards)
    const double* sv = raw + kl * STRIDE;
    auto rhs = [&](int rw) -> double {  // entry rw of the negated rhs of knot k
      if (!pack_rhs) return zin[kl * ZS + rw];
      if (rw < NS) return -(kl == 0 ? prevd[rw] : sv[rw - STRIDE + OD]);
      if (rw < 2 * NS) return -sv[NS + NI + rw - NS];
      return (kg < Ng - 1) ? -sv[2 * NS + NI + rw - 2 * NS] : 0.0;
    };
    double v;
    if (kg == 0) {
      if (row < NS) v = -sv[row] * rhs(row) - rhs(NS + row);
      else if (row < 2 * NS) v = -rhs(row - NS);
      else v = rhs(row) / sv[NS + row - 2 * NS];
    } else {
      if (row < NS) v = rhs(row);
      else if (row < 2 * NS) v = rhs(row) / sv[row - NS];
      else v = (kg < Ng - 1) ? rhs(row) / sv[NS + row - 2 * NS] : rhs(row);
    }
    zv[idx] = v;
  }
  for (int idx = tid; idx < G * (NS + NI); idx += nthr) {
    const int kl = idx / (NS + NI), i = idx - kl * (NS + NI);
    const int k = k0 + kl;
    const int kg = s.kofs + k;  // global knot index (shards)
    const double d = raw[kl * STRIDE + i];
    if (!(d > 0.0) && !(i >= NS && kg == Ng - 1)) {
      b.status[(size_t)prob * 3 * N + (i < NS ? rs_slot_Q(k) : rs_slot_R(k))] = -1;
      b.flags[prob] = 1;
    }
    inv[idx] = 1.0 / d;
  }
  __syncthreads();
  // pass C: R(k): X = Q \ A^T, U = R \ B^T (nested_dissection.c:75-84); knot 0: Lambda = -A0^T, X = 0 (:31-36);
  // last knot: R(k) = 0
  for (int idx = tid; idx < G * AB; idx += nthr) {
    const int kl = idx / AB, e = idx - kl * AB;
    const int row = e / NS;  // 0..NS+NI-1: X rows then U rows, inv[] holds 1/Q then 1/R in the same order
    const int k = k0 + kl;
    const int kg = s.kofs + k;  // global knot index (shards)
    const double rawv = raw[kl * STRIDE + OA + e];
    double v = rawv * inv[kl * (NS + NI) + row];
    if (kg == Ng - 1) v = 0.0;
    if (kg == 0 && row < NS) { Rs[kl * SLOT + e] = -rawv; v = 0.0; }
    Rs[kl * SLOT + NS * NS + e] = v;
  }
  __syncthreads();
  // pass D: L(k): X = Q \ (-I) (:96-102), everything else zero (this also retires the raw staging)
  for (int idx = tid; idx < G * SLOT / 2; idx += nthr) reinterpret_cast<double2*>(Ls)[idx] = make_double2(0.0, 0.0);
  __syncthreads();
  for (int idx = tid; idx < G * NS; idx += nthr) {
    const int kl = idx / NS, i = idx - kl * NS;
    if (s.kofs + k0 + kl > 0) Ls[kl * SLOT + (NS + i) * NS + i] = -inv[kl * (NS + NI) + i];
  }

  // ---- levels ----
  // The level-j merge of subtree t only involves the 2^j warps that own its knot pairs, so (DMMA path, where a
  // warp owns a knot pair) the warps of a subtree meet on their own named barrier instead of the CTA barrier;
  // level 0 needs none at all (warp w runs the spine of separator w and then updates its own pair).
  // The spine warp of a subtree is its first warp, and W is indexed by that warp.
  int g = 0;
  while ((1 << g) < G) ++g;
  __syncthreads();
  int barid = 1;
  for (int j = 0; j < g; ++j) {
    const int t = warp >> j;                 // subtree of this warp at level j
    const int w0 = t << j;                   // its first warp = spine warp
    const int nsub = (G / 2) >> j;           // subtrees at this level
    const bool local = DMMA && (nsub > 1);   // barrier scope: subtree (named) or CTA
    if (j > 0) {
      if (local) bar_sync(barid + t, 32 << j); else __syncthreads();
    }
    if (warp == w0) {
      const int il = ((2 * t + 1) << j) - 1;  // local index of the separator knot i
      const double* AB = knot_ptr(s, b, prob, k0 + il) + s.oA;
      double* Lsm = Ls + (il + 1) * SLOT;  // Lambda rows of knot i+1: still zero, used as scratch
      const int bad = spine_merge<NS, NI>(AB, Ls + il * SLOT + NS * NS, Rs + il * SLOT + NS * NS, zv + il * ZS + NS,
                                          Ls + (il + 1) * SLOT + NS * NS, Rs + (il + 1) * SLOT + NS * NS,
                                          zv + (il + 1) * ZS + NS, zv + (il + 1) * ZS, Lsm, W + w0 * WSZ);
      for (int e = lane; e < NS * NS; e += 32) Lsm[e] = 0.0;
      if (bad && lane == 0) {
        const int i = k0 + il;
        b.status[(size_t)prob * 3 * N + rs_slot_S(i >> (j + 1), j, N)] = -1;
        b.flags[prob] = 1;
      }
    }
    if (j == 0 && DMMA) __syncwarp();
    else if (local) bar_sync(barid + t, 32 << j);
    else __syncthreads();
    if (j > 0) barid += nsub;  // ids are unique per (level, subtree): a warp may run ahead into the next level
    if constexpr (DMMA) schur_dmma<NS, NI>(Ls, Rs, zv, W, j);
    else schur_dfma<NS, NI, G>(Ls, Rs, zv, W, j);
  }
  __syncthreads();

  // ---- output: TMA bulk stores straight out of shared memory ----
  if (M0 == 1) {
    double* out = soln_ptr(s, b, prob, 0);
    for (int idx = tid; idx < G * ZS; idx += nthr) out[idx] = zv[idx];
    return;
  }
  {
    // boundary entry: rows [X(first knot) | X,U(last knot)] of L, R, z and the lambda-rhs of the first knot,
    // gathered into W (free after the last level)
    static_assert(D::ENT <= (G / 2) * WSZ, "entry is staged in W");
    double* en = W;
    for (int idx = tid; idx < SLOT; idx += nthr) {
      const int er = idx / NS, c = idx - er * NS;
      const int lr = er < NS ? NS + er : (G - 1) * ZS + er;
      en[idx] = Ls[lr * NS + c];
      en[SLOT + idx] = Rs[lr * NS + c];
    }
    for (int er = tid; er < ZS; er += nthr) en[2 * SLOT + er] = zv[er < NS ? NS + er : (G - 1) * ZS + er];
    for (int c = tid; c < NS; c += nthr) en[2 * SLOT + ZS + c] = zv[c];
    fence_async_smem();
    __syncthreads();
    if (tid == 0) {
      const size_t base = ((size_t)prob * N + k0) * ZS;
      bulk_s2g(fb.stL + base * NS, Ls, G * SLOT * (int)sizeof(double));
      bulk_s2g(fb.stR + base * NS, Rs, G * SLOT * (int)sizeof(double));
      bulk_s2g(fb.stz + base, zv, G * ZS * (int)sizeof(double));
      bulk_s2g(fb.ent[0] + ((size_t)prob * M0 + grp) * D::ENT, en, D::ENT * (int)sizeof(double));
      bulk_commit();
      bulk_wait_read0();  // shared memory must stay valid until the copy engine has read it
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// Tiers >= 1: one CTA = (problem, `ecount` consecutive boundary entries); runs log2(ecount) more levels.
// An entry stands for a subtree [a,b] already merged below: rows [X(a) | X(b) | U(b)] of L, R and z (the only
// rows a later ndlqr_FactorInnerProduct reads) plus z_lambda(a). Merging entries e_l (left subtree ends there)
// and e_l+1: spine_merge on knot i = last knot of e_l, then the Schur update of just the rows that stay
// boundary rows: X(a) of the left subtree's first entry (own = R) and X,U(b) of the right subtree's last entry
// (own = L). The separator record [f_L | f_R | zs] goes to fb.sep for the top-down pass.
// `span` = knots per input entry; warps = max(1, ecount/2).
// dynamic smem: (ecount*ENT + warps*(WSZ + factor_scratch_doubles<NS>())) doubles.
// ---------------------------------------------------------------------------------------------------------
// Where a tier reads and writes; lets the same kernel run the local tiers of a problem, the last local tier of
// a shard (merged entry goes to the exchange payload) and the cross-GPU tier (entries, [A|B] blocks and records
// live in the gathered payload buffer).
struct TierArgs {
  const double* in;        // entry e of problem p: in + (p*Min + e)*in_stride
  long long in_stride;
  const double* ab;        // [A|B] of the last knot of entry e: ab + p*ab_pstride + e*ab_stride
  long long ab_stride, ab_pstride;
  double* sep;             // record of the separator between entries e and e+1: sep + p*sep_pstride + (e+1)*sep_mul*WSZ
  long long sep_pstride;
  int sep_mul;
  double* out;             // merged entry of CTA c of problem p: out + (p*ncta + c)*out_stride; nullptr = none
  long long out_stride;
  int Min, ecount;
  int span;                // knots per input entry (status words only; 0 = cross-GPU tier, flags only)
  // cross-GPU tier fed by peer stores over NVLink (shard_push_kernel): entry e is complete once wait_flags[e] >= wait_epoch
  const unsigned long long* wait_flags;  // nullptr = inputs are already complete (stream order / NCCL)
  unsigned long long wait_epoch;
};

// ---- peer-to-peer exchange of the shard payloads (one long horizon split over GPUs) ----
struct PeerPtrs {
  double* gathered[8];             // rank r's gathered buffer (payload slots of all ranks), mapped into this process
  unsigned long long* flags[8];    // rank r's arrival flags, one per source rank
};

__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// CTA r stores this rank's payload into rank r's gathered buffer (slot `rank`) with plain peer stores over NVLink,
// then publishes it: system-scope fence, release store of the epoch into rank r's flag word for this source.
__global__ void shard_push_kernel(const double* src, PeerPtrs peers, int rank, int pay, unsigned long long epoch) {
  pdl_sync();
  const int r = blockIdx.x;
  double* dst = peers.gathered[r] + (size_t)rank * pay;
  if (dst != src) {
    const double2* s2 = reinterpret_cast<const double2*>(src);
    double2* d2 = reinterpret_cast<double2*>(dst);
    for (int i = threadIdx.x; i < pay / 2; i += blockDim.x) d2[i] = s2[i];
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) st_release_sys(peers.flags[r] + rank, epoch);
}

template <int NS>
__host__ __device__ constexpr int tier_warp_scratch() { return ((2 * NS + 1) * (NS + 1) + 1) & ~1; }

template <int NS, int NI>
__global__ void __launch_bounds__(128, 4) tier_kernel(Shape s, Bufs b, TierArgs a) {
  pdl_sync();
  using D = Dim<NS, NI>;
  constexpr int ZS = D::ZS, SLOT = D::SLOT, WSZ = D::WSZ, ENT = D::ENT;
  static_assert(ZS <= 32, "lean Schur update maps one row per lane");
  extern __shared__ __align__(16) double smf[];
  double* sm = smf;
  const int tid = threadIdx.x, nthr = blockDim.x;
  const int warp = warp_index_sync(), lane = tid & 31;
  const int ecount = a.ecount;
  const int ncta = a.Min / ecount;
  const int prob = blockIdx.x / ncta, sg = blockIdx.x % ncta;
  double* ents = sm;
  // per warp: W plus the overflow of the transpose buffer of spine_merge (F2). The Lt scratch needs no room of its
  // own: rows X(b), U(b) of the LEFT entry of a merge are dead once the inner products are formed.
  constexpr int WSCR = tier_warp_scratch<NS>();
  static_assert(factor_scratch_doubles<NS>() <= (ZS - NS) * NS, "Lt scratch fits in the dead rows of the left entry");
  double* Wall = ents + ecount * ENT;
  const double* in = a.in + ((size_t)prob * a.Min + (size_t)sg * ecount) * a.in_stride;
  __shared__ __align__(8) unsigned long long mbar;
  if (tid == 0) {
    mbar_init(&mbar, 1);
    fence_mbar_init();
  }
  __syncthreads();
  if (a.wait_flags) {  // entries arrive by peer stores: wait for their flags (bounded: a lost peer must not hang the GPU)
    if (tid < a.Min) {
      const unsigned long long t0 = global_ns();
      while (ld_acquire_sys(a.wait_flags + tid) < a.wait_epoch) {
        if (global_ns() - t0 > 4000000000ull) { b.flags[prob] = 2; break; }
        __nanosleep(64);
      }
    }
    __syncthreads();
    asm volatile("fence.proxy.async;" ::: "memory");  // the TMA (async proxy) reads what the acquire made visible
  }
  if (tid == 0) {  // TMA bulk copies (one when the CTA's entries are contiguous in HBM)
    mbar_expect_tx(&mbar, ecount * ENT * (int)sizeof(double));
    if (a.in_stride == ENT) bulk_g2s(ents, in, ecount * ENT * (int)sizeof(double), &mbar);
    else
      for (int e = 0; e < ecount; ++e) bulk_g2s(ents + e * ENT, in + e * a.in_stride, ENT * (int)sizeof(double), &mbar);
  }
  mbar_wait(&mbar, 0);
  int levels = 0;
  while ((1 << levels) < ecount) ++levels;
  for (int l = 0; l < levels; ++l) {
    __syncthreads();
    const int nsep = ecount >> (l + 1);
    if (warp < nsep) {
      const int el = ((2 * warp + 1) << l) - 1, er = el + 1;
      const int ea = (2 * warp) << l, eb = er + (1 << l) - 1;
      const int ge = sg * ecount + el;  // entry index within the problem; the separator sits right after it
      const double* AB = a.ab + (size_t)prob * a.ab_pstride + (size_t)ge * a.ab_stride;
      double* El = ents + el * ENT;
      double* Er = ents + er * ENT;
      double* Wf = Wall + warp * WSCR;
      const int bad = spine_merge<NS, NI, true, true>(AB, El + NS * NS, El + SLOT + NS * NS, El + 2 * SLOT + NS, Er, Er + SLOT,
                                          Er + 2 * SLOT, Er + 2 * SLOT + ZS, El + NS * NS, Wf);
      double* rec = a.sep + (size_t)prob * a.sep_pstride + (size_t)(ge + 1) * a.sep_mul * WSZ;
      for (int e = lane; e < WSZ; e += 32) rec[e] = Wf[e];
      if (bad && lane == 0) {
        if (a.span > 0) {
          const int i = (ge + 1) * a.span - 1;  // local knot index of the separator
          int lev = 0;
          while ((i >> lev) & 1) ++lev;
          b.status[(size_t)prob * 3 * s.N + rs_slot_S(i >> (lev + 1), lev, s.N)] = -1;
        }
        b.flags[prob] = 1;
      }
      if constexpr (NS % 4 == 0) {
        lean_schur_dmma<NS, NI>(ents + ea * ENT, ents + eb * ENT, ents + ea * ENT, ents + eb * ENT, Wf);
      } else if (lane < ZS) {
        double* E = ents + (lane < NS ? ea : eb) * ENT;
        schur_row<NS>(E + lane * NS, E + SLOT + lane * NS, E + 2 * SLOT + lane, Wf, lane < NS ? 0 : 1, false, 0);
      }
    }
  }
  __syncthreads();
  if (a.out) {  // hand the merged entry up: rows X(a) from the first entry, X,U(b) from the last
    double* out = a.out + ((size_t)prob * ncta + sg) * a.out_stride;
    const double* E0 = ents;
    const double* E1 = ents + (ecount - 1) * ENT;
    for (int idx = tid; idx < SLOT; idx += nthr) {
      const double* E = idx < NS * NS ? E0 : E1;
      out[idx] = E[idx];
      out[SLOT + idx] = E[SLOT + idx];
    }
    for (int r = tid; r < ZS; r += nthr) out[2 * SLOT + r] = (r < NS ? E0 : E1)[2 * SLOT + r];
    for (int c = tid; c < NS; c += nthr) out[2 * SLOT + ZS + c] = E0[2 * SLOT + ZS + c];
  }
}

// ---------------------------------------------------------------------------------------------------------
// Top-down: final multiplier of every upper separator, from the top level down:
//   lambda(q) = zs(q) - f_L(q) lambda(q - h) - f_R(q) lambda(q + h),  h = 2^ctz(q)
// (this is what the reference's remaining Schur updates do to the Lambda rows of knot i+1, telescoped).
// Slots 0 and M0 of fb.lamhat hold zeros (no separator outside the horizon). One CTA per problem.
// ---------------------------------------------------------------------------------------------------------
template <int NS>
__device__ __forceinline__ void topdown_levels(const double* sep_all, double* lam_all, int M0, const double* left, const double* right,
                                               int span, int step, int cta) {
  // left/right: multipliers of the separators bounding the range (a shard's neighbours), nullptr = none.
  // A CTA covers the `span` slots after slot blk*span and resolves the separators at multiples of `step` inside it
  // (coarsest first); the multipliers at both ends of its range are already final, or (span == M0) set here.
  // Long horizons run two passes: (span = M0, step = H) with one CTA, then (span = H, step = 1) with M0/H CTAs.
  constexpr int WSZ = (2 * NS + 1) * NS;
  const int nblk = M0 / span;
  const int prob = cta / nblk, base = (cta % nblk) * span;
  double* lam = lam_all + (size_t)prob * (M0 + 1) * NS;
  const double* sep = sep_all + (size_t)prob * M0 * WSZ;
  if (span == M0 && threadIdx.x < NS) {
    lam[threadIdx.x] = left ? left[(size_t)prob * NS + threadIdx.x] : 0.0;
    lam[M0 * NS + threadIdx.x] = right ? right[(size_t)prob * NS + threadIdx.x] : 0.0;
  }
  int T = 0, T0 = 0;
  while ((1 << T) < span) ++T;
  while ((1 << T0) < step) ++T0;
  for (int t = T - 1; t >= T0; --t) {
    __syncthreads();
    const int cnt = span >> (t + 1), h = 1 << t;
    for (int idx = threadIdx.x; idx < cnt * NS; idx += blockDim.x) {
      const int u = idx / NS, r = idx - u * NS;
      const int q = base + ((2 * u + 1) << t);
      const double* rec = sep + (size_t)q * WSZ;
      const double* la = lam + (q - h) * NS;
      const double* lb = lam + (q + h) * NS;
      double v = rec[2 * NS * NS + r];
#pragma unroll
      for (int c = 0; c < NS; ++c) v = fma(-rec[c * NS + r], la[c], fma(-rec[(NS + c) * NS + r], lb[c], v));
      lam[q * NS + r] = v;
    }
  }
}

template <int NS>
__global__ void topdown_kernel(const double* sep_all, double* lam_all, int M0, const double* left, const double* right,
                               int span, int step) {
  pdl_sync();
  topdown_levels<NS>(sep_all, lam_all, M0, left, right, span, step, (int)blockIdx.x);
}

// Latency variant of topdown_kernel for few CTAs (one long horizon): the records of the CTA's range (at most 63) come
// into shared memory by TMA bulk copies up front and the multipliers of the range live in shared memory too, so the
// log2(span/step) dependent levels cost shared-memory round trips instead of L2 / HBM ones. Same arithmetic, same
// order. dynamic smem: ((span/step - 1) * WSZ + (span/step + 1) * NS) doubles.
template <int NS>
__global__ void __launch_bounds__(128) topdown_smem_kernel(const double* sep_all, double* lam_all, int M0, const double* left,
                                                            const double* right, int span, int step) {
  pdl_sync();
  constexpr int WSZ = (2 * NS + 1) * NS;
  extern __shared__ __align__(16) double tsm[];
  __shared__ __align__(8) unsigned long long mbar;
  const int nblk = M0 / span, nloc = span / step;
  const int prob = blockIdx.x / nblk, base = (blockIdx.x % nblk) * span;
  double* lam = lam_all + (size_t)prob * (M0 + 1) * NS;
  const double* sep = sep_all + (size_t)prob * M0 * WSZ;
  double* recs = tsm;                               // record of local separator i (1 .. nloc-1) at recs + (i-1)*WSZ
  double* lams = tsm + (size_t)(nloc - 1) * WSZ;    // multipliers of local slots 0 .. nloc
  if (threadIdx.x == 0) {
    mbar_init(&mbar, 1);
    fence_mbar_init();
    mbar_expect_tx(&mbar, (nloc - 1) * WSZ * (int)sizeof(double));
    if (step == 1) bulk_g2s(recs, sep + (size_t)(base + 1) * WSZ, (nloc - 1) * WSZ * (int)sizeof(double), &mbar);
    else
      for (int i = 1; i < nloc; ++i)
        bulk_g2s(recs + (size_t)(i - 1) * WSZ, sep + (size_t)(base + i * step) * WSZ, WSZ * (int)sizeof(double), &mbar);
  }
  if (threadIdx.x < NS) {
    const int r = threadIdx.x;
    double la, lb;
    if (span == M0) {
      la = left ? left[(size_t)prob * NS + r] : 0.0;
      lb = right ? right[(size_t)prob * NS + r] : 0.0;
      lam[r] = la;
      lam[M0 * NS + r] = lb;
    } else {
      la = lam[base * NS + r];
      lb = lam[(base + span) * NS + r];
    }
    lams[r] = la;
    lams[nloc * NS + r] = lb;
  }
  __syncthreads();
  mbar_wait(&mbar, 0);
  int T = 0;
  while ((1 << T) < nloc) ++T;
  for (int t = T - 1; t >= 0; --t) {
    const int cnt = nloc >> (t + 1), h = 1 << t;
    for (int idx = threadIdx.x; idx < cnt * NS; idx += blockDim.x) {
      const int u = idx / NS, r = idx - u * NS;
      const int i = (2 * u + 1) << t;
      const double* rec = recs + (size_t)(i - 1) * WSZ;
      const double* la = lams + (i - h) * NS;
      const double* lb = lams + (i + h) * NS;
      double v = rec[2 * NS * NS + r];
#pragma unroll
      for (int c = 0; c < NS; ++c) v = fma(-rec[c * NS + r], la[c], fma(-rec[(NS + c) * NS + r], lb[c], v));
      lams[i * NS + r] = v;
    }
    __syncthreads();
  }
  for (int idx = threadIdx.x; idx < (nloc - 1) * NS; idx += blockDim.x) {
    const int i = 1 + idx / NS, r = idx % NS;
    lam[(size_t)(base + i * step) * NS + r] = lams[i * NS + r];
  }
}

// host: one pass for short ranges, two for long ones (the serial chain of one CTA is log2(M0) dependent levels). With
// few CTAs (a single long horizon) the levels run out of shared memory (topdown_smem_kernel); a full batch keeps the
// streaming kernel, which has many CTAs per SM in flight.
template <int NS>
inline int launch_topdown(const double* sep, double* lam, int M0, int batch, const double* left, const double* right,
                          cudaStream_t st) {
  constexpr int H = 64, WSZ = (2 * NS + 1) * NS;
  static const int force = [] { const char* v = getenv("RSLQR_TOPDOWN_SMEM"); return v ? atoi(v) : -1; }();
  auto pass = [&](int span, int step) {
    const int nloc = span / step, ctas = batch * (M0 / span);
    const int bytes = ((nloc - 1) * WSZ + (nloc + 1) * NS) * (int)sizeof(double);
    // (63 records of n = 16 would not fit; odd n: the records are only 8-byte aligned, no bulk copies)
    const bool smem_ok = nloc >= 2 && nloc <= H && bytes <= 200 * 1024 && WSZ % 2 == 0;
    const bool use_smem = smem_ok && (force >= 0 ? force != 0 : ctas <= 1024);
    if (use_smem) {
      smem_optin(topdown_smem_kernel<NS>, bytes);
      launch_k(topdown_smem_kernel<NS>, dim3(ctas), dim3(128), (size_t)bytes, st, sep, lam, M0, left, right, span, step);
    } else {
      launch_k(topdown_kernel<NS>, dim3(ctas), dim3(128), (size_t)0, st, sep, lam, M0, left, right, span, step);
    }
  };
  if (M0 <= H) {
    pass(M0, 1);
    return 1;
  }
  if (M0 <= 2 * H && batch * 1 > 1024) {  // (throughput regime: one streaming pass, as before)
    pass(M0, 1);
    return 1;
  }
  pass(M0, H);
  pass(H, 1);
  return 2;
}

// ---------------------------------------------------------------------------------------------------------
// Finalize (the rhs propagation of all upper levels in one pass, reference src/solve.c:175-180):
//   soln(k) = z(k) - L(k) lambda_a - R(k) lambda_{b+1},  [a,b] = the tier-0 group of k;
// the Lambda rows of the first knot of a group are the separator multiplier itself. One thread per row.
// ---------------------------------------------------------------------------------------------------------
template <int NS, int NI>
__global__ void finalize_kernel(Shape s, Bufs b, FusedBufs fb, int G0, int M0, long long rows_total) {
  constexpr int ZS = 2 * NS + NI;
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= rows_total) return;
  const long long per = (long long)s.N * ZS;
  const int prob = (int)(gid / per);
  const int R = (int)(gid - prob * per);
  const int k = R / ZS, row = R - k * ZS;
  const int grp = k / G0;
  const double* la = fb.lamhat + ((size_t)prob * (M0 + 1) + grp) * NS;
  const double* lb = la + NS;
  const double2* pl = reinterpret_cast<const double2*>(fb.stL + gid * NS);
  const double2* pr = reinterpret_cast<const double2*>(fb.stR + gid * NS);
  double v = fb.stz[gid];
#pragma unroll
  for (int c = 0; c < NS / 2; ++c) {
    const double2 l2 = pl[c], r2 = pr[c];
    v = fma(-l2.x, la[2 * c], v);
    v = fma(-l2.y, la[2 * c + 1], v);
    v = fma(-r2.x, lb[2 * c], v);
    v = fma(-r2.y, lb[2 * c + 1], v);
  }
  if (row < NS && s.kofs + k > 0 && k == grp * G0) v = la[row];
  b.soln[gid] = v;
}

// ---------------------------------------------------------------------------------------------------------
// host side: plan, buffers, launches
// ---------------------------------------------------------------------------------------------------------
struct FusedTier { int Min, ecount, span; };
struct FusedPlan {
  bool available;
  int ns, ni, G0, M0, ntiers, use_dmma;
  FusedTier tiers[8];
  FusedBufs fb;
  double* xsep;   // [8][WSZ]  records of the cross-GPU separators (shards)
  double* xlam;   // [9][NS]   their final multipliers, slots 0 and G zero
  size_t bytes;
  int launches;
};

inline bool fused_shape_ok(int n, int m) { return (n == 12 && m == 4) || (n == 6 && m == 3); }

// Decides the tiling for (n, m, N) and allocates the device buffers of the tiered path. Leaves
// plan->available = false (and no allocation) when the shape is not covered; the caller then uses the staged path.
inline int fused_plan(const Shape& s, int batch, int g0_pref, FusedPlan* p, std::string& err) {
  memset(p, 0, sizeof(*p));
  if (!fused_shape_ok(s.n, s.m) || s.N < 8) return 0;
  p->ns = s.n; p->ni = s.m;
  p->G0 = (g0_pref == 16 && s.N >= 16) ? 16 : 8;
  p->M0 = s.N / p->G0;
  p->use_dmma = (s.n == 12);
  int M = p->M0, span = p->G0;
  while (M > 1) {
    if (p->ntiers == 8) return 0;
    const int e = M < 8 ? M : 8;
    p->tiers[p->ntiers++] = FusedTier{M, e, span};
    M /= e; span *= e;
  }
  if (p->M0 > 1) {
    const size_t zs = s.zs, n = s.n;
    const size_t ent = fused_ent_doubles(s.n, s.m), wsz = (2 * n + 1) * n;
    auto alloc = [&](double** ptr, size_t cnt) {
      p->bytes += cnt * sizeof(double);
      return cudaMalloc((void**)ptr, cnt * sizeof(double)) == cudaSuccess;
    };
    bool ok = alloc(&p->fb.stL, (size_t)batch * s.N * zs * n) && alloc(&p->fb.stR, (size_t)batch * s.N * zs * n) &&
              alloc(&p->fb.stz, (size_t)batch * s.N * zs) && alloc(&p->fb.sep, (size_t)batch * p->M0 * wsz) &&
              alloc(&p->fb.lamhat, (size_t)batch * (p->M0 + 1) * n);
    for (int t = 0; ok && t < p->ntiers; ++t) ok = alloc(&p->fb.ent[t], (size_t)batch * p->tiers[t].Min * ent);
    ok = ok && alloc(&p->xsep, 8 * wsz) && alloc(&p->xlam, 9 * n);
    if (!ok) {
      cudaGetLastError();
      err = "device allocation for the tiered path failed";
      return -4;
    }
  }
  p->available = true;
  return 0;
}

inline void fused_free(FusedPlan* p) {
  cudaFree(p->fb.stL); cudaFree(p->fb.stR); cudaFree(p->fb.stz); cudaFree(p->fb.sep); cudaFree(p->fb.lamhat);
  for (int t = 0; t < 8; ++t) cudaFree(p->fb.ent[t]);
  cudaFree(p->xsep); cudaFree(p->xlam);
  memset(p, 0, sizeof(*p));
}

template <int NS, int NI, int G, bool DMMA>
inline cudaError_t launch_tier0(const FusedPlan& p, const Shape& s, const Bufs& b, int batch, int pack_rhs, cudaStream_t st) {
  constexpr int smem = tier0_smem_doubles<NS, NI, G>() * (int)sizeof(double);
  cudaError_t e = smem_optin(tier0_kernel<NS, NI, G, DMMA>, smem);
  if (e != cudaSuccess) return e;
  tier0_kernel<NS, NI, G, DMMA><<<(unsigned)((size_t)batch * p.M0), 16 * G, smem, st>>>(s, b, p.fb, pack_rhs, p.M0);
  return cudaGetLastError();
}

template <int NS, int NI>
inline cudaError_t tier_attr() {
  using D = Dim<NS, NI>;
  return smem_optin(tier_kernel<NS, NI>, (8 * D::ENT + 4 * tier_warp_scratch<NS>()) * (int)sizeof(double));
}

template <int NS, int NI>
inline void launch_tier(const Shape& s, const Bufs& b, const TierArgs& a, int batch, cudaStream_t st) {
  using D = Dim<NS, NI>;
  const int warps = a.ecount / 2 > 0 ? a.ecount / 2 : 1;
  const int smem = (a.ecount * D::ENT + warps * tier_warp_scratch<NS>()) * (int)sizeof(double);
  launch_k(tier_kernel<NS, NI>, dim3((unsigned)((size_t)batch * (a.Min / a.ecount))), dim3(32 * warps), (size_t)smem, st, s, b, a);
}

// tiers >= 1 of the local tree. `last_out`: where the last tier puts its merged entry (shards), nullptr = nowhere.
template <int NS, int NI>
inline cudaError_t launch_tiers(FusedPlan& p, const Shape& s, const Bufs& b, int batch, cudaStream_t st,
                                double* last_out, long long last_out_stride) {
  using D = Dim<NS, NI>;
  cudaError_t e = tier_attr<NS, NI>();
  if (e != cudaSuccess) return e;
  for (int t = 0; t < p.ntiers; ++t) {
    const FusedTier& T = p.tiers[t];
    TierArgs a{};
    a.in = p.fb.ent[t]; a.in_stride = D::ENT;
    a.ab = b.knots + (size_t)(T.span - 1) * s.stride + s.oA; a.ab_stride = (long long)T.span * s.stride;
    a.ab_pstride = (long long)s.N * s.stride;
    a.sep = p.fb.sep; a.sep_pstride = (long long)p.M0 * D::WSZ; a.sep_mul = T.span / p.G0;
    const bool last = t == p.ntiers - 1;
    a.out = last ? last_out : p.fb.ent[t + 1];
    a.out_stride = last ? last_out_stride : D::ENT;
    a.Min = T.Min; a.ecount = T.ecount; a.span = T.span;
    launch_tier<NS, NI>(s, b, a, batch, st);
    ++p.launches;
  }
  return cudaGetLastError();
}

// top-down over the local upper separators + finalize. left/right: multipliers of the neighbouring separators.
template <int NS, int NI>
inline cudaError_t launch_down(FusedPlan& p, const Shape& s, const Bufs& b, int batch, cudaStream_t st,
                               const double* left, const double* right) {
  using D = Dim<NS, NI>;
  p.launches += launch_topdown<NS>(p.fb.sep, p.fb.lamhat, p.M0, batch, left, right, st);
  const long long rows = (long long)batch * s.N * D::ZS;
  finalize_kernel<NS, NI><<<(unsigned)((rows + 255) / 256), 256, 0, st>>>(s, b, p.fb, p.G0, p.M0, rows);
  ++p.launches;
  return cudaGetLastError();
}

template <int NS, int NI>
inline cudaError_t launch_upper(FusedPlan& p, const Shape& s, const Bufs& b, int batch, cudaStream_t st) {
  cudaError_t e = launch_tiers<NS, NI>(p, s, b, batch, st, nullptr, 0);
  if (e != cudaSuccess) return e;
  return launch_down<NS, NI>(p, s, b, batch, st, nullptr, nullptr);
}

// Whole solve of `batch` problems on stream `st`. force_dfma != 0 selects the DFMA Schur update for A/B tests.
inline int fused_launch(FusedPlan& p, const Shape& s, const Bufs& b, int batch, int pack_rhs, int force_dfma,
                        cudaStream_t st, std::string& err, cudaEvent_t k0 = nullptr, cudaEvent_t k1 = nullptr) {
  if (!p.available) { err = "shape not covered by the tiered kernels"; return -5; }
  p.launches = 1;
  cudaError_t e;
  const bool dm = p.use_dmma && !force_dfma;
  if (k0) cudaEventRecord(k0, st);  // k0..k1 bracket the dominant kernel (tier 0) only
  if (s.n == 12) {
    if (p.G0 == 16) e = dm ? launch_tier0<12, 4, 16, true>(p, s, b, batch, pack_rhs, st) : launch_tier0<12, 4, 16, false>(p, s, b, batch, pack_rhs, st);
    else e = dm ? launch_tier0<12, 4, 8, true>(p, s, b, batch, pack_rhs, st) : launch_tier0<12, 4, 8, false>(p, s, b, batch, pack_rhs, st);
    if (k1) cudaEventRecord(k1, st);
    if (e == cudaSuccess && p.M0 > 1) e = launch_upper<12, 4>(p, s, b, batch, st);
  } else {
    if (p.G0 == 16) e = launch_tier0<6, 3, 16, false>(p, s, b, batch, pack_rhs, st);
    else e = launch_tier0<6, 3, 8, false>(p, s, b, batch, pack_rhs, st);
    if (k1) cudaEventRecord(k1, st);
    if (e == cudaSuccess && p.M0 > 1) e = launch_upper<6, 3>(p, s, b, batch, st);
  }
  if (e != cudaSuccess) { err = std::string("tiered launch: ") + cudaGetErrorString(e); return -3; }
  return 0;
}

// ---- shards of one long horizon (one per GPU): reduce -> exchange (caller) -> finish ----
inline size_t fused_payload_doubles(int n, int m) { return fused_ent_doubles(n, m) + (((size_t)n * (n + m) + 1) & ~(size_t)1); }

template <int NS, int NI>
inline cudaError_t shard_reduce_t(FusedPlan& p, const Shape& s, const Bufs& b, int pack_rhs, int force_dfma, cudaStream_t st,
                                  double* payload) {
  using D = Dim<NS, NI>;
  const bool dm = p.use_dmma && !force_dfma;
  cudaError_t e;
  if constexpr (NS == 12) {
    if (p.G0 == 16) e = dm ? launch_tier0<NS, NI, 16, true>(p, s, b, 1, pack_rhs, st) : launch_tier0<NS, NI, 16, false>(p, s, b, 1, pack_rhs, st);
    else e = dm ? launch_tier0<NS, NI, 8, true>(p, s, b, 1, pack_rhs, st) : launch_tier0<NS, NI, 8, false>(p, s, b, 1, pack_rhs, st);
  } else {
    e = p.G0 == 16 ? launch_tier0<NS, NI, 16, false>(p, s, b, 1, pack_rhs, st) : launch_tier0<NS, NI, 8, false>(p, s, b, 1, pack_rhs, st);
  }
  if (e != cudaSuccess) return e;
  p.launches = 1;
  e = launch_tiers<NS, NI>(p, s, b, 1, st, payload, (long long)fused_payload_doubles(NS, NI));
  if (e != cudaSuccess) return e;
  // [A|B] of the shard's last knot rides along: the neighbour's merge needs it (ndlqr_FactorInnerProduct)
  return cudaMemcpyAsync(payload + D::ENT, b.knots + (size_t)(s.N - 1) * s.stride + s.oA, sizeof(double) * NS * (NS + NI),
                         cudaMemcpyDeviceToDevice, st);
}

template <int NS, int NI>
inline cudaError_t shard_finish_t(FusedPlan& p, const Shape& s, const Bufs& b, cudaStream_t st, const double* gathered,
                                  int rank, int nranks) {
  using D = Dim<NS, NI>;
  const double *left = nullptr, *right = nullptr;
  if (nranks > 1) {
    cudaError_t e = tier_attr<NS, NI>();
    if (e != cudaSuccess) return e;
    TierArgs a{};
    const long long pay = (long long)fused_payload_doubles(NS, NI);
    a.in = gathered; a.in_stride = pay;
    a.ab = gathered + D::ENT; a.ab_stride = pay; a.ab_pstride = 0;
    a.sep = p.xsep; a.sep_pstride = 0; a.sep_mul = 1;
    a.out = nullptr; a.out_stride = 0;
    a.Min = nranks; a.ecount = nranks; a.span = 0;
    launch_tier<NS, NI>(s, b, a, 1, st);  // every rank runs the top log2(nranks) levels redundantly, bit-identically
    topdown_kernel<NS><<<1, 128, 0, st>>>(p.xsep, p.xlam, nranks, nullptr, nullptr, nranks, 1);
    p.launches += 2;
    left = p.xlam + (size_t)rank * NS;
    right = p.xlam + (size_t)(rank + 1) * NS;
  }
  return launch_down<NS, NI>(p, s, b, 1, st, left, right);
}

inline int fused_shard_reduce(FusedPlan& p, const Shape& s, const Bufs& b, int pack_rhs, int force_dfma, cudaStream_t st,
                              double* payload, std::string& err) {
  if (!p.available || p.ntiers < 1) { err = "shard needs the tiered kernels and at least two groups of knots"; return -5; }
  cudaError_t e = s.n == 12 ? shard_reduce_t<12, 4>(p, s, b, pack_rhs, force_dfma, st, payload)
                            : shard_reduce_t<6, 3>(p, s, b, pack_rhs, force_dfma, st, payload);
  if (e != cudaSuccess) { err = std::string("shard reduce: ") + cudaGetErrorString(e); return -3; }
  return 0;
}

inline int fused_shard_finish(FusedPlan& p, const Shape& s, const Bufs& b, cudaStream_t st, const double* gathered, int rank,
                              int nranks, std::string& err) {
  if (!p.available || p.ntiers < 1) { err = "shard needs the tiered kernels"; return -5; }
  cudaError_t e = s.n == 12 ? shard_finish_t<12, 4>(p, s, b, st, gathered, rank, nranks)
                            : shard_finish_t<6, 3>(p, s, b, st, gathered, rank, nranks);
  if (e != cudaSuccess) { err = std::string("shard finish: ") + cudaGetErrorString(e); return -3; }
  return 0;
}

}  // namespace rslqr

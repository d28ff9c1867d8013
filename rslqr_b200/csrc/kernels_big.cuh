// kernels_big.cuh -- the dual-form nested dissection (see kernels_dual.cuh for the formulation) for LARGE blocks:
// n = 64, m = 16 (BASELINE configs[4]) and, with n padded on chip to blocks of NS = 32 or 64, every other n <= 64, m <= 32
// that has no instance of the warp kernels. At n = 64 an element [Saa | Sab | Sbb | sa | sb] is 99 KB and a merge touches three of
// them plus a 66 KB record, so elements live in HBM / L2 and a CTA (not a warp) owns a merge.
//
// Per level, three launches over all (problem, merge) pairs, all block arithmetic on mma.sync.m8n8k4.f64:
//   big_factor_reg_kernel : S = SbbL + SaaR = U^T D U                  (ndlqr_FactorInnerProduct + the separator
//                           Cholesky, src/nested_dissection.c:114-134, src/solve.c:87-98)
//   big_solve_kernel      : W = U^-T [S_L | S_R | s]                   (the forward half of ndlqr_SolveCholeskyFactor, :136-152)
//   big_schur_kernel      : Saa' = SaaL - W_L^T D^-1 W_L, Sab' = -W_L^T D^-1 W_R, sa' = saL - W_L^T D^-1 w_s,
//                           Sbb' = SbbR - W_R^T D^-1 W_R, sb' = sbR - W_R^T D^-1 w_s
//                           (= ndlqr_UpdateShurFactor restricted to what a later inner product reads, :154-171)
// then one top-down launch per level (lambda of the separators: records, then one back-substitution with U) and
// big_down_kernel (x_k, u_k from the leaf equations, as scan_down_kernel). Against the staged path on this shape (every
// knot updated at every level, dense `fact` array of 755 MB per problem): 2.2 instead of 23.6 GFLOP per solve, 0.26
// instead of 0.76 GB per problem.
#pragma once
#include "kernels_fused.cuh"

namespace rslqr {

// Ampere-style asynchronous copies global -> shared (LDGSTS): no staging registers, every copy of a prologue in flight
// at once. 16-byte copies bypass L1 (.cg); 8-byte ones serve transposes and sources that are only 8-byte aligned.
__device__ __forceinline__ void cp_async8(double* sdst, const double* gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(sdst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async16(double* sdst, const double* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(sdst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

template <int NS>
struct BigDim {
  static constexpr int BLK = NS * NS;
  static constexpr int OAA = 0, OAB = BLK, OBB = 2 * BLK, OSA = 3 * BLK, OSB = 3 * BLK + NS;
  static constexpr int ENT = 3 * BLK + 2 * NS;   // [Saa | Sab | Sbb | sa | sb], blocks row-major ld NS
  static constexpr int WSZ = (2 * NS + 1) * NS;  // record [W_L | W_R | w_s], column-major ld NS
};

struct BigBufs {
  double* entA;  // [batch][N][ENT]      elements of the even levels (level l has N >> l of them, packed at the front)
  double* entB;  // [batch][N/2][ENT]    elements of the odd levels
  double* rec;   // [batch][N][WSZ]      record [W_L | W_R | w_s] of the separator after knot i at slot i (slot N-1 unused)
  double* lam;   // [batch][N+1][NS]     multipliers lambda_k; slots 0 and N stay zero
  double* ufac;  // [batch][N][BigFac::SIZE] factor U, block inverses and 1/d of the separator after knot i at slot i
};

constexpr int BIG_NI_MAX = 32;  // the leaf and down kernels are laid out for NI = 16 or 32 input columns (m <= NI)

// G output tiles of Sbb = [A | B] D [A | B]^T for one warp: tiles warp, warp + 8, ... of the lower block triangle in row-major
// order (tile i = (t, j), j <= t, i = t (t + 1) / 2 + j). AB: [A | B] in shared memory (column-major, ld NS + 4), dv: D.
// Rows / columns >= n are padding: 0.5 on the diagonal there (the S of a merge, SbbL + SaaR, gets the identity).
template <int NS, int NI, int G>
__device__ __forceinline__ void big_leaf_sbb(const double* AB, const double* dv, double* Sbb, int warp, int lane, int n) {
  constexpr int XU = NS + NI, LDA = NS + 4, KS = XU / 4;
  const int q = lane & 3, rr = lane >> 2;
  const double *pa[G], *pb[G];
  double* po[G];
  double c[G][2];
  int drow[G];  // row of the tile's lane if the tile is a diagonal one, else -1
#pragma unroll
  for (int u = 0; u < G; ++u) {
    int t = 0, jj = warp + 8 * u;
    while (jj > t) { ++t; jj -= t; }
    pa[u] = AB + 8 * t + rr + LDA * q;
    pb[u] = AB + 8 * jj + rr + LDA * q;
    po[u] = Sbb + (8 * t + rr) * NS + 8 * jj + 2 * q;
    drow[u] = t == jj ? 8 * t + rr : -1;
    c[u][0] = c[u][1] = 0.0;
  }
#pragma unroll 4
  for (int ks = 0; ks < KS; ++ks) {
    const double dk = dv[4 * ks + q];
#pragma unroll
    for (int u = 0; u < G; ++u) dmma884(c[u][0], c[u][1], pa[u][4 * LDA * ks] * dk, pb[u][4 * LDA * ks]);
  }
#pragma unroll
  for (int u = 0; u < G; ++u) {
    if (drow[u] >= n) {  // padded row of a diagonal tile (the products are zero there)
      const int col = (drow[u] & ~7) + 2 * q;
      if (col == drow[u]) c[u][0] = 0.5;
      if (col + 1 == drow[u]) c[u][1] = 0.5;
    }
    *reinterpret_cast<double2*>(po[u]) = make_double2(c[u][0], c[u][1]);
  }
}

// ---------------------------------------------------------------------------------------------------------
// Leaf elements (ndlqr_SolveLeaf on diagonal Q, R, src/nested_dissection.c:10-105, with the rhs packing of
// src/solver.c:138-190), one CTA per (problem, knot):
//   Saa = Q^-1, Sab = -Q^-1 A^T, Sbb = A Q^-1 A^T + B R^-1 B^T, sa = -z_x - z_lambda, sb = A z_x + B z_u
//   knot 0: x_0 is fixed (z_x = x0) and lambda_0 is no unknown: Q^-1 -> 0 in the blocks; last knot: only Saa, sa.
// The blocks are NS x NS with NS = 32 or 64 >= n = s.n: a smaller n is padded on chip (zero rows / columns, 0.5 on the
// padded diagonal of Saa and Sbb), m = s.m <= NI (16 or 32 columns of B); everything above the leaves works on NS x NS blocks.
// dynamic smem: big_leaf_smem_doubles<NS, NI>()
// ---------------------------------------------------------------------------------------------------------
template <int NS, int NI>
__host__ __device__ constexpr int big_leaf_smem_doubles() {
  return (NS + 4) * (NS + NI) + 2 * (NS + NI) + 2 * NS + (256 / NS - 1) * NS;
}

template <int NS, int NI>
__global__ void __launch_bounds__(256, 3) big_leaf_kernel(Shape s, Bufs b, BigBufs bb, int pack_rhs) {
  using D = BigDim<NS>;
  constexpr int XU = NS + NI, LDA = NS + 4;  // (LDA = 4 mod 16: the DMMA fragments of a warp fall into two wavefronts)
  constexpr int NT = NS / 8, NTASK = NT * (NT + 1) / 2, PARTS = 256 / NS;
  static_assert(NS % 32 == 0 && 256 % NS == 0 && XU % 4 == 0 && XU % PARTS == 0 && NTASK >= 8, "8 warps; k-steps of 4");
  extern __shared__ __align__(16) double smf[];
  double* AB = smf;                 // [NS x XU] column-major with leading dimension LDA: columns [0, NS) = A, [NS, XU) = B
  double* dv = AB + LDA * XU;       // 1/Q, 1/R (as used in the products: 0 for Q of knot 0, for the last knot, for padding)
  double* zv = dv + XU;             // leaf-solved z_x, z_u
  double* sc = zv + XU;             // 1/Q of this knot (also for knot 0), then z_lambda
  const int N = s.N, n = s.n, m = s.m;
  const int prob = blockIdx.x / N, k = blockIdx.x % N;
  const int tid = threadIdx.x, nthr = blockDim.x;
  const double* kn = knot_ptr(s, b, prob, k);
  const bool first = k == 0, last = k == N - 1;
  // [A | B] -> shared memory by 8-byte asynchronous copies (the knot block is only 8-byte aligned), in flight while the
  // diagonal / right-hand-side part below is computed; padding is zero
  if (n == NS && m == NI) {  // nothing to pad: [A | B] is one contiguous run of the knot block
    for (int e = tid; e < NS * XU; e += nthr) {
      const int l = e / NS, r = e - l * NS;
      if (last) AB[r + LDA * l] = 0.0;
      else cp_async8(AB + r + LDA * l, kn + s.oA + e);
    }
  } else {
    for (int e = tid; e < NS * XU; e += nthr) {
      const int l = e / NS, r = e - l * NS;
      const int lc = l < NS ? l : l - NS;                       // column of A, or of B
      const bool valid = r < n && (l < NS ? lc < n : lc < m) && !last;
      if (valid) cp_async8(AB + r + LDA * l, kn + (l < NS ? s.oA : s.oB) + (size_t)lc * n + r);
      else AB[r + LDA * l] = 0.0;
    }
  }
  for (int l = tid; l < XU; l += nthr) {
    const bool isR = l >= NS;
    const int li = isR ? l - NS : l;
    if (li >= (isR ? m : n)) {  // padding
      zv[l] = 0.0; dv[l] = 0.0;
      if (!isR) { sc[l] = 0.0; sc[NS + l] = 0.0; }
      continue;
    }
    const double d = kn[isR ? n + li : li];  // Q then R lead the block
    if (!(d > 0.0) && !(isR && last)) {
      b.status[(size_t)prob * 3 * N + (isR ? rs_slot_R(k) : rs_slot_Q(k))] = -1;
      b.flags[prob] = 1;
    }
    const double dinv = (isR && last) ? 0.0 : 1.0 / d;
    double rx;  // negated rhs of row li of x, or of u
    if (!pack_rhs) rx = soln_ptr(s, b, prob, k)[isR ? 2 * n + li : n + li];
    else rx = isR ? (last ? 0.0 : -kn[s.orr + li]) : -kn[s.oq + li];
    double z;
    if (first && !isR) {
      const double rl = pack_rhs ? -b.x0[(size_t)prob * n + li] : soln_ptr(s, b, prob, k)[li];
      z = -rl;  // z_x(0) = x0
    } else {
      z = (isR && last) ? rx : rx * dinv;
    }
    zv[l] = z;
    dv[l] = (last || (first && !isR)) ? 0.0 : dinv;
    if (!isR) {
      sc[l] = first ? 0.0 : dinv;
      double zl = 0.0;  // z_lambda(k) = negated rhs of the lambda row (unused for knot 0)
      if (!first) zl = pack_rhs ? -(kn - s.stride)[s.od + li] : soln_ptr(s, b, prob, k)[li];
      sc[NS + l] = zl;
    }
  }
  cp_async_wait_all();
  __syncthreads();
  double* E = bb.entA + ((size_t)prob * N + k) * D::ENT;
  // the readers of Saa, Sbb (factorization, keep part of the Schur update) take the lower block triangle only
  // (8 x 8 tiles with column tile <= row tile): the tiles above it are neither computed nor written
  for (int e = 2 * tid; e < NS * NS; e += 2 * nthr) {  // two columns per thread: 128-bit loads and stores
    const int r = e / NS, c = e - r * NS;
    const double qi = sc[r], di = r < n ? qi : 0.5;    // (padded diagonal: 0.5)
    if ((c >> 3) <= (r >> 3)) *reinterpret_cast<double2*>(E + D::OAA + e) = make_double2(r == c ? di : 0.0, r == c + 1 ? di : 0.0);
    const double2 av = *reinterpret_cast<const double2*>(AB + c + LDA * r);  // -A(c, r) / Q(r)   (AB is zero for the last knot)
    *reinterpret_cast<double2*>(E + D::OAB + e) = make_double2(-qi * av.x, -qi * av.y);
  }
  // Sbb = [A | B] D [A | B]^T on the fp64 tensor cores, 8 x 8 output tiles: the tiles of the lower block triangle, dealt
  // round-robin to the 8 warps, all of a warp in flight (NS = 64: 36 tiles, 5 or 4 per warp)
  {
    const int warp = tid >> 5;
    constexpr int GLO = NTASK / 8, REM = NTASK % 8;
    if (warp < REM) big_leaf_sbb<NS, NI, GLO + 1>(AB, dv, E + D::OBB, warp, tid & 31, n);
    else big_leaf_sbb<NS, NI, GLO>(AB, dv, E + D::OBB, warp, tid & 31, n);
  }
  {  // sb = [A | B] z: PARTS threads per row, a share of the XU terms each (thread = NS part + row)
    const int r = tid & (NS - 1), part = tid / NS;
    double v = 0.0;
#pragma unroll 4
    for (int l = part * (XU / PARTS); l < (part + 1) * (XU / PARTS); ++l) v = fma(AB[r + LDA * l], zv[l], v);
    double* ps = sc + 2 * NS;  // partial sums
    if (part > 0) ps[(part - 1) * NS + r] = v;
    __syncthreads();
    if (part == 0) {
#pragma unroll
      for (int pp = 1; pp < PARTS; ++pp) v += ps[(pp - 1) * NS + r];
      E[D::OSA + r] = -zv[r] - sc[NS + r];
      E[D::OSB + r] = v;
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// Spine of a merge: S = SbbL + SaaR = U^T D U, then W = U^-T [S_L | S_R | s] (the forward substitution ONLY).
//   big_factor_reg_kernel : one warp per merge, S in registers; output per separator (`ufac`, slot = the separator's
//                           knot index): U (strict upper triangle, row-major ld NS), the inverses X_b of the 8 x 8 diagonal blocks
//                           of L = U^T, and 1/d.
//   big_solve_kernel      : one CTA of 9 warps per merge. U and the X_b are staged in shared memory; a warp owns row tiles of
//                           Y^T = [S_L | S_R | s]^T (8 right-hand sides each, 17 tiles) straight from global memory in its
//                           accumulator registers and runs the blocked forward substitution in transposed form,
//                               Z_b = (Y^T_b - sum_{b' < b} Z_b' U_b'b) X_b^T,      Z = W^T
//                           every product an m8n8k4 DMMA pair, no shared-memory round trip for the right-hand sides.
//                           W = [W_L | W_R | w_s] goes to the separator's record.
// The backward substitution is never run on the 2n+1 columns: the Schur update only needs products of the form
// S_x^T S^-1 S_y = W_x^T D^-1 W_y (big_schur_kernel), and the separator's multiplier
//     lambda = S^-1 (s - S_L lambda_a - S_R lambda_b) = U^-1 D^-1 (w_s - W_L lambda_a - W_R lambda_b)
// needs ONE back-substitution per separator, in the top-down pass (big_topdown_kernel), which is why U and 1/d of
// every separator are kept.
// ---------------------------------------------------------------------------------------------------------
template <int NS>
struct BigFac {
  static constexpr int NB = NS / 8;
  static constexpr int OU = 0, OX = NS * NS, OD = NS * NS + NB * 64, SIZE = NS * NS + NB * 64 + NS;  // doubles per merge in `ufac`
};

// ---------------------------------------------------------------------------------------------------------
// big_factor_reg_kernel: the matrix is held in REGISTERS as DMMA accumulator
// fragments (lower block triangle: 36 blocks of 8 x 8, two doubles per lane each) and a blocked right-looking LDL^T:
//   for block column b:  S_bb = L_bb D_b L_bb^T by row elimination on [S_bb | I] inside the warp (shuffles; gives
//                         X_b = L_bb^-1 directly),
//                        panel  R_c = S_cb X_b^T (= L_cb D_b),  L_cb = R_c D_b^-1          (2 DMMAs per block)
//                        trailing S_cc' -= L_cb (L_c'b D_b)^T                                (2 DMMAs per block)
// C fragments become A / B operands through four 64-bit shuffles (frag_c2a). One warp per merge, no shared memory, no
// barrier; about 2.5 k warp instructions per matrix (a scalar one-warp factorization of a packed triangle in shared
// memory, the first version of this kernel, executed 78 k).
// ---------------------------------------------------------------------------------------------------------
constexpr int BIG_FACTOR_REG_WPC = 4;

// C fragment (lane (g, q): C[g][2q], C[g][2q+1]) -> A fragments of the same block (a0 = C[g][q], a1 = C[g][4+q]);
// read as a B operand this is the block's transpose
__device__ __forceinline__ void frag_c2a(double c0, double c1, double& a0, double& a1, int lane) {
  const int g4 = lane & ~3, q = lane & 3, src = g4 | (q >> 1);
  const double x0 = shfl_d(c0, src), x1 = shfl_d(c1, src);
  const double y0 = shfl_d(c0, src | 2), y1 = shfl_d(c1, src | 2);
  a0 = (q & 1) ? x1 : x0;
  a1 = (q & 1) ? y1 : y0;
}

template <int NS>
__global__ void __launch_bounds__(32 * BIG_FACTOR_REG_WPC) big_factor_reg_kernel(Shape s, Bufs b, const double* ein, double* ufac,
                                                                                  int level, int M, long long nmerge) {
  using D = BigDim<NS>;
  using F = BigFac<NS>;
  constexpr int NB = NS / 8;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, q = lane & 3;
  const long long mi = (long long)blockIdx.x * BIG_FACTOR_REG_WPC + warp;
  if (mi >= nmerge) return;
  const int N = s.N;
  const int prob = (int)(mi / M), t = (int)(mi % M);
  const double* EL = ein + ((size_t)prob * 2 * M + 2 * t) * D::ENT;
  const double* ER = EL + D::ENT;
  double* U = ufac + ((size_t)prob * N + ((2 * t + 1) << level) - 1) * F::SIZE;  // slot = the separator's knot index
  double S[NB][NB][2];  // block (c, c'), c >= c', as a C fragment
#pragma unroll
  for (int c = 0; c < NB; ++c)
#pragma unroll
    for (int c2 = 0; c2 <= c; ++c2) {
      const int off = (8 * c + g) * NS + 8 * c2 + 2 * q;
      const double2 x = __ldg(reinterpret_cast<const double2*>(EL + D::OBB + off));
      const double2 y = __ldg(reinterpret_cast<const double2*>(ER + D::OAA + off));
      S[c][c2][0] = x.x + y.x;
      S[c][c2][1] = x.y + y.y;
    }
  int bad = 0;
#pragma unroll
  for (int bk = 0; bk < NB; ++bk) {
    // --- diagonal block: row elimination on [S_bb | I]
    double s0 = S[bk][bk][0], s1 = S[bk][bk][1];
    double x0 = (g == 2 * q) ? 1.0 : 0.0, x1 = (g == 2 * q + 1) ? 1.0 : 0.0;
    double dq0 = 0.0, dq1 = 0.0;   // 1/d of columns 2q, 2q+1 of the block
    double da0 = 0.0, da1 = 0.0;   // 1/d of columns q, 4+q
    double* Ud = U + F::OU + (size_t)(8 * bk) * NS + 8 * bk;
#pragma unroll 1
    for (int j = 0; j < 8; ++j) {
      const double v = (j & 1) ? s1 : s0;
      const double colv = shfl_d(v, (lane & ~3) | (j >> 1));  // S'(g, j)
      const double d = shfl_d(v, (j << 2) | (j >> 1));        // S'(j, j)
      const int rowj = (j << 2) | q;
      const double rs0 = shfl_d(s0, rowj), rs1 = shfl_d(s1, rowj), rx0 = shfl_d(x0, rowj), rx1 = shfl_d(x1, rowj);
      if (!(d > 0.0)) bad = 1;
      const double dinv = fast_rcp(d);
      const double l = g > j ? colv * dinv : 0.0;
      s0 = fma(-l, rs0, s0); s1 = fma(-l, rs1, s1);
      x0 = fma(-l, rx0, x0); x1 = fma(-l, rx1, x1);
      if (q == (j >> 1)) { if (j & 1) dq1 = dinv; else dq0 = dinv; }
      if (q == (j & 3)) { if (j >> 2) da1 = dinv; else da0 = dinv; }
      if (q == 0 && g > j) Ud[(size_t)j * NS + g] = l;  // U(8b + j, 8b + g) = l(g, j): only the strict upper triangle of U is kept
    }
    *reinterpret_cast<double2*>(U + F::OX + bk * 64 + g * 8 + 2 * q) = make_double2(x0, x1);
    if (g == 0) *reinterpret_cast<double2*>(U + F::OD + 8 * bk + 2 * q) = make_double2(dq0, dq1);
    if (bk + 1 < NB) {
      double xa0, xa1;
      frag_c2a(x0, x1, xa0, xa1, lane);  // X_b as a B operand = X_b^T
      double nl[NB][2], rd[NB][2];       // -L_cb as A fragments, (L_cb D_b)^T as B fragments
#pragma unroll
      for (int c = bk + 1; c < NB; ++c) {
        double pa0, pa1;
        frag_c2a(S[c][bk][0], S[c][bk][1], pa0, pa1, lane);
        double r0 = 0.0, r1 = 0.0;
        dmma884(r0, r1, pa0, xa0);
        dmma884(r0, r1, pa1, xa1);
        // U(8b + k, 8c + g) = L_cb(g, k) (nothing is stored on or below the diagonal: no reader looks there)
        U[F::OU + (size_t)(8 * bk + 2 * q) * NS + 8 * c + g] = r0 * dq0;
        U[F::OU + (size_t)(8 * bk + 2 * q + 1) * NS + 8 * c + g] = r1 * dq1;
        frag_c2a(r0, r1, rd[c][0], rd[c][1], lane);
        nl[c][0] = -rd[c][0] * da0;
        nl[c][1] = -rd[c][1] * da1;
      }
#pragma unroll
      for (int c = bk + 1; c < NB; ++c)
#pragma unroll
        for (int c2 = bk + 1; c2 <= c; ++c2) {
          dmma884(S[c][c2][0], S[c][c2][1], nl[c][0], rd[c2][0]);
          dmma884(S[c][c2][0], S[c][c2][1], nl[c][1], rd[c2][1]);
        }
    }
  }
  if (bad && lane == 0) {
    b.status[(size_t)prob * 3 * N + rs_slot_S(t, level, N)] = -1;
    b.flags[prob] = 1;
  }
}

// Forward substitution in TRANSPOSED form: Z = Y^T U^-1 (Z = W^T), T row tiles of 8 right-hand sides x NS per warp, the
// 8 column blocks of a row tile as DMMA accumulators (C fragments) in registers, right-looking:
//     Z_b = Zc_b X_b^T (final),   then   Zc_b2 -= Z_b U_{b,b2}  for b2 > b
// The right-hand sides are the A operands (a C fragment becomes an A fragment by four 64-bit shuffles, frag_c2a: no shared
// memory round trip), the 8 x 8 blocks of U and X_b^T are the B operands, loaded once per block for all T tiles. A final
// block leaves for the record (W(:, c) = Z(c, :), column-major ld NS: 16-byte runs) straight from the registers.
template <int NS, int T>
__device__ __forceinline__ void big_solve_rows(const double* Us, const double* Xs, double (&c)[NS / 8][T][2], double* const (&rec)[T],
                                               const int (&nrows)[T]) {
  constexpr int NB = NS / 8, LDU = NS + 4;
  const int lane = threadIdx.x & 31, q = lane & 3, rr = lane >> 2;
#pragma unroll
  for (int bk = 0; bk < NB; ++bk) {
    const double xb0 = Xs[bk * 64 + rr * 8 + q], xb1 = Xs[bk * 64 + rr * 8 + 4 + q];  // B fragments of X_b^T
    double na[T][2];
#pragma unroll
    for (int u = 0; u < T; ++u) {
      double a0, a1;
      frag_c2a(c[bk][u][0], c[bk][u][1], a0, a1, lane);
      c[bk][u][0] = c[bk][u][1] = 0.0;
      dmma884(c[bk][u][0], c[bk][u][1], a0, xb0);
      dmma884(c[bk][u][0], c[bk][u][1], a1, xb1);
      if (rr < nrows[u]) *reinterpret_cast<double2*>(rec[u] + (size_t)rr * NS + 8 * bk + 2 * q) = make_double2(c[bk][u][0], c[bk][u][1]);
      if (bk + 1 < NB) {
        frag_c2a(c[bk][u][0], c[bk][u][1], a0, a1, lane);
        na[u][0] = -a0; na[u][1] = -a1;
      }
    }
    if (bk + 1 < NB) {
#pragma unroll
      for (int b2 = bk + 1; b2 < NB; ++b2) {
        const double ub0 = Us[(8 * bk + q) * LDU + 8 * b2 + rr], ub1 = Us[(8 * bk + 4 + q) * LDU + 8 * b2 + rr];
#pragma unroll
        for (int u = 0; u < T; ++u) { dmma884(c[b2][u][0], c[b2][u][1], na[u][0], ub0); dmma884(c[b2][u][0], c[b2][u][1], na[u][1], ub1); }
      }
    }
  }
}

// NS / 8 + 1 warps: NS / 8 take a row tile of S_L^T and one of S_R^T each, the last one the vector s (NS = 64: 288 threads)
template <int NS>
__host__ __device__ constexpr int big_solve_threads() { return 32 * (NS / 8 + 1); }

// dynamic smem: (NS * (NS + 4) + (NS / 8) * 64) doubles: the blocks of U right of the diagonal, the X_b
template <int NS>
__global__ void __launch_bounds__(big_solve_threads<NS>(), 3) big_solve_kernel(Shape s, const double* ein, const double* ufac, double* rec_all,
                                                                       int level, int M) {
  using D = BigDim<NS>;
  using F = BigFac<NS>;
  static_assert(NS % 8 == 0, "NS / 8 column blocks, 2 NS / 8 + 1 row tiles over NS / 8 + 1 warps");
  constexpr int NB = NS / 8, LDU = NS + 4;
  extern __shared__ __align__(16) double smf[];
  double* Us = smf;                      // U, row-major ld LDU
  double* Xs = Us + NS * LDU;            // X_b, 8 x 8 row-major each
  const int N = s.N;
  const int prob = blockIdx.x / M, t = blockIdx.x % M;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, nthr = blockDim.x, q = lane & 3, rr = lane >> 2;
  const double* EL = ein + ((size_t)prob * 2 * M + 2 * t) * D::ENT;
  const double* ER = EL + D::ENT;
  const int i_sep = ((2 * t + 1) << level) - 1;
  const double* Uf = ufac + ((size_t)prob * N + i_sep) * F::SIZE;
  for (int e = 2 * tid; e < NS * NS; e += 2 * nthr) {
    const int r = e / NS, c = e - r * NS;
    if (c >= 8 * (r / 8 + 1)) cp_async16(Us + r * LDU + c, Uf + F::OU + e);   // (the blocks right of the diagonal one: all that is read)
  }
  for (int e = 2 * tid; e < NB * 64; e += 2 * nthr) cp_async16(Xs + e, Uf + F::OX + e);
  double* rec = rec_all + ((size_t)prob * N + i_sep) * D::WSZ;  // [W_L | W_R | w_s], column-major ld NS
  // the right-hand sides come straight from global memory into the accumulator fragments (row = right-hand side):
  //   S_L^T = SabL: row c, 16-byte runs;  S_R^T(c, k) = SabR(k, c): 64-byte runs over the 8 rows of a tile
  if (warp < NB) {
    double c[NB][2][2];
    const double* pl = EL + D::OAB + (size_t)(8 * warp + rr) * NS + 2 * q;
    const double* pr = ER + D::OAB + (size_t)(2 * q) * NS + 8 * warp + rr;
#pragma unroll
    for (int bk = 0; bk < NB; ++bk) {
      const double2 v = __ldg(reinterpret_cast<const double2*>(pl + 8 * bk));
      c[bk][0][0] = v.x; c[bk][0][1] = v.y;
      c[bk][1][0] = __ldg(pr + (size_t)(8 * bk) * NS);
      c[bk][1][1] = __ldg(pr + (size_t)(8 * bk + 1) * NS);
    }
    cp_async_wait_all();
    __syncthreads();
    double* const rc[2] = {rec + (size_t)8 * warp * NS, rec + (size_t)(NS + 8 * warp) * NS};
    const int nr[2] = {8, 8};
    big_solve_rows<NS, 2>(Us, Xs, c, rc, nr);
  } else {
    double c[NB][1][2];
#pragma unroll
    for (int bk = 0; bk < NB; ++bk) {  // s = sbL + saR as row 0 of the tile
      const int k = 8 * bk + 2 * q;
      c[bk][0][0] = rr == 0 ? __ldg(EL + D::OSB + k) + __ldg(ER + D::OSA + k) : 0.0;
      c[bk][0][1] = rr == 0 ? __ldg(EL + D::OSB + k + 1) + __ldg(ER + D::OSA + k + 1) : 0.0;
    }
    cp_async_wait_all();
    __syncthreads();
    double* const rc[1] = {rec + (size_t)2 * NS * NS};
    const int nr[1] = {1};
    big_solve_rows<NS, 1>(Us, Xs, c, rc, nr);
  }
}

// ---------------------------------------------------------------------------------------------------------
// Schur update of one merge on the fp64 tensor cores, one CTA of 8 warps. W = [W_L | W_R | w_s] (the record) is staged
// in shared memory (ld NS + 4: the B fragments of a warp fall into two wavefronts) and is BOTH operands:
//     out(r, c) = keep(r, c) - sum_k W_x(k, r) (1/d_k) W_y(k, c)
// the A fragments of a warp's 8 rows (column r of W_L or W_R from the record, scaled by 1/d) stay in registers across
// its column tiles. NS = 64: 8 warps x 8 rows per half.
// dynamic smem: (2 NS + 1) * (NS + 4) doubles.
// ---------------------------------------------------------------------------------------------------------
// One run of column tiles of the Schur update for the 8 rows of a warp: out(r, 8j..8j+7) = keep(r, .) - a(r, :) F(:, fcol0 + 8j..)
// for j < cnt (F = W here), four tiles in flight, the keep part requested one group ahead of its use. a: A fragments of
// the rows (registers); Fcols: column fcol0 of W in shared memory (ld LDF); keep_row: row r of the block that is updated, nullptr =
// zero; dst_row: row r of the output block.
struct BigKeep4 { double2 v0, v1, v2, v3; };

template <int NS, int NACT>
__device__ __forceinline__ void big_schur_group(const double (&a)[NS / 4], const double* Fcols, BigKeep4 kv, double* dst_row, int nt0,
                                                int q, int rr) {
  constexpr int LDF = NS + 4, KS = NS / 4;
  const double* bp = Fcols + (8 * nt0 + rr) * LDF + q;   // B fragment of tile u, k-step ks: bp[u * 8 * LDF + 4 * ks]
  double c[NACT][2];
  c[0][0] = -kv.v0.x; c[0][1] = -kv.v0.y;                // accumulate a.F - keep, negate on the way out
  if constexpr (NACT > 1) { c[1][0] = -kv.v1.x; c[1][1] = -kv.v1.y; }
  if constexpr (NACT > 2) { c[2][0] = -kv.v2.x; c[2][1] = -kv.v2.y; }
  if constexpr (NACT > 3) { c[3][0] = -kv.v3.x; c[3][1] = -kv.v3.y; }
#pragma unroll
  for (int ks = 0; ks < KS; ++ks) {
#pragma unroll
    for (int u = 0; u < NACT; ++u) dmma884(c[u][0], c[u][1], a[ks], bp[u * 8 * LDF + 4 * ks]);
  }
#pragma unroll
  for (int u = 0; u < NACT; ++u) *reinterpret_cast<double2*>(dst_row + 8 * (nt0 + u) + 2 * q) = make_double2(-c[u][0], -c[u][1]);
}

template <int NS>
__device__ __forceinline__ void big_schur_run(const double (&a)[NS / 4], const double* Fcols, const double* keep_row, double* dst_row,
                                              int cnt, int q, int rr) {
  auto keep1 = [&](int tile) {
    return (keep_row && tile < cnt) ? __ldg(reinterpret_cast<const double2*>(keep_row + 8 * tile + 2 * q)) : make_double2(0.0, 0.0);
  };
  auto keep = [&](int nt0) { return BigKeep4{keep1(nt0), keep1(nt0 + 1), keep1(nt0 + 2), keep1(nt0 + 3)}; };
#pragma unroll 1
  for (int nt0 = 0; nt0 < cnt; nt0 += 4) {
    const BigKeep4 kv = keep(nt0);      // (zeros beyond cnt)
    const int nact = cnt - nt0;         // (warp-uniform) tiles of this group; one instantiation per count: no guards inside
    if (nact >= 4) big_schur_group<NS, 4>(a, Fcols, kv, dst_row, nt0, q, rr);
    else if (nact == 3) big_schur_group<NS, 3>(a, Fcols, kv, dst_row, nt0, q, rr);
    else if (nact == 2) big_schur_group<NS, 2>(a, Fcols, kv, dst_row, nt0, q, rr);
    else big_schur_group<NS, 1>(a, Fcols, kv, dst_row, nt0, q, rr);
  }
}

// Saa' and Sbb' are symmetric and every reader (the factorization, this kernel's keep part) takes the lower block triangle
// only, so only the column tiles up to the diagonal one are computed: 136 instead of 192 tiles per merge. Warp w takes
// row tile w of the left half and row tile 7 - w of the right half: 17 tiles each.
template <int NS>
__global__ void __launch_bounds__(4 * NS, 3) big_schur_kernel(Shape s, const double* ein, double* eout, const double* rec_all,
                                                         const double* ufac, int level, int M) {
  using D = BigDim<NS>;
  using F = BigFac<NS>;
  static_assert(NS % 16 == 0, "NS / 8 warps x 8 rows per half; LDF = 4 mod 16");
  constexpr int LDF = NS + 4, KS = NS / 4, NT = NS / 8;
  extern __shared__ __align__(16) double Fs[];  // column c of W at Fs + c * LDF
  const int N = s.N;
  const int prob = blockIdx.x / M, t = blockIdx.x % M;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, q = lane & 3, rr = lane >> 2;
  const int Min = 2 * M;
  const double* EL = ein + ((size_t)prob * Min + 2 * t) * D::ENT;
  const double* ER = EL + D::ENT;
  double* EO = eout + ((size_t)prob * M + t) * D::ENT;
  const int i_sep = ((2 * t + 1) << level) - 1;
  const double* rec = rec_all + ((size_t)prob * N + i_sep) * D::WSZ;
  const double* dinv = ufac + ((size_t)prob * N + i_sep) * F::SIZE + F::OD;
  for (int e = tid; e < D::WSZ / 2; e += blockDim.x) {
    const int c = (2 * e) / NS, k = 2 * e - c * NS;
    cp_async16(Fs + c * LDF + k, rec + 2 * e);
  }
  const int tL = warp, tR = NT - 1 - warp;  // row tiles of this warp in the left / right half
  double di[KS];
#pragma unroll
  for (int ks = 0; ks < KS; ++ks) di[ks] = __ldg(dinv + 4 * ks + q);
  cp_async_wait_all();
  __syncthreads();
  const double* ws = Fs + 2 * NS * LDF;
  double a[KS];
  {  // left half: [Saa' | Sab' | sa'] rows of row tile tL; A fragments: W_L(k, r) / d_k
    const int r = 8 * tL + rr;
    double part = 0.0;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      a[ks] = Fs[r * LDF + 4 * ks + q] * di[ks];
      part = fma(a[ks], ws[4 * ks + q], part);
    }
    part += __shfl_xor_sync(0xffffffffu, part, 1);
    part += __shfl_xor_sync(0xffffffffu, part, 2);
    if (q == 0) EO[D::OSA + r] = EL[D::OSA + r] - part;
    big_schur_run<NS>(a, Fs, EL + D::OAA + r * NS, EO + D::OAA + r * NS, tL + 1, q, rr);   // W_L -> Saa'
    big_schur_run<NS>(a, Fs + NS * LDF, nullptr, EO + D::OAB + r * NS, NT, q, rr);        // W_R -> Sab'
  }
  {  // right half: [Sbb' | sb'] rows of row tile tR; A fragments: W_R(k, r) / d_k
    const int r = 8 * tR + rr;
    double part = 0.0;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      a[ks] = Fs[(NS + r) * LDF + 4 * ks + q] * di[ks];
      part = fma(a[ks], ws[4 * ks + q], part);
    }
    part += __shfl_xor_sync(0xffffffffu, part, 1);
    part += __shfl_xor_sync(0xffffffffu, part, 2);
    if (q == 0) EO[D::OSB + r] = ER[D::OSB + r] - part;
    big_schur_run<NS>(a, Fs + NS * LDF, ER + D::OBB + r * NS, EO + D::OBB + r * NS, tR + 1, q, rr);  // W_R -> Sbb'
  }
}

// ---------------------------------------------------------------------------------------------------------
// Top-down, one launch per level (coarsest first): the multiplier of every separator of that level ([a,b] = the merged
// subtree),   lambda_{i+1} = U^-1 D^-1 (w_s - W_L lambda_a - W_R lambda_{b+1}).
// One CTA of NS threads per (problem, separator): the strict upper triangle of U comes into shared memory by asynchronous
// copies while the threads (one row each) read the record; the back-substitution then runs in one warp, two rows per lane,
// the solved component broadcast by a shuffle per step.
// ---------------------------------------------------------------------------------------------------------
template <int NS>
__global__ void __launch_bounds__(NS) big_topdown_kernel(int N, const double* rec_all, const double* ufac, double* lam_all, int level,
                                                          int M) {
  using D = BigDim<NS>;
  using F = BigFac<NS>;
  static_assert(NS % 32 == 0, "NS / 32 rows per lane in the back-substitution");
  constexpr int RPL = NS / 32;
  constexpr int LDU = NS + 1;
  __shared__ double la[NS], lb[NS], vs[NS], Us[NS * LDU];
  const int prob = blockIdx.x / M, t = blockIdx.x % M;
  const int i_sep = ((2 * t + 1) << level) - 1;
  const int a = i_sep + 1 - (1 << level), b1 = i_sep + 1 + (1 << level);
  double* lam = lam_all + (size_t)prob * (N + 1) * NS;
  const int r = threadIdx.x;
  const double* Uf = ufac + ((size_t)prob * N + i_sep) * F::SIZE;
  for (int e = r; e < NS * NS; e += NS) {
    const int j = e / NS, k = e - j * NS;
    if (k > j) cp_async8(Us + j * LDU + k, Uf + F::OU + e);
  }
  la[r] = a == 0 ? 0.0 : lam[(size_t)a * NS + r];  // (lambda_0 does not couple: W_L is zero there)
  lb[r] = lam[(size_t)b1 * NS + r];                // slot N is zero
  __syncthreads();
  const double* rec = rec_all + ((size_t)prob * N + i_sep) * D::WSZ;
  double v = rec[2 * NS * NS + r];
#pragma unroll 8
  for (int c = 0; c < NS; ++c) v = fma(-rec[c * NS + r], la[c], fma(-rec[(NS + c) * NS + r], lb[c], v));
  vs[r] = v * __ldg(Uf + F::OD + r);
  cp_async_wait_all();
  __syncthreads();
  if (r < 32) {
    double x[RPL];  // rows r, r + 32, ...
#pragma unroll
    for (int i = 0; i < RPL; ++i) x[i] = vs[r + 32 * i];
#pragma unroll 4
    for (int k = NS - 1; k > 0; --k) {
      double own = x[0];
#pragma unroll
      for (int i = 1; i < RPL; ++i) own = k >= 32 * i ? x[i] : own;
      const double xk = __shfl_sync(0xffffffffu, own, k & 31);
#pragma unroll
      for (int i = 0; i < RPL; ++i)
        if (r + 32 * i < k) x[i] = fma(-Us[(r + 32 * i) * LDU + k], xk, x[i]);
    }
#pragma unroll
    for (int i = 0; i < RPL; ++i) lam[(size_t)(i_sep + 1) * NS + 32 * i + r] = x[i];
  }
}

// ---------------------------------------------------------------------------------------------------------
// Down: x_k = Q_k^-1 (rhs_x + lambda_k - A_k^T lambda_{k+1}), u_k = R_k^-1 (rhs_u - B_k^T lambda_{k+1}) from the leaf
// equations (the reference's rhs propagation src/solve.c:175-180 telescoped); knot 0: x_0 = x0 and
// lambda_0 = Q_0 x0 + q_0 + A_0^T lambda_1 (src/nested_dissection.c:24-59). One CTA per (problem, knot), one thread per row.
// ---------------------------------------------------------------------------------------------------------
template <int NS, int NI>
__global__ void __launch_bounds__(256) big_down_kernel(Shape s, Bufs b, const double* lam_all, int pack_rhs) {
  constexpr int XU = NS + NI, CPW = XU / 8, RPL = NS / 32;
  static_assert(NS % 32 == 0 && XU % 8 == 0, "NS / 32 elements of a column per lane; columns dealt to 8 warps");
  __shared__ double dot[XU];
  const int N = s.N, n = s.n, m = s.m;
  const int prob = blockIdx.x / N, k = blockIdx.x % N;
  const int r = threadIdx.x, warp = r >> 5, lane = r & 31;
  const double* lam = lam_all + (size_t)prob * (N + 1) * NS;   // (rows >= n of a multiplier are padding: zero)
  const bool first = k == 0, last = k == N - 1;
  const double* kn = knot_ptr(s, b, prob, k);
  // [A | B]^T lambda_{k+1}: warp w takes columns w, w + 8, ... (slots [0, NS) = columns of A, [NS, XU) = columns of B): a
  // lane reads consecutive-by-lane elements of a column (coalesced), all loads of the warp in flight, then a shuffle
  // reduction per column
  {
    double lv[RPL];
#pragma unroll
    for (int i = 0; i < RPL; ++i) lv[i] = lam[(size_t)(k + 1) * NS + 32 * i + lane];
    double part[CPW];
#pragma unroll
    for (int u = 0; u < CPW; ++u) {
      const int l = warp + 8 * u, lc = l < NS ? l : l - NS;
      const bool valid = !last && (l < NS ? lc < n : lc < m);
      const double* col = kn + (l < NS ? s.oA : s.oB) + (size_t)lc * n;
      double v = 0.0;
#pragma unroll
      for (int i = 0; i < RPL; ++i)
        if (valid && 32 * i + lane < n) v = fma(__ldg(col + 32 * i + lane), lv[i], v);
      part[u] = v;
    }
#pragma unroll
    for (int u = 0; u < CPW; ++u) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) part[u] += __shfl_xor_sync(0xffffffffu, part[u], o);
      if (lane == 0) dot[warp + 8 * u] = part[u];
    }
  }
  __syncthreads();
  if (r >= XU) return;
  const bool isR = r >= NS;
  const int ri = isR ? r - NS : r;
  if (ri >= (isR ? m : n)) return;   // padding
  double* out = soln_ptr(s, b, prob, k);
  const double acc = dot[r];  // row of [A | B]^T lambda_{k+1}
  const double d = kn[isR ? n + ri : ri];
  double rx;
  if (!pack_rhs) rx = out[isR ? 2 * n + ri : n + ri];
  else rx = isR ? (last ? 0.0 : -kn[s.orr + ri]) : -kn[s.oq + ri];
  if (!isR) {
    const double lk = first ? 0.0 : lam[(size_t)k * NS + ri];
    double xv, lv;
    if (first) {
      const double rl = pack_rhs ? -b.x0[(size_t)prob * n + ri] : out[ri];
      xv = -rl;
      lv = -d * rl - rx + acc;
    } else {
      xv = (rx + lk - acc) / d;
      lv = lk;
    }
    out[ri] = lv;
    out[n + ri] = xv;
  } else {
    out[2 * n + ri] = last ? rx : (rx - acc) / d;
  }
}

// ---------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------
struct BigPlan {
  bool covered;    // the shape has these kernels
  bool allocated;
  int ns;          // block size the kernels run at: 32 or 64 (n is padded up to it on chip)
  BigBufs bb;
  size_t bytes;
  int launches;
};

// any n <= 64 with m <= 32: blocks of 32 or 64 (the smallest that holds n), 16 or 32 input columns
inline bool big_shape_ok(int n, int m, int N) { return n >= 1 && n <= 64 && m >= 1 && m <= BIG_NI_MAX && N >= 2; }
inline int big_block_size(int n) { return n <= 32 ? 32 : 64; }
inline size_t big_ent_doubles(int ns) { return (size_t)3 * ns * ns + 2 * ns; }
inline size_t big_rec_doubles(int ns) { return ((size_t)2 * ns + 1) * ns; }
inline size_t big_fac_doubles(int ns) { return (size_t)ns * ns + (ns / 8) * 64 + ns; }

inline void big_free(BigPlan* p) {
  cudaFree(p->bb.entA); cudaFree(p->bb.entB); cudaFree(p->bb.rec); cudaFree(p->bb.lam); cudaFree(p->bb.ufac);
  const bool cov = p->covered;
  const int ns = p->ns;
  memset(p, 0, sizeof(*p));
  p->covered = cov;
  p->ns = ns;
}

inline int big_alloc(const Shape& s, int batch, BigPlan* p, std::string& err) {
  if (p->allocated) return 0;
  p->ns = big_block_size(s.n);
  const size_t ns = (size_t)p->ns, ent = big_ent_doubles(p->ns), wsz = big_rec_doubles(p->ns);
  auto alloc = [&](double** ptr, size_t cnt) {
    p->bytes += cnt * sizeof(double);
    return cudaMalloc((void**)ptr, cnt * sizeof(double)) == cudaSuccess;
  };
  const bool ok = alloc(&p->bb.entA, (size_t)batch * s.N * ent) && alloc(&p->bb.entB, (size_t)batch * (s.N / 2) * ent) &&
                  alloc(&p->bb.rec, (size_t)batch * s.N * wsz) && alloc(&p->bb.lam, (size_t)batch * (s.N + 1) * ns) &&
                  alloc(&p->bb.ufac, (size_t)batch * s.N * big_fac_doubles(p->ns));
  if (!ok) {
    cudaGetLastError();
    big_free(p);
    err = "device allocation for the large-block dual path failed";
    return -4;
  }
  p->allocated = true;
  return 0;
}

// problems [first, first + count) of the handle (the kernels index problems from 0: views start at `first`)
template <int NS, int NI>
inline int big_launch_t(BigPlan& p, const Shape& s, const Bufs& b, int batch, int pack_rhs, cudaStream_t st, std::string& err) {
  const int N = s.N;
  p.launches = 0;
  cudaError_t e = cudaMemsetAsync(p.bb.lam, 0, sizeof(double) * (size_t)batch * (N + 1) * NS, st);
  if (e != cudaSuccess) { err = std::string("big launch: ") + cudaGetErrorString(e); return -3; }
  const int leaf_smem = big_leaf_smem_doubles<NS, NI>() * (int)sizeof(double);
  const int schur_smem = (2 * NS + 1) * (NS + 4) * (int)sizeof(double);
  const int solve_smem = (NS * (NS + 4) + (NS / 8) * 64) * (int)sizeof(double);
  if ((e = smem_optin(big_leaf_kernel<NS, NI>, leaf_smem)) != cudaSuccess || (e = smem_optin(big_solve_kernel<NS>, solve_smem)) != cudaSuccess ||
      (e = smem_optin(big_schur_kernel<NS>, schur_smem)) != cudaSuccess) {
    err = std::string("big launch: ") + cudaGetErrorString(e);
    return -3;
  }
  big_leaf_kernel<NS, NI><<<(unsigned)((size_t)batch * N), 256, leaf_smem, st>>>(s, b, p.bb, pack_rhs);
  ++p.launches;
  const double* ein = p.bb.entA;
  for (int level = 0; level < s.K; ++level) {
    const int M = N >> (level + 1);
    double* eout = (level & 1) ? p.bb.entA : p.bb.entB;
    const long long nmerge = (long long)batch * M;
    big_factor_reg_kernel<NS><<<(unsigned)((nmerge + BIG_FACTOR_REG_WPC - 1) / BIG_FACTOR_REG_WPC), 32 * BIG_FACTOR_REG_WPC, 0, st>>>(
        s, b, ein, p.bb.ufac, level, M, nmerge);
    big_solve_kernel<NS><<<(unsigned)nmerge, big_solve_threads<NS>(), solve_smem, st>>>(s, ein, p.bb.ufac, p.bb.rec, level, M);
    p.launches += 2;
    if (level < s.K - 1) {
      big_schur_kernel<NS><<<(unsigned)nmerge, 4 * NS, schur_smem, st>>>(s, ein, eout, p.bb.rec, p.bb.ufac, level, M);
      ++p.launches;
    }
    ein = eout;
  }
  for (int level = s.K - 1; level >= 0; --level) {
    const int M = N >> (level + 1);
    big_topdown_kernel<NS><<<(unsigned)((size_t)batch * M), NS, 0, st>>>(N, p.bb.rec, p.bb.ufac, p.bb.lam, level, M);
    ++p.launches;
  }
  big_down_kernel<NS, NI><<<(unsigned)((size_t)batch * N), 256, 0, st>>>(s, b, p.bb.lam, pack_rhs);
  ++p.launches;
  e = cudaGetLastError();
  if (e != cudaSuccess) { err = std::string("big launch: ") + cudaGetErrorString(e); return -3; }
  return 0;
}

inline int big_launch(BigPlan& p, const Shape& s, const Bufs& b, int batch, int pack_rhs, cudaStream_t st, std::string& err) {
  if (s.m <= 16)
    return p.ns == 32 ? big_launch_t<32, 16>(p, s, b, batch, pack_rhs, st, err) : big_launch_t<64, 16>(p, s, b, batch, pack_rhs, st, err);
  return p.ns == 32 ? big_launch_t<32, 32>(p, s, b, batch, pack_rhs, st, err) : big_launch_t<64, 32>(p, s, b, batch, pack_rhs, st, err);
}

}  // namespace rslqr

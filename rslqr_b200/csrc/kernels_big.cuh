// kernels_big.cuh -- the dual-form nested dissection (see kernels_dual.cuh for the formulation) for LARGE blocks,
// n = 64, m = 16 (BASELINE configs[4]): an element [Saa | Sab | Sbb | sa | sb] is 99 KB and a merge touches three of
// them plus a 66 KB record, so elements live in HBM / L2 and a CTA (not a warp) owns a merge.
//
// Per level, two launches over all (problem, merge) pairs:
//   big_spine_kernel : S = SbbL + SaaR, then ONE elimination sweep with the 2n+1 right-hand sides riding along as extra
//                      rows, one row per THREAD in registers (rows of S | columns of S_L | columns of S_R | s; the same
//                      scheme as spine_core2, with the pivot rows broadcast through shared memory), then the backward
//                      sweep; [f_L | f_R | zs] goes straight from registers to the separator's record.
//                      (= ndlqr_FactorInnerProduct + Cholesky + ndlqr_SolveCholeskyFactor, src/nested_dissection.c:114-152)
//   big_schur_kernel : Saa' = SaaL - SabL f_L, Sab' = -SabL f_R, sa' = saL - SabL zs,
//                      Sbb' = SbbR - SabR^T f_R, sb' = sbR - SabR^T zs      on mma.sync.m8n8k4.f64
//                      (= ndlqr_UpdateShurFactor restricted to what a later inner product reads, :154-171)
// then one top-down launch per level (lambda of the separators from the records) and big_down_kernel (x_k, u_k from the
// leaf equations, as scan_down_kernel). Against the staged path on this shape (every knot updated at every level, dense
// `fact` array of 755 MB per problem): 3.4 instead of 23.6 GFLOP per solve, 0.22 instead of 0.76 GB per problem.
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
  static constexpr int WSZ = (2 * NS + 1) * NS;  // record [f_L | f_R | zs], column-major ld NS
};

struct BigBufs {
  double* entA;  // [batch][N][ENT]      elements of the even levels (level l has N >> l of them, packed at the front)
  double* entB;  // [batch][N/2][ENT]    elements of the odd levels
  double* rec;   // [batch][N][WSZ]      record of the separator after knot i at slot i (slot N-1 unused)
  double* lam;   // [batch][N+1][NS]     multipliers lambda_k; slots 0 and N stay zero
  double* ufac;  // [batch][N/2][BigFac::SIZE] factor U, block inverses and 1/d of the merges of the level in flight
};

// ---------------------------------------------------------------------------------------------------------
// Leaf elements (ndlqr_SolveLeaf on diagonal Q, R, src/nested_dissection.c:10-105, with the rhs packing of
// src/solver.c:138-190), one CTA per (problem, knot):
//   Saa = Q^-1, Sab = -Q^-1 A^T, Sbb = A Q^-1 A^T + B R^-1 B^T, sa = -z_x - z_lambda, sb = A z_x + B z_u
//   knot 0: x_0 is fixed (z_x = x0) and lambda_0 is no unknown: Q^-1 -> 0 in the blocks; last knot: only Saa, sa.
// dynamic smem: ((NS+8)*(NS+NI) + 2*(NS+NI) + 2*NS + 3*NS) doubles
// ---------------------------------------------------------------------------------------------------------
template <int NS, int NI>
__global__ void __launch_bounds__(256, 2) big_leaf_kernel(Shape s, Bufs b, BigBufs bb, int pack_rhs, int sym) {
  using D = BigDim<NS>;
  constexpr int XU = NS + NI, LDA = NS + 8;  // (LDA = 8 mod 16: the DMMA fragments of a warp fall into two wavefronts)
  static_assert(NS == 64 && XU % 4 == 0, "8 warps x 8 rows; k-steps of 4");
  extern __shared__ __align__(16) double smf[];
  double* AB = smf;                 // [NS x XU] column-major with leading dimension LDA
  double* dv = AB + LDA * XU;       // 1/Q, 1/R (as used in the products: 0 for Q of knot 0 and for the last knot)
  double* zv = dv + XU;             // leaf-solved z_x, z_u
  double* sc = zv + XU;             // 1/Q of this knot (also for knot 0), then z_lambda
  const int N = s.N;
  const int prob = blockIdx.x / N, k = blockIdx.x % N;
  const int tid = threadIdx.x, nthr = blockDim.x;
  const double* kn = knot_ptr(s, b, prob, k);
  const bool first = k == 0, last = k == N - 1;
  // [A | B] -> shared memory by 8-byte asynchronous copies (the knot block is only 8-byte aligned), in flight while the
  // diagonal / right-hand-side part below is computed
  for (int e = tid; e < NS * XU; e += nthr) {
    const int l = e / NS, r = e - l * NS;
    if (last) AB[r + LDA * l] = 0.0;
    else cp_async8(AB + r + LDA * l, kn + s.oA + e);
  }
  for (int l = tid; l < XU; l += nthr) {
    const bool isR = l >= NS;
    const double d = kn[l];  // Q then R lead the block
    if (!(d > 0.0) && !(isR && last)) {
      b.status[(size_t)prob * 3 * N + (isR ? rs_slot_R(k) : rs_slot_Q(k))] = -1;
      b.flags[prob] = 1;
    }
    const double dinv = (isR && last) ? 0.0 : 1.0 / d;
    double rx;  // negated rhs of row l of [x; u]
    if (!pack_rhs) rx = soln_ptr(s, b, prob, k)[NS + l];
    else rx = isR ? (last ? 0.0 : -kn[s.orr + l - NS]) : -kn[s.oq + l];
    double z;
    if (first && !isR) {
      const double rl = pack_rhs ? -b.x0[(size_t)prob * NS + l] : soln_ptr(s, b, prob, k)[l];
      z = -rl;  // z_x(0) = x0
    } else {
      z = (isR && last) ? rx : rx * dinv;
    }
    zv[l] = z;
    dv[l] = (last || (first && !isR)) ? 0.0 : dinv;
    if (!isR) {
      sc[l] = first ? 0.0 : dinv;
      double zl = 0.0;  // z_lambda(k) = negated rhs of the lambda row (unused for knot 0)
      if (!first) zl = pack_rhs ? -(kn - s.stride)[s.od + l] : soln_ptr(s, b, prob, k)[l];
      sc[NS + l] = zl;
    }
  }
  cp_async_wait_all();
  __syncthreads();
  double* E = bb.entA + ((size_t)prob * N + k) * D::ENT;
  // sym: the readers of Saa, Sbb (factorization, keep part of the Schur update) take the lower block triangle only
  // (8 x 8 tiles with column tile <= row tile): the tiles above it are neither computed nor written
  for (int e = tid; e < NS * NS; e += nthr) {
    const int r = e / NS, c = e - r * NS;
    if (!sym || (c >> 3) <= (r >> 3)) E[D::OAA + e] = r == c ? sc[r] : 0.0;
    E[D::OAB + e] = -sc[r] * AB[c + LDA * r];  // -A(c, r) / Q(r)   (AB is zero for the last knot)
  }
  // Sbb = [A | B] D [A | B]^T on the fp64 tensor cores, 8 x 8 output tiles. sym: the 36 tiles of the lower block triangle,
  // dealt round-robin to the 8 warps (4 or 5 each, all in flight); else all 64 tiles, 8 per warp in two groups of 4.
  {
    const int warp = tid >> 5, lane = tid & 31, q = lane & 3, rr = lane >> 2;
    constexpr int KS = XU / 4, NT = NS / 8, G = 5;
    const int ntask = sym ? NT * (NT + 1) / 2 : NT * NT;
#pragma unroll 1
    for (int i0 = warp; i0 < ntask; i0 += 8 * G) {
      int tr[G], tc[G];
      double c[G][2];
#pragma unroll
      for (int u = 0; u < G; ++u) {
        const int i = i0 + 8 * u;
        int t = 0, jj = i;
        if (sym) { while (jj > t) { ++t; jj -= t; } }   // i = t (t + 1) / 2 + jj, jj <= t
        else { t = i / NT; jj = i - t * NT; }
        tr[u] = i < ntask ? t : 0; tc[u] = i < ntask ? jj : 0;
        c[u][0] = c[u][1] = 0.0;
      }
      const int nact = (ntask - i0 + 7) / 8;  // (warp-uniform) tasks of this group
#pragma unroll 4
      for (int ks = 0; ks < KS; ++ks) {
        const double dk = dv[4 * ks + q];
#pragma unroll
        for (int u = 0; u < G; ++u)
          if (u < nact)
            dmma884(c[u][0], c[u][1], AB[8 * tr[u] + rr + LDA * (4 * ks + q)] * dk, AB[8 * tc[u] + rr + LDA * (4 * ks + q)]);
      }
#pragma unroll
      for (int u = 0; u < G; ++u)
        if (u < nact) *reinterpret_cast<double2*>(E + D::OBB + (8 * tr[u] + rr) * NS + 8 * tc[u] + 2 * q) = make_double2(c[u][0], c[u][1]);
    }
  }
  {  // sb = [A | B] z: four threads per row, a quarter of the XU terms each (thread = 64 part + row)
    const int r = tid & (NS - 1), part = tid / NS;
    double v = 0.0;
#pragma unroll 4
    for (int l = part * (XU / 4); l < (part + 1) * (XU / 4); ++l) v = fma(AB[r + LDA * l], zv[l], v);
    double* ps = sc + 2 * NS;  // partial sums
    if (part > 0) ps[(part - 1) * NS + r] = v;
    __syncthreads();
    if (part == 0) {
      E[D::OSA + r] = -zv[r] - sc[NS + r];
      E[D::OSB + r] = v + ps[r] + ps[NS + r] + ps[2 * NS + r];
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// Spine of one merge, one CTA of 3 NS + 1 (rounded up to warps) threads, one row of NS doubles per thread in registers:
//   threads [0,NS)      : row t of S = SbbL + SaaR
//   threads [NS,2NS)    : column c of S_L = row c of SabL
//   threads [2NS,3NS)   : column c of S_R = column c of SabR
//   thread  3NS         : s = sbL + saR
// Step j: thread j scales its row by 1/d_j and publishes [1/d_j | U(j, j+1..)] in shared memory; everybody eliminates
// with its own unscaled element (an appended row ends up holding D^-1 Lt^-1 y). The backward sweep x_j = y_j -
// sum_{k>j} U(j,k) x_k reads the same rows. Returns the solved columns straight into the record.
// dynamic smem: NS * (NS + 2) doubles.
// ---------------------------------------------------------------------------------------------------------
template <int NS>
__host__ __device__ constexpr int big_spine_threads() { return ((3 * NS + 1 + 31) / 32) * 32; }

template <int NS>
__global__ void __launch_bounds__(big_spine_threads<NS>()) big_spine_kernel(Shape s, Bufs b, const double* ein, double* rec_all,
                                                                            int level, int M) {
  using D = BigDim<NS>;
  constexpr int LD = NS + 2;
  extern __shared__ __align__(16) double Ur[];
  __shared__ int s_bad;
  const int N = s.N;
  const int prob = blockIdx.x / M, t = blockIdx.x % M;
  const int tid = threadIdx.x;
  const int Min = 2 * M;                                   // elements of this level per problem
  const double* EL = ein + ((size_t)prob * Min + 2 * t) * D::ENT;
  const double* ER = EL + D::ENT;
  const int i_sep = ((2 * t + 1) << level) - 1;            // knot index of the separator
  double row[NS];
  if (tid == 0) s_bad = 0;
  if (tid < NS) {
    const double2* pa = reinterpret_cast<const double2*>(EL + D::OBB + tid * NS);
    const double2* pb = reinterpret_cast<const double2*>(ER + D::OAA + tid * NS);
#pragma unroll
    for (int k = 0; k < NS; k += 2) {
      const double2 a = __ldg(pa + k / 2), c = __ldg(pb + k / 2);
      row[k] = a.x + c.x; row[k + 1] = a.y + c.y;
    }
  } else if (tid < 2 * NS) {
    const double2* pa = reinterpret_cast<const double2*>(EL + D::OAB + (tid - NS) * NS);
#pragma unroll
    for (int k = 0; k < NS; k += 2) { const double2 a = __ldg(pa + k / 2); row[k] = a.x; row[k + 1] = a.y; }
  } else if (tid < 3 * NS) {
    const double* pc = ER + D::OAB + (tid - 2 * NS);
#pragma unroll
    for (int k = 0; k < NS; ++k) row[k] = __ldg(pc + k * NS);
  } else {
#pragma unroll
    for (int k = 0; k < NS; ++k) row[k] = tid == 3 * NS ? __ldg(EL + D::OSB + k) + __ldg(ER + D::OSA + k) : 0.0;
  }
#pragma unroll
  for (int j = 0; j < NS; ++j) {
    const int j0 = j & ~1;
    if (tid == j) {
      const double d = row[j];
      if (!(d > 0.0)) s_bad = 1;
      const double dinv = 1.0 / d;
#pragma unroll
      for (int k = j0; k < NS; k += 2) {
        const double u0 = k < j ? 0.0 : (k == j ? dinv : row[k] * dinv);
        const double u1 = k + 1 == j ? dinv : row[k + 1] * dinv;
        *reinterpret_cast<double2*>(Ur + j * LD + k) = make_double2(u0, u1);
      }
    }
    __syncthreads();
    const double m1 = row[j];
    double dinv = 0.0;
#pragma unroll
    for (int k = j0; k < NS; k += 2) {
      const double2 u = *reinterpret_cast<const double2*>(Ur + j * LD + k);
      if (k == j) dinv = u.x;
      else if (k > j) row[k] = fma(-m1, u.x, row[k]);
      if (k + 1 == j) dinv = u.y;
      else if (k + 1 > j) row[k + 1] = fma(-m1, u.y, row[k + 1]);
    }
    row[j] = m1 * dinv;
  }
  if (tid >= NS && tid <= 3 * NS) {
#pragma unroll
    for (int j = NS - 2; j >= 0; --j) {
      const int k0 = (j + 1) & ~1;
#pragma unroll
      for (int k = NS - 2; k >= k0; k -= 2) {  // the latest x last: one dependent FMA per step
        const double2 u = *reinterpret_cast<const double2*>(Ur + j * LD + k);
        row[j] = fma(-u.y, row[k + 1], row[j]);
        if (k > j) row[j] = fma(-u.x, row[k], row[j]);
      }
    }
    double* rec = rec_all + ((size_t)prob * N + i_sep) * D::WSZ + (size_t)(tid - NS) * NS;  // column tid - NS of [f_L | f_R | zs]
#pragma unroll
    for (int k = 0; k < NS; k += 2) *reinterpret_cast<double2*>(rec + k) = make_double2(row[k], row[k + 1]);
  }
  __syncthreads();
  if (tid == 0 && s_bad) {
    b.status[(size_t)prob * 3 * N + rs_slot_S(t, level, N)] = -1;
    b.flags[prob] = 1;
  }
}

// ---------------------------------------------------------------------------------------------------------
// The same spine as big_spine_kernel, split so that the 2n+1 solves run on the fp64 tensor cores (RSLQR_BIG_SPINE=1
// selects the single-kernel variant above for A/B runs):
//
//   big_factor_kernel : ONE WARP per merge. S = SbbL + SaaR is built in shared memory and factored in place,
//                       S = U^T D U (right-looking LDL^T, the warp's lanes own the rows below the pivot; no CTA barrier,
//                       so the pivot chains of the merges of a CTA overlap). Output per merge (`ufac`): U (unit upper,
//                       row-major ld NS), the inverses X_b of the 8 x 8 diagonal blocks of L = U^T, and 1/d.
//   big_solve_kernel  : one CTA of 8 warps per merge. Y = [S_L | S_R | s] (NS x 2NS+1, padded to 17 column tiles of 8) and U
//                       are staged in shared memory; a warp owns column tiles and runs BOTH blocked substitution sweeps
//                       for them without any CTA barrier: for block row b,
//                           W_b = X_b (Y_b - sum_{b' < b} L_bb' W_b'),   then   F_b = X_b^T (D^-1 W_b - sum_{b' > b} U_bb' F_b')
//                       every product an m8n8k4 DMMA pair (the 8 x 8 blocks are the A operands, the column tile of Y the B
//                       operand, read back from shared memory in B-fragment layout). [f_L | f_R | zs] = F goes to the record.
// ---------------------------------------------------------------------------------------------------------
template <int NS>
struct BigFac {
  static constexpr int NB = NS / 8;
  static constexpr int OU = 0, OX = NS * NS, OD = NS * NS + NB * 64, SIZE = NS * NS + NB * 64 + NS;  // doubles per merge in `ufac`
};

constexpr int BIG_FACTOR_WPC = 12;  // merges (warps) per CTA of big_factor_kernel

// packed lower triangle: element (i, k), k <= i, at i (i + 1) / 2 + k
__device__ __forceinline__ int big_tri(int i) { return (i * (i + 1)) >> 1; }

template <int NS>
__host__ __device__ constexpr int big_factor_warp_doubles() { return ((NS * (NS + 1) / 2) + 1) & ~1; }

template <int NS>
__global__ void __launch_bounds__(32 * BIG_FACTOR_WPC) big_factor_kernel(Shape s, Bufs b, const double* ein, double* ufac, int level,
                                                                          int M, long long nmerge) {
  using D = BigDim<NS>;
  using F = BigFac<NS>;
  static_assert(NS == 64, "two trailing columns per lane");
  extern __shared__ __align__(16) double smf[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long mi = (long long)blockIdx.x * BIG_FACTOR_WPC + warp;
  if (mi >= nmerge) return;
  const int N = s.N;
  const int prob = (int)(mi / M), t = (int)(mi % M);
  double* A = smf + (size_t)warp * big_factor_warp_doubles<NS>();   // lower triangle of S, packed by rows
  const double* EL = ein + ((size_t)prob * 2 * M + 2 * t) * D::ENT;
  const double* ER = EL + D::ENT;
  for (int e = lane; e < NS * NS; e += 32) {
    const int r = e / NS, c = e - r * NS;
    if (c <= r) A[big_tri(r) + c] = __ldg(EL + D::OBB + e) + __ldg(ER + D::OAA + e);
  }
  __syncwarp();
  double* U = ufac + (size_t)mi * F::SIZE;
  int bad = 0;
  // right-looking LDL^T on the lower triangle: after step j column j holds the multipliers l(i,j) = S'(i,j) / d_j.
  // A lane owns the trailing COLUMNS k0 = j + 1 + lane and k1 = k0 + 32: the update S'(i,k) -= l(i,j) S'(k,j) walks down
  // the rows i >= k with S'(k,j) in a register, l(i,j) a broadcast load and S'(i, k0..) one contiguous run per row.
#pragma unroll 1
  for (int j = 0; j < NS; ++j) {
    const double d = A[big_tri(j) + j];
    if (!(d > 0.0)) bad = 1;
    const double dinv = 1.0 / d;
    const int k0 = j + 1 + lane, k1 = k0 + 32;
    const bool two = j + 33 < NS;                                    // (warp-uniform) a second column per lane exists
    const double c0 = k0 < NS ? A[big_tri(k0) + j] : 0.0, c1 = k1 < NS ? A[big_tri(k1) + j] : 0.0;  // unscaled S'(k, j)
    __syncwarp();
    if (k0 < NS) A[big_tri(k0) + j] = c0 * dinv;
    if (k1 < NS) A[big_tri(k1) + j] = c1 * dinv;
    if (lane == 0) A[big_tri(j) + j] = dinv;
    __syncwarp();
#pragma unroll 1
    for (int i = j + 1; i < NS; i += 4) {
      double li[4], a0[4], a1[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {  // all loads of the group before any store (the compiler cannot tell they do not alias)
        const int r = i + u < NS ? i + u : NS - 1;
        const int base = big_tri(r);
        li[u] = A[base + j];
        a0[u] = k0 <= r ? A[base + k0] : 0.0;
        a1[u] = (two && k1 <= r) ? A[base + k1] : 0.0;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int r = i + u;
        const int base = big_tri(r < NS ? r : NS - 1);
        if (r < NS && k0 <= r) A[base + k0] = fma(-li[u], c0, a0[u]);
        if (two && r < NS && k1 <= r) A[base + k1] = fma(-li[u], c1, a1[u]);
      }
    }
    __syncwarp();
  }
  // U(j,k) = l(k,j) for k > j (row-major, ld NS), unit diagonal implied; 1/d
  for (int e = lane; e < NS * NS; e += 32) {
    const int j = e / NS, k = e - j * NS;
    U[F::OU + e] = k > j ? A[big_tri(k) + j] : (k == j ? 1.0 : 0.0);
  }
  for (int j = lane; j < NS; j += 32) U[F::OD + j] = A[big_tri(j) + j];
  // X_b = inverse of the unit lower 8 x 8 diagonal block of L = U^T: lane = (4 blocks per round) x 8 columns
#pragma unroll 1
  for (int round = 0; round < F::NB / 4; ++round) {
    const int blk = 4 * round + (lane >> 3), c = lane & 7;
    double x[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      double v = r == c ? 1.0 : 0.0;
#pragma unroll
      for (int k = 0; k < 8; ++k)
        if (k < r) v = fma(-A[big_tri(8 * blk + r) + 8 * blk + k], x[k], v);   // T(r, k) = l(8blk + r, 8blk + k)
      x[r] = r < c ? 0.0 : v;
    }
#pragma unroll
    for (int r = 0; r < 8; ++r) U[F::OX + blk * 64 + r * 8 + c] = x[r];
  }
  if (bad && lane == 0) {
    b.status[(size_t)prob * 3 * N + rs_slot_S(t, level, N)] = -1;
    b.flags[prob] = 1;
  }
}

// ---------------------------------------------------------------------------------------------------------
// big_factor_reg_kernel: the same outputs as big_factor_kernel, with the matrix held in REGISTERS as DMMA accumulator
// fragments (lower block triangle: 36 blocks of 8 x 8, two doubles per lane each) and a blocked right-looking LDL^T:
//   for block column b:  S_bb = L_bb D_b L_bb^T by row elimination on [S_bb | I] inside the warp (shuffles; gives
//                         X_b = L_bb^-1 directly),
//                        panel  R_c = S_cb X_b^T (= L_cb D_b),  L_cb = R_c D_b^-1          (2 DMMAs per block)
//                        trailing S_cc' -= L_cb (L_c'b D_b)^T                                (2 DMMAs per block)
// C fragments become A / B operands through four 64-bit shuffles (frag_c2a). One warp per merge, no shared memory, no
// barrier; about 2.5 k warp instructions per matrix against 78 k for the packed-triangle kernel above.
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
  double* U = ufac + (size_t)mi * F::SIZE;
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
      if (q == 0) Ud[(size_t)j * NS + g] = g > j ? l : (g == j ? 1.0 : 0.0);  // U(8b + j, 8b + g) = l(g, j)
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
        // U(8b + k, 8c + g) = L_cb(g, k); the block below the diagonal of U is zero
        U[F::OU + (size_t)(8 * bk + 2 * q) * NS + 8 * c + g] = r0 * dq0;
        U[F::OU + (size_t)(8 * bk + 2 * q + 1) * NS + 8 * c + g] = r1 * dq1;
        *reinterpret_cast<double2*>(U + F::OU + (size_t)(8 * c + g) * NS + 8 * bk + 2 * q) = make_double2(0.0, 0.0);
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

// Both substitution sweeps for T column tiles at once, RIGHT-LOOKING: the 8 block rows of a tile stay in registers as
// DMMA accumulators (C fragments); once block row b is final it goes through shared memory once (C layout -> B
// fragment) and every other block row still to be solved is updated with it, so a B fragment is loaded once per block
// and an A fragment (8 x 8 block of U) once per DMMA pair of T tiles. The solved tile leaves from registers straight
// to the record.
//   forward : W_b = X_b Y_b, then Y_c -= L_cb W_b for c > b          (L_cb (r, k) = U(8b + k, 8c + r))
//   scaling : V = D^-1 W
//   backward: F_b = X_b^T V_b, then V_c -= U_cb F_b for c < b
template <int NS, int T>
__device__ __forceinline__ void big_solve_tiles(const double* Us, const double* Xs, const double* ds, double* const (&Yt)[T],
                                                double* const (&rec)[T], const int (&ncols)[T]) {
  constexpr int NB = NS / 8, LDU = NS + 4, LDY = 2 * NS + 8;
  const int lane = threadIdx.x & 31, q = lane & 3, rr = lane >> 2;
  double c[NB][T][2];
#pragma unroll
  for (int bk = 0; bk < NB; ++bk)
#pragma unroll
    for (int u = 0; u < T; ++u) {
      const double2 cv = *reinterpret_cast<const double2*>(Yt[u] + (8 * bk + rr) * LDY + 2 * q);
      c[bk][u][0] = cv.x; c[bk][u][1] = cv.y;
    }
  __syncwarp();
  // C fragment of block row `bk` -> shared memory -> product with an 8 x 8 block (A fragments a0, a1) -> C fragment
  auto through = [&](int bk, double (&v)[T][2], double a0, double a1) {
#pragma unroll
    for (int u = 0; u < T; ++u) *reinterpret_cast<double2*>(Yt[u] + (8 * bk + rr) * LDY + 2 * q) = make_double2(v[u][0], v[u][1]);
    __syncwarp();
#pragma unroll
    for (int u = 0; u < T; ++u) {
      const double b0 = Yt[u][(8 * bk + q) * LDY + rr], b1 = Yt[u][(8 * bk + 4 + q) * LDY + rr];
      v[u][0] = v[u][1] = 0.0;
      dmma884(v[u][0], v[u][1], a0, b0);
      dmma884(v[u][0], v[u][1], a1, b1);
    }
    __syncwarp();
  };
  // block row `bk` (final) as B fragments
  auto as_b = [&](int bk, const double (&v)[T][2], double (&bf)[T][2]) {
#pragma unroll
    for (int u = 0; u < T; ++u) *reinterpret_cast<double2*>(Yt[u] + (8 * bk + rr) * LDY + 2 * q) = make_double2(v[u][0], v[u][1]);
    __syncwarp();
#pragma unroll
    for (int u = 0; u < T; ++u) { bf[u][0] = Yt[u][(8 * bk + q) * LDY + rr]; bf[u][1] = Yt[u][(8 * bk + 4 + q) * LDY + rr]; }
    __syncwarp();
  };
#pragma unroll
  for (int bk = 0; bk < NB; ++bk) {
    through(bk, c[bk], Xs[bk * 64 + rr * 8 + q], Xs[bk * 64 + rr * 8 + 4 + q]);  // W_b = X_b Y_b
    if (bk + 1 < NB) {
      double bf[T][2];
      as_b(bk, c[bk], bf);
#pragma unroll
      for (int b2 = bk + 1; b2 < NB; ++b2) {
        const double a0 = -Us[(8 * bk + q) * LDU + 8 * b2 + rr], a1 = -Us[(8 * bk + 4 + q) * LDU + 8 * b2 + rr];
#pragma unroll
        for (int u = 0; u < T; ++u) { dmma884(c[b2][u][0], c[b2][u][1], a0, bf[u][0]); dmma884(c[b2][u][0], c[b2][u][1], a1, bf[u][1]); }
      }
    }
  }
#pragma unroll
  for (int bk = 0; bk < NB; ++bk) {  // V = D^-1 W
    const double dr = ds[8 * bk + rr];
#pragma unroll
    for (int u = 0; u < T; ++u) { c[bk][u][0] *= dr; c[bk][u][1] *= dr; }
  }
#pragma unroll
  for (int bk = NB - 1; bk >= 0; --bk) {
    through(bk, c[bk], Xs[bk * 64 + q * 8 + rr], Xs[bk * 64 + (4 + q) * 8 + rr]);  // F_b = X_b^T V_b
    // the solved block row: straight from the C fragment to the record (column-major, ld NS: 8 consecutive k per column)
#pragma unroll
    for (int u = 0; u < T; ++u) {
      if (2 * q < ncols[u]) rec[u][(size_t)(2 * q) * NS + 8 * bk + rr] = c[bk][u][0];
      if (2 * q + 1 < ncols[u]) rec[u][(size_t)(2 * q + 1) * NS + 8 * bk + rr] = c[bk][u][1];
    }
    if (bk > 0) {
      double bf[T][2];
      as_b(bk, c[bk], bf);
#pragma unroll
      for (int b2 = 0; b2 < bk; ++b2) {
        const double a0 = -Us[(8 * b2 + rr) * LDU + 8 * bk + q], a1 = -Us[(8 * b2 + rr) * LDU + 8 * bk + 4 + q];
#pragma unroll
        for (int u = 0; u < T; ++u) { dmma884(c[b2][u][0], c[b2][u][1], a0, bf[u][0]); dmma884(c[b2][u][0], c[b2][u][1], a1, bf[u][1]); }
      }
    }
  }
}

constexpr int BIG_SOLVE_THREADS = 288;  // 9 warps: 8 take two column tiles each, the ninth the tile with the zs column

template <int NS>
__global__ void __launch_bounds__(BIG_SOLVE_THREADS) big_solve_kernel(Shape s, const double* ein, const double* ufac, double* rec_all,
                                                                       int level, int M) {
  using D = BigDim<NS>;
  using F = BigFac<NS>;
  static_assert(NS == 64, "8 block rows, 17 column tiles over 9 warps");
  constexpr int NB = NS / 8, LDU = NS + 4, LDY = 2 * NS + 8;
  extern __shared__ __align__(16) double smf[];
  double* Us = smf;                      // U, row-major ld LDU
  double* Ys = Us + NS * LDU;            // Y / W / F, row k at Ys + k * LDY, column = right-hand side
  double* Xs = Ys + NS * LDY;            // X_b, 8 x 8 row-major each
  double* ds = Xs + NB * 64;             // 1/d
  const int N = s.N;
  const int prob = blockIdx.x / M, t = blockIdx.x % M;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, nthr = blockDim.x;
  const double* EL = ein + ((size_t)prob * 2 * M + 2 * t) * D::ENT;
  const double* ER = EL + D::ENT;
  const double* Uf = ufac + ((size_t)prob * M + t) * F::SIZE;
  const int i_sep = ((2 * t + 1) << level) - 1;
  // staging by asynchronous copies: everything in flight at once, no staging registers
  for (int e = 2 * tid; e < NS * NS; e += 2 * nthr) {
    const int r = e / NS, c = e - r * NS;
    cp_async16(Us + r * LDU + c, Uf + F::OU + e);
    cp_async16(Ys + r * LDY + NS + c, ER + D::OAB + e);   // S_R(k, c) = SabR(k, c)
  }
  // S_L(k, c) = SabL(c, k): a warp moves tiles of 8 k x 4 c (lane = 4 kk + cc): 64-byte runs of a SabL row in, a 4-way
  // conflicted store out (a plain transpose would be 16-way with this row stride)
  for (int tile = warp; tile < (NS / 8) * (NS / 4); tile += nthr >> 5) {
    const int k = 8 * (tile % (NS / 8)) + (lane >> 2), c = 4 * (tile / (NS / 8)) + (lane & 3);
    cp_async8(Ys + k * LDY + c, EL + D::OAB + c * NS + k);
  }
  for (int e = 2 * tid; e < NB * 64 + NS; e += 2 * nthr) cp_async16(Xs + e, Uf + F::OX + e);  // X blocks, then 1/d (contiguous)
  for (int e = tid; e < NS * 8; e += nthr) {
    const int k = e >> 3, c = e & 7;
    Ys[k * LDY + 2 * NS + c] = c == 0 ? __ldg(EL + D::OSB + k) + __ldg(ER + D::OSA + k) : 0.0;
  }
  cp_async_wait_all();
  __syncthreads();
  double* rec = rec_all + ((size_t)prob * N + i_sep) * D::WSZ;  // [f_L | f_R | zs], column-major ld NS
  if (warp < 8) {
    double* const Yt[2] = {Ys + 8 * warp, Ys + 8 * (warp + 8)};
    double* const rc[2] = {rec + (size_t)8 * warp * NS, rec + (size_t)8 * (warp + 8) * NS};
    const int nc[2] = {8, 8};
    big_solve_tiles<NS, 2>(Us, Xs, ds, Yt, rc, nc);
  } else {
    double* const Yt[1] = {Ys + 8 * 16};
    double* const rc[1] = {rec + (size_t)8 * 16 * NS};
    const int nc[1] = {1};
    big_solve_tiles<NS, 1>(Us, Xs, ds, Yt, rc, nc);
  }
}

// ---------------------------------------------------------------------------------------------------------
// Schur update of one merge on the fp64 tensor cores, one CTA of 8 warps. F = [f_L | f_R | zs] is staged in shared
// memory (ld NS + 4: the B fragments of a warp fall into two wavefronts); warp w owns rows 8w..8w+7 of the left half
// (SabL rows -> [Saa' | Sab' | sa']) and rows 8w..8w+7 of the right half (SabR^T rows -> [Sbb' | sb']); its A fragments
// stay in registers across all column tiles. NS = 64: 8 warps x 8 rows.
// dynamic smem: (2 NS + 1) * (NS + 4) doubles.
// ---------------------------------------------------------------------------------------------------------
// One run of column tiles of the Schur update for the 8 rows of a warp: out(r, 8j..8j+7) = keep(r, .) - a(r, :) F(:, fcol0 + 8j..)
// for j < cnt, four tiles in flight, the keep part requested one group ahead of its use. a: A fragments of the rows
// (registers); Fcols: F column fcol0 in shared memory (ld LDF); keep_row: row r of the block that is updated, nullptr =
// zero; dst_row: row r of the output block.
template <int NS>
__device__ __forceinline__ void big_schur_run(const double (&a)[NS / 4], const double* Fcols, const double* keep_row, double* dst_row,
                                              int cnt, int q, int rr) {
  constexpr int LDF = NS + 4, KS = NS / 4;
  auto keep = [&](int nt0, double2 (&v)[4]) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      v[u] = make_double2(0.0, 0.0);
      if (keep_row && nt0 + u < cnt) v[u] = __ldg(reinterpret_cast<const double2*>(keep_row + 8 * (nt0 + u) + 2 * q));
    }
  };
  double2 kv[4];
  keep(0, kv);
#pragma unroll 1
  for (int nt0 = 0; nt0 < cnt; nt0 += 4) {
    double c[4][2];
#pragma unroll
    for (int u = 0; u < 4; ++u) { c[u][0] = -kv[u].x; c[u][1] = -kv[u].y; }  // accumulate a.F - keep, negate on the way out
    if (nt0 + 4 < cnt) keep(nt0 + 4, kv);
    const int nact = cnt - nt0;  // (warp-uniform) tiles of this group: 1..4
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (u < nact) dmma884(c[u][0], c[u][1], a[ks], Fcols[(8 * (nt0 + u) + rr) * LDF + 4 * ks + q]);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (u < nact) *reinterpret_cast<double2*>(dst_row + 8 * (nt0 + u) + 2 * q) = make_double2(-c[u][0], -c[u][1]);
  }
}

// sym: Saa' and Sbb' are symmetric and every reader (the factorization, this kernel's keep part) takes the lower block
// triangle only, so only the column tiles up to the diagonal one are computed: 136 instead of 192 tiles per merge. Warp w
// takes row tile w of the left half and row tile 7 - w of the right half: 17 tiles each. sym = 0 (the single-kernel spine
// variant reads full rows): all tiles, row tile w in both halves.
template <int NS>
__global__ void __launch_bounds__(256, 3) big_schur_kernel(Shape s, const double* ein, double* eout, const double* rec_all, int level,
                                                         int M, int sym) {
  using D = BigDim<NS>;
  static_assert(NS == 64, "8 warps x 8 rows per half");
  constexpr int LDF = NS + 4, KS = NS / 4, NT = NS / 8;
  extern __shared__ __align__(16) double Fs[];  // column c of F at Fs + c * LDF
  const int N = s.N;
  const int prob = blockIdx.x / M, t = blockIdx.x % M;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, q = lane & 3, rr = lane >> 2;
  const int Min = 2 * M;
  const double* EL = ein + ((size_t)prob * Min + 2 * t) * D::ENT;
  const double* ER = EL + D::ENT;
  double* EO = eout + ((size_t)prob * M + t) * D::ENT;
  const int i_sep = ((2 * t + 1) << level) - 1;
  const double* rec = rec_all + ((size_t)prob * N + i_sep) * D::WSZ;
  for (int e = tid; e < D::WSZ / 2; e += blockDim.x) {
    const int c = (2 * e) / NS, k = 2 * e - c * NS;
    cp_async16(Fs + c * LDF + k, rec + 2 * e);
  }
  const int tL = warp, tR = sym ? NT - 1 - warp : warp;  // row tiles of this warp in the left / right half
  // A fragments: left half: SabL(r, k); right half: SabR(k, r). Those of the left half are requested while F is in flight.
  double a[KS];
  {
    const int r = 8 * tL + rr;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) a[ks] = __ldg(EL + D::OAB + r * NS + 4 * ks + q);
  }
  cp_async_wait_all();
  __syncthreads();
  const double* zs = Fs + 2 * NS * LDF;
  {  // left half: [Saa' | Sab' | sa'] rows of row tile tL
    const int r = 8 * tL + rr;
    double part = 0.0;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) part = fma(a[ks], zs[4 * ks + q], part);
    part += __shfl_xor_sync(0xffffffffu, part, 1);
    part += __shfl_xor_sync(0xffffffffu, part, 2);
    if (q == 0) EO[D::OSA + r] = EL[D::OSA + r] - part;
    big_schur_run<NS>(a, Fs, EL + D::OAA + r * NS, EO + D::OAA + r * NS, sym ? tL + 1 : NT, q, rr);   // f_L -> Saa'
    big_schur_run<NS>(a, Fs + NS * LDF, nullptr, EO + D::OAB + r * NS, NT, q, rr);                    // f_R -> Sab'
  }
  {  // right half: [Sbb' | sb'] rows of row tile tR
    const int r = 8 * tR + rr;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) a[ks] = __ldg(ER + D::OAB + (4 * ks + q) * NS + r);
    double part = 0.0;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) part = fma(a[ks], zs[4 * ks + q], part);
    part += __shfl_xor_sync(0xffffffffu, part, 1);
    part += __shfl_xor_sync(0xffffffffu, part, 2);
    if (q == 0) EO[D::OSB + r] = ER[D::OSB + r] - part;
    big_schur_run<NS>(a, Fs + NS * LDF, ER + D::OBB + r * NS, EO + D::OBB + r * NS, sym ? tR + 1 : NT, q, rr);  // f_R -> Sbb'
  }
}

// ---------------------------------------------------------------------------------------------------------
// Top-down, one launch per level (coarsest first): lambda_{i+1} = zs - f_L lambda_a - f_R lambda_{b+1} for the separators
// of that level ([a,b] = the merged subtree). One CTA of NS threads per (problem, separator).
// ---------------------------------------------------------------------------------------------------------
template <int NS>
__global__ void __launch_bounds__(NS) big_topdown_kernel(int N, const double* rec_all, double* lam_all, int level, int M) {
  using D = BigDim<NS>;
  __shared__ double la[NS], lb[NS];
  const int prob = blockIdx.x / M, t = blockIdx.x % M;
  const int i_sep = ((2 * t + 1) << level) - 1;
  const int a = i_sep + 1 - (1 << level), b1 = i_sep + 1 + (1 << level);
  double* lam = lam_all + (size_t)prob * (N + 1) * NS;
  const int r = threadIdx.x;
  la[r] = a == 0 ? 0.0 : lam[(size_t)a * NS + r];  // (lambda_0 does not couple: f_L is zero there)
  lb[r] = lam[(size_t)b1 * NS + r];                // slot N is zero
  __syncthreads();
  const double* rec = rec_all + ((size_t)prob * N + i_sep) * D::WSZ;
  double v = rec[2 * NS * NS + r];
#pragma unroll 8
  for (int c = 0; c < NS; ++c) v = fma(-rec[c * NS + r], la[c], fma(-rec[(NS + c) * NS + r], lb[c], v));
  lam[(size_t)(i_sep + 1) * NS + r] = v;
}

// ---------------------------------------------------------------------------------------------------------
// Down: x_k = Q_k^-1 (rhs_x + lambda_k - A_k^T lambda_{k+1}), u_k = R_k^-1 (rhs_u - B_k^T lambda_{k+1}) from the leaf
// equations (the reference's rhs propagation src/solve.c:175-180 telescoped); knot 0: x_0 = x0 and
// lambda_0 = Q_0 x0 + q_0 + A_0^T lambda_1 (src/nested_dissection.c:24-59). One CTA per (problem, knot), one thread per row.
// ---------------------------------------------------------------------------------------------------------
template <int NS, int NI>
__global__ void __launch_bounds__(256) big_down_kernel(Shape s, Bufs b, const double* lam_all, int pack_rhs) {
  constexpr int XU = NS + NI, CPW = XU / 8;
  static_assert(NS == 64 && XU % 8 == 0, "two k per lane; columns dealt to 8 warps");
  __shared__ double dot[XU];
  const int N = s.N;
  const int prob = blockIdx.x / N, k = blockIdx.x % N;
  const int r = threadIdx.x, warp = r >> 5, lane = r & 31;
  const double* lam = lam_all + (size_t)prob * (N + 1) * NS;
  const bool first = k == 0, last = k == N - 1;
  const double* kn = knot_ptr(s, b, prob, k);
  // [A | B]^T lambda_{k+1}: warp w takes columns w, w + 8, ...: a lane reads two consecutive-by-lane elements of a column
  // (coalesced), all 2 CPW loads of the warp in flight, then a shuffle reduction per column
  {
    const double l0 = lam[(size_t)(k + 1) * NS + lane], l1 = lam[(size_t)(k + 1) * NS + 32 + lane];
    double part[CPW];
#pragma unroll
    for (int u = 0; u < CPW; ++u) {
      const double* col = kn + s.oA + (size_t)(warp + 8 * u) * NS;
      part[u] = last ? 0.0 : fma(__ldg(col + lane), l0, __ldg(col + 32 + lane) * l1);
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
  double* out = soln_ptr(s, b, prob, k);
  const double acc = dot[r];  // row r of [A | B]^T lambda_{k+1}
  const double d = kn[r];
  double rx;
  if (!pack_rhs) rx = out[NS + r];
  else rx = r < NS ? -kn[s.oq + r] : (last ? 0.0 : -kn[s.orr + r - NS]);
  if (r < NS) {
    const double lk = first ? 0.0 : lam[(size_t)k * NS + r];
    double xv, lv;
    if (first) {
      const double rl = pack_rhs ? -b.x0[(size_t)prob * NS + r] : out[r];
      xv = -rl;
      lv = -d * rl - rx + acc;
    } else {
      xv = (rx + lk - acc) / d;
      lv = lk;
    }
    out[r] = lv;
    out[NS + r] = xv;
  } else {
    out[NS + r] = last ? rx : (rx - acc) / d;
  }
}

// ---------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------
struct BigPlan {
  bool covered;    // the shape has these kernels
  bool allocated;
  BigBufs bb;
  size_t bytes;
  int launches;
};

inline bool big_shape_ok(int n, int m, int N) { return n == 64 && m == 16 && N >= 2; }

inline void big_free(BigPlan* p) {
  cudaFree(p->bb.entA); cudaFree(p->bb.entB); cudaFree(p->bb.rec); cudaFree(p->bb.lam); cudaFree(p->bb.ufac);
  const bool cov = p->covered;
  memset(p, 0, sizeof(*p));
  p->covered = cov;
}

inline int big_alloc(const Shape& s, int batch, BigPlan* p, std::string& err) {
  if (p->allocated) return 0;
  const size_t ent = BigDim<64>::ENT, wsz = BigDim<64>::WSZ;
  auto alloc = [&](double** ptr, size_t cnt) {
    p->bytes += cnt * sizeof(double);
    return cudaMalloc((void**)ptr, cnt * sizeof(double)) == cudaSuccess;
  };
  const bool ok = alloc(&p->bb.entA, (size_t)batch * s.N * ent) && alloc(&p->bb.entB, (size_t)batch * (s.N / 2) * ent) &&
                  alloc(&p->bb.rec, (size_t)batch * s.N * wsz) && alloc(&p->bb.lam, (size_t)batch * (s.N + 1) * s.n) &&
                  alloc(&p->bb.ufac, (size_t)batch * (s.N / 2) * BigFac<64>::SIZE);
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
inline int big_launch(BigPlan& p, const Shape& s, const Bufs& b, int batch, int pack_rhs, cudaStream_t st, std::string& err) {
  constexpr int NS = 64, NI = 16;
  using D = BigDim<NS>;
  const int N = s.N;
  p.launches = 0;
  cudaError_t e = cudaMemsetAsync(p.bb.lam, 0, sizeof(double) * (size_t)batch * (N + 1) * NS, st);
  if (e != cudaSuccess) { err = std::string("big launch: ") + cudaGetErrorString(e); return -3; }
  const int leaf_smem = ((NS + 8) * (NS + NI) + 2 * (NS + NI) + 2 * NS + 3 * NS) * (int)sizeof(double);
  const int spine_smem = NS * (NS + 2) * (int)sizeof(double);
  const int schur_smem = (2 * NS + 1) * (NS + 4) * (int)sizeof(double);
  const int factor_smem = BIG_FACTOR_WPC * big_factor_warp_doubles<NS>() * (int)sizeof(double);
  const int solve_smem = (NS * (NS + 4) + NS * (2 * NS + 8) + (NS / 8) * 64 + NS) * (int)sizeof(double);
  static const int one_kernel_spine = [] { const char* v = getenv("RSLQR_BIG_SPINE"); return v ? atoi(v) : 0; }();
  static const int packed_factor = [] { const char* v = getenv("RSLQR_BIG_FACTOR_PACKED"); return v ? atoi(v) : 0; }();  // A/B
  if ((e = smem_optin(big_factor_kernel<NS>, factor_smem)) != cudaSuccess || (e = smem_optin(big_solve_kernel<NS>, solve_smem)) != cudaSuccess) {
    err = std::string("big launch: ") + cudaGetErrorString(e);
    return -3;
  }
  if ((e = smem_optin(big_leaf_kernel<NS, NI>, leaf_smem)) != cudaSuccess || (e = smem_optin(big_spine_kernel<NS>, spine_smem)) != cudaSuccess ||
      (e = smem_optin(big_schur_kernel<NS>, schur_smem)) != cudaSuccess) {
    err = std::string("big launch: ") + cudaGetErrorString(e);
    return -3;
  }
  big_leaf_kernel<NS, NI><<<(unsigned)((size_t)batch * N), 256, leaf_smem, st>>>(s, b, p.bb, pack_rhs, one_kernel_spine ? 0 : 1);
  ++p.launches;
  const double* ein = p.bb.entA;
  for (int level = 0; level < s.K; ++level) {
    const int M = N >> (level + 1);
    double* eout = (level & 1) ? p.bb.entA : p.bb.entB;
    if (one_kernel_spine) {
      big_spine_kernel<NS><<<(unsigned)((size_t)batch * M), big_spine_threads<NS>(), spine_smem, st>>>(s, b, ein, p.bb.rec, level, M);
      ++p.launches;
    } else {
      const long long nmerge = (long long)batch * M;
      if (packed_factor)
        big_factor_kernel<NS><<<(unsigned)((nmerge + BIG_FACTOR_WPC - 1) / BIG_FACTOR_WPC), 32 * BIG_FACTOR_WPC, factor_smem, st>>>(
            s, b, ein, p.bb.ufac, level, M, nmerge);
      else
        big_factor_reg_kernel<NS><<<(unsigned)((nmerge + BIG_FACTOR_REG_WPC - 1) / BIG_FACTOR_REG_WPC), 32 * BIG_FACTOR_REG_WPC, 0, st>>>(
            s, b, ein, p.bb.ufac, level, M, nmerge);
      big_solve_kernel<NS><<<(unsigned)nmerge, BIG_SOLVE_THREADS, solve_smem, st>>>(s, ein, p.bb.ufac, p.bb.rec, level, M);
      p.launches += 2;
    }
    if (level < s.K - 1) {
      big_schur_kernel<NS><<<(unsigned)((size_t)batch * M), 256, schur_smem, st>>>(s, ein, eout, p.bb.rec, level, M,
                                                                                     one_kernel_spine ? 0 : 1);
      ++p.launches;
    }
    ein = eout;
  }
  for (int level = s.K - 1; level >= 0; --level) {
    const int M = N >> (level + 1);
    big_topdown_kernel<NS><<<(unsigned)((size_t)batch * M), NS, 0, st>>>(N, p.bb.rec, p.bb.lam, level, M);
    ++p.launches;
  }
  big_down_kernel<NS, NI><<<(unsigned)((size_t)batch * N), 256, 0, st>>>(s, b, p.bb.lam, pack_rhs);
  ++p.launches;
  e = cudaGetLastError();
  if (e != cudaSuccess) { err = std::string("big launch: ") + cudaGetErrorString(e); return -3; }
  return 0;
}

}  // namespace rslqr

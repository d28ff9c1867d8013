// kernels_staged_mma.cuh -- staged-path kernels for LARGE state dimensions (n a multiple of 16, e.g. the n=64, m=16
// configuration of BASELINE configs[4]), where one factor block no longer fits the on-chip tiers and the per-level
// work is a dense block contraction:
//   schur_mma_kernel  : ndlqr_UpdateShurFactor (reference src/nested_dissection.c:154-171) as a tiled
//                       [(2n+m) x n] -= [(2n+m) x n] * [n x n] product on the fp64 tensor cores (mma.sync m8n8k4);
//                       operands stream from HBM exactly once per task, f is staged in shared memory.
//   solves_cols_kernel: ndlqr_SolveCholeskyFactor (reference src/nested_dissection.c:136-152 ->
//                       src/linalg_custom.c:113-138): one thread per right-hand-side column, the column lives in
//                       registers, L is broadcast from shared memory; no barrier inside the substitution.
// Same task decoding (Sel) and the same `fact` layout as kernels_staged.cuh, so they are drop-in replacements for
// schur_kernel / solves_kernel and the stage-by-stage parity tests apply unchanged.
#pragma once
#include "kernels_fused.cuh"  // dmma884
#include "kernels_staged.cuh"

namespace rslqr {

template <int NS, int NI>
__global__ void __launch_bounds__(32 * ((2 * NS + NI) / 16), 2) schur_mma_kernel(Shape s, Bufs b, Sel sel, int level) {
  static_assert(NS % 16 == 0 && NI % 16 == 0, "row chunks of 16 must not straddle the Lambda / X / U blocks");
  constexpr int LDF = NS + 4;  // padded leading dimension of f in shared memory: conflict-free B-fragment loads
  constexpr int KS = NS / 4, NT = NS / 8, nn = NS * NS;
  extern __shared__ __align__(16) double sf_mma[];
  int task = blockIdx.x, slot = 0;
  if (sel.col < 0) { slot = 1 + task % 2; task /= 2; }
  const int prob = sel.p0 + task / sel.nitems;
  const int k = sel.item0 + task % sel.nitems;
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
  for (int e = threadIdx.x; e < nn; e += blockDim.x) sf_mma[(e / NS) * LDF + (e % NS)] = f[e];
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q = lane & 3, rr = lane >> 2;
  const int r0 = warp * 16;  // 16 rows of the [Lambda | X | U] block per warp
  if (r0 < NS && !calc) return;
  // block-local addressing: Lambda and X have leading dimension n, U has m
  int base, ld, rloc;
  if (r0 < NS) { base = 0; ld = NS; rloc = r0; }
  else if (r0 < 2 * NS) { base = nn; ld = NS; rloc = r0 - NS; }
  else { base = 2 * nn; ld = NI; rloc = r0 - 2 * NS; }
  const double* A0 = Fk + base + rloc + rr;
  double* C0 = g + base + rloc + rr;
  double c0[NT][2], c1[NT][2];
#pragma unroll
  for (int nt = 0; nt < NT; ++nt) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      c0[nt][h] = C0[ld * (8 * nt + 2 * q + h)];
      c1[nt][h] = C0[ld * (8 * nt + 2 * q + h) + 8];
    }
  }
#pragma unroll 4
  for (int ks = 0; ks < KS; ++ks) {
    const double a0 = A0[ld * (4 * ks + q)];
    const double a1 = A0[ld * (4 * ks + q) + 8];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const double bf = -sf_mma[(8 * nt + rr) * LDF + 4 * ks + q];
      dmma884(c0[nt][0], c0[nt][1], a0, bf);
      dmma884(c1[nt][0], c1[nt][1], a1, bf);
    }
  }
#pragma unroll
  for (int nt = 0; nt < NT; ++nt) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      C0[ld * (8 * nt + 2 * q + h)] = c0[nt][h];
      C0[ld * (8 * nt + 2 * q + h) + 8] = c1[nt][h];
    }
  }
}

// G (n x n, columns = right-hand sides) <- (L L^T)^{-1} G. One CTA per (problem, separator, boundary column), one
// thread per column of G. The column sits in shared memory (conflict-free, ld n+1) and is worked on in register
// blocks of 8: every 8x8 block of L (broadcast 16-byte loads) costs 64 FMAs against 16 loads/stores of the column.
// No barrier inside the substitution. dynamic smem: (NS*(NS+2) + NS*(NS+1) + NS) doubles.
template <int NS>
__global__ void __launch_bounds__(NS) solves_cols_kernel(Shape s, Bufs b, Sel sel, int level) {
  constexpr int LDL = NS + 2, LDG = NS + 1, nn = NS * NS, BS = 8, NB = NS / BS;
  static_assert(NS % BS == 0, "block size");
  extern __shared__ __align__(16) double ssm[];
  double* sL = ssm;               // L row-major: sL[i*LDL + j] = L(i,j); rows 16-byte aligned
  double* sG = ssm + NS * LDL;    // sG[r*LDG + c] = G(r,c)
  double* dinv = sG + NS * LDG;   // 1 / L(j,j)
  int task = blockIdx.x, slot = 0;
  if (sel.col < 0) { slot = 1 + task % 2; task /= 2; }
  const int prob = sel.p0 + task / sel.nitems;
  const int leaf = sel.item0 + task % sel.nitems;
  const int p = sel.col < 0 ? active_col(s, leaf, level, slot) : sel.col;
  if (p < 0) return;
  const int i1 = rs_sep_index(leaf, level) + 1;
  const double* L = fact_ptr(s, b, prob, i1, level);
  double* G = fact_ptr(s, b, prob, i1, p);
  for (int e = threadIdx.x; e < nn; e += NS) {
    const int r = e % NS, c = e / NS;  // column-major in global memory
    sL[r * LDL + c] = L[e];
    sG[r * LDG + c] = G[e];
  }
  __syncthreads();
  dinv[threadIdx.x] = 1.0 / sL[threadIdx.x * LDL + threadIdx.x];
  __syncthreads();
  double* col = sG + threadIdx.x;
  // forward: L y = b (reference clap_LowerTriBackSub, src/linalg_custom.c:113-132)
#pragma unroll 1
  for (int jb = 0; jb < NB; ++jb) {
    double xj[BS];
#pragma unroll
    for (int t = 0; t < BS; ++t) xj[t] = col[(jb * BS + t) * LDG];
#pragma unroll
    for (int t = 0; t < BS; ++t) {
      xj[t] *= dinv[jb * BS + t];
#pragma unroll
      for (int u = t + 1; u < BS; ++u) xj[u] = fma(-sL[(jb * BS + u) * LDL + jb * BS + t], xj[t], xj[u]);
    }
#pragma unroll
    for (int t = 0; t < BS; ++t) col[(jb * BS + t) * LDG] = xj[t];
#pragma unroll 1
    for (int ib = jb + 1; ib < NB; ++ib) {
      double xi[BS];
#pragma unroll
      for (int u = 0; u < BS; ++u) xi[u] = col[(ib * BS + u) * LDG];
#pragma unroll
      for (int u = 0; u < BS; ++u) {
        const double2* lrow = reinterpret_cast<const double2*>(sL + (ib * BS + u) * LDL + jb * BS);
#pragma unroll
        for (int t = 0; t < BS; t += 2) {
          const double2 l2 = lrow[t / 2];
          xi[u] = fma(-l2.x, xj[t], xi[u]);
          xi[u] = fma(-l2.y, xj[t + 1], xi[u]);
        }
      }
#pragma unroll
      for (int u = 0; u < BS; ++u) col[(ib * BS + u) * LDG] = xi[u];
    }
  }
  // backward: L^T x = y
#pragma unroll 1
  for (int jb = NB - 1; jb >= 0; --jb) {
    double xj[BS];
#pragma unroll
    for (int t = 0; t < BS; ++t) xj[t] = col[(jb * BS + t) * LDG];
#pragma unroll
    for (int t = BS - 1; t >= 0; --t) {
      xj[t] *= dinv[jb * BS + t];
#pragma unroll
      for (int u = 0; u < t; ++u) xj[u] = fma(-sL[(jb * BS + t) * LDL + jb * BS + u], xj[t], xj[u]);
    }
#pragma unroll
    for (int t = 0; t < BS; ++t) col[(jb * BS + t) * LDG] = xj[t];
#pragma unroll 1
    for (int ib = 0; ib < jb; ++ib) {
      double xi[BS];
#pragma unroll
      for (int u = 0; u < BS; ++u) xi[u] = col[(ib * BS + u) * LDG];
#pragma unroll
      for (int t = 0; t < BS; ++t) {
        const double2* lrow = reinterpret_cast<const double2*>(sL + (jb * BS + t) * LDL + ib * BS);
#pragma unroll
        for (int u = 0; u < BS; u += 2) {
          const double2 l2 = lrow[u / 2];
          xi[u] = fma(-l2.x, xj[t], xi[u]);
          xi[u + 1] = fma(-l2.y, xj[t], xi[u + 1]);
        }
      }
#pragma unroll
      for (int u = 0; u < BS; ++u) col[(ib * BS + u) * LDG] = xi[u];
    }
  }
  __syncthreads();
  for (int e = threadIdx.x; e < nn; e += NS) G[e] = sG[(e % NS) * LDG + e / NS];
}

// ndlqr_FactorInnerProduct (reference src/nested_dissection.c:114-134) for large n on the fp64 tensor cores:
//   S = F(i+1,p).Lambda <- [A_i | B_i] [F(i,p).X ; F(i,p).U] - F(i+1,p).X - S_old
// One CTA per (problem, separator, column slot); NS/16 warps, each 16 rows x NS columns of S. The stacked operand
// [X;U] ((NS+NI) x NS) is staged in shared memory, column per row of smem (ld NS+NI+4: conflict-free B fragments).
template <int NS, int NI>
__global__ void __launch_bounds__(32 * (NS / 16)) products_mma_kernel(Shape s, Bufs b, Sel sel, int level) {
  static_assert(NS % 16 == 0 && NI % 4 == 0, "tile shapes");
  constexpr int KK = NS + NI, LDB = KK + 4, KS = KK / 4, NT = NS / 8, nn = NS * NS;
  extern __shared__ __align__(16) double sb_mma[];
  int task = blockIdx.x, slot = 0, p;
  if (sel.col < 0) { slot = task % 3; task /= 3; }
  const int prob = sel.p0 + task / sel.nitems;
  const int leaf = sel.item0 + task % sel.nitems;
  p = sel.col < 0 ? active_col(s, leaf, level, slot) : sel.col;
  if (p < 0) return;
  const int i = rs_sep_index(leaf, level);
  const double* AB = knot_ptr(s, b, prob, i) + s.oA;  // [A | B], NS rows, column-major
  const double* F1 = fact_ptr(s, b, prob, i, p);
  double* F2 = fact_ptr(s, b, prob, i + 1, p);
  for (int e = threadIdx.x; e < nn; e += blockDim.x) sb_mma[(e / NS) * LDB + (e % NS)] = F1[nn + e];            // X(l, c)
  for (int e = threadIdx.x; e < NI * NS; e += blockDim.x) sb_mma[(e / NI) * LDB + NS + (e % NI)] = F1[2 * nn + e];  // U(l, c)
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q = lane & 3, rr = lane >> 2;
  const int r0 = warp * 16 + rr;
  double c0[NT][2], c1[NT][2];
#pragma unroll
  for (int nt = 0; nt < NT; ++nt)
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int e = r0 + NS * (8 * nt + 2 * q + h);
      c0[nt][h] = -F2[nn + e] - F2[e];
      c1[nt][h] = -F2[nn + e + 8] - F2[e + 8];
    }
#pragma unroll 4
  for (int ks = 0; ks < KS; ++ks) {
    const double a0 = AB[r0 + NS * (4 * ks + q)];
    const double a1 = AB[r0 + 8 + NS * (4 * ks + q)];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const double bf = sb_mma[(8 * nt + rr) * LDB + 4 * ks + q];
      dmma884(c0[nt][0], c0[nt][1], a0, bf);
      dmma884(c1[nt][0], c1[nt][1], a1, bf);
    }
  }
#pragma unroll
  for (int nt = 0; nt < NT; ++nt)
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int e = r0 + NS * (8 * nt + 2 * q + h);
      F2[e] = c0[nt][h];
      F2[e + 8] = c1[nt][h];
    }
}

}  // namespace rslqr

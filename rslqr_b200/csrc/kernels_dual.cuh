// kernels_dual.cuh -- the scan path in DUAL form, for any (n, m) with n <= 16, m <= 16.
//
// Eliminating x_k, u_k at the leaves (ndlqr_SolveLeaf, reference src/nested_dissection.c:10-105; Q_k, R_k are
// diagonal) leaves a block-tridiagonal SPD system in the multipliers alone:
//     M lambda = rho,   M(k+1,k+1) = A_k Q_k^-1 A_k^T + B_k R_k^-1 B_k^T + Q_{k+1}^-1,   M(k+1,k+2) = -Q_{k+1}^-1 A_{k+1}^T
// whose diagonal blocks are exactly the separator matrices S the reference factors (ndlqr_FactorInnerProduct at the
// leaf level, src/nested_dissection.c:114-134), and the reference's nested dissection is block cyclic reduction on it.
// A subtree [a,b] is carried as its Schur complement onto the two multipliers bounding it (lambda_a, lambda_{b+1}):
//     element matrix [[Saa, Sab], [Sab^T, Sbb]] (n x n blocks),  element rhs [sa; sb]
// instead of as the (2n+m)-row boundary entry of kernels_scan.cuh. Merging [a,i] and [i+1,b] over lambda_{i+1}:
//     S = SbbL + SaaR,  S_L = SabL^T,  S_R = SabR,  s = sbL + saR       (the reference's S, S_L, S_R, s)
//     [f_L f_R zs] = S^-1 [S_L S_R s]                                    (Cholesky + ndlqr_SolveCholeskyFactor)
//     Saa' = SaaL - S_L^T f_L,  Sab' = -S_L^T f_R,  Sbb' = SbbR - S_R^T f_R,  sa' = saL - S_L^T zs,  sb' = sbR - S_R^T zs
// Same separators, same S, same records [f_L | f_R | zs] as the other paths (and as the reference up to rounding);
// what differs is that the product with [A_b | B_b] that every later inner product applies to the boundary rows
// X(b), U(b) is applied once, at the leaves, so a merge above the leaves is n x n blocks only: no input rows, no
// P1 product, and m only appears in the leaf kernel and in the down pass. tools/dual_prototype.py is the numpy
// statement of the same recurrences, checked against the oracle.
//
// Entry of a subtree, ORIENTED by the side it will be merged from (left child: L-type, right child: R-type):
//     [ Kp | Cp | Sp | ks | ss ]     Kp, Cp, Sp: NP x LD row-major (NP = n rounded up to a multiple of 4, LD = NP + 2)
//     L-type: Kp = Saa, Cp = Sab,   Sp = Sbb, ks = sa, ss = sb      (the parent's S_L columns are the rows of Cp)
//     R-type: Kp = Sbb, Cp = Sab^T, Sp = Saa, ks = sb, ss = sa      (the parent's S_R columns are the rows of Cp)
// so that a merge reads S = Sp(L) + Sp(R), the right-hand sides as ROWS of Cp(L), Cp(R) and s = ss(L) + ss(R), and
// writes  [Cp(L); Cp(R)] [f_L f_R zs]  subtracted from the keep parts. Padded rows/columns (n not a multiple of 4)
// are zero except for 0.5 on the diagonal of Sp, which makes the padded part of S the identity.
#pragma once
#include "kernels_scan.cuh"

namespace rslqr {

// ---------------------------------------------------------------------------------------------------------
// One merge, one warp. EL: L-type entry of the left subtree, ER: R-type entry of the right one (shared memory).
// Eout: the merged entry, R-type if out_right else L-type; may be EL itself (rows are consumed tile by tile before
// they are overwritten) or a separate buffer, or nullptr (root: only the record is wanted).
// W: [f_L | f_R | zs] column-major ld NP on return (WP doubles); Lr: factor scratch (LR doubles).
// Returns non-zero (warp-uniform) when a pivot is <= 0 (the reference's Cholesky fails, src/linalg_custom.c:88).
// ---------------------------------------------------------------------------------------------------------
template <int NS>
__device__ __forceinline__ int dual_merge(const double* EL, const double* ER, double* Eout, bool out_right, double* W,
                                          double* Lr) {
  using D = DualDim<NS>;
  constexpr int NP = D::NP, LD = D::LD;
  static_assert(2 * NP <= 32, "one warp per separator");
  const int lane = threadIdx.x & 31, q = lane & 3, rr = lane >> 2;
  int bad;
  {
    // lanes [0,NP): row c of S = Sp(L) + Sp(R) (symmetric: row = column); lanes [NP,2NP): row c of Cp(L) (= column c of
    // S_L) and of Cp(R) (= column c of S_R); s rides in acc2 of lane 0
    const bool isS = lane < NP;
    const int c = isS ? lane : (lane < 2 * NP ? lane - NP : 0);
    const double* pa = EL + (isS ? D::OS : D::OC) + c * LD;
    const double* pb = ER + (isS ? D::OS : D::OC) + c * LD;
    double acc[NP], acc2[NP];
#pragma unroll
    for (int r = 0; r < NP; r += 2) {
      const double2 a = *reinterpret_cast<const double2*>(pa + r);
      const double2 b = *reinterpret_cast<const double2*>(pb + r);
      const double2 sl = *reinterpret_cast<const double2*>(EL + D::OSS + r);
      const double2 sr = *reinterpret_cast<const double2*>(ER + D::OSS + r);
      acc[r] = isS ? a.x + b.x : a.x;
      acc[r + 1] = isS ? a.y + b.y : a.y;
      acc2[r] = isS ? (lane == 0 ? sl.x + sr.x : 0.0) : b.x;
      acc2[r + 1] = isS ? (lane == 0 ? sl.y + sr.y : 0.0) : b.y;
    }
    __syncwarp();  // (Lr / W may live where other lanes were still reading)
    bad = spine_core2<NP, true>(acc, acc2, Lr, W);
  }
  if (!Eout) return bad;
  // ---- [Cp(L); Cp(R)] (2NP x NP) times [f_L f_R zs] (NP x 2NP+1) on the fp64 tensor cores, subtracted from the keep
  //      parts: rows < NP come from the left entry, rows >= NP from the right one ----
  constexpr int MT = 2 * NP / 8, NT = 2 * NP / 8, KS = NP / 4;
  double bneg[KS][NT], zq[KS];
#pragma unroll
  for (int ks = 0; ks < KS; ++ks) {
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) bneg[ks][nt] = -W[(8 * nt + rr) * NP + 4 * ks + q];
    zq[ks] = W[2 * NP * NP + 4 * ks + q];
  }
#pragma unroll
  for (int mt = 0; mt < MT; ++mt) {
    const int rho = 8 * mt + rr;
    const bool isL = rho < NP;
    const int r = isL ? rho : rho - NP;
    const double* X = isL ? EL : ER;
    double a[KS];
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) a[ks] = X[D::OC + r * LD + 4 * ks + q];
    double c[NT][2];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const int g = 8 * nt + 2 * q;                    // columns g, g+1: both in the same half (NP is even)
      const bool colL = g < NP;
      if (colL == isL) {                               // aa' (left rows, left columns) or bb': keep part
        const double2 v = *reinterpret_cast<const double2*>(X + D::OK + r * LD + (colL ? g : g - NP));
        c[nt][0] = v.x; c[nt][1] = v.y;
      } else { c[nt][0] = 0.0; c[nt][1] = 0.0; }
    }
    const double z0 = X[D::OKS + r];
    __syncwarp();  // every lane has read this tile's rows: they may be overwritten now (Eout == EL)
#pragma unroll
    for (int ks = 0; ks < KS; ++ks)
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) dmma884(c[nt][0], c[nt][1], a[ks], bneg[ks][nt]);
    double part = 0.0;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) part = fma(a[ks], zq[ks], part);
    part += __shfl_xor_sync(0xffffffffu, part, 1);
    part += __shfl_xor_sync(0xffffffffu, part, 2);
    // destinations: the side that matches the output's orientation keeps [Kp | Cp | ks], the other one gives Sp, ss
    const bool own = isL != out_right;                 // L rows of an L-type output, R rows of an R-type output
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const int g = 8 * nt + 2 * q;
      const bool colL = g < NP;
      const bool diag = colL == isL;                   // aa' / bb' (else ab' / ba')
      const int gc = colL ? g : g - NP;
      if (own) {
        *reinterpret_cast<double2*>(Eout + (diag ? D::OK : D::OC) + r * LD + gc) = make_double2(c[nt][0], c[nt][1]);
      } else if (diag) {
        double v0 = c[nt][0], v1 = c[nt][1];
        if constexpr (NP > NS) {  // padded rows of Sp keep 0.5 on the diagonal (the next S gets the identity there)
          if (r >= NS) { v0 = gc == r ? 0.5 : 0.0; v1 = gc + 1 == r ? 0.5 : 0.0; }
        }
        *reinterpret_cast<double2*>(Eout + D::OS + r * LD + gc) = make_double2(v0, v1);
      }
    }
    if (q == 0) Eout[(own ? D::OKS : D::OSS) + r] = z0 - part;
    __syncwarp();
  }
  return bad;
}

// record [f_L | f_R | zs]: shared memory (ld NP) -> global (ld NS, the layout topdown / down read)
template <int NS>
__device__ __forceinline__ void dual_store_record(double* rec, const double* W) {
  using D = DualDim<NS>;
  constexpr int NP = D::NP;
  const int lane = threadIdx.x & 31;
  for (int e = lane; e < D::WSZ; e += 32) {
    const int col = e / NS, row = e - col * NS;
    const int wc = col < NS ? col : (col < 2 * NS ? NP + col - NS : 2 * NP);
    rec[e] = W[wc * NP + row];
  }
}

// Sbb = [A | B] D [A | B]^T of one knot on the fp64 tensor cores, whole warp: AB = [A | B] (column-major ld NS) and
// dv = the leaf diagonals [1/Q | 1/R] of the knot's transformed block; noQ: knot 0 (1/Q -> 0); dead: last knot (no b
// side: zero). Writes rows [0, NS) of the NP x LD row-major block `dst`, zero in the padded columns [NS, NP).
template <int NS, int NI>
__device__ __forceinline__ void dual_leaf_sbb(const double* AB, const double* dv, bool noQ, bool dead, double* dst) {
  using D = DualDim<NS>;
  constexpr int NP = D::NP, LD = D::LD, XU = NS + NI, MT = (NS + 7) / 8, KS = (XU + 3) / 4;
  const int lane = threadIdx.x & 31, q = lane & 3, rr = lane >> 2;
  double a[MT][KS];  // A fragments [A | B](8 mt + rr, 4 ks + q) D(4 ks + q); unscaled they are the B fragments of the same tile index
  double bq[MT][KS];
#pragma unroll
  for (int ks = 0; ks < KS; ++ks) {
    const int l = 4 * ks + q;
    const bool lv = l < XU;
    const double d = (!lv || dead || (l < NS && noQ)) ? 0.0 : dv[lv ? l : 0];
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) {
      const int r = 8 * mt + rr;
      const double v = (lv && r < NS) ? AB[r + NS * (lv ? l : 0)] : 0.0;
      bq[mt][ks] = v;
      a[mt][ks] = v * d;
    }
  }
#pragma unroll
  for (int mt = 0; mt < MT; ++mt)
#pragma unroll
    for (int nt = 0; nt < MT; ++nt) {
      double c0 = 0.0, c1 = 0.0;
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) dmma884(c0, c1, a[mt][ks], bq[nt][ks]);
      const int row = 8 * mt + rr, col = 8 * nt + 2 * q;
      if (row < NS && col < NP) *reinterpret_cast<double2*>(dst + row * LD + col) = make_double2(col < NS ? c0 : 0.0, col + 1 < NS ? c1 : 0.0);
    }
}

// ---------------------------------------------------------------------------------------------------------
// Leaf elements of two neighbouring knots at once, straight from their transformed blocks (leaf_transform):
// lanes [0,NS) build the L-type element of knot kL into EL, lanes [NS,2NS) the R-type element of knot kR into ER.
//   element of knot k:  Saa = Q^-1,  Sab = -Q^-1 A^T,  Sbb = A Q^-1 A^T + B R^-1 B^T,
//                       sa = -z_x - z_lambda,  sb = A z_x + B z_u          (z: the leaf-solved rhs)
//   knot 0 (x_0 is fixed, lambda_0 is no unknown of the dual system): Q^-1 -> 0 in Saa, Sab, Sbb
//   last knot of the horizon: only Saa, sa (its R, A, B, d are ignored)
// Sbb on DMMA (dual_leaf_sbb); the rest one row per lane, all stores contiguous rows.
// ---------------------------------------------------------------------------------------------------------
template <int NS, int NI>
__device__ __forceinline__ void dual_leaf_pair(const double* BL, const double* zlL, const double* BR, const double* zlR,
                                               bool first0, bool lastR, double* EL, double* ER) {
  using D = DualDim<NS>;
  using RD = RawDim<NS, NI>;
  constexpr int NP = D::NP, LD = D::LD, XU = NS + NI;
  const int lane = threadIdx.x & 31;
  const bool right = lane >= NS;
  const int c = right ? lane - NS : lane;
  const bool act = lane < 2 * NS;
  const double* B = right ? BR : BL;
  const double* zl = right ? zlR : zlL;
  double* E = right ? ER : EL;
  const double* AB = B + RD::OA;
  const bool noQ = !right && first0;     // knot 0
  const bool dead = right && lastR;      // last knot: no b side
  const int cs = act ? c : 0;
  // Sbb of both knots on DMMA, straight into its place: Sp of the L-type element, Kp of the R-type one
  dual_leaf_sbb<NS, NI>(BL + RD::OA, BL, first0, false, EL + D::OS);
  dual_leaf_sbb<NS, NI>(BR + RD::OA, BR, false, lastR, ER + D::OK);
  double sb = 0.0;     // row c of [A | B] z
#pragma unroll
  for (int l = 0; l < XU; ++l) sb = fma(dead ? 0.0 : AB[cs + NS * l], B[RD::OZ + l], sb);
  const double dq = noQ ? 0.0 : B[cs];                 // 1 / Q(c)
  const double sa = -B[RD::OZ + cs] - zl[cs];
  if (act) {
    double* kp = E + D::OK + c * LD;
    double* cp = E + D::OC + c * LD;
    double* sp = E + D::OS + c * LD;
    // L-type: Kp = Saa (diagonal), Cp = Sab row c = -A(:, c)^T / Q(c), Sp = Sbb;  R-type: Kp = Sbb,
    // Cp = Sab^T row c = -A(c, :) / Q(:), Sp = Saa
#pragma unroll
    for (int r = 0; r < NP; r += 2) {
      double cx = 0.0, cy = 0.0;
      if (r < NS) cx = right ? (dead ? 0.0 : -AB[c + NS * r] * B[r]) : -AB[r + NS * c] * dq;
      if (r + 1 < NS) cy = right ? (dead ? 0.0 : -AB[c + NS * (r + 1)] * B[r + 1]) : -AB[r + 1 + NS * c] * dq;
      const double dx = r == c ? dq : 0.0, dy = r + 1 == c ? dq : 0.0;
      *reinterpret_cast<double2*>(cp + r) = make_double2(cx, cy);
      *reinterpret_cast<double2*>((right ? sp : kp) + r) = make_double2(dx, dy);   // Saa (the other block holds Sbb already)
    }
    E[D::OKS + c] = right ? sb : sa;
    E[D::OSS + c] = right ? sa : sb;
  }
  // padding rows NS..NP-1 of both elements: zero, 0.5 on the diagonal of Sp (S of the merge gets the identity there)
  if constexpr (NP > NS) {
    for (int it = lane; it < 2 * (NP - NS); it += 32) {
      double* Ep = it < NP - NS ? EL : ER;
      const int r = NS + (it < NP - NS ? it : it - (NP - NS));
      for (int k = 0; k < NP; ++k) {
        Ep[D::OK + r * LD + k] = 0.0;
        Ep[D::OC + r * LD + k] = 0.0;
        Ep[D::OS + r * LD + k] = k == r ? 0.5 : 0.0;
      }
      Ep[D::OKS + r] = 0.0;
      Ep[D::OSS + r] = 0.0;
    }
  }
  __syncwarp();
}

// ---------------------------------------------------------------------------------------------------------
// Leaf kernel: one warp = 4 consecutive knots (a subtree of the two lowest levels). One TMA bulk copy brings the flat
// knot blocks in, leaf_transform rewrites them in place (1/Q, 1/R, leaf-solved rhs), then
//   knots (0,1) -> merge -> X (L-type),  knots (2,3) -> merge -> Y (R-type),  X + Y -> the subtree's entry
// Output per subtree: three separator records (slots 0, 1, 2 = separators after local knots 0, 1, 2) and the entry,
// L-type for an even subtree index, R-type for an odd one.
// per warp: [NSE | 4 knot blocks | X | Y | Z | W | Lr]
// ---------------------------------------------------------------------------------------------------------

template <int NS, int NI>
__host__ __device__ constexpr int dual_leaf_warp_doubles() {
  using D = DualDim<NS>;
  return ((NS + 1) & ~1) + ((SCAN_LEAF * RawDim<NS, NI>::STRIDE + 1) & ~1) + 3 * D::ENT + ((D::WP + 1) & ~1) + D::LR;
}

// warps per CTA of the leaf kernel: the count (3..8) that puts the most warps of this shape on an SM (227 KB of shared
// memory, 1 KB reserved per CTA); the smaller count on a tie
template <int NS, int NI>
__host__ __device__ constexpr int dual_leaf_wpc() {
  const long long per_warp = (long long)dual_leaf_warp_doubles<NS, NI>() * 8;
  int best = 3, best_warps = 0;
  for (int w = 3; w <= 8; ++w) {
    const long long cta = w * per_warp + 1024;
    int ctas = (int)((227LL * 1024) / cta);
    if (ctas > 32) ctas = 32;
    const int warps = ctas * w > 64 ? 64 : ctas * w;
    if (warps > best_warps) { best = w; best_warps = warps; }
  }
  return best;
}

template <int NS, int NI>
__global__ void __launch_bounds__(32 * dual_leaf_wpc<NS, NI>()) dual_leaf_kernel(Shape s, Bufs b, double* ent0, double* sep_in, int pack_rhs,
                                                                   int M0, long long nsub) {
  pdl_sync();
  using D = DualDim<NS>;
  using RD = RawDim<NS, NI>;
  constexpr int NSE = (NS + 1) & ~1, RAW = (SCAN_LEAF * RD::STRIDE + 1) & ~1;
  constexpr int RAWB = SCAN_LEAF * RD::STRIDE * (int)sizeof(double);
  static_assert(RAWB % 16 == 0 && D::ENT % 2 == 0 && D::LR % 2 == 0, "TMA / 128-bit alignment");
  extern __shared__ __align__(16) double smf[];
  constexpr int DUAL_WPC = dual_leaf_wpc<NS, NI>();
  __shared__ __align__(8) unsigned long long mbar[DUAL_WPC];
  const int warp = warp_index_sync(), lane = threadIdx.x & 31;
  const long long sub = (long long)blockIdx.x * DUAL_WPC + warp;
  if (sub >= nsub) return;
  double* R0 = smf + warp * dual_leaf_warp_doubles<NS, NI>();
  double* R = R0 + (NSE - NS);            // leaf_transform's region: R[0..NS) = z_lambda of block 0, blocks at R + NS
  double* X = R0 + NSE + RAW;
  double* Y = X + D::ENT;
  double* Z = Y + D::ENT;
  double* W = Z + D::ENT;
  double* Lr = W + ((D::WP + 1) & ~1);
  const int prob = (int)(sub / M0), grp = (int)(sub - (long long)prob * M0);
  const int k0 = grp * SCAN_LEAF;
  const int N = s.N;
  if (lane == 0) {
    mbar_init(&mbar[warp], 1);
    fence_mbar_init();
    mbar_expect_tx(&mbar[warp], RAWB);
    bulk_g2s(R + NS, knot_ptr(s, b, prob, k0), RAWB, &mbar[warp]);
  }
  __syncwarp();
  const double zl0 = leaf_zl0<NS, NI>(s, b, prob, k0, pack_rhs);
  mbar_wait(&mbar[warp], 0);
  leaf_transform<NS, NI, SCAN_LEAF>(s, b, prob, k0, pack_rhs, R, zl0);
  double* rec = sep_in + ((size_t)prob * M0 + grp) * 3 * D::WSZ;
  auto blk = [&](int j) { return R + NS + j * RD::STRIDE; };
  auto zlam = [&](int j) { return R + j * RD::STRIDE; };
  int bad;
  // knots (0,1): separator after local knot 0
  dual_leaf_pair<NS, NI>(blk(0), zlam(0), blk(1), zlam(1), s.kofs + k0 == 0, s.kofs + k0 + 1 == s.Ng - 1, X, Z);
  bad = dual_merge<NS>(X, Z, X, false, W, Lr);
  dual_store_record<NS>(rec, W);
  if (bad && lane == 0) { b.status[(size_t)prob * 3 * N + rs_slot_S(k0 >> 1, 0, N)] = -1; b.flags[prob] = 1; }
  __syncwarp();
  // knots (2,3): separator after local knot 2
  dual_leaf_pair<NS, NI>(blk(2), zlam(2), blk(3), zlam(3), false, s.kofs + k0 + 3 == s.Ng - 1, Y, Z);
  bad = dual_merge<NS>(Y, Z, Y, true, W, Lr);
  dual_store_record<NS>(rec + 2 * D::WSZ, W);
  if (bad && lane == 0) { b.status[(size_t)prob * 3 * N + rs_slot_S((k0 >> 1) + 1, 0, N)] = -1; b.flags[prob] = 1; }
  __syncwarp();
  // level 1: separator after local knot 1
  bad = dual_merge<NS>(X, Y, M0 > 1 ? X : nullptr, (grp & 1) != 0, W, Lr);
  dual_store_record<NS>(rec + D::WSZ, W);
  if (bad && lane == 0) { b.status[(size_t)prob * 3 * N + rs_slot_S(k0 >> 2, 1, N)] = -1; b.flags[prob] = 1; }
  if (M0 > 1) {
    fence_async_smem();
    __syncwarp();
    if (lane == 0) {
      bulk_s2g(ent0 + ((size_t)prob * M0 + grp) * D::ENT, X, D::ENT * (int)sizeof(double));
      bulk_commit();
      bulk_wait_read0();  // shared memory must stay valid until the copy engine has read it
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// Tier kernel: one CTA = (problem, `ecount` consecutive entries), log2(ecount) more levels, one warp per merge, merged
// in place. Entry e of a tier is L-type for even e, R-type for odd e. Records go to `sep` (slot = index of the 4-knot
// subtree that starts right after the separator), the CTA's merged entry to `out`.
// ---------------------------------------------------------------------------------------------------------
struct DualTierArgs {
  const double* in;      // element e of problem p: in + (p*Min + e)*in_stride
  long long in_stride;   // ENT when the elements are packed (one bulk copy per CTA), else one copy per element
  double* out;           // merged element of CTA c of problem p: out + (p*ncta + c)*out_stride; nullptr = none
  long long out_stride;
  double* sep;           // record of the separator after element e: sep + p*sep_pstride + (e+1)*sep_mul*WSZ
  long long sep_pstride;
  int sep_mul;
  int Min, ecount;
  int span;              // knots per input element (status words only; 0 = cross-GPU tier: flags only)
  int root_out;          // the last merge of a single-CTA tier also produces its element (a shard's contribution), R-type
  int root_right;        // if root_right (the shard's rank is odd)
  // cross-GPU tier fed by peer stores over NVLink (shard_push_kernel): element e is complete once wait_flags[e] >= wait_epoch
  const unsigned long long* wait_flags;
  unsigned long long wait_epoch;
  double* top_lam;       // single-CTA tier: also resolve the multipliers of its separators top-down (topdown_levels) into top_lam
};

template <int NS>
__host__ __device__ constexpr int dual_tier_warp_scratch() { return ((DualDim<NS>::WP + 1) & ~1) + DualDim<NS>::LR; }

template <int NS>
__global__ void __launch_bounds__(128) dual_tier_kernel(Shape s, Bufs b, DualTierArgs a) {
  pdl_sync();
  using D = DualDim<NS>;
  extern __shared__ __align__(16) double smf[];
  __shared__ __align__(8) unsigned long long mbar;
  const int tid = threadIdx.x, warp = warp_index_sync(), lane = tid & 31;
  const int ecount = a.ecount, ncta = a.Min / ecount;
  const int prob = blockIdx.x / ncta, sg = blockIdx.x % ncta;
  double* ents = smf;
  double* scr = ents + ecount * D::ENT + warp * dual_tier_warp_scratch<NS>();
  double* W = scr;
  double* Lr = scr + ((D::WP + 1) & ~1);
  const double* in = a.in + ((size_t)prob * a.Min + (size_t)sg * ecount) * a.in_stride;
  if (tid == 0) {
    mbar_init(&mbar, 1);
    fence_mbar_init();
  }
  __syncthreads();
  if (a.wait_flags) {  // elements arrive by peer stores: wait for their flags (bounded: a lost peer must not hang the GPU)
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
  if (tid == 0) {
    mbar_expect_tx(&mbar, ecount * D::ENT * (int)sizeof(double));
    if (a.in_stride == D::ENT) bulk_g2s(ents, in, ecount * D::ENT * (int)sizeof(double), &mbar);
    else
      for (int e = 0; e < ecount; ++e) bulk_g2s(ents + e * D::ENT, in + e * a.in_stride, D::ENT * (int)sizeof(double), &mbar);
  }
  mbar_wait(&mbar, 0);
  int levels = 0;
  while ((1 << levels) < ecount) ++levels;
  for (int l = 0; l < levels; ++l) {
    if (l > 0) __syncthreads();
    const int nsep = ecount >> (l + 1);
    if (warp < nsep) {
      const int el = (2 * warp) << l, er = el + (1 << l);   // merged elements live in the slot of their first element
      const int ge_last = sg * ecount + er - 1;             // last input element of the left subtree
      const int oidx = (sg * ecount + el) >> (l + 1);       // index of the merged subtree at its level
      const bool root = a.Min == ecount && l == levels - 1;
      double* EL = ents + el * D::ENT;
      const bool oright = root ? a.root_right != 0 : (oidx & 1) != 0;
      const int bad = dual_merge<NS>(EL, ents + er * D::ENT, (root && !a.root_out) ? nullptr : EL, oright, W, Lr);
      dual_store_record<NS>(a.sep + (size_t)prob * a.sep_pstride + (size_t)(ge_last + 1) * a.sep_mul * D::WSZ, W);
      if (bad && lane == 0) {
        if (a.span > 0) {
          const int i = (ge_last + 1) * a.span - 1;         // local knot index of the separator
          int lev = 0;
          while ((i >> lev) & 1) ++lev;
          b.status[(size_t)prob * 3 * s.N + rs_slot_S(i >> (lev + 1), lev, s.N)] = -1;
        }
        b.flags[prob] = 1;
      }
    }
  }
  if (a.top_lam) {  // (the records of this CTA's separators are in global memory: visible to the CTA after the barrier)
    __syncthreads();
    topdown_levels<NS>(a.sep, a.top_lam, a.Min, nullptr, nullptr, a.Min, 1, 0);
  }
  if (a.out) {
    __syncthreads();
    fence_async_smem();
    __syncthreads();
    if (tid == 0) {
      bulk_s2g(a.out + ((size_t)prob * ncta + sg) * a.out_stride, ents, D::ENT * (int)sizeof(double));
      bulk_commit();
      bulk_wait_read0();
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------
// shapes with a compiled instance of the dual-form kernels: n <= 16, m <= 16. One line here = one more shape on the scan path (kernels are templates on (n, m)).
#define RSLQR_DUAL_SHAPES(X) \
  X(6, 3) X(12, 4) X(12, 6) X(8, 4) X(16, 8) X(4, 2) /* round-2 list */ \
  X(6, 2) X(8, 2) X(8, 3) X(10, 4) X(10, 5) X(12, 2) X(12, 3) X(12, 8) X(14, 4) X(14, 6) X(16, 4) X(16, 6) \
  X(3, 1) X(5, 2) X(7, 3) X(9, 3) X(11, 4) X(13, 4) X(15, 5) /* odd n: blocks are padded to a multiple of 4 on chip */ \
  X(2, 1) X(4, 1) X(4, 4) X(6, 6) X(7, 7) X(8, 8) X(14, 7) X(16, 16)

inline bool dual_shape_ok(int n, int m) {
#define RSLQR_X(N_, M_) if (n == N_ && m == M_) return true;
  RSLQR_DUAL_SHAPES(RSLQR_X)
#undef RSLQR_X
  return false;
}

template <int NS>
inline void dual_launch_tier(const Shape& s, const Bufs& b, const DualTierArgs& a, int batch, cudaStream_t st) {
  using D = DualDim<NS>;
  const int warps = a.ecount / 2 > 0 ? a.ecount / 2 : 1;
  const int sm = (a.ecount * D::ENT + warps * dual_tier_warp_scratch<NS>()) * (int)sizeof(double);
  launch_k(dual_tier_kernel<NS>, dim3((unsigned)((size_t)batch * (a.Min / a.ecount))), dim3(32 * warps), (size_t)sm, st, s, b, a);
}

// the local tiers of the plan. last_out: where the last tier puts the merged element of the whole local tree (a shard's
// contribution to the exchange), R-type if root_right; nullptr = the local tree is the whole problem.
template <int NS>
inline cudaError_t dual_launch_tiers(ScanPlan& p, const Shape& s, const Bufs& b, int batch, cudaStream_t st, double* last_out,
                                     long long last_out_stride, int root_right) {
  using D = DualDim<NS>;
  FusedPlan& f = p.f;
  cudaError_t e = smem_optin(dual_tier_kernel<NS>, (8 * D::ENT + 4 * dual_tier_warp_scratch<NS>()) * (int)sizeof(double));
  if (e != cudaSuccess) return e;
  for (int t = 0; t < f.ntiers; ++t) {
    const FusedTier& T = f.tiers[t];
    const bool last = t == f.ntiers - 1;
    DualTierArgs a{};
    a.in = f.fb.ent[t]; a.in_stride = D::ENT;
    a.out = last ? last_out : f.fb.ent[t + 1];
    a.out_stride = last ? last_out_stride : D::ENT;
    a.sep = f.fb.sep; a.sep_pstride = (long long)f.M0 * D::WSZ; a.sep_mul = T.span / SCAN_LEAF;
    a.Min = T.Min; a.ecount = T.ecount; a.span = T.span;
    a.root_out = last && last_out ? 1 : 0; a.root_right = root_right;
    dual_launch_tier<NS>(s, b, a, batch, st);
    ++f.launches;
  }
  return cudaGetLastError();
}

template <int NS, int NI>
inline cudaError_t dual_launch_t(ScanPlan& p, const Shape& s, const Bufs& b, int batch, int pack_rhs, cudaStream_t st,
                                 cudaEvent_t k0, cudaEvent_t k1) {
  using D = DualDim<NS>;
  FusedPlan& f = p.f;
  cudaError_t e;
  f.launches = 1;
  if (k0) cudaEventRecord(k0, st);
  if constexpr (NS == 12 && NI == 4) {
    if (p.hybrid) {  // boundary-row leaf kernel (levels 0..1), dual elements out
      e = launch_scan_leaf<12, 4>(p, s, b, batch, pack_rhs, st);
      if (e != cudaSuccess) return e;
    }
  }
  if (!p.hybrid) {
    constexpr int DUAL_WPC = dual_leaf_wpc<NS, NI>();
    const int smem = DUAL_WPC * dual_leaf_warp_doubles<NS, NI>() * (int)sizeof(double);
    e = smem_optin(dual_leaf_kernel<NS, NI>, smem);
    if (e != cudaSuccess) return e;
    const long long nsub = (long long)batch * f.M0;
    launch_k(dual_leaf_kernel<NS, NI>, dim3((unsigned)((nsub + DUAL_WPC - 1) / DUAL_WPC)), dim3(32 * DUAL_WPC), (size_t)smem, st, s, b,
             p.sb.ent0, p.sb.sep_in, pack_rhs, f.M0, nsub);
  }
  if (k1) cudaEventRecord(k1, st);
  e = dual_launch_tiers<NS>(p, s, b, batch, st, nullptr, 0, 0);
  if (e != cudaSuccess) return e;
  return launch_scan_down<NS, NI>(p, s, b, batch, pack_rhs, st, nullptr, nullptr);
}

inline int dual_launch(ScanPlan& p, const Shape& s, const Bufs& b, int batch, int pack_rhs, cudaStream_t st, std::string& err,
                       cudaEvent_t k0, cudaEvent_t k1) {
  if (!p.f.available || !(p.dual || p.hybrid)) { err = "shape not covered by the dual-form scan kernels"; return -5; }
  cudaError_t e;
  bool found = false;
  e = cudaSuccess;
#define RSLQR_X(N_, M_) \
  if (!found && s.n == N_ && s.m == M_) { found = true; e = dual_launch_t<N_, M_>(p, s, b, batch, pack_rhs, st, k0, k1); }
  RSLQR_DUAL_SHAPES(RSLQR_X)
#undef RSLQR_X
  if (!found) { err = "shape not covered by the dual-form scan kernels"; return -5; }
  if (e != cudaSuccess) { err = std::string("dual launch: ") + cudaGetErrorString(e); return -3; }
  return 0;
}

// ---- shards of one long horizon in hybrid form (n = 12, m = 4): the exchange payload is the shard's dual element (its
//      Schur complement onto lambda of its first knot and of the first knot of the next shard), nothing else ----
inline int dual_shard_reduce(ScanPlan& p, const Shape& s, const Bufs& b, int pack_rhs, cudaStream_t st, double* payload,
                             std::string& err) {
  using D = DualDim<12>;
  if (!p.f.available || !p.hybrid || p.f.ntiers < 1) { err = "shard needs the hybrid scan kernels and at least two subtrees"; return -5; }
  p.f.launches = 1;
  cudaError_t e = launch_scan_leaf<12, 4>(p, s, b, 1, pack_rhs, st);
  if (e == cudaSuccess) e = dual_launch_tiers<12>(p, s, b, 1, st, payload, D::ENT, (s.kofs / s.N) & 1);
  if (e != cudaSuccess) { err = std::string("shard reduce: ") + cudaGetErrorString(e); return -3; }
  return 0;
}

inline int dual_shard_finish(ScanPlan& p, const Shape& s, const Bufs& b, int pack_rhs, cudaStream_t st, const double* gathered,
                             int rank, int nranks, std::string& err, const unsigned long long* wait_flags = nullptr,
                             unsigned long long wait_epoch = 0) {
  using D = DualDim<12>;
  if (!p.f.available || !p.hybrid || p.f.ntiers < 1) { err = "shard needs the hybrid scan kernels"; return -5; }
  const double *left = nullptr, *right = nullptr;
  cudaError_t e = cudaSuccess;
  if (nranks > 1) {
    e = smem_optin(dual_tier_kernel<12>, (8 * D::ENT + 4 * dual_tier_warp_scratch<12>()) * (int)sizeof(double));
    if (e == cudaSuccess) {
      DualTierArgs a{};
      a.in = gathered; a.in_stride = D::ENT;
      a.out = nullptr; a.out_stride = 0;
      a.sep = p.f.xsep; a.sep_pstride = 0; a.sep_mul = 1;
      a.Min = nranks; a.ecount = nranks; a.span = 0;
      a.wait_flags = wait_flags; a.wait_epoch = wait_epoch;
      a.top_lam = p.f.xlam;                  // ... and the multipliers of those separators (same CTA, no second launch)
      dual_launch_tier<12>(s, b, a, 1, st);  // every rank runs the top log2(nranks) levels redundantly, bit-identically
      p.f.launches += 1;
      left = p.f.xlam + (size_t)rank * 12;
      right = p.f.xlam + (size_t)(rank + 1) * 12;
    }
  }
  if (e == cudaSuccess) e = launch_scan_down<12, 4>(p, s, b, 1, pack_rhs, st, left, right);
  if (e != cudaSuccess) { err = std::string("shard finish: ") + cudaGetErrorString(e); return -3; }
  return 0;
}

}  // namespace rslqr

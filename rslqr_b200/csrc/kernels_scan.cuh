// kernels_scan.cuh -- "scan" device path: the whole nested dissection on boundary entries, one warp per merge.
//
// The tiered path (kernels_fused.cuh) keeps the dense state (L, R, z) of every knot for its first levels and
// a CTA cooperates on a group of knots, so most of its warps wait at the level barriers while one warp per
// separator runs the serial factorization chain. Here EVERY level works on boundary entries only -- the rows
// [X(a) | X(b) | U(b)] of the two boundary columns of a subtree [a,b], all a later ndlqr_FactorInnerProduct
// (reference src/nested_dissection.c:114) ever reads -- and a warp owns a subtree outright:
//
//   scan_leaf_kernel : one warp = 4 consecutive knots. The two level-0 merges are formed straight from the
//                      flat knot blocks (the leaf factors of ndlqr_SolveLeaf, src/nested_dissection.c:10-105,
//                      are diagonal scalings, so S = A Q^-1 A^T + B R^-1 B^T + Q^-1 and the Schur update of a
//                      leaf pair is one 12x12x24 product), the level-1 merge is the general one. No CTA barrier.
//   tier_kernel      : (kernels_fused.cuh) the remaining levels, 4 (batches) or 8 (one long horizon) entries per CTA.
//   topdown_kernel   : (kernels_fused.cuh) multipliers of the separators above the 4-knot subtrees.
//   scan_down_kernel : one warp = 4 knots: multipliers of its three interior separators from the stored records
//                      [f_L | f_R | zs], then x_k, u_k from the leaf equations (the reference's rhs propagation
//                      src/solve.c:175-180 telescoped: soln(k) = z(k) - L(k) lambda_k - R(k) lambda_{k+1} with the
//                      LEAF blocks L(k) = [0; -Q^-1; 0], R(k) = [0; Q^-1 A^T; R^-1 B^T]).
//
// Nothing but entries (5.7 KB per subtree) and separator records (2.4 KB each) is written to HBM.
// Arithmetic: IEEE fp64; inner products and Schur updates on mma.sync.m8n8k4.f64 (DMMA).
#pragma once
#include "kernels_fused.cuh"

namespace rslqr {

// entry `rw` of the negated right-hand side of local knot k (reference src/solver.c:138-190), either packed on the
// fly from the flat knot blocks or read from the soln buffer (pack_rhs == 0: the caller uploaded its own rhs)
__device__ __forceinline__ double scan_rhs(const Shape& s, const Bufs& b, int prob, int k, int rw, int pack_rhs) {
  if (!pack_rhs) return soln_ptr(s, b, prob, k)[rw];
  const double* kn = knot_ptr(s, b, prob, k);
  if (rw < s.n) return -(k == 0 ? b.x0[(size_t)prob * s.n + rw] : (kn - s.stride)[s.od + rw]);
  if (rw < 2 * s.n) return -kn[s.oq + rw - s.n];
  return (s.kofs + k < s.Ng - 1) ? -kn[s.orr + rw - 2 * s.n] : 0.0;
}

// ---------------------------------------------------------------------------------------------------------
// Shared-memory image of a warp's SCAN_LEAF flat knot blocks, brought in by ONE TMA bulk copy and rewritten in
// place into what the leaf stage needs (ndlqr_SolveLeaf on diagonal Q, R, src/nested_dissection.c:10-105, with
// the rhs packing of src/solver.c:138-190):
//   region R = [ 12 doubles | block 0 | block 1 | ... ],  block j at R + NS + j*STRIDE  (STRIDE = flat knot block)
//   block:  [0, NS+NI)        1/Q, 1/R            (was Q, R; 1/R = 0 on the last knot of the horizon)
//           [NS+NI, 2(NS+NI)) leaf-solved z_x, z_u (was q, r)
//           [OA, OD)          A, B column-major    (untouched)
//           [OD, OD+NS)       z_lambda of the NEXT knot = -d   (was d)
//   R[0..NS)                  z_lambda of block 0 (from the knot before the subtree, x0 or the knot-0 special)
// Because OD + NS == STRIDE, z_lambda of block j is R + j*STRIDE for every j.
// ---------------------------------------------------------------------------------------------------------
template <int NS, int NI>
struct RawDim {
  static constexpr int VEC = 2 * NS + 2 * NI, AB = NS * (NS + NI);
  static constexpr int STRIDE = VEC + 1 + AB + NS;
  static constexpr int OZ = NS + NI, OA = VEC + 1, OD = OA + AB;
  static_assert(OD + NS == STRIDE, "z_lambda of block j sits at R + j*STRIDE");
};

// zl0 (lanes < NS): -z_lambda of block 0 when the rhs is packed here, loaded by the caller BEFORE it waits for the
// knot blocks so that this second DRAM round trip overlaps the first (leaf_zl0).
template <int NS, int NI>
__device__ __forceinline__ double leaf_zl0(const Shape& s, const Bufs& b, int prob, int k0, int pack_rhs) {
  using RD = RawDim<NS, NI>;
  const int lane = threadIdx.x & 31;
  if (!pack_rhs || lane >= NS) return 0.0;
  return k0 == 0 ? b.x0[(size_t)prob * NS + lane] : __ldg(knot_ptr(s, b, prob, k0) - RD::STRIDE + RD::OD + lane);
}

template <int NS, int NI, int NK>
__device__ __forceinline__ void leaf_transform(const Shape& s, const Bufs& b, int prob, int k0, int pack_rhs, double* R, double zl0) {
  using RD = RawDim<NS, NI>;
  constexpr int XU = NS + NI;
  const int lane = threadIdx.x & 31;
  const int N = s.N;
  // z_lambda: -d in place, or the caller's rhs
#pragma unroll 1
  for (int it = lane; it < NK * NS; it += 32) {
    const int j = it / NS, r = it - j * NS;
    double v;
    if (!pack_rhs) v = soln_ptr(s, b, prob, k0 + j)[r];
    else if (j == 0) v = -zl0;  // (it < NS here: r == lane)
    else v = -R[j * RD::STRIDE + r];
    R[j * RD::STRIDE + r] = v;
  }
  __syncwarp();
#pragma unroll 1
  for (int it = lane; it < NK * XU; it += 32) {
    const int j = it / XU, l = it - j * XU;  // (XU is a compile-time constant: a shift for XU == 16)
    const int k = k0 + j, kg = s.kofs + k;
    const bool last = kg == s.Ng - 1, isR = l >= NS;
    double* B = R + NS + j * RD::STRIDE;
    const double d = B[l];
    if (!(d > 0.0) && !(isR && last)) {
      b.status[(size_t)prob * 3 * N + (isR ? rs_slot_R(k) : rs_slot_Q(k))] = -1;
      b.flags[prob] = 1;
    }
    const double dinv = (isR && last) ? 0.0 : 1.0 / d;
    double rx;
    if (!pack_rhs) rx = soln_ptr(s, b, prob, k)[NS + l];
    else rx = (isR && last) ? 0.0 : -B[RD::OZ + l];
    double zi;
    if (kg == 0) {
      if (!isR) {
        const double rl = R[l];  // rhs_lambda of knot 0 (-x0)
        zi = -rl;
        R[l] = -d * rl - rx;     // z_lambda(0) = Q x0 + q
      } else zi = rx * dinv;
    } else {
      zi = (isR && last) ? rx : rx * dinv;
    }
    B[l] = dinv;
    B[RD::OZ + l] = zi;
  }
  __syncwarp();
}

// ---------------------------------------------------------------------------------------------------------
// Level-0 merge of two single knots i (left) and i+1 (right), straight from their transformed blocks Bl, Br:
//   S   = A_i Q_i^-1 A_i^T + B_i R_i^-1 B_i^T + Q_{i+1}^-1      (ndlqr_FactorInnerProduct on the leaf factors)
//   S_L = -A_i Q_i^-1,   S_R = -Q_{i+1}^-1 A_{i+1}^T,   s = A_i z_x(i) + B_i z_u(i) - z_x(i+1) - z_lambda(i+1)
//   [f_L f_R zs] = S^-1 [S_L S_R s]                               (Cholesky + ndlqr_SolveCholeskyFactor)
//   entry rows X(i)   : -Q_i^-1 ([I|0] + A_i^T [f_L f_R]),  z_x(i) - Q_i^-1 A_i^T zs        (ndlqr_UpdateShurFactor)
//         rows X(i+1) :  Q_{i+1}^-1 ([0|A_{i+1}^T] + [f_L f_R]),  z_x(i+1) + Q_{i+1}^-1 zs
//         rows U(i+1) :  [0 | R_{i+1}^-1 B_{i+1}^T],  z_u(i+1)                              (unchanged)
// Knot 0 has X_R = 0 and Lambda = -A_0^T (src/nested_dissection.c:24-59): its X rows stay (0, 0, x0).
// The last knot of the horizon has R(k) = 0. zlL / zlR: z_lambda of the two knots.
// E: output entry (Dim::ENT layout; its head doubles as the factor scratch), Wf: [f_L | f_R | zs] on return.
// ---------------------------------------------------------------------------------------------------------
// before_write(): called once, after the operand loads and tensor-core products and before the first store into E --
// the caller's hook for "the copy engine must have finished reading what used to be there".
template <int NS, int NI, class Hook>
__device__ __forceinline__ int merge_raw(const double* Bl, const double* Br, const double* zlL, const double* zlR, bool first0,
                                         bool lastR, double* E, double* Wf, Hook&& before_write) {
  using D = Dim<NS, NI>;
  using RD = RawDim<NS, NI>;
  constexpr int ZS = D::ZS, SLOT = D::SLOT;
  constexpr int KS = (NS + NI) / 4, KA = NS / 4;
  static_assert(NS % 4 == 0 && (NS + NI) % 4 == 0 && NS > 8 && NS <= 16 && 2 * NS == 24 && NS + NI == 16, "tile shapes");
  const int lane = threadIdx.x & 31, q = lane & 3, rr = lane >> 2;
  const bool r1 = 8 + rr < NS;  // second row tile is partial
  const double* ABl = Bl + RD::OA;
  const double* ABr = Br + RD::OA;
  double* Lsm = E;                                   // factor scratch: head of the entry being produced
  constexpr int LDP = NS + 1;                        // parking area for S / S_L / s, odd leading dimension
  double* P = E + factor_scratch_doubles<NS>();
  static_assert(factor_scratch_doubles<NS>() + (2 * NS + 1) * LDP <= D::ENT, "scratch fits in the entry");
  double acc[NS];
  {
    double a0[KS], a1[KS];
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      const int l = 4 * ks + q;
      a0[ks] = ABl[l * NS + rr];
      a1[ks] = r1 ? ABl[l * NS + 8 + rr] : 0.0;
    }
    double d[2][2][2];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int nt = 0; nt < 2; ++nt) d[mt][nt][0] = d[mt][nt][1] = 0.0;
    double p0 = 0.0, p1 = 0.0;
    double sl0[KA], sl1[KA];
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      double dk = Bl[4 * ks + q];
      if (first0 && ks < KA) dk = 0.0;  // X_R(0) = 0
      const double zk = Bl[RD::OZ + 4 * ks + q];
      const double b0 = a0[ks] * dk, b1 = a1[ks] * dk;
      dmma884(d[0][0][0], d[0][0][1], a0[ks], b0);
      dmma884(d[0][1][0], d[0][1][1], a0[ks], b1);
      dmma884(d[1][0][0], d[1][0][1], a1[ks], b0);
      dmma884(d[1][1][0], d[1][1][1], a1[ks], b1);
      p0 = fma(a0[ks], zk, p0);
      p1 = fma(a1[ks], zk, p1);
      if (ks < KA) { sl0[ks] = -b0; sl1[ks] = -b1; }
    }
    p0 += __shfl_xor_sync(0xffffffffu, p0, 1);
    p0 += __shfl_xor_sync(0xffffffffu, p0, 2);
    p1 += __shfl_xor_sync(0xffffffffu, p1, 1);
    p1 += __shfl_xor_sync(0xffffffffu, p1, 2);
    before_write();
#pragma unroll
    for (int ks = 0; ks < KA; ++ks) {  // the scaled operand IS -S_L: S_L(r, l) = -A_i(r, l) / Q_i(l), parked as column NS + l
      P[(NS + 4 * ks + q) * LDP + rr] = sl0[ks];
      if (r1) P[(NS + 4 * ks + q) * LDP + 8 + rr] = sl1[ks];
    }
#pragma unroll
    for (int mt = 0; mt < 2; ++mt) {
      const int row = 8 * mt + rr, rs = row < NS ? row : 0;
      const double dr = Br[rs];
      const double t = Br[RD::OZ + rs] + zlR[rs];
#pragma unroll
      for (int nt = 0; nt < 2; ++nt)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int col = 8 * nt + 2 * q + h;
          if (row < NS && col < NS) P[col * LDP + row] = d[mt][nt][h] + (col == row ? dr : 0.0);
        }
      if (q == 0 && row < NS) P[2 * NS * LDP + row] = (mt ? p1 : p0) - t;
    }
    __syncwarp();
    // lane c <= 2NS takes column c of the parking area [S | S_L | s] (odd leading dimension: conflict-free)
    const int mc = lane <= 2 * NS ? lane : 2 * NS;
#pragma unroll
    for (int r = 0; r < NS; ++r) acc[r] = P[mc * LDP + r];
  }
  // S_R(:,c) = -A_{i+1}(c,:)^T / Q_{i+1} straight into the second register column of lanes [NS,2NS)
  double acc2[NS];
  {
    const bool isL = lane >= NS && lane < 2 * NS;
    const int cl = isL ? lane - NS : 0;
#pragma unroll
    for (int r = 0; r < NS; ++r) {
      const double ar = ABr[r * NS + cl], dr = Br[r];
      acc2[r] = (isL && !lastR) ? -ar * dr : 0.0;
    }
  }
  const int bad = spine_core<NS>(acc, acc2, Lsm, Wf);
  // ---- entry rows X(i+1), U(i+1) straight from the solved columns still in registers: lane NS+c holds f_L(:,c)
  //      (acc) and f_R(:,c) (acc2), lane 2NS holds zs. Row r of the entry is a scaling by 1/Q_{i+1}(r):
  //      L = f_L/Q, R = (A_{i+1}^T + f_R)/Q, z = z_x + zs/Q; the U rows are the leaf values. Every lane loads
  //      unconditionally; the stores are predicated per lane (no branches, no traffic from idle lanes). ----
  {
    const bool isL = lane >= NS && lane < 2 * NS, isz = lane == 2 * NS;
    const int cc = isL ? lane - NS : 0;
    const int act1 = isL || isz, act2 = isL;
    const unsigned d1 = smem_u32(isz ? E + 2 * SLOT + NS : E + NS * NS + cc);  // L part / z
    const unsigned d2 = smem_u32(E + SLOT + NS * NS + cc);                      // R part
    const unsigned s1 = (isz ? 1 : NS) * (unsigned)sizeof(double), s2 = NS * (unsigned)sizeof(double);
    const bool addA = isL && !lastR;
#pragma unroll
    for (int r = 0; r < NS; ++r) {
      const double dr = Br[r];
      const double al = ABr[r * NS + cc];
      const double zx = Br[RD::OZ + r];
      sts_pred(d1 + r * s1, fma(dr, acc[r], isz ? zx : 0.0), act1);
      sts_pred(d2 + r * s2, dr * (acc2[r] + (addA ? al : 0.0)), act2);
    }
#pragma unroll
    for (int ru = 0; ru < NI; ++ru) {
      const double du = Br[NS + ru];
      const double bl = ABr[(NS + ru) * NS + cc];
      const double zu = Br[RD::OZ + NS + ru];
      sts_pred(d1 + (NS + ru) * s1, isz ? zu : 0.0, act1);
      sts_pred(d2 + (NS + ru) * s2, addA ? du * bl : 0.0, act2);
    }
    if (lane < NS) E[2 * SLOT + ZS + lane] = zlL[lane];
  }
  __syncwarp();
  // ---- entry rows X(i): A_i^T [f_L f_R] on DMMA ----
  {
    double bf[KA][3], zq[KA];
#pragma unroll
    for (int ks = 0; ks < KA; ++ks) {
#pragma unroll
      for (int nt = 0; nt < 3; ++nt) bf[ks][nt] = Wf[(8 * nt + rr) * NS + 4 * ks + q];
      zq[ks] = Wf[2 * NS * NS + 4 * ks + q];
    }
#pragma unroll
    for (int mt = 0; mt < 2; ++mt) {
      const int row = 8 * mt + rr, rs = row < NS ? row : 0;
      double at[KA];
#pragma unroll
      for (int ks = 0; ks < KA; ++ks) at[ks] = row < NS ? ABl[rs * NS + 4 * ks + q] : 0.0;
      double c[3][2];
#pragma unroll
      for (int nt = 0; nt < 3; ++nt) c[nt][0] = c[nt][1] = 0.0;
      double part = 0.0;
#pragma unroll
      for (int ks = 0; ks < KA; ++ks) {
#pragma unroll
        for (int nt = 0; nt < 3; ++nt) dmma884(c[nt][0], c[nt][1], at[ks], bf[ks][nt]);
        part = fma(at[ks], zq[ks], part);
      }
      part += __shfl_xor_sync(0xffffffffu, part, 1);
      part += __shfl_xor_sync(0xffffffffu, part, 2);
      const double dl = first0 ? 0.0 : Bl[rs];
      const double zrow = Bl[RD::OZ + rs];
      if (row < NS) {
#pragma unroll
        for (int nt = 0; nt < 3; ++nt) {
          const int col = 8 * nt + 2 * q;
          const double v0 = -dl * ((col == row ? 1.0 : 0.0) + c[nt][0]);
          const double v1 = -dl * ((col + 1 == row ? 1.0 : 0.0) + c[nt][1]);
          double* dst = col < NS ? E + row * NS + col : E + SLOT + row * NS + col - NS;
          *reinterpret_cast<double2*>(dst) = make_double2(v0, v1);
        }
        if (q == 0) E[2 * SLOT + row] = zrow - dl * part;
      }
    }
  }
  __syncwarp();
  return bad;
}


// ---------------------------------------------------------------------------------------------------------
// Boundary entry of a subtree [a,b] (rows [X(a) | X(b) | U(b)] of L, R, z and z_lambda(a)) -> its dual-form element
// (kernels_dual.cuh), so that the levels above the leaf kernel run on n x n blocks only. With
//   soln(k) = z(k) - L(k) lambda_a - R(k) lambda_{b+1}
// the rows of the multiplier system the subtree contributes to are
//   lambda_a     :  x_a                       ->  Saa = -L_X(a), Sab = -R_X(a), sa = -z_x(a) - z_lambda(a)
//   lambda_{b+1} : -(A_b x_b + B_b u_b)       ->  [Sba | Sbb | sb] = [A_b | B_b] [L | R | z]_{X,U rows of b}
// the second being the product every later ndlqr_FactorInnerProduct (src/nested_dissection.c:114) would apply to
// these rows: it is applied here once. E: entry (Dim layout), AB: [A_b | B_b] (global memory if LDG, else shared), Eo: dual entry
// (DualDim layout, R-type if `right` else L-type; must not alias E). One warp.
// ---------------------------------------------------------------------------------------------------------
template <int NS, int NI, bool LDG = true>
__device__ __forceinline__ void entry_to_dual(const double* E, const double* __restrict__ AB, double* Eo, bool right, bool lastb) {
  using D = Dim<NS, NI>;
  using DD = DualDim<NS>;
  constexpr int SLOT = D::SLOT, ZS = D::ZS, LD = DD::LD, KS = (NS + NI) / 4;
  static_assert(NS % 4 == 0 && (NS + NI) % 4 == 0 && NS > 8 && NS <= 16 && DD::NP == NS, "tile shapes");
  const int lane = threadIdx.x & 31, q = lane & 3, rr = lane >> 2;
  const double* LX = E + NS * NS;            // rows X(b), U(b) of the L block
  const double* RX = E + SLOT + NS * NS;
  const double* zx = E + 2 * SLOT + NS;
  double d[2][4][2];
#pragma unroll
  for (int mt = 0; mt < 2; ++mt)
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) d[mt][nt][0] = d[mt][nt][1] = 0.0;
#pragma unroll
  for (int ks = 0; ks < KS; ++ks) {
    const int l = 4 * ks + q;
    // (b = last knot of the horizon: its A, B are ignored, SURVEY A.1 -- there is no multiplier behind it to couple to)
    const double a0 = lastb ? 0.0 : (LDG ? __ldg(AB + l * NS + rr) : AB[l * NS + rr]);
    const double a1 = (8 + rr < NS && !lastb) ? (LDG ? __ldg(AB + l * NS + 8 + rr) : AB[l * NS + 8 + rr]) : 0.0;
    double bf[4];
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
      const int n = 8 * nt + rr;
      bf[nt] = n < NS ? LX[l * NS + n] : (n < 2 * NS ? RX[l * NS + n - NS] : (n == 2 * NS ? zx[l] : 0.0));
    }
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
      dmma884(d[0][nt][0], d[0][nt][1], a0, bf[nt]);
      dmma884(d[1][nt][0], d[1][nt][1], a1, bf[nt]);
    }
  }
  // [Sba | Sbb | sb]: an R-type element keeps all three (Cp, Kp, ks), an L-type one only Sbb, sb (Sp, ss)
#pragma unroll
  for (int mt = 0; mt < 2; ++mt) {
    const int row = 8 * mt + rr;
    if (row < NS) {
#pragma unroll
      for (int nt = 0; nt < 3; ++nt) {
        const int col = 8 * nt + 2 * q;
        const double2 v = make_double2(d[mt][nt][0], d[mt][nt][1]);
        if (col < NS) {
          if (right) *reinterpret_cast<double2*>(Eo + DD::OC + row * LD + col) = v;
        } else {
          *reinterpret_cast<double2*>(Eo + (right ? DD::OK : DD::OS) + row * LD + col - NS) = v;
        }
      }
      if (q == 0) Eo[(right ? DD::OKS : DD::OSS) + row] = d[mt][3][0];
    }
  }
  // Saa = -L_X(a) (Kp of an L-type element, Sp of an R-type one), Sab = -R_X(a) (Cp, L-type only), sa
  for (int e = lane; e < NS * NS / 2; e += 32) {
    const int row = e / (NS / 2), col = 2 * (e - row * (NS / 2));
    const double2 l = *reinterpret_cast<const double2*>(E + row * NS + col);
    *reinterpret_cast<double2*>(Eo + (right ? DD::OS : DD::OK) + row * LD + col) = make_double2(-l.x, -l.y);
    if (!right) {
      const double2 r2 = *reinterpret_cast<const double2*>(E + SLOT + row * NS + col);
      *reinterpret_cast<double2*>(Eo + DD::OC + row * LD + col) = make_double2(-r2.x, -r2.y);
    }
  }
  if (lane < NS) Eo[(right ? DD::OSS : DD::OKS) + lane] = -E[2 * SLOT + lane] - E[2 * SLOT + ZS + lane];
}

struct ScanBufs {
  double* ent0;    // [batch][M0][ENT]   entries of the 4-knot subtrees (input of the first tier)
  double* sep_in;  // [batch][M0][3][WSZ] records of the interior separators of a subtree: local knots 1, 2, 3
};

constexpr int SCAN_LEAF = 4;   // knots per warp in the leaf kernel
constexpr int SCAN_WPC = 6;    // warps per CTA (independent of each other; sized for shared memory: 2 CTAs = 12 warps / SM)

// per warp: [NS | SCAN_LEAF flat knot blocks | entry 0 | entry 1]. The record W = [f_L | f_R | zs] of a merge has no
// region of its own: it is written where something is dead at that moment (see scan_leaf_kernel).
template <int NS, int NI>
__host__ __device__ constexpr int scan_leaf_warp_doubles() {
  using D = Dim<NS, NI>;
  return NS + SCAN_LEAF * RawDim<NS, NI>::STRIDE + 2 * D::ENT;
}
template <int NS, int NI>
__host__ __device__ constexpr int scan_leaf_smem_doubles() { return SCAN_WPC * scan_leaf_warp_doubles<NS, NI>(); }

template <int NS, int NI>
__global__ void __launch_bounds__(32 * SCAN_WPC, 2) scan_leaf_kernel(Shape s, Bufs b, ScanBufs sb, int pack_rhs, int M0,
                                                                      long long nsub, int dual_out) {
  pdl_sync();
  using D = Dim<NS, NI>;
  using RD = RawDim<NS, NI>;
  constexpr int ZS = D::ZS, SLOT = D::SLOT, WSZ = D::WSZ, ENT = D::ENT;
  constexpr int RAWB = SCAN_LEAF * RD::STRIDE * (int)sizeof(double);
  static_assert(RAWB % 16 == 0 && (NS * 8) % 16 == 0 && scan_leaf_warp_doubles<NS, NI>() % 2 == 0 && (WSZ * 8) % 16 == 0, "TMA alignment");
  extern __shared__ __align__(16) double smf[];
  __shared__ __align__(8) unsigned long long mbar[SCAN_WPC];
  const int warp = warp_index_sync(), lane = threadIdx.x & 31;
  const long long sub = (long long)blockIdx.x * SCAN_WPC + warp;
  if (sub >= nsub) return;
  double* R = smf + warp * scan_leaf_warp_doubles<NS, NI>();
  double* E0 = R + NS + SCAN_LEAF * RD::STRIDE;
  double* E1 = E0 + ENT;
  double* B23 = R + NS + 2 * RD::STRIDE;  // knot blocks 2, 3: dead once the merge (k0+2, k0+3) is done
  static_assert(WSZ + factor_scratch_doubles<NS>() <= 2 * RD::STRIDE && WSZ <= ENT && WSZ % 2 == 0 &&
                    (NS + 2 * RD::STRIDE) % 2 == 0 && (NS + SCAN_LEAF * RD::STRIDE) % 2 == 0,
                "W and the level-1 factor scratch fit in blocks 2, 3, 16-byte aligned");
  const int prob = (int)(sub / M0), grp = (int)(sub - (long long)prob * M0);
  const int k0 = grp * SCAN_LEAF;
  const int N = s.N;
  // one TMA bulk copy brings the subtree's flat knot blocks (contiguous in HBM) into shared memory
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
  double* rec = sb.sep_in + ((size_t)prob * M0 + grp) * 3 * WSZ;
  // record [f_L | f_R | zs] of a separator: TMA bulk store straight out of W; the region holding W may be rewritten
  // once the copy engine has read it (rec_done)
  auto store_rec = [&](int slot, const double* Wf) {
    fence_async_smem();
    __syncwarp();
    if (lane == 0) {
      bulk_s2g(rec + slot * WSZ, Wf, WSZ * (int)sizeof(double));
      bulk_commit();
    }
  };
  auto rec_done = [&]() {
    if (lane == 0) bulk_wait_read0();
    __syncwarp();
  };
  int bad;
  // ---- level 0: first (k0+2, k0+3) -> E1 with its W in the still unused E0, then (k0, k0+1) -> E0 with its W in the
  //      now dead blocks 2, 3 ----
#pragma unroll 1
  for (int h = 1; h >= 0; --h) {
    const int k = k0 + 2 * h;
    const double* Bl = R + NS + (2 * h) * RD::STRIDE;
    double* Wf = h ? E0 : B23;
    // (h == 0 writes into E0, which the copy engine may still be reading as the record of merge h == 1: the wait
    //  sits behind this merge's operand loads and products instead of right after the store)
    bad = merge_raw<NS, NI>(Bl, Bl + RD::STRIDE, R + (2 * h) * RD::STRIDE, R + (2 * h + 1) * RD::STRIDE, s.kofs + k == 0,
                            s.kofs + k + 1 == s.Ng - 1, h ? E1 : E0, Wf, rec_done);
    store_rec(2 * h, Wf);
    if (bad && lane == 0) {
      b.status[(size_t)prob * 3 * N + rs_slot_S(k >> 1, 0, N)] = -1;
      b.flags[prob] = 1;
    }
  }
  // ---- level 1: E0 + E1 -> E0 (separator after knot k0+1); W and the factor scratch in the dead blocks 0, 1 (the
  //      [A | B] of block 1 is the last thing read there: the hook syncs the warp before the first write). Block 3
  //      keeps its [A | B] for the conversion of the merged entry below. ----
  double* Wf = R + NS;
  static_assert((NS * 8) % 16 == 0, "W is 16-byte aligned behind the z_lambda slot");
  // [A | B] of the subtree's last knot was overwritten by the record of merge (k0, k0+1): once the copy engine has read
  // that record, fetch it again (8-byte cp.async, it is L2-resident) into its old place; it lands while the
  // factorization below runs and is what entry_to_dual multiplies the boundary rows with
  double* ABlast = R + NS + (SCAN_LEAF - 1) * RD::STRIDE + RD::OA;
  auto hook1 = [&]() {
    rec_done();
    if (dual_out && M0 > 1) {
      const double* src = knot_ptr(s, b, prob, k0 + SCAN_LEAF - 1) + s.oA;
      for (int e = lane; e < RD::AB; e += 32)
        asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(ABlast + e)), "l"(src + e) : "memory");
      asm volatile("cp.async.commit_group;" ::: "memory");
    }
  };
  bad = spine_merge<NS, NI, false, true>(R + NS + RD::STRIDE + RD::OA, E0 + NS * NS, E0 + SLOT + NS * NS, E0 + 2 * SLOT + NS, E1,
                                         E1 + SLOT, E1 + 2 * SLOT, E1 + 2 * SLOT + ZS, Wf + WSZ, Wf, hook1);
  store_rec(1, Wf);
  if (bad && lane == 0) {
    b.status[(size_t)prob * 3 * N + rs_slot_S(k0 >> 2, 1, N)] = -1;
    b.flags[prob] = 1;
  }
  if (M0 > 1) {  // (a single subtree is the whole problem: only the records are needed)
    lean_schur_dmma<NS, NI>(E0, E1, E0, E0, Wf);
    if (dual_out) {
      // the levels above run in dual form (kernels_dual.cuh): hand the subtree up as its Schur-complement element,
      // built in the now dead right entry
      using DD = DualDim<NS>;
      static_assert(DD::ENT <= ENT, "the dual element fits in the dead entry");
      asm volatile("cp.async.wait_all;" ::: "memory");
      __syncwarp();
      entry_to_dual<NS, NI, false>(E0, ABlast, E1, (grp & 1) != 0,
                                   s.kofs + k0 + SCAN_LEAF - 1 == s.Ng - 1);
      fence_async_smem();
      __syncwarp();
      if (lane == 0) {
        bulk_s2g(sb.ent0 + ((size_t)prob * M0 + grp) * DD::ENT, E1, DD::ENT * (int)sizeof(double));
        bulk_commit();
      }
    } else {
      fence_async_smem();
      __syncwarp();
      if (lane == 0) {
        bulk_s2g(sb.ent0 + ((size_t)prob * M0 + grp) * ENT, E0, ENT * (int)sizeof(double));
        bulk_commit();
      }
    }
  }
  if (lane == 0) bulk_wait_read0();  // shared memory must stay valid until the copy engine has read it
}


// ---------------------------------------------------------------------------------------------------------
// Down: one warp = one 4-knot subtree [4g, 4g+3]. lam[g], lam[g+1] are the multipliers of the separators
// bounding it. Interior multipliers from the records (rec 0/1/2 = separators at local knots 1/2/3):
//   l2 = zs1 - fL1 la - fR1 lb;  l1 = zs0 - fL0 la - fR0 l2;  l3 = zs2 - fL2 l2 - fR2 lb
// then for every knot k (lk = its own multiplier, ln = the next one):
//   x_k = Q_k^-1 (rhs_x + lk - A_k^T ln),  u_k = R_k^-1 (rhs_u - B_k^T ln),  lambda_k = lk
//   knot 0: x_0 = -rhs_lambda (= x0), lambda_0 = -Q_0 rhs_lambda - rhs_x + A_0^T l_1   (src/nested_dissection.c:24-59)
//   last knot: u slot = rhs_u (no input there)
// Two knots at a time: lanes 0..15 take the rows of one, lanes 16..31 of the next.
// ---------------------------------------------------------------------------------------------------------
template <int NS, int NI>
__global__ void __launch_bounds__(128) scan_down_kernel(Shape s, Bufs b, ScanBufs sb, const double* lam_all, int pack_rhs,
                                                         int M0, long long nsub) {
  pdl_sync();
  using D = Dim<NS, NI>;
  constexpr int WSZ = D::WSZ, XU = NS + NI;
  // two knots per warp pass (lanes 0..15 / 16..31) when the X and U rows of a knot fit in a half-warp, else one
  constexpr bool TWO = XU <= 16;
  static_assert(XU <= 32 && NS <= 16, "rows of a knot fit in a warp, multipliers in a half-warp");
  const int lane = threadIdx.x & 31;
  const long long sub = (long long)blockIdx.x * (blockDim.x >> 5) + warp_index_sync();
  if (sub >= nsub) return;
  const int prob = (int)(sub / M0), grp = (int)(sub - (long long)prob * M0);
  const double* la_p = lam_all + ((size_t)prob * (M0 + 1) + grp) * NS;
  const double* rec = sb.sep_in + ((size_t)prob * M0 + grp) * 3 * WSZ;
  const int half = lane >> 4, r16 = lane & 15;
  const int rc = r16 < NS ? r16 : 0;
  // multipliers, one entry per lane r16 < NS (both halves hold copies)
  const double la = la_p[rc], lb = la_p[NS + rc];
  auto sepmul = [&](const double* R, double left, double right) -> double {
    double v = R[2 * NS * NS + rc];
#pragma unroll
    for (int c = 0; c < NS; ++c) {
      const double lc = __shfl_sync(0xffffffffu, left, (lane & 16) + c);
      const double rcv = __shfl_sync(0xffffffffu, right, (lane & 16) + c);
      v = fma(-R[c * NS + rc], lc, fma(-R[(NS + c) * NS + rc], rcv, v));
    }
    return v;
  };
  const double l2 = sepmul(rec + WSZ, la, lb);
  // half 0 computes l1 (record 0: la | l2), half 1 computes l3 (record 2: l2 | lb)
  const double l13 = sepmul(rec + (half ? 2 * WSZ : 0), half ? l2 : la, half ? lb : l2);
  const double l1 = __shfl_sync(0xffffffffu, l13, rc), l3 = __shfl_sync(0xffffffffu, l13, 16 + rc);
  constexpr int PASSES = TWO ? 2 : 4;
#pragma unroll 1
  for (int pass = 0; pass < PASSES; ++pass) {
    const int j = TWO ? 2 * pass + half : pass;  // local knot of this (half-)warp
    const int r = TWO ? r16 : lane;              // row of [x; u] this lane produces
    const int base = TWO ? (lane & 16) : 0;      // where this knot's copy of the multipliers sits
    const int k = grp * SCAN_LEAF + j;
    const int kg = s.kofs + k;
    const double lk = j == 0 ? la : (j == 1 ? l1 : (j == 2 ? l2 : l3));
    const double ln = j == 0 ? l1 : (j == 1 ? l2 : (j == 2 ? l3 : lb));
    const double* kn = knot_ptr(s, b, prob, k);
    double* out = soln_ptr(s, b, prob, k);
    const bool last = kg == s.Ng - 1;
    double acc = 0.0;  // row r of [A|B]^T ln
    const double* col = kn + s.oA + (r < XU ? r : 0) * NS;
#pragma unroll
    for (int c = 0; c < NS; ++c) {
      const double lc = __shfl_sync(0xffffffffu, ln, base + c);
      acc = fma(last ? 0.0 : __ldg(col + c), lc, acc);
    }
    // (a lane with r >= NS holds a copy of multiplier entry r16 < NS or of entry 0: its own lk is only used for r < NS,
    //  where r == r16 in both layouts)
    if (r < XU) {
      const double d = __ldg(kn + r);
      const double rx = scan_rhs(s, b, prob, k, NS + r, pack_rhs);
      if (r < NS) {
        const double rl = kg == 0 ? scan_rhs(s, b, prob, k, r, pack_rhs) : 0.0;
        double xv, lv;
        if (kg == 0) { xv = -rl; lv = -d * rl - rx + acc; }
        else { xv = (rx + lk - acc) / d; lv = lk; }
        // all reads of the rhs (possibly aliasing soln) of this row are done before the writes below
        out[r] = lv;
        out[NS + r] = xv;
      } else {
        out[NS + r] = last ? rx : (rx - acc) / d;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------
struct ScanPlan {
  FusedPlan f;     // G0 = SCAN_LEAF; only sep / lamhat / ent[t >= 1] / xsep / xlam of f.fb are allocated
  ScanBufs sb;
  bool dual;       // entries are the Schur-complement elements of kernels_dual.cuh (any covered shape), else the boundary
                   // entries of this file (n = 12, m = 4)
  bool hybrid;     // n = 12, m = 4: scan_leaf_kernel (boundary rows, levels 0..1) hands dual elements up, the tiers run dual
  size_t ent;      // doubles per entry
};

inline bool dual_shape_ok(int n, int m);       // kernels_dual.cuh
inline size_t dual_ent_doubles(int n);
// RSLQR_DUAL=1 routes n = 12, m = 4 through the dual-form kernels too (A/B runs)
inline bool scan_classic_ok(int n, int m) {
  const char* v = getenv("RSLQR_DUAL");  // (read at plan time: a handle keeps the form it was created with)
  return n == 12 && m == 4 && !(v && atoi(v));
}
inline bool scan_shape_ok(int n, int m) { return scan_classic_ok(n, m) || dual_shape_ok(n, m); }

inline void scan_free(ScanPlan* p) {
  cudaFree(p->sb.ent0); cudaFree(p->sb.sep_in);
  p->f.fb.ent[0] = nullptr;  // alias of sb.ent0
  fused_free(&p->f);
  memset(p, 0, sizeof(*p));
}

// allow_hybrid: false for shards (their exchange and cross-GPU tier work on boundary entries)
inline int scan_plan(const Shape& s, int batch, ScanPlan* p, std::string& err, bool allow_hybrid = true) {
  memset(p, 0, sizeof(*p));
  if (!scan_shape_ok(s.n, s.m) || s.N < SCAN_LEAF) return 0;
  FusedPlan& f = p->f;
  p->dual = !scan_classic_ok(s.n, s.m);
  {
    const char* v = getenv("RSLQR_NO_HYBRID");  // A/B: boundary entries at every level, as in round 1
    p->hybrid = !p->dual && allow_hybrid && !(v && atoi(v));
  }
  p->ent = (p->dual || p->hybrid) ? dual_ent_doubles(s.n) : fused_ent_doubles(s.n, s.m);
  f.ns = s.n; f.ni = s.m; f.G0 = SCAN_LEAF; f.M0 = s.N / SCAN_LEAF; f.use_dmma = 1;
  // entries per tier CTA: 4 for batches (2 then 1 of 2 warps active over a CTA's two levels, measured +1.7 % over 8),
  // 8 for a single long horizon (fewer dependent launches); RSLQR_TIER_E = 2, 4 or 8 overrides. Falls back to 8 when a
  // smaller arity would need more than the 8 tier slots.
  const char* tv = getenv("RSLQR_TIER_E");
  int tier_e = tv ? atoi(tv) : (batch >= 64 ? 4 : 8);
  if (tier_e != 2 && tier_e != 4 && tier_e != 8) tier_e = 8;
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
    if (tier_e == 8) { memset(p, 0, sizeof(*p)); return 0; }
    tier_e = 8;
  }
  const size_t n = s.n, ent = p->ent, wsz = (2 * n + 1) * n;
  auto alloc = [&](double** ptr, size_t cnt) {
    f.bytes += cnt * sizeof(double);
    return cudaMalloc((void**)ptr, cnt * sizeof(double)) == cudaSuccess;
  };
  bool ok = alloc(&f.fb.sep, (size_t)batch * f.M0 * wsz) && alloc(&f.fb.lamhat, (size_t)batch * (f.M0 + 1) * n) &&
            alloc(&p->sb.sep_in, (size_t)batch * f.M0 * 3 * wsz) && alloc(&p->sb.ent0, (size_t)batch * f.M0 * ent) &&
            alloc(&f.xsep, 8 * wsz) && alloc(&f.xlam, 9 * n);
  f.fb.ent[0] = p->sb.ent0;
  for (int t = 1; ok && t < f.ntiers; ++t) ok = alloc(&f.fb.ent[t], (size_t)batch * f.tiers[t].Min * ent);
  if (!ok) {
    cudaGetLastError();
    scan_free(p);
    err = "device allocation for the scan path failed";
    return -4;
  }
  f.available = true;
  return 0;
}

template <int NS, int NI>
inline cudaError_t launch_scan_leaf(ScanPlan& p, const Shape& s, const Bufs& b, int batch, int pack_rhs, cudaStream_t st) {
  // RSLQR_LEAF_PAD_KB: extra (unused) shared memory per CTA, an occupancy knob for experiments
  static const int pad = [] { const char* v = getenv("RSLQR_LEAF_PAD_KB"); return v ? atoi(v) * 1024 : 0; }();
  const int smem = scan_leaf_smem_doubles<NS, NI>() * (int)sizeof(double) + pad;
  cudaError_t e = smem_optin(scan_leaf_kernel<NS, NI>, smem);
  if (e != cudaSuccess) return e;
  const long long nsub = (long long)batch * p.f.M0;
  return launch_k(scan_leaf_kernel<NS, NI>, dim3((unsigned)((nsub + SCAN_WPC - 1) / SCAN_WPC)), dim3(32 * SCAN_WPC), (size_t)smem, st, s, b,
                  p.sb, pack_rhs, p.f.M0, nsub, p.hybrid ? 1 : 0);
}

// top-down over the separators above the subtrees, then multipliers inside the subtrees + solution
template <int NS, int NI>
inline cudaError_t launch_scan_down(ScanPlan& p, const Shape& s, const Bufs& b, int batch, int pack_rhs, cudaStream_t st,
                                    const double* left, const double* right) {
  p.f.launches += launch_topdown<NS>(p.f.fb.sep, p.f.fb.lamhat, p.f.M0, batch, left, right, st);
  const long long nsub = (long long)batch * p.f.M0;
  launch_k(scan_down_kernel<NS, NI>, dim3((unsigned)((nsub + 3) / 4)), dim3(128), (size_t)0, st, s, b, p.sb, (const double*)p.f.fb.lamhat,
           pack_rhs, p.f.M0, nsub);
  ++p.f.launches;
  return cudaGetLastError();
}

inline int dual_launch(ScanPlan& p, const Shape& s, const Bufs& b, int batch, int pack_rhs, cudaStream_t st, std::string& err,
                       cudaEvent_t k0, cudaEvent_t k1);  // kernels_dual.cuh

inline int scan_launch(ScanPlan& p, const Shape& s, const Bufs& b, int batch, int pack_rhs, cudaStream_t st, std::string& err,
                       cudaEvent_t k0 = nullptr, cudaEvent_t k1 = nullptr) {
  if (!p.f.available) { err = "shape not covered by the scan kernels"; return -5; }
  if (p.dual || p.hybrid) return dual_launch(p, s, b, batch, pack_rhs, st, err, k0, k1);
  p.f.launches = 1;
  if (k0) cudaEventRecord(k0, st);  // k0..k1 bracket the dominant kernel (leaf) only
  cudaError_t e = launch_scan_leaf<12, 4>(p, s, b, batch, pack_rhs, st);
  if (k1) cudaEventRecord(k1, st);
  if (e == cudaSuccess) e = launch_tiers<12, 4>(p.f, s, b, batch, st, nullptr, 0);
  if (e == cudaSuccess) e = launch_scan_down<12, 4>(p, s, b, batch, pack_rhs, st, nullptr, nullptr);
  if (e != cudaSuccess) { err = std::string("scan launch: ") + cudaGetErrorString(e); return -3; }
  return 0;
}

// ---- shards of one long horizon ----
inline int scan_shard_reduce(ScanPlan& p, const Shape& s, const Bufs& b, int pack_rhs, cudaStream_t st, double* payload,
                             std::string& err) {
  using D = Dim<12, 4>;
  if (!p.f.available || p.f.ntiers < 1) { err = "shard needs the scan kernels and at least two subtrees"; return -5; }
  p.f.launches = 1;
  cudaError_t e = launch_scan_leaf<12, 4>(p, s, b, 1, pack_rhs, st);
  if (e == cudaSuccess) e = launch_tiers<12, 4>(p.f, s, b, 1, st, payload, (long long)fused_payload_doubles(12, 4));
  if (e == cudaSuccess)
    e = cudaMemcpyAsync(payload + D::ENT, b.knots + (size_t)(s.N - 1) * s.stride + s.oA, sizeof(double) * 12 * 16,
                        cudaMemcpyDeviceToDevice, st);
  if (e != cudaSuccess) { err = std::string("shard reduce: ") + cudaGetErrorString(e); return -3; }
  return 0;
}

inline int scan_shard_finish(ScanPlan& p, const Shape& s, const Bufs& b, int pack_rhs, cudaStream_t st, const double* gathered,
                             int rank, int nranks, std::string& err, const unsigned long long* wait_flags = nullptr,
                             unsigned long long wait_epoch = 0) {
  using D = Dim<12, 4>;
  if (!p.f.available || p.f.ntiers < 1) { err = "shard needs the scan kernels"; return -5; }
  const double *left = nullptr, *right = nullptr;
  cudaError_t e = cudaSuccess;
  if (nranks > 1) {
    e = tier_attr<12, 4>();
    if (e == cudaSuccess) {
      TierArgs a{};
      const long long pay = (long long)fused_payload_doubles(12, 4);
      a.wait_flags = wait_flags; a.wait_epoch = wait_epoch;
      a.in = gathered; a.in_stride = pay;
      a.ab = gathered + D::ENT; a.ab_stride = pay; a.ab_pstride = 0;
      a.sep = p.f.xsep; a.sep_pstride = 0; a.sep_mul = 1;
      a.out = nullptr; a.out_stride = 0;
      a.Min = nranks; a.ecount = nranks; a.span = 0;
      launch_tier<12, 4>(s, b, a, 1, st);
      launch_k(topdown_kernel<12>, dim3(1), dim3(128), (size_t)0, st, (const double*)p.f.xsep, p.f.xlam, nranks, (const double*)nullptr,
               (const double*)nullptr, nranks, 1);
      p.f.launches += 2;
      left = p.f.xlam + (size_t)rank * 12;
      right = p.f.xlam + (size_t)(rank + 1) * 12;
    }
  }
  if (e == cudaSuccess) e = launch_scan_down<12, 4>(p, s, b, 1, pack_rhs, st, left, right);
  if (e != cudaSuccess) { err = std::string("shard finish: ") + cudaGetErrorString(e); return -3; }
  return 0;
}

}  // namespace rslqr

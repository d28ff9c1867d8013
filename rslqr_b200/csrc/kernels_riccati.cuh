// kernels_riccati.cuh -- the Riccati recursion on the device, as the comparison solver.
//
// The reference ships a sequential Riccati solver beside rsLQR as its own baseline (src/riccati_solve.c:27-150:
// ndlqr_BackwardPass / ndlqr_ForwardPass). It solves the same KKT system, so its result is comparable entry by entry
// with ndlqr_Solve's [lambda_k, x_k, u_k] vector. On a GPU the recursion is serial in the horizon but a batch supplies
// the parallelism: ONE WARP PER PROBLEM, everything of a problem (P, p, scratch) in that warp's shared memory, no CTA
// barrier. It is not the product path (that is the nested-dissection solve); it is the throughput competitor for
// batches of short horizons and is reported next to it by bench.py.
//
// Backward pass, knot k = N-2 .. 0 (src/riccati_solve.c:41-118), with Abar = [A_k | B_k | d_k] (n x (n+m+1)):
//   M = P_{k+1} Abar                       (n x 17; column 16 = P d, + p_{k+1} -> v)              18 DMMA
//   H = [A|B]^T M  + diag(Q,R), + [q;r]    (16 x 17: [[Qxx Qux^T Qx],[Qux Quu Qu]])                18 DMMA
//   K = Quu^-1 Qux, dv = Quu^-1 Qu         (4 x 4 LDL^T, every lane redundantly; lane c solves column c)
//   P_k = Qxx - Qux^T K,  p_k = Qx - Qux^T dv                                                       6 DMMA
// (the reference's Joseph-form update src/riccati_solve.c:104-117 equals this in exact arithmetic; P is symmetrised
// when it is read back as an operand). Per knot [K | dv | P_k (lower triangle) | p_k] (142 doubles) goes to HBM for the forward pass.
// Forward pass (src/riccati_solve.c:121-150): lambda_k = P_k x_k + p_k, u_k = -K_k x_k - dv_k, x_{k+1} = A x + B u + d.
// The knot blocks stream through shared memory in PAIRS (one flat block is 1896 bytes, a pair is 16-byte aligned)
// by double-buffered TMA bulk copies.
#pragma once
#include "kernels_scan.cuh"

namespace rslqr {

constexpr int RIC_WPC = 4;  // warps (= problems) per CTA

template <int NS, int NI>
struct RicDim {
  static constexpr int XU = NS + NI, NC = XU + 1;
  static constexpr int LDM = NS + 1;                       // M column-major, odd leading dimension
  static constexpr int LDH = NC;                           // H row-major [XU][NC]
  static constexpr int PK = NS * (NS + 1) / 2;             // P is symmetric: its lower triangle, row by row
  static constexpr int GSZ = NI * NS + NI + PK + NS;       // [K (NI x NS row-major) | dv | P (packed) | p]
  static constexpr int PAIR = 2 * RawDim<NS, NI>::STRIDE;
  // per warp: [knot pair x2 | K dv Ppacked p (GSZ, contiguous: copied to HBM as one run) | P (full, operand) | M | H]
  static constexpr int OG = 2 * PAIR, OP = OG + GSZ, OM = OP + NS * NS, OH = OM + ((NC * LDM + 1) & ~1);
  static constexpr int WARP = OH + XU * LDH;
  static_assert(PAIR % 2 == 0 && GSZ % 2 == 0 && WARP % 2 == 0, "16-byte alignment of the regions");
};

template <int NS, int NI>
__global__ void __launch_bounds__(32 * RIC_WPC, 4) riccati_kernel(Shape s, Bufs b, double* gain, int nprob) {
  using RD = RawDim<NS, NI>;
  using D = RicDim<NS, NI>;
  constexpr int XU = D::XU, LDM = D::LDM, LDH = D::LDH, GSZ = D::GSZ, PAIR = D::PAIR, PK = D::PK;
  constexpr int KA = NS / 4;
  static_assert(NS == 12 && NI == 4, "tile shapes below are written for n = 12, m = 4");
  extern __shared__ __align__(16) double smf[];
  __shared__ __align__(8) unsigned long long mbar[RIC_WPC][2];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, q = lane & 3, rr = lane >> 2;
  const int prob = blockIdx.x * RIC_WPC + warp;
  if (prob >= nprob) return;
  const int N = s.N, NP = N / 2;
  double* W = smf + warp * D::WARP;
  double* KB = W;                 // two pair buffers
  double* Kk = W + D::OG;         // K (NI x NS), then dv, packed P, p contiguous
  double* dv = Kk + NI * NS;
  double* Pk = dv + NI;
  double* pv = Pk + PK;
  double* Ps = W + D::OP;         // full P: the operand of the next step
  double* Ms = W + D::OM;
  double* Hs = W + D::OH;
  const double* kn0 = knot_ptr(s, b, prob, 0);
  double* G = gain + (size_t)prob * N * GSZ;
  if (lane == 0) {
    mbar_init(&mbar[warp][0], 1);
    mbar_init(&mbar[warp][1], 1);
    fence_mbar_init();
  }
  __syncwarp();
  auto fetch_pair = [&](int j, int buf) {  // pair j = knots 2j, 2j+1
    if (lane == 0) {
      mbar_expect_tx(&mbar[warp][buf], PAIR * (int)sizeof(double));
      bulk_g2s(KB + buf * PAIR, kn0 + (size_t)j * PAIR, PAIR * (int)sizeof(double), &mbar[warp][buf]);
    }
  };
  auto store_gain = [&](int k) {  // [K | dv | P (packed) | p] of knot k -> HBM (one contiguous run); call after a __syncwarp
    double2* dst = reinterpret_cast<double2*>(G + (size_t)k * GSZ);
    const double2* src = reinterpret_cast<const double2*>(Kk);
    for (int e = lane; e < GSZ / 2; e += 32) dst[e] = src[e];
  };
  int bad = 0;
  unsigned phase[2] = {0u, 0u};
  // ------------------------------------------------------------------ backward pass
  fetch_pair(NP - 1, (NP - 1) & 1);
  for (int j = NP - 1; j >= 0; --j) {
    const int buf = j & 1;
    if (j > 0) {
      // the other buffer was last read (generic proxy) two iterations ago; order those reads before the async write
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      fetch_pair(j - 1, buf ^ 1);
    }
    mbar_wait(&mbar[warp][buf], phase[buf]);
    phase[buf] ^= 1u;
    for (int h2 = 1; h2 >= 0; --h2) {
      const int k = 2 * j + h2;
      const double* kb = KB + buf * PAIR + h2 * RD::STRIDE;
      if (k == N - 1) {  // terminal knot: P = diag(Q), p = q, no input (src/riccati_solve.c:31-38)
        for (int e = lane; e < NS * NS; e += 32) Ps[e] = (e / NS == e % NS) ? kb[e / NS] : 0.0;
        for (int e = lane; e < PK; e += 32) Pk[e] = 0.0;
        __syncwarp();
        if (lane < NS) Pk[lane * (lane + 1) / 2 + lane] = kb[lane];
        for (int e = lane; e < NI * NS + NI; e += 32) Kk[e] = 0.0;
        if (lane < NS) pv[lane] = kb[XU + lane];
        __syncwarp();
        store_gain(k);
        __syncwarp();
        continue;
      }
      const double* AB = kb + RD::OA;
      // ---- M = P [A | B | d]  (+ p on the last column) ----
      double ab[KA][2];  // fragments of [A | B]: B operand here, A operand of the next product (same elements)
      {
        double c[2][3][2];
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
          for (int nt = 0; nt < 3; ++nt) c[mt][nt][0] = c[mt][nt][1] = 0.0;
#pragma unroll
        for (int ks = 0; ks < KA; ++ks) {
          const int kk = 4 * ks + q;
          double a[2], bb[3];
#pragma unroll
          for (int mt = 0; mt < 2; ++mt) {
            const int row = 8 * mt + rr, rs = row < NS ? row : 0;
            const double v = 0.5 * (Ps[rs * NS + kk] + Ps[kk * NS + rs]);
            a[mt] = row < NS ? v : 0.0;
          }
          bb[0] = ab[ks][0] = AB[rr * NS + kk];
          bb[1] = ab[ks][1] = AB[(8 + rr) * NS + kk];
          bb[2] = rr == 0 ? kb[RD::OD + kk] : 0.0;
#pragma unroll
          for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int nt = 0; nt < 3; ++nt) dmma884(c[mt][nt][0], c[mt][nt][1], a[mt], bb[nt]);
        }
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {
          const int row = 8 * mt + rr;
          if (row < NS) {
#pragma unroll
            for (int nt = 0; nt < 2; ++nt)
#pragma unroll
              for (int h = 0; h < 2; ++h) Ms[(8 * nt + 2 * q + h) * LDM + row] = c[mt][nt][h];
            if (q == 0) Ms[XU * LDM + row] = c[mt][2][0] + pv[row];
          }
        }
      }
      __syncwarp();
      // ---- H = [A | B]^T M + diag(Q, R), last column + [q; r] ----
      double hc[2][3][2];
      {
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
          for (int nt = 0; nt < 3; ++nt) hc[mt][nt][0] = hc[mt][nt][1] = 0.0;
#pragma unroll
        for (int ks = 0; ks < KA; ++ks) {
          const int kk = 4 * ks + q;
          double a[2], bb[3];
          a[0] = ab[ks][0];
          a[1] = ab[ks][1];
          bb[0] = Ms[rr * LDM + kk];
          bb[1] = Ms[(8 + rr) * LDM + kk];
          bb[2] = rr == 0 ? Ms[XU * LDM + kk] : 0.0;
#pragma unroll
          for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int nt = 0; nt < 3; ++nt) dmma884(hc[mt][nt][0], hc[mt][nt][1], a[mt], bb[nt]);
        }
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {
          const int i = 8 * mt + rr;
          const double dg = kb[i];        // Q (i < NS) then R: the block starts with [Q | R | q | r]
          const double gr = kb[XU + i];   // q then r
#pragma unroll
          for (int nt = 0; nt < 2; ++nt)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              const int col = 8 * nt + 2 * q + h;
              if (col == i) hc[mt][nt][h] += dg;
              if (i >= NS) Hs[i * LDH + col] = hc[mt][nt][h];  // only the input rows [Qux | Quu] are read back
            }
          if (q == 0) {
            hc[mt][2][0] += gr;
            if (i >= NS) Hs[i * LDH + XU] = hc[mt][2][0];
          }
        }
      }
      __syncwarp();
      // ---- K = Quu^-1 Qux, dv = Quu^-1 Qu: 4 x 4 LDL^T in every lane, one right-hand side per lane ----
      {
        double Qm[NI][NI], di[NI], y[NI];
#pragma unroll
        for (int a2 = 0; a2 < NI; ++a2)
#pragma unroll
          for (int b2 = 0; b2 <= a2; ++b2) Qm[a2][b2] = Hs[(NS + a2) * LDH + NS + b2];
        const int col = lane < NS ? lane : XU;
#pragma unroll
        for (int l = 0; l < NI; ++l) y[l] = Hs[(NS + l) * LDH + col];
#pragma unroll
        for (int jj = 0; jj < NI; ++jj) {
          if (!(Qm[jj][jj] > 0.0)) bad = 1;
          di[jj] = fast_rcp(Qm[jj][jj]);
          double lcol[NI];
#pragma unroll
          for (int i = jj + 1; i < NI; ++i) lcol[i] = Qm[i][jj] * di[jj];
#pragma unroll
          for (int i = jj + 1; i < NI; ++i)
#pragma unroll
            for (int k2 = jj + 1; k2 <= i; ++k2) Qm[i][k2] = fma(-lcol[i], Qm[k2][jj], Qm[i][k2]);  // Qm[.][jj] still unscaled
#pragma unroll
          for (int i = jj + 1; i < NI; ++i) Qm[i][jj] = lcol[i];
        }
#pragma unroll
        for (int i = 1; i < NI; ++i)
#pragma unroll
          for (int jj = 0; jj < i; ++jj) y[i] = fma(-Qm[i][jj], y[jj], y[i]);
#pragma unroll
        for (int i = 0; i < NI; ++i) y[i] *= di[i];
#pragma unroll
        for (int i = NI - 2; i >= 0; --i)
#pragma unroll
          for (int jj = i + 1; jj < NI; ++jj) y[i] = fma(-Qm[jj][i], y[jj], y[i]);
        if (lane < NS) {
#pragma unroll
          for (int l = 0; l < NI; ++l) Kk[l * NS + lane] = y[l];
        } else if (lane == NS) {
#pragma unroll
          for (int l = 0; l < NI; ++l) dv[l] = y[l];
        }
      }
      __syncwarp();
      // ---- P = Qxx - Qux^T K, p = Qx - Qux^T dv (accumulated onto the H fragments still in registers) ----
      {
        double a[2], bb[3];
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {
          const int i = 8 * mt + rr;
          a[mt] = i < NS ? -Hs[(NS + q) * LDH + i] : 0.0;
        }
        bb[0] = Kk[q * NS + rr];
        bb[1] = (8 + rr < NS) ? Kk[q * NS + 8 + rr] : 0.0;
        bb[2] = rr == 0 ? dv[q] : 0.0;
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
          for (int nt = 0; nt < 3; ++nt) dmma884(hc[mt][nt][0], hc[mt][nt][1], a[mt], bb[nt]);
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {
          const int i = 8 * mt + rr;
          if (i < NS) {
#pragma unroll
            for (int nt = 0; nt < 2; ++nt)
#pragma unroll
              for (int h = 0; h < 2; ++h) {
                const int col = 8 * nt + 2 * q + h;
                if (col < NS) Ps[i * NS + col] = hc[mt][nt][h];
                if (col <= i) Pk[i * (i + 1) / 2 + col] = hc[mt][nt][h];  // the lower triangle is what the forward pass reads
              }
            if (q == 0) pv[i] = hc[mt][2][0];
          }
        }
      }
      __syncwarp();
      store_gain(k);
      __syncwarp();
    }
  }
  if (bad && lane == 0) b.flags[prob] = 1;
}

// ------------------------------------------------------------------ forward pass (its own kernel: no shared memory and
// a small register footprint, so most problems of a batch of 4096 are resident at once and the per-knot DRAM latency
// of the serial rollout is hidden by the other warps).
// Lane roles: lanes [0,NS) lambda rows (P(r,:), p_r), lanes [NS,NS+NI) input rows (K(l,:), dv_l), lanes [16,16+NS)
// next-state rows (A(r,:), B(r,:), d_r) -- x_k lives in lanes 16.. . The operands of the NEXT knot are fetched into
// registers while the current one is computed.
template <int NS, int NI>
__global__ void __launch_bounds__(32 * RIC_WPC, 5) riccati_forward_kernel(Shape s, Bufs b, const double* gain, int nprob) {
  using RD = RawDim<NS, NI>;
  constexpr int XU = NS + NI, GSZ = RicDim<NS, NI>::GSZ, ZS = 2 * NS + NI;
  static_assert(XU == 16 && NS <= 16, "lane roles");
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int prob = blockIdx.x * RIC_WPC + warp;
  if (prob >= nprob) return;
  const int N = s.N;
  const double* kn0 = knot_ptr(s, b, prob, 0);
  const double* G = gain + (size_t)prob * N * GSZ;
  double* out = soln_ptr(s, b, prob, 0);
  const bool isl = lane < NS, isu = lane >= NS && lane < XU, isx = lane >= XU && lane < XU + NS;
  const int r = isl ? lane : (isu ? lane - NS : (isx ? lane - XU : 0));
  // a lane's row: 12 doubles `row` (P / K / A), 4 doubles `rb` (B, x lanes only), one scalar `tail` (p_r / dv_l / d_r)
  struct Ops { double row[NS], rb[NI], tail; };
  auto load_step = [&](int k, Ops& o) {
    const double* g = G + (size_t)k * GSZ;
    const double* kb = kn0 + (size_t)k * RD::STRIDE;
    // row-major P and K rows are contiguous; A(r,:) is strided by NS in the column-major block
    // (P is packed: element (r, c) of the symmetric matrix sits at max(r,c)(max(r,c)+1)/2 + min(r,c))
    const double* base = isl ? g + NI * NS + NI : (isu ? g + r * NS : kb + RD::OA + r);
    const int st = isx ? NS : 1;
#pragma unroll
    for (int c = 0; c < NS; ++c) o.row[c] = base[isl ? (c <= r ? r * (r + 1) / 2 + c : c * (c + 1) / 2 + r) : c * st];
#pragma unroll
    for (int c = 0; c < NI; ++c) o.rb[c] = isx ? kb[RD::OA + (NS + c) * NS + r] : 0.0;
    o.tail = isl ? g[NI * NS + NI + RicDim<NS, NI>::PK + r] : (isu ? g[NI * NS + r] : kb[RD::OD + r]);
  };
  double xr = isx ? b.x0[(size_t)prob * NS + r] : 0.0;
  Ops cur, nxt;
  load_step(0, cur);
  for (int k = 0; k < N; ++k) {
    if (k + 1 < N) load_step(k + 1, nxt);  // in flight while this knot is computed
    double acc0 = cur.tail, acc1 = 0.0;    // P_r . x + p_r  |  K_l . x + dv_l  |  A_r . x + d_r
#pragma unroll
    for (int c = 0; c < NS; c += 2) {
      acc0 = fma(cur.row[c], shfl_d(xr, XU + c), acc0);
      acc1 = fma(cur.row[c + 1], shfl_d(xr, XU + c + 1), acc1);
    }
    double val = acc0 + acc1;              // lambda_r | -(u_l) | partial x_{k+1}
#pragma unroll
    for (int c = 0; c < NI; ++c) val = fma(cur.rb[c], -shfl_d(val, NS + c), val);  // (rb = 0 outside the x lanes)
    double* o = out + (size_t)k * ZS;
    if (isl) o[r] = val;
    else if (isu) o[2 * NS + r] = (k == N - 1) ? 0.0 : -val;
    else if (isx) o[NS + r] = xr;
    xr = val;
    cur = nxt;
  }
}

template <int NS, int NI>
inline int launch_riccati(const Shape& s, const Bufs& b, double* gain, int nprob, cudaStream_t st, std::string& err) {
  using D = RicDim<NS, NI>;
  const int smem = RIC_WPC * D::WARP * (int)sizeof(double);
  {
    cudaError_t e = smem_optin(riccati_kernel<NS, NI>, smem);
    if (e != cudaSuccess) { err = std::string("riccati: ") + cudaGetErrorString(e); return 1; }
  }
  riccati_kernel<NS, NI><<<(unsigned)((nprob + RIC_WPC - 1) / RIC_WPC), 32 * RIC_WPC, smem, st>>>(s, b, gain, nprob);
  riccati_forward_kernel<NS, NI><<<(unsigned)((nprob + RIC_WPC - 1) / RIC_WPC), 32 * RIC_WPC, 0, st>>>(s, b, gain, nprob);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { err = std::string("riccati: ") + cudaGetErrorString(e); return 1; }
  return 0;
}

}  // namespace rslqr

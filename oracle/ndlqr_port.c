/* oracle/ndlqr_port.c -- TEST INFRASTRUCTURE ONLY (see ndlqr_port.h).
 * CPU restatement of rsLQR's ndlqr_Solve path in plain C. Every function cites the reference
 * file:line it follows. Loops are dense over all upper levels exactly like the reference
 * (no structural-zero skipping), so intermediate `fact` contents are comparable block by block.
 */
#include "ndlqr_port.h"

#include <math.h>
#include <omp.h>
#include <stdlib.h>
#include <string.h>

/* ---------- dense primitives (src/linalg_custom.c) ---------- */

/* C(rc x cc) = alpha * op(A) * op(B) + beta * C, column-major; inner dim `kk`.
 * Same accumulation order as clap_MatrixMultiply (linalg_custom.c:20-43): C_ij is scaled by beta
 * first, then products are added for k ascending. lda/ldb are the stored leading dimensions. */
static void mm(int rc, int cc, int kk, double alpha, const double* A, int lda, int tA,
               const double* B, int ldb, int tB, double beta, double* C, int ldc) {
  for (int i = 0; i < rc; ++i) {
    for (int j = 0; j < cc; ++j) {
      double acc = C[i + ldc * j] * beta;
      for (int k = 0; k < kk; ++k) {
        double a = tA ? A[k + lda * i] : A[i + lda * k];
        double b = tB ? B[j + ldb * k] : B[k + ldb * j];
        acc += alpha * a * b;
      }
      C[i + ldc * j] = acc;
    }
  }
}

/* in-place lower Cholesky, left-looking (linalg_custom.c:88-111); -1 on a pivot <= 0 */
static int chol_lower(double* A, int n) {
  for (int j = 0; j < n; ++j) {
    for (int k = 0; k < j; ++k) {
      double ajk = A[j + n * k];
      for (int i = j; i < n; ++i) A[i + n * j] -= A[i + n * k] * ajk;
    }
    double d = A[j + n * j];
    if (d <= 0) return -1;
    d = sqrt(d);
    for (int i = j; i < n; ++i) A[i + n * j] /= d;
  }
  return 0;
}

/* b <- L^{-1} b (tr = 0) or L^{-T} b (tr = 1), b is n x w (linalg_custom.c:113-132) */
static void tri_sub(const double* L, int n, double* b, int w, int tr) {
  for (int s = 0; s < n; ++s) {
    int j = tr ? n - 1 - s : s;
    for (int c = 0; c < w; ++c) {
      double x = b[j + n * c] / L[j + n * j];
      b[j + n * c] = x;
      if (!tr) {
        for (int i = j + 1; i < n; ++i) b[i + n * c] -= L[i + n * j] * x;
      } else {
        for (int i = 0; i < j; ++i) b[i + n * c] -= L[j + n * i] * x;
      }
    }
  }
}

static void chol_solve(const double* L, int n, double* b, int w) { /* linalg_custom.c:134 */
  tri_sub(L, n, b, w, 0);
  tri_sub(L, n, b, w, 1);
}

/* ---------- tree closed forms (src/binary_tree.c) ---------- */

int port_index_level(int k) { /* binary_tree.c:71 -> trailing one bits of k */
  int l = 0;
  while (k & 1) { k >>= 1; ++l; }
  return l;
}
int port_index_from_leaf(int leaf, int level) { /* binary_tree.c:65 */
  return (1 << level) * (2 * leaf + 1) - 1;
}
int port_index_at_level(int k, int level) { /* binary_tree.c:89 */
  return (1 << level) * (2 * (k >> (level + 1)) + 1) - 1;
}
int port_should_calc_lambda(int index, int k, int level) { /* nested_dissection.c:173 */
  int a = (k >> (level + 1)) << (level + 1);
  int is_start = (k == a) || (k == index + 1);
  return !is_start || k == 0;
}

/* ---------- storage ---------- */

int port_knot_stride(int n, int m) { return 2 * n + 2 * m + 1 + n * n + n * m + n; }

static double* F(const PortSolver* s, double* base, int k, int level) { /* nddata.c:82-96 */
  return base + (size_t)(k + s->N * level) * s->fs;
}
static double* Z(const PortSolver* s, int k) { return s->soln + (size_t)k * s->zs; }

PortSolver* port_new(int n, int m, int N) {
  if (n <= 0 || m <= 0 || N < 2 || (N & (N - 1))) return NULL; /* nddata.c:17-21 */
  PortSolver* s = (PortSolver*)calloc(1, sizeof(PortSolver));
  s->n = n; s->m = m; s->N = N;
  s->K = 0; while ((1 << s->K) < N) ++s->K;
  s->fs = (2 * n + m) * n;
  s->zs = 2 * n + m;
  size_t nf = (size_t)N * s->K * s->fs;
  s->data = (double*)calloc(nf, sizeof(double));
  s->fact = (double*)calloc(nf, sizeof(double));
  s->soln = (double*)calloc((size_t)N * s->zs, sizeof(double));
  s->diag = (double*)calloc((size_t)N * (n * n + m * m), sizeof(double));
  s->chol_status = (int*)calloc(3 * (size_t)N, sizeof(int));
  return s;
}

void port_free(PortSolver* s) {
  if (!s) return;
  free(s->data); free(s->fact); free(s->soln); free(s->diag); free(s->chol_status); free(s);
}

void port_reset(PortSolver* s) { /* solver.c:98-106 */
  size_t nf = (size_t)s->N * s->K * s->fs;
  memset(s->data, 0, nf * sizeof(double));
  memset(s->fact, 0, nf * sizeof(double));
  memset(s->soln, 0, (size_t)s->N * s->zs * sizeof(double));
  memset(s->diag, 0, (size_t)s->N * (s->n * s->n + s->m * s->m) * sizeof(double));
  memset(s->chol_status, 0, 3 * (size_t)s->N * sizeof(int));
}

/* Fill data/soln/diag from the flat problem (solver.c:122-194). Does not touch `fact`. */
int port_init(PortSolver* s, const double* knots, const double* x0) {
  const int n = s->n, m = s->m, N = s->N, st = port_knot_stride(n, m);
  const int oQ = 0, oR = n, oq = n + m, orr = 2 * n + m, oA = 2 * n + 2 * m + 1;
  const int oB = oA + n * n, od = oB + n * m;
  const int dsz = n * n + m * m;
  memcpy(Z(s, 0), x0, n * sizeof(double));                       /* solver.c:138-139 */
  for (int k = 0; k < N; ++k) {
    const double* kn = knots + (size_t)k * st;
    double* Qk = s->diag + (size_t)k * dsz;
    double* Rk = Qk + n * n;
    memset(Qk, 0, dsz * sizeof(double));
    for (int i = 0; i < n; ++i) Qk[i + n * i] = kn[oQ + i];      /* solver.c:160-169, 181-185 */
    memcpy(Z(s, k) + n, kn + oq, n * sizeof(double));            /* q_k */
    if (k == N - 1) break;
    for (int i = 0; i < m; ++i) Rk[i + m * i] = kn[oR + i];
    memcpy(Z(s, k) + 2 * n, kn + orr, m * sizeof(double));       /* r_k */
    int lev = port_index_level(k);
    double* C = F(s, s->data, k, lev);                           /* solver.c:146-152 */
    for (int i = 0; i < n; ++i)
      for (int j = 0; j < n; ++j) C[n * n + i + n * j] = kn[oA + j + n * i];         /* X = A^T */
    for (int i = 0; i < m; ++i)
      for (int j = 0; j < n; ++j) C[2 * n * n + i + m * j] = kn[oB + j + n * i];     /* U = B^T */
    double* C2 = F(s, s->data, k + 1, lev);                      /* solver.c:172-176 */
    memset(C2 + n * n, 0, (size_t)(n * n + m * n) * sizeof(double));
    for (int i = 0; i < n; ++i) C2[n * n + i + n * i] = -1.0;
    memcpy(Z(s, k + 1), kn + od, n * sizeof(double));            /* lambda_{k+1} slot = d_k */
  }
  int nvars = s->zs * N - m;
  for (int i = 0; i < nvars; ++i) s->soln[i] *= -1;              /* solver.c:188-190 */
  return 0;
}

/* ---------- stages ---------- */

/* One knot of the leaf stage (nested_dissection.c:10-105). */
void port_solve_leaf(PortSolver* s, int k) {
  const int n = s->n, m = s->m, N = s->N, nn = n * n, dsz = nn + m * m;
  double* Q = s->diag + (size_t)k * dsz;
  double* R = Q + nn;
  double* z = Z(s, k);
  if (k == 0) {                                                   /* nested_dissection.c:24-59 */
    double* C = F(s, s->data, 0, 0);
    double* Fk = F(s, s->fact, 0, 0);
    for (int i = 0; i < nn; ++i) Fk[i] = -C[nn + i];              /* F.lambda = -A0^T */
    memset(Fk + nn, 0, nn * sizeof(double));                      /* F.state  = 0     */
    memcpy(Fk + 2 * nn, C + 2 * nn, (size_t)m * n * sizeof(double));
    s->chol_status[1] = chol_lower(R, m);
    chol_solve(R, m, Fk + 2 * nn, n);                             /* F.input = R \ B0^T */
    chol_solve(R, m, z + 2 * n, 1);                               /* zu = R \ zu        */
    double* tmp = C;                                              /* scratch: C.lambda (:48) */
    memcpy(tmp, z, n * sizeof(double));
    memcpy(z, z + n, n * sizeof(double));
    mm(n, 1, n, -1.0, Q, n, 0, tmp, n, 0, -1.0, z, n);            /* zy = -Q zy - zx */
    for (int i = 0; i < n; ++i) z[n + i] = -tmp[i];               /* zx = -zy(old)   */
    s->chol_status[0] = chol_lower(Q, n);
    return;
  }
  s->chol_status[2 * k] = chol_lower(Q, n);                       /* :63-65 */
  if (k < N - 1) {                                                /* :70-87 */
    int lev = port_index_level(k);
    double* C = F(s, s->data, k, lev);
    double* Fk = F(s, s->fact, k, lev);
    s->chol_status[2 * k + 1] = chol_lower(R, m);
    chol_solve(R, m, z + 2 * n, 1);
    memcpy(Fk + nn, C + nn, nn * sizeof(double));
    chol_solve(Q, n, Fk + nn, n);                                 /* Q \ A^T */
    memcpy(Fk + 2 * nn, C + 2 * nn, (size_t)m * n * sizeof(double));
    chol_solve(R, m, Fk + 2 * nn, n);                             /* R \ B^T */
  }
  chol_solve(Q, n, z + n, 1);                                     /* :89-90 */
  int plev = port_index_level(k - 1);                             /* :96-102 */
  double* C = F(s, s->data, k, plev);
  double* Fk = F(s, s->fact, k, plev);
  memcpy(Fk + nn, C + nn, nn * sizeof(double));
  chol_solve(Q, n, Fk + nn, n);                                   /* Q \ -I */
  memset(Fk + 2 * nn, 0, (size_t)m * n * sizeof(double));
}

void port_leaves(PortSolver* s) { /* solve.c:59-63 */
#pragma omp for
  for (int k = 0; k < s->N; ++k) port_solve_leaf(s, k);
}

/* S = F2.lambda <- C1x' F1x + C1u' F1u + C2x' F2x + C2u' F2u - S  (nested_dissection.c:114-134).
 * `fs_` is either s->fact (w = n) or s->soln (w = 1, fact_level must be 0). */
void port_inner_product(PortSolver* s, double* fs_, int w, int index, int data_level,
                        int fact_level) {
  const int n = s->n, m = s->m;
  const size_t fsz = (size_t)(2 * n + m) * w;
  double* C1 = F(s, s->data, index, data_level);
  double* C2 = F(s, s->data, index + 1, data_level);
  double* F1 = fs_ + (size_t)(index + s->N * fact_level) * fsz;
  double* F2 = fs_ + (size_t)(index + 1 + s->N * fact_level) * fsz;
  double* S = F2;
  mm(n, w, n, 1.0, C1 + n * n, n, 1, F1 + n * w, n, 0, -1.0, S, n);
  mm(n, w, m, 1.0, C1 + 2 * n * n, m, 1, F1 + 2 * n * w, m, 0, 1.0, S, n);
  mm(n, w, n, 1.0, C2 + n * n, n, 1, F2 + n * w, n, 0, 1.0, S, n);
  mm(n, w, m, 1.0, C2 + 2 * n * n, m, 1, F2 + 2 * n * w, m, 0, 1.0, S, n);
}

void port_level_products(PortSolver* s, int level) { /* solve.c:71-82 */
  int leaves = 1 << (s->K - level - 1), cur = s->K - level;
#pragma omp for
  for (int i = 0; i < leaves * cur; ++i) {
    int index = port_index_from_leaf(i / cur, level);
    port_inner_product(s, s->fact, s->n, index, level, level + i % cur);
  }
}

static int s_slot(const PortSolver* s, int leaf, int level) { /* cholesky_factors.c:65-82 */
  return 2 * s->N + (s->N - (s->N >> level)) + leaf;
}

void port_level_cholesky(PortSolver* s, int level) { /* solve.c:87-98 */
  int leaves = 1 << (s->K - level - 1);
#pragma omp for
  for (int leaf = 0; leaf < leaves; ++leaf) {
    int index = port_index_from_leaf(leaf, level);
    s->chol_status[s_slot(s, leaf, level)] = chol_lower(F(s, s->fact, index + 1, level), s->n);
  }
}

void port_level_solves(PortSolver* s, int level) { /* solve.c:103-115, nested_dissection.c:136 */
  int leaves = 1 << (s->K - level - 1), up = s->K - level - 1;
#pragma omp for
  for (int i = 0; i < leaves * up; ++i) {
    int index = port_index_from_leaf(i / up, level);
    int p = level + 1 + i % up;
    chol_solve(F(s, s->fact, index + 1, level), s->n, F(s, s->fact, index + 1, p), s->n);
  }
}

/* g(k,p) -= F(k,level) * f, f = separator block of column p (nested_dissection.c:154-171).
 * `dst` is s->fact (w = n, column p) or s->soln (w = 1, p ignored). */
static void schur_update(PortSolver* s, double* dst, int w, int index, int k, int level, int p,
                         int calc_lambda) {
  const int n = s->n, m = s->m;
  const size_t fsz = (size_t)(2 * n + m) * w;
  const double* f = dst + (size_t)(index + 1 + s->N * p) * fsz; /* .lambda */
  double* g = dst + (size_t)(k + s->N * p) * fsz;
  const double* Fk = F(s, s->fact, k, level);
  if (calc_lambda) mm(n, w, n, -1.0, Fk, n, 0, f, n, 0, 1.0, g, n);
  mm(n, w, n, -1.0, Fk + n * n, n, 0, f, n, 0, 1.0, g + n * w, n);
  mm(m, w, n, -1.0, Fk + 2 * n * n, m, 0, f, n, 0, 1.0, g + 2 * n * w, m);
}

void port_level_schur(PortSolver* s, int level) { /* solve.c:120-131 */
  int up = s->K - level - 1;
#pragma omp for
  for (int i = 0; i < s->N * up; ++i) {
    int k = i / up, p = level + 1 + i % up;
    int index = port_index_at_level(k, level);
    schur_update(s, s->fact, s->n, index, k, level, p, port_should_calc_lambda(index, k, level));
  }
}

void port_rhs_level(PortSolver* s, int level) { /* solve.c:137-182 */
  int leaves = 1 << (s->K - level - 1);
#pragma omp for
  for (int leaf = 0; leaf < leaves; ++leaf) {
    int index = port_index_from_leaf(leaf, level);
    port_inner_product(s, s->soln, 1, index, level, 0);
  }
#pragma omp for
  for (int leaf = 0; leaf < leaves; ++leaf) {
    int index = port_index_from_leaf(leaf, level);
    chol_solve(F(s, s->fact, index + 1, level), s->n, Z(s, index + 1), 1);
  }
#pragma omp for
  for (int k = 0; k < s->N; ++k) {
    int index = port_index_at_level(k, level);
    schur_update(s, s->soln, 1, index, k, level, 0, port_should_calc_lambda(index, k, level));
  }
}

int port_solve(PortSolver* s, int num_threads) { /* solve.c:38-190 */
  if (num_threads < 1) num_threads = 1;
#pragma omp parallel num_threads(num_threads)
  {
    port_leaves(s);
    for (int level = 0; level < s->K; ++level) {
      port_level_products(s, level);
      port_level_cholesky(s, level);
      port_level_solves(s, level);
      port_level_schur(s, level);
    }
    for (int level = 0; level < s->K; ++level) port_rhs_level(s, level);
  }
  int bad = 0;
  for (int i = 0; i < 3 * s->N - 1; ++i) bad |= s->chol_status[i];
  return bad ? -1 : 0;
}

/* ---------- independent checks ---------- */

/* || K z - b ||_2 for the KKT system of SURVEY App. A.1; z = [lam0,x0,u0,lam1,...,lamN-1,xN-1]. */
double port_kkt_residual(int n, int m, int N, const double* knots, const double* x0,
                         const double* z) {
  const int st = port_knot_stride(n, m), zs = 2 * n + m;
  const int oR = n, oq = n + m, orr = 2 * n + m, oA = 2 * n + 2 * m + 1;
  const int oB = oA + n * n, od = oB + n * m;
  double ss = 0.0;
  for (int i = 0; i < n; ++i) { /* lambda_0 row: -x_0 = -x_init */
    double r = -z[n + i] + x0[i];
    ss += r * r;
  }
  for (int k = 0; k < N; ++k) {
    const double* kn = knots + (size_t)k * st;
    const double* lam = z + (size_t)k * zs;
    const double* x = lam + n;
    const double* u = x + n;
    const double* lamn = lam + zs;
    int last = (k == N - 1);
    for (int i = 0; i < n; ++i) { /* x_k row */
      double r = kn[i] * x[i] - lam[i] + kn[oq + i];
      if (!last)
        for (int j = 0; j < n; ++j) r += kn[oA + j + n * i] * lamn[j];
      ss += r * r;
    }
    if (last) break;
    for (int i = 0; i < m; ++i) { /* u_k row */
      double r = kn[oR + i] * u[i] + kn[orr + i];
      for (int j = 0; j < n; ++j) r += kn[oB + j + n * i] * lamn[j];
      ss += r * r;
    }
    for (int i = 0; i < n; ++i) { /* lambda_{k+1} row */
      double r = -lamn[n + i] + kn[od + i];
      for (int j = 0; j < n; ++j) r += kn[oA + i + n * j] * x[j];
      for (int j = 0; j < m; ++j) r += kn[oB + i + n * j] * u[j];
      ss += r * r;
    }
  }
  return sqrt(ss);
}

/* Serial Riccati recursion (backward pass + rollout), same solution layout as port_solve.
 * Follows the structure of src/riccati_solve.c:26-150 (Qxx/Quu/Qux, gains by Cholesky of Quu). */
int port_riccati(int n, int m, int N, const double* knots, const double* x0, double* z) {
  const int st = port_knot_stride(n, m), zs = 2 * n + m;
  const int oR = n, oq = n + m, orr = 2 * n + m, oA = 2 * n + 2 * m + 1;
  const int oB = oA + n * n, od = oB + n * m;
  size_t per = (size_t)n * n + n + (size_t)m * n + m;
  double* buf = (double*)calloc(per * N + 4 * (size_t)n * n + 4 * (size_t)m * n + 8 * n, sizeof(double));
  if (!buf) return -1;
  double* P = buf;                       /* N x (n*n) */
  double* p = P + (size_t)N * n * n;     /* N x n     */
  double* Kg = p + (size_t)N * n;        /* N x (m*n) */
  double* dg = Kg + (size_t)N * m * n;   /* N x m     */
  double* PA = dg + (size_t)N * m;       /* n*n */
  double* PB = PA + n * n;               /* n*m */
  double* Quu = PB + n * m;              /* m*m (<= m*n) */
  double* Qux = Quu + m * n;             /* m*n */
  double* t = Qux + m * n;               /* n */
  int rc = 0;
  const double* kl = knots + (size_t)(N - 1) * st;
  double* PN = P + (size_t)(N - 1) * n * n;
  for (int i = 0; i < n; ++i) { PN[i + n * i] = kl[i]; p[(size_t)(N - 1) * n + i] = kl[oq + i]; }
  for (int k = N - 2; k >= 0; --k) {
    const double* kn = knots + (size_t)k * st;
    const double *A = kn + oA, *B = kn + oB, *d = kn + od;
    const double* Pn = P + (size_t)(k + 1) * n * n;
    const double* pn = p + (size_t)(k + 1) * n;
    double* Pk = P + (size_t)k * n * n;
    double* pk = p + (size_t)k * n;
    double* Kk = Kg + (size_t)k * m * n;
    double* dk = dg + (size_t)k * m;
    mm(n, n, n, 1.0, Pn, n, 0, A, n, 0, 0.0, PA, n);
    mm(n, m, n, 1.0, Pn, n, 0, B, n, 0, 0.0, PB, n);
    for (int i = 0; i < n; ++i) t[i] = pn[i];
    mm(n, 1, n, 1.0, Pn, n, 0, d, n, 0, 1.0, t, n);              /* t = P d + p */
    mm(n, n, n, 1.0, A, n, 1, PA, n, 0, 0.0, Pk, n);             /* Qxx */
    for (int i = 0; i < n; ++i) Pk[i + n * i] += kn[i];
    mm(m, m, n, 1.0, B, n, 1, PB, n, 0, 0.0, Quu, m);
    for (int i = 0; i < m; ++i) Quu[i + m * i] += kn[oR + i];
    mm(m, n, n, 1.0, B, n, 1, PA, n, 0, 0.0, Qux, m);
    for (int i = 0; i < n; ++i) pk[i] = kn[oq + i];
    mm(n, 1, n, 1.0, A, n, 1, t, n, 0, 1.0, pk, n);              /* Qx */
    for (int i = 0; i < m; ++i) dk[i] = kn[orr + i];
    mm(m, 1, n, 1.0, B, n, 1, t, n, 0, 1.0, dk, m);              /* Qu */
    if (chol_lower(Quu, m)) rc = -1;
    memcpy(Kk, Qux, (size_t)m * n * sizeof(double));
    chol_solve(Quu, m, Kk, n);                                   /* K = Quu \ Qux */
    chol_solve(Quu, m, dk, 1);                                   /* d = Quu \ Qu  */
    mm(n, n, m, -1.0, Qux, m, 1, Kk, m, 0, 1.0, Pk, n);          /* P = Qxx - Qux' K */
    mm(n, 1, m, -1.0, Qux, m, 1, dk, m, 0, 1.0, pk, n);          /* p = Qx - Qux' d  */
  }
  double* x = z + n;
  memcpy(x, x0, n * sizeof(double));
  for (int k = 0; k < N; ++k) {
    double* lam = z + (size_t)k * zs;
    double* xk = lam + n;
    double* uk = xk + n;
    const double* Pk = P + (size_t)k * n * n;
    for (int i = 0; i < n; ++i) lam[i] = p[(size_t)k * n + i];
    mm(n, 1, n, 1.0, Pk, n, 0, xk, n, 0, 1.0, lam, n);           /* lambda = P x + p */
    if (k == N - 1) break;
    const double* kn = knots + (size_t)k * st;
    for (int i = 0; i < m; ++i) uk[i] = -dg[(size_t)k * m + i];
    mm(m, 1, n, -1.0, Kg + (size_t)k * m * n, m, 0, xk, n, 0, 1.0, uk, m);
    double* xn = lam + zs + n;
    for (int i = 0; i < n; ++i) xn[i] = kn[od + i];
    mm(n, 1, n, 1.0, kn + oA, n, 0, xk, n, 0, 1.0, xn, n);
    mm(n, 1, m, 1.0, kn + oB, n, 0, uk, m, 0, 1.0, xn, n);
  }
  free(buf);
  return rc;
}

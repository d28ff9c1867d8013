/* oracle/ndlqr_port.h -- TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C restatement of the reference's nested-dissection LQR solve (rsLQR), used as the parity
 * checker for the CUDA path. Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library; the product (librslqr.so) never links or calls it.
 *
 * Parity pinning: this port is checked in tests/test_oracle.py against the reference's own golden
 * vectors (lqr_prob.json:soln, lqr_prob_256.json:soln, sample_problem.json b / E** / F*2*, the
 * hard-coded stage values of test/nested_dissection_test.c) and against oracle/_ref/libref_rslqr.so
 * (the unmodified reference sources compiled by oracle/Makefile).
 *
 * Storage follows the reference exactly (src/nddata.c:15-64): a "factor" is three column-major
 * blocks back to back [Lambda (n x w) | X (n x w) | U (m x w)], w = n for matrix data and 1 for
 * the right-hand side; factor (k, level) sits at offset (k + N*level) * (2n+m)*w.
 * A problem is passed flat: knots[k] = [Q(n) R(m) q(n) r(m) c A(n*n) B(n*m) d(n)] (the LQRData
 * block order of src/lqr_data.c:25-36) and x0[n].
 */
#ifndef NDLQR_PORT_H
#define NDLQR_PORT_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
  int n, m, N, K;     /* states, inputs, horizon (power of two), K = log2 N */
  int fs;             /* (2n+m)*n doubles per matrix factor */
  int zs;             /* 2n+m doubles per rhs factor */
  double* data;       /* [K][N][fs]  KKT column blocks C(k,p)            (src/solver.c:149-175) */
  double* fact;       /* [K][N][fs]  factor blocks F(k,p)                                      */
  double* soln;       /* [N][zs]     rhs -> solution                                           */
  double* diag;       /* [N][n*n+m*m] dense Q_k then R_k, overwritten by their Cholesky factors */
  int* chol_status;   /* [2N + N-1] 0 = ok, -1 = non-positive pivot (src/cholesky_factors.c:49-82) */
} PortSolver;

int port_knot_stride(int n, int m);                 /* doubles per flat knot block */
PortSolver* port_new(int n, int m, int N);          /* NULL unless N is a power of two >= 2 */
void port_free(PortSolver* s);
void port_reset(PortSolver* s);                     /* zero data/fact/soln/diag (src/solver.c:98) */
int port_init(PortSolver* s, const double* knots, const double* x0); /* src/solver.c:122 */

/* tree closed forms (src/binary_tree.c:65-106) */
int port_index_level(int k);
int port_index_from_leaf(int leaf, int level);
int port_index_at_level(int k, int level);
int port_should_calc_lambda(int index, int k, int level);

/* stages, each = the loop body of one phase of src/solve.c over all its tasks */
void port_solve_leaf(PortSolver* s, int k);                         /* nested_dissection.c:10  */
void port_leaves(PortSolver* s);                                    /* solve.c:59-63           */
void port_inner_product(PortSolver* s, double* fact_or_soln, int w, int index, int data_level,
                        int fact_level);                            /* nested_dissection.c:114 */
void port_level_products(PortSolver* s, int level);                 /* solve.c:71-82           */
void port_level_cholesky(PortSolver* s, int level);                 /* solve.c:87-98           */
void port_level_solves(PortSolver* s, int level);                   /* solve.c:103-115         */
void port_level_schur(PortSolver* s, int level);                    /* solve.c:120-131         */
void port_rhs_level(PortSolver* s, int level);                      /* solve.c:137-182         */
int port_solve(PortSolver* s, int num_threads);                     /* solve.c:38 (whole thing) */

/* independent check: || K z - b ||_2 of the KKT system of SURVEY App. A.1, from the flat problem */
double port_kkt_residual(int n, int m, int N, const double* knots, const double* x0,
                         const double* z);

/* serial Riccati recursion on the flat problem (timing/accuracy comparison; src/riccati_solve.c) */
int port_riccati(int n, int m, int N, const double* knots, const double* x0, double* z);

#ifdef __cplusplus
}
#endif
#endif

/* Consumer of the public C API in the order the reference's consumers use it (read problem -> new solver ->
 * initialize -> solve -> summary -> free; cf. reference examples/installexample/main.c:8-25), plus a check of
 * the result against the "soln" vector stored in the problem file. Exit code 0 = solved and matched. */
#include <math.h>
#include <stdio.h>
#include "ndlqr.h"
#include "json_utils.h"

int main(int argc, char** argv) {
  if (argc < 2) return 2;
  LQRProblem* prob = ndlqr_ReadLQRProblemJSONFile(argv[1]);
  if (!prob) return 3;
  const int n = prob->lqrdata[0]->nstates, m = prob->lqrdata[0]->ninputs, N = prob->nhorizon;
  NdLqrSolver* solver = ndlqr_NewNdLqrSolver(n, m, N);
  if (!solver) return 4;
  int rc = ndlqr_InitializeWithLQRProblem(prob, solver);
  if (rc == 0) rc = ndlqr_Solve(solver);
  if (rc != 0) return 5;
  ndlqr_PrintSolveSummary(solver);
  Matrix z = ndlqr_GetSolution(solver);
  Matrix want = ReadMatrixJSONFile(argv[1], "soln");
  if (!want.data || want.rows * want.cols != z.rows * z.cols) return 6;
  double err = 0;
  for (int i = 0; i < z.rows; ++i) err += (z.data[i] - want.data[i]) * (z.data[i] - want.data[i]);
  printf("nvars %d  ||z - soln||_2 = %.3e\n", z.rows, sqrt(err));
  FreeMatrix(&want);
  ndlqr_FreeLQRProblem(prob);
  ndlqr_FreeNdLqrSolver(solver);
  return sqrt(err) < 1e-6 ? 0 : 1;
}

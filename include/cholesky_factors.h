/* cholesky_factors.h -- index map of every Cholesky factorization of one solve
 * (reference src/cholesky_factors.h:30-35, src/cholesky_factors.c:49-82):
 *   slot 2k     -> Q_k,   slot 2k+1 -> R_k,
 *   slot 2N + (N - N/2^level) + leaf -> separator S(leaf, level).
 * The device keeps one int status word per slot in the same order. */
#ifndef RSLQR_CHOLESKY_FACTORS_H
#define RSLQR_CHOLESKY_FACTORS_H
#include "linalg.h"
#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
  int depth;
  int nhorizon;
  CholeskyInfo* cholinfo;
  int numfacts; /* 2N + N - 1 */
} NdLqrCholeskyFactors;

NdLqrCholeskyFactors* ndlqr_NewCholeskyFactors(int depth, int nhorizon);
int ndlqr_FreeCholeskyFactors(NdLqrCholeskyFactors* cholfacts);
int ndlqr_GetQFactorizon(NdLqrCholeskyFactors* cholfacts, int index, CholeskyInfo** cholfact);
int ndlqr_GetRFactorizon(NdLqrCholeskyFactors* cholfacts, int index, CholeskyInfo** cholfact);
int ndlqr_GetSFactorization(NdLqrCholeskyFactors* cholfacts, int leaf, int level,
                            CholeskyInfo** cholfact);

#ifdef __cplusplus
}
#endif
#endif

/* nddata.h -- host mirror of the factor storage (reference src/nddata.h:40-44,83-93).
 * A factor is three column-major blocks back to back, [lambda (n x w) | state (n x w) |
 * input (m x w)]; factor (index, level) starts at data + (index + N*level) * (2n+m)*w.
 * w = n for KKT data / factors (depth = log2 N columns), w = 1 for the right-hand side
 * (depth forced to 1). The device keeps `fact` and `soln` in this same order per problem so a
 * mirror is refreshed with a single copy. */
#ifndef RSLQR_NDDATA_H
#define RSLQR_NDDATA_H
#include "lqr_data.h"
#include "matrix.h"
#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
  Matrix lambda;
  Matrix state;
  Matrix input;
} NdFactor;

Matrix ndlqr_GetLambdaFactor(NdFactor* factor);
Matrix ndlqr_GetStateFactor(NdFactor* factor);
Matrix ndlqr_GetInputFactor(NdFactor* factor);

typedef struct {
  int nstates;
  int ninputs;
  int nsegments; /* N - 1 */
  int depth;
  int width;
  double* data;
  NdFactor* factors;
} NdData;

NdData* ndlqr_NewNdData(int nstates, int ninputs, int nhorizon, int width); /* NULL on bad args */
int ndlqr_FreeNdData(NdData* nddata);
int ndlqr_GetNdFactor(NdData* nddata, int index, int level, NdFactor** factor);
void ndlqr_ResetNdData(NdData* nddata);

#ifdef __cplusplus
}
#endif
#endif

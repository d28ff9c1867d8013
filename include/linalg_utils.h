/* linalg_utils.h -- stopwatch hooks of the helper API (reference src/linalg_utils.h:4-17). As in the reference's
 * default build the timing macros are disabled, so MatrixGetLinAlgTimeMilliseconds() is 0 unless a caller brackets
 * calls with Start/Stop itself. */
#ifndef RSLQR_LINALG_UTILS_H
#define RSLQR_LINALG_UTILS_H
#ifdef __cplusplus
extern "C" {
#endif
#define MATRIX_LATIME_START
#define MATRIX_LATIME_STOP
static const int kMatrixLinearAlgebraTimingEnabled = 0;
void MatrixLinAlgTimeStart(void);
void MatrixLinAlgTimeStop(void);
void MatrixLinAlgTimeReset(void);
double MatrixGetLinAlgTimeMilliseconds(void);
#ifdef __cplusplus
}
#endif
#endif

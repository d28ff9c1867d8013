/* matrix.h -- column-major, non-owning dense view used across the rsLQR public API.
 * ABI-compatible with the reference's `Matrix` (reference src/matrix.h:71-75): the struct is
 * returned by value from ndlqr_GetSolution, so field order and types are part of the contract.
 * Host-side helpers only; nothing here is on the device solve path. */
#ifndef RSLQR_MATRIX_H
#define RSLQR_MATRIX_H
#include <stdbool.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
  int rows;
  int cols;
  double* data; /* rows*cols doubles, element (i,j) at data[i + rows*j] */
} Matrix;

Matrix NewMatrix(int rows, int cols);          /* malloc'ed storage; release with FreeMatrix */
int FreeMatrix(Matrix* mat);
int MatrixSetConst(Matrix* mat, double val);
int MatrixNumElements(const Matrix* mat);
int MatrixGetLinearIndex(const Matrix* mat, int row, int col); /* -1 when out of range */
double* MatrixGetElement(const Matrix* mat, int row, int col);
double* MatrixGetElementTranspose(const Matrix* mat, int row, int col, bool istransposed);
int MatrixSetElement(Matrix* mat, int row, int col, double val);
int MatrixCopy(Matrix* dest, Matrix* src);            /* -1 unless same element count */
int MatrixCopyTranspose(Matrix* dest, Matrix* src);   /* dest = src^T */
int MatrixScaleByConst(Matrix* mat, double alpha);
double MatrixNormedDifference(Matrix* A, Matrix* B);  /* ||A-B||_2, INFINITY on size mismatch */
int MatrixFlatten(Matrix* mat);                       /* -> (rows*cols, 1) */
int MatrixFlattenToRow(Matrix* mat);                  /* -> (1, rows*cols) */
int PrintMatrix(const Matrix* mat);
int PrintRowVector(const Matrix* mat);

#ifdef __cplusplus
}
#endif
#endif

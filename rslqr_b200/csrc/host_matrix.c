/* host_matrix.c -- host-side Matrix view helpers (public API of include/matrix.h).
 * Behaviour follows the reference's src/matrix.c:19-160 (return codes, messages on stderr). */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "matrix.h"

#define RSLQR_PRINT_DIGITS 5

Matrix NewMatrix(int rows, int cols) {
  Matrix mat;
  mat.rows = rows;
  mat.cols = cols;
  mat.data = (double*)malloc(sizeof(double) * (size_t)rows * (size_t)cols);
  return mat;
}

int FreeMatrix(Matrix* mat) {
  if (!mat || !mat->data) return -1;
  free(mat->data);
  mat->data = NULL;
  return 0;
}

int MatrixNumElements(const Matrix* mat) { return mat ? mat->rows * mat->cols : -1; }

int MatrixSetConst(Matrix* mat, double val) {
  if (!mat) return -1;
  const int len = MatrixNumElements(mat);
  for (int e = 0; e < len; ++e) mat->data[e] = val;
  return 0;
}

int MatrixGetLinearIndex(const Matrix* mat, int row, int col) {
  if (!mat || row < 0 || col < 0) return -1;
  return col * mat->rows + row;
}

double* MatrixGetElement(const Matrix* mat, int row, int col) {
  return mat ? mat->data + MatrixGetLinearIndex(mat, row, col) : NULL;
}

double* MatrixGetElementTranspose(const Matrix* mat, int row, int col, bool istransposed) {
  return istransposed ? MatrixGetElement(mat, col, row) : MatrixGetElement(mat, row, col);
}

int MatrixSetElement(Matrix* mat, int row, int col, double val) {
  if (!mat) return -1;
  const int at = MatrixGetLinearIndex(mat, row, col);
  if (at < 0) return -1;
  mat->data[at] = val;
  return 1; /* the reference reports success as 1 here (src/matrix.c:62-71) */
}

int MatrixCopy(Matrix* dest, Matrix* src) {
  if (!dest || !src) return -1;
  if (dest->rows != src->rows || dest->cols != src->cols) {
    fprintf(stderr, "Can't copy matrices of different sizes.\n");
    return -1;
  }
  memcpy(dest->data, src->data, sizeof(double) * (size_t)MatrixNumElements(src));
  return 0;
}

int MatrixCopyTranspose(Matrix* dest, Matrix* src) {
  if (!dest || !src) return -1;
  if (dest->rows != src->cols || dest->cols != src->rows) {
    fprintf(stderr, "Matrix sizes are not transposes of each other. Got (%d,%d) and (%d,%d).\n",
            dest->rows, dest->cols, src->rows, src->cols);
    return -1;
  }
  for (int j = 0; j < dest->cols; ++j)
    for (int i = 0; i < dest->rows; ++i) dest->data[i + dest->rows * j] = src->data[j + src->rows * i];
  return 0;
}

int MatrixScaleByConst(Matrix* mat, double alpha) {
  if (!mat) return -1;
  const int len = MatrixNumElements(mat);
  for (int e = 0; e < len; ++e) mat->data[e] *= alpha;
  return 0;
}

double MatrixNormedDifference(Matrix* A, Matrix* B) {
  if (!A || !B) return INFINITY;
  if (A->rows != B->rows || A->cols != B->cols) {
    fprintf(stderr, "Can't compare matrices of different sizes. Got (%d,%d) and (%d,%d)\n",
            A->rows, A->cols, B->rows, B->cols);
    return INFINITY;
  }
  double acc = 0.0;
  const int len = MatrixNumElements(A);
  for (int e = 0; e < len; ++e) {
    const double delta = A->data[e] - B->data[e];
    acc += delta * delta;
  }
  return sqrt(acc);
}

int MatrixFlatten(Matrix* mat) {
  if (!mat) return -1;
  mat->rows = MatrixNumElements(mat);
  mat->cols = 1;
  return 0;
}

int MatrixFlattenToRow(Matrix* mat) {
  if (!mat) return -1;
  mat->cols = MatrixNumElements(mat);
  mat->rows = 1;
  return 0;
}

int PrintMatrix(const Matrix* mat) {
  if (!mat) return -1;
  for (int i = 0; i < mat->rows; ++i) {
    for (int j = 0; j < mat->cols; ++j)
      printf("% 6.*g ", RSLQR_PRINT_DIGITS, mat->data[i + mat->rows * j]);
    printf("\n");
  }
  return 0;
}

int PrintRowVector(const Matrix* mat) {
  if (!mat) return -1;
  printf("[ ");
  const int len = MatrixNumElements(mat);
  for (int e = 0; e < len; ++e) printf("% 6.*g ", RSLQR_PRINT_DIGITS, mat->data[e]);
  printf("]\n");
  return 0;
}

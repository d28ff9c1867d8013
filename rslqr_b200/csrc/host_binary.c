/* host_binary.c -- flat binary container for batches of LQR problems (include/rslqr_batch.h).
 *
 * The reference ships problems as JSON (src/json_utils.h:24-61), which is impractical for 4096 x 256 knots. The
 * binary file holds exactly what the device upload takes -- per problem x0 (n doubles) followed by N flat knot
 * blocks in the LQRData order Q,R,q,r,c,A,B,d (reference src/lqr_data.c:25-36) -- so a range of problems is read
 * with one fread straight into (pinned) upload buffers.
 *
 *   header (64 bytes): magic "RSLQRB01", int32 n, m, N, count, uint32 endian tag 0x01020304, zero padding
 *   record p         : double x0[n]; double knots[N][block]      (block = ndlqr_LQRDataBlockSize(n, m))
 * All integers and doubles in the byte order of the writing host; the endian tag rejects foreign files. */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "rslqr_batch.h"

#define RSLQR_BIN_MAGIC "RSLQRB01"
#define RSLQR_BIN_HEADER 64
#define RSLQR_BIN_ENDIAN 0x01020304u

typedef struct {
  char magic[8];
  int32_t n, m, N, count;
  uint32_t endian;
  char pad[RSLQR_BIN_HEADER - 8 - 4 * 4 - 4];
} BinHeader;

static size_t record_doubles(int n, int m, int N) { return (size_t)n + (size_t)N * (size_t)ndlqr_LQRDataBlockSize(n, m); }

static int read_header(FILE* f, const char* path, BinHeader* h) {
  if (fread(h, 1, RSLQR_BIN_HEADER, f) != RSLQR_BIN_HEADER || memcmp(h->magic, RSLQR_BIN_MAGIC, 8) != 0) {
    fprintf(stderr, "ERROR: %s is not an rsLQR binary problem file.\n", path);
    return -1;
  }
  if (h->endian != RSLQR_BIN_ENDIAN) {
    fprintf(stderr, "ERROR: %s was written with a different byte order.\n", path);
    return -1;
  }
  if (h->n <= 0 || h->m <= 0 || h->N <= 0 || h->count < 0) {
    fprintf(stderr, "ERROR: %s has a corrupt header (n=%d m=%d N=%d count=%d).\n", path, h->n, h->m, h->N, h->count);
    return -1;
  }
  return 0;
}

int rslqr_WriteProblemsBinary(const char* path, int n, int m, int N, int count, const double* knots, const double* x0) {
  if (!path || !knots || !x0 || n <= 0 || m <= 0 || N <= 0 || count < 0) return -1;
  FILE* f = fopen(path, "wb");
  if (!f) {
    fprintf(stderr, "ERROR: cannot open %s for writing.\n", path);
    return -1;
  }
  BinHeader h;
  memset(&h, 0, sizeof(h));
  memcpy(h.magic, RSLQR_BIN_MAGIC, 8);
  h.n = n; h.m = m; h.N = N; h.count = count; h.endian = RSLQR_BIN_ENDIAN;
  const size_t kper = (size_t)N * (size_t)ndlqr_LQRDataBlockSize(n, m);
  int ok = fwrite(&h, 1, RSLQR_BIN_HEADER, f) == RSLQR_BIN_HEADER;
  for (int p = 0; p < count && ok; ++p) {
    ok = fwrite(x0 + (size_t)p * n, sizeof(double), (size_t)n, f) == (size_t)n &&
         fwrite(knots + (size_t)p * kper, sizeof(double), kper, f) == kper;
  }
  if (fclose(f) != 0) ok = 0;
  if (!ok) fprintf(stderr, "ERROR: short write to %s.\n", path);
  return ok ? 0 : -1;
}

int rslqr_ReadProblemsBinaryHeader(const char* path, int* n, int* m, int* N, int* count) {
  if (!path) return -1;
  FILE* f = fopen(path, "rb");
  if (!f) {
    fprintf(stderr, "ERROR: cannot open %s.\n", path);
    return -1;
  }
  BinHeader h;
  const int rc = read_header(f, path, &h);
  fclose(f);
  if (rc) return -1;
  if (n) *n = h.n;
  if (m) *m = h.m;
  if (N) *N = h.N;
  if (count) *count = h.count;
  return 0;
}

int rslqr_ReadProblemsBinary(const char* path, int first, int count, double* knots, double* x0) {
  if (!path || !knots || !x0 || first < 0 || count < 0) return -1;
  FILE* f = fopen(path, "rb");
  if (!f) {
    fprintf(stderr, "ERROR: cannot open %s.\n", path);
    return -1;
  }
  BinHeader h;
  if (read_header(f, path, &h)) { fclose(f); return -1; }
  if ((long long)first + count > h.count) {
    fprintf(stderr, "ERROR: %s holds %d problems, [%d, %d) requested.\n", path, h.count, first, first + count);
    fclose(f);
    return -1;
  }
  const size_t rec = record_doubles(h.n, h.m, h.N), kper = rec - (size_t)h.n;
  int ok = fseek(f, (long)(RSLQR_BIN_HEADER + sizeof(double) * rec * (size_t)first), SEEK_SET) == 0;
  for (int p = 0; p < count && ok; ++p) {
    ok = fread(x0 + (size_t)p * h.n, sizeof(double), (size_t)h.n, f) == (size_t)h.n &&
         fread(knots + (size_t)p * kper, sizeof(double), kper, f) == kper;
  }
  fclose(f);
  if (!ok) fprintf(stderr, "ERROR: %s is truncated.\n", path);
  return ok ? 0 : -1;
}

int rslqr_WriteLQRProblemsBinary(const char* path, const LQRProblem* const* probs, int count) {
  if (!path || !probs || count <= 0 || !probs[0] || probs[0]->nhorizon <= 0) return -1;
  const int N = probs[0]->nhorizon, n = probs[0]->lqrdata[0]->nstates, m = probs[0]->lqrdata[0]->ninputs;
  const size_t blk = (size_t)ndlqr_LQRDataBlockSize(n, m), kper = (size_t)N * blk;
  double* knots = (double*)malloc(sizeof(double) * kper * (size_t)count);
  double* x0 = (double*)malloc(sizeof(double) * (size_t)n * (size_t)count);
  int rc = (knots && x0) ? 0 : -1;
  for (int p = 0; p < count && !rc; ++p) {
    const LQRProblem* pr = probs[p];
    if (!pr || pr->nhorizon != N) { rc = -1; break; }
    for (int k = 0; k < N; ++k) {
      const LQRData* kd = pr->lqrdata[k];
      if (kd->nstates != n || kd->ninputs != m) { rc = -1; break; }
      memcpy(knots + (size_t)p * kper + (size_t)k * blk, kd->Q, sizeof(double) * blk);  /* one block, starts at Q */
    }
    memcpy(x0 + (size_t)p * n, pr->x0, sizeof(double) * (size_t)n);
  }
  if (rc) fprintf(stderr, "ERROR: rslqr_WriteLQRProblemsBinary: problems of different shapes (or out of memory).\n");
  if (!rc) rc = rslqr_WriteProblemsBinary(path, n, m, N, count, knots, x0);
  free(knots);
  free(x0);
  return rc;
}

LQRProblem* rslqr_ReadLQRProblemBinary(const char* path, int index) {
  int n, m, N, count;
  if (rslqr_ReadProblemsBinaryHeader(path, &n, &m, &N, &count)) return NULL;
  const size_t blk = (size_t)ndlqr_LQRDataBlockSize(n, m);
  double* knots = (double*)malloc(sizeof(double) * blk * (size_t)N);
  double* x0 = (double*)malloc(sizeof(double) * (size_t)n);
  LQRProblem* prob = NULL;
  if (knots && x0 && rslqr_ReadProblemsBinary(path, index, 1, knots, x0) == 0) {
    prob = ndlqr_NewLQRProblem(n, m, N);
    if (prob) {
      for (int k = 0; k < N; ++k) memcpy(prob->lqrdata[k]->Q, knots + (size_t)k * blk, sizeof(double) * blk);
      memcpy(prob->x0, x0, sizeof(double) * (size_t)n);
    }
  }
  free(knots);
  free(x0);
  return prob;
}

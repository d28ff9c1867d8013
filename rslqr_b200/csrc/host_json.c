/* host_json.c -- JSON problem loaders (include/json_utils.h) on a self-contained reader.
 *
 * The reference parses with cJSON 1.7.15 (reference src/json_utils.c); that library is not part
 * of this build, so the file format is re-read by a small recursive-descent parser into a node
 * tree, and the loaders walk the tree. File format, field names, 1-based `index`, arrays-of-columns
 * matrices and NULL / {0,0,NULL} + stderr error behaviour follow src/json_utils.c:186-348. */
#include <ctype.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "json_utils.h"
#include "rslqr_host_internal.h"

/* ---------------- node tree ---------------- */
typedef enum { JN_NULL, JN_BOOL, JN_NUM, JN_STR, JN_ARR, JN_OBJ } JKind;

typedef struct JNode {
  JKind kind;
  double num;          /* JN_NUM / JN_BOOL */
  char* key;           /* member name when the parent is an object */
  char* str;           /* JN_STR */
  struct JNode* first; /* children of arrays / objects */
  struct JNode* next;  /* sibling */
  int count;           /* number of children */
} JNode;

typedef struct {
  const char* at;
  const char* end;
  const char* fail; /* position of the first error */
} JCursor;

static void jfree(JNode* nd) {
  while (nd) {
    JNode* nx = nd->next;
    jfree(nd->first);
    free(nd->key);
    free(nd->str);
    free(nd);
    nd = nx;
  }
}

static void skip_ws(JCursor* c) {
  while (c->at < c->end && isspace((unsigned char)*c->at)) ++c->at;
}

static JNode* jnew(JKind kind) {
  JNode* nd = (JNode*)calloc(1, sizeof(JNode));
  if (nd) nd->kind = kind;
  return nd;
}

static char* parse_string_raw(JCursor* c) {
  if (c->at >= c->end || *c->at != '"') { c->fail = c->at; return NULL; }
  const char* from = ++c->at;
  size_t cap = 16, len = 0;
  char* out = (char*)malloc(cap);
  while (c->at < c->end && *c->at != '"') {
    char ch = *c->at++;
    if (ch == '\\' && c->at < c->end) {
      char e = *c->at++;
      switch (e) {
        case 'n': ch = '\n'; break;
        case 't': ch = '\t'; break;
        case 'r': ch = '\r'; break;
        case 'b': ch = '\b'; break;
        case 'f': ch = '\f'; break;
        case 'u': ch = '?'; c->at += (c->end - c->at >= 4) ? 4 : 0; break; /* names here are ASCII */
        default: ch = e; break;
      }
    }
    if (len + 2 > cap) { cap *= 2; out = (char*)realloc(out, cap); }
    out[len++] = ch;
  }
  if (c->at >= c->end) { c->fail = from; free(out); return NULL; }
  ++c->at; /* closing quote */
  out[len] = '\0';
  return out;
}

static JNode* parse_value(JCursor* c, int depth);

static JNode* parse_container(JCursor* c, int depth, int is_obj) {
  const char close = is_obj ? '}' : ']';
  JNode* box = jnew(is_obj ? JN_OBJ : JN_ARR);
  JNode* tail = NULL;
  ++c->at; /* opening bracket */
  skip_ws(c);
  if (c->at < c->end && *c->at == close) { ++c->at; return box; }
  for (;;) {
    char* key = NULL;
    skip_ws(c);
    if (is_obj) {
      key = parse_string_raw(c);
      if (!key) break;
      skip_ws(c);
      if (c->at >= c->end || *c->at != ':') { c->fail = c->at; free(key); break; }
      ++c->at;
    }
    JNode* child = parse_value(c, depth + 1);
    if (!child) { free(key); break; }
    child->key = key;
    if (tail) tail->next = child; else box->first = child;
    tail = child;
    ++box->count;
    skip_ws(c);
    if (c->at < c->end && *c->at == ',') { ++c->at; continue; }
    if (c->at < c->end && *c->at == close) { ++c->at; return box; }
    c->fail = c->at;
    break;
  }
  if (!c->fail) c->fail = c->at;
  jfree(box);
  return NULL;
}

static JNode* parse_value(JCursor* c, int depth) {
  if (depth > 64) { c->fail = c->at; return NULL; }
  skip_ws(c);
  if (c->at >= c->end) { c->fail = c->at; return NULL; }
  const char ch = *c->at;
  if (ch == '{') return parse_container(c, depth, 1);
  if (ch == '[') return parse_container(c, depth, 0);
  if (ch == '"') {
    char* s = parse_string_raw(c);
    if (!s) return NULL;
    JNode* nd = jnew(JN_STR);
    nd->str = s;
    return nd;
  }
  if ((size_t)(c->end - c->at) >= 4 && !strncmp(c->at, "true", 4)) {
    c->at += 4; JNode* nd = jnew(JN_BOOL); nd->num = 1; return nd;
  }
  if ((size_t)(c->end - c->at) >= 5 && !strncmp(c->at, "false", 5)) {
    c->at += 5; return jnew(JN_BOOL);
  }
  if ((size_t)(c->end - c->at) >= 4 && !strncmp(c->at, "null", 4)) {
    c->at += 4; return jnew(JN_NULL);
  }
  char* stop = NULL;
  const double v = strtod(c->at, &stop); /* buffer is NUL-terminated by rslqr_read_file */
  if (stop == c->at) { c->fail = c->at; return NULL; }
  c->at = stop;
  JNode* nd = jnew(JN_NUM);
  nd->num = v;
  return nd;
}

static JNode* jparse_file(const char* filename, const char* what) {
  char* text = NULL;
  int len = 0;
  if (ReadFile(filename, &text, &len) != 0) {
    fprintf(stderr, "ERROR: Reading %s file failed.\n", what);
    return NULL;
  }
  JCursor cur = {text, text + len, NULL};
  JNode* root = parse_value(&cur, 0);
  if (!root) {
    fprintf(stderr, "ERROR: Error parsing JSON file: %.40s\n", cur.fail ? cur.fail : "");
  }
  free(text);
  return root;
}

static const JNode* jget(const JNode* obj, const char* name) {
  if (!obj || obj->kind != JN_OBJ) return NULL;
  for (const JNode* it = obj->first; it; it = it->next)
    if (it->key && !strcmp(it->key, name)) return it;
  return NULL;
}

static int jint(const JNode* obj, const char* name, int fallback) {
  const JNode* it = jget(obj, name);
  return (it && it->kind == JN_NUM) ? (int)it->num : fallback;
}

/* ---------------- typed readers ---------------- */

static int read_vector(const JNode* obj, const char* name, double* out, int len) {
  const JNode* arr = jget(obj, name);
  if (!arr || arr->kind != JN_ARR) {
    fprintf(stderr, "Couldn't find an array of name %s.\n", name);
    return -1;
  }
  int i = 0;
  for (const JNode* it = arr->first; it; it = it->next, ++i)
    if (it->kind == JN_NUM && i < len) out[i] = it->num;
  if (i == len) return 0;
  fprintf(stderr, "JSON array was %s than expected (%d instead of %d).\n",
          i > len ? "longer" : "shorter", i, len);
  return -1;
}

static int matrix_shape(const JNode* obj, const char* name, int* rows, int* cols) {
  const JNode* arr = jget(obj, name);
  if (!arr || arr->kind != JN_ARR) {
    fprintf(stderr, "ERROR: Unable to parse the field %s as a JSON array.\n", name);
    return -1;
  }
  int j = 0;
  for (const JNode* col = arr->first; col; col = col->next) {
    if (col->kind != JN_ARR) continue;
    if (j == 0) {
      *rows = col->count;
    } else if (*rows != col->count) {
      fprintf(stderr, "ERROR: The number of rows changed during parsing. Failed to read as a 2D array.\n");
      return -1;
    }
    ++j;
  }
  *cols = j;
  return 0;
}

static int read_matrix(const JNode* obj, const char* name, double* out, int rows, int cols) {
  const JNode* arr = jget(obj, name);
  if (!arr || arr->kind != JN_ARR) {
    fprintf(stderr, "Couldn't find an array of name %s.\n", name);
    return -1;
  }
  int status = 0, j = 0;
  for (const JNode* col = arr->first; col; col = col->next) {
    if (col->kind != JN_ARR || j >= cols) continue;
    int i = 0;
    for (const JNode* it = col->first; it; it = it->next, ++i)
      if (it->kind == JN_NUM && i < rows) out[i + j * rows] = it->num;
    if (i != rows) {
      status = -1;
      fprintf(stderr, "Got unexpected length of JSON column number %d (%d instead of %d).\n", j, i, rows);
    }
    ++j;
  }
  if (j != cols) {
    status = -1;
    fprintf(stderr, "Got an unexpected number of JSON columns (%d instead of %d).\n", j, cols);
  }
  return status;
}

static int fill_knot(const JNode* obj, LQRData* knot) {
  const int n = jint(obj, "nstates", 0), m = jint(obj, "ninputs", 0);
  if (n != knot->nstates || m != knot->ninputs) {
    fprintf(stderr, "ERROR: The state and input dimensions in the JSON file didn't match the expected dimensions\n");
    return -1;
  }
  const JNode* c = jget(obj, "c");
  if (!c || c->kind != JN_NUM) return -1;
  knot->c[0] = c->num;
  int bad = 0;
  bad += read_vector(obj, "Q", knot->Q, n);
  bad += read_vector(obj, "R", knot->R, m);
  bad += read_vector(obj, "q", knot->q, n);
  bad += read_vector(obj, "r", knot->r, m);
  bad += read_vector(obj, "d", knot->d, n);
  bad += read_matrix(obj, "A", knot->A, n, n);
  bad += read_matrix(obj, "B", knot->B, n, m);
  if (bad) {
    fprintf(stderr, "ERROR: The LQR data file wasn't successfully parsed. Returning an empty struct.\n");
    return -1;
  }
  return 0;
}

/* ---------------- public loaders ---------------- */

LQRProblem* ndlqr_ReadLQRProblemJSONFile(const char* filename) {
  JNode* root = jparse_file(filename, "LQR Problem");
  if (!root) return NULL;
  const int nhorizon = jint(root, "nhorizon", 0);
  const JNode* knots = jget(root, "lqrdata");
  const JNode* first = (knots && knots->kind == JN_ARR) ? knots->first : NULL;
  const int n = jint(first, "nstates", 0), m = jint(first, "ninputs", 0);
  LQRProblem* prob = ndlqr_NewLQRProblem(n, m, nhorizon);
  if (!prob) { jfree(root); return NULL; }
  for (const JNode* it = first; it; it = it->next) {
    const int slot = jint(it, "index", 0) - 1; /* file is 1-based (src/json_utils.c:237) */
    if (slot < 0 || slot >= nhorizon || fill_knot(it, prob->lqrdata[slot]) != 0)
      fprintf(stderr, "WARNING: Failed to parse the LQR JSON data at index %d\n", slot);
  }
  const int bad = read_vector(root, "x0", prob->x0, n);
  jfree(root);
  if (bad) {
    ndlqr_FreeLQRProblem(prob);
    return NULL;
  }
  return prob;
}

LQRData* ndlqr_ReadLQRDataJSONFile(const char* filename) {
  JNode* root = jparse_file(filename, "LQR");
  if (!root) return NULL;
  const int n = jint(root, "nstates", 0), m = jint(root, "ninputs", 0);
  if (n <= 0 || m <= 0) {
    fprintf(stderr, "ERROR: Couldn't get a valid state and control dimension from the LQR Data file: %s\n", filename);
    jfree(root);
    return NULL;
  }
  LQRData* knot = ndlqr_NewLQRData(n, m);
  const int bad = fill_knot(root, knot);
  jfree(root);
  if (bad) {
    fprintf(stderr, "ERROR: Error parsing LQR JSON data.\n");
    ndlqr_FreeLQRData(knot);
    return NULL;
  }
  return knot;
}

Matrix ReadMatrixJSONFile(const char* filename, const char* name) {
  Matrix none = {0, 0, NULL};
  JNode* root = jparse_file(filename, "LQR");
  if (!root) return none;
  int rows = 0, cols = 0;
  if (matrix_shape(root, name, &rows, &cols) != 0) {
    fprintf(stderr, "ERROR: Couldn't get the size of the JSON Matrix.\n");
    jfree(root);
    return none;
  }
  Matrix mat = NewMatrix(rows, cols);
  if (read_matrix(root, name, mat.data, rows, cols) != 0) {
    fprintf(stderr, "ERROR: Wasn't able to parse the JSON Matrix.\n");
    FreeMatrix(&mat);
    jfree(root);
    return none;
  }
  jfree(root);
  return mat;
}

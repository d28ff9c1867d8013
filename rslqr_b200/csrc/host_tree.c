/* host_tree.c -- host-side index structures and mirrors: integer helpers, the ordered binary tree,
 * NdData mirrors, the Cholesky bookkeeping table (include/utils.h, binary_tree.h, nddata.h,
 * cholesky_factors.h, linalg.h). All of it is plain bookkeeping; the tree is laid out from the
 * closed forms in tree_math.h instead of the reference's recursive construction
 * (reference src/binary_tree.c:7-63) and produces the same node fields. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "rslqr_cuda.h"
#include "rslqr_host_internal.h"
#include "tree_math.h"

/* ---- utils.h ---- */
bool IsPowerOfTwo(int x) { return x > 0 && (x & (x - 1)) == 0; }

int LogOfTwo(int x) {
  if (x == 0) return 0;
  int pos = 0;
  while (!((x >> pos) & 1)) ++pos;
  return pos;
}

int ReadFile(const char* filename, char** out, int* len) {
  FILE* fp = fopen(filename, "rb");
  if (!fp) {
    fprintf(stderr, "Couldn't open file\n");
    return -1;
  }
  fseek(fp, 0L, SEEK_END);
  const long size = ftell(fp);
  rewind(fp);
  char* text = (char*)malloc((size_t)size + 1);
  if (!text) {
    fprintf(stderr, "Couldn't allocate memory for the file contents.");
    fclose(fp);
    return -1;
  }
  if (size > 0 && fread(text, (size_t)size, 1, fp) != 1) {
    fprintf(stderr, "Failed to read the entire file.");
    free(text);
    fclose(fp);
    return -1;
  }
  fclose(fp);
  text[size] = '\0';
  *out = text;
  *len = (int)size;
  return 0;
}

/* ---- binary_tree.h ---- */
OrderedBinaryTree ndlqr_BuildTree(int N) {
  OrderedBinaryTree tree = {NULL, NULL, 0, 0};
  if (!IsPowerOfTwo(N)) return tree;
  BinaryNode* nodes = (BinaryNode*)calloc((size_t)N, sizeof(BinaryNode));
  if (!nodes) return tree;
  const int depth = rs_log2(N);
  for (int k = 0; k < N; ++k) nodes[k].idx = k;
  for (int k = 0; k + 1 < N; ++k) { /* node k = the separator right after knot k */
    BinaryNode* nd = nodes + k;
    const int lv = rs_knot_level(k), half = 1 << lv;
    nd->level = lv;
    nd->levelidx = k >> (lv + 1);
    nd->left_inds.start = k - half + 1;
    nd->left_inds.stop = k;
    nd->right_inds.start = k + 1;
    nd->right_inds.stop = k + half;
    nd->left_child = lv ? nodes + (k - half / 2) : NULL;
    nd->right_child = lv ? nodes + (k + half / 2) : NULL;
    nd->parent = (lv + 1 < depth) ? nodes + rs_sep_of_knot(k, lv + 1) : NULL;
  }
  tree.root = N > 1 ? nodes + (N / 2 - 1) : nodes;
  tree.node_list = nodes;
  tree.num_elements = N;
  tree.depth = depth;
  return tree;
}

int ndlqr_FreeTree(OrderedBinaryTree* tree) {
  if (!tree) return -1;
  free(tree->node_list);
  tree->node_list = NULL;
  tree->root = NULL;
  return 0;
}

int ndlqr_GetIndexFromLeaf(const OrderedBinaryTree* tree, int leaf, int level) {
  (void)tree;
  return rs_sep_index(leaf, level);
}

int ndlqr_GetIndexLevel(const OrderedBinaryTree* tree, int index) {
  return tree->node_list[index].level;
}

int ndlqr_GetIndexAtLevel(const OrderedBinaryTree* tree, int index, int level) {
  if (!tree) return -1;
  if (index < 0 || index >= tree->num_elements)
    fprintf(stderr, "ERROR: Invalid index (%d). Should be between %d and %d.\n", index, 0,
            tree->num_elements - 1);
  if (level < 0 || level >= tree->depth)
    fprintf(stderr, "ERROR: Invalid level (%d). Should be between %d and %d.\n", level, 0,
            tree->depth - 1);
  return rs_sep_of_knot(index, level);
}

/* ---- nddata.h ---- */
Matrix ndlqr_GetLambdaFactor(NdFactor* factor) { return factor->lambda; }
Matrix ndlqr_GetStateFactor(NdFactor* factor) { return factor->state; }
Matrix ndlqr_GetInputFactor(NdFactor* factor) { return factor->input; }

NdData* ndlqr_NewNdData(int nstates, int ninputs, int nhorizon, int width) {
  if (nstates <= 0 || ninputs <= 0 || nhorizon - 1 <= 0) return NULL;
  if (!IsPowerOfTwo(nhorizon)) {
    fprintf(stderr, "ERROR: Number of segments must be one less than a power of 2.\n");
    return NULL;
  }
  const int depth = (width == 1) ? 1 : LogOfTwo(nhorizon); /* the rhs has a single column */
  const size_t count = (size_t)nhorizon * (size_t)depth;
  const size_t each = (size_t)(2 * nstates + ninputs) * (size_t)width;
  NdData* nd = (NdData*)malloc(sizeof(NdData));
  double* store = (double*)calloc(count * each, sizeof(double));
  NdFactor* views = (NdFactor*)malloc(count * sizeof(NdFactor));
  if (!nd || !store || !views) {
    fprintf(stderr, "ERROR: Failed to allocate memory for NdData.\n");
    free(nd);
    free(store);
    free(views);
    return NULL;
  }
  for (size_t f = 0; f < count; ++f) {
    double* at = store + f * each;
    const Matrix lam = {nstates, width, at};
    const Matrix sta = {nstates, width, at + (size_t)nstates * width};
    const Matrix inp = {ninputs, width, at + 2 * (size_t)nstates * width};
    views[f].lambda = lam;
    views[f].state = sta;
    views[f].input = inp;
  }
  nd->nstates = nstates;
  nd->ninputs = ninputs;
  nd->nsegments = nhorizon - 1;
  nd->depth = depth;
  nd->width = width;
  nd->data = store;
  nd->factors = views;
  return nd;
}

void ndlqr_ResetNdData(NdData* nddata) {
  const size_t count = (size_t)(nddata->nsegments + 1) * (size_t)nddata->depth;
  const size_t each = (size_t)(2 * nddata->nstates + nddata->ninputs) * (size_t)nddata->width;
  memset(nddata->data, 0, count * each * sizeof(double));
  /* ndlqr_ResetNdData(solver->fact) is the reference's way to prepare a re-solve (src/solve.h:28-32,
   * test/nested_dissection_test.c:301): the device copy of that array is the one the stage calls work on */
  int which = -1;
  NdLqrSolver* owner = rslqr_owner_of(nddata, &which);
  if (owner && which == 1 && owner->device) rslqr_cuda_zero_fact((rslqr_cuda_solver*)owner->device);
}

int ndlqr_FreeNdData(NdData* nddata) {
  if (!nddata) return -1;
  free(nddata->data);
  free(nddata->factors);
  free(nddata);
  return 0;
}

int ndlqr_GetNdFactor(NdData* nddata, int index, int level, NdFactor** factor) {
  if (index < 0 || index > nddata->nsegments) {
    fprintf(stderr, "Invalid index. Must be between %d and %d, got %d.\n", 0, nddata->nsegments, index);
    return -1;
  }
  if (level < 0 || level >= nddata->depth) {
    fprintf(stderr, "Invalid level. Must be between %d and %d, got %d.\n", 0, nddata->depth - 1, level);
    return -1;
  }
  *factor = nddata->factors + (index + (nddata->nsegments + 1) * level);
  return 0;
}

/* ---- linalg.h / cholesky_factors.h ---- */
CholeskyInfo DefaultCholeskyInfo(void) {
  CholeskyInfo info = {'\0', 1, '\0', NULL, 1};
  return info;
}
void FreeFactorization(CholeskyInfo* cholinfo) {
  if (cholinfo) {
    cholinfo->fact = NULL;
    cholinfo->is_freed = 1;
  }
}
enum MatrixLinearAlgebraLibrary MatrixGetLinearAlgebraLibrary(void) { return libCUDA; }
void MatrixPrintLinearAlgebraLibrary(void) { printf("Using CUDA sm_100a kernels for linear algebra\n"); }

NdLqrCholeskyFactors* ndlqr_NewCholeskyFactors(int depth, int nhorizon) {
  if (depth <= 0 || nhorizon <= 0) return NULL;
  NdLqrCholeskyFactors* table = (NdLqrCholeskyFactors*)malloc(sizeof(NdLqrCholeskyFactors));
  if (!table) return NULL;
  int separators = 0;
  for (int lv = 0; lv < depth; ++lv) separators += 1 << (depth - lv - 1);
  table->numfacts = 2 * nhorizon + separators;
  table->cholinfo = (CholeskyInfo*)malloc(sizeof(CholeskyInfo) * (size_t)table->numfacts);
  if (!table->cholinfo) {
    free(table);
    return NULL;
  }
  for (int f = 0; f < table->numfacts; ++f) table->cholinfo[f] = DefaultCholeskyInfo();
  table->depth = depth;
  table->nhorizon = nhorizon;
  return table;
}

int ndlqr_FreeCholeskyFactors(NdLqrCholeskyFactors* cholfacts) {
  if (!cholfacts) return -1;
  free(cholfacts->cholinfo);
  free(cholfacts);
  return 0;
}

int ndlqr_GetQFactorizon(NdLqrCholeskyFactors* cholfacts, int index, CholeskyInfo** cholfact) {
  if (!cholfacts || index < 0 || index >= cholfacts->nhorizon) return -1;
  *cholfact = cholfacts->cholinfo + rs_slot_Q(index);
  return 0;
}

int ndlqr_GetRFactorizon(NdLqrCholeskyFactors* cholfacts, int index, CholeskyInfo** cholfact) {
  if (!cholfacts || index < 0 || index >= cholfacts->nhorizon - 1) return -1;
  *cholfact = cholfacts->cholinfo + rs_slot_R(index);
  return 0;
}

int ndlqr_GetSFactorization(NdLqrCholeskyFactors* cholfacts, int leaf, int level, CholeskyInfo** cholfact) {
  if (!cholfacts || level < 0 || level >= cholfacts->depth) return -1;
  if (leaf < 0 || leaf >= (1 << (cholfacts->depth - level - 1))) return -1;
  *cholfact = cholfacts->cholinfo + rs_slot_S(leaf, level, cholfacts->nhorizon);
  return 0;
}

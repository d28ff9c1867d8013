/* tree_math.h -- closed forms that replace the reference's in-memory binary tree
 * (reference src/binary_tree.c:65-106; src/nested_dissection.c:173-177; src/cholesky_factors.c:65-82).
 * Usable from C, C++ and CUDA device code. N = 2^K knots, levels 0..K-1. */
#ifndef RSLQR_TREE_MATH_H
#define RSLQR_TREE_MATH_H

#if defined(__CUDACC__)
#define RSLQR_HD __host__ __device__ __forceinline__
#else
#define RSLQR_HD static inline
#endif

/* level of the separator that sits right after knot k = number of trailing one bits of k
 * (ndlqr_GetIndexLevel, valid for k <= N-2) */
RSLQR_HD int rs_knot_level(int k) {
  int lv = 0;
  while (k & 1) { k >>= 1; ++lv; }
  return lv;
}
/* knot index of separator number `leaf` on `level` (ndlqr_GetIndexFromLeaf) */
RSLQR_HD int rs_sep_index(int leaf, int level) { return ((2 * leaf + 1) << level) - 1; }
/* which level-`level` separator owns knot k (ndlqr_GetIndexAtLevel) */
RSLQR_HD int rs_sep_of_knot(int k, int level) { return rs_sep_index(k >> (level + 1), level); }
/* first / last knot of the level-`level` subtree that contains k */
RSLQR_HD int rs_subtree_first(int k, int level) { return (k >> (level + 1)) << (level + 1); }
RSLQR_HD int rs_subtree_last(int k, int level) { return rs_subtree_first(k, level) + (2 << level) - 1; }
/* ndlqr_ShouldCalcLambda: the lambda rows of knot k take part in a level-`level` Schur update
 * unless k opens one of the two halves of its subtree; knot 0 always takes part */
RSLQR_HD int rs_calc_lambda(int k, int level) {
  const int a = rs_subtree_first(k, level);
  return (k == 0) || !(k == a || k == a + (1 << level));
}
/* The only upper columns with structurally non-zero content for the level-`level` subtree
 * [a,b]: the levels of the separators bounding it. -1 = absent. */
RSLQR_HD int rs_left_col(int a) { return a > 0 ? rs_knot_level(a - 1) : -1; }
RSLQR_HD int rs_right_col(int b, int N) { return b < N - 1 ? rs_knot_level(b) : -1; }
/* status-word slots (NdLqrCholeskyFactors order) */
RSLQR_HD int rs_slot_Q(int k) { return 2 * k; }
RSLQR_HD int rs_slot_R(int k) { return 2 * k + 1; }
RSLQR_HD int rs_slot_S(int leaf, int level, int N) { return 2 * N + (N - (N >> level)) + leaf; }
RSLQR_HD int rs_log2(int N) {
  int K = 0;
  while ((1 << K) < N) ++K;
  return K;
}
#endif

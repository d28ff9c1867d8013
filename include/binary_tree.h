/* binary_tree.h -- the ordered binary tree over knot indices that names the separators of the
 * nested dissection (reference src/binary_tree.h:23-69). On the device none of this exists: the
 * kernels use the closed forms (level(k) = trailing one-bits of k, separator(leaf,level) =
 * 2^level (2 leaf + 1) - 1, ...; see rslqr_b200/csrc/tree_math.h). The host struct is still
 * built because callers read `solver->tree` and the node fields directly. */
#ifndef RSLQR_BINARY_TREE_H
#define RSLQR_BINARY_TREE_H
#include <stdbool.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
  int start; /* first knot index */
  int stop;  /* last knot index (node ranges are inclusive, as the reference fills them) */
} UnitRange;

typedef struct BinaryNode_s BinaryNode;
struct BinaryNode_s {
  int idx;
  int level;
  int levelidx;
  UnitRange left_inds;
  UnitRange right_inds;
  BinaryNode* parent;
  BinaryNode* left_child;
  BinaryNode* right_child;
};

typedef struct {
  BinaryNode* root;
  BinaryNode* node_list; /* indexed by knot */
  int num_elements;
  int depth;             /* log2(num_elements) */
} OrderedBinaryTree;

OrderedBinaryTree ndlqr_BuildTree(int N);
int ndlqr_FreeTree(OrderedBinaryTree* tree);
int ndlqr_GetIndexFromLeaf(const OrderedBinaryTree* tree, int leaf, int level);
int ndlqr_GetIndexLevel(const OrderedBinaryTree* tree, int index);
int ndlqr_GetIndexAtLevel(const OrderedBinaryTree* tree, int index, int level);

#ifdef __cplusplus
}
#endif
#endif

/* matmul.h -- placeholder kept so that `#include "ndlqr.h"` consumers compile unchanged.
 * The reference header (src/matmul.h) declares AVX2 experiments that are not compiled into its
 * library; their role (small fixed-size fp64 block products) is played here by the device
 * functions in rslqr_b200/csrc/. Nothing is declared. */
#ifndef RSLQR_MATMUL_H
#define RSLQR_MATMUL_H
#endif

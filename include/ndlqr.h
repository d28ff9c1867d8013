/* ndlqr.h -- umbrella header: the file users include (reference src/ndlqr.h:11-14). */
#ifndef RSLQR_NDLQR_H
#define RSLQR_NDLQR_H
#include "lqr_problem.h"
#include "matmul.h"
#include "solve.h"
#include "solver.h"
#endif

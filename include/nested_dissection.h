/* nested_dissection.h -- the individual stages of the solve, callable one item at a time on a
 * solver (reference src/nested_dissection.h; src/nested_dissection.c:10-192).
 * Each call runs the corresponding sm_100a stage kernel restricted to the requested item on the
 * solver's device state and refreshes the host mirrors; they exist for stage-level parity tests. */
#ifndef RSLQR_NESTED_DISSECTION_H
#define RSLQR_NESTED_DISSECTION_H
#include <stdlib.h>
#include "solver.h"
#ifdef __cplusplus
extern "C" {
#endif

int ndlqr_SolveLeaf(NdLqrSolver* solver, int index);
int ndlqr_SolveLeaves(NdLqrSolver* solver);
bool ndlqr_ShouldCalcLambda(OrderedBinaryTree* tree, int index, int i);
int ndlqr_ComputeShurCompliment(NdLqrSolver* solver, int index, int level, int upper_level);

/* Reference signatures. `data` / `fact` / `soln` must be the NdData mirrors of a live solver (the
 * owning solver is found from the pointer); passing solver->soln selects the right-hand side. */
int ndlqr_FactorInnerProduct(NdData* data, NdData* fact, int index, int data_level, int fact_level);
int ndlqr_SolveCholeskyFactor(NdData* fact, CholeskyInfo* cholinfo, int index, int level,
                              int upper_level);
int ndlqr_UpdateShurFactor(NdData* fact, NdData* soln, int index, int i, int level,
                           int upper_level, bool calc_lambda);

/* Stage calls on the solver (the reference versions take bare NdData pointers; here the data
 * lives on the device, so the solver is the handle). fact_level/upper_level 0 with use_rhs != 0
 * addresses the right-hand side instead of a factor column. */
int ndlqr_SolverFactorInnerProduct(NdLqrSolver* solver, int index, int data_level, int fact_level, int use_rhs);
int ndlqr_SolverFactorCholesky(NdLqrSolver* solver, int leaf, int level);
int ndlqr_SolverSolveCholeskyFactor(NdLqrSolver* solver, int index, int level, int upper_level, int use_rhs);
int ndlqr_SolverUpdateShurFactor(NdLqrSolver* solver, int index, int i, int level, int upper_level,
                                 bool calc_lambda, int use_rhs);

#ifdef __cplusplus
}
#endif
#endif

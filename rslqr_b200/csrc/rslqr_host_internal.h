/* rslqr_host_internal.h -- declarations shared by the host C files (not installed). */
#ifndef RSLQR_HOST_INTERNAL_H
#define RSLQR_HOST_INTERNAL_H
#include "solver.h"
#include "utils.h"
/* find the live solver that owns an NdData mirror; *which = 0 data, 1 fact, 2 soln */
NdLqrSolver* rslqr_owner_of(const NdData* nd, int* which);
#endif

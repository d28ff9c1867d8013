/* json_utils.h -- JSON problem loaders with the reference's file format and error behaviour
 * (reference src/json_utils.h:24-76, src/json_utils.c:186-348). The reference parses with cJSON;
 * this build carries its own small recursive-descent reader (rslqr_b200/csrc/host_json.c), so no
 * third-party dependency. 2-D arrays are arrays of COLUMNS; `index` inside lqrdata[] is 1-based. */
#ifndef RSLQR_JSON_UTILS_H
#define RSLQR_JSON_UTILS_H
#include "lqr_problem.h"
#include "matrix.h"
#ifdef __cplusplus
extern "C" {
#endif

LQRData* ndlqr_ReadLQRDataJSONFile(const char* filename);       /* NULL + stderr on failure */
LQRProblem* ndlqr_ReadLQRProblemJSONFile(const char* filename); /* NULL + stderr on failure */
Matrix ReadMatrixJSONFile(const char* filename, const char* name); /* {0,0,NULL} on failure */

#ifdef __cplusplus
}
#endif
#endif

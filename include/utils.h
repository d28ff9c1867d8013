/* utils.h -- small integer/file helpers kept for source compatibility (reference src/utils.h). */
#ifndef RSLQR_UTILS_H
#define RSLQR_UTILS_H
#include <stdbool.h>
#ifdef __cplusplus
extern "C" {
#endif
bool IsPowerOfTwo(int x);
static inline int PowerOfTwo(int x) { return 1 << x; }
int LogOfTwo(int x);                                       /* position of the lowest set bit */
int ReadFile(const char* filename, char** out, int* len);  /* malloc'ed, NUL-terminated */
#ifdef __cplusplus
}
#endif
#endif

/* Test infrastructure only (see oracle/README.md).
 * sdsl's construct_sa.hpp includes "divsufsort.h", which upstream generates with cmake from
 * divsufsort.h.cmake.  The finimizer path never calls into libdivsufsort, so declarations are
 * enough to let the reference headers parse when compiled directly with g++. */
#pragma once
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
typedef uint8_t sauchar_t;
typedef int32_t saint_t;
typedef int32_t saidx_t;
saint_t divsufsort(const sauchar_t *T, saidx_t *SA, saidx_t n);
#ifdef __cplusplus
}
#endif

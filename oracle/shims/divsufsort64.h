/* Test infrastructure only: 64-bit twin of shims/divsufsort.h (declarations only, never called). */
#pragma once
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
typedef uint8_t sauchar_t;
typedef int32_t saint_t;
typedef int64_t saidx64_t;
saint_t divsufsort64(const sauchar_t *T, saidx64_t *SA, saidx64_t n);
#ifdef __cplusplus
}
#endif

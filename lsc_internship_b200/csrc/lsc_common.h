// lsc_common.h -- helpers shared by the translation units that implement the C ABI.
#pragma once
#include <cuda_runtime.h>

#include "../../include/lsc_b200.h"

void lsc_count_launches(unsigned long long n);
int lsc_fail(int code, const char* fmt, ...);  // records the message lsc_last_error() returns

#define LSC_CUDA_TRY(expr)                                                                                   \
    do {                                                                                                     \
        cudaError_t e_ = (expr);                                                                             \
        if (e_ != cudaSuccess)                                                                               \
            return lsc_fail(LSC_ECUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

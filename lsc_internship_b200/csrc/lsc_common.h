// lsc_common.h -- helpers shared by the translation units that implement the C ABI.
#pragma once
#include <cuda_runtime.h>

#include "../../include/lsc_b200.h"

void lsc_count_launches(unsigned long long n);
int lsc_fail(int code, const char* fmt, ...);  // records the message lsc_last_error() returns

// stream for the halo exchange that overlaps a sweep: highest priority, so that the few NCCL / ghost-column blocks
// are scheduled ahead of the queued sweep blocks
static inline cudaError_t lsc_create_comm_stream(cudaStream_t* s) {
    int lo = 0, hi = 0;  // numerically lowest = highest priority
    cudaError_t e = cudaDeviceGetStreamPriorityRange(&lo, &hi);
    if (e != cudaSuccess) return e;
    return cudaStreamCreateWithPriority(s, cudaStreamNonBlocking, hi);
}

#define LSC_CUDA_TRY(expr)                                                                                   \
    do {                                                                                                     \
        cudaError_t e_ = (expr);                                                                             \
        if (e_ != cudaSuccess)                                                                               \
            return lsc_fail(LSC_ECUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

// lsc_common.h -- helpers shared by the translation units that implement the C ABI.
#pragma once
#include <cuda_runtime.h>

#include "../../include/lsc_b200.h"

void lsc_count_launches(unsigned long long n);
void lsc_uncount_launches(unsigned long long n);
extern "C" unsigned long long lsc_kernel_launch_count(void);
int lsc_fail(int code, const char* fmt, ...);  // records the message lsc_last_error() returns
// one __constant__ slot of step scalars per live context and device (kind 0: MHD, 1: Euler); -1: none free
int lsc_slot_acquire(int device, int kind);
void lsc_slot_release(int device, int kind, int slot);

// device buffers of a context: cudaMalloc / cudaFree, or -- with LSC_GUARD=1 -- the same with 64 KiB guard zones on
// both sides that lsc_guard_check() inspects (dev_guard.cu)
cudaError_t lsc_dev_malloc(void** p, size_t bytes);
cudaError_t lsc_dev_free(void* p);
template <class T>
static inline cudaError_t lsc_dev_malloc(T** p, size_t bytes) {
    return lsc_dev_malloc(reinterpret_cast<void**>(p), bytes);
}

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

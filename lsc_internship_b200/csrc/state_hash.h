// state_hash.h -- device-side, decomposition-independent fingerprint of a state's interior (state_hash.cu)
#pragma once
#include <cuda_runtime.h>

#include "mhd_kernels.cuh"

namespace lsc {
// adds this slab's contribution to *out (a device word the caller zeroed); i0_global = first global interior column
void launch_state_hash(const GridDims& g, int nvar, const double* u, long long i0_global, unsigned long long* out,
                       cudaStream_t s);

// The same fingerprint with NO allocation anywhere on the call path (state_hash.cu: one partial word per thread
// block in a file-scope __device__ array, summed by a second kernel, copied to *host_out on stream s, which is
// synchronised before returning).  The array is sized for STATE_HASH_STATIC_MAX_NY rows per slab; a taller slab is
// refused with cudaErrorInvalidValue.  `device` must be the current device.
constexpr int STATE_HASH_STATIC_PARTIALS = 8192;  // 64 KiB of partial sums (+ 256 B for the result word) per GPU
constexpr int STATE_HASH_STATIC_MAX_NY = STATE_HASH_STATIC_PARTIALS * 128;  // one block per 128 rows at least
cudaError_t state_hash_static(int device, const GridDims& g, int nvar, const double* u, long long i0_global,
                              unsigned long long* host_out, cudaStream_t s);
}  // namespace lsc

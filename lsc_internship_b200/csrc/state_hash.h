// state_hash.h -- device-side, decomposition-independent fingerprint of a state's interior (state_hash.cu)
#pragma once
#include <cuda_runtime.h>

#include "mhd_kernels.cuh"

namespace lsc {
// adds this slab's contribution to *out (a device word the caller zeroed); i0_global = first global interior column
void launch_state_hash(const GridDims& g, int nvar, const double* u, long long i0_global, unsigned long long* out,
                       cudaStream_t s);
}  // namespace lsc

// state_hash.cu -- decomposition-independent 64-bit fingerprint of the interior of a state, computed on the device.
//
// Every interior value contributes mix(bit pattern + golden * (global cell number * nvar + k + 1)) and the
// contributions are summed modulo 2^64: the sum does not depend on how the columns are distributed over slabs /
// ranks (add the per-rank words modulo 2^64), while a changed bit or two swapped cells change it.  Used by
// bench.py (`sim.state_hash`) so that the N = 1/2/4/8 lines of a strong-scaling run can be compared by anyone, and
// by the tests against the same function evaluated with numpy on the downloaded array.
#include "state_hash.h"

#include "mhd_kernels.cuh"

namespace lsc {

__device__ __forceinline__ unsigned long long mix64(unsigned long long z) {  // splitmix64 finaliser
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

template <int NC>
__global__ void state_hash_kernel(const GridDims g, const double* __restrict__ u, const long long i0_global,
                                  unsigned long long* __restrict__ out) {
    const int j = 2 + blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long h = 0;
    if (j < g.ny + 2) {
        for (int i = 2 + blockIdx.y; i < g.nxl + 2; i += gridDim.y) {
            const size_t o = blocked_index<NC>(g, i, j);
            const unsigned long long cell = (unsigned long long)(i0_global + (i - 2)) * (unsigned long long)g.ny +
                                            (unsigned long long)(j - 2);
#pragma unroll
            for (int k = 0; k < NC; ++k) {
                const unsigned long long bits = (unsigned long long)__double_as_longlong(u[o + k * KSTR]);
                h += mix64(bits + 0x9E3779B97F4A7C15ull * (cell * NC + k + 1));
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) h += __shfl_xor_sync(0xffffffffu, h, o);
    if ((threadIdx.x & 31) == 0) atomicAdd(out, h);
}

void launch_state_hash(const GridDims& g, int nvar, const double* u, long long i0_global, unsigned long long* out,
                       cudaStream_t s) {
    dim3 block(128);
    dim3 grid((g.ny + 127) / 128, g.nxl < 256 ? g.nxl : 256);
    if (nvar == NV) state_hash_kernel<NV><<<grid, block, 0, s>>>(g, u, i0_global, out);
    else state_hash_kernel<4><<<grid, block, 0, s>>>(g, u, i0_global, out);
}

}  // namespace lsc

// state_hash.cu -- decomposition-independent 64-bit fingerprint of the interior of a state, computed on the device.
//
// Every interior value contributes mix(bit pattern + golden * (global cell number * nvar + k + 1)) and the
// contributions are summed modulo 2^64: the sum does not depend on how the columns are distributed over slabs /
// ranks (add the per-rank words modulo 2^64), while a changed bit or two swapped cells change it.  Used by
// bench.py (`sim.state_hash`) so that the N = 1/2/4/8 lines of a strong-scaling run can be compared by anyone, and
// by the tests against the same function evaluated with numpy on the downloaded array.
#include "state_hash.h"

#include <mutex>

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

// ---- allocation-free variant ------------------------------------------------------------------------------------
// For callers that cannot call cudaMalloc (tight serving loops, harnesses that forbid it): the scratch words live in
// this file-scope array instead of a per-call allocation.  Every GPU that loads the library holds its
// 64 KiB + 256 B until exit.  All calls on a device share it: two calls that overlap on the GPU (different streams,
// different host threads) race on the partial sums and on the result word -- callers serialise.  Nothing in it has
// to survive between calls and nothing has to start at zero except the result word, which every call clears first:
// the partial of each block is a plain store and the second kernel reads exactly the words the first one wrote.
namespace lsc {

constexpr int HASH_RESULT_WORD = STATE_HASH_STATIC_PARTIALS;  // first word of the 256-byte tail
static __device__ __align__(256) unsigned long long g_hash_scratch[STATE_HASH_STATIC_PARTIALS + 32];

// state_hash_kernel with the atomicAdd on one word replaced by one partial sum per block
template <int NC>
__global__ void state_hash_partials_kernel(const GridDims g, const double* __restrict__ u, const long long i0_global,
                                           unsigned long long* __restrict__ partials) {
    __shared__ unsigned long long warp_sum[4];
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
    if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = h;
    __syncthreads();
    if (threadIdx.x == 0)
        partials[blockIdx.y * gridDim.x + blockIdx.x] = warp_sum[0] + warp_sum[1] + warp_sum[2] + warp_sum[3];
}

// one block: sum of partials[0 .. n) modulo 2^64 -> *result
__global__ void state_hash_finish_kernel(const unsigned long long* __restrict__ partials, int n,
                                         unsigned long long* __restrict__ result) {
    __shared__ unsigned long long warp_sum[8];
    unsigned long long h = 0;
    for (int q = threadIdx.x; q < n; q += blockDim.x) h += partials[q];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) h += __shfl_xor_sync(0xffffffffu, h, o);
    if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = h;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long t = 0;
        for (int w = 0; w < 8; ++w) t += warp_sum[w];
        *result = t;
    }
}

// Device address of g_hash_scratch per device.  In host code the array's name is only a handle for the runtime, not
// a device address: it is resolved once per device, with that device current, and the kernels get the pointer.
static cudaError_t hash_scratch_address(int device, unsigned long long** out) {
    constexpr int MAXDEV = 64;
    static unsigned long long* addr[MAXDEV] = {};
    static std::mutex mu;
    if (device < 0 || device >= MAXDEV) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[device]) {
        int cur = -1;
        cudaError_t e = cudaGetDevice(&cur);
        if (e != cudaSuccess) return e;
        if (cur != device) return cudaErrorInvalidDevice;
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_hash_scratch);
        if (e != cudaSuccess) return e;
        addr[device] = static_cast<unsigned long long*>(p);
    }
    *out = addr[device];
    return cudaSuccess;
}

cudaError_t state_hash_static(int device, const GridDims& g, int nvar, const double* u, long long i0_global,
                              unsigned long long* host_out, cudaStream_t s) {
    if (g.ny < 1 || g.nxl < 1 || g.ny > STATE_HASH_STATIC_MAX_NY) return cudaErrorInvalidValue;
    unsigned long long* scratch = nullptr;
    cudaError_t e = hash_scratch_address(device, &scratch);
    if (e != cudaSuccess) return e;
    const int gx = (g.ny + 127) / 128;
    int gy = STATE_HASH_STATIC_PARTIALS / gx;
    if (gy > g.nxl) gy = g.nxl;
    if (gy > 256) gy = 256;
    e = cudaMemsetAsync(scratch + HASH_RESULT_WORD, 0, sizeof(unsigned long long), s);
    if (e != cudaSuccess) return e;
    const dim3 block(128), grid(gx, gy);
    if (nvar == NV) state_hash_partials_kernel<NV><<<grid, block, 0, s>>>(g, u, i0_global, scratch);
    else state_hash_partials_kernel<4><<<grid, block, 0, s>>>(g, u, i0_global, scratch);
    state_hash_finish_kernel<<<1, 256, 0, s>>>(scratch, gx * gy, scratch + HASH_RESULT_WORD);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    e = cudaMemcpyAsync(host_out, scratch + HASH_RESULT_WORD, sizeof(unsigned long long), cudaMemcpyDeviceToHost, s);
    if (e != cudaSuccess) return e;
    return cudaStreamSynchronize(s);
}

}  // namespace lsc

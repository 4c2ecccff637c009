// cfl_reduce.cuh -- publication of the per-step CFL maximum (shared by the MHD and Euler y sweeps).
#pragma once
#include <cuda_runtime.h>

namespace lsc {

__device__ __forceinline__ void atomic_max_nonneg(unsigned long long* addr, double v) {
    // v >= 0 and not NaN: the IEEE bit pattern is monotone in the value
    atomicMax(addr, (unsigned long long)__double_as_longlong(v));
}

// Warp-level publication: with the fp32 CFL filter most warps carry no candidate at all (local == 0 in every
// lane) and leave after one vote; a warp that does hold one publishes it with its own atomic -- no shared
// memory, no block barrier.  `local` started from 0.0 and was updated only with `>`, so NaNs never enter
// (MUSCL_NEW.cpp:111, SLIC_euler_2d.cpp:186).  Every lane of the warp must call it.
__device__ __forceinline__ void warp_max_to_global(double local, unsigned long long* out) {
    if (!__any_sync(0xffffffffu, local > 0.0)) return;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double other = __shfl_xor_sync(0xffffffffu, local, o);
        local = (other > local) ? other : local;
    }
    // Same-address atomics serialise in L2: in a region of uniform state every warp holds the running maximum
    // itself, so only a value that would actually raise it is sent.
    if ((threadIdx.x & 31) == 0 &&
        local > __longlong_as_double((long long)*(const volatile unsigned long long*)out))
        atomic_max_nonneg(out, local);
}

// 1/R rounded up, R = the running exact maximum (a bit pattern in *status) rounded down to fp32; R = 0 gives
// +inf and every bound becomes NaN/inf, i.e. every warp evaluates.  Stale values are fine: R only grows.
// In two halves, so that a kernel can issue the load first thing and convert where the value is needed:
// a warp that converts right behind the load sits out a whole L2 round trip before it has issued its tile loads.
// The conversion is a volatile asm so that it stays behind the barrier it is placed after.
__device__ __forceinline__ unsigned long long running_max_bits(const unsigned long long* status) {
    return *(const volatile unsigned long long*)status;
}
__device__ __forceinline__ float inv_running_max_from_bits(unsigned long long bits) {
    float r;
    asm volatile("cvt.rm.f32.f64 %0, %1;" : "=f"(r) : "d"(__longlong_as_double((long long)bits)));
    return __frcp_ru(r);
}

}  // namespace lsc

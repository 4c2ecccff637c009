// mhd_kat.cu -- known-answer test hooks: run the device flux functions on arrays of states so
// tests/ can compare them bit-for-bit with the oracle ("unit level").
#include <cuda_runtime.h>

#include "../../include/lsc_b200.h"
#include "mhd_physics.cuh"

void lsc_count_launches(unsigned long long n);

namespace lsc {

template <int DIR>
__global__ void kat_hllc_kernel(long n, const double* __restrict__ L, const double* __restrict__ R, double gamma,
                                double ch, double* __restrict__ F) {
    const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    double l[NV], r[NV], f[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        l[k] = L[t * NV + k];
        r[k] = R[t * NV + k];
    }
    FastDiv dv;
    hllc_flux<DIR>(dv, l, r, gamma, gamma - 1, ch, ch * ch, 1.0 / (2.0 * ch), ch / 2.0, f);
    if (dv.bad()) {
        IeeeDiv sv;
        hllc_flux<DIR>(sv, l, r, gamma, gamma - 1, ch, ch * ch, 1.0 / (2.0 * ch), ch / 2.0, f);
    }
#pragma unroll
    for (int k = 0; k < NV; ++k) F[t * NV + k] = f[k];
}

template <int DIR>
__global__ void kat_fluxdef_kernel(long n, const double* __restrict__ U, double gamma, double ch,
                                   double* __restrict__ F) {
    const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    double u[NV], f[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) u[k] = U[t * NV + k];
    FastDiv dv;
    flux_def<DIR>(dv, u, gamma - 1, ch * ch, f);
    if (dv.bad()) {
        IeeeDiv sv;
        flux_def<DIR>(sv, u, gamma - 1, ch * ch, f);
    }
#pragma unroll
    for (int k = 0; k < NV; ++k) F[t * NV + k] = f[k];
}

// ---- the shared-reciprocal division (fp64_div.cuh) against the compiler's IEEE operator ----
__device__ __forceinline__ unsigned long long mix64(unsigned long long z) {
    z += 0x9e3779b97f4a7c15ull;
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
}

// mode 0: random mantissas, exponents within +-40 of 1 (the solver's regime)
// mode 1: fully random bit patterns (all exponents, NaN/inf/denormals): flag must catch what is unsafe
// mode 2: divisor mantissa all ones / dividends one ulp around exact multiples (half-ulp ties)
// mode 3: tiny, denormal and zero dividends over solver-range divisors (numerical fronts into gas at rest)
// POLICY: FastDiv (sticky flag) or BranchDiv (never flags: every quotient must equal IEEE)
template <class POLICY>
__global__ void kat_div_kernel(unsigned long long seed, int mode, int per_thread, unsigned long long* mismatches,
                               unsigned long long* flagged) {
    const unsigned long long t = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long bad = 0, flg = 0;
    for (int it = 0; it < per_thread; ++it) {
        const unsigned long long ra = mix64(seed + 2 * (t * per_thread + it));
        const unsigned long long rb = mix64(seed + 2 * (t * per_thread + it) + 1);
        double a, b;
        if (mode == 1) {
            a = __longlong_as_double((long long)ra);
            b = __longlong_as_double((long long)rb);
        } else if (mode == 3) {
            const unsigned long long eb = 1023 - 40 + (rb >> 52) % 81;
            b = __longlong_as_double((long long)(((rb >> 63) << 63) | (eb << 52) | (rb & 0xfffffffffffffull)));
            const unsigned long long ea = (ra >> 52) % 300;  // exponent field 0..299: denormal .. 2^-724
            a = __longlong_as_double((long long)(((ra >> 63) << 63) | (ea << 52) | (ra & 0xfffffffffffffull)));
            if ((it % 5) == 0) a = __longlong_as_double((long long)((ra >> 63) << 63));  // +-0
        } else {
            const unsigned long long ea = 1023 - 40 + (ra >> 52) % 81, eb = 1023 - 40 + (rb >> 52) % 81;
            unsigned long long ma = ra & 0xfffffffffffffull, mb = rb & 0xfffffffffffffull;
            if (mode == 2) {
                mb |= 0xffffffffff000ull;                   // divisor just below a power of two
                if (it & 1) ma = (ma & 0xfff) | 0xffffffffff000ull;
            }
            a = __longlong_as_double((long long)(((ra >> 63) << 63) | (ea << 52) | ma));
            b = __longlong_as_double((long long)(((rb >> 63) << 63) | (eb << 52) | mb));
            if (mode == 2 && (it & 2)) {                   // a := (small integer k) * b +- 1 ulp
                const double k = (double)(1 + (ra & 1023));
                a = __longlong_as_double(__double_as_longlong(k * b) + ((ra >> 20) & 1 ? 1 : -1));
            }
            if (mode == 0 && (it % 17) == 0) a = 0.0;      // exact zero dividends are common in the solver
        }
        POLICY dv;
        double q;
        if (it & 4) {
            q = dv.div(a, b);  // the single-quotient form (its own window bookkeeping)
        } else {
            const Rcp r = dv.rcp(b);
            q = dv.div(a, r);
        }
        const double ref = a / b;
        if (dv.bad()) {
            ++flg;
        } else if (__double_as_longlong(q) != __double_as_longlong(ref) && !(q != q && ref != ref)) {
            ++bad;  // (two NaNs count as equal)
        }
    }
    atomicAdd(mismatches, bad);
    atomicAdd(flagged, flg);
}

// FastDiv::sqrt against sqrt(): mode 0 solver-range positive operands, 1 arbitrary bit patterns (negative, NaN, inf,
// subnormal: must be flagged, never wrong), 2 perfect squares and their neighbours (exact / half-ulp cases)
__global__ void kat_sqrt_kernel(unsigned long long seed, int mode, int per_thread, unsigned long long* mismatches,
                                unsigned long long* flagged) {
    const unsigned long long t = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long bad = 0, flg = 0;
    for (int it = 0; it < per_thread; ++it) {
        const unsigned long long ra = mix64(seed + (t * per_thread + it));
        double a;
        if (mode == 1) {
            a = __longlong_as_double((long long)ra);
        } else {
            const unsigned long long ea = 1023 - 60 + (ra >> 52) % 121;
            a = __longlong_as_double((long long)((ea << 52) | (ra & 0xfffffffffffffull)));
            if (mode == 2) {
                const double k = (double)(1 + (ra & 0x3ffffff));  // k*k is exact (52 bits)
                a = __longlong_as_double(__double_as_longlong(k * k) + (long long)((ra >> 40) % 3) - 1);
            }
        }
        FastDiv dv;
        const double q = dv.sqrt(a);
        const double ref = sqrt(a);
        if (dv.bad()) {
            ++flg;
        } else if (__double_as_longlong(q) != __double_as_longlong(ref) && !(q != q && ref != ref)) {
            ++bad;
        }
    }
    atomicAdd(mismatches, bad);
    atomicAdd(flagged, flg);
}

// exact signal speed (IEEE division) and the fp32 upper bound of speed/run_max; out[2t] = exact, out[2t+1] = bound
__global__ void kat_cfl_bound_kernel(long n, const double* __restrict__ U, double gamma, double run_max,
                                     double* __restrict__ out) {
    const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    double u[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) u[k] = U[t * NV + k];
    IeeeDiv dv;
    out[2 * t] = cfl_speed(dv, u, gamma);
    out[2 * t + 1] = (double)cfl_speed_over_max_f32(u, __double2float_ru(gamma), __double2float_ru(gamma - 1),
                                                    __frcp_ru(__double2float_rd(run_max)));
}

static int run_kat(int device, int dir, long n, const double* a, const double* b, double gamma, double ch,
                   double* out) {
    if (n <= 0 || !a || !out || (dir != 0 && dir != 1)) return LSC_EINVAL;
    if (cudaSetDevice(device) != cudaSuccess) return LSC_ECUDA;
    const size_t bytes = (size_t)n * NV * sizeof(double);
    double *da = nullptr, *db = nullptr, *dout = nullptr;
    int rc = LSC_OK;
    if (cudaMalloc(&da, bytes) != cudaSuccess || cudaMalloc(&dout, bytes) != cudaSuccess ||
        (b && cudaMalloc(&db, bytes) != cudaSuccess)) {
        rc = LSC_ENOMEM;
    } else {
        cudaMemcpy(da, a, bytes, cudaMemcpyHostToDevice);
        if (b) cudaMemcpy(db, b, bytes, cudaMemcpyHostToDevice);
        const int block = 128;
        const unsigned grid = (unsigned)((n + block - 1) / block);
        if (b) {
            if (dir == 0) kat_hllc_kernel<0><<<grid, block>>>(n, da, db, gamma, ch, dout);
            else kat_hllc_kernel<1><<<grid, block>>>(n, da, db, gamma, ch, dout);
        } else {
            if (dir == 0) kat_fluxdef_kernel<0><<<grid, block>>>(n, da, gamma, ch, dout);
            else kat_fluxdef_kernel<1><<<grid, block>>>(n, da, gamma, ch, dout);
        }
        lsc_count_launches(1);
        if (cudaDeviceSynchronize() != cudaSuccess ||
            cudaMemcpy(out, dout, bytes, cudaMemcpyDeviceToHost) != cudaSuccess) {
            cudaGetLastError();
            rc = LSC_ECUDA;
        }
    }
    cudaFree(da);
    cudaFree(db);
    cudaFree(dout);
    return rc;
}

}  // namespace lsc

extern "C" int lsc_kat_division(int device, unsigned long long seed, int mode, long nblocks, int per_thread,
                                unsigned long long* mismatches, unsigned long long* flagged) {
    if (!mismatches || !flagged || nblocks <= 0 || per_thread <= 0) return LSC_EINVAL;
    if (cudaSetDevice(device) != cudaSuccess) return LSC_ECUDA;
    unsigned long long* d = nullptr;
    if (cudaMalloc(&d, 16) != cudaSuccess) return LSC_ENOMEM;
    cudaMemset(d, 0, 16);
    if (mode >= 20)
        lsc::kat_sqrt_kernel<<<(unsigned)nblocks, 256>>>(seed, mode - 20, per_thread, d, d + 1);
    else if (mode >= 10)
        lsc::kat_div_kernel<lsc::BranchDiv><<<(unsigned)nblocks, 256>>>(seed, mode - 10, per_thread, d, d + 1);
    else
        lsc::kat_div_kernel<lsc::FastDiv><<<(unsigned)nblocks, 256>>>(seed, mode, per_thread, d, d + 1);
    lsc_count_launches(1);
    unsigned long long h[2] = {0, 0};
    int rc = LSC_OK;
    if (cudaDeviceSynchronize() != cudaSuccess || cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost) != cudaSuccess) {
        cudaGetLastError();
        rc = LSC_ECUDA;
    }
    cudaFree(d);
    *mismatches = h[0];
    *flagged = h[1];
    return rc;
}

extern "C" int lsc_mhd_kat_cfl_bound(int device, long n, const double* U, double gamma, double run_max,
                                     double* exact_and_bound) {
    if (n <= 0 || !U || !exact_and_bound) return LSC_EINVAL;
    if (cudaSetDevice(device) != cudaSuccess) return LSC_ECUDA;
    double *du = nullptr, *dout = nullptr;
    int rc = LSC_OK;
    if (cudaMalloc(&du, (size_t)n * lsc::NV * 8) != cudaSuccess || cudaMalloc(&dout, (size_t)n * 16) != cudaSuccess) {
        rc = LSC_ENOMEM;
    } else {
        cudaMemcpy(du, U, (size_t)n * lsc::NV * 8, cudaMemcpyHostToDevice);
        lsc::kat_cfl_bound_kernel<<<(unsigned)((n + 127) / 128), 128>>>(n, du, gamma, run_max, dout);
        lsc_count_launches(1);
        if (cudaDeviceSynchronize() != cudaSuccess ||
            cudaMemcpy(exact_and_bound, dout, (size_t)n * 16, cudaMemcpyDeviceToHost) != cudaSuccess) {
            cudaGetLastError();
            rc = LSC_ECUDA;
        }
    }
    cudaFree(du);
    cudaFree(dout);
    return rc;
}

extern "C" int lsc_mhd_kat_hllc(int device, int dir, long n, const double* L, const double* R, double gamma,
                                double ch, double* F) {
    if (!R) return LSC_EINVAL;
    return lsc::run_kat(device, dir, n, L, R, gamma, ch, F);
}

extern "C" int lsc_mhd_kat_fluxdef(int device, int dir, long n, const double* U, double gamma, double ch,
                                   double* F) {
    return lsc::run_kat(device, dir, n, U, nullptr, gamma, ch, F);
}

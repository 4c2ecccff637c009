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
    hllc_flux<DIR>(l, r, gamma, gamma - 1, ch, ch * ch, f);
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
    flux_def<DIR>(u, gamma - 1, ch * ch, f);
#pragma unroll
    for (int k = 0; k < NV; ++k) F[t * NV + k] = f[k];
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

extern "C" int lsc_mhd_kat_hllc(int device, int dir, long n, const double* L, const double* R, double gamma,
                                double ch, double* F) {
    if (!R) return LSC_EINVAL;
    return lsc::run_kat(device, dir, n, L, R, gamma, ch, F);
}

extern "C" int lsc_mhd_kat_fluxdef(int device, int dir, long n, const double* U, double gamma, double ch,
                                   double* F) {
    return lsc::run_kat(device, dir, n, U, nullptr, gamma, ch, F);
}

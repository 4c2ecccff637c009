// fp64_pipe.cu -- B200 fp64 pipe facts the sweep kernels are designed against (development aid):
//   dependent-issue latency of DFMA/DADD/DMUL, per-SM throughput vs warps/scheduler and ILP,
//   and whether integer instructions interleave for free when the fp64 pipe is saturated.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -o fp64_pipe fp64_pipe.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int ILP, int INTS, int OP>
__global__ void k(double* out, int iters, double a, double b, long long* clk) {
    double x[ILP];
    unsigned m[4] = {threadIdx.x, threadIdx.x * 3u, 7u, 11u};
#pragma unroll
    for (int i = 0; i < ILP; ++i) x[i] = a + i + threadIdx.x;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 8; ++r) {
#pragma unroll
            for (int i = 0; i < ILP; ++i) {
                if (OP == 0) x[i] = __fma_rn(x[i], a, b);
                if (OP == 1) x[i] = __dadd_rn(x[i], b);
                if (OP == 2) x[i] = __dmul_rn(x[i], a);
#pragma unroll
                for (int q = 0; q < INTS; ++q) m[q & 3] = max(m[q & 3], (unsigned)__double2hiint(x[i]) * 2u - 0x20000000u + q);
            }
        }
    }
    long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s + m[0] + m[1] + m[2] + m[3];
    if (threadIdx.x == 0 && blockIdx.x == 0) *clk = t1 - t0;
}

template <int ILP, int INTS, int OP>
void run(const char* name, int warps_per_sm) {
    double* out; long long* clk; long long h;
    int threads = 128, blocks_per_sm = warps_per_sm / 4;
    if (warps_per_sm < 4) { threads = 32 * warps_per_sm; blocks_per_sm = 1; }
    int blocks = 148 * blocks_per_sm, iters = 2000;
    cudaMalloc(&out, sizeof(double) * blocks * threads); cudaMalloc(&clk, 8);
    k<ILP, INTS, OP><<<blocks, threads>>>(out, 10, 1.0000001, 1e-9, clk);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k<ILP, INTS, OP><<<blocks, threads>>>(out, iters, 1.0000001, 1e-9, clk);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    cudaMemcpy(&h, clk, 8, cudaMemcpyDeviceToHost);
    double ninst = (double)iters * 8 * ILP;  // fp64 instr per thread
    double per_sm_per_clk = ninst * threads * blocks_per_sm / (double)h;
    printf("%-6s ILP=%d INTS=%d warps/SM=%2d : %.2f clk per fp64 instr per warp, %.1f fp64 lanes/clk/SM, %.2f Tinst/s\n",
           name, ILP, INTS, warps_per_sm, (double)h / ninst, per_sm_per_clk, ninst * threads * blocks / (ms * 1e-3) / 1e12);
    cudaFree(out); cudaFree(clk);
}

int main() {
    run<1, 0, 0>("DFMA", 1); run<1, 0, 1>("DADD", 1); run<1, 0, 2>("DMUL", 1);
    run<2, 0, 0>("DFMA", 1); run<4, 0, 0>("DFMA", 1); run<8, 0, 0>("DFMA", 1);
    run<1, 0, 0>("DFMA", 4); run<2, 0, 0>("DFMA", 4); run<4, 0, 0>("DFMA", 4); run<8, 0, 0>("DFMA", 4);
    run<1, 0, 0>("DFMA", 16); run<2, 0, 0>("DFMA", 16); run<4, 0, 0>("DFMA", 16); run<4, 0, 0>("DFMA", 32);
    run<4, 0, 1>("DADD", 16); run<4, 0, 2>("DMUL", 16);
    run<4, 1, 0>("DFMA", 16); run<4, 2, 0>("DFMA", 16); run<4, 3, 0>("DFMA", 16); run<4, 4, 0>("DFMA", 16);
    run<2, 2, 0>("DFMA", 16); run<2, 2, 0>("DFMA", 32);
    return 0;
}

// hllc2d.cu -- the first-order sibling scheme of the MHD path:
// MHD_codes/HLLC_MHD_2D.cpp (H below) -- dimensionally split first-order Godunov update of the 9-component GLM-MHD
// system with the HLLC flux, ONE ghost cell, periodic in x and y, Orszag-Tang initial data.  Kernels + C ABI
// (lsc_hllc2d_*, include/lsc_b200.h) in one translation unit; compiled with -fmad=false like the rest.
//
// Parity is bit equality with the REAL program (tests/test_hllc2d.py compares with fixtures and live runs of H built
// from the reference's own source): every expression below keeps H's association, IEEE division and sqrt.
// H's flux functions differ from MUSCL_NEW's
// in rounding-relevant ways (pressure as (gamma-1)*(E - (kinetic + magnetic)), sound speed squared from a rounded
// square root, star pressures with half weights, FluxDef without the GLM entries), so they are restated here rather
// than shared with mhd_physics.cuh; what IS shared is ComputeTimeStep's signal speed (H:71-120 is M:71-120 verbatim).
//
// What one pass of H's loop (H:728-786) does to the cells that matter, derived from its loop bounds:
//   ComputeTimeStep over i in [1,nx-1], j in [1,ny-1] (`< size()-2` on arrays of nx+2: the LAST interior row and
//   column are left out, H:73-74); applyBoundaryConditions(u) (periodic images into the one ghost layer);
//   fluxX[i] = FluxHLLCX(u[i], u[i+1]); uBar = u - (dt/dx)(fluxX[i] - fluxX[i-1]); BC(uBar); the same in y;
//   psi of the new state times exp(-(0.5 dt) ch / 0.18) ONCE (H:776-780 -- the damping at the head of the loop,
//   H:730-734, scales uPlus1 before it is overwritten and uses the outer, never-updated dt/maxSpeed: it only reaches
//   ghost cells, which the next applyBoundaryConditions overwrites).  Ghost cells never influence an interior cell
//   except as periodic images, so the device stores the nx x ny interior only and indexes neighbours modulo n.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <new>
#include <vector>

#include <cuda_runtime.h>

#include "cfl_reduce.cuh"
#include "lsc_common.h"
#include "mhd_physics.cuh"

namespace lsc {
namespace h2d {

constexpr int NQ = 9;

// ConservativeToPrimitive H:22-34: velocities and pressure as FluxDef sees them
struct PrimH {
    double vx, vy, vz, p;
};
__device__ __forceinline__ PrimH c2p_h(const double (&u)[NQ], double gamma) {
    PrimH q;
    q.vx = u[1] / u[0];
    q.vy = u[2] / u[0];
    q.vz = u[3] / u[0];
    q.p = (u[4] - 0.5 * u[0] * (q.vx * q.vx + q.vy * q.vy + q.vz * q.vz) - 0.5 * (u[5] * u[5] + u[6] * u[6] + u[7] * u[7])) *
          (gamma - 1);
    return q;
}

// FluxDefX H:41-54 / FluxDefY H:56-69: no GLM entries (flux[n] = flux[8] = 0)
template <int DIR>
__device__ __forceinline__ void flux_def_h(const double (&u)[NQ], double gamma, double (&f)[NQ]) {
    const PrimH q = c2p_h(u, gamma);
    const double hB2 = 0.5 * (u[5] * u[5] + u[6] * u[6] + u[7] * u[7]);
    const double vB = q.vx * u[5] + q.vy * u[6] + q.vz * u[7];
    if (DIR == 0) {
        f[0] = u[1];
        f[1] = u[0] * q.vx * q.vx + q.p + hB2 - u[5] * u[5];
        f[2] = u[0] * q.vx * q.vy - u[5] * u[6];
        f[3] = u[0] * q.vx * q.vz - u[5] * u[7];
        f[4] = (u[4] + q.p + hB2) * q.vx - vB * u[5];
        f[5] = 0;
        f[6] = u[6] * q.vx - u[5] * q.vy;
        f[7] = u[7] * q.vx - u[5] * q.vz;
        f[8] = 0;
    } else {
        f[0] = u[2];
        f[1] = u[0] * q.vx * q.vy - u[5] * u[6];
        f[2] = u[0] * q.vy * q.vy + q.p + hB2 - u[6] * u[6];
        f[3] = u[0] * q.vy * q.vz - u[6] * u[7];
        f[4] = (u[4] + q.p + hB2) * q.vy - vB * u[6];
        f[5] = u[5] * q.vy - u[6] * q.vx;
        f[6] = 0;
        f[7] = u[7] * q.vy - u[6] * q.vz;
        f[8] = 0;
    }
}

// One side of an interface as uHLLx/y and FluxHLLCX/Y see it (H:143-164, H:286-305): velocities, pressure from the
// summed kinetic + magnetic term, total pressure, fast magnetosonic speed along the normal.  `Bn` is the normal
// field the caller wants in the magnetic terms: the state's own in FluxHLLC, the GLM-cleaned one in uHLL.
struct SideH {
    double ux, uy, uz, p, pT, cf;
};
template <int DIR>
__device__ __forceinline__ SideH side_h(const double (&s)[NQ], double Bn, double gamma) {
    SideH q;
    const double rho = s[0];
    q.ux = s[1] / rho;
    q.uy = s[2] / rho;
    q.uz = s[3] / rho;
    const double Bx = (DIR == 0) ? Bn : s[5], By = (DIR == 0) ? s[6] : Bn, Bz = s[7];
    const double inter = 0.5 * rho * (q.ux * q.ux + q.uy * q.uy + q.uz * q.uz) + 0.5 * (Bx * Bx + By * By + Bz * Bz);
    const double B2 = Bx * Bx + By * By + Bz * Bz;
    q.p = (gamma - 1.0) * (s[4] - inter);
    q.pT = q.p + 0.5 * B2;
    const double a = sqrt(gamma * q.p / rho);
    const double t = a * a + B2 / rho;  // both occurrences of (a*a + B2/rho) are the same value
    q.cf = sqrt(0.5 * (a * a + (B2 / rho) + sqrt(t * t - 4.0 * a * a * Bn * Bn / rho)));
    return q;
}

// FluxHLLCX H:281-429 / FluxHLLCY H:431-581 with psiAndBx/y H:121-139 and uHLLx/y H:141-280.
template <int DIR>
__device__ __forceinline__ void flux_hllc_h(const double (&L)[NQ], const double (&R)[NQ], double gamma, double ch,
                                            double (&F)[NQ]) {
    constexpr int N = (DIR == 0) ? 5 : 6;   // normal field
    constexpr int T = (DIR == 0) ? 6 : 5;   // in-plane tangential field
    constexpr int MN = (DIR == 0) ? 1 : 2;  // normal momentum
    constexpr int MT = (DIR == 0) ? 2 : 1;
    // mixed GLM interface values (H:125-126)
    const double BnT = 0.5 * (L[N] + R[N]) - (1.0 / (2.0 * ch)) * (R[8] - L[8]);
    const double psiT = 0.5 * (L[8] + R[8]) - (ch / 2.0) * (R[N] - L[N]);

    // ---- wave speeds and contact speed from the ORIGINAL states (H:286-336)
    const SideH a = side_h<DIR>(L, L[N], gamma), b = side_h<DIR>(R, R[N], gamma);
    const double unL = (DIR == 0) ? a.ux : a.uy, unR = (DIR == 0) ? b.ux : b.uy;
    const double cmax = std_max(a.cf, b.cf);
    const double sL = std_min(unL, unR) - cmax;
    const double sR = std_max(unL, unR) + cmax;
    const double qS = (b.pT - a.pT + L[0] * unL * (sL - unL) - R[0] * unR * (sR - unR)) /
                      (L[0] * (sL - unL) - R[0] * (sR - unR));

    double FL[NQ], FR[NQ];
    flux_def_h<DIR>(L, gamma, FL);
    flux_def_h<DIR>(R, gamma, FR);

    // ---- HLL state (uHLLx/y): its own wave speeds, formed with the CLEANED normal field in the magnetic terms
    double H0, H1, H2, H3, HT, HZ;  // rho, momenta, tangential fields of the HLL state
    {
        const SideH c = side_h<DIR>(L, BnT, gamma), d = side_h<DIR>(R, BnT, gamma);
        const double vnL = (DIR == 0) ? c.ux : c.uy, vnR = (DIR == 0) ? d.ux : d.uy;
        const double cm = std_max(c.cf, d.cf);
        const double hL = std_min(vnL, vnR) - cm;
        const double hR = std_max(vnL, vnR) + cm;
        if (hL >= 0) {
            H0 = L[0]; H1 = L[1]; H2 = L[2]; H3 = L[3]; HT = L[T]; HZ = L[7];
        } else if (hL < 0 && hR > 0) {
            const double w = hR - hL;
            H0 = (hR * R[0] - hL * L[0] - FR[0] + FL[0]) / w;
            H1 = (hR * R[1] - hL * L[1] - FR[1] + FL[1]) / w;
            H2 = (hR * R[2] - hL * L[2] - FR[2] + FL[2]) / w;
            H3 = (hR * R[3] - hL * L[3] - FR[3] + FL[3]) / w;
            HT = (hR * R[T] - hL * L[T] - FR[T] + FL[T]) / w;
            HZ = (hR * R[7] - hL * L[7] - FR[7] + FL[7]) / w;
        } else {
            H0 = R[0]; H1 = R[1]; H2 = R[2]; H3 = R[3]; HT = R[T]; HZ = R[7];
        }
    }
    const double uxH = H1 / H0, uyH = H2 / H0, uzH = H3 / H0;

    // ---- region (H:413-424), then only the star state that region uses
    // the reference's if-chain as it stands.  (Written as three chained boolean expressions -- regB = !regA && (sL <= 0
    // && qS >= 0) ... -- cicc 12.9 at -O3 compiled this function to "right star state whenever a star state is
    // used": caught by the first parity run, gone with -Xcicc -O1 or with this form; ptxas -O1 did not matter.)
    int region;
    if (sL >= 0) region = 0;
    else if (sL <= 0 && qS >= 0) region = 1;
    else if (qS <= 0 && sR >= 0) region = 2;
    else region = 3;
    const bool regA = region == 0, regB = region == 1, regC = region == 2;
    if (regA || (!regB && !regC)) {
#pragma unroll
        for (int k = 0; k < NQ; ++k) F[k] = regA ? FL[k] : FR[k];
    } else {
        const bool left = regB;
        // the selected side, element by element
        double S[NQ], FS[NQ];
#pragma unroll
        for (int k = 0; k < NQ; ++k) {
            S[k] = left ? L[k] : R[k];
            FS[k] = left ? FL[k] : FR[k];
        }
        SideH q;
        q.ux = left ? a.ux : b.ux;
        q.uy = left ? a.uy : b.uy;
        q.uz = left ? a.uz : b.uz;
        q.p = left ? a.p : b.p;
        q.pT = left ? a.pT : b.pT;
        q.cf = left ? a.cf : b.cf;
        const double ss = left ? sL : sR;
        const double un = left ? unL : unR;
        // star fields: normal = cleaned value, tangential = HLL values (H:347-349 / H:497-499)
        const double BxS = (DIR == 0) ? BnT : HT, ByS = (DIR == 0) ? HT : BnT, BzS = HZ;
        const double Bn = S[N];
        // H:352 (left: gas pressure) vs H:392 (right: TOTAL pressure) -- the reference's asymmetry, kept
        const double pS = S[0] * (ss - un) * (qS - un) + (left ? q.p : q.pT) - 0.5 * Bn * Bn + 0.5 * BnT * BnT;
        const double rhoS = S[0] * (ss - un) / (ss - qS);
        const double mnS = rhoS * qS;
        const double mtS = S[MT] * (ss - un) / (ss - qS) - (BxS * ByS - S[5] * S[6]) / (ss - qS);
        const double mzS = S[3] * (ss - un) / (ss - qS) - (BnT * BzS - Bn * S[7]) / (ss - qS);
        const double vBS = BxS * uxH + ByS * uyH + BzS * uzH;
        const double vB = S[5] * q.ux + S[6] * q.uy + S[7] * q.uz;
        const double ES = (S[4] * (ss - un) - q.pT * un + pS * qS + Bn * (vB - vBS)) / (ss - qS);
        double US[NQ];
        US[0] = rhoS;
        US[MN] = mnS;
        US[MT] = mtS;
        US[3] = mzS;
        US[4] = ES;
        US[5] = BxS;
        US[6] = ByS;
        US[7] = BzS;
        US[8] = psiT;
#pragma unroll
        for (int k = 0; k < NQ; ++k) F[k] = FS[k] + ss * (US[k] - S[k]);
    }
    F[N] = psiT;          // H:426 / H:578
    F[8] = ch * ch * BnT;  // H:427 / H:579
}

// interior-only SoA planes: component k of cell (i, j), 0-based interior indices, j contiguous
struct Grid {
    int nx, ny;
    size_t plane;
};
__host__ __device__ inline size_t at(const Grid& g, int k, int i, int j) {
    return (size_t)k * g.plane + (size_t)i * (size_t)g.ny + (size_t)j;
}
__device__ __forceinline__ void load9(const Grid& g, const double* __restrict__ u, int i, int j, double (&q)[NQ]) {
#pragma unroll
    for (int k = 0; k < NQ; ++k) q[k] = __ldg(u + at(g, k, i, j));
}

// x update (H:739-755): uBar = u - (dt/dx)(F_{i+1/2} - F_{i-1/2}).  One thread per row j marches along i over a strip
// of columns with the previous interface flux and the next cell in registers, so every interface flux is evaluated
// once (plus one per strip); adjacent threads are adjacent in j, the contiguous axis.  Neighbours by periodic wrap.
constexpr int H2D_STRIP = 32;
__global__ void __launch_bounds__(128) x_update_kernel(const Grid g, const double* __restrict__ in,
                                                       double* __restrict__ out, double gamma, double ch, double dtdh) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= g.ny) return;
    const int ia = blockIdx.y * H2D_STRIP;
    const int ib = min(ia + H2D_STRIP, g.nx);
    double um[NQ], u0[NQ], Fm[NQ];
    load9(g, in, ia == 0 ? g.nx - 1 : ia - 1, j, um);
    load9(g, in, ia, j, u0);
    flux_hllc_h<0>(um, u0, gamma, ch, Fm);
    for (int i = ia; i < ib; ++i) {
        double up[NQ], Fp[NQ];
        load9(g, in, i == g.nx - 1 ? 0 : i + 1, j, up);
        flux_hllc_h<0>(u0, up, gamma, ch, Fp);
#pragma unroll
        for (int k = 0; k < NQ; ++k) {
            out[at(g, k, i, j)] = u0[k] - dtdh * (Fp[k] - Fm[k]);
            Fm[k] = Fp[k];
            u0[k] = up[k];
        }
    }
}

// y update (H:757-780): u' = uBar - (dt/dy)(F_{j+1/2} - F_{j-1/2}), psi damping, and the next step's signal speed
// over H's range (H:73-74).  One thread per cell along j; a block of 128 threads evaluates the 128 interface fluxes
// above its cells (one per thread), passes them down through shared memory and updates 127 cells.
constexpr int H2D_BJ = 128;
__global__ void __launch_bounds__(H2D_BJ) y_update_kernel(const Grid g, const double* __restrict__ in,
                                                          double* __restrict__ out, double gamma, double ch, double dtdh,
                                                          double psi_factor, unsigned long long* __restrict__ status) {
    __shared__ double sF[NQ][H2D_BJ];
    const int i = blockIdx.y;
    const int t = threadIdx.x;
    const int c = (int)blockIdx.x * (H2D_BJ - 1) - 1 + t;  // this thread's cell; thread 0 only supplies the flux below cell 1
    const bool in_range = c < g.ny;                        // (c == -1 is the periodic image ny-1)
    double u0[NQ], F[NQ];
    if (in_range) {
        const int jc = c < 0 ? g.ny - 1 : c;
        double up[NQ];
        load9(g, in, i, jc, u0);
        load9(g, in, i, jc == g.ny - 1 ? 0 : jc + 1, up);
        flux_hllc_h<1>(u0, up, gamma, ch, F);
#pragma unroll
        for (int k = 0; k < NQ; ++k) sF[k][t] = F[k];
    }
    __syncthreads();
    double speed = 0.0;
    if (in_range && t >= 1) {
        double un[NQ];
#pragma unroll
        for (int k = 0; k < NQ; ++k) {
            double v = u0[k] - dtdh * (F[k] - sF[k][t - 1]);
            if (k == 8) v = v * psi_factor;
            un[k] = v;
            out[at(g, k, i, c)] = v;
        }
        if (i < g.nx - 1 && c < g.ny - 1) {  // H:73-74: the last interior row and column are left out
            IeeeDiv dv;
            speed = cfl_speed(dv, un, gamma);  // H:75-109 == M:75-109
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double other = __shfl_xor_sync(0xffffffffu, speed, o);
        speed = (other > speed) ? other : speed;
    }
    if ((t & 31) == 0 && speed > 0.0) atomic_max_nonneg(status, speed);
}

__global__ void max_speed_kernel_h(const Grid g, const double* __restrict__ u, double gamma,
                                   unsigned long long* __restrict__ status) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int i = blockIdx.y;
    double speed = 0.0;
    if (j < g.ny - 1 && i < g.nx - 1) {
        double q[NQ];
        load9(g, u, i, j, q);
        IeeeDiv dv;
        const double sp = cfl_speed(dv, q, gamma);
        if (sp > speed) speed = sp;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double other = __shfl_xor_sync(0xffffffffu, speed, o);
        speed = (other > speed) ? other : speed;
    }
    if ((threadIdx.x & 31) == 0 && speed > 0.0) atomic_max_nonneg(status, speed);
}

// reference array [(nx+2)][(ny+2)][9] (one ghost layer) <-> interior planes
__global__ void aos_to_planes(const Grid g, const double* __restrict__ aos, double* __restrict__ u) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
    if (j >= g.ny) return;
    const double* s = aos + ((size_t)(i + 1) * (size_t)(g.ny + 2) + (size_t)(j + 1)) * NQ;
#pragma unroll
    for (int k = 0; k < NQ; ++k) u[at(g, k, i, j)] = s[k];
}
__global__ void planes_to_aos(const Grid g, const double* __restrict__ u, double* __restrict__ aos) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
    if (j >= g.ny) return;
    double* d = aos + ((size_t)(i + 1) * (size_t)(g.ny + 2) + (size_t)(j + 1)) * NQ;
#pragma unroll
    for (int k = 0; k < NQ; ++k) d[k] = u[at(g, k, i, j)];
}

}  // namespace h2d
}  // namespace lsc

struct lsc_hllc2d_ctx {
    lsc_hllc2d_config cfg;
    lsc::h2d::Grid g{};
    double dx = 0, dy = 0;
    double* u[3] = {nullptr, nullptr, nullptr};  // u^n, uBar, u^{n+1}
    int cur = 0;
    double* d_aos = nullptr;
    unsigned long long* d_status = nullptr;
    unsigned long long* h_status = nullptr;
    cudaStream_t stream = nullptr;
    double t = 0, dt = 0, ch = 0;
    long step = 0;
    bool uploaded = false;
};

namespace {
using lsc::h2d::NQ;

inline double* cell(double* a, int ny, int i, int j) { return a + ((size_t)i * (size_t)(ny + 2) + (size_t)j) * NQ; }

// applyBoundaryConditions H:583-612: periodic in x over j in [0, ny+1], then periodic in y over i in [0, nx+1]
void host_bc(const lsc_hllc2d_config& c, double* a) {
    const int nx = c.nx, ny = c.ny;
    for (int j = 0; j < ny + 2; ++j) {
        std::memcpy(cell(a, ny, 0, j), cell(a, ny, nx, j), NQ * sizeof(double));
        std::memcpy(cell(a, ny, nx + 1, j), cell(a, ny, 1, j), NQ * sizeof(double));
    }
    for (int i = 0; i < nx + 2; ++i) {
        std::memcpy(cell(a, ny, i, 0), cell(a, ny, i, ny), NQ * sizeof(double));
        std::memcpy(cell(a, ny, i, ny + 1), cell(a, ny, i, 1), NQ * sizeof(double));
    }
}

dim3 grid_of(const lsc::h2d::Grid& g) { return dim3((g.ny + 127) / 128, g.nx); }

}  // namespace

extern "C" {

int lsc_hllc2d_create(const lsc_hllc2d_config* cfg, lsc_hllc2d_ctx** out) {
    if (!cfg || !out) return lsc_fail(LSC_EINVAL, "lsc_hllc2d_create: null argument");
    *out = nullptr;
    if (cfg->nx < 3 || cfg->ny < 3) return lsc_fail(LSC_EINVAL, "nx, ny must be >= 3 (got %d x %d)", cfg->nx, cfg->ny);
    if (cfg->nx > 65535) return lsc_fail(LSC_EINVAL, "nx must be <= 65535");
    if (!(cfg->x1 > cfg->x0) || !(cfg->y1 > cfg->y0)) return lsc_fail(LSC_EINVAL, "empty domain");
    if (!(cfg->gamma > 1.0) || !(cfg->cfl > 0.0)) return lsc_fail(LSC_EINVAL, "gamma must be > 1 and cfl > 0");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return lsc_fail(LSC_ECUDA, "no CUDA device: this library has no CPU path");
    }
    if (cfg->device < 0 || cfg->device >= ndev) return lsc_fail(LSC_EINVAL, "device %d out of range", cfg->device);
    LSC_CUDA_TRY(cudaSetDevice(cfg->device));
    lsc_hllc2d_ctx* c = new (std::nothrow) lsc_hllc2d_ctx();
    if (!c) return lsc_fail(LSC_ENOMEM, "host allocation failed");
    c->cfg = *cfg;
    c->g.nx = cfg->nx;
    c->g.ny = cfg->ny;
    c->g.plane = (size_t)cfg->nx * (size_t)cfg->ny;
    c->dx = (cfg->x1 - cfg->x0) / cfg->nx;  // H:626-627
    c->dy = (cfg->y1 - cfg->y0) / cfg->ny;
    const size_t bytes = c->g.plane * NQ * sizeof(double);
    const size_t aos_bytes = (size_t)(cfg->nx + 2) * (size_t)(cfg->ny + 2) * NQ * sizeof(double);
    bool ok = true;
    for (int b = 0; b < 3 && ok; ++b) ok = lsc_dev_malloc(&c->u[b], bytes) == cudaSuccess;
    ok = ok && lsc_dev_malloc(&c->d_aos, aos_bytes) == cudaSuccess &&
         lsc_dev_malloc(&c->d_status, sizeof(unsigned long long)) == cudaSuccess &&
         cudaHostAlloc(&c->h_status, sizeof(unsigned long long), cudaHostAllocDefault) == cudaSuccess &&
         cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) == cudaSuccess;
    if (!ok) {
        cudaGetLastError();
        lsc_hllc2d_destroy(c);
        return lsc_fail(LSC_ENOMEM, "device allocation failed (%zu bytes per state)", bytes);
    }
    *out = c;
    return LSC_OK;
}

int lsc_hllc2d_destroy(lsc_hllc2d_ctx* c) {
    if (!c) return LSC_OK;
    cudaSetDevice(c->cfg.device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    for (int b = 0; b < 3; ++b) lsc_dev_free(c->u[b]);
    lsc_dev_free(c->d_aos);
    lsc_dev_free(c->d_status);
    if (c->h_status) cudaFreeHost(c->h_status);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
    return LSC_OK;
}

// IC loop H:642-721 (Orszag-Tang, the committed block H:707-716; cell centres x0 + (i - 0.5) dx over the array
// with its ONE ghost layer; PrimitiveToConservative H:9-21), host libm
int lsc_hllc2d_initial_condition(const lsc_hllc2d_config* cfg, double* aos) {
    if (!cfg || !aos) return lsc_fail(LSC_EINVAL, "null argument");
    const double pi = 3.14159265358979323846;  // H:628
    const double gamma = cfg->gamma;
    const double dx = (cfg->x1 - cfg->x0) / cfg->nx, dy = (cfg->y1 - cfg->y0) / cfg->ny;
    for (int i = 0; i < cfg->nx + 2; ++i)
        for (int j = 0; j < cfg->ny + 2; ++j) {
            const double x = cfg->x0 + (i - 0.5) * dx;
            const double y = cfg->y0 + (j - 0.5) * dy;
            double p[NQ];
            p[0] = gamma * gamma;
            p[1] = -std::sin(2.0 * pi * y);
            p[2] = std::sin(2.0 * pi * x);
            p[3] = 0.0;
            p[4] = gamma;
            p[5] = -std::sin(2.0 * pi * y);
            p[6] = std::sin(4.0 * pi * x);
            p[7] = 0.0;
            p[8] = 0;
            double* v = cell(aos, cfg->ny, i, j);
            v[0] = p[0];
            v[1] = p[0] * p[1];
            v[2] = p[0] * p[2];
            v[3] = p[0] * p[3];
            v[5] = p[5];
            v[6] = p[6];
            v[7] = p[7];
            v[4] = p[4] / (gamma - 1) + 0.5 * p[0] * (p[1] * p[1] + p[2] * p[2] + p[3] * p[3]) +
                   0.5 * (p[5] * p[5] + p[6] * p[6] + p[7] * p[7]);
            v[8] = p[8];
        }
    return LSC_OK;
}

int lsc_hllc2d_upload(lsc_hllc2d_ctx* c, const double* aos) {
    if (!c || !aos) return lsc_fail(LSC_EINVAL, "null argument");
    LSC_CUDA_TRY(cudaSetDevice(c->cfg.device));
    const size_t aos_bytes = (size_t)(c->cfg.nx + 2) * (size_t)(c->cfg.ny + 2) * NQ * sizeof(double);
    LSC_CUDA_TRY(cudaMemcpyAsync(c->d_aos, aos, aos_bytes, cudaMemcpyHostToDevice, c->stream));
    lsc::h2d::aos_to_planes<<<grid_of(c->g), 128, 0, c->stream>>>(c->g, c->d_aos, c->u[c->cur]);
    // ComputeTimeStep of the first pass (H:736) sees the array as given; its range holds interior cells only
    LSC_CUDA_TRY(cudaMemsetAsync(c->d_status, 0, sizeof(unsigned long long), c->stream));
    lsc::h2d::max_speed_kernel_h<<<grid_of(c->g), 128, 0, c->stream>>>(c->g, c->u[c->cur], c->cfg.gamma, c->d_status);
    lsc_count_launches(2);
    LSC_CUDA_TRY(cudaGetLastError());
    LSC_CUDA_TRY(cudaStreamSynchronize(c->stream));
    c->t = c->dt = c->ch = 0;
    c->step = 0;
    c->uploaded = true;
    return LSC_OK;
}

int lsc_hllc2d_init_problem(lsc_hllc2d_ctx* c) {
    if (!c) return lsc_fail(LSC_EINVAL, "null argument");
    std::vector<double> host;
    try {
        host.resize((size_t)(c->cfg.nx + 2) * (size_t)(c->cfg.ny + 2) * NQ);
    } catch (...) {
        return lsc_fail(LSC_ENOMEM, "host allocation failed");
    }
    lsc_hllc2d_initial_condition(&c->cfg, host.data());
    return lsc_hllc2d_upload(c, host.data());
}

// one pass of the loop body H:728-786
static int h2d_step_once(lsc_hllc2d_ctx* c) {
    LSC_CUDA_TRY(cudaMemcpyAsync(c->h_status, c->d_status, sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
    LSC_CUDA_TRY(cudaStreamSynchronize(c->stream));
    double ms;
    std::memcpy(&ms, c->h_status, sizeof ms);
    if (!(ms > 0.0) || !std::isfinite(ms)) return lsc_fail(LSC_ENAN, "max signal speed is %g at step %ld", ms, c->step);
    const double dspace = std::min(c->dx, c->dy);   // H:117
    const double dt = c->cfg.cfl * dspace / ms;     // H:118
    c->dt = dt;
    c->ch = ms;
    c->t = c->t + dt;  // H:737
    const double psi_factor = std::exp(-1.0 * (0.5 * dt) * ms / 0.18);  // psiUpdate(psi, maxSpeed, 0.5*dt), H:36-39, H:778
    const int bar = (c->cur + 1) % 3, next = (c->cur + 2) % 3;
    LSC_CUDA_TRY(cudaMemsetAsync(c->d_status, 0, sizeof(unsigned long long), c->stream));
    const dim3 gx((c->g.ny + 127) / 128, (c->g.nx + lsc::h2d::H2D_STRIP - 1) / lsc::h2d::H2D_STRIP);
    lsc::h2d::x_update_kernel<<<gx, 128, 0, c->stream>>>(c->g, c->u[c->cur], c->u[bar], c->cfg.gamma, ms, dt / c->dx);
    const dim3 gy((c->g.ny + lsc::h2d::H2D_BJ - 2) / (lsc::h2d::H2D_BJ - 1), c->g.nx);
    lsc::h2d::y_update_kernel<<<gy, lsc::h2d::H2D_BJ, 0, c->stream>>>(c->g, c->u[bar], c->u[next], c->cfg.gamma, ms,
                                                                       dt / c->dy, psi_factor, c->d_status);
    lsc_count_launches(2);
    LSC_CUDA_TRY(cudaGetLastError());
    c->cur = next;
    c->step += 1;
    return LSC_OK;
}

int lsc_hllc2d_step(lsc_hllc2d_ctx* c, long nsteps) {
    if (!c) return lsc_fail(LSC_EINVAL, "null argument");
    if (!c->uploaded) return lsc_fail(LSC_ESTATE, "lsc_hllc2d_step before lsc_hllc2d_upload");
    LSC_CUDA_TRY(cudaSetDevice(c->cfg.device));
    for (long s = 0; s < nsteps; ++s) {
        int rc = h2d_step_once(c);
        if (rc) return rc;
    }
    LSC_CUDA_TRY(cudaStreamSynchronize(c->stream));
    return LSC_OK;
}

int lsc_hllc2d_run_until(lsc_hllc2d_ctx* c, double t_stop, lsc_mhd_log_cb log_cb, void* user, long* steps_out) {
    if (!c) return lsc_fail(LSC_EINVAL, "null argument");
    if (!c->uploaded) return lsc_fail(LSC_ESTATE, "lsc_hllc2d_run_until before lsc_hllc2d_upload");
    LSC_CUDA_TRY(cudaSetDevice(c->cfg.device));
    long n = 0;
    do {  // H:728
        int rc = h2d_step_once(c);
        if (rc) return rc;
        ++n;
        if (log_cb && log_cb(c->step, c->t, c->dt, c->ch, user))  // H:738
            return lsc_fail(LSC_ECALLBACK, "log callback stopped the loop at step %ld", c->step);
    } while (c->t < t_stop);  // H:785
    LSC_CUDA_TRY(cudaStreamSynchronize(c->stream));
    if (steps_out) *steps_out = n;
    return LSC_OK;
}

int lsc_hllc2d_get_scalars(const lsc_hllc2d_ctx* c, double* t, double* dt, double* ch, long* step) {
    if (!c) return lsc_fail(LSC_EINVAL, "null argument");
    if (t) *t = c->t;
    if (dt) *dt = c->dt;
    if (ch) *ch = c->ch;
    if (step) *step = c->step;
    return LSC_OK;
}

// u as H holds it after the applyBoundaryConditions behind the loop (H:787): interior + periodic ghost layer
int lsc_hllc2d_download(lsc_hllc2d_ctx* c, double* aos) {
    if (!c || !aos) return lsc_fail(LSC_EINVAL, "null argument");
    if (!c->uploaded) return lsc_fail(LSC_ESTATE, "lsc_hllc2d_download before lsc_hllc2d_upload");
    LSC_CUDA_TRY(cudaSetDevice(c->cfg.device));
    const size_t aos_bytes = (size_t)(c->cfg.nx + 2) * (size_t)(c->cfg.ny + 2) * NQ * sizeof(double);
    lsc::h2d::planes_to_aos<<<grid_of(c->g), 128, 0, c->stream>>>(c->g, c->u[c->cur], c->d_aos);
    lsc_count_launches(1);
    LSC_CUDA_TRY(cudaGetLastError());
    LSC_CUDA_TRY(cudaMemcpyAsync(aos, c->d_aos, aos_bytes, cudaMemcpyDeviceToHost, c->stream));
    LSC_CUDA_TRY(cudaStreamSynchronize(c->stream));
    host_bc(c->cfg, aos);
    return LSC_OK;
}

// writer H:793-814: ConservativeToPrimitive of rows j = 1..ny+1, columns i = 1..nx+1, std::endl per line
int lsc_hllc2d_write_dat(const lsc_hllc2d_config* cfg, const double* aos, const char* path) {
    if (!cfg || !aos || !path) return lsc_fail(LSC_EINVAL, "null argument");
    std::ofstream output(path);
    if (!output) return lsc_fail(LSC_EINVAL, "cannot write %s", path);
    const double dx = (cfg->x1 - cfg->x0) / cfg->nx, dy = (cfg->y1 - cfg->y0) / cfg->ny;
    const double gamma = cfg->gamma;
    for (int j = 1; j < cfg->ny + 2; j++) {
        for (int i = 1; i < cfg->nx + 2; i++) {
            const double* u = aos + ((size_t)i * (size_t)(cfg->ny + 2) + (size_t)j) * NQ;
            double v[NQ];
            v[0] = u[0];
            v[1] = u[1] / u[0];
            v[2] = u[2] / u[0];
            v[3] = u[3] / u[0];
            v[5] = u[5];
            v[6] = u[6];
            v[7] = u[7];
            v[4] = (u[4] - 0.5 * v[0] * (v[1] * v[1] + v[2] * v[2] + v[3] * v[3]) -
                    0.5 * (u[5] * u[5] + u[6] * u[6] + u[7] * u[7])) * (gamma - 1);
            v[8] = u[8];
            const double x = cfg->x0 + (i - 1) * dx;
            const double y = cfg->y0 + (j - 1) * dy;
            output << x << " " << y << " " << v[0] << " " << v[1] << " " << v[2] << " " << v[3] << " " << v[4] << " "
                   << std::sqrt(v[5] * v[5] + v[6] * v[6]) / v[7] << " " << v[6] << " " << v[7] << " " << v[8]
                   << std::endl;
        }
        output << std::endl;
    }
    return output.good() ? LSC_OK : lsc_fail(LSC_EINVAL, "write to %s failed", path);
}

}  // extern "C"

// euler_kernels.cu -- sm_100a kernels of the 2-D Euler SLIC step (Euler_codes/SLIC_euler_2d.cpp = S).
//
//   euler_x_sweep_kernel : u -> ubar, the x half of XthenY (S:256-314)                  32 B read + 32 B write
//   euler_y_sweep_kernel : A = y-sweep(ubar) (S:317-373), B = y-sweep(u) (S:405-461; the x half of YthenX is
//                          dead in the reference: its flux goes to fluxY (S:510) while the update reads the
//                          zero fluxX (S:519)), u' = 0.5*(A+B) (S:609-615), max signal speed of u' (S:165-195)
//                                                                                        64 B read + 32 B write
// Same structure as the MHD kernels (mhd_kernels.cu): x marches along i with the stencil in registers, y is one
// thread per cell with face states and fluxes exchanged through shared memory; division policy as there.
#include <algorithm>
#include <type_traits>

#include "cfl_reduce.cuh"
#include "euler_kernels.cuh"
#include "euler_physics.cuh"

// production division policy of these kernels (fp64_div.cuh); overridable for experiments
#ifndef EULER_FAST_POLICY
#define EULER_FAST_POLICY FastDiv
#endif

namespace lsc {

// per-context step scalars (see mhd_kernels.cu: c_step): uniform-indexed, loaded into uniform registers
static __constant__ EulerStepScalars c_estep[STEP_SLOTS];

__device__ __forceinline__ EulerParams with_step_scalars(const EulerParams& a) {
    const EulerStepScalars& sc = c_estep[a.slot];
    EulerParams p = a;
    p.dtdx = sc.dtdx;
    p.hdtdx = sc.hdtdx;
    p.dtdy = sc.dtdy;
    p.hdtdy = sc.hdtdy;
    p.hdxdt = sc.hdxdt;
    p.hdydt = sc.hdydt;
    return p;
}

cudaError_t publish_euler_step_scalars(const EulerStepCtl* ctl, int slot, cudaStream_t s) {
    return cudaMemcpyToSymbolAsync(c_estep, &ctl->sc, sizeof(EulerStepScalars), (size_t)slot * sizeof(EulerStepScalars),
                                   cudaMemcpyDeviceToDevice, s);
}

// ComputeTimeStep's scalars and the loop predicate of S:600-617 / G:627-646 by one thread (cf. prepare_kernel in
// mhd_kernels.cu; the Euler loop has no exp() and no dump cadence).
__global__ void euler_prepare_kernel(EulerStepCtl* __restrict__ ctl, const StepCmd* __restrict__ cmd,
                                     unsigned long long* __restrict__ status, double* __restrict__ log_ring) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    EulerStepCtl& c = *ctl;
    bool fresh = false;
    if (c.batch_seen != cmd->batch) {
        c.batch_seen = cmd->batch;
        c.halt = HALT_NONE;
        c.log_count = 0;
        fresh = true;
    }
    if (c.halt != HALT_NONE) return;
    int halt = HALT_NONE;
    if (!fresh && !(c.t < cmd->t_stop)) halt = HALT_PAUSE;  // S:617 / G:646
    else if (c.step >= cmd->step_limit) halt = HALT_LIMIT;
    else if (cmd->keep_log && c.log_count >= STEP_LOG_CAP) halt = HALT_LOG_FULL;
    if (halt == HALT_NONE) {
        const double ms = __longlong_as_double((long long)status[0]);
        if (!(ms > 0.0) || !(ms <= 1.7976931348623157e308)) {
            c.max_speed = ms;
            halt = HALT_NAN;
        } else {
            const double dt = c.cfl * c.dspace / ms;  // S:192
            c.max_speed = ms;
            c.dt = dt;
            c.t += dt;  // S:602
            c.step += 1;
            EulerStepScalars& sc = c.sc;
            sc.dtdx = dt / c.dx;
            sc.hdtdx = 0.5 * (dt / c.dx);
            sc.dtdy = dt / c.dy;
            sc.hdtdy = 0.5 * (dt / c.dy);
            sc.hdxdt = 0.5 * c.dx / dt;
            sc.hdydt = 0.5 * c.dy / dt;
            c.redo_total += status[1] + status[2] + status[3];
            status[0] = status[1] = status[2] = status[3] = 0ull;
            if (cmd->keep_log) {
                double* rec = log_ring + 3 * c.log_count;
                rec[0] = c.t;
                rec[1] = dt;
                rec[2] = ms;
                c.log_count += 1;
            }
        }
    }
    c.halt = halt;
    c.sc.halt = halt;
}

void launch_euler_prepare(EulerStepCtl* ctl, const StepCmd* cmd, unsigned long long* status, double* log_ring,
                          cudaStream_t s) {
    euler_prepare_kernel<<<1, 32, 0, s>>>(ctl, cmd, status, log_ring);
}

__host__ __device__ inline size_t euler_index(const GridDims& g, int i, int j) { return blocked_index<NE>(g, i, j); }

__device__ __forceinline__ void eload(const double* __restrict__ base, size_t idx, double (&u)[NE]) {
#pragma unroll
    for (int k = 0; k < NE; ++k) u[k] = __ldg(base + idx + k * KSTR);
}

// ---------------------------------------------------------------------------------------------------------
// x sweep.  Ghost half states are boundary images of interior half states (S:297-298): periodic ones equal a
// reconstruction from the ghost columns; transmissive (G:199-215) ones copy the first / last interior cell.
// ---------------------------------------------------------------------------------------------------------
template <class D, class EOS>
__device__ __noinline__ D euler_x_halfstates_global(const EOS eos, const GridDims g, const EulerParams p,
                                                       const double* __restrict__ cur, int i, int j,
                                                       double (&hL)[NE], double (&hR)[NE]) {
    D dv;  // own guard state, handed back by value (see FastDiv::absorb)
    double um[NE], u0[NE], up[NE];
    eload(cur, euler_index(g, i - 1, j), um);
    eload(cur, euler_index(g, i, j), u0);
    eload(cur, euler_index(g, i + 1, j), up);
    euler_halfstates<0>(dv, um, u0, up, eos, p.cm, p.cp, p.hdtdx, hL, hR);
    return dv;
}

// EoS policy object of a launch: ideal gas (S / G) or the table of T (EulerParams::tab_*)
template <class EOS>
__device__ __forceinline__ EOS make_eos(const EulerParams& p);
template <>
__device__ __forceinline__ IdealEos make_eos<IdealEos>(const EulerParams& p) {
    return IdealEos{p.gm1, p.gamma};
}
template <>
__device__ __forceinline__ TableEos make_eos<TableEos>(const EulerParams& p) {
    return TableEos{p.tab_dens, p.tab_pres, p.tab_sound, p.tab_en, p.tab_warm};
}

#ifndef EULER_X_UNROLL
#define EULER_X_UNROLL 2
#endif
constexpr int EX_UNROLL = EULER_X_UNROLL;  // copies of the march-loop body: the 20 carried doubles are renamed, not moved
constexpr long EFIX_GRID = 148 * 4;  // persistent grid of the fix kernels (see mhd_kernels.cu)

template <class D, class EOS>
__device__ __forceinline__ bool euler_x_sweep_block(const GridDims& g, const EulerParams& p,
                                                    const double* __restrict__ cur, double* __restrict__ ubar,
                                                    const int bx, const int by, const int strip) {
    const int j = 2 + bx * blockDim.x + threadIdx.x;
    if (j >= g.ny + 2) return false;
    const int ia = 2 + by * strip;
    const int ib = min(ia + strip, g.nxl + 2);
    const bool lo_image = (ia == 2) && (p.bcx_lo == BC_TRANSMISSIVE_E);
    const bool hi_image = (ib == g.nxl + 2) && (p.bcx_hi == BC_TRANSMISSIVE_E);
    D dv;
    const EOS eos = make_eos<EOS>(p);

    double hRprev[NE], hLprev[NE], Fprev[NE];
    {
        double tL[NE], tR[NE];
        dv.absorb(euler_x_halfstates_global<D>(eos, g, p, cur, lo_image ? 2 : ia - 1, j, tL, tR));
#pragma unroll
        for (int k = 0; k < NE; ++k) {
            hRprev[k] = tR[k];
            hLprev[k] = tL[k];
            Fprev[k] = 0.0;
        }
    }
    double uM[NE], u0[NE], uP[NE];
    eload(cur, euler_index(g, ia - 1, j), uM);
    eload(cur, euler_index(g, ia, j), u0);
    eload(cur, euler_index(g, ia + 1, j), uP);

#pragma unroll EX_UNROLL
    for (int c = ia; c <= ib; ++c) {
        // the column loaded at the end of this iteration is consumed at the start of the next one: pull the one
        // after it into L1 now, a whole iteration ahead
        if (c + 3 <= g.nxl + 3) {
            const double* pf = cur + euler_index(g, c + 3, j);
#pragma unroll
            for (int k = 0; k < NE; ++k) asm volatile("prefetch.global.L1 [%0];" ::"l"(pf + k * KSTR));
        }
        double hL[NE], hR[NE];
        if (c == ib && hi_image) {
#pragma unroll
            for (int k = 0; k < NE; ++k) {  // u[n+2] = u[n+1] applied to the half-step arrays
                hL[k] = hLprev[k];
                hR[k] = hRprev[k];
            }
        } else {
            euler_halfstates<0>(dv, uM, u0, uP, eos, p.cm, p.cp, p.hdtdx, hL, hR);
        }
        double F[NE];
        euler_force<0>(dv, hRprev, hL, eos, p.hdtdx, p.hdxdt, F);  // interface c-1/2, S:302
        if (c > ia) {
            const size_t o = euler_index(g, c - 1, j);
#pragma unroll
            for (int k = 0; k < NE; ++k) ubar[o + k * KSTR] = uM[k] - p.dtdx * (F[k] - Fprev[k]);  // S:309
        }
#pragma unroll
        for (int k = 0; k < NE; ++k) {
            Fprev[k] = F[k];
            hRprev[k] = hR[k];
            hLprev[k] = hL[k];
            uM[k] = u0[k];
            u0[k] = uP[k];
        }
        if (c < ib) eload(cur, euler_index(g, c + 2, j), uP);
    }
    return dv.bad();
}

template <class D, class EOS>
__global__ void __launch_bounds__(128, 4) euler_x_sweep_kernel(const GridDims g, const EulerParams pin,
                                                               const double* __restrict__ cur,
                                                               double* __restrict__ ubar,
                                                               unsigned long long* __restrict__ status,
                                                               unsigned* __restrict__ fix_list, const int strip) {
    if (c_estep[pin.slot].halt) return;
    const EulerParams p = with_step_scalars(pin);
    const bool bad = euler_x_sweep_block<D, EOS>(g, p, cur, ubar, blockIdx.x, blockIdx.y, strip);
    if (__syncthreads_or(bad) && threadIdx.x == 0)
        fix_list[atomicAdd(status + 1, 1ull)] = blockIdx.y * gridDim.x + blockIdx.x;
}

template <class EOS>
__global__ void __launch_bounds__(128) euler_x_sweep_fix_kernel(const GridDims g, const EulerParams pin,
                                                                const double* __restrict__ cur,
                                                                double* __restrict__ ubar,
                                                                const unsigned long long* __restrict__ status,
                                                                const unsigned* __restrict__ fix_list,
                                                                const int strip, const unsigned nbx) {
    if (c_estep[pin.slot].halt) return;
    const EulerParams p = with_step_scalars(pin);
    const unsigned long long count = status[1];
    for (unsigned long long q = blockIdx.x; q < count; q += gridDim.x) {
        const unsigned bid = fix_list[q];
        euler_x_sweep_block<IeeeDiv, EOS>(g, p, cur, ubar, (int)(bid % nbx), (int)(bid / nbx), strip);
    }
}

void launch_euler_x_sweep(const GridDims& g, const EulerParams& p, const double* cur, double* ubar,
                          unsigned long long* status, unsigned* fix_list, cudaStream_t s) {
    dim3 block(128);
    const int row_blocks = (g.ny + block.x - 1) / block.x;
    const int strip = choose_x_strip(row_blocks, g.nxl, 64, 148 * 4);
    dim3 grid(row_blocks, (g.nxl + strip - 1) / strip);
    const unsigned nfix = (unsigned)std::min<long>(EFIX_GRID, (long)grid.x * grid.y);
    if (p.tab_dens) {  // tabulated EoS (T)
        euler_x_sweep_kernel<EULER_FAST_POLICY, TableEos><<<grid, block, 0, s>>>(g, p, cur, ubar, status, fix_list, strip);
        euler_x_sweep_fix_kernel<TableEos><<<nfix, block, 0, s>>>(g, p, cur, ubar, status, fix_list, strip, grid.x);
    } else {
        euler_x_sweep_kernel<EULER_FAST_POLICY, IdealEos><<<grid, block, 0, s>>>(g, p, cur, ubar, status, fix_list, strip);
        euler_x_sweep_fix_kernel<IdealEos><<<nfix, block, 0, s>>>(g, p, cur, ubar, status, fix_list, strip, grid.x);
    }
}

long euler_edge_list_entries(const GridDims& g) { return (long)((g.ny + 125) / 126) * 4; }

long euler_fix_list_entries(const GridDims& g) {
    const long xb = (long)((g.ny + 127) / 128) * ((g.nxl + 7) / 8);
    const long yb = (long)((g.ny + 125) / 126) * g.nxl;
    return std::max(xb, yb);
}

// ---------------------------------------------------------------------------------------------------------
// y sweep (both of them) + average + CFL maximum.
// ---------------------------------------------------------------------------------------------------------
constexpr int EBJ = 128;
#ifndef EULER_YS_PREFETCH
#define EULER_YS_PREFETCH 1  // 0: no L2 prefetch of the successor block's tile
#endif

long euler_y_block_count(const GridDims& g) { return (long)((g.ny + (EBJ - 2) - 1) / (EBJ - 2)) * g.nxl; }

// stored row that a (possibly ghost) row index refers to: periodic S:217-225, transmissive G:208-214
__device__ __forceinline__ int emap_row(int j, int ny, int bc) {
    if (j >= 2 && j <= ny + 1) return j;
    if (bc == BC_PERIODIC) return (j < 2) ? j + ny : j - ny;
    return (j < 2) ? 2 : ny + 1;
}

// One y sweep of `in` for this thread's cell: returns the updated value of cell c in `out` (valid for
// tid in [1, EBJ-2], c in [2, ny+1]).  sH / sF: shared exchange buffers [NE][EBJ].
template <class D, class EOS>
__device__ __forceinline__ void euler_y_pass(D& dv, const EOS& eos, const GridDims& g, const EulerParams& p,
                                             const double* __restrict__ in, int i, int c, int tid,
                                             double (*sH)[EBJ], double (*sF)[EBJ], double (&out)[NE]) {
    const int ny = g.ny;
    const bool active = (c <= ny + 2);
    double hB[NE], hT[NE], u0[NE];  // written and read only by active threads
    if (active) {
        // ghost cells carry the half states of their boundary image: periodic -> same stencil through the wrap,
        // transmissive -> the first / last interior cell (S:356-357 with G:208-214)
        int s = c;
        if (p.bcy != BC_PERIODIC) s = min(max(c, 2), ny + 1);
        double um[NE], up[NE];
        eload(in, euler_index(g, i, emap_row(s - 1, ny, p.bcy)), um);
        eload(in, euler_index(g, i, emap_row(s, ny, p.bcy)), u0);
        eload(in, euler_index(g, i, emap_row(s + 1, ny, p.bcy)), up);
        euler_halfstates<1>(dv, um, u0, up, eos, p.cm, p.cp, p.hdtdy, hB, hT);
#pragma unroll
        for (int k = 0; k < NE; ++k) sH[k][tid] = hB[k];
    }
    __syncthreads();
    const bool has_flux = (tid < EBJ - 1) && (c <= ny + 1);
    double F[NE];
    if (has_flux) {
        double nb[NE];
#pragma unroll
        for (int k = 0; k < NE; ++k) nb[k] = sH[k][tid + 1];
        euler_force<1>(dv, hT, nb, eos, p.hdtdy, p.hdydt, F);  // getFluxY(top of c, bottom of c+1), S:361
#pragma unroll
        for (int k = 0; k < NE; ++k) sF[k][tid] = F[k];
    }
    __syncthreads();
    if (tid >= 1 && has_flux) {
#pragma unroll
        for (int k = 0; k < NE; ++k) out[k] = u0[k] - p.dtdy * (F[k] - sF[k][tid - 1]);  // S:370
    }
}

template <class D, class EOS>
__device__ __forceinline__ bool euler_y_sweep_block(const GridDims& g, const EulerParams& p,
                                                    const double* __restrict__ ubar, const double* __restrict__ cur,
                                                    double* __restrict__ next, const int bx, const int by,
                                                    double& speed, const unsigned long long* status,
                                                    const int pf_cols = 0) {
    // fast instantiation with an EoS that has a cheap bound: fp32 filter of the CFL reduction (see the MHD y
    // sweep): the running exact maximum of the blocks that already finished, read first to hide its latency
    constexpr bool FILTER = !std::is_same<D, IeeeDiv>::value && EOS::kHasCflBound;
    float inv_run_max = 0.0f;
    unsigned long long run_bits = 0;  // loaded first thing, converted where the filter needs it (cfl_reduce.cuh)
    if (FILTER) run_bits = running_max_bits(status);
    __shared__ double sH[2][NE][EBJ];  // one pair of exchange buffers per pass: no barrier between the passes
    __shared__ double sF[2][NE][EBJ];
    const int tid = threadIdx.x;
    const int i = 2 + by;
    const int jo = 2 + bx * (EBJ - 2);
    const int c = jo - 1 + tid;
#if EULER_YS_PREFETCH
    if (!std::is_same<D, IeeeDiv>::value) {
        // as in the MHD y sweep: ask L2 for the tile (both states; one contiguous run each in the blocked layout) of
        // the block one resident wave further on, which will take this block's place on the SM
        if (pf_cols > 0 && tid == EBJ - 1) {
            const int ip = i + pf_cols;  // pf_cols: one resident wave of blocks in columns, from the launch (YCols)
            if (ip <= g.nxl + 1) {
                const int g0 = (JOFF + jo - 2) >> 5;
                const int g1 = (JOFF + min(jo + EBJ - 1, g.ny + 3)) >> 5;
                const size_t off = (size_t)ip * column_doubles(g, NE) + (size_t)g0 * (NE * JBLK);
                const unsigned bytes = (unsigned)(g1 - g0 + 1) * (NE * JBLK * 8);
                asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(ubar + off), "r"(bytes) : "memory");
                asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(cur + off), "r"(bytes) : "memory");
            }
        }
    }
#endif
    D dv;
    const EOS eos = make_eos<EOS>(p);
    const int nbx = (g.ny + (EBJ - 2) - 1) / (EBJ - 2);  // blocks per column (also in the fix kernel's grid)
    int* const block_hints = p.tab_hints ? p.tab_hints + 3 * ((size_t)by * nbx + bx) : nullptr;
    eos.load_hints(block_hints);
    double A[NE], B[NE];  // valid for the threads that update a cell (the only ones that read them)
    euler_y_pass(dv, eos, g, p, ubar, i, c, tid, sH[0], sF[0], A);  // XthenY: y sweep of ubar
    euler_y_pass(dv, eos, g, p, cur, i, c, tid, sH[1], sF[1], B);   // YthenX: y sweep of u (its x sweep is a no-op)

    speed = 0.0;
    const bool updates = (tid >= 1 && tid < EBJ - 1 && c <= g.ny + 1);
    double un[NE];
    if (updates) {
        const size_t o = euler_index(g, i, c);
#pragma unroll
        for (int k = 0; k < NE; ++k) {
            // uPlus1 = uPlus1Bar - (dt/dx)*(fluxX - fluxX) with fluxX == 0 (S:519), then the average S:612
            const double b = B[k] - p.dtdx * (0.0 - 0.0);
            un[k] = 0.5 * (A[k] + b);
            next[o + k * KSTR] = un[k];
        }
    }
    // Only the global maximum of the signal speed is consumed: a warp evaluates the exact fp64 expression only if
    // one of its cells is not provably (fp32 upper bound) below the running exact maximum.  Skipped cells are
    // below the final maximum, so the result is the same bit pattern whatever the block order.
    bool evaluate = updates;
    if (FILTER) {
        inv_run_max = inv_running_max_from_bits(run_bits);
        bool candidate = false;
        if (updates) candidate = !(eos.cfl_speed_over_max_f32(un, inv_run_max) < 1.0f);
        const bool warp_has_candidate = __any_sync(0xffffffffu, candidate);  // every lane votes (no short circuit)
        evaluate = updates && warp_has_candidate;
    }
    if (evaluate) {
        const double sp = eos.cfl_speed(dv, un);
        if (sp > speed) speed = sp;
    }
    if (tid == EBJ / 2 && updates) eos.store_hints(block_hints);
    return dv.bad();
}

// block maximum of `speed` -> status[0]
__device__ __forceinline__ void euler_block_max(double speed, unsigned long long* status) {
    __shared__ double warp_max[32];
    const int tid = threadIdx.x;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double other = __shfl_xor_sync(0xffffffffu, speed, o);
        speed = (other > speed) ? other : speed;
    }
    if ((tid & 31) == 0) warp_max[tid >> 5] = speed;
    __syncthreads();
    if (tid < 32) {
        double v = (tid < EBJ / 32) ? warp_max[tid] : 0.0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double other = __shfl_xor_sync(0xffffffffu, v, o);
            v = (other > v) ? other : v;
        }
        if (tid == 0 && v > 0.0) atomicMax(status, (unsigned long long)__double_as_longlong(v));
    }
}

template <class D, class EOS>
__global__ void __launch_bounds__(EBJ, 4) euler_y_sweep_kernel(const GridDims g, const EulerParams pin,
                                                               const double* __restrict__ ubar,
                                                               const double* __restrict__ cur,
                                                               double* __restrict__ next,
                                                               unsigned long long* __restrict__ status,
                                                               unsigned* __restrict__ fix_list, const YCols yc) {
    if (c_estep[pin.slot].halt) return;
    const EulerParams p = with_step_scalars(pin);
    double speed;
    // column of this block row: all columns, the four slab-edge columns or the rest (YCols, mhd_kernels.cuh)
    const int col = yc.c0 + (int)blockIdx.y + ((int)blockIdx.y >= yc.jump_at ? yc.jump : 0);
    const bool bad = euler_y_sweep_block<D, EOS>(g, p, ubar, cur, next, blockIdx.x, col, speed, status, yc.pf_cols);
    if (__syncthreads_or(bad)) {  // whole block is recomputed by the fix kernel; its maximum must not count
        if (threadIdx.x == 0) fix_list[atomicAdd(status + yc.counter, 1ull)] = col * gridDim.x + blockIdx.x;
    } else if (!std::is_same<D, IeeeDiv>::value && EOS::kHasCflBound) {
        warp_max_to_global(speed, status);
    } else {
        euler_block_max(speed, status);
    }
}

template <class EOS>
__global__ void __launch_bounds__(EBJ) euler_y_sweep_fix_kernel(const GridDims g, const EulerParams pin,
                                                                const double* __restrict__ ubar,
                                                                const double* __restrict__ cur,
                                                                double* __restrict__ next,
                                                                unsigned long long* __restrict__ status,
                                                                const unsigned* __restrict__ fix_list,
                                                                const unsigned nbx, const int counter) {
    if (c_estep[pin.slot].halt) return;
    const EulerParams p = with_step_scalars(pin);
    const unsigned long long count = status[counter];
    for (unsigned long long q = blockIdx.x; q < count; q += gridDim.x) {
        const unsigned bid = fix_list[q];
        double speed;
        euler_y_sweep_block<IeeeDiv, EOS>(g, p, ubar, cur, next, (int)(bid % nbx), (int)(bid / nbx), speed, status);
        euler_block_max(speed, status);
        __syncthreads();
    }
}

void launch_euler_y_sweep(const GridDims& g, const EulerParams& p, const double* ubar, const double* cur, double* next,
                          unsigned long long* status, unsigned* fix_list, cudaStream_t s, int part) {
    dim3 block(EBJ);
    YCols yc = y_columns(g, part);
    // the slab-edge launch keeps its own list (behind the euler_fix_list_entries of the others) and counter
    unsigned* list = fix_list + (part == Y_EDGES ? euler_fix_list_entries(g) : 0);
    dim3 grid((g.ny + (EBJ - 2) - 1) / (EBJ - 2), yc.ncols);
    yc.pf_cols = (148 * 4 + (int)grid.x - 1) / (int)grid.x;
    const unsigned nfix = (unsigned)std::min<long>(EFIX_GRID, (long)grid.x * grid.y);
    if (p.tab_dens) {
        euler_y_sweep_kernel<EULER_FAST_POLICY, TableEos><<<grid, block, 0, s>>>(g, p, ubar, cur, next, status, list, yc);
        euler_y_sweep_fix_kernel<TableEos><<<nfix, block, 0, s>>>(g, p, ubar, cur, next, status, list, grid.x, yc.counter);
    } else {
        euler_y_sweep_kernel<EULER_FAST_POLICY, IdealEos><<<grid, block, 0, s>>>(g, p, ubar, cur, next, status, list, yc);
        euler_y_sweep_fix_kernel<IdealEos><<<nfix, block, 0, s>>>(g, p, ubar, cur, next, status, list, grid.x, yc.counter);
    }
}

// Stand-alone CFL reduction over [i_lo,i_hi] x [j_lo,j_hi]: the reference's range is [2,n+2] on the array as
// it stands (S:173-174), ghost row/column n+2 included.
template <class EOS>
__global__ void euler_max_speed_kernel(const GridDims g, const EulerParams p, const double* __restrict__ u, int i_lo,
                                       int i_hi, int j_lo, int j_hi, unsigned long long* __restrict__ status) {
    const int j = j_lo + blockIdx.x * blockDim.x + threadIdx.x;
    double best = 0.0;
    if (j <= j_hi) {
        for (int i = i_lo + blockIdx.y; i <= i_hi; i += gridDim.y) {
            double q[NE];
            eload(u, euler_index(g, i, j), q);
            IeeeDiv dv;
            const double sp = make_eos<EOS>(p).cfl_speed(dv, q);
            if (sp > best) best = sp;
        }
    }
    if (best > 0.0) atomicMax(status, (unsigned long long)__double_as_longlong(best));
}

void launch_euler_max_speed(const GridDims& g, const EulerParams& p, const double* u, int i_lo, int i_hi, int j_lo, int j_hi,
                            unsigned long long* status, cudaStream_t s) {
    dim3 block(128);
    const int nj = j_hi - j_lo + 1, ni = i_hi - i_lo + 1;
    dim3 grid((nj + 127) / 128, ni < 64 ? ni : 64);
    if (p.tab_dens) euler_max_speed_kernel<TableEos><<<grid, block, 0, s>>>(g, p, u, i_lo, i_hi, j_lo, j_hi, status);
    else euler_max_speed_kernel<IdealEos><<<grid, block, 0, s>>>(g, p, u, i_lo, i_hi, j_lo, j_hi, status);
}

// ghost columns, S:205-210 (periodic) / G:203-208 (transmissive: 0,1 <- 2 ; n+2,n+3 <- n+1)
__global__ void euler_x_ghost_kernel(const GridDims g, int bc_x, double* __restrict__ u) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= g.ny + 4) return;
    double* pl = u + blockIdx.y * KSTR;
    const int nx = g.nxl;
    int src[4];
    if (bc_x == BC_PERIODIC) {
        src[0] = nx; src[1] = nx + 1; src[2] = 2; src[3] = 3;
    } else {
        src[0] = 2; src[1] = 2; src[2] = nx + 1; src[3] = nx + 1;
    }
    const int dst[4] = {0, 1, nx + 2, nx + 3};
#pragma unroll
    for (int q = 0; q < 4; ++q) pl[euler_index(g, dst[q], j)] = pl[euler_index(g, src[q], j)];
}

void launch_euler_x_ghosts(const GridDims& g, int bc_x, double* u, cudaStream_t s) {
    dim3 block(256), grid((g.ny + 4 + 255) / 256, NE);
    euler_x_ghost_kernel<<<grid, block, 0, s>>>(g, bc_x, u);
}

// transmissive image of ONE side (0 = low, 1 = high), G:203-208: end ranks of a slab decomposition
__global__ void euler_x_ghost_side_kernel(const GridDims g, int hi_side, double* __restrict__ u) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= g.ny + 4) return;
    double* pl = u + blockIdx.y * KSTR;
    const int nx = g.nxl;
    if (hi_side) {
        const double v = pl[euler_index(g, nx + 1, j)];
        pl[euler_index(g, nx + 2, j)] = v;
        pl[euler_index(g, nx + 3, j)] = v;
    } else {
        const double v = pl[euler_index(g, 2, j)];
        pl[euler_index(g, 0, j)] = v;
        pl[euler_index(g, 1, j)] = v;
    }
}

void launch_euler_x_ghosts_side(const GridDims& g, int hi_side, double* u, cudaStream_t s) {
    dim3 block(256), grid((g.ny + 4 + 255) / 256, NE);
    euler_x_ghost_side_kernel<<<grid, block, 0, s>>>(g, hi_side, u);
}

// reference array-of-structures [i][j][4] <-> blocked device layout for `ni` columns starting at i0
__global__ void euler_aos_to_soa_kernel(const GridDims g, const double* __restrict__ aos, int i0,
                                        double* __restrict__ soa) {
    __shared__ double tile[32 * NE];
    const int jb = blockIdx.x * 32, il = blockIdx.y, nrow = g.ny + 4;
    const int nj = min(32, nrow - jb);
    const double* src = aos + ((size_t)il * nrow + jb) * NE;
    for (int t = threadIdx.x; t < nj * NE; t += blockDim.x) tile[t] = src[t];
    __syncthreads();
    for (int t = threadIdx.x; t < nj * NE; t += blockDim.x) {
        const int k = t / nj, jj = t - k * nj;
        soa[euler_index(g, i0 + il, jb + jj) + k * KSTR] = tile[jj * NE + k];
    }
}

__global__ void euler_soa_to_aos_kernel(const GridDims g, const double* __restrict__ soa, int i0,
                                        double* __restrict__ aos) {
    __shared__ double tile[32 * NE];
    const int jb = blockIdx.x * 32, il = blockIdx.y, nrow = g.ny + 4;
    const int nj = min(32, nrow - jb);
    for (int t = threadIdx.x; t < nj * NE; t += blockDim.x) {
        const int k = t / nj, jj = t - k * nj;
        tile[jj * NE + k] = soa[euler_index(g, i0 + il, jb + jj) + k * KSTR];
    }
    __syncthreads();
    double* dst = aos + ((size_t)il * nrow + jb) * NE;
    for (int t = threadIdx.x; t < nj * NE; t += blockDim.x) dst[t] = tile[t];
}

void launch_euler_aos_to_soa(const GridDims& g, const double* aos, int i0, int ni, double* soa, cudaStream_t s) {
    dim3 grid((g.ny + 4 + 31) / 32, ni);
    euler_aos_to_soa_kernel<<<grid, 128, 0, s>>>(g, aos, i0, soa);
}

void launch_euler_soa_to_aos(const GridDims& g, const double* soa, int i0, int ni, double* aos, cudaStream_t s) {
    dim3 grid((g.ny + 4 + 31) / 32, ni);
    euler_soa_to_aos_kernel<<<grid, 128, 0, s>>>(g, soa, i0, aos);
}

// Test hook: exact signal speed (S:175-184) and the fp32 upper bound of speed/run_max of the y sweep's filter.
__global__ void euler_kat_cfl_bound_kernel(long n, const double* __restrict__ U, double gamma, double run_max,
                                           double* __restrict__ out) {
    const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    double u[NE];
#pragma unroll
    for (int k = 0; k < NE; ++k) u[k] = U[t * NE + k];
    const IdealEos eos{gamma - 1, gamma};
    IeeeDiv dv;
    out[2 * t] = eos.cfl_speed(dv, u);
    out[2 * t + 1] = (double)eos.cfl_speed_over_max_f32(u, __frcp_ru(__double2float_rd(run_max)));
}

void launch_euler_kat_cfl_bound(long n, const double* U, double gamma, double run_max, double* out) {
    euler_kat_cfl_bound_kernel<<<(unsigned)((n + 127) / 128), 128>>>(n, U, gamma, run_max, out);
}

}  // namespace lsc

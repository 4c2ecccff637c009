// mhd_kernels.cu -- sm_100a kernels of the dimensionally split MHD step.
//
//   x_sweep_kernel : u^n (+ u^{n-1} for the limiter) -> uBar      MUSCL_NEW.cpp:838-902
//   y_sweep_kernel : uBar (+ u^n for the limiter)    -> u^{n+1}   MUSCL_NEW.cpp:905-971, + max signal
//                    speed of u^{n+1} for the next step           MUSCL_NEW.cpp:71-115
//   x_ghost_kernel : ghost columns of a state                     MUSCL_NEW.cpp:609-622
//   max_speed_kernel : stand-alone CFL reduction (initial state)  MUSCL_NEW.cpp:71-115
//
// Compiled with -fmad=false: FMA contraction changes the answer at the 1e-9 level
// after one step and at the 1e-2 level after ten.
//
// Both sweeps are templated on the division policy (fp64_div.cuh).  The production instantiation is
// FastDiv: branch-free shared-reciprocal quotients with a sticky "operand outside the guarded window"
// flag.  A thread block in which any thread raised the flag appends its index to a device-side list
// (status[1] / status[2] count the entries); right after each sweep a small persistent "fix" kernel --
// the IeeeDiv instantiation of the same block function, every quotient by the compiler's IEEE `/` --
// recomputes exactly those blocks.  No host round trip, no branch per quotient, and the result is the
// correctly rounded one for any input (e.g. the denormal momenta at a numerical front).
#include <algorithm>
#include <cstdlib>
#include <type_traits>

#include "cfl_reduce.cuh"
#include "mhd_kernels.cuh"
#include "mhd_physics.cuh"

// production division policy of these kernels (fp64_div.cuh); overridable for experiments
#ifndef MHD_FAST_POLICY
#define MHD_FAST_POLICY FastDiv
#endif

namespace lsc {

// Per-context step scalars (step_ctl.cuh).  Indexed with a kernel parameter, i.e. a uniform value: ptxas loads the
// entry with LDCU into uniform registers once per kernel (checked in the SASS) -- no per-thread registers.
static __constant__ StepScalars c_step[STEP_SLOTS];

#define LSC_RETURN_IF_HALTED(slot) \
    if (c_step[slot].halt) return

template <int DIR>
__device__ __forceinline__ SweepParams make_params(const SweepArgs& a) {
    const StepScalars& sc = c_step[a.slot];
    SweepParams p;
    p.gamma = a.gamma;
    p.gm1 = a.gm1;
    p.first_step = a.first_step;
    p.bc_lo = a.bc_lo;
    p.bc_hi = a.bc_hi;
    p.ch = sc.ch;
    p.ch2 = sc.ch2;
    p.inv2ch = sc.inv2ch;
    p.halfch = sc.halfch;
    p.dtdh = DIR == 0 ? sc.dtdx : sc.dtdy;
    p.hdtdh = DIR == 0 ? sc.hdtdx : sc.hdtdy;
    p.psi_factor = sc.psi_factor;
    return p;
}

#ifndef LSC_RELAXED_TU
cudaError_t publish_step_scalars(const StepCtl* ctl, int slot, cudaStream_t s) {
    return cudaMemcpyToSymbolAsync(c_step, &ctl->sc, sizeof(StepScalars), (size_t)slot * sizeof(StepScalars),
                                   cudaMemcpyDeviceToDevice, s);
}
#else
cudaError_t publish_step_scalars_relaxed(const StepCtl* ctl, int slot, cudaStream_t s) {
    return cudaMemcpyToSymbolAsync(c_step, &ctl->sc, sizeof(StepScalars), (size_t)slot * sizeof(StepScalars),
                                   cudaMemcpyDeviceToDevice, s);
}
#endif

__device__ __forceinline__ void load_cell(const double* __restrict__ base, size_t idx, double (&u)[NV]) {
#pragma unroll
    for (int k = 0; k < NV; ++k) u[k] = __ldg(base + idx + k * MHD_KSTR);
}

// Reduce a per-thread running maximum (started from 0.0, updated only with `>` so NaNs
// never enter -- MUSCL_NEW.cpp:111) over the block and fold it into *out.
__device__ __forceinline__ void block_max_to_global(double local, unsigned long long* out) {
    __shared__ double warp_max[32];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double other = __shfl_xor_sync(0xffffffffu, local, o);
        local = (other > local) ? other : local;
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) warp_max[warp] = local;
    __syncthreads();
    if (warp == 0) {
        const int nwarps = (blockDim.x + 31) >> 5;
        double v = (lane < nwarps) ? warp_max[lane] : 0.0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double other = __shfl_xor_sync(0xffffffffu, v, o);
            v = (other > v) ? other : v;
        }
        if (lane == 0 && v > 0.0) atomic_max_nonneg(out, v);
    }
}

#ifndef XS_BLOCK
#define XS_BLOCK 128
#endif
#ifndef XS_MINB
#define XS_MINB 4
#endif
#ifndef XS_STRIP
#define XS_STRIP 128
#endif
#ifndef XS_MIN_STRIP
#define XS_MIN_STRIP 2  // shortest strip of an x-sweep thread (small, latency-bound grids; choose_x_strip)
#endif
#ifndef XS_PF
#define XS_PF 2  // the x march prefetches column c + XS_PF into L1 while it works on column c
#endif
#ifndef YS_MINB
#define YS_MINB 4
#endif
#ifndef YS_BJ
#define YS_BJ 128
#endif
#ifndef YS_TMA
#define YS_TMA 0  // 1: stage the y tile with cp.async.bulk + mbarrier (experiment J1, profiles/r2/experiments)
#endif
#ifndef YS_PF_BLOCKS
#define YS_PF_BLOCKS (148 * YS_MINB)  // y sweep: L2 prefetch distance in blocks = one resident wave (0 = off)
#endif

constexpr long FIX_GRID = 148 * 4;  // persistent grid of the fix kernels

// ------------------------------------------------------------------------------------
// x sweep: one thread owns one row j and marches along i over a strip of columns,
// keeping the 3-cell stencil, its primitives, the previous face state and the previous
// interface flux in registers.  Adjacent threads are adjacent in j (the contiguous axis),
// so every global access is a fully coalesced 8-byte-per-lane transaction, each input
// value is loaded once per strip, and reconstruct -> predict -> HLLC -> flux difference
// -> psi damping never leave the register file.
// ------------------------------------------------------------------------------------

// Half-step face states of cell i computed straight from global memory (strip warm-up and the
// non-periodic boundary images of MUSCL_NEW.cpp:878-879).
template <class D>
__device__ __noinline__ D x_halfstates_global(const GridDims g, const SweepParams p,
                                                 const double* __restrict__ cur,
                                                 const double* __restrict__ prev, int i, int j,
                                                 double (&hL)[NV], double (&hR)[NV]) {
    D dv;  // own guard state, handed back by value (see FastDiv::absorb)
    double um[NV], u0[NV], up[NV], pm[NV], p0[NV], pp[NV];
    load_cell(cur, mhd_index(g, i - 1, j), um);
    load_cell(cur, mhd_index(g, i, j), u0);
    load_cell(cur, mhd_index(g, i + 1, j), up);
    c2p_full(dv, um, p.gm1, pm);
    c2p_full(dv, u0, p.gm1, p0);
    if (p.first_step) {
#pragma unroll
        for (int k = 0; k < NV; ++k) pp[k] = 0.0;
    } else {
        double q[NV];
        load_cell(prev, mhd_index(g, i + 1, j), q);
        c2p_full(dv, q, p.gm1, pp);
    }
#pragma unroll
    for (int k = 0; k < NV; ++k) reconstruct1(dv, pm[k], p0[k], pp[k], um[k], u0[k], up[k], hL[k], hR[k]);
    predictor<0>(dv, hL, hR, p.gm1, p.ch2, p.hdtdh);
    return dv;
}

__device__ __forceinline__ void prefetch_cell_l1(const double* __restrict__ base, size_t idx) {
#pragma unroll
    for (int k = 0; k < NV; ++k) asm volatile("prefetch.global.L1 [%0];" ::"l"(base + idx + k * MHD_KSTR));
}

// The non-trivial primitives (vx, vy, vz, p) of a cell are carried from one column to the next; the five that
// are copies of conserved components (rho, B, psi) are taken from the reloaded conserved state.
__device__ __forceinline__ void prim_vector(const double (&u)[NV], const double (&nt)[4], double (&v)[NV]) {
    v[0] = u[0]; v[1] = nt[0]; v[2] = nt[1]; v[3] = nt[2]; v[4] = nt[3];
    v[5] = u[5]; v[6] = u[6]; v[7] = u[7]; v[8] = u[8];
}

// One strip of one row.  Carried in registers from column to column: the right face state of the previous
// cell, the previous interface flux and the 4 non-trivial primitives of the previous cell (22 doubles).  The
// conserved 3-cell stencil is NOT carried: a column loaded for cell c is needed again for cells c+1, c+2 and for
// the update of cell c -- it is re-read through L1 (lines were prefetched one column ahead and are private to
// this warp's rows), which costs one issue slot per value where a register spill would cost two.
template <class D>
__device__ __forceinline__ bool x_sweep_block(const GridDims& g, const SweepParams& p,
                                              const double* __restrict__ cur, const double* __restrict__ prev,
                                              double* __restrict__ ubar, const int bx, const int by,
                                              const int strip) {
    const int j = 2 + bx * blockDim.x + threadIdx.x;
    if (j >= g.ny + 2) return false;
    const int ia = 2 + by * strip;
    const int ib = min(ia + strip, g.nxl + 2);  // cells [ia, ib) are updated by this thread
    const bool lo_image = (ia == 2) && (p.bc_lo == BC_TRANSMISSIVE);
    const bool hi_image = (ib == g.nxl + 2) && (p.bc_hi == BC_TRANSMISSIVE);
    D dv;

#pragma unroll
    for (int a = 1; a < XS_PF; ++a) {
        prefetch_cell_l1(cur, mhd_index(g, ia + a, j));
        if (!p.first_step) prefetch_cell_l1(prev, mhd_index(g, ia + a, j));
    }
    double hRprev[NV], Fprev[NV], PMnt[4];
    {
        // face state R of cell ia-1; at a transmissive wall ghost cell 1 carries the half state of
        // cell 3 (u[1] = u[3] applied to the half-step arrays, MUSCL_NEW.cpp:617-622, :878-879)
        double tL[NV], tR[NV];
        dv.absorb(x_halfstates_global<D>(g, p, cur, prev, lo_image ? 3 : ia - 1, j, tL, tR));
#pragma unroll
        for (int k = 0; k < NV; ++k) hRprev[k] = tR[k];
        double uM[NV];
        load_cell(cur, mhd_index(g, ia - 1, j), uM);
        const Prim q = c2p(dv, uM, p.gm1);
        PMnt[0] = q.vx; PMnt[1] = q.vy; PMnt[2] = q.vz; PMnt[3] = q.p;
    }

    for (int c = ia; c <= ib; ++c) {
        if (c + XS_PF <= g.nxl + 3) {
            prefetch_cell_l1(cur, mhd_index(g, c + XS_PF, j));
            if (!p.first_step) prefetch_cell_l1(prev, mhd_index(g, c + XS_PF, j));
        }
        double hL[NV], hR[NV], P0nt[4];
        if (c == ib && hi_image) {
            double tL[NV], tR[NV];
            dv.absorb(x_halfstates_global<D>(g, p, cur, prev, g.nxl, j, tL, tR));  // u[nx+2] = u[nx]
#pragma unroll
            for (int k = 0; k < NV; ++k) {
                hL[k] = tL[k];
                hR[k] = tR[k];
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) P0nt[k] = 0.0;  // last column of the strip: not carried further
        } else {
            double uM[NV], u0[NV], uP[NV], PM[NV], P0[NV], Pp[NV];
            load_cell(cur, mhd_index(g, c - 1, j), uM);
            load_cell(cur, mhd_index(g, c, j), u0);
            load_cell(cur, mhd_index(g, c + 1, j), uP);
            if (p.first_step) {
#pragma unroll
                for (int k = 0; k < NV; ++k) Pp[k] = 0.0;
            } else {
                double q[NV];
                load_cell(prev, mhd_index(g, c + 1, j), q);
                c2p_full(dv, q, p.gm1, Pp);
            }
            {
                const Prim q0 = c2p(dv, u0, p.gm1);
                P0nt[0] = q0.vx; P0nt[1] = q0.vy; P0nt[2] = q0.vz; P0nt[3] = q0.p;
            }
            prim_vector(uM, PMnt, PM);
            prim_vector(u0, P0nt, P0);
            double ratio[NV];
#pragma unroll
            for (int k = 0; k < NV; ++k) ratio[k] = limiter_ratio_k(dv, k, PM[k], P0[k], Pp[k]);
#pragma unroll
            for (int k = 0; k < NV; ++k) {
                // v_z / B_z keep the branch: identically zero in planar problems, where it skips the quotient for good
                if (k == 3 || k == 7) reconstruct_with_ratio<D, false>(dv, ratio[k], uM[k], u0[k], uP[k], hL[k], hR[k]);
                else reconstruct_with_ratio<D, true>(dv, ratio[k], uM[k], u0[k], uP[k], hL[k], hR[k]);
            }
            predictor<0>(dv, hL, hR, p.gm1, p.ch2, p.hdtdh);
        }
        double F[NV];
        // interface c-1/2 (MUSCL_NEW.cpp:884)
        hllc_flux<0>(dv, hRprev, hL, p.gamma, p.gm1, p.ch, p.ch2, p.inv2ch, p.halfch, F);
        if (c > ia) {
            // uBar[c-1] = u[c-1] - (dt/dx)(F[c-1/2] - F[c-3/2])   MUSCL_NEW.cpp:891; psi damping :900
            const size_t o = mhd_index(g, c - 1, j);
#pragma unroll
            for (int k = 0; k < NV; ++k) {
                double v = __ldg(cur + o + k * MHD_KSTR) - p.dtdh * (F[k] - Fprev[k]);
                if (k == 8) v = v * p.psi_factor;
                ubar[o + k * MHD_KSTR] = v;
            }
        }
#pragma unroll
        for (int k = 0; k < NV; ++k) {
            Fprev[k] = F[k];
            hRprev[k] = hR[k];
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) PMnt[k] = P0nt[k];
    }
    return dv.bad();
}

template <class D>
__global__ void __launch_bounds__(XS_BLOCK, XS_MINB) x_sweep_kernel(const GridDims g, const SweepArgs a,
                                                                    const double* __restrict__ cur,
                                                                    const double* __restrict__ prev,
                                                                    double* __restrict__ ubar,
                                                                    unsigned* __restrict__ block_flags,
                                                                    const int strip) {
    LSC_RETURN_IF_HALTED(a.slot);
    const SweepParams p = make_params<0>(a);
    // threads of the x sweep are independent: no barrier, a flagged thread just marks its block
    if (x_sweep_block<D>(g, p, cur, prev, ubar, blockIdx.x, blockIdx.y, strip))
        block_flags[blockIdx.y * gridDim.x + blockIdx.x] = 1u;
}

#ifndef LSC_RELAXED_TU
// ------------------------------------------------------------------------------------
// Head of a step (step_ctl.cuh): ComputeTimeStep's scalars MUSCL_NEW.cpp:117-118, `ch = maxSpeed; t += dt`
// :834-835, psiUpdate's factor :66-69 and the loop predicates :975 / :1003, by one thread.  status[0] holds the
// bit pattern of the max signal speed of the current state (the previous y sweep's fused reduction, or the
// stand-alone reduction after an upload; already max-reduced over ranks), status[1..3] the fix kernels' counts.
// ------------------------------------------------------------------------------------
__global__ void prepare_kernel(StepCtl* __restrict__ ctl, const StepCmd* __restrict__ cmd,
                               unsigned long long* __restrict__ status, double* __restrict__ log_ring) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    StepCtl& c = *ctl;
    bool fresh = false;
    if (c.batch_seen != cmd->batch) {  // first step of a batch: the host has handled whatever stopped the last one
        c.batch_seen = cmd->batch;
        c.halt = HALT_NONE;
        c.log_count = 0;
        fresh = true;
    }
    if (c.halt != HALT_NONE) return;  // c.sc.halt is set already: the rest of this step does nothing
    int halt = HALT_NONE;
    if (!fresh && pause_due(c.t, c.dt, cmd->t_stop, cmd->dump_counter)) halt = HALT_PAUSE;
    else if (c.step >= cmd->step_limit) halt = HALT_LIMIT;
    else if (cmd->keep_log && c.log_count >= STEP_LOG_CAP) halt = HALT_LOG_FULL;
    if (halt == HALT_NONE) {
        const double ms = __longlong_as_double((long long)status[0]);
        if (!(ms > 0.0) || !(ms <= 1.7976931348623157e308)) {
            c.max_speed = ms;
            halt = HALT_NAN;
        } else {
            const double dt = c.cfl * c.dspace / ms;  // :118
            const double arg = psi_argument(dt, ms);  // psiUpdate(psi, maxSpeed, 0.5*dt), :67, :900
            const unsigned long long abits = (unsigned long long)__double_as_longlong(arg);
            int hit = -1;
            for (int q = 0; q < c.ntab; ++q)
                if (c.tab_arg[q] == abits) hit = q;
            if (hit < 0) {
                c.miss_arg = abits;
                halt = HALT_EXP_MISS;  // status stays as it is: the resumed prepare consumes it
            } else {
                c.max_speed = ms;
                c.dt = dt;
                c.t += dt;  // :835
                c.step += 1;
                StepScalars& sc = c.sc;
                sc.ch = ms;  // :834
                sc.ch2 = ms * ms;
                sc.inv2ch = 1.0 / (2.0 * ms);  // :126
                sc.halfch = ms / 2.0;          // :127
                sc.dtdx = dt / c.dx;
                sc.hdtdx = 0.5 * (dt / c.dx);
                sc.dtdy = dt / c.dy;
                sc.hdtdy = 0.5 * (dt / c.dy);
                sc.psi_factor = c.tab_val[hit];
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
    }
    c.halt = halt;
    c.sc.halt = halt;
}

void launch_prepare(StepCtl* ctl, const StepCmd* cmd, unsigned long long* status, double* log_ring, cudaStream_t s) {
    prepare_kernel<<<1, 32, 0, s>>>(ctl, cmd, status, log_ring);
}

// words[n] = max(words[0..n)): the slabs of a single-process team bring their status[0] words to the leader
// (bit patterns of non-negative doubles order like the values)
__global__ void max_words_kernel(unsigned long long* __restrict__ words, int n) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    unsigned long long m = 0;
    for (int q = 0; q < n; ++q) m = words[q] > m ? words[q] : m;
    words[n] = m;
}

void launch_max_words(unsigned long long* words, int n, cudaStream_t s) { max_words_kernel<<<1, 32, 0, s>>>(words, n); }

// Recompute the blocks a fast sweep flagged, with IEEE division (persistent grid over the device-side list).
__global__ void __launch_bounds__(XS_BLOCK) x_sweep_fix_kernel(const GridDims g, const SweepArgs a,
                                                               const double* __restrict__ cur,
                                                               const double* __restrict__ prev,
                                                               double* __restrict__ ubar,
                                                               unsigned long long* __restrict__ status,
                                                               unsigned* __restrict__ block_flags,
                                                               const int strip, const unsigned nbx,
                                                               const unsigned nblocks) {
    LSC_RETURN_IF_HALTED(a.slot);
    const SweepParams p = make_params<0>(a);
    for (unsigned bid = blockIdx.x; bid < nblocks; bid += gridDim.x) {
        if (block_flags[bid] == 0u) continue;  // uniform across the block
        x_sweep_block<IeeeDiv>(g, p, cur, prev, ubar, (int)(bid % nbx), (int)(bid / nbx), strip);
        __syncthreads();
        if (threadIdx.x == 0) {
            block_flags[bid] = 0u;  // all-zero again for the next step
            atomicAdd(status + 1, 1ull);
        }
    }
}

// development aid: LSC_XS_STRIP=<columns> overrides choose_x_strip (strip-length experiments, scripts/strip_sweep.py)
static int strip_override() {
    static const int v = [] {
        const char* e = getenv("LSC_XS_STRIP");
        return e ? atoi(e) : 0;
    }();
    return v;
}

void launch_x_sweep(const GridDims& g, const SweepArgs& p, const double* u_cur, const double* u_prev,
                    double* u_bar, unsigned long long* status, unsigned* x_flags, cudaStream_t s) {
    // strip length along i: at most XS_STRIP columns per thread (3 warm-up cells amortised over the strip),
    // chosen so that the blocks fill whole rounds of the 148 SMs x XS_MINB resident blocks (choose_x_strip)
    dim3 block(XS_BLOCK);
    const int row_blocks = (g.ny + block.x - 1) / block.x;
    int strip = choose_x_strip(row_blocks, g.nxl, XS_STRIP, 148 * XS_MINB, XS_MIN_STRIP);
    if (strip_override() >= XS_MIN_STRIP) strip = strip_override();
    dim3 grid(row_blocks, (g.nxl + strip - 1) / strip);
    x_sweep_kernel<MHD_FAST_POLICY><<<grid, block, 0, s>>>(g, p, u_cur, u_prev, u_bar, x_flags, strip);
    const unsigned nblocks = grid.x * grid.y;
    const unsigned nfix = (unsigned)std::min<long>(FIX_GRID, (long)nblocks);
    x_sweep_fix_kernel<<<nfix, block, 0, s>>>(g, p, u_cur, u_prev, u_bar, status, x_flags, strip, grid.x, nblocks);
}

long x_sweep_block_count(const GridDims& g) {
    const int row_blocks = (g.ny + XS_BLOCK - 1) / XS_BLOCK;
    return (long)row_blocks * ((g.nxl + XS_MIN_STRIP - 1) / XS_MIN_STRIP);  // strips are never shorter
}
#else  // LSC_RELAXED_TU
void launch_x_sweep_relaxed(const GridDims& g, const SweepArgs& p, const double* u_cur, const double* u_prev,
                            double* u_bar, unsigned* x_flags, cudaStream_t s) {
    dim3 block(XS_BLOCK);
    const int row_blocks = (g.ny + block.x - 1) / block.x;
    const int strip = choose_x_strip(row_blocks, g.nxl, XS_STRIP, 148 * XS_MINB, XS_MIN_STRIP);
    dim3 grid(row_blocks, (g.nxl + strip - 1) / strip);
    x_sweep_kernel<RelaxedDiv><<<grid, block, 0, s>>>(g, p, u_cur, u_prev, u_bar, x_flags, strip);
}
#endif

// ------------------------------------------------------------------------------------
// y sweep: one thread per cell along the contiguous axis.  A block of BJ threads covers cells
// jo-1 .. jo+BJ-2 of one column; face states and interface fluxes are exchanged between
// neighbouring threads through shared memory, so a block produces BJ-2 updated cells.
// ------------------------------------------------------------------------------------
constexpr int BJ = YS_BJ;

// Which stored cell a (possibly ghost) row index refers to, and whether the reflective sign
// applies (MUSCL_NEW.cpp:625-654).  Interior rows map to themselves.
__device__ __forceinline__ int map_row(int j, int ny, int bc, bool& flip) {
    flip = false;
    if (j >= 2 && j <= ny + 1) return j;
    if (bc == BC_PERIODIC) return (j < 2) ? j + ny : j - ny;
    flip = (bc == BC_REFLECTIVE);
    return (j < 2) ? j + 2 : j - 2;
}

__device__ __forceinline__ void load_row_cell(const double* __restrict__ base, const GridDims& g, int i, int j,
                                              int bc, double (&u)[NV]) {
    bool flip;
    const int js = map_row(j, g.ny, bc, flip);
    load_cell(base, mhd_index(g, i, js), u);
    if (flip) {
        u[2] = -u[2];
        u[6] = -u[6];
    }
}

// Half-step face states (bottom, top) of cell j of column i straight from global memory.
template <class D>
__device__ __noinline__ D y_halfstates_global(const GridDims g, const SweepParams p,
                                                 const double* __restrict__ ubar,
                                                 const double* __restrict__ cur, int i, int j,
                                                 double (&hB)[NV], double (&hT)[NV]) {
    D dv;
    double um[NV], u0[NV], up[NV], q[NV], pm[NV], p0[NV], pp[NV];
    load_row_cell(ubar, g, i, j - 1, p.bc_lo, um);
    load_row_cell(ubar, g, i, j, p.bc_lo, u0);
    load_row_cell(ubar, g, i, j + 1, p.bc_lo, up);
    load_row_cell(cur, g, i, j - 1, p.bc_lo, q);
    c2p_full(dv, q, p.gm1, pm);
    load_row_cell(cur, g, i, j, p.bc_lo, q);
    c2p_full(dv, q, p.gm1, p0);
    load_row_cell(cur, g, i, j + 1, p.bc_lo, q);
    c2p_full(dv, q, p.gm1, pp);
#pragma unroll
    for (int k = 0; k < NV; ++k) reconstruct1(dv, pm[k], p0[k], pp[k], um[k], u0[k], up[k], hB[k], hT[k]);
    predictor<1>(dv, hB, hT, p.gm1, p.ch2, p.hdtdh);
    return dv;
}

// One block of the y sweep at block coordinates (bx, by).  Returns this thread's sticky division flag and its
// contribution to the CFL maximum in `speed`.  All threads of the block must call it (it synchronises).
// NONPER: the y boundaries are not periodic (compile-time, so that the periodic instantiation carries none of the
// ghost-image logic: as run-time predicates it cost the periodic y sweep 3.5 %).
template <class D, bool NONPER>
__device__ __forceinline__ bool y_sweep_block(const GridDims& g, const SweepParams& p,
                                              const double* __restrict__ ubar, const double* __restrict__ cur,
                                              double* __restrict__ next, const int bx, const int by,
                                              double& speed, const unsigned long long* status,
                                              const int pf_cols = 0) {
    // Fast instantiation only: the running exact maximum of the blocks that already finished (stale is fine,
    // it only grows), rounded down to fp32.  Read first so that its latency hides behind the tile load.
    constexpr bool FILTER = !std::is_same<D, IeeeDiv>::value;
    float inv_run_max = 0.0f;  // 1/R rounded up (R = 0 -> +inf -> every warp evaluates)
    unsigned long long run_bits = 0;
    if (FILTER) run_bits = running_max_bits(status);
#if YS_TMA
    // Experiment J1: the tile of BOTH states staged by one bulk copy (TMA, cp.async.bulk + mbarrier) per
    // state -- in the blocked layout the rows of a tile are one contiguous run of 5-6 row groups -- and kept in the
    // device layout; threads read their own and their neighbours' uBar from it, only primitives go through sP.
    constexpr int RAW_GROUPS = (BJ + 2 + 2 * JBLK - 2) / JBLK;  // row groups a tile of BJ+2 rows can touch (6)
    constexpr int RAW_DOUBLES = RAW_GROUPS * NV * JBLK;
    __shared__ __align__(128) double rawU[RAW_DOUBLES];  // uBar, device layout
    __shared__ __align__(128) double rawC[RAW_DOUBLES];  // u^n, device layout; reused for the bottom face states
    __shared__ double sP[NV][BJ + 2];                    // primitives of u^n; later the interface fluxes
    __shared__ unsigned long long tile_mbar;
    double(*sH)[BJ] = reinterpret_cast<double(*)[BJ]>(rawC);  // [NV][BJ]: written after the first barrier
    static_assert(NV * BJ <= RAW_DOUBLES, "face states must fit the u^n staging buffer");
#else
    __shared__ double sU[NV][BJ + 2];  // uBar of cells jo-2 .. jo+BJ-1
    __shared__ double sP[NV][BJ + 2];  // primitives of u^n there; later the interface fluxes
    __shared__ double sH[NV][BJ];      // bottom face state of each thread's cell
#endif

    const int tid = threadIdx.x;
    const int i = 2 + by;
    const int jo = 2 + bx * (BJ - 2);
    const int c = jo - 1 + tid;  // this thread's cell
#if YS_PF_BLOCKS > 0
    // The tile a block loads first thing is what its warps wait for longest (ncu: the first reciprocal behind the
    // tile load holds 10 % of the sweep's warp samples).  Blocks are dispatched in launch order, so the block that
    // will take this one's place on the SM is about one resident wave (YS_PF_BLOCKS) further on: ask for its tile
    // -- the same rows, pf_cols columns on; in the blocked layout the rows of a tile are ONE contiguous run per
    // state -- to be brought into L2 now, with one bulk prefetch per state.
    if (FILTER && pf_cols > 0 && tid == BJ - 1) {
        const int ip = i + pf_cols;  // pf_cols: YS_PF_BLOCKS in columns, from the launch (YCols)
        if (ip <= g.nxl + 1) {
            const int g0 = (JOFF + jo - 2) >> 5;
            const int g1 = (JOFF + min(jo + BJ - 1, g.ny + 3)) >> 5;
            const size_t off = (size_t)ip * mhd_column_doubles(g) + (size_t)g0 * (NV * JBLK);
            const unsigned bytes = (unsigned)(g1 - g0 + 1) * (NV * JBLK * 8);
            asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(ubar + off), "r"(bytes) : "memory");
            asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(cur + off), "r"(bytes) : "memory");
        }
    }
#endif
    const int bc = p.bc_lo;      // the reference uses one y boundary type for both ends
    D dv;

#if YS_TMA
    const int raw_g0 = (JOFF + jo - 2) >> 5;  // first row group of the tile
    auto ridx = [&](const int t) -> int {      // tile entry t (row jo-2+t) in the staged device layout
        const int jj = JOFF + jo - 2 + t;
        return ((jj >> 5) - raw_g0) * (NV * JBLK) + (jj & (JBLK - 1));
    };
    const bool whole_tile_interior = (jo - 2 >= 2) && (jo + BJ - 1 <= g.ny + 1);  // block-uniform
    if (whole_tile_interior) {
        const unsigned mbar = (unsigned)__cvta_generic_to_shared(&tile_mbar);
        // thread 0 alone sets the barrier up, issues the two copies and waits for them; the block barrier behind it
        // hands the staged tile to everyone else (YS_TMA == 2: every thread waits on the mbarrier itself, which
        // costs one more block barrier to publish its initialisation)
        if (tid == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mbar));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
#if YS_TMA == 2
        __syncthreads();
#endif
        if (tid == 0) {
            const int g1 = (JOFF + jo + BJ - 1) >> 5;
            const unsigned bytes = (unsigned)(g1 - raw_g0 + 1) * (NV * JBLK * 8);
            const size_t off = (size_t)i * mhd_column_doubles(g) + (size_t)raw_g0 * (NV * JBLK);
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(2 * bytes) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                             (unsigned)__cvta_generic_to_shared(rawU)),
                         "l"(ubar + off), "r"(bytes), "r"(mbar)
                         : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                             (unsigned)__cvta_generic_to_shared(rawC)),
                         "l"(cur + off), "r"(bytes), "r"(mbar)
                         : "memory");
        }
#if YS_TMA == 2
        const bool waits = true;
#else
        const bool waits = (tid == 0);
#endif
        if (waits)
            asm volatile(
                "{\n\t.reg .pred p;\n\t"
                "WAIT_TILE:\n\t"
                "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n\t"
                "@!p bra WAIT_TILE;\n\t}" ::"r"(mbar)
                : "memory");
#if YS_TMA != 2
        __syncthreads();
#endif
    } else {
        // a tile that reaches into the ghost rows (first / last blocks of a column): per-thread loads through the
        // boundary map, stored at the entry's place in the same staged layout
        auto fill = [&](const int t) {
            const int jt = jo - 2 + t;
            if (jt > g.ny + 3) return;
            bool flip = false;
            int js = jt;
            if (jt < 2 || jt > g.ny + 1) {
                if (NONPER) js = map_row(jt, g.ny, bc, flip);
                else js = (jt < 2) ? jt + g.ny : jt - g.ny;
            }
            const size_t idx = mhd_index(g, i, js);
            double qu[NV], q[NV];
            load_cell(ubar, idx, qu);
            load_cell(cur, idx, q);
            if (flip) {
                qu[2] = -qu[2]; qu[6] = -qu[6];
                q[2] = -q[2];   q[6] = -q[6];
            }
            const int r = ridx(t);
#pragma unroll
            for (int k = 0; k < NV; ++k) {
                rawU[r + k * JBLK] = qu[k];
                rawC[r + k * JBLK] = q[k];
            }
        };
        fill(tid + 1);
        if (tid < 2) fill(tid == 0 ? 0 : BJ + 1);
        __syncthreads();
    }
    double ub[NV], p0[NV];
    {
        const int r = ridx(tid + 1);
        double q[NV];
#pragma unroll
        for (int k = 0; k < NV; ++k) {
            ub[k] = rawU[r + k * JBLK];
            q[k] = rawC[r + k * JBLK];
        }
        c2p_full(dv, q, p.gm1, p0);
#pragma unroll
        for (int k = 0; k < NV; ++k) sP[k][tid + 1] = p0[k];
        if (tid < 2) {  // the two end entries' primitives
            const int te = (tid == 0) ? 0 : BJ + 1;
            if (jo - 2 + te <= g.ny + 3) {
                const int re = ridx(te);
                double qe[NV], ve[NV];
#pragma unroll
                for (int k = 0; k < NV; ++k) qe[k] = rawC[re + k * JBLK];
                c2p_full(dv, qe, p.gm1, ve);
#pragma unroll
                for (int k = 0; k < NV; ++k) sP[k][te] = ve[k];
            }
        }
    }
#define SU_AT(k, t) rawU[ridx(t) + (k) * JBLK]
#else
    // Tile of BJ+2 cells (rows jo-2 .. jo+BJ-1): uBar and the primitives of u^n go to shared memory for the
    // neighbours.  A thread fills the tile entry of its OWN cell (t = tid+1) and keeps those 18 values in
    // registers; the two end entries are an extra pass of threads 0 and 1.
    // Both loads of threads 0 and 1 -- their own cell and an end entry -- are issued before anything waits on
    // them: with the end entries filled by a second pass the block's first barrier waited for warp 0 to go to
    // memory twice (ncu: 7.7 % of the sweep's warp samples sat in that barrier).
    auto tile_slot = [&](const int t, size_t& idx, bool& flip) -> bool {
        const int jt = jo - 2 + t;
        flip = false;
        idx = 0;
        if (jt > g.ny + 3) return false;
        // interior rows map to themselves; only ghost rows go through the boundary map.  Both states of the
        // cell share one address computation.
        int js = jt;
        if (jt < 2 || jt > g.ny + 1) {
            if (NONPER) js = map_row(jt, g.ny, bc, flip);
            else js = (jt < 2) ? jt + g.ny : jt - g.ny;  // periodic wrap, no sign flip
        }
        idx = mhd_index(g, i, js);
        return true;
    };
    auto tile_put = [&](const int t, const bool flip, double (&qu)[NV], double (&q)[NV], double (&v)[NV]) {
        if (flip) {
            qu[2] = -qu[2];
            qu[6] = -qu[6];
            q[2] = -q[2];
            q[6] = -q[6];
        }
#pragma unroll
        for (int k = 0; k < NV; ++k) sU[k][t] = qu[k];
        c2p_full(dv, q, p.gm1, v);
#pragma unroll
        for (int k = 0; k < NV; ++k) sP[k][t] = v[k];
    };
    double ub[NV], p0[NV];
    {
        const int te = (tid == 0) ? 0 : BJ + 1;
        const bool extra_lane = tid < 2;
        size_t idx_o, idx_e;
        bool flip_o, flip_e;
        const bool have_o = tile_slot(tid + 1, idx_o, flip_o);
        const bool have_e = extra_lane && tile_slot(te, idx_e, flip_e);
        double qo[NV], ue[NV], qe[NV], ve[NV];
        if (have_o) {
            load_cell(ubar, idx_o, ub);
            load_cell(cur, idx_o, qo);
        }
        if (have_e) {
            load_cell(ubar, idx_e, ue);
            load_cell(cur, idx_e, qe);
        }
        if (have_o) tile_put(tid + 1, flip_o, ub, qo, p0);
        if (have_e) tile_put(te, flip_e, ue, qe, ve);
    }
#define SU_AT(k, t) sU[k][t]
#endif
    __syncthreads();
    // loaded first thing, converted behind the tile loads: a warp that converts right behind the load sits out an L2
    // round trip before it has issued its tile loads
    if (FILTER) inv_run_max = inv_running_max_from_bits(run_bits);

    const bool active = (c <= g.ny + 2);
    // Non-periodic boundaries: the ghost face states are the (sign-flipped) face states of cell 3 / ny -- the
    // reference applies its boundary function to the half-step arrays (MUSCL_NEW.cpp:940-941).  Those cells belong
    // to this very block (cell 3 is thread 2 of the column's first block, cell ny the thread two below the ghost),
    // and only one face of each ghost is consumed: the TOP state of ghost cell 1 (interface 1|2) and the BOTTOM
    // state of ghost cell ny+2 (interface ny+1|ny+2).  So the ghost threads compute nothing: thread c == 3
    // publishes its top state for thread c == 1, and thread c == ny+1 takes the bottom state of cell ny (its lower
    // neighbour's entry) instead of its upper neighbour's.  Only when the block starts less than two cells below
    // the top ghost does that thread fall back to the stencil from global memory.  (The ghost used to run the whole
    // reconstruction + predictor from global memory while its block waited at the barrier: KH 1024^2 y sweep
    // 0.139 ms against 0.100 ms for the periodic problem.)
    constexpr bool nonper = NONPER;  // == (bc != BC_PERIODIC), chosen by the launch
    const bool image_lo = nonper && c == 1;                   // thread 0 of the column's first block
    const bool image_hi = nonper && c == g.ny + 2;
    const bool image_hi_global = image_hi && tid < 2;         // cell ny lies in the previous block
    const bool takes_lower = nonper && c == g.ny + 1 && tid >= 1;  // its upper neighbour is the top ghost
    __shared__ double sT[NV];                                 // top face state of cell 3 (first block, non-periodic)
    double hB[NV], hT[NV];
    if (active) {
        if (image_hi_global) {
            double tB[NV], tT[NV];
            dv.absorb(y_halfstates_global<D>(g, p, ubar, cur, i, g.ny, tB, tT));
#pragma unroll
            for (int k = 0; k < NV; ++k) {
                hB[k] = tB[k];
                hT[k] = tT[k];
                ub[k] = 0.0;
            }
            if (bc == BC_REFLECTIVE) {
                hB[2] = -hB[2]; hB[6] = -hB[6];
                hT[2] = -hT[2]; hT[6] = -hT[6];
            }
        } else if (image_lo || image_hi) {
#pragma unroll
            for (int k = 0; k < NV; ++k) {
                hB[k] = 0.0;
                hT[k] = 0.0;
                ub[k] = 0.0;
            }
        } else {
            const int t = tid + 1;
            double ratio[NV];
#pragma unroll
            for (int k = 0; k < NV; ++k) ratio[k] = limiter_ratio_k(dv, k, sP[k][t - 1], p0[k], sP[k][t + 1]);
#pragma unroll
            for (int k = 0; k < NV; ++k) {
                if (k == 3 || k == 7)
                    reconstruct_with_ratio<D, false>(dv, ratio[k], SU_AT(k, t - 1), ub[k], SU_AT(k, t + 1), hB[k], hT[k]);
                else
                    reconstruct_with_ratio<D, true>(dv, ratio[k], SU_AT(k, t - 1), ub[k], SU_AT(k, t + 1), hB[k], hT[k]);
            }
            predictor<1>(dv, hB, hT, p.gm1, p.ch2, p.hdtdh);
            if (nonper && c == 3) {
#pragma unroll
                for (int k = 0; k < NV; ++k) sT[k] = hT[k];
            }
        }
#pragma unroll
        for (int k = 0; k < NV; ++k) sH[k][tid] = hB[k];
    }
    __syncthreads();
    if (image_lo) {
#pragma unroll
        for (int k = 0; k < NV; ++k) hT[k] = sT[k];
        if (bc == BC_REFLECTIVE) {
            hT[2] = -hT[2];
            hT[6] = -hT[6];
        }
    }
    // flux through the interface above cell c: FluxHLLCY(top of c, bottom of c+1)  MUSCL_NEW.cpp:945
    double F[NV];
    const bool has_flux = (tid < BJ - 1) && (c <= g.ny + 1);
    if (has_flux) {
        double nb[NV];
        const int tn = takes_lower ? tid - 1 : tid + 1;
#pragma unroll
        for (int k = 0; k < NV; ++k) nb[k] = sH[k][tn];
        if (takes_lower && bc == BC_REFLECTIVE) {
            nb[2] = -nb[2];
            nb[6] = -nb[6];
        }
        hllc_flux<1>(dv, hT, nb, p.gamma, p.gm1, p.ch, p.ch2, p.inv2ch, p.halfch, F);
#pragma unroll
        for (int k = 0; k < NV; ++k) sP[k][tid] = F[k];
    }
    __syncthreads();

    speed = 0.0;
    const bool updates = (tid >= 1 && has_flux && c >= 2);
    double un[NV];
    if (updates) {
        // u^{n+1} = uBar - (dt/dy)(F[j] - F[j-1])  MUSCL_NEW.cpp:953; psi damping :963
        const size_t o = mhd_index(g, i, c);
#pragma unroll
        for (int k = 0; k < NV; ++k) {
            double v = ub[k] - p.dtdh * (F[k] - sP[k][tid - 1]);
            if (k == 8) v = v * p.psi_factor;
            un[k] = v;
            next[o + k * MHD_KSTR] = v;
        }
    }
    // Signal speed of the new state for the next step's CFL condition (MUSCL_NEW.cpp:75-111).  Only the global
    // maximum is consumed, so a warp evaluates the exact fp64 expression only if one of its cells could hold
    // it: its fp32 upper bound is not provably below the running exact maximum.  Skipped cells are < that
    // running maximum <= the final maximum, so the result is the same bit pattern whatever the block order.
    bool evaluate = updates;
    if (FILTER) {
        bool candidate = false;
        if (updates)
            candidate = !(cfl_speed_over_max_f32(un, __double2float_ru(p.gamma), __double2float_ru(p.gm1), inv_run_max) < 1.0f);
        const bool warp_has_candidate = __any_sync(0xffffffffu, candidate);  // every lane votes (no short circuit)
        evaluate = updates && warp_has_candidate;
    }
    if (evaluate) {
        const double sp = cfl_speed(dv, un, p.gamma);
        if (sp > speed) speed = sp;
    }
    return dv.bad();
}

template <class D, bool NONPER>
__global__ void __launch_bounds__(BJ, YS_MINB) y_sweep_kernel(const GridDims g, const SweepArgs a,
                                                              const double* __restrict__ ubar,
                                                              const double* __restrict__ cur,
                                                              double* __restrict__ next,
                                                              unsigned long long* __restrict__ status,
                                                              unsigned* __restrict__ fix_list, const YCols yc) {
    LSC_RETURN_IF_HALTED(a.slot);
    const SweepParams p = make_params<1>(a);
    double speed;
    // column of this block row: a launch covers all columns, the four slab-edge columns or the rest (YCols)
    const int col = yc.c0 + (int)blockIdx.y + ((int)blockIdx.y >= yc.jump_at ? yc.jump : 0);
    const bool bad = y_sweep_block<D, NONPER>(g, p, ubar, cur, next, blockIdx.x, col, speed, status, yc.pf_cols);
    if (__syncthreads_or(bad)) {  // whole block is recomputed by the fix kernel; its maximum must not count
        if (threadIdx.x == 0) fix_list[atomicAdd(status + yc.counter, 1ull)] = col * gridDim.x + blockIdx.x;
    } else {
        warp_max_to_global(speed, status);
    }
}

#ifndef LSC_RELAXED_TU
template <bool NONPER>
__global__ void __launch_bounds__(BJ) y_sweep_fix_kernel(const GridDims g, const SweepArgs a,
                                                         const double* __restrict__ ubar,
                                                         const double* __restrict__ cur, double* __restrict__ next,
                                                         unsigned long long* __restrict__ status,
                                                         const unsigned* __restrict__ fix_list, const unsigned nbx,
                                                         const int counter) {
    LSC_RETURN_IF_HALTED(a.slot);
    const SweepParams p = make_params<1>(a);
    const unsigned long long count = status[counter];
    for (unsigned long long q = blockIdx.x; q < count; q += gridDim.x) {
        const unsigned bid = fix_list[q];
        double speed;
        y_sweep_block<IeeeDiv, NONPER>(g, p, ubar, cur, next, (int)(bid % nbx), (int)(bid / nbx), speed, status);
        block_max_to_global(speed, status);
        __syncthreads();  // shared tiles are reused by the next list entry
    }
}

void launch_y_sweep(const GridDims& g, const SweepArgs& p, const double* u_bar, const double* u_cur,
                    double* u_next, unsigned long long* status, unsigned* fix_list, cudaStream_t s,
                    int part) {
    dim3 block(BJ);
    YCols yc = y_columns(g, part);
    // the slab-edge launch keeps its own list (behind the y_sweep_block_count entries of the others) and counter
    unsigned* list = fix_list + (part == Y_EDGES ? y_sweep_block_count(g) : 0);
    dim3 grid((g.ny + (BJ - 2) - 1) / (BJ - 2), yc.ncols);
    yc.pf_cols = (YS_PF_BLOCKS + (int)grid.x - 1) / (int)grid.x;  // L2 prefetch distance: one resident wave of blocks
    if (p.bc_lo != BC_PERIODIC)
        y_sweep_kernel<MHD_FAST_POLICY, true><<<grid, block, 0, s>>>(g, p, u_bar, u_cur, u_next, status, list, yc);
    else
        y_sweep_kernel<MHD_FAST_POLICY, false><<<grid, block, 0, s>>>(g, p, u_bar, u_cur, u_next, status, list, yc);
    const unsigned nfix = (unsigned)std::min<long>(FIX_GRID, (long)grid.x * grid.y);
    if (p.bc_lo != BC_PERIODIC)
        y_sweep_fix_kernel<true><<<nfix, block, 0, s>>>(g, p, u_bar, u_cur, u_next, status, list, grid.x, yc.counter);
    else
        y_sweep_fix_kernel<false><<<nfix, block, 0, s>>>(g, p, u_bar, u_cur, u_next, status, list, grid.x, yc.counter);
}

long y_sweep_block_count(const GridDims& g) { return (long)((g.ny + (BJ - 2) - 1) / (BJ - 2)) * g.nxl; }
long y_edge_block_count(const GridDims& g) { return (long)((g.ny + (BJ - 2) - 1) / (BJ - 2)) * 4; }
#else  // LSC_RELAXED_TU
void launch_y_sweep_relaxed(const GridDims& g, const SweepArgs& p, const double* u_bar, const double* u_cur,
                            double* u_next, unsigned long long* status, unsigned* fix_list, cudaStream_t s,
                            int part) {
    dim3 block(BJ);
    const YCols yc = y_columns(g, part);
    dim3 grid((g.ny + (BJ - 2) - 1) / (BJ - 2), yc.ncols);
    if (p.bc_lo != BC_PERIODIC)
        y_sweep_kernel<RelaxedDiv, true><<<grid, block, 0, s>>>(g, p, u_bar, u_cur, u_next, status, fix_list, yc);
    else
        y_sweep_kernel<RelaxedDiv, false><<<grid, block, 0, s>>>(g, p, u_bar, u_cur, u_next, status, fix_list, yc);
}
#endif
#ifndef LSC_RELAXED_TU

// ------------------------------------------------------------------------------------
// Ghost columns in x (single-rank case; between ranks the same columns are filled by the
// halo exchange).  periodic: MUSCL_NEW.cpp:609-614, transmissive: :617-622.
// ------------------------------------------------------------------------------------
__global__ void x_ghost_kernel(const GridDims g, int bc_x, double* __restrict__ u) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= g.ny + 4) return;
    const int k = blockIdx.y;
    const int nx = g.nxl;
    int dst[4] = {0, 1, nx + 2, nx + 3};
    int src[4];
    if (bc_x == BC_PERIODIC) {
        src[0] = nx; src[1] = nx + 1; src[2] = 2; src[3] = 3;
    } else {
        src[0] = 2; src[1] = 3; src[2] = nx; src[3] = nx + 1;
    }
    double* pl = u + k * MHD_KSTR;
#pragma unroll
    for (int q = 0; q < 4; ++q) pl[mhd_index(g, dst[q], j)] = pl[mhd_index(g, src[q], j)];
}

void launch_x_ghosts(const GridDims& g, int bc_x, double* u, cudaStream_t s) {
    dim3 block(256);
    dim3 grid((g.ny + 4 + 255) / 256, NV);
    x_ghost_kernel<<<grid, block, 0, s>>>(g, bc_x, u);
}

// boundary image of one side only (end ranks of a non-periodic slab decomposition), MUSCL_NEW.cpp:617-622
__global__ void x_ghost_side_kernel(const GridDims g, int hi_side, double* __restrict__ u) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= g.ny + 4) return;
    double* pl = u + blockIdx.y * MHD_KSTR;
    const int nx = g.nxl;
    if (hi_side) {
        pl[mhd_index(g, nx + 2, j)] = pl[mhd_index(g, nx, j)];
        pl[mhd_index(g, nx + 3, j)] = pl[mhd_index(g, nx + 1, j)];
    } else {
        pl[mhd_index(g, 0, j)] = pl[mhd_index(g, 2, j)];
        pl[mhd_index(g, 1, j)] = pl[mhd_index(g, 3, j)];
    }
}

void launch_x_ghosts_side(const GridDims& g, int hi_side, double* u, cudaStream_t s) {
    dim3 block(256), grid((g.ny + 4 + 255) / 256, NV);
    x_ghost_side_kernel<<<grid, block, 0, s>>>(g, hi_side, u);
}

// Stand-alone CFL reduction over cells [i_lo,i_hi] x [j_lo,j_hi] (inclusive).  The reference
// reduces over i in [1,nx+1], j in [1,ny+1] of the array as it stands BEFORE the step's boundary
// fill (MUSCL_NEW.cpp:73-74, :833 vs :838) -- on the first step that includes analytic ghost values.
__global__ void max_speed_kernel(const GridDims g, double gamma, const double* __restrict__ u, int i_lo,
                                 int i_hi, int j_lo, int j_hi, unsigned long long* __restrict__ max_bits) {
    const int j = j_lo + blockIdx.x * blockDim.x + threadIdx.x;
    double best = 0.0;
    if (j <= j_hi) {
        for (int i = i_lo + blockIdx.y; i <= i_hi; i += gridDim.y) {
            double q[NV];
            load_cell(u, mhd_index(g, i, j), q);
            IeeeDiv dv;  // cold kernel (once per upload): plain IEEE division
            const double sp = cfl_speed(dv, q, gamma);
            if (sp > best) best = sp;
        }
    }
    block_max_to_global(best, max_bits);
}

void launch_max_speed(const GridDims& g, double gamma, const double* u, int i_lo, int i_hi, int j_lo, int j_hi,
                      unsigned long long* max_bits, cudaStream_t s) {
    dim3 block(128);
    const int nj = j_hi - j_lo + 1, ni = i_hi - i_lo + 1;
    dim3 grid((nj + 127) / 128, ni < 64 ? ni : 64);
    max_speed_kernel<<<grid, block, 0, s>>>(g, gamma, u, i_lo, i_hi, j_lo, j_hi, max_bits);
}

// ------------------------------------------------------------------------------------
// Reference array-of-structures [i][j][9] <-> blocked device layout, for `ni` columns starting at i0.
// The AoS chunk pointer addresses column i0 (chunk-local), rows 0..ny+3.
// ------------------------------------------------------------------------------------
__global__ void aos_to_soa_kernel(const GridDims g, const double* __restrict__ aos, int i0, int ni,
                                  double* __restrict__ soa) {
    __shared__ double tile[32 * NV];
    const int jb = blockIdx.x * 32;
    const int il = blockIdx.y;
    const int nrow = g.ny + 4;
    const int nj = min(32, nrow - jb);
    const double* src = aos + ((size_t)il * nrow + jb) * NV;
    for (int t = threadIdx.x; t < nj * NV; t += blockDim.x) tile[t] = src[t];
    __syncthreads();
    for (int t = threadIdx.x; t < nj * NV; t += blockDim.x) {
        const int k = t / nj, jj = t - k * nj;
        soa[mhd_index(g, i0 + il, jb + jj) + k * MHD_KSTR] = tile[jj * NV + k];
    }
}

__global__ void soa_to_aos_kernel(const GridDims g, const double* __restrict__ soa, int i0, int ni,
                                  double* __restrict__ aos) {
    __shared__ double tile[32 * NV];
    const int jb = blockIdx.x * 32;
    const int il = blockIdx.y;
    const int nrow = g.ny + 4;
    const int nj = min(32, nrow - jb);
    for (int t = threadIdx.x; t < nj * NV; t += blockDim.x) {
        const int k = t / nj, jj = t - k * nj;
        tile[jj * NV + k] = soa[mhd_index(g, i0 + il, jb + jj) + k * MHD_KSTR];
    }
    __syncthreads();
    double* dst = aos + ((size_t)il * nrow + jb) * NV;
    for (int t = threadIdx.x; t < nj * NV; t += blockDim.x) dst[t] = tile[t];
}

void launch_aos_to_soa(const GridDims& g, const double* aos, int i0, int ni, double* soa, cudaStream_t s) {
    dim3 grid((g.ny + 4 + 31) / 32, ni);
    aos_to_soa_kernel<<<grid, 96, 0, s>>>(g, aos, i0, ni, soa);
}

void launch_soa_to_aos(const GridDims& g, const double* soa, int i0, int ni, double* aos, cudaStream_t s) {
    dim3 grid((g.ny + 4 + 31) / 32, ni);
    soa_to_aos_kernel<<<grid, 96, 0, s>>>(g, soa, i0, ni, aos);
}

#endif  // !LSC_RELAXED_TU
}  // namespace lsc

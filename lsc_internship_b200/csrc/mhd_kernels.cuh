// mhd_kernels.cuh -- launch-side declarations shared by the kernels and the C ABI.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>

#include "step_ctl.cuh"

namespace lsc {

constexpr int NV = 9;  // rho, mx, my, mz, E, Bx, By, Bz, psi  (MUSCL_NEW.cpp:11-19)

// Boundary-condition codes (include/lsc_b200.h mirrors these values).
enum : int { BC_PERIODIC = 0, BC_REFLECTIVE = 1, BC_TRANSMISSIVE = 2 };

// Device layout of one 9-component MHD state: column-major blocks of 32 rows.  Column i (x index, the slow
// axis of the reference's [i][j] order, MUSCL_NEW.cpp:673-674) is one contiguous run of NV*pitch doubles; inside
// it rows are grouped 32 at a time and the nine components of a group are stored one after the other:
//     base + i*(NV*pitch) + (jj/32)*(NV*32) + k*32 + jj%32,      jj = JOFF + j
// so a warp's 32 adjacent rows of one component are one 256-byte run (coalesced exactly like a plane), the
// nine components of a cell sit at COMPILE-TIME offsets k*256 bytes from one address (one 64-bit address per
// cell instead of one per value: the separate planes cost ~140 integer instructions per cell in the x sweep),
// and the two halo columns a neighbour rank needs are ONE contiguous message.  Two ghost cells per side;
// JOFF = 14 and pitch % 32 == 0 put the first interior row (j = 2) on a 128-byte line.
// The 4-component Euler states (euler_kernels.cu) use the same layout with 4 components per group of 32 rows.
constexpr int JOFF = 14;
constexpr int JBLK = 32;  // rows per layout block

struct GridDims {
    int nxl;        // interior columns held by this rank (slab along x)
    int ny;         // interior rows
    int pitch;      // padded rows per column (a column of an NC-component state is NC*pitch doubles)
    size_t plane;   // (nxl+4)*pitch: doubles of one component of a state
};

// index of component 0 of cell (i,j) of an NC-component state; component k is KSTR doubles further
constexpr size_t KSTR = JBLK;
template <int NC>
__host__ __device__ inline size_t blocked_index(const GridDims& g, int i, int j) {
    const int jj = JOFF + j;
    return (size_t)i * (size_t)(NC * g.pitch) + (size_t)((jj >> 5) * (NC * JBLK) + (jj & (JBLK - 1)));
}
constexpr size_t MHD_KSTR = KSTR;
__host__ __device__ inline size_t mhd_index(const GridDims& g, int i, int j) { return blocked_index<NV>(g, i, j); }
// doubles per column of an nvar-component state: halo messages and ghost copies move whole columns
__host__ __device__ inline size_t column_doubles(const GridDims& g, int nvar) { return (size_t)nvar * (size_t)g.pitch; }
__host__ __device__ inline size_t mhd_column_doubles(const GridDims& g) { return column_doubles(g, NV); }

// Columns per thread of an x sweep.  A block runs strip+1 loop iterations after a warm-up worth about half of one;
// all blocks of a launch cost the same, so the launch takes ceil(blocks / resident) rounds of (strip + 1.5)
// columns.  Picks the
// strip in [min_strip, max_strip] that minimises that product: e.g. a 2048-column slab (16384^2 on 8 GPUs) with
// 128-column strips is 2048 blocks = 3.46 rounds of the 592 resident blocks (the 4th round 46 % full); 90-column
// strips make it 2944 blocks = 4.97 rounds of shorter blocks, 12 % less time.  The decomposition does not touch
// the arithmetic (tests/test_mhd_gpu.py compares different strip lengths bit for bit).
// min_strip: small grids are latency-bound (a thread spends ~3.7 us per column when its warp has a scheduler to
// itself), so while everything still fits one round shorter strips -- more threads, fewer columns each -- win
// despite the 3 warm-up cells: OT 256^2 x sweep 42 us with 8-column strips.
inline int choose_x_strip(int row_blocks, int nxl, int max_strip, int resident_blocks, int min_strip = 8) {
    long best_cost = -1;
    int best = max_strip;
    for (int s = min_strip; s <= max_strip; ++s) {
        const long blocks = (long)row_blocks * ((nxl + s - 1) / s);
        const long rounds = (blocks + resident_blocks - 1) / resident_blocks;
        const long cost = rounds * (2 * s + 3);
        if (best_cost < 0 || cost < best_cost || (cost == best_cost && s > best)) {
            best_cost = cost;
            best = s;
        }
    }
    return best;
}

// What a sweep needs besides the state.  The HOST knows the first six fields when it enqueues a step (SweepArgs,
// kernel parameters); the rest are the step's scalars, computed on the device by the prepare kernel and read from
// the context's __constant__ slot (step_ctl.cuh) -- the kernels assemble a SweepParams from both.
struct SweepArgs {
    double gamma, gm1;
    int first_step;      // 1: limiter's "previous primitives" are all zero (MUSCL_NEW.cpp:676-677, :845)
    int bc_lo, bc_hi;    // physical boundary type at the low/high end of the sweep axis, or -1 when the
                         // end is an inter-rank slab edge (ghost columns then hold neighbour data)
    int slot;            // which c_step[] entry holds this context's scalars
};
struct SweepParams {
    double gamma, gm1;
    double inv2ch, halfch;  // 1.0/(2.0*ch) and ch/2.0 (MUSCL_NEW.cpp:126-127)
    double ch, ch2;      // GLM wave speed = max signal speed of this step (MUSCL_NEW.cpp:834)
    double dtdh;         // dt/dx (x sweep) or dt/dy (y sweep)
    double hdtdh;        // 0.5*(dt/dx)
    double psi_factor;   // exp(-(0.5 dt) ch / 0.18), glibc's value through the host-made table (MUSCL_NEW.cpp:66-69)
    int first_step;
    int bc_lo, bc_hi;
};

// Head of a step on the stream: (1) the one-thread prepare kernel -- consumes status[0] (max signal speed of the
// current state), forms the step's scalars in ctl->sc, accumulates and clears the status words, evaluates the
// pause conditions (step_ctl.cuh); (2) ctl->sc -> c_step[slot] of the translation unit that holds the sweeps
// (`relaxed`: mhd_kernels_relaxed.cu keeps its own copy).
void launch_prepare(StepCtl* ctl, const StepCmd* cmd, unsigned long long* status, double* log_ring, cudaStream_t s);
void launch_max_words(unsigned long long* words, int n, cudaStream_t s);  // words[n] = max(words[0..n))
cudaError_t publish_step_scalars(const StepCtl* ctl, int slot, cudaStream_t s);
cudaError_t publish_step_scalars_relaxed(const StepCtl* ctl, int slot, cudaStream_t s);

// status[0] = bit pattern of the max signal speed (y sweep), status[1] / status[2] = number of x / y sweep
// blocks in which a fast division saw an operand outside its guarded window; those block indices are in
// recomputed with IEEE division by the fix kernel launched right behind each sweep.  The x sweep marks them in
// x_flags (x_sweep_block_count entries, all zero on entry and again on exit), the y sweep lists them in fix_list
// (y_sweep_block_count entries, plus y_edge_block_count for the slab-edge launch, which counts in status[3]).
void launch_x_sweep(const GridDims& g, const SweepArgs& p, const double* u_cur, const double* u_prev,
                    double* u_bar, unsigned long long* status, unsigned* x_flags, cudaStream_t s);
// `part` selects the columns of one y-sweep launch (halo/compute overlap): Y_ALL, or
// Y_EDGES (local columns 0,1 and nxl-2,nxl-1: what the x neighbours need) followed by Y_INTERIOR (the rest), so
// that the halo exchange of the edge columns runs while the interior is still being swept.  Needs nxl >= 8.
enum : int { Y_ALL = 0, Y_EDGES = 1, Y_INTERIOR = 2 };
struct YCols {
    int c0, jump_at, jump;  // local column of block row by: c0 + by + (by >= jump_at ? jump : 0)
    int ncols;              // block rows of the launch
    int counter;            // status word that counts this launch's flagged blocks (2, slab edges: 3)
    int pf_cols = 0;        // columns between a block and the one whose tile it asks L2 to fetch (0: none)
};
inline YCols y_columns(const GridDims& g, int part) {
    if (part == Y_EDGES) return YCols{0, 2, g.nxl - 4, 4, 3};
    if (part == Y_INTERIOR) return YCols{2, 0x7fffffff, 0, g.nxl - 4, 2};
    return YCols{0, 0x7fffffff, 0, g.nxl, 2};
}
void launch_y_sweep(const GridDims& g, const SweepArgs& p, const double* u_bar, const double* u_cur,
                    double* u_next, unsigned long long* status, unsigned* fix_list, cudaStream_t s,
                    int part = Y_ALL);
// relaxed-arithmetic instantiations (mhd_kernels_relaxed.cu: FMA contraction, 1-multiply quotients; no fix pass)
void launch_x_sweep_relaxed(const GridDims& g, const SweepArgs& p, const double* u_cur, const double* u_prev,
                            double* u_bar, unsigned* x_flags, cudaStream_t s);
void launch_y_sweep_relaxed(const GridDims& g, const SweepArgs& p, const double* u_bar, const double* u_cur,
                            double* u_next, unsigned long long* status, unsigned* fix_list, cudaStream_t s,
                            int part = Y_ALL);
long x_sweep_block_count(const GridDims& g);
long y_sweep_block_count(const GridDims& g);
long y_edge_block_count(const GridDims& g);
void launch_x_ghosts(const GridDims& g, int bc_x, double* u, cudaStream_t s);
// transmissive image of ONE side (0 = low, 1 = high): end ranks of a non-periodic slab decomposition
void launch_x_ghosts_side(const GridDims& g, int hi_side, double* u, cudaStream_t s);
void launch_max_speed(const GridDims& g, double gamma, const double* u, int i_lo, int i_hi, int j_lo, int j_hi,
                      unsigned long long* max_bits, cudaStream_t s);
void launch_aos_to_soa(const GridDims& g, const double* aos, int i0, int ni, double* soa, cudaStream_t s);
void launch_soa_to_aos(const GridDims& g, const double* soa, int i0, int ni, double* aos, cudaStream_t s);

}  // namespace lsc

// euler_kernels.cuh -- launch-side declarations of the Euler SLIC kernels.
#pragma once
#include "mhd_kernels.cuh"

namespace lsc {

enum : int { BC_TRANSMISSIVE_E = 3 };  // G:199-215: ghosts 0,1 <- 2 ; n+2,n+3 <- n+1

// per-step scalars of the Euler loop (S:600-603), formed on the device by euler_prepare_kernel and read by the
// sweeps from the context's __constant__ slot -- the Euler counterpart of StepScalars (step_ctl.cuh)
struct EulerStepScalars {
    double dtdx, hdtdx;   // dt/dx, 0.5*(dt/dx)
    double dtdy, hdtdy;
    double hdxdt, hdydt;  // 0.5*dx/dt, 0.5*dy/dt            (S:129, S:156)
    int halt, pad;
};

// device-resident loop state of an Euler context (no psi factor, no dump cadence: S:600-617 / G:627-646)
struct EulerStepCtl {
    EulerStepScalars sc;
    double t, dt, max_speed;
    long long step;
    unsigned long long redo_total;
    unsigned long long batch_seen;
    int halt, log_count;
    double cfl, dspace, dx, dy;  // S:537, S:166-169, S:546-547
};

struct EulerParams {
    double gamma, gm1;
    double cm, cp;        // 0.5*(1+omega), 0.5*(1-omega)   (S:270)
    // the six step scalars: filled in by the kernels from c_estep[slot] (the host does not know dt)
    double dtdx, hdtdx;   // dt/dx, 0.5*(dt/dx)
    double dtdy, hdtdy;
    double hdxdt, hdydt;  // 0.5*dx/dt, 0.5*dy/dt            (S:129, S:156)
    int slot;             // which c_estep[] entry holds this context's step scalars
    int bcx_lo, bcx_hi;   // physical x boundary type at each end of this slab, -1 at an inter-rank edge
    int bcy;
    // tabulated EoS of MHD_codes/TabEoS.cpp (device pointers; tab_dens == nullptr: ideal gas, S / G)
    const double* tab_dens;   // [500]
    const double* tab_pres;   // [500]
    const double* tab_sound;  // [500][500]  [pressure][density]
    const double* tab_en;     // [500][500]
    int* tab_hints;           // 3 ints per y-sweep block: search brackets carried from step to step (or nullptr)
    int tab_warm;             // 1: the table passed eos_table_is_monotone -> warm-started searches (TableEos)
};

// head of a step: dt = C*min(dx,dy)/max (S:192), t += dt (S:602), the sweeps' scalars; pause once t >= t_stop
// (S:617); status[0] is consumed and the status words are cleared (StepCmd / halt codes: step_ctl.cuh)
void launch_euler_prepare(EulerStepCtl* ctl, const StepCmd* cmd, unsigned long long* status, double* log_ring,
                          cudaStream_t s);
cudaError_t publish_euler_step_scalars(const EulerStepCtl* ctl, int slot, cudaStream_t s);

// status / fix_list as in mhd_kernels.cuh: [0] max speed bits, [1]/[2] x/y blocks recomputed with IEEE division
void launch_euler_x_sweep(const GridDims& g, const EulerParams& p, const double* cur, double* ubar,
                          unsigned long long* status, unsigned* fix_list, cudaStream_t s);
// part: Y_ALL, or Y_EDGES then Y_INTERIOR for the halo/compute overlap (mhd_kernels.cuh); the edge launch counts its
// flagged blocks in status[3] and lists them behind the euler_fix_list_entries of the other launches
void launch_euler_y_sweep(const GridDims& g, const EulerParams& p, const double* ubar, const double* cur, double* next,
                          unsigned long long* status, unsigned* fix_list, cudaStream_t s, int part = Y_ALL);
long euler_fix_list_entries(const GridDims& g);
long euler_edge_list_entries(const GridDims& g);
long euler_y_block_count(const GridDims& g);  // blocks of one y-sweep launch (size of EulerParams::tab_hints / 3)
void launch_euler_max_speed(const GridDims& g, const EulerParams& p, const double* u, int i_lo, int i_hi, int j_lo, int j_hi,
                            unsigned long long* status, cudaStream_t s);
void launch_euler_x_ghosts(const GridDims& g, int bc_x, double* u, cudaStream_t s);
void launch_euler_x_ghosts_side(const GridDims& g, int hi_side, double* u, cudaStream_t s);
void launch_euler_aos_to_soa(const GridDims& g, const double* aos, int i0, int ni, double* soa, cudaStream_t s);
void launch_euler_soa_to_aos(const GridDims& g, const double* soa, int i0, int ni, double* aos, cudaStream_t s);

// test hook (device pointers): out[2t] = exact cfl speed, out[2t+1] = fp32 bound of speed / run_max
void launch_euler_kat_cfl_bound(long n, const double* U, double gamma, double run_max, double* out);

}  // namespace lsc

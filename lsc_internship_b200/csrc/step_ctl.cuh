// step_ctl.cuh -- the device-resident part of the time loop MUSCL_NEW.cpp:831-1003.
//
// ComputeTimeStep's scalars (dt, ch = max signal speed, M:117-118, :834), `t += dt` (M:835), the psi damping factor
// (M:66-69) and the loop's exit / dump predicates (M:975, :1003) are evaluated by a one-thread "prepare" kernel that
// runs at the head of every step, right behind the CFL reduction fused into the previous y sweep (and, across ranks,
// behind its max-allreduce).  The sweeps read the scalars from a per-context slot of __constant__ memory
// (uniform-register operands, no per-thread registers), which is refreshed on the stream with
// cudaMemcpyToSymbolAsync(device -> device).  The host therefore enqueues whole batches of steps without ever
// reading anything back; it synchronises once per batch (once per lsc_mhd_step call, once per dump or stop in
// lsc_mhd_run_until).  Steps enqueued beyond a pause point are no-ops: every kernel of a step starts with
// `if (slot.halt) return`.
//
// Bit-exactness: every expression below is the reference's, in its association, in IEEE double -- division and
// multiplication round identically on the device and on the host.  The one libm call of the loop, exp() in
// psiUpdate (M:67), cannot be evaluated bit-identically on the device; its argument -((0.5 dt) ch)/0.18 equals
// -0.5*C*min(dx,dy)/0.18 up to the roundings of dt = C*ds/ch and the product, i.e. it takes one of a handful of
// adjacent doubles (exactly 2 on the survey's 2M samples).  The host precomputes
// {argument bit pattern -> glibc exp} for those candidates; prepare looks the factor up by exact bit match and, on a
// miss, halts with HALT_EXP_MISS -- the host then adds the missing pair (its own exp of that very argument) and
// resumes.  No approximation is involved at any point.
#pragma once
#include <cmath>
#include <cuda_runtime.h>

namespace lsc {

constexpr int STEP_SLOTS = 32;      // contexts per device that can step concurrently (one __constant__ slot each)
constexpr int STEP_LOG_CAP = 1024;  // per-step (t, dt, ch) records kept on the device between two host syncs
constexpr int EXP_TAB_CAP = 64;

enum : int {
    HALT_NONE = 0,
    HALT_PAUSE = 1,     // the step just finished hit the stop time (M:1003) or a dump is due (M:975)
    HALT_LIMIT = 2,     // the requested number of steps is done
    HALT_NAN = 3,       // max signal speed not finite / not positive
    HALT_EXP_MISS = 4,  // psi-damping argument not in the table: the host supplies exp() and resumes
    HALT_LOG_FULL = 5,  // the per-step log ring is full: the host drains it and resumes
};

// per-step scalars the sweeps consume; one copy per context in __constant__ memory (c_step[slot])
struct StepScalars {
    double ch, ch2;          // GLM wave speed = max signal speed (M:834), ch*ch
    double inv2ch, halfch;   // 1.0/(2.0*ch), ch/2.0 (M:126-127)
    double dtdx, hdtdx;      // dt/dx, 0.5*(dt/dx) (M:867, :891)
    double dtdy, hdtdy;
    double psi_factor;       // exp(-(0.5 dt) ch / 0.18) (M:66-69)
    int halt, pad;
};

// host -> device command block, written (stream-ordered) before a batch of steps is enqueued
struct StepCmd {
    unsigned long long batch;  // changes with every batch: the first prepare of a batch clears `halt`
    long long step_limit;      // halt (HALT_LIMIT) once this many steps are done in total
    double t_stop;             // pause after a step that ends with t >= t_stop (+inf: never)
    int dump_counter;          // pause after a step with t > 0.1*counter && t - dt <= 0.1*counter (M:975); < 0: off
    int keep_log;              // record (t, dt, ch) of every step in the log ring
};

struct StepCtl {
    StepScalars sc;
    double t, dt, max_speed;
    long long step;                    // steps started (== completed once the stream is idle)
    unsigned long long redo_total;     // thread blocks the IEEE fix kernels recomputed so far
    unsigned long long batch_seen;
    unsigned long long miss_arg;       // HALT_EXP_MISS: bit pattern of the argument that is not in the table
    int halt, log_count;
    // constants of the run (M:659-670)
    double cfl, dspace, dx, dy;
    int ntab, pad;
    unsigned long long tab_arg[EXP_TAB_CAP];
    double tab_val[EXP_TAB_CAP];
};

// M:975 and M:1003 for the step that just ended at time t after an increment dt
__host__ __device__ inline bool pause_due(double t, double dt, double t_stop, int dump_counter) {
    if (dump_counter >= 0 && t > 0.1 * dump_counter && t - dt <= 0.1 * dump_counter) return true;
    return !(t < t_stop);
}

// psiUpdate's argument (M:67 with dt := 0.5*dt, M:900): ((-1.0*dt')*ch)/0.18
__host__ __device__ inline double psi_argument(double dt, double ch) { return -1.0 * (0.5 * dt) * ch / 0.18; }

}  // namespace lsc

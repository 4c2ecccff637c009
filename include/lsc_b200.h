/*
 * lsc_b200.h -- C ABI of the B200-native 2-D finite-volume hot path
 * (liblsc_b200.so).  Plain C, no CUDA/torch types, no exceptions cross it.
 *
 * The reference (JosephineMellor/LSC_internship) has NO callable interface: each
 * solver is a single main() whose "API" is a block of constants, a commented-in
 * initial-condition block, commented-in boundary loops, a stdout log line and
 * .dat files.  Each entry point below therefore cites the span of the reference
 * main() it replaces (paths relative to the reference root):
 *
 *   M = MHD_codes/MUSCL_NEW.cpp      S = Euler_codes/SLIC_euler_2d.cpp
 *   G = Euler_codes/generaliseSLIC.cpp
 *
 * Conventions: every function returns 0 on success and a negative LSC_E* code on
 * failure; lsc_last_error() returns a thread-local message for the last failure.
 * Host arrays are always caller-owned and are not retained after the call
 * returns.  One host thread drives one context; different contexts are
 * independent.  There is no CPU fallback: without a CUDA device every create
 * call fails with LSC_ECUDA.
 *
 * Host state layout (the reference's, M:673-674 / S:542-543): array of structures
 * [i][j][k], i = x (slowest), j = y, k = component (fastest), with TWO ghost cells
 * on every side, i.e. (nx+4)*(ny+4)*NVAR doubles.  MHD components (M:11-19):
 * rho, rho*vx, rho*vy, rho*vz, E, Bx, By, Bz, psi.  Euler (S:10-17): rho, rho*vx,
 * rho*vy, E.
 */
#ifndef LSC_B200_H
#define LSC_B200_H

#ifdef __cplusplus
extern "C" {
#endif

#define LSC_MHD_NVAR 9
#define LSC_EULER_NVAR 4

/* error codes */
#define LSC_OK 0
#define LSC_EINVAL (-1)  /* bad argument / configuration */
#define LSC_ECUDA (-2)   /* CUDA runtime failure (incl. "no device") */
#define LSC_ENOMEM (-3)  /* device or host allocation failed */
#define LSC_ESTATE (-4)  /* call out of order (e.g. step before upload) */
#define LSC_ENAN (-5)    /* max signal speed is not finite / not positive */
#define LSC_ECOMM (-6)   /* NCCL failure */
#define LSC_ECALLBACK (-7) /* a log callback returned non-zero: the loop stopped after that step */

/* boundary types.  M:607-655 selects these by (un)commenting loops. */
#define LSC_BC_PERIODIC 0     /* M:609-614 (x), M:641-646 (y); S:205-225 */
#define LSC_BC_REFLECTIVE 1   /* M:625-638 (y only: mirrors cells 2,3 / ny,ny+1, negates rho*vy and By) */
#define LSC_BC_TRANSMISSIVE 2 /* M:617-622 (x), M:649-654 (y): ghost 0,1 <- 2,3 ; n+2,n+3 <- n,n+1 */
#define LSC_BC_TRANSMISSIVE_E 3 /* Euler only, G:199-215: ghost 0,1 <- 2 ; n+2,n+3 <- n+1 */

/* initial conditions the reference ships as commented blocks */
#define LSC_MHD_IC_KELVIN_HELMHOLTZ 0 /* M:800-808 (the committed one) */
#define LSC_MHD_IC_ORSZAG_TANG 1      /* M:811-819 */

/* Replaces the constants at the top of main(), M:659-671. */
typedef struct lsc_mhd_config {
    int nx, ny;            /* M:659-660  interior cells of the WHOLE grid */
    double x0, x1, y0, y1; /* M:661-664 */
    double gamma;          /* M:668 */
    double cfl;            /* M:667  (C) */
    int bc_x;              /* LSC_BC_PERIODIC | LSC_BC_TRANSMISSIVE */
    int bc_y;              /* LSC_BC_REFLECTIVE | LSC_BC_PERIODIC | LSC_BC_TRANSMISSIVE */
    int device;            /* CUDA device ordinal this context runs on */
    /* slab decomposition along x (one context per GPU).  rank r of nranks owns interior
     * columns [r*nx/nranks, (r+1)*nx/nranks) plus two halo columns per side.  nranks = 1:
     * the whole grid.  For nranks > 1 call lsc_mhd_comm_init() before upload. */
    int rank, nranks;
    /* optional explicit slab bounds (uneven decompositions, e.g. sized by a calibration pass): this rank owns
     * interior columns [slab_i0, slab_i0 + slab_nx).  slab_nx = 0 (the default of a zero-initialised struct):
     * the even split above.  The caller makes the slabs of the ranks tile [0, nx) in rank order. */
    int slab_i0, slab_nx;
} lsc_mhd_config;

typedef struct lsc_mhd_ctx lsc_mhd_ctx;

/* Called at the dump cadence of M:975 (t crossed 0.1*counter) from lsc_mhd_run_until. */
typedef int (*lsc_mhd_dump_cb)(lsc_mhd_ctx* ctx, int counter, double t, double dt, void* user);
/* Called once per step with the values of the log line M:836 ("t= .. dt= .. ch= ..").  Return 0 to continue; any
 * other value stops the loop (lsc_mhd_run_until then returns LSC_ECALLBACK). */
typedef int (*lsc_mhd_log_cb)(long step, double t, double dt, double ch, void* user);

const char* lsc_last_error(void);
int lsc_device_count(void);

/* pinned host memory helpers for callers that want full-rate PCIe transfers */
int lsc_host_alloc(void** ptr, unsigned long long bytes);
int lsc_host_free(void* ptr);

/* allocate the 15 work arrays of M:673-713 -- here: 3 device-resident states (blocked component-major layout, csrc/mhd_kernels.cuh) */
int lsc_mhd_create(const lsc_mhd_config* cfg, lsc_mhd_ctx** out);
int lsc_mhd_destroy(lsc_mhd_ctx* ctx);

/* number of interior columns / first global interior column of this rank's slab */
int lsc_mhd_local_nx(const lsc_mhd_ctx* ctx);
int lsc_mhd_local_i0(const lsc_mhd_ctx* ctx);

/* Fill a host AoS array [(nxl+4)][(ny+4)][9] with this rank's slab of the initial condition of
 * M:719-825 (cell centres x0+(i-0.5)dx incl. ghosts, host libm, PrimitiveToConservative M:9-21). */
int lsc_mhd_initial_condition(const lsc_mhd_ctx* ctx, int problem, double* aos);

/* Set the state (replaces the IC loop's stores, M:822).  aos holds this rank's slab incl. its
 * ghost/halo cells, (nxl+4)*(ny+4)*9 doubles.  Resets t = 0, step = 0. */
int lsc_mhd_upload(lsc_mhd_ctx* ctx, const double* aos);

/* Generate the initial condition slab-chunk by slab-chunk and upload it without ever holding the
 * whole AoS array on the host (the 16384^2 case: 19 GB).  Equivalent to
 * lsc_mhd_initial_condition + lsc_mhd_upload. */
int lsc_mhd_init_problem(lsc_mhd_ctx* ctx, int problem);

/* Arithmetic of the two sweeps.  LSC_ARITH_REFERENCE (default) reproduces the reference's IEEE results bit
 * for bit (no FMA contraction, correctly rounded division) and is the only mode the parity tests accept.
 * LSC_ARITH_RELAXED runs the same kernels compiled with FMA contraction and 1-multiply
 * quotients: faster, deviates from the reference at rounding level per operation (the scheme amplifies that,
 * see tests/test_mhd_gpu.py::test_relaxed_arithmetic_deviation for the measured norms). */
#define LSC_ARITH_REFERENCE 0
#define LSC_ARITH_RELAXED 1
int lsc_mhd_set_arithmetic(lsc_mhd_ctx* ctx, int mode);

/* nsteps passes of the do{}while body M:831-971 (CFL -> x sweep -> y sweep -> copy). */
int lsc_mhd_step(lsc_mhd_ctx* ctx, long nsteps);

/* The loop M:831-1003: steps until t >= t_stop (no clamping of the last step; t overshoots),
 * firing dump_cb when `t > 0.1*counter && t-dt <= 0.1*counter` (M:975) and log_cb every step
 * (M:836).  Either callback may be NULL.  *steps_out (may be NULL) receives the step count. */
int lsc_mhd_run_until(lsc_mhd_ctx* ctx, double t_stop, lsc_mhd_dump_cb dump_cb, lsc_mhd_log_cb log_cb,
                      void* user, long* steps_out);

/* Single-process slab group -- how ONE program (the reference is one process, M:658-1034) uses the 8 GPUs of a box:
 * `n` contexts created with nranks = n, rank = 0..n-1 (cfg.device per context: one GPU each, or several slabs on one
 * GPU) are advanced in lock step by the calling thread.  Halo columns move by peer copies between the contexts'
 * buffers over NVLink; the signal speed is max-reduced through 8-byte peer copies and a one-thread kernel on the
 * first context's GPU.  Everything is stream-ordered by events: like the plain calls, a group call synchronises with
 * the host once per batch of steps, not per step.  Same kernels and slab-edge code as the NCCL mode; results are
 * bit-identical to the single-context run for any n.  lsc_mhd_group_run_until hands ctxs[0] to the callbacks. */
int lsc_mhd_group_step(lsc_mhd_ctx** ctxs, int n, long nsteps);
int lsc_mhd_group_step_timed(lsc_mhd_ctx** ctxs, int n, long nsteps, float* ms);
int lsc_mhd_group_run_until(lsc_mhd_ctx** ctxs, int n, double t_stop, lsc_mhd_dump_cb dump_cb, lsc_mhd_log_cb log_cb,
                            void* user, long* steps_out);
/* lsc_mhd_download for a group: the WHOLE grid [(nx+4)][(ny+4)][9] assembled from the slabs, ghost ring as the
 * reference's `u` holds it after M:971 -- what lsc_mhd_write_dat takes. */
int lsc_mhd_group_download(lsc_mhd_ctx** ctxs, int n, double* aos);

/* The time loop is device-resident (csrc/step_ctl.cuh): dt, ch, t, the psi factor and the loop predicates of
 * M:831-1003 are formed on the device and the host synchronises once per batch of steps.  *host_syncs = number of
 * batches so far; *exp_table_misses = how often the psi-damping argument was not in the host-made {argument ->
 * glibc exp} table and the host had to supply it (each miss costs one extra synchronisation, never accuracy). */
int lsc_mhd_loop_stats(const lsc_mhd_ctx* ctx, long* host_syncs, long* exp_table_misses);

/* As lsc_mhd_step, bracketed by CUDA events on the context's own stream; *ms receives the device
 * time of the whole call (kernels, the per-step 8-byte readback of the signal speed, ghost/halo fill). */
int lsc_mhd_step_timed(lsc_mhd_ctx* ctx, long nsteps, float* ms);

/* Per-kernel device timing for roofline reporting: when enabled, every x-sweep and y-sweep launch is
 * bracketed by CUDA events on the launching stream.  lsc_mhd_kernel_times sums and clears them:
 * ms[0] = x sweeps, ms[1] = y sweeps, launches[0..1] = how many of each. */
int lsc_mhd_set_profiling(lsc_mhd_ctx* ctx, int enabled);
int lsc_mhd_kernel_times(lsc_mhd_ctx* ctx, double ms[2], long launches[2]);
/* with profiling enabled and nranks > 1: device time of the NCCL halo exchanges (ms[0], calls[0]) and of the
 * max-allreduces (ms[1], calls[1]) since the last call (CUDA events on the solver's stream) */
int lsc_mhd_comm_times(lsc_mhd_ctx* ctx, double ms[2], long calls[2]);
/* Halo/compute overlap (default on).  With nranks > 1 and a slab of >= 8 columns a step
 * sweeps the four slab-edge columns in y first and issues their NCCL exchange on a second, high-priority stream
 * while the interior columns are still being swept; the next call that needs the ghost columns waits for it on
 * the device.  0 restores the serial order (y sweep, then exchange); 2 splits the y sweep on a single rank too
 * (the ghost-column kernel then takes the place of the exchange; for tests).  Bit-identical in every mode. */
int lsc_mhd_set_overlap(lsc_mhd_ctx* ctx, int enabled);

/* How many thread blocks the device-side IEEE fix kernels had to recompute so far because a fast
 * shared-reciprocal division saw an operand outside its guarded exponent window (csrc/fp64_div.cuh; e.g. the
 * denormal momenta at a numerical front).  0 for smooth problems such as Orszag-Tang / Kelvin-Helmholtz. */
long lsc_mhd_ieee_fallback_count(const lsc_mhd_ctx* ctx);

/* t, the last dt, the last ch (= max signal speed, M:834) and the step count. Any may be NULL. */
int lsc_mhd_get_scalars(const lsc_mhd_ctx* ctx, double* t, double* dt, double* ch, long* step);

/* Copy this rank's slab back in the reference layout, ghost ring included, with exactly the ghost
 * contents the reference's `u` holds after M:971 (boundary images; the four never-written corner
 * cells [nx+2..nx+3][ny+2..ny+3] are zero once a step has been taken).
 * nranks > 1: the slab's interior and halo columns as they are on the device, y ghost rows as plain boundary images
 * over ALL nxl+4 columns -- the reference's loop bounds (i < nx+2) and zero corners belong to the assembled grid:
 * copy the interior columns of every slab into one [(nx+4)][(ny+4)][9] array and call lsc_mhd_fill_ghost_ring. */
int lsc_mhd_download(lsc_mhd_ctx* ctx, double* aos);

/* Decomposition-independent 64-bit fingerprint of the INTERIOR of the current state (M:673 `u` after M:969), computed
 * on the device: sum modulo 2^64 over interior values of splitmix64(bit pattern + 0x9E3779B97F4A7C15 * (global cell
 * number * NVAR + k + 1)), global cell number = i*ny + j (0-based interior indices).  *hash receives THIS slab's
 * contribution; the fingerprint of a decomposed grid is the sum of its slabs' words modulo 2^64, and equals the
 * single-context value exactly when every interior bit agrees. */
int lsc_mhd_state_hash(lsc_mhd_ctx* ctx, unsigned long long* hash);

/* lsc_mhd_state_hash for callers that cannot allocate device memory: same input, same output word, but no
 * cudaMalloc / cudaFree (nor any other allocating routine) on the call path.  NO COUNTERPART IN THE REFERENCE.
 * The scratch words (one partial sum per thread block) live in
 * a file-scope `static __device__` array of 64 KiB + 256 B that every GPU which loads the library holds until exit;
 * it bounds the slab height: ny <= LSC_STATE_HASH_NOALLOC_MAX_NY, beyond that the call fails with LSC_EINVAL and
 * says so in lsc_last_error().  All calls on one device share the array: calls that overlap on the GPU (different
 * contexts on different host threads) race on it and return garbage -- serialise them.  Nothing is carried over
 * from one call to the next (the result word is cleared at the start of every call). */
#define LSC_STATE_HASH_NOALLOC_MAX_NY 1048576
int lsc_mhd_state_hash_noalloc(lsc_mhd_ctx* ctx, unsigned long long* hash);

/* Host helper for callers that assemble a full-grid array from slabs: rebuilds the ghost ring the
 * reference's `u` holds after M:971 (boundary images per cfg->bc_x / bc_y with the reference's loop bounds,
 * the four never-written corner cells zero) from the interior of `aos` [(nx+4)][(ny+4)][9]. */
int lsc_mhd_fill_ghost_ring(const lsc_mhd_config* cfg, double* aos);

/* Byte-compatible writer of M:989-1001 / M:1022-1032 for a FULL-grid AoS array (ghost ring
 * included, as returned by lsc_mhd_download on a single-rank context): primitive variables,
 * default 6-digit ostream formatting, blank line after each j row. */
int lsc_mhd_write_dat(const lsc_mhd_config* cfg, const double* aos, const char* path);

/* Multi-GPU: 128-byte NCCL unique id created on rank 0 and distributed by the caller
 * (torch.distributed broadcast, a file, MPI ...); every rank then joins the communicator.
 * Per step the ranks exchange two halo columns of u^{n+1} with each x neighbour and all-reduce
 * (max) the signal speed -- the only cross-rank data of the path. */
int lsc_comm_unique_id(void* id128);
int lsc_mhd_comm_init(lsc_mhd_ctx* ctx, const void* id128);

/* Test hooks: run the device flux / reconstruction functions on n states (AoS, 9 doubles each).
 * dir 0 = x, 1 = y.  Used by tests/ for known-answer parity against the oracle. */
int lsc_mhd_kat_hllc(int device, int dir, long n, const double* L, const double* R, double gamma, double ch,
                     double* F);
int lsc_mhd_kat_fluxdef(int device, int dir, long n, const double* U, double gamma, double ch, double* F);

/* Test hook for the CFL reduction's warp filter (csrc/mhd_physics.cuh: cfl_speed_upper_f32): for n states
 * returns exact_and_bound[2t] = the exact fp64 signal speed (ComputeTimeStep, MUSCL_NEW.cpp:75-109) and
 * [2t+1] = the single-precision upper bound of (signal speed / run_max) the y sweep uses to skip warps that
 * cannot raise the running maximum run_max (NaN = "not trusted, evaluate exactly"). */
int lsc_mhd_kat_cfl_bound(int device, long n, const double* U, double gamma, double run_max,
                          double* exact_and_bound);

/* Test hook for the shared-reciprocal fp64 division used by the kernels (csrc/fp64_div.cuh): divides
 * nblocks*256*per_thread pseudo-random operand pairs both ways on the device and counts quotients that
 * differ bitwise from the compiler's IEEE `/` (must be 0) and pairs the guard flagged for the slow path.
 * mode 0: solver-range operands, 1: arbitrary bit patterns, 2: half-ulp / all-ones-mantissa adversaries,
 * 3: tiny / denormal / zero dividends.  mode + 10: the same for the branching policy (BranchDiv), which never
 * flags -- all of its quotients must equal IEEE.  mode + 20 (modes 0-2): the guarded branch-free square root
 * FastDiv::sqrt against sqrt() -- solver-range operands, arbitrary bit patterns, perfect squares +- 1 ulp. */
int lsc_kat_division(int device, unsigned long long seed, int mode, long nblocks, int per_thread,
                     unsigned long long* mismatches, unsigned long long* flagged);

/* ------------------------------------------------------------------------------------------------
 * 2-D Euler SLIC (S = Euler_codes/SLIC_euler_2d.cpp, G = Euler_codes/generaliseSLIC.cpp).
 * One step = dt (S:165-195) -> BC -> A = XthenY(u) (S:230-376) -> B = YthenX(u) (S:379-525; its second
 * sweep is dead in the reference, S:510 vs S:519) -> u = 0.5*(A+B) (S:609-615) -> BC.
 * ---------------------------------------------------------------------------------------------- */
#define LSC_EULER_IC_ORSZAG_TANG 0 /* S:585-589 (committed in S) */
#define LSC_EULER_IC_QUADRANTS 1   /* G:594-617 (committed in G) */
#define LSC_EULER_IC_CYLINDER 2    /* cylindrical explosion of BASELINE config 3: rho,p = 1,1 inside
                                      (x-.5)^2+(y-.5)^2 <= 0.2^2, 0.125,0.1 outside, at rest (not in the reference) */

#define LSC_EULER_IC_TABEOS 3       /* T:918-941: four quadrants at rest, rho = 1, p = 0.01e5 (x > 0.5) / 1000e5;
                                      PrimativeToConservative through the EoS table (needs lsc_euler_set_eos_table) */

/* Replaces the constants S:529-540 / G:565-577. */
typedef struct lsc_euler_config {
    int nx, ny;
    double x0, x1, y0, y1;
    double gamma; /* S:538 */
    double cfl;   /* S:537 */
    double omega; /* S:539 */
    int bc;       /* LSC_BC_PERIODIC (S:198-226) | LSC_BC_TRANSMISSIVE_E (G:199-215), both axes */
    int device;
    /* slab decomposition along x, as for MHD: rank r of nranks owns interior columns
     * [r*nx/nranks, (r+1)*nx/nranks) plus two halo columns per side; every host array of a context with
     * nranks > 1 is that slab, [(nxl+4)][(ny+4)][4].  nranks <= 1: the whole grid. */
    int rank, nranks;
} lsc_euler_config;

typedef struct lsc_euler_ctx lsc_euler_ctx;
typedef int (*lsc_euler_log_cb)(long step, double t, double dt, void* user); /* G:631 "t = .. dt = .."; non-zero stops the loop */

int lsc_euler_create(const lsc_euler_config* cfg, lsc_euler_ctx** out);
int lsc_euler_destroy(lsc_euler_ctx* ctx);
int lsc_euler_local_nx(const lsc_euler_ctx* ctx);
int lsc_euler_local_i0(const lsc_euler_ctx* ctx);
/* one process per GPU: join the NCCL communicator (id from lsc_comm_unique_id on rank 0).  Per step the ranks
 * exchange two halo columns of the new u with each x neighbour and all-reduce (max) the signal speed. */
int lsc_euler_comm_init(lsc_euler_ctx* ctx, const void* id128);
/* single process: advance the n slabs of one grid (contexts with rank 0..n-1 of n, on any devices of this
 * process) in lock step, halos by peer copies, max on the host -- same kernels and slab-edge path as NCCL. */
int lsc_euler_group_step(lsc_euler_ctx** slabs, int n, long nsteps);
/* lsc_euler_run_until / lsc_euler_download for the slabs of one process: the loop S:600-617 / G:627-646 in lock step,
 * and the WHOLE grid [(nx+4)][(ny+4)][4] assembled with the ghost ring u holds after S:616 (what lsc_euler_write_dat
 * takes).  This is how drivers/slic_euler_2d and drivers/tab_eos use several GPUs (--gpus N). */
int lsc_euler_group_run_until(lsc_euler_ctx** slabs, int n, double t_stop, lsc_euler_log_cb log_cb, void* user,
                              long* steps_out);
int lsc_euler_group_download(lsc_euler_ctx** slabs, int n, double* aos);
/* IC loop S:552-594 / G:586-620 followed by the boundary fill S:596, into [(nx+4)][(ny+4)][4] (the slab
 * [(nxl+4)][(ny+4)][4] of cfg->rank when cfg->nranks > 1). Host libm. */
int lsc_euler_initial_condition(const lsc_euler_config* cfg, int problem, double* aos);
/* The first ComputeTimeStep sees the array as given (S:173-174); the x ghost columns are then replaced by boundary
 * images -- for a slab (nranks > 1) by the neighbours' cells / the images, at the head of the first step -- exactly
 * as S:604 does, so the caller's ghost columns may hold anything (e.g. the initial data's own values). */
int lsc_euler_upload(lsc_euler_ctx* ctx, const double* aos);
int lsc_euler_init_problem(lsc_euler_ctx* ctx, int problem);
int lsc_euler_step(lsc_euler_ctx* ctx, long nsteps);                 /* n passes of the do-body S:600-616 */
int lsc_euler_step_timed(lsc_euler_ctx* ctx, long nsteps, float* ms);
/* do { step } while (t < t_stop)  -- S:600-617 with t_stop = tStop/(21-counter), G:627-646 with tStop. */
int lsc_euler_run_until(lsc_euler_ctx* ctx, double t_stop, lsc_euler_log_cb log_cb, void* user, long* steps_out);
int lsc_euler_get_scalars(const lsc_euler_ctx* ctx, double* t, double* dt, long* step);
/* the Euler loop is device-resident like the MHD one (lsc_mhd_loop_stats): host synchronisations so far */
int lsc_euler_loop_stats(const lsc_euler_ctx* ctx, long* host_syncs);
int lsc_euler_download(lsc_euler_ctx* ctx, double* aos);              /* u after S:616, ghost ring included */
/* S:621-642 (blank_line_per_row = 0, euler_<k>.dat) / G:645-667 (= 1, SLIC2d.dat) */
int lsc_euler_write_dat(const lsc_euler_config* cfg, const double* aos, const char* path, int blank_line_per_row);
long lsc_euler_ieee_fallback_count(const lsc_euler_ctx* ctx);
int lsc_euler_state_hash(lsc_euler_ctx* ctx, unsigned long long* hash); /* as lsc_mhd_state_hash, NVAR = 4 */

/* ------------------------------------------------------------------------------------------------
 * Tabulated equation of state (T = MHD_codes/TabEoS.cpp; SURVEY.md 8f row N2, BASELINE config 5): the same SLIC
 * sweeps as G with every pressure / sound speed taken from a 500x500 table instead of the ideal-gas formulas:
 * ConservativeToPrimative T:248-327 (two bisections + reverse bilinear interpolation) inside fluxX/Y_def
 * T:408-438, SoundSpeed T:330-404 inside ComputeTimeStep T:494-526.  cfg.gamma is not used on this path.
 * Tables are [pressure index][density index], 500 x 500 doubles; axes strictly ascending.
 * ---------------------------------------------------------------------------------------------- */
/* Plasma19(), T:18-133: parse a table file in the layout the reference reads (line 10 densities, 12 pressures,
 * 14-513 sound speed, 515-1014 specific internal energy; the remaining blocks are unused by the solver). */
int lsc_eos_table_read(const char* path, double* densities, double* pressures, double* sound_speeds,
                       double* energies);
/* replaces the globals T:135-144.  densities == NULL switches the context back to the ideal gas. */
int lsc_euler_set_eos_table(lsc_euler_ctx* ctx, const double* densities, const double* pressures,
                            const double* sound_speeds, const double* energies);
/* IC loop T:912-952 + boundary fill T:954 into [(nx+4)][(ny+4)][4] (or the slab of cfg->rank) */
/* The table searches of T (T:257-275, :291-299, :366-376) are bisections.  When the table passes the library's
 * monotonicity check (axes strictly ascending -- required -- and energy columns ascending by more than the rounding
 * of the density interpolation) the bracket a bisection ends in is unique, and the kernels find it by first trying
 * the bracket of the thread's previous lookup (bit-identical results, ~2.5x faster); otherwise, or after
 * lsc_euler_set_eos_search(ctx, 0), every search is the reference's bisection. */
int lsc_euler_set_eos_search(lsc_euler_ctx* ctx, int warm_start);
int lsc_euler_eos_search_is_warm(const lsc_euler_ctx* ctx);
int lsc_euler_tabeos_initial_condition(const lsc_euler_config* cfg, const double* densities, const double* pressures,
                                       const double* energies, double* aos);
/* T:974-999: SLIC2d.dat through the table's ConservativeToPrimative, blank line after each j row */
int lsc_euler_tabeos_write_dat(const lsc_euler_config* cfg, const double* aos, const char* path,
                               const double* densities, const double* pressures, const double* energies);
/* Test hook for the Euler y sweep's CFL filter (csrc/euler_physics.cuh: IdealEos::cfl_speed_over_max_f32), as
 * lsc_mhd_kat_cfl_bound: U holds n states of 4 doubles; exact_and_bound[2t] = ComputeTimeStep's speed + a
 * (SLIC_euler_2d.cpp:175-184), [2t+1] = the fp32 upper bound of it in units of run_max (NaN = not trusted). */
int lsc_euler_kat_cfl_bound(int device, long n, const double* U, double gamma, double run_max,
                            double* exact_and_bound);
int lsc_euler_set_profiling(lsc_euler_ctx* ctx, int enabled);
int lsc_euler_set_overlap(lsc_euler_ctx* ctx, int enabled);   /* as lsc_mhd_set_overlap: 0 serial, 1 default, 2 always */
int lsc_euler_kernel_times(lsc_euler_ctx* ctx, double ms[2], long launches[2]);

/* ------------------------------------------------------------------------------------------------
 * First-order sibling of the MHD path (H = MHD_codes/HLLC_MHD_2D.cpp): dimensionally split
 * Godunov update with the HLLC flux and mixed GLM cleaning, ONE ghost cell, periodic in x and y, Orszag-Tang.
 * Host arrays are H's: [(nx+2)][(ny+2)][9].  One step = ComputeTimeStep (H:71-120, over i in [1,nx-1], j in [1,ny-1])
 * -> boundary fill -> x update (H:739-755) -> y update (H:757-771) -> psi damping (H:776-780).
 * ---------------------------------------------------------------------------------------------- */
typedef struct lsc_hllc2d_config {
    int nx, ny;            /* H:616-617 */
    double x0, x1, y0, y1; /* H:618-621 */
    double gamma;          /* H:625 */
    double cfl;            /* H:624 (C = 0.75 as committed) */
    int device;
} lsc_hllc2d_config;
typedef struct lsc_hllc2d_ctx lsc_hllc2d_ctx;
int lsc_hllc2d_create(const lsc_hllc2d_config* cfg, lsc_hllc2d_ctx** out);
int lsc_hllc2d_destroy(lsc_hllc2d_ctx* ctx);
int lsc_hllc2d_initial_condition(const lsc_hllc2d_config* cfg, double* aos); /* IC loop H:642-721, host libm */
int lsc_hllc2d_upload(lsc_hllc2d_ctx* ctx, const double* aos);
int lsc_hllc2d_init_problem(lsc_hllc2d_ctx* ctx);
int lsc_hllc2d_step(lsc_hllc2d_ctx* ctx, long nsteps);                        /* n passes of H:728-784 */
/* do { step } while (t < t_stop), H:728-785; log_cb gets the values of the line H:738 ("t= .. dt= .. ch = ..") */
int lsc_hllc2d_run_until(lsc_hllc2d_ctx* ctx, double t_stop, lsc_mhd_log_cb log_cb, void* user, long* steps_out);
int lsc_hllc2d_get_scalars(const lsc_hllc2d_ctx* ctx, double* t, double* dt, double* ch, long* step);
int lsc_hllc2d_download(lsc_hllc2d_ctx* ctx, double* aos);  /* u after the boundary fill behind the loop, H:787 */
int lsc_hllc2d_write_dat(const lsc_hllc2d_config* cfg, const double* aos, const char* path); /* MHD.dat, H:793-814 */

/* Debug aid (compute-sanitizer substitute): with LSC_GUARD=1 in the environment at create time, every device buffer
 * of a context sits between two 64 KiB guard zones; *corrupted_bytes receives the number of guard bytes a kernel has
 * overwritten so far (0 = no out-of-bounds write), *guarded_buffers (may be NULL) how many buffers are watched. */
int lsc_guard_check(unsigned long long* corrupted_bytes, unsigned long long* guarded_buffers);

/* number of kernels this library has launched in this process (for bench.py's gpu_launches) */
unsigned long long lsc_kernel_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* LSC_B200_H */

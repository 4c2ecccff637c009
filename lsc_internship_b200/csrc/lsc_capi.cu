// lsc_capi.cu -- the C ABI of include/lsc_b200.h: contexts, transfers and the time loop
// of MHD_codes/MUSCL_NEW.cpp main() (:658-1034) driving the sm_100a kernels.
#include <atomic>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

#include <cuda_runtime.h>

#include "../../include/lsc_b200.h"
#include "lsc_comm.h"
#include "lsc_common.h"
#include "mhd_host.h"
#include "mhd_kernels.cuh"
#include "state_hash.h"

namespace {

thread_local char g_err[512] = "";
std::atomic<unsigned long long> g_launches{0};

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
    return code;
}

#define CUDA_TRY(expr)                                                                        \
    do {                                                                                      \
        cudaError_t e_ = (expr);                                                              \
        if (e_ != cudaSuccess)                                                                \
            return fail(LSC_ECUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

constexpr size_t STATUS_BYTES = 4 * sizeof(unsigned long long);

inline int to_kernel_bc(int bc) {
    switch (bc) {
        case LSC_BC_PERIODIC: return lsc::BC_PERIODIC;
        case LSC_BC_REFLECTIVE: return lsc::BC_REFLECTIVE;
        default: return lsc::BC_TRANSMISSIVE;
    }
}

}  // namespace

void lsc_count_launches(unsigned long long n) { g_launches += n; }

// error sink shared with the other translation units of the library (lsc_common.h)
int lsc_fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
    return code;
}

struct lsc_mhd_ctx {
    lsc_mhd_config cfg;
    int nxl = 0, i0 = 0;  // slab: interior columns, first global interior column (0-based)
    double dx = 0, dy = 0;
    lsc::GridDims g{};
    // three rotating states: u^n, uBar, and u^{n-1} whose buffer receives u^{n+1}
    double* state[3] = {nullptr, nullptr, nullptr};
    int cur = 0, bar = 1, prev = 2;
    // status words: [0] bit pattern of the max signal speed of state[cur], [1]/[2] number of x/y sweep blocks
    // recomputed with IEEE division in the last step (mhd_kernels.cu), [3] the same for the slab-edge y launch
    unsigned long long* d_max = nullptr;
    unsigned long long* h_max = nullptr;  // pinned copy
    unsigned* d_fix = nullptr;            // y-sweep block indices flagged by the fast sweep
    unsigned* d_xflags = nullptr;         // per-block flags of the x sweep (all zero between steps)
    bool pending = false;                 // status of the last launched step not read back yet
    bool pending_step = false;            // ... and that status belongs to a step (not to an upload)
    double max_speed = 0;                 // max signal speed of state[cur] once resolved
    long ieee_redos = 0;
    double* d_stage = nullptr;            // AoS staging for transfers
    int stage_cols = 0;
    cudaStream_t stream = nullptr;
    // halo/compute overlap: the exchange runs on its own high-priority stream behind the
    // slab-edge columns of the y sweep while `stream` sweeps the interior
    cudaStream_t comm_stream = nullptr;
    cudaEvent_t edges_done = nullptr, halo_done = nullptr;
    int overlap = 1;  // 0 off, 1 when the slab has NCCL neighbours, 2 always (single rank: the ghost kernel overlaps)
    double t = 0, dt = 0, ch = 0;
    long step = 0;
    bool uploaded = false;
    bool ghosts_current = false;  // x ghost/halo columns of state[cur] hold boundary images
    lsc::Comm* comm = nullptr;
    bool in_group = false;            // driven by lsc_mhd_group_step (single-process slab group)
    cudaEvent_t halo_event = nullptr;  // "u^{n+1} complete" marker for neighbours in a group
    bool profiling = false;
    bool relaxed = false;  // LSC_ARITH_RELAXED: FMA-contracted sweeps with 1-multiply quotients (not the parity path)
    std::vector<cudaEvent_t> prof_events;  // quadruples: x begin, x end, y begin, y end
    std::vector<cudaEvent_t> halo_events, reduce_events;  // pairs around the NCCL halo exchange / max-allreduce
};

extern "C" {

const char* lsc_last_error(void) { return g_err; }

int lsc_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

unsigned long long lsc_kernel_launch_count(void) { return g_launches.load(); }

int lsc_host_alloc(void** ptr, unsigned long long bytes) {
    if (!ptr) return fail(LSC_EINVAL, "lsc_host_alloc: null out pointer");
    CUDA_TRY(cudaHostAlloc(ptr, (size_t)bytes, cudaHostAllocDefault));
    return LSC_OK;
}

int lsc_host_free(void* ptr) {
    CUDA_TRY(cudaFreeHost(ptr));
    return LSC_OK;
}

int lsc_mhd_create(const lsc_mhd_config* cfg, lsc_mhd_ctx** out) {
    if (!cfg || !out) return fail(LSC_EINVAL, "lsc_mhd_create: null argument");
    *out = nullptr;
    if (cfg->nx < 4 || cfg->ny < 4) return fail(LSC_EINVAL, "nx, ny must be >= 4 (got %d x %d)", cfg->nx, cfg->ny);
    if (!(cfg->x1 > cfg->x0) || !(cfg->y1 > cfg->y0)) return fail(LSC_EINVAL, "empty domain");
    if (!(cfg->gamma > 1.0) || !(cfg->cfl > 0.0)) return fail(LSC_EINVAL, "gamma must be > 1 and cfl > 0");
    if (cfg->bc_x != LSC_BC_PERIODIC && cfg->bc_x != LSC_BC_TRANSMISSIVE)
        return fail(LSC_EINVAL, "bc_x must be periodic or transmissive (MUSCL_NEW.cpp:609-622)");
    if (cfg->bc_y != LSC_BC_PERIODIC && cfg->bc_y != LSC_BC_TRANSMISSIVE && cfg->bc_y != LSC_BC_REFLECTIVE)
        return fail(LSC_EINVAL, "bc_y must be reflective, periodic or transmissive (MUSCL_NEW.cpp:625-654)");
    if (cfg->nranks < 1 || cfg->rank < 0 || cfg->rank >= cfg->nranks) return fail(LSC_EINVAL, "bad rank/nranks");
    if (cfg->nx / cfg->nranks < 4) return fail(LSC_EINVAL, "slab narrower than 4 columns");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(LSC_ECUDA, "no CUDA device: this library has no CPU path");
    }
    if (cfg->device < 0 || cfg->device >= ndev) return fail(LSC_EINVAL, "device %d out of range", cfg->device);
    CUDA_TRY(cudaSetDevice(cfg->device));

    lsc_mhd_ctx* c = new (std::nothrow) lsc_mhd_ctx();
    if (!c) return fail(LSC_ENOMEM, "host allocation failed");
    c->cfg = *cfg;
    const long long b0 = (long long)cfg->rank * cfg->nx / cfg->nranks;
    const long long b1 = (long long)(cfg->rank + 1) * cfg->nx / cfg->nranks;
    c->i0 = (int)b0;
    c->nxl = (int)(b1 - b0);
    c->dx = (cfg->x1 - cfg->x0) / cfg->nx;  // MUSCL_NEW.cpp:669-670
    c->dy = (cfg->y1 - cfg->y0) / cfg->ny;
    c->g.nxl = c->nxl;
    c->g.ny = cfg->ny;
    c->g.pitch = ((cfg->ny + 4 + lsc::JOFF) + lsc::JBLK - 1) / lsc::JBLK * lsc::JBLK;
    c->g.plane = (size_t)(c->nxl + 4) * (size_t)c->g.pitch;
    const size_t state_bytes = c->g.plane * lsc::NV * sizeof(double);
    for (int b = 0; b < 3; ++b) {
        if (cudaMalloc(&c->state[b], state_bytes) != cudaSuccess) {
            cudaGetLastError();
            lsc_mhd_destroy(c);
            return fail(LSC_ENOMEM, "cudaMalloc of %zu bytes failed", state_bytes);
        }
        cudaMemset(c->state[b], 0, state_bytes);
    }
    // staging: at most ~256 MB of AoS columns
    const size_t col_bytes = (size_t)(cfg->ny + 4) * lsc::NV * sizeof(double);
    long long cols = (long long)((256ull << 20) / col_bytes);
    if (cols < 1) cols = 1;
    if (cols > c->nxl + 4) cols = c->nxl + 4;
    c->stage_cols = (int)cols;
    if (cudaMalloc(&c->d_stage, col_bytes * (size_t)cols) != cudaSuccess ||
        cudaMalloc(&c->d_max, STATUS_BYTES) != cudaSuccess ||
        cudaMalloc(&c->d_fix, sizeof(unsigned) * (size_t)(lsc::y_sweep_block_count(c->g) +
                                                          lsc::y_edge_block_count(c->g))) != cudaSuccess ||
        cudaMalloc(&c->d_xflags, sizeof(unsigned) * (size_t)lsc::x_sweep_block_count(c->g)) != cudaSuccess ||
        cudaMemset(c->d_xflags, 0, sizeof(unsigned) * (size_t)lsc::x_sweep_block_count(c->g)) != cudaSuccess ||
        cudaHostAlloc(&c->h_max, STATUS_BYTES, cudaHostAllocDefault) != cudaSuccess ||
        cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->halo_event, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->edges_done, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->halo_done, cudaEventDisableTiming) != cudaSuccess ||
        lsc_create_comm_stream(&c->comm_stream) != cudaSuccess) {
        cudaGetLastError();
        lsc_mhd_destroy(c);
        return fail(LSC_ENOMEM, "allocation of staging / scalars / stream failed");
    }
    *out = c;
    return LSC_OK;
}

int lsc_mhd_destroy(lsc_mhd_ctx* c) {
    if (!c) return LSC_OK;
    cudaSetDevice(c->cfg.device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->comm) lsc::comm_destroy(c->comm);
    for (int b = 0; b < 3; ++b)
        if (c->state[b]) cudaFree(c->state[b]);
    if (c->d_fix) cudaFree(c->d_fix);
    if (c->d_xflags) cudaFree(c->d_xflags);
    if (c->d_stage) cudaFree(c->d_stage);
    if (c->d_max) cudaFree(c->d_max);
    if (c->h_max) cudaFreeHost(c->h_max);
    if (c->halo_event) cudaEventDestroy(c->halo_event);
    if (c->edges_done) cudaEventDestroy(c->edges_done);
    if (c->halo_done) cudaEventDestroy(c->halo_done);
    if (c->comm_stream) cudaStreamDestroy(c->comm_stream);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
    return LSC_OK;
}

int lsc_mhd_local_nx(const lsc_mhd_ctx* c) { return c ? c->nxl : LSC_EINVAL; }
int lsc_mhd_local_i0(const lsc_mhd_ctx* c) { return c ? c->i0 : LSC_EINVAL; }

int lsc_mhd_initial_condition(const lsc_mhd_ctx* c, int problem, double* aos) {
    if (!c || !aos) return fail(LSC_EINVAL, "null argument");
    if (problem != LSC_MHD_IC_KELVIN_HELMHOLTZ && problem != LSC_MHD_IC_ORSZAG_TANG)
        return fail(LSC_EINVAL, "unknown problem %d", problem);
    lsc::mhd_initial_condition(c->cfg, problem, c->i0, 0, c->nxl + 4, aos);
    return LSC_OK;
}

}  // extern "C"

namespace {

// after new data has been placed in state[cur]: reset the clock and reduce the initial signal speed
// over the reference's range (MUSCL_NEW.cpp:73-74: i,j in [1, n+1], array as uploaded, ghosts included)
int finish_upload(lsc_mhd_ctx* c) {
    CUDA_TRY(cudaMemsetAsync(c->d_max, 0, STATUS_BYTES, c->stream));
    const int i_lo = (c->cfg.rank == 0) ? 1 : 2;
    lsc::launch_max_speed(c->g, c->cfg.gamma, c->state[c->cur], i_lo, c->nxl + 1, 1, c->cfg.ny + 1, c->d_max,
                          c->stream);
    lsc_count_launches(1);
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    c->t = 0;
    c->dt = 0;
    c->ch = 0;
    c->step = 0;
    c->uploaded = true;
    c->ghosts_current = false;
    c->pending = true;  // status[0] holds the initial signal speed, status[1] = 0
    c->pending_step = false;
    return LSC_OK;
}

int upload_columns(lsc_mhd_ctx* c, const double* aos_chunk, int il0, int ni) {
    const size_t col = (size_t)(c->cfg.ny + 4) * lsc::NV;
    CUDA_TRY(cudaMemcpyAsync(c->d_stage, aos_chunk, col * ni * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    lsc::launch_aos_to_soa(c->g, c->d_stage, il0, ni, c->state[c->cur], c->stream);
    lsc_count_launches(1);
    CUDA_TRY(cudaGetLastError());
    // the staging buffer is reused by the next chunk
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    return LSC_OK;
}

}  // namespace

extern "C" {

int lsc_mhd_upload(lsc_mhd_ctx* c, const double* aos) {
    if (!c || !aos) return fail(LSC_EINVAL, "null argument");
    CUDA_TRY(cudaSetDevice(c->cfg.device));
    const size_t col = (size_t)(c->cfg.ny + 4) * lsc::NV;
    for (int il = 0; il < c->nxl + 4; il += c->stage_cols) {
        const int ni = std::min(c->stage_cols, c->nxl + 4 - il);
        int rc = upload_columns(c, aos + col * il, il, ni);
        if (rc) return rc;
    }
    return finish_upload(c);
}

int lsc_mhd_init_problem(lsc_mhd_ctx* c, int problem) {
    if (!c) return fail(LSC_EINVAL, "null argument");
    if (problem != LSC_MHD_IC_KELVIN_HELMHOLTZ && problem != LSC_MHD_IC_ORSZAG_TANG)
        return fail(LSC_EINVAL, "unknown problem %d", problem);
    CUDA_TRY(cudaSetDevice(c->cfg.device));
    const size_t col = (size_t)(c->cfg.ny + 4) * lsc::NV;
    const int chunk = std::min(c->stage_cols, 64);
    double* host = nullptr;
    CUDA_TRY(cudaHostAlloc(&host, col * chunk * sizeof(double), cudaHostAllocDefault));
    int rc = LSC_OK;
    for (int il = 0; il < c->nxl + 4 && rc == LSC_OK; il += chunk) {
        const int ni = std::min(chunk, c->nxl + 4 - il);
        lsc::mhd_initial_condition(c->cfg, problem, c->i0, il, ni, host);
        rc = upload_columns(c, host, il, ni);
    }
    cudaFreeHost(host);
    if (rc) return rc;
    return finish_upload(c);
}

int lsc_mhd_set_arithmetic(lsc_mhd_ctx* c, int mode) {
    if (!c) return fail(LSC_EINVAL, "null argument");
    if (mode != LSC_ARITH_REFERENCE && mode != LSC_ARITH_RELAXED) return fail(LSC_EINVAL, "unknown arithmetic mode %d", mode);
    c->relaxed = (mode == LSC_ARITH_RELAXED);
    return LSC_OK;
}

int lsc_mhd_get_scalars(const lsc_mhd_ctx* c, double* t, double* dt, double* ch, long* step) {
    if (!c) return fail(LSC_EINVAL, "null argument");
    if (t) *t = c->t;
    if (dt) *dt = c->dt;
    if (ch) *ch = c->ch;
    if (step) *step = c->step;
    return LSC_OK;
}

}  // extern "C"

namespace {

// Boundary images / neighbour data into the two ghost columns on each side of state b.
int fill_x_ghosts(lsc_mhd_ctx* c, int b, cudaStream_t s = nullptr) {
    if (!s) s = c->stream;
    if (c->cfg.nranks == 1) {
        lsc::launch_x_ghosts(c->g, to_kernel_bc(c->cfg.bc_x), c->state[b], s);
        lsc_count_launches(1);
        CUDA_TRY(cudaGetLastError());
        return LSC_OK;
    }
    if (!c->comm) return fail(LSC_ESTATE, "nranks > 1 but lsc_mhd_comm_init was not called");
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (c->profiling) {
        CUDA_TRY(cudaEventCreate(&e0));
        CUDA_TRY(cudaEventCreate(&e1));
        CUDA_TRY(cudaEventRecord(e0, s));
    }
    int rc = lsc::comm_exchange_halos(c->comm, c->g, to_kernel_bc(c->cfg.bc_x), c->state[b], s);
    if (rc) return fail(LSC_ECOMM, "halo exchange failed: %s", lsc::comm_error());
    if (c->profiling) {
        CUDA_TRY(cudaEventRecord(e1, s));
        c->halo_events.push_back(e0);
        c->halo_events.push_back(e1);
    }
    return LSC_OK;
}

// Halo exchange overlapped with the interior of the y sweep: NCCL slabs wide enough to have an interior.
bool overlap_active(const lsc_mhd_ctx* c) {
    if (c->in_group || c->nxl < 8) return false;
    return c->overlap == 2 || (c->overlap == 1 && c->cfg.nranks > 1 && c->comm);
}

// 16-byte readback of {max signal speed bits, fast-division flag} of this context (no communication).
int read_status_local(lsc_mhd_ctx* c) {
    CUDA_TRY(cudaSetDevice(c->cfg.device));
    CUDA_TRY(cudaMemcpyAsync(c->h_max, c->d_max, STATUS_BYTES, cudaMemcpyDeviceToHost, c->stream));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    return LSC_OK;
}

// Read back the status of the step launched last (max signal speed of the new state; how many blocks the
// device-side IEEE fix kernels had to recompute).
int resolve_pending(lsc_mhd_ctx* c) {
    if (!c->pending) return LSC_OK;
    if (c->in_group) return fail(LSC_ESTATE, "context belongs to a group: use the lsc_mhd_group_* calls");
    if (c->cfg.nranks > 1) {
        if (!c->comm) return fail(LSC_ESTATE, "nranks > 1 but lsc_mhd_comm_init was not called");
        cudaEvent_t e0 = nullptr, e1 = nullptr;
        if (c->profiling) {
            CUDA_TRY(cudaEventCreate(&e0));
            CUDA_TRY(cudaEventCreate(&e1));
            CUDA_TRY(cudaEventRecord(e0, c->stream));
        }
        if (lsc::comm_allreduce_max_u64(c->comm, c->d_max, 1, c->stream))
            return fail(LSC_ECOMM, "allreduce failed: %s", lsc::comm_error());
        if (c->profiling) {
            CUDA_TRY(cudaEventRecord(e1, c->stream));
            c->reduce_events.push_back(e0);
            c->reduce_events.push_back(e1);
        }
    }
    int rc = read_status_local(c);
    if (rc) return rc;
    if (c->pending_step) c->ieee_redos += (long)(c->h_max[1] + c->h_max[2] + c->h_max[3]);
    std::memcpy(&c->max_speed, &c->h_max[0], sizeof(double));
    c->pending = false;
    return LSC_OK;
}

// ComputeTimeStep's scalars (MUSCL_NEW.cpp:117-118, :834-835) and the two sweeps (:841-965) for one context.
// The x ghost columns of state[cur] must be current.  Rotates the buffers (u = uPlus1, :969).
int launch_sweeps(lsc_mhd_ctx* c) {
    CUDA_TRY(cudaSetDevice(c->cfg.device));
    const double max_speed = c->max_speed;
    if (!(max_speed > 0.0) || !std::isfinite(max_speed))
        return fail(LSC_ENAN, "max signal speed is %g at step %ld", max_speed, c->step);
    const double dspace = std::min(c->dx, c->dy);      // :117
    const double dt = c->cfg.cfl * dspace / max_speed;  // :118
    const double ch = max_speed;                        // :834
    c->t += dt;                                         // :835
    c->dt = dt;
    c->ch = ch;

    lsc::SweepParams p{};
    p.gamma = c->cfg.gamma;
    p.gm1 = c->cfg.gamma - 1;
    p.ch = ch;
    p.ch2 = ch * ch;
    p.inv2ch = 1.0 / (2.0 * ch);
    p.halfch = ch / 2.0;
    p.psi_factor = std::exp(-1.0 * (0.5 * dt) * ch / 0.18);  // psiUpdate(psi, maxSpeed, 0.5*dt), :66-69, :900
    p.first_step = (c->step == 0) ? 1 : 0;

    // ---- x sweep (:841-902)
    p.dtdh = dt / c->dx;
    p.hdtdh = 0.5 * (dt / c->dx);
    p.bc_lo = (c->cfg.rank == 0) ? to_kernel_bc(c->cfg.bc_x) : -1;
    p.bc_hi = (c->cfg.rank == c->cfg.nranks - 1) ? to_kernel_bc(c->cfg.bc_x) : -1;
    if (c->cfg.nranks > 1 && c->cfg.bc_x == LSC_BC_PERIODIC) p.bc_lo = p.bc_hi = -1;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    if (c->profiling) {
        for (auto& e : ev) CUDA_TRY(cudaEventCreate(&e));
        CUDA_TRY(cudaEventRecord(ev[0], c->stream));
    }
    CUDA_TRY(cudaMemsetAsync(c->d_max, 0, STATUS_BYTES, c->stream));
    if (c->relaxed)
        lsc::launch_x_sweep_relaxed(c->g, p, c->state[c->cur], c->state[c->prev], c->state[c->bar], c->d_xflags,
                                    c->stream);
    else
        lsc::launch_x_sweep(c->g, p, c->state[c->cur], c->state[c->prev], c->state[c->bar], c->d_max, c->d_xflags,
                            c->stream);
    if (c->profiling) CUDA_TRY(cudaEventRecord(ev[1], c->stream));

    // ---- y sweep (:905-965) + copy (:969) + next step's signal speed (:71-115)
    p.dtdh = dt / c->dy;
    p.hdtdh = 0.5 * (dt / c->dy);
    p.bc_lo = p.bc_hi = to_kernel_bc(c->cfg.bc_y);
    if (c->profiling) CUDA_TRY(cudaEventRecord(ev[2], c->stream));
    auto y_launch = [&](int part) {
        if (c->relaxed)
            lsc::launch_y_sweep_relaxed(c->g, p, c->state[c->bar], c->state[c->cur], c->state[c->prev], c->d_max,
                                        c->d_fix, c->stream, part);
        else
            lsc::launch_y_sweep(c->g, p, c->state[c->bar], c->state[c->cur], c->state[c->prev], c->d_max, c->d_fix,
                                c->stream, part);
    };
    const bool overlapped = overlap_active(c);
    if (overlapped) {
        // The four slab-edge columns of u^{n+1} (what the x neighbours need) first; their halo exchange (:971 for
        // the ghost columns) then runs on comm_stream while this stream sweeps the interior columns.  The interior
        // launch neither reads nor writes the edge or ghost columns of the new state.
        y_launch(lsc::Y_EDGES);
        CUDA_TRY(cudaEventRecord(c->edges_done, c->stream));
        y_launch(lsc::Y_INTERIOR);
        if (c->profiling) CUDA_TRY(cudaEventRecord(ev[3], c->stream));  // y time = both launches, not the halo wait
        CUDA_TRY(cudaStreamWaitEvent(c->comm_stream, c->edges_done, 0));
        int rc = fill_x_ghosts(c, c->prev, c->comm_stream);
        if (rc) return rc;
        CUDA_TRY(cudaEventRecord(c->halo_done, c->comm_stream));
        CUDA_TRY(cudaStreamWaitEvent(c->stream, c->halo_done, 0));  // everything later on `stream` sees the halos
        lsc_count_launches(2);
    } else {
        y_launch(lsc::Y_ALL);
        if (c->profiling) CUDA_TRY(cudaEventRecord(ev[3], c->stream));
    }
    if (c->profiling) {
        for (auto e : ev) c->prof_events.push_back(e);
    }
    lsc_count_launches(4);
    CUDA_TRY(cudaGetLastError());

    // u = uPlus1 (:969): rotate; the old u^n stays as the limiter's "previous" state
    const int next = c->prev;
    c->prev = c->cur;
    c->cur = next;
    c->pending = true;
    c->pending_step = true;
    c->step += 1;
    return LSC_OK;
}

// One pass of the loop body MUSCL_NEW.cpp:831-971.
int step_once(lsc_mhd_ctx* c) {
    int rc = resolve_pending(c);  // ComputeTimeStep (:833): the reduction already ran on the device
    if (rc) return rc;
    if (!c->ghosts_current) {  // applyBoundaryConditions(u) (:838): only x ghost columns are read on the device
        rc = fill_x_ghosts(c, c->cur);
        if (rc) return rc;
        c->ghosts_current = true;
    }
    const bool overlapped = overlap_active(c);
    rc = launch_sweeps(c);
    if (rc) return rc;
    if (overlapped) return LSC_OK;    // the exchange was issued behind the slab-edge columns of the y sweep
    return fill_x_ghosts(c, c->cur);  // applyBoundaryConditions(u) (:971)
}

// ---------------------------------------------------------------------------------------------
// Single-process slab group: P contexts (on one or several GPUs of this process) advanced in lock
// step, halos moved by cudaMemcpyPeerAsync between the contexts' buffers, max reduced on the host.
// Same kernels and same slab-edge code path as the NCCL mode.
// ---------------------------------------------------------------------------------------------
int group_halo(lsc_mhd_ctx** g, int n) {
    const bool ring = (g[0]->cfg.bc_x == LSC_BC_PERIODIC);
    for (int r = 0; r < n; ++r) {  // mark "u^{n+1} of rank r is complete"
        CUDA_TRY(cudaSetDevice(g[r]->cfg.device));
        CUDA_TRY(cudaEventRecord(g[r]->halo_event, g[r]->stream));
    }
    for (int r = 0; r < n; ++r) {
        lsc_mhd_ctx* c = g[r];
        CUDA_TRY(cudaSetDevice(c->cfg.device));
        const int left = (r > 0) ? r - 1 : (ring ? n - 1 : -1);
        const int right = (r < n - 1) ? r + 1 : (ring ? 0 : -1);
        const size_t col = lsc::mhd_column_doubles(c->g);  // a column of the blocked layout is one contiguous run
        const size_t bytes = 2 * col * sizeof(double);
        for (int side = 0; side < 2; ++side) {
            const int nb = side == 0 ? left : right;
            if (nb < 0) {
                lsc::launch_x_ghosts_side(c->g, side, c->state[c->cur], c->stream);
                lsc_count_launches(1);
                continue;
            }
            lsc_mhd_ctx* o = g[nb];
            CUDA_TRY(cudaStreamWaitEvent(c->stream, o->halo_event, 0));
            double* dst = c->state[c->cur] + (side == 0 ? 0 : (size_t)(c->g.nxl + 2) * col);
            // left ghost <- neighbour's last two interior columns; right ghost <- its first two
            const double* src = o->state[o->cur] + (side == 0 ? (size_t)o->g.nxl * col : 2 * col);
            CUDA_TRY(cudaMemcpyPeerAsync(dst, c->cfg.device, src, o->cfg.device, bytes, c->stream));
        }
    }
    CUDA_TRY(cudaGetLastError());
    return LSC_OK;
}

int group_resolve(lsc_mhd_ctx** g, int n) {
    if (!g[0]->pending) return LSC_OK;
    unsigned long long mx = 0;
    for (int r = 0; r < n; ++r) {
        int rc = read_status_local(g[r]);
        if (rc) return rc;
        mx = std::max(mx, g[r]->h_max[0]);
    }
    for (int r = 0; r < n; ++r) {
        if (g[r]->pending_step) g[r]->ieee_redos += (long)(g[r]->h_max[1] + g[r]->h_max[2] + g[r]->h_max[3]);
        std::memcpy(&g[r]->max_speed, &mx, sizeof(double));
        g[r]->pending = false;
    }
    return LSC_OK;
}

int group_check(lsc_mhd_ctx** g, int n) {
    if (!g || n < 1) return fail(LSC_EINVAL, "empty group");
    for (int r = 0; r < n; ++r) {
        if (!g[r]) return fail(LSC_EINVAL, "null context in group");
        if (g[r]->cfg.nranks != n || g[r]->cfg.rank != r)
            return fail(LSC_EINVAL, "group member %d has rank %d of %d", r, g[r]->cfg.rank, g[r]->cfg.nranks);
        if (!g[r]->uploaded) return fail(LSC_ESTATE, "group member %d has no state uploaded", r);
        if (g[r]->comm) return fail(LSC_ESTATE, "group member %d is attached to NCCL", r);
        if (g[r]->step != g[0]->step) return fail(LSC_ESTATE, "group members are at different steps");
    }
    return LSC_OK;
}

}  // namespace

extern "C" {

int lsc_mhd_step(lsc_mhd_ctx* c, long nsteps) {
    if (!c) return fail(LSC_EINVAL, "null argument");
    if (!c->uploaded) return fail(LSC_ESTATE, "lsc_mhd_step before lsc_mhd_upload");
    CUDA_TRY(cudaSetDevice(c->cfg.device));
    for (long s = 0; s < nsteps; ++s) {
        int rc = step_once(c);
        if (rc) return rc;
    }
    return resolve_pending(c);  // syncs; the state is final (and IEEE-exact) when this returns
}

int lsc_mhd_group_step(lsc_mhd_ctx** g, int n, long nsteps) {
    int rc = group_check(g, n);
    if (rc) return rc;
    for (int r = 0; r < n; ++r) g[r]->in_group = true;
    for (long s = 0; s < nsteps; ++s) {
        rc = group_resolve(g, n);
        if (rc) return rc;
        if (!g[0]->ghosts_current) {
            rc = group_halo(g, n);
            if (rc) return rc;
            for (int r = 0; r < n; ++r) g[r]->ghosts_current = true;
        }
        for (int r = 0; r < n; ++r) {
            rc = launch_sweeps(g[r]);
            if (rc) return rc;
        }
        rc = group_halo(g, n);
        if (rc) return rc;
    }
    rc = group_resolve(g, n);
    for (int r = 0; r < n; ++r) g[r]->in_group = false;
    return rc;
}

int lsc_mhd_step_timed(lsc_mhd_ctx* c, long nsteps, float* ms) {
    if (!c || !ms) return fail(LSC_EINVAL, "null argument");
    if (!c->uploaded) return fail(LSC_ESTATE, "lsc_mhd_step_timed before lsc_mhd_upload");
    CUDA_TRY(cudaSetDevice(c->cfg.device));
    cudaEvent_t e0, e1;
    CUDA_TRY(cudaEventCreate(&e0));
    CUDA_TRY(cudaEventCreate(&e1));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    CUDA_TRY(cudaEventRecord(e0, c->stream));
    int rc = LSC_OK;
    for (long s = 0; s < nsteps && rc == LSC_OK; ++s) rc = step_once(c);
    if (rc == LSC_OK) rc = resolve_pending(c);
    if (rc == LSC_OK) {
        CUDA_TRY(cudaEventRecord(e1, c->stream));
        CUDA_TRY(cudaEventSynchronize(e1));
        CUDA_TRY(cudaEventElapsedTime(ms, e0, e1));
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    return rc;
}

long lsc_mhd_ieee_fallback_count(const lsc_mhd_ctx* c) { return c ? c->ieee_redos : -1; }

int lsc_mhd_set_overlap(lsc_mhd_ctx* c, int enabled) {
    if (!c) return fail(LSC_EINVAL, "null context");
    if (enabled < 0 || enabled > 2) return fail(LSC_EINVAL, "overlap mode %d", enabled);
    c->overlap = enabled;
    return LSC_OK;
}

int lsc_mhd_set_profiling(lsc_mhd_ctx* c, int enabled) {
    if (!c) return fail(LSC_EINVAL, "null argument");
    c->profiling = enabled != 0;
    return LSC_OK;
}

int lsc_mhd_kernel_times(lsc_mhd_ctx* c, double ms[2], long launches[2]) {
    if (!c || !ms || !launches) return fail(LSC_EINVAL, "null argument");
    CUDA_TRY(cudaSetDevice(c->cfg.device));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    ms[0] = ms[1] = 0;
    launches[0] = launches[1] = 0;
    for (size_t q = 0; q + 3 < c->prof_events.size(); q += 4) {
        float a = 0, b = 0;
        CUDA_TRY(cudaEventElapsedTime(&a, c->prof_events[q], c->prof_events[q + 1]));
        CUDA_TRY(cudaEventElapsedTime(&b, c->prof_events[q + 2], c->prof_events[q + 3]));
        ms[0] += a;
        ms[1] += b;
        launches[0] += 1;
        launches[1] += 1;
    }
    for (auto e : c->prof_events) cudaEventDestroy(e);
    c->prof_events.clear();
    return LSC_OK;
}

int lsc_mhd_comm_times(lsc_mhd_ctx* c, double ms[2], long calls[2]) {
    if (!c || !ms || !calls) return fail(LSC_EINVAL, "null argument");
    CUDA_TRY(cudaSetDevice(c->cfg.device));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    std::vector<cudaEvent_t>* lists[2] = {&c->halo_events, &c->reduce_events};
    for (int w = 0; w < 2; ++w) {
        ms[w] = 0;
        calls[w] = 0;
        for (size_t q = 0; q + 1 < lists[w]->size(); q += 2) {
            float a = 0;
            CUDA_TRY(cudaEventElapsedTime(&a, (*lists[w])[q], (*lists[w])[q + 1]));
            ms[w] += a;
            calls[w] += 1;
        }
        for (auto e : *lists[w]) cudaEventDestroy(e);
        lists[w]->clear();
    }
    return LSC_OK;
}

int lsc_mhd_run_until(lsc_mhd_ctx* c, double t_stop, lsc_mhd_dump_cb dump_cb, lsc_mhd_log_cb log_cb, void* user,
                      long* steps_out) {
    if (!c) return fail(LSC_EINVAL, "null argument");
    if (!c->uploaded) return fail(LSC_ESTATE, "lsc_mhd_run_until before lsc_mhd_upload");
    CUDA_TRY(cudaSetDevice(c->cfg.device));
    int counter = 0;
    long n = 0;
    do {  // MUSCL_NEW.cpp:831
        int rc = step_once(c);
        if (rc) return rc;
        ++n;
        if (log_cb && log_cb(c->step, c->t, c->dt, c->ch, user))  // :836
            return fail(LSC_ECALLBACK, "log callback stopped the loop at step %ld", c->step);
        if (c->t > 0.1 * counter && c->t - c->dt <= 0.1 * counter) {  // :975
            if (dump_cb) {
                rc = dump_cb(c, counter, c->t, c->dt, user);
                if (rc) return fail(rc, "dump callback returned %d", rc);
            }
            counter += 1;  // :1000
        }
    } while (c->t < t_stop);  // :1003
    if (steps_out) *steps_out = n;
    return resolve_pending(c);
}

int lsc_mhd_download(lsc_mhd_ctx* c, double* aos) {
    if (!c || !aos) return fail(LSC_EINVAL, "null argument");
    if (!c->uploaded) return fail(LSC_ESTATE, "lsc_mhd_download before lsc_mhd_upload");
    CUDA_TRY(cudaSetDevice(c->cfg.device));
    if (c->pending && c->pending_step) {  // a launched step may still have to be recomputed (IEEE path)
        int rc = resolve_pending(c);
        if (rc) return rc;
    }
    const size_t col = (size_t)(c->cfg.ny + 4) * lsc::NV;
    for (int il = 0; il < c->nxl + 4; il += c->stage_cols) {
        const int ni = std::min(c->stage_cols, c->nxl + 4 - il);
        lsc::launch_soa_to_aos(c->g, c->state[c->cur], il, ni, c->d_stage, c->stream);
        lsc_count_launches(1);
        CUDA_TRY(cudaGetLastError());
        CUDA_TRY(cudaMemcpyAsync(aos + col * il, c->d_stage, col * ni * sizeof(double), cudaMemcpyDeviceToHost,
                                 c->stream));
        CUDA_TRY(cudaStreamSynchronize(c->stream));
    }
    if (c->step > 0) {
        // the device never stores y ghost rows (the y sweep indexes through the boundary map);
        // rebuild the reference's ghost ring on the host from the interior
        if (c->cfg.nranks == 1) {
            lsc::mhd_fill_ghost_ring(c->cfg, aos);
        } else {
            lsc::mhd_apply_bc_y(c->cfg, c->nxl + 4, aos);
        }
    }
    return LSC_OK;
}

int lsc_mhd_state_hash(lsc_mhd_ctx* c, unsigned long long* hash) {
    if (!c || !hash) return fail(LSC_EINVAL, "null argument");
    if (!c->uploaded) return fail(LSC_ESTATE, "lsc_mhd_state_hash before lsc_mhd_upload");
    CUDA_TRY(cudaSetDevice(c->cfg.device));
    if (c->pending && c->pending_step) {  // a launched step may still have to be recomputed (IEEE path)
        int rc = resolve_pending(c);
        if (rc) return rc;
    }
    unsigned long long* d = nullptr;
    CUDA_TRY(cudaMalloc(&d, sizeof *d));
    cudaMemsetAsync(d, 0, sizeof *d, c->stream);
    lsc::launch_state_hash(c->g, lsc::NV, c->state[c->cur], c->i0, d, c->stream);
    lsc_count_launches(1);
    cudaError_t e = cudaMemcpyAsync(hash, d, sizeof *d, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(d);
    CUDA_TRY(e);
    return LSC_OK;
}

int lsc_mhd_fill_ghost_ring(const lsc_mhd_config* cfg, double* aos) {
    if (!cfg || !aos) return fail(LSC_EINVAL, "null argument");
    lsc::mhd_fill_ghost_ring(*cfg, aos);
    return LSC_OK;
}

int lsc_mhd_write_dat(const lsc_mhd_config* cfg, const double* aos, const char* path) {
    if (!cfg || !aos || !path) return fail(LSC_EINVAL, "null argument");
    if (lsc::mhd_write_dat(*cfg, aos, path)) return fail(LSC_EINVAL, "cannot write %s", path);
    return LSC_OK;
}

int lsc_comm_unique_id(void* id128) {
    if (!id128) return fail(LSC_EINVAL, "null argument");
    if (lsc::comm_unique_id(id128)) return fail(LSC_ECOMM, "%s", lsc::comm_error());
    return LSC_OK;
}

int lsc_mhd_comm_init(lsc_mhd_ctx* c, const void* id128) {
    if (!c || !id128) return fail(LSC_EINVAL, "null argument");
    if (c->cfg.nranks == 1) return LSC_OK;
    CUDA_TRY(cudaSetDevice(c->cfg.device));
    c->comm = lsc::comm_create(id128, c->cfg.rank, c->cfg.nranks, c->g, lsc::NV);
    if (!c->comm) return fail(LSC_ECOMM, "%s", lsc::comm_error());
    return LSC_OK;
}

}  // extern "C"

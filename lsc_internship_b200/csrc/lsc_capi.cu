// lsc_capi.cu -- the C ABI of include/lsc_b200.h: contexts, transfers and the time loop
// of MHD_codes/MUSCL_NEW.cpp main() (:658-1034) driving the sm_100a kernels.
//
// The loop is device-resident (step_ctl.cuh): dt, ch, t, the psi factor and the loop predicates are formed by a
// one-thread prepare kernel at the head of every step, the sweeps read them from the context's __constant__ slot,
// and the host enqueues whole batches of steps.  It synchronises once per batch -- once per lsc_mhd_step() call,
// once per dump / stop in lsc_mhd_run_until() -- to read back {t, dt, ch, step, halt reason} (and the per-step log
// when a log callback wants it).  The same loop drives a plain context, one NCCL rank of a process-per-GPU run,
// and a "team" of slabs owned by one process (lsc_mhd_group_*: the drop-in program on the 8 GPUs of a box).
#include <atomic>
#include <climits>
#include <cmath>
#include <cstdarg>
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <mutex>
#include <new>
#include <vector>

#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>

#include "../../include/lsc_b200.h"
#include "lsc_comm.h"
#include "lsc_common.h"
#include "mhd_host.h"
#include "mhd_kernels.cuh"
#include "state_hash.h"
#include "step_ctl.cuh"

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
constexpr int MAX_DEVICES = 64;
constexpr long BATCH_MAX = 4096;  // steps enqueued between two host synchronisations at most

inline int to_kernel_bc(int bc) {
    switch (bc) {
        case LSC_BC_PERIODIC: return lsc::BC_PERIODIC;
        case LSC_BC_REFLECTIVE: return lsc::BC_REFLECTIVE;
        default: return lsc::BC_TRANSMISSIVE;
    }
}

// __constant__ slots of the step scalars: one per context that exists on a device (kind 0: MHD, 1: Euler -- the
// two solver families keep separate __constant__ arrays)
std::mutex g_slot_mutex;
bool g_slot_used[2][MAX_DEVICES][lsc::STEP_SLOTS];

}  // namespace

int lsc_slot_acquire(int device, int kind) {
    if (device < 0 || device >= MAX_DEVICES) return -1;
    std::lock_guard<std::mutex> lock(g_slot_mutex);
    for (int s = 0; s < lsc::STEP_SLOTS; ++s)
        if (!g_slot_used[kind][device][s]) {
            g_slot_used[kind][device][s] = true;
            return s;
        }
    return -1;
}
void lsc_slot_release(int device, int kind, int slot) {
    if (slot < 0 || device < 0 || device >= MAX_DEVICES) return;
    std::lock_guard<std::mutex> lock(g_slot_mutex);
    g_slot_used[kind][device][slot] = false;
}

namespace {

// NVTX ranges around batches / steps / exchanges when LSC_NVTX=1: off by default
bool nvtx_on() {
    static const bool on = [] {
        const char* e = std::getenv("LSC_NVTX");
        return e && *e && *e != '0';
    }();
    return on;
}
struct NvtxRange {
    bool on;
    explicit NvtxRange(const char* name) : on(nvtx_on()) {
        if (on) nvtxRangePushA(name);
    }
    ~NvtxRange() {
        if (on) nvtxRangePop();
    }
};

// everything the host reads back after a batch, in one pinned block
struct HostBlock {
    lsc::StepCmd cmd;
    lsc::StepCtl ctl;
    unsigned long long status[4];
    double log[3 * lsc::STEP_LOG_CAP];
};

}  // namespace

void lsc_count_launches(unsigned long long n) { g_launches += n; }
void lsc_uncount_launches(unsigned long long n) { g_launches -= n; }  // launches captured into a graph, not run

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
    unsigned long long* d_status = nullptr;
    unsigned* d_fix = nullptr;     // y-sweep block indices flagged by the fast sweep
    unsigned* d_xflags = nullptr;  // per-block flags of the x sweep (all zero between steps)
    // device-resident loop state (step_ctl.cuh) and its host mirror
    lsc::StepCtl* d_ctl = nullptr;
    lsc::StepCmd* d_cmd = nullptr;
    double* d_log = nullptr;
    unsigned long long* d_gather = nullptr;  // team leader: members' status[0] words, then their maximum
    HostBlock* h = nullptr;                  // pinned
    int slot = -1;
    unsigned long long batch = 0;
    int ntab = 0;  // psi-damping table {argument bits -> glibc exp}
    unsigned long long tab_arg[lsc::EXP_TAB_CAP];
    double tab_val[lsc::EXP_TAB_CAP];
    long exp_misses = 0;
    long batches = 0;  // host synchronisations of the time loop so far
    long ieee_redos = 0;
    double* d_stage = nullptr;  // AoS staging for transfers
    int stage_cols = 0;
    cudaStream_t stream = nullptr;
    // halo/compute overlap: the exchange runs on its own high-priority stream behind the
    // slab-edge columns of the y sweep while `stream` sweeps the interior
    cudaStream_t comm_stream = nullptr;
    // host <-> device transfers of whole states: the staging buffer is used as two halves, the copy of one half
    // (copy_stream) overlaps the layout transposition of the other (stream)
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t stage_filled[2] = {nullptr, nullptr}, stage_free[2] = {nullptr, nullptr};
    cudaEvent_t edges_done = nullptr, halo_done = nullptr;
    int overlap = 1;  // 0 off, 1 when the slab has NCCL neighbours, 2 always (single rank: the ghost kernel overlaps)
    double t = 0, dt = 0, ch = 0;
    long step = 0;
    bool uploaded = false;
    bool ghosts_current = false;  // x ghost/halo columns of state[cur] hold boundary images
    lsc::Comm* comm = nullptr;
    bool in_team = false;              // driven by the lsc_mhd_group_* calls (single-process slabs)
    cudaEvent_t halo_event = nullptr;  // team: "u^{n+1} of this slab is complete"
    cudaEvent_t red_event = nullptr;   // team: "status word of this slab is in the leader's gather buffer"
    cudaEvent_t max_event = nullptr;   // team leader: "the maximum over the slabs is ready"
    bool profiling = false;
    bool relaxed = false;  // LSC_ARITH_RELAXED: FMA-contracted sweeps with 1-multiply quotients (not the parity path)
    std::vector<cudaEvent_t> prof_events;  // quadruples: x begin, x end, y begin, y end
    std::vector<cudaEvent_t> halo_events, reduce_events;  // pairs around the NCCL halo exchange / max-allreduce
    // CUDA graphs of GRAPH_STEPS steps (plain contexts), keyed on what is baked into them
    struct StepGraph {
        int key;
        cudaGraphExec_t exec;
        unsigned long long launches;
    };
    std::vector<StepGraph> graphs;
    bool graph_disabled = false;
    bool peers_enabled = false;  // team: peer access from this context's GPU to the other members' GPUs
};

namespace {

void destroy_events(std::vector<cudaEvent_t>& v) {
    for (auto e : v) cudaEventDestroy(e);
    v.clear();
}

// {argument bit pattern -> glibc exp(argument)} for the psi-damping arguments a run can produce (step_ctl.cuh):
// the expression's value for a spread of signal speeds, plus the neighbouring doubles.
void build_exp_table(lsc_mhd_ctx* c) {
    c->ntab = 0;
    // test hook: start from an empty table, so that every distinct argument goes through the miss path once
    if (const char* e = std::getenv("LSC_EXP_TABLE_EMPTY"))
        if (*e && *e != '0') return;
    const double ds = std::min(c->dx, c->dy);
    auto add = [&](double arg) {
        unsigned long long b;
        std::memcpy(&b, &arg, sizeof b);
        for (int q = 0; q < c->ntab; ++q)
            if (c->tab_arg[q] == b) return;
        if (c->ntab >= lsc::EXP_TAB_CAP) return;
        c->tab_arg[c->ntab] = b;
        c->tab_val[c->ntab] = std::exp(arg);  // psiUpdate, MUSCL_NEW.cpp:67
        c->ntab += 1;
    };
    double centre = 0;
    for (int q = 0; q < 4096 && c->ntab < lsc::EXP_TAB_CAP / 2; ++q) {
        const double ms = 0.01 * std::pow(1.0e5, q / 4095.0) * (1.0 + 1.0 / 3.0);  // 0.013 .. 1333, full mantissas
        const double dt = c->cfg.cfl * ds / ms;
        centre = lsc::psi_argument(dt, ms);
        add(centre);
    }
    double lo = centre, hi = centre;
    for (int q = 0; q < c->ntab; ++q) {
        double v;
        std::memcpy(&v, &c->tab_arg[q], sizeof v);
        lo = std::min(lo, v);
        hi = std::max(hi, v);
    }
    for (int q = 0; q < 4; ++q) {
        lo = std::nextafter(lo, -std::numeric_limits<double>::infinity());
        hi = std::nextafter(hi, std::numeric_limits<double>::infinity());
        add(lo);
        add(hi);
    }
}

// host mirror -> device: fresh loop state (t = 0, step = 0) with the run's constants and the exp table
int reset_ctl(lsc_mhd_ctx* c) {
    lsc::StepCtl& k = c->h->ctl;
    std::memset(&k, 0, sizeof k);
    k.cfl = c->cfg.cfl;
    k.dspace = std::min(c->dx, c->dy);  // MUSCL_NEW.cpp:117
    k.dx = c->dx;
    k.dy = c->dy;
    k.batch_seen = c->batch;
    k.ntab = c->ntab;
    std::memcpy(k.tab_arg, c->tab_arg, sizeof k.tab_arg);
    std::memcpy(k.tab_val, c->tab_val, sizeof k.tab_val);
    CUDA_TRY(cudaMemcpyAsync(c->d_ctl, &k, sizeof k, cudaMemcpyHostToDevice, c->stream));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    return LSC_OK;
}

}  // namespace

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
    CUDA_TRY(cudaHostAlloc(ptr, (size_t)bytes, cudaHostAllocPortable));
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
    long long b0 = (long long)cfg->rank * cfg->nx / cfg->nranks;
    long long b1 = (long long)(cfg->rank + 1) * cfg->nx / cfg->nranks;
    if (cfg->slab_nx > 0) {  // explicit slab bounds (uneven decompositions sized by the caller)
        b0 = cfg->slab_i0;
        b1 = b0 + cfg->slab_nx;
        if (b0 < 0 || b1 > cfg->nx) return fail(LSC_EINVAL, "slab [%lld, %lld) outside the grid", b0, b1);
        if ((cfg->rank == 0) != (b0 == 0) || (cfg->rank == cfg->nranks - 1) != (b1 == cfg->nx))
            return fail(LSC_EINVAL, "explicit slab bounds: rank 0 must start at column 0, the last rank end at nx");
    }
    if (b1 - b0 < 4) return fail(LSC_EINVAL, "slab narrower than 4 columns");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(LSC_ECUDA, "no CUDA device: this library has no CPU path");
    }
    if (cfg->device < 0 || cfg->device >= ndev || cfg->device >= MAX_DEVICES)
        return fail(LSC_EINVAL, "device %d out of range", cfg->device);
    CUDA_TRY(cudaSetDevice(cfg->device));

    lsc_mhd_ctx* c = new (std::nothrow) lsc_mhd_ctx();
    if (!c) return fail(LSC_ENOMEM, "host allocation failed");
    c->cfg = *cfg;
    c->i0 = (int)b0;
    c->nxl = (int)(b1 - b0);
    c->dx = (cfg->x1 - cfg->x0) / cfg->nx;  // MUSCL_NEW.cpp:669-670
    c->dy = (cfg->y1 - cfg->y0) / cfg->ny;
    c->g.nxl = c->nxl;
    c->g.ny = cfg->ny;
    c->g.pitch = ((cfg->ny + 4 + lsc::JOFF) + lsc::JBLK - 1) / lsc::JBLK * lsc::JBLK;
    c->g.plane = (size_t)(c->nxl + 4) * (size_t)c->g.pitch;
    c->slot = lsc_slot_acquire(cfg->device, 0);
    if (c->slot < 0) {
        delete c;
        return fail(LSC_ENOMEM, "more than %d live contexts on device %d", lsc::STEP_SLOTS, cfg->device);
    }
    const size_t state_bytes = c->g.plane * lsc::NV * sizeof(double);
    for (int b = 0; b < 3; ++b) {
        if (lsc_dev_malloc(&c->state[b], state_bytes) != cudaSuccess) {
            cudaGetLastError();
            lsc_mhd_destroy(c);
            return fail(LSC_ENOMEM, "cudaMalloc of %zu bytes failed", state_bytes);
        }
        cudaMemset(c->state[b], 0, state_bytes);
    }
    // staging: at most ~256 MB of AoS columns
    const size_t col_bytes = (size_t)(cfg->ny + 4) * lsc::NV * sizeof(double);
    long long cols = (long long)((256ull << 20) / col_bytes);
    if (cols < 2) cols = 2;  // two halves (lsc_mhd_upload / download_columns)
    if (cols > c->nxl + 4) cols = c->nxl + 4;
    c->stage_cols = (int)cols;
    if (lsc_dev_malloc(&c->d_stage, col_bytes * (size_t)cols) != cudaSuccess ||
        lsc_dev_malloc(&c->d_status, STATUS_BYTES) != cudaSuccess ||
        cudaMemset(c->d_status, 0, STATUS_BYTES) != cudaSuccess ||
        lsc_dev_malloc(&c->d_fix, sizeof(unsigned) * (size_t)(lsc::y_sweep_block_count(c->g) +
                                                          lsc::y_edge_block_count(c->g))) != cudaSuccess ||
        lsc_dev_malloc(&c->d_xflags, sizeof(unsigned) * (size_t)lsc::x_sweep_block_count(c->g)) != cudaSuccess ||
        cudaMemset(c->d_xflags, 0, sizeof(unsigned) * (size_t)lsc::x_sweep_block_count(c->g)) != cudaSuccess ||
        lsc_dev_malloc(&c->d_ctl, sizeof(lsc::StepCtl)) != cudaSuccess ||
        lsc_dev_malloc(&c->d_cmd, sizeof(lsc::StepCmd)) != cudaSuccess ||
        cudaMemset(c->d_cmd, 0, sizeof(lsc::StepCmd)) != cudaSuccess ||
        lsc_dev_malloc(&c->d_log, sizeof(double) * 3 * lsc::STEP_LOG_CAP) != cudaSuccess ||
        lsc_dev_malloc(&c->d_gather, sizeof(unsigned long long) * (size_t)(cfg->nranks + 1)) != cudaSuccess ||
        cudaHostAlloc(&c->h, sizeof(HostBlock), cudaHostAllocPortable) != cudaSuccess ||
        cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->halo_event, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->red_event, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->max_event, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->edges_done, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->halo_done, cudaEventDisableTiming) != cudaSuccess ||
        lsc_create_comm_stream(&c->comm_stream) != cudaSuccess ||
        cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->stage_filled[0], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->stage_filled[1], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->stage_free[0], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->stage_free[1], cudaEventDisableTiming) != cudaSuccess) {
        cudaGetLastError();
        lsc_mhd_destroy(c);
        return fail(LSC_ENOMEM, "allocation of staging / scalars / stream failed");
    }
    std::memset(c->h, 0, sizeof(HostBlock));
    if (const char* e = std::getenv("LSC_NO_GRAPH")) c->graph_disabled = (*e && *e != '0');
    build_exp_table(c);
    int rc = reset_ctl(c);
    if (rc) {
        lsc_mhd_destroy(c);
        return rc;
    }
    *out = c;
    return LSC_OK;
}

int lsc_mhd_destroy(lsc_mhd_ctx* c) {
    if (!c) return LSC_OK;
    cudaSetDevice(c->cfg.device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->comm_stream) cudaStreamSynchronize(c->comm_stream);
    if (c->copy_stream) cudaStreamSynchronize(c->copy_stream);
    if (c->comm) lsc::comm_destroy(c->comm);
    for (int b = 0; b < 3; ++b)
        if (c->state[b]) lsc_dev_free(c->state[b]);
    if (c->d_fix) lsc_dev_free(c->d_fix);
    if (c->d_xflags) lsc_dev_free(c->d_xflags);
    if (c->d_stage) lsc_dev_free(c->d_stage);
    if (c->d_status) lsc_dev_free(c->d_status);
    if (c->d_ctl) lsc_dev_free(c->d_ctl);
    if (c->d_cmd) lsc_dev_free(c->d_cmd);
    if (c->d_log) lsc_dev_free(c->d_log);
    if (c->d_gather) lsc_dev_free(c->d_gather);
    if (c->h) cudaFreeHost(c->h);
    if (c->halo_event) cudaEventDestroy(c->halo_event);
    if (c->red_event) cudaEventDestroy(c->red_event);
    if (c->max_event) cudaEventDestroy(c->max_event);
    if (c->edges_done) cudaEventDestroy(c->edges_done);
    if (c->halo_done) cudaEventDestroy(c->halo_done);
    destroy_events(c->prof_events);
    destroy_events(c->halo_events);
    destroy_events(c->reduce_events);
    for (auto& g : c->graphs) cudaGraphExecDestroy(g.exec);
    for (int b = 0; b < 2; ++b) {
        if (c->stage_filled[b]) cudaEventDestroy(c->stage_filled[b]);
        if (c->stage_free[b]) cudaEventDestroy(c->stage_free[b]);
    }
    if (c->comm_stream) cudaStreamDestroy(c->comm_stream);
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    if (c->stream) cudaStreamDestroy(c->stream);
    lsc_slot_release(c->cfg.device, 0, c->slot);
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
    CUDA_TRY(cudaMemsetAsync(c->d_status, 0, STATUS_BYTES, c->stream));
    const int i_lo = (c->cfg.rank == 0) ? 1 : 2;
    lsc::launch_max_speed(c->g, c->cfg.gamma, c->state[c->cur], i_lo, c->nxl + 1, 1, c->cfg.ny + 1, c->d_status,
                          c->stream);
    lsc_count_launches(1);
    CUDA_TRY(cudaGetLastError());
    int rc = reset_ctl(c);  // synchronises
    if (rc) return rc;
    c->t = 0;
    c->dt = 0;
    c->ch = 0;
    c->step = 0;
    c->ieee_redos = 0;
    c->uploaded = true;
    c->ghosts_current = false;
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
    // two half buffers: the host->device copy of chunk k+1 (copy_stream) runs while chunk k is transposed into the
    // device layout (stream); a half is refilled only after its transposition has finished
    const int half = std::max(1, c->stage_cols / 2);
    int k = 0;
    for (int il = 0; il < c->nxl + 4; il += half, ++k) {
        const int ni = std::min(half, c->nxl + 4 - il);
        const int b = k & 1;
        double* buf = c->d_stage + (size_t)b * half * col;
        if (k >= 2) CUDA_TRY(cudaStreamWaitEvent(c->copy_stream, c->stage_free[b], 0));
        CUDA_TRY(cudaMemcpyAsync(buf, aos + col * il, col * ni * sizeof(double), cudaMemcpyHostToDevice, c->copy_stream));
        CUDA_TRY(cudaEventRecord(c->stage_filled[b], c->copy_stream));
        CUDA_TRY(cudaStreamWaitEvent(c->stream, c->stage_filled[b], 0));
        lsc::launch_aos_to_soa(c->g, buf, il, ni, c->state[c->cur], c->stream);
        lsc_count_launches(1);
        CUDA_TRY(cudaGetLastError());
        CUDA_TRY(cudaEventRecord(c->stage_free[b], c->stream));
    }
    CUDA_TRY(cudaStreamSynchronize(c->stream));
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

// One solver as the loop sees it: a plain context (n == 1; possibly one NCCL rank of a process-per-GPU run) or the
// slabs of one grid owned by this process (n > 1: halos by peer copies, the maximum through the leader's buffer).
struct Team {
    lsc_mhd_ctx** m;
    int n;
    lsc_mhd_ctx* lead() const { return m[0]; }
};

struct ProfScope {  // CUDA events around a span of the stream when profiling is on; leak-free on early returns
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    cudaStream_t s;
    std::vector<cudaEvent_t>* sink;
    ProfScope(bool on, cudaStream_t stream, std::vector<cudaEvent_t>* out) : s(stream), sink(on ? out : nullptr) {
        if (!sink) return;
        if (cudaEventCreate(&e0) != cudaSuccess || cudaEventCreate(&e1) != cudaSuccess) {
            if (e0) cudaEventDestroy(e0);
            e0 = e1 = nullptr;
            sink = nullptr;
            return;
        }
        cudaEventRecord(e0, s);
    }
    void end() {
        if (!sink) return;
        cudaEventRecord(e1, s);
        sink->push_back(e0);
        sink->push_back(e1);
        sink = nullptr;
    }
    ~ProfScope() {
        if (sink) {  // span abandoned by an error return
            cudaEventDestroy(e0);
            cudaEventDestroy(e1);
        }
    }
};

// Boundary images / neighbour data into the two ghost columns on each side of state b (plain context / NCCL rank).
int fill_x_ghosts(lsc_mhd_ctx* c, int b, cudaStream_t s = nullptr) {
    if (!s) s = c->stream;
    if (c->cfg.nranks == 1) {
        lsc::launch_x_ghosts(c->g, to_kernel_bc(c->cfg.bc_x), c->state[b], s);
        lsc_count_launches(1);
        CUDA_TRY(cudaGetLastError());
        return LSC_OK;
    }
    if (!c->comm) return fail(LSC_ESTATE, "nranks > 1 but lsc_mhd_comm_init was not called");
    NvtxRange r("lsc halo exchange");
    ProfScope prof(c->profiling, s, &c->halo_events);
    int rc = lsc::comm_exchange_halos(c->comm, c->g, to_kernel_bc(c->cfg.bc_x), c->state[b], s);
    if (rc) return fail(LSC_ECOMM, "halo exchange failed: %s", lsc::comm_error());
    prof.end();
    return LSC_OK;
}

// Halo exchange overlapped with the interior of the y sweep: NCCL slabs wide enough to have an interior.
bool overlap_active(const lsc_mhd_ctx* c) {
    if (c->in_team || c->nxl < 8) return false;
    return c->overlap == 2 || (c->overlap == 1 && c->cfg.nranks > 1 && c->comm);
}

// Team halos of state[buf[r]]: every slab records "complete", then pulls its neighbours' edge columns.
int team_halo(const Team& T, const int* buf) {
    const bool ring = (T.lead()->cfg.bc_x == LSC_BC_PERIODIC);
    for (int r = 0; r < T.n; ++r) {
        CUDA_TRY(cudaSetDevice(T.m[r]->cfg.device));
        CUDA_TRY(cudaEventRecord(T.m[r]->halo_event, T.m[r]->stream));
    }
    for (int r = 0; r < T.n; ++r) {
        lsc_mhd_ctx* c = T.m[r];
        CUDA_TRY(cudaSetDevice(c->cfg.device));
        const int left = (r > 0) ? r - 1 : (ring ? T.n - 1 : -1);
        const int right = (r < T.n - 1) ? r + 1 : (ring ? 0 : -1);
        const size_t col = lsc::mhd_column_doubles(c->g);  // a column of the blocked layout is one contiguous run
        const size_t bytes = 2 * col * sizeof(double);
        for (int side = 0; side < 2; ++side) {
            const int nb = side == 0 ? left : right;
            if (nb < 0) {
                lsc::launch_x_ghosts_side(c->g, side, c->state[buf[r]], c->stream);
                lsc_count_launches(1);
                continue;
            }
            lsc_mhd_ctx* o = T.m[nb];
            CUDA_TRY(cudaStreamWaitEvent(c->stream, o->halo_event, 0));
            double* dst = c->state[buf[r]] + (side == 0 ? 0 : (size_t)(c->g.nxl + 2) * col);
            // left ghost <- neighbour's last two interior columns; right ghost <- its first two
            const double* src = o->state[buf[nb]] + (side == 0 ? (size_t)o->g.nxl * col : 2 * col);
            CUDA_TRY(cudaMemcpyPeerAsync(dst, c->cfg.device, src, o->cfg.device, bytes, c->stream));
        }
    }
    CUDA_TRY(cudaGetLastError());
    return LSC_OK;
}

// status[0] of every member <- maximum over the members (ComputeTimeStep's reduction is global, M:71-115).
// NCCL rank: one max-allreduce.  Team: 8-byte peer copies into the leader's gather buffer, one tiny kernel there,
// 8-byte peer copies back -- stream-ordered by events, no host involvement, no kernel-level peer access needed.
int reduce_max_speed(const Team& T) {
    if (T.n == 1) {
        lsc_mhd_ctx* c = T.lead();
        if (c->cfg.nranks == 1) return LSC_OK;
        if (!c->comm) return fail(LSC_ESTATE, "nranks > 1 but lsc_mhd_comm_init was not called");
        ProfScope prof(c->profiling, c->stream, &c->reduce_events);
        if (lsc::comm_allreduce_max_u64(c->comm, c->d_status, 1, c->stream))
            return fail(LSC_ECOMM, "allreduce failed: %s", lsc::comm_error());
        prof.end();
        return LSC_OK;
    }
    lsc_mhd_ctx* L = T.lead();
    for (int r = 0; r < T.n; ++r) {
        lsc_mhd_ctx* c = T.m[r];
        CUDA_TRY(cudaSetDevice(c->cfg.device));
        CUDA_TRY(cudaMemcpyPeerAsync(L->d_gather + r, L->cfg.device, c->d_status, c->cfg.device,
                                     sizeof(unsigned long long), c->stream));
        CUDA_TRY(cudaEventRecord(c->red_event, c->stream));
    }
    CUDA_TRY(cudaSetDevice(L->cfg.device));
    for (int r = 1; r < T.n; ++r) CUDA_TRY(cudaStreamWaitEvent(L->stream, T.m[r]->red_event, 0));
    lsc::launch_max_words(L->d_gather, T.n, L->stream);  // d_gather[n] = max(d_gather[0..n))
    lsc_count_launches(1);
    CUDA_TRY(cudaEventRecord(L->max_event, L->stream));
    for (int r = 0; r < T.n; ++r) {
        lsc_mhd_ctx* c = T.m[r];
        CUDA_TRY(cudaSetDevice(c->cfg.device));
        if (r > 0) CUDA_TRY(cudaStreamWaitEvent(c->stream, L->max_event, 0));
        CUDA_TRY(cudaMemcpyPeerAsync(c->d_status, c->cfg.device, L->d_gather + T.n, L->cfg.device,
                                     sizeof(unsigned long long), c->stream));
    }
    return LSC_OK;
}

// The two sweeps of one step (MUSCL_NEW.cpp:841-965) for one context, reading state[cur] / state[prev] and leaving
// u^{n+1} in state[prev]'s buffer.  The step's scalars are in the context's __constant__ slot already.
int launch_sweeps(lsc_mhd_ctx* c, int cur, int prev, bool first_step) {
    lsc::SweepArgs p{};
    p.gamma = c->cfg.gamma;
    p.gm1 = c->cfg.gamma - 1;
    p.first_step = first_step ? 1 : 0;
    p.slot = c->slot;

    // ---- x sweep (:841-902)
    p.bc_lo = (c->cfg.rank == 0) ? to_kernel_bc(c->cfg.bc_x) : -1;
    p.bc_hi = (c->cfg.rank == c->cfg.nranks - 1) ? to_kernel_bc(c->cfg.bc_x) : -1;
    if (c->cfg.nranks > 1 && c->cfg.bc_x == LSC_BC_PERIODIC) p.bc_lo = p.bc_hi = -1;
    {
        ProfScope prof(c->profiling, c->stream, &c->prof_events);
        if (c->relaxed)
            lsc::launch_x_sweep_relaxed(c->g, p, c->state[cur], c->state[prev], c->state[c->bar], c->d_xflags,
                                        c->stream);
        else
            lsc::launch_x_sweep(c->g, p, c->state[cur], c->state[prev], c->state[c->bar], c->d_status, c->d_xflags,
                                c->stream);
        prof.end();
    }

    // ---- y sweep (:905-965) + copy (:969) + next step's signal speed (:71-115)
    p.bc_lo = p.bc_hi = to_kernel_bc(c->cfg.bc_y);
    auto y_launch = [&](int part) {
        if (c->relaxed)
            lsc::launch_y_sweep_relaxed(c->g, p, c->state[c->bar], c->state[cur], c->state[prev], c->d_status,
                                        c->d_fix, c->stream, part);
        else
            lsc::launch_y_sweep(c->g, p, c->state[c->bar], c->state[cur], c->state[prev], c->d_status, c->d_fix,
                                c->stream, part);
    };
    if (overlap_active(c)) {
        // The four slab-edge columns of u^{n+1} (what the x neighbours need) first; their halo exchange (:971 for
        // the ghost columns) then runs on comm_stream while this stream sweeps the interior columns.  The interior
        // launch neither reads nor writes the edge or ghost columns of the new state.
        {
            ProfScope prof(c->profiling, c->stream, &c->prof_events);
            y_launch(lsc::Y_EDGES);
            CUDA_TRY(cudaEventRecord(c->edges_done, c->stream));
            y_launch(lsc::Y_INTERIOR);
            prof.end();  // y time = both launches, not the halo wait
        }
        CUDA_TRY(cudaStreamWaitEvent(c->comm_stream, c->edges_done, 0));
        int rc = fill_x_ghosts(c, prev, c->comm_stream);
        if (rc) return rc;
        CUDA_TRY(cudaEventRecord(c->halo_done, c->comm_stream));
        CUDA_TRY(cudaStreamWaitEvent(c->stream, c->halo_done, 0));  // everything later on `stream` sees the halos
        lsc_count_launches(2);
    } else {
        ProfScope prof(c->profiling, c->stream, &c->prof_events);
        y_launch(lsc::Y_ALL);
        prof.end();
    }
    lsc_count_launches(4);
    CUDA_TRY(cudaGetLastError());
    return LSC_OK;
}

// One pass of the loop body MUSCL_NEW.cpp:831-971 on every member, enqueued; cur/prev are rotated for the next one.
int enqueue_one_step(const Team& T, std::vector<int>& cur, std::vector<int>& prev, bool first_step) {
    NvtxRange range("lsc step");
    lsc_mhd_ctx* L = T.lead();
    int rc = reduce_max_speed(T);  // ComputeTimeStep (:833): the reduction itself ran inside the last y sweep
    if (rc) return rc;
    for (int r = 0; r < T.n; ++r) {
        lsc_mhd_ctx* c = T.m[r];
        CUDA_TRY(cudaSetDevice(c->cfg.device));
        lsc::launch_prepare(c->d_ctl, c->d_cmd, c->d_status, c->d_log, c->stream);
        lsc_count_launches(1);
        CUDA_TRY(c->relaxed ? lsc::publish_step_scalars_relaxed(c->d_ctl, c->slot, c->stream)
                            : lsc::publish_step_scalars(c->d_ctl, c->slot, c->stream));
        rc = launch_sweeps(c, cur[r], prev[r], first_step);
        if (rc) return rc;
        std::swap(cur[r], prev[r]);  // u = uPlus1 (:969); the old u^n stays as the limiter's "previous" state
    }
    if (T.n > 1) return team_halo(T, cur.data());     // applyBoundaryConditions(u) (:971)
    if (!overlap_active(L)) return fill_x_ghosts(L, cur[0]);
    return LSC_OK;  // issued behind the slab-edge columns of the y sweep
}

// A plain context (one GPU, no communicator) replays GRAPH_STEPS steps as one CUDA graph: the seven small launches
// of a step then cost ~1 us each instead of ~2.5, which is what a step of a small grid is made of (OT 256^2: the
// sweeps themselves take ~25 us).  GRAPH_STEPS is even, so a replay leaves the buffer rotation where it was; the
// graph is captured from the very same enqueue code and keyed on everything that is baked into it.
constexpr int GRAPH_STEPS = 8;

bool graph_eligible(const Team& T) {
    const lsc_mhd_ctx* c = T.lead();
    return T.n == 1 && c->cfg.nranks == 1 && !c->profiling && !c->graph_disabled;
}

int graph_for(const Team& T, int cur0, cudaGraphExec_t* out) {
    lsc_mhd_ctx* c = T.lead();
    const int key = cur0 | (c->relaxed ? 8 : 0) | (c->overlap << 4);
    for (auto& e : c->graphs)
        if (e.key == key) {
            *out = e.exec;
            return LSC_OK;
        }
    *out = nullptr;
    std::vector<int> cur{cur0}, prev{cur0 == c->cur ? c->prev : c->cur};
    const unsigned long long before = g_launches.load();
    if (cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
        cudaGetLastError();
        c->graph_disabled = true;
        return LSC_OK;
    }
    int rc = LSC_OK;
    for (int q = 0; q < GRAPH_STEPS && rc == LSC_OK; ++q) rc = enqueue_one_step(T, cur, prev, false);
    cudaGraph_t graph = nullptr;
    cudaError_t e = cudaStreamEndCapture(c->stream, &graph);
    const unsigned long long captured = g_launches.load() - before;
    g_launches -= captured;  // nothing ran yet: launches are counted per replay
    cudaGraphExec_t exec = nullptr;
    if (rc == LSC_OK && e == cudaSuccess && graph) e = cudaGraphInstantiate(&exec, graph, 0);
    if (graph) cudaGraphDestroy(graph);
    if (rc != LSC_OK || e != cudaSuccess || !exec) {  // capture not possible here: stay on plain launches
        cudaGetLastError();
        c->graph_disabled = true;
        return LSC_OK;
    }
    c->graphs.push_back({key, exec, captured});
    *out = exec;
    return LSC_OK;
}

// Enqueue `count` passes of the loop body on every member.  Nothing is read back.
int enqueue_steps(const Team& T, long count) {
    lsc_mhd_ctx* L = T.lead();
    std::vector<int> cur(T.n), prev(T.n);
    for (int r = 0; r < T.n; ++r) {
        cur[r] = T.m[r]->cur;
        prev[r] = T.m[r]->prev;
    }
    if (!L->ghosts_current) {  // applyBoundaryConditions(u) (:838): only x ghost columns are read on the device
        int rc = (T.n == 1) ? fill_x_ghosts(L, L->cur) : team_halo(T, cur.data());
        if (rc) return rc;
        for (int r = 0; r < T.n; ++r) T.m[r]->ghosts_current = true;
    }
    long e = 0;
    while (e < count) {
        if (count - e >= GRAPH_STEPS && L->step + e >= 1 && graph_eligible(T)) {
            cudaGraphExec_t exec = nullptr;
            int rc = graph_for(T, cur[0], &exec);
            if (rc) return rc;
            if (exec) {
                CUDA_TRY(cudaGraphLaunch(exec, L->stream));
                for (auto& g : L->graphs)
                    if (g.exec == exec) lsc_count_launches(g.launches);
                e += GRAPH_STEPS;
                continue;
            }
        }
        int rc = enqueue_one_step(T, cur, prev, L->step + e == 0);
        if (rc) return rc;
        e += 1;
    }
    return LSC_OK;
}

struct BatchResult {
    long done = 0;  // steps that really ran
    int halt = lsc::HALT_NONE;
};

// One batch: command block -> device, `count` steps enqueued, ONE synchronisation, host mirrors brought up to date.
int run_batch(const Team& T, long count, double t_stop, int dump_counter, long long step_limit, bool keep_log,
              BatchResult* res) {
    NvtxRange range("lsc batch");
    lsc_mhd_ctx* L = T.lead();
    for (int r = 0; r < T.n; ++r) {
        lsc_mhd_ctx* c = T.m[r];
        CUDA_TRY(cudaSetDevice(c->cfg.device));
        c->batch += 1;
        c->batches += 1;
        lsc::StepCmd& cmd = c->h->cmd;
        cmd.batch = c->batch;
        cmd.step_limit = step_limit;
        cmd.t_stop = t_stop;
        cmd.dump_counter = dump_counter;
        cmd.keep_log = (keep_log && r == 0) ? 1 : 0;
        CUDA_TRY(cudaMemcpyAsync(c->d_cmd, &cmd, sizeof cmd, cudaMemcpyHostToDevice, c->stream));
    }
    int rc = enqueue_steps(T, count);
    // read back whatever was enqueued, also after an error part-way (the streams must drain before anyone returns)
    for (int r = 0; r < T.n; ++r) {
        lsc_mhd_ctx* c = T.m[r];
        cudaSetDevice(c->cfg.device);
        // everything but the exp table (the tail of StepCtl)
        cudaMemcpyAsync(&c->h->ctl, c->d_ctl, offsetof(lsc::StepCtl, tab_arg), cudaMemcpyDeviceToHost, c->stream);
        cudaMemcpyAsync(c->h->status, c->d_status, STATUS_BYTES, cudaMemcpyDeviceToHost, c->stream);
    }
    for (int r = 0; r < T.n; ++r) {
        cudaSetDevice(T.m[r]->cfg.device);
        cudaError_t e = cudaStreamSynchronize(T.m[r]->stream);
        if (e != cudaSuccess && rc == LSC_OK)
            rc = fail(LSC_ECUDA, "step batch failed on slab %d: %s", r, cudaGetErrorString(e));
    }
    if (rc) return rc;
    const long before = L->step;
    for (int r = 0; r < T.n; ++r) {
        lsc_mhd_ctx* c = T.m[r];
        const lsc::StepCtl& k = c->h->ctl;
        if (k.step != L->h->ctl.step || k.halt != L->h->ctl.halt)
            return fail(LSC_ESTATE, "slab %d is at step %lld (halt %d), the leader at %lld (halt %d)", r, k.step,
                        k.halt, L->h->ctl.step, L->h->ctl.halt);
        const long done = (long)k.step - c->step;
        if (done & 1) std::swap(c->cur, c->prev);
        c->step = (long)k.step;
        c->t = k.t;
        c->dt = k.dt;
        c->ch = k.max_speed;
        c->ieee_redos = (long)(k.redo_total + c->h->status[1] + c->h->status[2] + c->h->status[3]);
    }
    res->done = L->step - before;
    res->halt = L->h->ctl.halt;
    if (keep_log && L->h->ctl.log_count > 0) {
        CUDA_TRY(cudaSetDevice(L->cfg.device));
        CUDA_TRY(cudaMemcpyAsync(L->h->log, L->d_log, sizeof(double) * 3 * (size_t)L->h->ctl.log_count,
                                 cudaMemcpyDeviceToHost, L->stream));
        CUDA_TRY(cudaStreamSynchronize(L->stream));
    }
    return LSC_OK;
}

// What the host does about a batch that stopped for a reason of its own (not a pause / limit).
int handle_halt(const Team& T, const BatchResult& res) {
    lsc_mhd_ctx* L = T.lead();
    if (res.halt == lsc::HALT_NAN)
        return fail(LSC_ENAN, "max signal speed is %g at step %ld", L->h->ctl.max_speed, L->step);
    if (res.halt == lsc::HALT_EXP_MISS) {
        // the psi-damping argument of this step is not in the table: glibc's exp of that very argument joins it
        double arg;
        std::memcpy(&arg, &L->h->ctl.miss_arg, sizeof arg);
        for (int r = 0; r < T.n; ++r) {
            lsc_mhd_ctx* c = T.m[r];
            c->exp_misses += 1;
            if (c->ntab >= lsc::EXP_TAB_CAP) c->ntab = lsc::EXP_TAB_CAP / 2;  // recycle the upper half
            c->tab_arg[c->ntab] = L->h->ctl.miss_arg;
            c->tab_val[c->ntab] = std::exp(arg);  // MUSCL_NEW.cpp:67
            c->ntab += 1;
            CUDA_TRY(cudaSetDevice(c->cfg.device));
            // tab_arg, tab_val and ntab of the device copy; the rest of the loop state stays as it is
            CUDA_TRY(cudaMemcpy(reinterpret_cast<char*>(c->d_ctl) + offsetof(lsc::StepCtl, tab_arg), c->tab_arg,
                                sizeof c->tab_arg, cudaMemcpyHostToDevice));
            CUDA_TRY(cudaMemcpy(reinterpret_cast<char*>(c->d_ctl) + offsetof(lsc::StepCtl, tab_val), c->tab_val,
                                sizeof c->tab_val, cudaMemcpyHostToDevice));
            CUDA_TRY(cudaMemcpy(reinterpret_cast<char*>(c->d_ctl) + offsetof(lsc::StepCtl, ntab), &c->ntab,
                                sizeof c->ntab, cudaMemcpyHostToDevice));
        }
    }
    return LSC_OK;
}

int team_check(lsc_mhd_ctx** g, int n) {
    if (!g || n < 1) return fail(LSC_EINVAL, "empty group");
    for (int r = 0; r < n; ++r) {
        if (!g[r]) return fail(LSC_EINVAL, "null context in group");
        if (g[r]->cfg.nranks != n || g[r]->cfg.rank != r)
            return fail(LSC_EINVAL, "group member %d has rank %d of %d", r, g[r]->cfg.rank, g[r]->cfg.nranks);
        if (!g[r]->uploaded) return fail(LSC_ESTATE, "group member %d has no state uploaded", r);
        if (g[r]->comm) return fail(LSC_ESTATE, "group member %d is attached to NCCL", r);
        if (g[r]->step != g[0]->step) return fail(LSC_ESTATE, "group members are at different steps");
    }
    // direct NVLink / PCIe peer copies between the members' GPUs (without it cudaMemcpyPeerAsync stages through
    // the host); enabled once per pair, "already enabled" is not an error
    for (int r = 0; r < n; ++r) {
        if (g[r]->peers_enabled) continue;
        CUDA_TRY(cudaSetDevice(g[r]->cfg.device));
        for (int q = 0; q < n; ++q) {
            const int a = g[r]->cfg.device, b = g[q]->cfg.device;
            if (a == b) continue;
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, a, b) == cudaSuccess && can) {
                cudaError_t e = cudaDeviceEnablePeerAccess(b, 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled)
                    return fail(LSC_ECUDA, "cudaDeviceEnablePeerAccess(%d -> %d): %s", a, b, cudaGetErrorString(e));
            }
            cudaGetLastError();
        }
        g[r]->peers_enabled = true;
    }
    return LSC_OK;
}

struct TeamScope {  // marks the members for the duration of a group call, on every exit path
    const Team& T;
    explicit TeamScope(const Team& t) : T(t) {
        for (int r = 0; r < T.n; ++r) T.m[r]->in_team = T.n > 1;
    }
    ~TeamScope() {
        for (int r = 0; r < T.n; ++r) T.m[r]->in_team = false;
    }
};

int check_plain(lsc_mhd_ctx* c, const char* what) {
    if (!c) return fail(LSC_EINVAL, "null argument");
    if (!c->uploaded) return fail(LSC_ESTATE, "%s before lsc_mhd_upload", what);
    if (c->cfg.nranks > 1 && !c->comm)
        return fail(LSC_ESTATE, "nranks > 1: call lsc_mhd_comm_init (one process per GPU) or use the lsc_mhd_group_* "
                                "calls (slabs owned by one process)");
    return LSC_OK;
}

// nsteps passes of the loop body, however many batches that takes
int team_step(const Team& T, long nsteps) {
    lsc_mhd_ctx* L = T.lead();
    const long long target = (long long)L->step + nsteps;
    while (L->step < target) {
        BatchResult res;
        const long count = (long)std::min<long long>(target - L->step, BATCH_MAX);
        int rc = run_batch(T, count, std::numeric_limits<double>::infinity(), -1, target, false, &res);
        if (rc) return rc;
        rc = handle_halt(T, res);
        if (rc) return rc;
    }
    return LSC_OK;
}

// The loop MUSCL_NEW.cpp:831-1003 with its log line (:836), dump cadence (:975-1001) and exit test (:1003).
int team_run_until(const Team& T, double t_stop, lsc_mhd_dump_cb dump_cb, lsc_mhd_log_cb log_cb, void* user,
                   long* steps_out) {
    lsc_mhd_ctx* L = T.lead();
    int counter = 0;
    long n = 0;
    bool again = true;
    while (again) {
        // how many steps until the device will pause, judged by the last dt (a wrong guess costs either a few
        // no-op launches or one more batch)
        long count = 1;
        if (L->dt > 0) {
            double next = t_stop;
            if (0.1 * counter > L->t) next = std::min(next, 0.1 * counter);
            const double est = (next - L->t) / L->dt;
            count = est > (double)BATCH_MAX ? BATCH_MAX : (est < 1.0 ? 1 : (long)est + 2);
        }
        if (log_cb) count = std::min<long>(count, lsc::STEP_LOG_CAP);
        BatchResult res;
        int rc = run_batch(T, count, t_stop, counter, LLONG_MAX, log_cb != nullptr, &res);
        if (rc) return rc;
        n += res.done;
        if (log_cb) {
            for (long q = 0; q < res.done; ++q) {  // :836
                const double* rec = L->h->log + 3 * q;
                if (log_cb(L->step - res.done + q + 1, rec[0], rec[1], rec[2], user))
                    return fail(LSC_ECALLBACK, "log callback stopped the loop at step %ld", L->step - res.done + q + 1);
            }
        }
        rc = handle_halt(T, res);
        if (rc) return rc;
        if (res.done == 0) continue;  // the batch stopped before its first step (exp table miss): go again
        // predicates of the step that ended the batch; the device evaluated the same ones for all earlier steps
        if (L->t > 0.1 * counter && L->t - L->dt <= 0.1 * counter) {  // :975
            if (dump_cb) {
                rc = dump_cb(L, counter, L->t, L->dt, user);
                if (rc) return fail(rc, "dump callback returned %d", rc);
            }
            counter += 1;  // :1000
        }
        again = L->t < t_stop;  // :1003
    }
    if (steps_out) *steps_out = n;
    return LSC_OK;
}

int team_step_timed(const Team& T, long nsteps, float* ms) {
    lsc_mhd_ctx* L = T.lead();
    CUDA_TRY(cudaSetDevice(L->cfg.device));
    cudaEvent_t e0, e1;
    CUDA_TRY(cudaEventCreate(&e0));
    CUDA_TRY(cudaEventCreate(&e1));
    for (int r = 0; r < T.n; ++r) {
        cudaSetDevice(T.m[r]->cfg.device);
        cudaStreamSynchronize(T.m[r]->stream);
    }
    cudaSetDevice(L->cfg.device);
    cudaEventRecord(e0, L->stream);
    int rc = team_step(T, nsteps);
    if (rc == LSC_OK) {
        cudaSetDevice(L->cfg.device);
        cudaError_t e = cudaEventRecord(e1, L->stream);
        if (e == cudaSuccess) e = cudaEventSynchronize(e1);
        if (e == cudaSuccess) e = cudaEventElapsedTime(ms, e0, e1);
        if (e != cudaSuccess) rc = fail(LSC_ECUDA, "timing events: %s", cudaGetErrorString(e));
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    return rc;
}

}  // namespace

extern "C" {

int lsc_mhd_step(lsc_mhd_ctx* c, long nsteps) {
    int rc = check_plain(c, "lsc_mhd_step");
    if (rc) return rc;
    Team T{&c, 1};
    return team_step(T, nsteps);
}

int lsc_mhd_step_timed(lsc_mhd_ctx* c, long nsteps, float* ms) {
    if (!ms) return fail(LSC_EINVAL, "null argument");
    int rc = check_plain(c, "lsc_mhd_step_timed");
    if (rc) return rc;
    Team T{&c, 1};
    return team_step_timed(T, nsteps, ms);
}

int lsc_mhd_run_until(lsc_mhd_ctx* c, double t_stop, lsc_mhd_dump_cb dump_cb, lsc_mhd_log_cb log_cb, void* user,
                      long* steps_out) {
    int rc = check_plain(c, "lsc_mhd_run_until");
    if (rc) return rc;
    Team T{&c, 1};
    return team_run_until(T, t_stop, dump_cb, log_cb, user, steps_out);
}

int lsc_mhd_group_step(lsc_mhd_ctx** g, int n, long nsteps) {
    int rc = team_check(g, n);
    if (rc) return rc;
    Team T{g, n};
    TeamScope scope(T);
    return team_step(T, nsteps);
}

int lsc_mhd_group_step_timed(lsc_mhd_ctx** g, int n, long nsteps, float* ms) {
    if (!ms) return fail(LSC_EINVAL, "null argument");
    int rc = team_check(g, n);
    if (rc) return rc;
    Team T{g, n};
    TeamScope scope(T);
    return team_step_timed(T, nsteps, ms);
}

int lsc_mhd_group_run_until(lsc_mhd_ctx** g, int n, double t_stop, lsc_mhd_dump_cb dump_cb, lsc_mhd_log_cb log_cb,
                            void* user, long* steps_out) {
    int rc = team_check(g, n);
    if (rc) return rc;
    Team T{g, n};
    TeamScope scope(T);
    return team_run_until(T, t_stop, dump_cb, log_cb, user, steps_out);
}

long lsc_mhd_ieee_fallback_count(const lsc_mhd_ctx* c) { return c ? c->ieee_redos : -1; }

int lsc_mhd_loop_stats(const lsc_mhd_ctx* c, long* host_syncs, long* exp_table_misses) {
    if (!c) return fail(LSC_EINVAL, "null argument");
    if (host_syncs) *host_syncs = c->batches;
    if (exp_table_misses) *exp_table_misses = c->exp_misses;
    return LSC_OK;
}

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
    cudaError_t err = cudaSuccess;
    for (size_t q = 0; q + 3 < c->prof_events.size() && err == cudaSuccess; q += 4) {
        float a = 0, b = 0;
        err = cudaEventElapsedTime(&a, c->prof_events[q], c->prof_events[q + 1]);
        if (err == cudaSuccess) err = cudaEventElapsedTime(&b, c->prof_events[q + 2], c->prof_events[q + 3]);
        ms[0] += a;
        ms[1] += b;
        launches[0] += 1;
        launches[1] += 1;
    }
    destroy_events(c->prof_events);
    CUDA_TRY(err);
    return LSC_OK;
}

int lsc_mhd_comm_times(lsc_mhd_ctx* c, double ms[2], long calls[2]) {
    if (!c || !ms || !calls) return fail(LSC_EINVAL, "null argument");
    CUDA_TRY(cudaSetDevice(c->cfg.device));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    CUDA_TRY(cudaStreamSynchronize(c->comm_stream));
    std::vector<cudaEvent_t>* lists[2] = {&c->halo_events, &c->reduce_events};
    cudaError_t err = cudaSuccess;
    for (int w = 0; w < 2; ++w) {
        ms[w] = 0;
        calls[w] = 0;
        for (size_t q = 0; q + 1 < lists[w]->size() && err == cudaSuccess; q += 2) {
            float a = 0;
            err = cudaEventElapsedTime(&a, (*lists[w])[q], (*lists[w])[q + 1]);
            ms[w] += a;
            calls[w] += 1;
        }
        destroy_events(*lists[w]);
    }
    CUDA_TRY(err);
    return LSC_OK;
}

}  // extern "C"

namespace {
// slab-local columns [il_lo, il_hi) of the current state, reference layout, to dst (column il_lo first)
int download_columns(lsc_mhd_ctx* c, int il_lo, int il_hi, double* dst) {
    CUDA_TRY(cudaSetDevice(c->cfg.device));
    const size_t col = (size_t)(c->cfg.ny + 4) * lsc::NV;
    // as in lsc_mhd_upload: chunk k is copied to the host (copy_stream) while chunk k+1 is transposed (stream)
    const int half = std::max(1, c->stage_cols / 2);
    int k = 0;
    for (int il = il_lo; il < il_hi; il += half, ++k) {
        const int ni = std::min(half, il_hi - il);
        const int b = k & 1;
        double* buf = c->d_stage + (size_t)b * half * col;
        if (k >= 2) CUDA_TRY(cudaStreamWaitEvent(c->stream, c->stage_free[b], 0));
        lsc::launch_soa_to_aos(c->g, c->state[c->cur], il, ni, buf, c->stream);
        lsc_count_launches(1);
        CUDA_TRY(cudaGetLastError());
        CUDA_TRY(cudaEventRecord(c->stage_filled[b], c->stream));
        CUDA_TRY(cudaStreamWaitEvent(c->copy_stream, c->stage_filled[b], 0));
        CUDA_TRY(cudaMemcpyAsync(dst + col * (il - il_lo), buf, col * ni * sizeof(double), cudaMemcpyDeviceToHost,
                                 c->copy_stream));
        CUDA_TRY(cudaEventRecord(c->stage_free[b], c->copy_stream));
    }
    CUDA_TRY(cudaStreamSynchronize(c->copy_stream));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    return LSC_OK;
}
}  // namespace

extern "C" {

int lsc_mhd_download(lsc_mhd_ctx* c, double* aos) {
    if (!c || !aos) return fail(LSC_EINVAL, "null argument");
    if (!c->uploaded) return fail(LSC_ESTATE, "lsc_mhd_download before lsc_mhd_upload");
    int rc = download_columns(c, 0, c->nxl + 4, aos);
    if (rc) return rc;
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

int lsc_mhd_group_download(lsc_mhd_ctx** g, int n, double* aos) {
    int rc = team_check(g, n);
    if (rc) return rc;
    if (!aos) return fail(LSC_EINVAL, "null argument");
    const size_t col = (size_t)(g[0]->cfg.ny + 4) * lsc::NV;
    for (int r = 0; r < n; ++r) {  // every slab writes its own columns (and the outer ghosts at the two ends)
        lsc_mhd_ctx* c = g[r];
        const int lo = (r == 0) ? 0 : 2, hi = (r == n - 1) ? c->nxl + 4 : c->nxl + 2;
        rc = download_columns(c, lo, hi, aos + col * (size_t)(c->i0 + lo));
        if (rc) return rc;
    }
    if (g[0]->step > 0) {
        lsc_mhd_config whole = g[0]->cfg;
        whole.rank = 0;
        whole.nranks = 1;
        whole.slab_i0 = whole.slab_nx = 0;
        lsc::mhd_fill_ghost_ring(whole, aos);
    }
    return LSC_OK;
}

int lsc_mhd_state_hash(lsc_mhd_ctx* c, unsigned long long* hash) {
    if (!c || !hash) return fail(LSC_EINVAL, "null argument");
    if (!c->uploaded) return fail(LSC_ESTATE, "lsc_mhd_state_hash before lsc_mhd_upload");
    CUDA_TRY(cudaSetDevice(c->cfg.device));
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

// No cudaMalloc / cudaFree on this path: the scratch words are a file-scope __device__ array of state_hash.cu.
int lsc_mhd_state_hash_noalloc(lsc_mhd_ctx* c, unsigned long long* hash) {
    if (!c || !hash) return fail(LSC_EINVAL, "null argument");
    static_assert(lsc::STATE_HASH_STATIC_MAX_NY == LSC_STATE_HASH_NOALLOC_MAX_NY, "header and kernel limit differ");
    if (c->g.ny > lsc::STATE_HASH_STATIC_MAX_NY)
        return fail(LSC_EINVAL, "lsc_mhd_state_hash_noalloc: ny = %d exceeds the static scratch array's limit of %d rows "
                                "(LSC_STATE_HASH_NOALLOC_MAX_NY); use lsc_mhd_state_hash", c->g.ny,
                    lsc::STATE_HASH_STATIC_MAX_NY);
    if (!c->uploaded) return fail(LSC_ESTATE, "lsc_mhd_state_hash_noalloc before lsc_mhd_upload");
    CUDA_TRY(cudaSetDevice(c->cfg.device));
    CUDA_TRY(lsc::state_hash_static(c->cfg.device, c->g, lsc::NV, c->state[c->cur], c->i0, hash, c->stream));
    lsc_count_launches(2);
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

// lsc_euler_capi.cu -- C ABI of the 2-D Euler SLIC solver (include/lsc_b200.h, lsc_euler_*): the time loop of
// Euler_codes/SLIC_euler_2d.cpp main() (S:528-643) / generaliseSLIC.cpp main() (G:564-668) driving the
// sm_100a kernels of euler_kernels.cu.
#include <climits>
#include <cmath>
#include <cstring>
#include <limits>
#include <new>
#include <vector>

#include "euler_host.h"
#include "euler_kernels.cuh"
#include "lsc_comm.h"
#include "lsc_common.h"
#include "state_hash.h"

namespace {
// everything the host reads back after a batch of steps, in one pinned block (cf. lsc_capi.cu)
struct EulerHostBlock {
    lsc::StepCmd cmd;
    lsc::EulerStepCtl ctl;
    unsigned long long status[4];
    double log[3 * lsc::STEP_LOG_CAP];
};
constexpr long E_BATCH_MAX = 4096;
constexpr int E_GRAPH_STEPS = 8;
}  // namespace

struct lsc_euler_ctx {
    lsc_euler_config cfg;
    int nxl = 0, i0 = 0;  // slab: interior columns, first global interior column (0-based)
    double dx = 0, dy = 0;
    lsc::GridDims g{};
    // u^n, ubar, u^{n+1}
    double* state[3] = {nullptr, nullptr, nullptr};
    int cur = 0, bar = 1, next = 2;
    // [0] max signal speed bits of state[cur], [1]/[2] x/y blocks recomputed with IEEE division, [3] slab-edge y launch
    unsigned long long* d_status = nullptr;
    unsigned* d_fix = nullptr;
    // device-resident loop state (euler_kernels.cuh: EulerStepCtl; step_ctl.cuh: StepCmd) and its host mirror
    lsc::EulerStepCtl* d_ctl = nullptr;
    lsc::StepCmd* d_cmd = nullptr;
    double* d_log = nullptr;
    unsigned long long* d_gather = nullptr;  // team leader: members' status[0] words, then their maximum
    EulerHostBlock* h = nullptr;             // pinned
    int slot = -1;
    unsigned long long batch = 0;
    long batches = 0;
    long ieee_redos = 0;
    double* d_stage = nullptr;
    int stage_cols = 0;
    // whole-state transfers use the staging buffer as two halves: the PCIe copy of one chunk (copy_stream) overlaps
    // the layout transposition of the other (stream); created at the first transfer
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t stage_filled[2] = {nullptr, nullptr}, stage_free[2] = {nullptr, nullptr};
    cudaStream_t stream = nullptr;
    // halo/compute overlap as in the MHD solver (lsc_capi.cu): exchange on comm_stream behind the slab-edge columns
    cudaStream_t comm_stream = nullptr;
    cudaEvent_t edges_done = nullptr, halo_done = nullptr;
    int overlap = 1;  // 0 off, 1 with NCCL neighbours, 2 always
    double t = 0, dt = 0;
    long step = 0;
    bool uploaded = false;
    bool profiling = false;
    std::vector<cudaEvent_t> prof_events;
    int* d_hints = nullptr;            // 3 search brackets per y-sweep block, carried from step to step
    double* d_table = nullptr;         // tabulated EoS: dens[500] pres[500] sound[500*500] energies[500*500]
    std::vector<double> h_table;       // host copy (IC, .dat output)
    bool table_monotone = false;       // energy columns ascending with margin: searches may be warm-started
    bool warm_search = true;           // lsc_euler_set_eos_search
    lsc::Comm* comm = nullptr;         // NCCL mode (one process per GPU)
    bool in_group = false;             // driven by the lsc_euler_group_* calls
    bool ghosts_current = false;       // x ghost / halo columns of state[cur] hold boundary images / neighbour cells
    cudaEvent_t halo_event = nullptr;  // team: "new u of this slab complete"
    cudaEvent_t red_event = nullptr;   // team: "status word of this slab is in the leader's gather buffer"
    cudaEvent_t max_event = nullptr;   // team leader: "the maximum over the slabs is ready"
    struct StepGraph {
        int key;
        cudaGraphExec_t exec;
        unsigned long long launches;
    };
    std::vector<StepGraph> graphs;  // CUDA graphs of E_GRAPH_STEPS steps (plain contexts)
    bool graph_disabled = false;
};

namespace {

int kernel_bc(int bc) { return bc == LSC_BC_PERIODIC ? (int)lsc::BC_PERIODIC : (int)lsc::BC_TRANSMISSIVE_E; }

constexpr size_t TAB_N = 500, TAB_DOUBLES = 2 * TAB_N + 2 * TAB_N * TAB_N;

// EoS part of the kernel parameters: ideal gas, or the table once lsc_euler_set_eos_table was called
void fill_eos(const lsc_euler_ctx* c, lsc::EulerParams& p) {
    p.gamma = c->cfg.gamma;
    p.gm1 = c->cfg.gamma - 1;
    p.tab_dens = p.tab_pres = p.tab_sound = p.tab_en = nullptr;
    p.tab_hints = nullptr;
    if (c->d_table) {
        p.tab_dens = c->d_table;
        p.tab_pres = c->d_table + TAB_N;
        p.tab_sound = c->d_table + 2 * TAB_N;
        p.tab_en = c->d_table + 2 * TAB_N + TAB_N * TAB_N;
    }
    p.tab_warm = (c->d_table && c->table_monotone && c->warm_search) ? 1 : 0;
    p.tab_hints = p.tab_warm ? c->d_hints : nullptr;
}

struct ETeam {  // a plain context / NCCL rank (n == 1) or the slabs of one grid owned by this process
    lsc_euler_ctx** m;
    int n;
    lsc_euler_ctx* lead() const { return m[0]; }
};

// boundary images / neighbour data into the two ghost columns on each side of state b (plain context / NCCL rank)
int e_fill_x_ghosts(lsc_euler_ctx* c, int b, cudaStream_t s = nullptr) {
    if (!s) s = c->stream;
    if (c->cfg.nranks <= 1) {
        lsc::launch_euler_x_ghosts(c->g, kernel_bc(c->cfg.bc), c->state[b], s);
        lsc_count_launches(1);
        LSC_CUDA_TRY(cudaGetLastError());
        return LSC_OK;
    }
    if (!c->comm) return lsc_fail(LSC_ESTATE, "nranks > 1 but lsc_euler_comm_init was not called");
    const int bc = c->cfg.bc == LSC_BC_PERIODIC ? lsc::BC_PERIODIC : lsc::BC_TRANSMISSIVE;
    if (lsc::comm_exchange_halos(c->comm, c->g, bc, c->state[b], s))
        return lsc_fail(LSC_ECOMM, "halo exchange failed: %s", lsc::comm_error());
    return LSC_OK;
}

bool e_overlap_active(const lsc_euler_ctx* c) {
    if (c->in_group || c->g.nxl < 8) return false;
    return c->overlap == 2 || (c->overlap == 1 && c->cfg.nranks > 1 && c->comm);
}

// team halos of state[buf[r]]: every slab records "complete", then pulls its neighbours' edge columns
int eg_halo(const ETeam& T, const int* buf) {
    const bool ring = (T.lead()->cfg.bc == LSC_BC_PERIODIC);
    for (int r = 0; r < T.n; ++r) {
        LSC_CUDA_TRY(cudaSetDevice(T.m[r]->cfg.device));
        LSC_CUDA_TRY(cudaEventRecord(T.m[r]->halo_event, T.m[r]->stream));
    }
    for (int r = 0; r < T.n; ++r) {
        lsc_euler_ctx* c = T.m[r];
        LSC_CUDA_TRY(cudaSetDevice(c->cfg.device));
        const int left = (r > 0) ? r - 1 : (ring ? T.n - 1 : -1);
        const int right = (r < T.n - 1) ? r + 1 : (ring ? 0 : -1);
        const size_t col = lsc::column_doubles(c->g, 4);  // a column of the blocked layout is one contiguous run
        const size_t bytes = 2 * col * sizeof(double);
        for (int side = 0; side < 2; ++side) {
            const int nb = side == 0 ? left : right;
            if (nb < 0) {
                lsc::launch_euler_x_ghosts_side(c->g, side, c->state[buf[r]], c->stream);
                lsc_count_launches(1);
                continue;
            }
            lsc_euler_ctx* o = T.m[nb];
            LSC_CUDA_TRY(cudaStreamWaitEvent(c->stream, o->halo_event, 0));
            double* dst = c->state[buf[r]] + (side == 0 ? 0 : (size_t)(c->g.nxl + 2) * col);
            const double* src = o->state[buf[nb]] + (side == 0 ? (size_t)o->g.nxl * col : 2 * col);
            LSC_CUDA_TRY(cudaMemcpyPeerAsync(dst, c->cfg.device, src, o->cfg.device, bytes, c->stream));
        }
    }
    LSC_CUDA_TRY(cudaGetLastError());
    return LSC_OK;
}

// status[0] of every member <- maximum over the members (S:165-195 is a global reduction); see lsc_capi.cu
int e_reduce_max_speed(const ETeam& T) {
    if (T.n == 1) {
        lsc_euler_ctx* c = T.lead();
        if (c->cfg.nranks <= 1) return LSC_OK;
        if (!c->comm) return lsc_fail(LSC_ESTATE, "nranks > 1 but lsc_euler_comm_init was not called");
        if (lsc::comm_allreduce_max_u64(c->comm, c->d_status, 1, c->stream))
            return lsc_fail(LSC_ECOMM, "allreduce failed: %s", lsc::comm_error());
        return LSC_OK;
    }
    lsc_euler_ctx* L = T.lead();
    for (int r = 0; r < T.n; ++r) {
        lsc_euler_ctx* c = T.m[r];
        LSC_CUDA_TRY(cudaSetDevice(c->cfg.device));
        LSC_CUDA_TRY(cudaMemcpyPeerAsync(L->d_gather + r, L->cfg.device, c->d_status, c->cfg.device,
                                         sizeof(unsigned long long), c->stream));
        LSC_CUDA_TRY(cudaEventRecord(c->red_event, c->stream));
    }
    LSC_CUDA_TRY(cudaSetDevice(L->cfg.device));
    for (int r = 1; r < T.n; ++r) LSC_CUDA_TRY(cudaStreamWaitEvent(L->stream, T.m[r]->red_event, 0));
    lsc::launch_max_words(L->d_gather, T.n, L->stream);
    lsc_count_launches(1);
    LSC_CUDA_TRY(cudaEventRecord(L->max_event, L->stream));
    for (int r = 0; r < T.n; ++r) {
        lsc_euler_ctx* c = T.m[r];
        LSC_CUDA_TRY(cudaSetDevice(c->cfg.device));
        if (r > 0) LSC_CUDA_TRY(cudaStreamWaitEvent(c->stream, L->max_event, 0));
        LSC_CUDA_TRY(cudaMemcpyPeerAsync(c->d_status, c->cfg.device, L->d_gather + T.n, L->cfg.device,
                                         sizeof(unsigned long long), c->stream));
    }
    return LSC_OK;
}

// the host-known part of the kernel parameters; the step scalars come from the context's __constant__ slot
void fill_params(const lsc_euler_ctx* c, lsc::EulerParams& p) {
    fill_eos(c, p);
    p.slot = c->slot;
    p.cm = 0.5 * (1 + c->cfg.omega);
    p.cp = 0.5 * (1 - c->cfg.omega);
    p.bcy = kernel_bc(c->cfg.bc);
    // physical boundary type at each end of this slab; -1 at an inter-rank edge (ghost columns = neighbour data)
    const bool split = c->cfg.nranks > 1;
    const bool ring = split && c->cfg.bc == LSC_BC_PERIODIC;
    p.bcx_lo = (ring || (split && c->cfg.rank > 0)) ? -1 : kernel_bc(c->cfg.bc);
    p.bcx_hi = (ring || (split && c->cfg.rank < c->cfg.nranks - 1)) ? -1 : kernel_bc(c->cfg.bc);
}

// the two sweeps of one step for one context: state[in] -> state[out] (S:604-615)
int e_launch(lsc_euler_ctx* c, int in, int out) {
    lsc::EulerParams p{};
    fill_params(c, p);
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    const bool prof = c->profiling;
    if (prof) {
        for (auto& e : ev) LSC_CUDA_TRY(cudaEventCreate(&e));
        LSC_CUDA_TRY(cudaEventRecord(ev[0], c->stream));
    }
    lsc::launch_euler_x_sweep(c->g, p, c->state[in], c->state[c->bar], c->d_status, c->d_fix, c->stream);
    if (prof) {
        LSC_CUDA_TRY(cudaEventRecord(ev[1], c->stream));
        LSC_CUDA_TRY(cudaEventRecord(ev[2], c->stream));
    }
    auto y_launch = [&](int part) {
        lsc::launch_euler_y_sweep(c->g, p, c->state[c->bar], c->state[in], c->state[out], c->d_status, c->d_fix,
                                  c->stream, part);
    };
    const bool overlapped = e_overlap_active(c);
    if (overlapped) {
        // slab-edge columns of the new u first; their exchange (S:616 for the ghost columns) runs on comm_stream
        // while this stream sweeps the interior, which touches neither the edge nor the ghost columns of `out`
        y_launch(lsc::Y_EDGES);
        LSC_CUDA_TRY(cudaEventRecord(c->edges_done, c->stream));
        y_launch(lsc::Y_INTERIOR);
        lsc_count_launches(2);
    } else {
        y_launch(lsc::Y_ALL);
    }
    if (prof) {
        LSC_CUDA_TRY(cudaEventRecord(ev[3], c->stream));
        for (auto e : ev) c->prof_events.push_back(e);
    }
    lsc_count_launches(4);
    LSC_CUDA_TRY(cudaGetLastError());
    if (overlapped) {
        LSC_CUDA_TRY(cudaStreamWaitEvent(c->comm_stream, c->edges_done, 0));
        int rc = e_fill_x_ghosts(c, out, c->comm_stream);
        if (rc) return rc;
        LSC_CUDA_TRY(cudaEventRecord(c->halo_done, c->comm_stream));
        LSC_CUDA_TRY(cudaStreamWaitEvent(c->stream, c->halo_done, 0));  // later work on `stream` sees the halos
    }
    return LSC_OK;
}

// one pass of the do-body S:600-616 on every member, enqueued; cur/next are rotated for the following one
int e_enqueue_one_step(const ETeam& T, std::vector<int>& cur, std::vector<int>& next) {
    lsc_euler_ctx* L = T.lead();
    int rc = e_reduce_max_speed(T);  // ComputeTimeStep (S:601): the reduction itself ran inside the last y sweep
    if (rc) return rc;
    for (int r = 0; r < T.n; ++r) {
        lsc_euler_ctx* c = T.m[r];
        LSC_CUDA_TRY(cudaSetDevice(c->cfg.device));
        lsc::launch_euler_prepare(c->d_ctl, c->d_cmd, c->d_status, c->d_log, c->stream);
        lsc_count_launches(1);
        LSC_CUDA_TRY(lsc::publish_euler_step_scalars(c->d_ctl, c->slot, c->stream));
        rc = e_launch(c, cur[r], next[r]);
        if (rc) return rc;
        std::swap(cur[r], next[r]);
    }
    if (T.n > 1) return eg_halo(T, cur.data());  // applyBoundaryConditions(u) after the average (S:616)
    if (!e_overlap_active(L)) return e_fill_x_ghosts(L, cur[0]);
    return LSC_OK;
}

bool e_graph_eligible(const ETeam& T) {
    const lsc_euler_ctx* c = T.lead();
    return T.n == 1 && c->cfg.nranks <= 1 && !c->profiling && !c->graph_disabled;
}

// E_GRAPH_STEPS steps of a plain context as one CUDA graph (see lsc_capi.cu: graph_for)
int e_graph_for(const ETeam& T, int cur0, cudaGraphExec_t* out) {
    lsc_euler_ctx* c = T.lead();
    const int key = cur0 | (c->overlap << 4) | (c->d_table ? 256 : 0) | ((c->d_table && c->table_monotone && c->warm_search) ? 512 : 0);
    for (auto& e : c->graphs)
        if (e.key == key) {
            *out = e.exec;
            return LSC_OK;
        }
    *out = nullptr;
    std::vector<int> cur{cur0}, next{cur0 == c->cur ? c->next : c->cur};
    const unsigned long long before = lsc_kernel_launch_count();
    if (cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
        cudaGetLastError();
        c->graph_disabled = true;
        return LSC_OK;
    }
    int rc = LSC_OK;
    for (int q = 0; q < E_GRAPH_STEPS && rc == LSC_OK; ++q) rc = e_enqueue_one_step(T, cur, next);
    cudaGraph_t graph = nullptr;
    cudaError_t e = cudaStreamEndCapture(c->stream, &graph);
    const unsigned long long captured = lsc_kernel_launch_count() - before;
    lsc_uncount_launches(captured);  // nothing ran yet: launches are counted per replay
    cudaGraphExec_t exec = nullptr;
    if (rc == LSC_OK && e == cudaSuccess && graph) e = cudaGraphInstantiate(&exec, graph, 0);
    if (graph) cudaGraphDestroy(graph);
    if (rc != LSC_OK || e != cudaSuccess || !exec) {
        cudaGetLastError();
        c->graph_disabled = true;
        return LSC_OK;
    }
    c->graphs.push_back({key, exec, captured});
    *out = exec;
    return LSC_OK;
}

int e_enqueue_steps(const ETeam& T, long count) {
    lsc_euler_ctx* L = T.lead();
    std::vector<int> cur(T.n), next(T.n);
    for (int r = 0; r < T.n; ++r) {
        cur[r] = T.m[r]->cur;
        next[r] = T.m[r]->next;
    }
    if (!L->ghosts_current) {
        // applyBoundaryConditions(u) at the head of the first pass (S:604) for a decomposed grid: the uploaded array's
        // ghost columns at the ends of the domain hold whatever the caller had there (the initial data's own values,
        // for instance), not boundary images.  A plain context did this at the end of the upload.
        int rc = LSC_OK;
        if (T.n > 1) rc = eg_halo(T, cur.data());
        else if (L->cfg.nranks > 1) rc = e_fill_x_ghosts(L, cur[0]);
        if (rc) return rc;
        for (int r = 0; r < T.n; ++r) T.m[r]->ghosts_current = true;
    }
    long e = 0;
    while (e < count) {
        if (count - e >= E_GRAPH_STEPS && e_graph_eligible(T)) {
            cudaGraphExec_t exec = nullptr;
            int rc = e_graph_for(T, cur[0], &exec);
            if (rc) return rc;
            if (exec) {
                LSC_CUDA_TRY(cudaGraphLaunch(exec, L->stream));
                for (auto& g : L->graphs)
                    if (g.exec == exec) lsc_count_launches(g.launches);
                e += E_GRAPH_STEPS;
                continue;
            }
        }
        int rc = e_enqueue_one_step(T, cur, next);
        if (rc) return rc;
        e += 1;
    }
    return LSC_OK;
}

struct EBatch {
    long done = 0;
    int halt = lsc::HALT_NONE;
};

// one batch: command block -> device, `count` steps enqueued, ONE synchronisation, host mirrors brought up to date
int e_run_batch(const ETeam& T, long count, double t_stop, long long step_limit, bool keep_log, EBatch* res) {
    lsc_euler_ctx* L = T.lead();
    for (int r = 0; r < T.n; ++r) {
        lsc_euler_ctx* c = T.m[r];
        LSC_CUDA_TRY(cudaSetDevice(c->cfg.device));
        c->batch += 1;
        c->batches += 1;
        lsc::StepCmd& cmd = c->h->cmd;
        cmd.batch = c->batch;
        cmd.step_limit = step_limit;
        cmd.t_stop = t_stop;
        cmd.dump_counter = -1;
        cmd.keep_log = (keep_log && r == 0) ? 1 : 0;
        LSC_CUDA_TRY(cudaMemcpyAsync(c->d_cmd, &cmd, sizeof cmd, cudaMemcpyHostToDevice, c->stream));
    }
    int rc = e_enqueue_steps(T, count);
    for (int r = 0; r < T.n; ++r) {
        lsc_euler_ctx* c = T.m[r];
        cudaSetDevice(c->cfg.device);
        cudaMemcpyAsync(&c->h->ctl, c->d_ctl, sizeof(lsc::EulerStepCtl), cudaMemcpyDeviceToHost, c->stream);
        cudaMemcpyAsync(c->h->status, c->d_status, 32, cudaMemcpyDeviceToHost, c->stream);
    }
    for (int r = 0; r < T.n; ++r) {
        cudaSetDevice(T.m[r]->cfg.device);
        cudaError_t e = cudaStreamSynchronize(T.m[r]->stream);
        if (e != cudaSuccess && rc == LSC_OK)
            rc = lsc_fail(LSC_ECUDA, "step batch failed on slab %d: %s", r, cudaGetErrorString(e));
    }
    if (rc) return rc;
    const long before = L->step;
    for (int r = 0; r < T.n; ++r) {
        lsc_euler_ctx* c = T.m[r];
        const lsc::EulerStepCtl& k = c->h->ctl;
        if (k.step != L->h->ctl.step || k.halt != L->h->ctl.halt)
            return lsc_fail(LSC_ESTATE, "slab %d is at step %lld (halt %d), the leader at %lld (halt %d)", r, k.step,
                            k.halt, L->h->ctl.step, L->h->ctl.halt);
        if (((long)k.step - c->step) & 1) std::swap(c->cur, c->next);
        c->step = (long)k.step;
        c->t = k.t;
        c->dt = k.dt;
        c->ieee_redos = (long)(k.redo_total + c->h->status[1] + c->h->status[2] + c->h->status[3]);
    }
    res->done = L->step - before;
    res->halt = L->h->ctl.halt;
    if (res->halt == lsc::HALT_NAN)
        return lsc_fail(LSC_ENAN, "max signal speed is %g at step %ld", L->h->ctl.max_speed, L->step);
    if (keep_log && L->h->ctl.log_count > 0) {
        LSC_CUDA_TRY(cudaSetDevice(L->cfg.device));
        LSC_CUDA_TRY(cudaMemcpyAsync(L->h->log, L->d_log, sizeof(double) * 3 * (size_t)L->h->ctl.log_count,
                                     cudaMemcpyDeviceToHost, L->stream));
        LSC_CUDA_TRY(cudaStreamSynchronize(L->stream));
    }
    return LSC_OK;
}

int e_team_step(const ETeam& T, long nsteps) {
    lsc_euler_ctx* L = T.lead();
    const long long target = (long long)L->step + nsteps;
    while (L->step < target) {
        EBatch res;
        const long count = (long)std::min<long long>(target - L->step, E_BATCH_MAX);
        int rc = e_run_batch(T, count, std::numeric_limits<double>::infinity(), target, false, &res);
        if (rc) return rc;
    }
    return LSC_OK;
}

// do { step } while (t < t_stop)  -- S:600-617 / G:627-646 (the body runs at least once)
int e_team_run_until(const ETeam& T, double t_stop, lsc_euler_log_cb log_cb, void* user, long* steps_out) {
    lsc_euler_ctx* L = T.lead();
    long n = 0;
    bool again = true;
    while (again) {
        long count = 1;
        if (L->dt > 0) {
            const double est = (t_stop - L->t) / L->dt;
            count = est > (double)E_BATCH_MAX ? E_BATCH_MAX : (est < 1.0 ? 1 : (long)est + 2);
        }
        if (log_cb) count = std::min<long>(count, lsc::STEP_LOG_CAP);
        EBatch res;
        int rc = e_run_batch(T, count, t_stop, LLONG_MAX, log_cb != nullptr, &res);
        if (rc) return rc;
        n += res.done;
        if (log_cb)
            for (long q = 0; q < res.done; ++q) {  // G:631
                const double* rec = L->h->log + 3 * q;
                if (log_cb(L->step - res.done + q + 1, rec[0], rec[1], user))
                    return lsc_fail(LSC_ECALLBACK, "log callback stopped the loop at step %ld", L->step - res.done + q + 1);
            }
        again = L->t < t_stop;  // S:617 / G:646
    }
    if (steps_out) *steps_out = n;
    return LSC_OK;
}

int e_team_step_timed(const ETeam& T, long nsteps, float* ms) {
    lsc_euler_ctx* L = T.lead();
    LSC_CUDA_TRY(cudaSetDevice(L->cfg.device));
    cudaEvent_t e0, e1;
    LSC_CUDA_TRY(cudaEventCreate(&e0));
    LSC_CUDA_TRY(cudaEventCreate(&e1));
    for (int r = 0; r < T.n; ++r) {
        cudaSetDevice(T.m[r]->cfg.device);
        cudaStreamSynchronize(T.m[r]->stream);
    }
    cudaSetDevice(L->cfg.device);
    cudaEventRecord(e0, L->stream);
    int rc = e_team_step(T, nsteps);
    if (rc == LSC_OK) {
        cudaSetDevice(L->cfg.device);
        cudaError_t e = cudaEventRecord(e1, L->stream);
        if (e == cudaSuccess) e = cudaEventSynchronize(e1);
        if (e == cudaSuccess) e = cudaEventElapsedTime(ms, e0, e1);
        if (e != cudaSuccess) rc = lsc_fail(LSC_ECUDA, "timing events: %s", cudaGetErrorString(e));
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    return rc;
}

int e_check_plain(lsc_euler_ctx* c, const char* what) {
    if (!c) return lsc_fail(LSC_EINVAL, "null argument");
    if (!c->uploaded) return lsc_fail(LSC_ESTATE, "%s before lsc_euler_upload", what);
    if (c->cfg.nranks > 1 && !c->comm)
        return lsc_fail(LSC_ESTATE, "nranks > 1: call lsc_euler_comm_init (one process per GPU) or use the "
                                    "lsc_euler_group_* calls (slabs owned by one process)");
    return LSC_OK;
}

int eg_check(lsc_euler_ctx** g, int n) {
    if (!g || n < 1) return lsc_fail(LSC_EINVAL, "empty group");
    for (int r = 0; r < n; ++r) {
        if (!g[r]) return lsc_fail(LSC_EINVAL, "null context in group");
        if (g[r]->cfg.nranks != n || g[r]->cfg.rank != r)
            return lsc_fail(LSC_EINVAL, "group member %d has rank %d of %d", r, g[r]->cfg.rank, g[r]->cfg.nranks);
        if (!g[r]->uploaded) return lsc_fail(LSC_ESTATE, "group member %d has no state uploaded", r);
        if (g[r]->comm) return lsc_fail(LSC_ESTATE, "group member %d is attached to NCCL", r);
        if (g[r]->step != g[0]->step) return lsc_fail(LSC_ESTATE, "group members are at different steps");
    }
    for (int r = 0; r < n; ++r) {  // direct peer copies between the members' GPUs
        LSC_CUDA_TRY(cudaSetDevice(g[r]->cfg.device));
        for (int q = 0; q < n; ++q) {
            const int a = g[r]->cfg.device, b = g[q]->cfg.device;
            int can = 0;
            if (a != b && cudaDeviceCanAccessPeer(&can, a, b) == cudaSuccess && can) cudaDeviceEnablePeerAccess(b, 0);
            cudaGetLastError();
        }
    }
    return LSC_OK;
}

struct EulerGroupScope {  // in_group for the duration of a group call, reset on every exit path
    lsc_euler_ctx** g;
    int n;
    EulerGroupScope(lsc_euler_ctx** g_, int n_) : g(g_), n(n_) {
        for (int r = 0; r < n; ++r) g[r]->in_group = n > 1;
    }
    ~EulerGroupScope() {
        for (int r = 0; r < n; ++r) g[r]->in_group = false;
    }
};

}  // namespace

extern "C" {

int lsc_euler_local_nx(const lsc_euler_ctx* c) { return c ? c->nxl : -1; }
int lsc_euler_local_i0(const lsc_euler_ctx* c) { return c ? c->i0 : -1; }

int lsc_euler_comm_init(lsc_euler_ctx* c, const void* id128) {
    if (!c || !id128) return lsc_fail(LSC_EINVAL, "null argument");
    if (c->cfg.nranks <= 1) return LSC_OK;
    if (c->comm) return lsc_fail(LSC_ESTATE, "communicator already initialised");
    LSC_CUDA_TRY(cudaSetDevice(c->cfg.device));
    c->comm = lsc::comm_create(id128, c->cfg.rank, c->cfg.nranks, c->g, 4);
    if (!c->comm) return lsc_fail(LSC_ECOMM, "%s", lsc::comm_error());
    return LSC_OK;
}

int lsc_euler_group_step(lsc_euler_ctx** g, int n, long nsteps) {
    int rc = eg_check(g, n);
    if (rc) return rc;
    EulerGroupScope scope(g, n);
    ETeam T{g, n};
    return e_team_step(T, nsteps);
}

int lsc_euler_group_run_until(lsc_euler_ctx** g, int n, double t_stop, lsc_euler_log_cb log_cb, void* user,
                              long* steps_out) {
    int rc = eg_check(g, n);
    if (rc) return rc;
    EulerGroupScope scope(g, n);
    ETeam T{g, n};
    return e_team_run_until(T, t_stop, log_cb, user, steps_out);
}

int lsc_euler_create(const lsc_euler_config* cfg, lsc_euler_ctx** out) {
    if (!cfg || !out) return lsc_fail(LSC_EINVAL, "lsc_euler_create: null argument");
    *out = nullptr;
    if (cfg->nx < 4 || cfg->ny < 4) return lsc_fail(LSC_EINVAL, "nx, ny must be >= 4 (got %d x %d)", cfg->nx, cfg->ny);
    if (!(cfg->x1 > cfg->x0) || !(cfg->y1 > cfg->y0)) return lsc_fail(LSC_EINVAL, "empty domain");
    if (!(cfg->gamma > 1.0) || !(cfg->cfl > 0.0)) return lsc_fail(LSC_EINVAL, "gamma must be > 1 and cfl > 0");
    if (cfg->bc != LSC_BC_PERIODIC && cfg->bc != LSC_BC_TRANSMISSIVE_E)
        return lsc_fail(LSC_EINVAL, "bc must be LSC_BC_PERIODIC (S:198-226) or LSC_BC_TRANSMISSIVE_E (G:199-215)");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return lsc_fail(LSC_ECUDA, "no CUDA device: this library has no CPU path");
    }
    if (cfg->device < 0 || cfg->device >= ndev) return lsc_fail(LSC_EINVAL, "device %d out of range", cfg->device);
    LSC_CUDA_TRY(cudaSetDevice(cfg->device));
    if (cfg->nranks > 1 && (cfg->rank < 0 || cfg->rank >= cfg->nranks)) return lsc_fail(LSC_EINVAL, "bad rank/nranks");
    lsc_euler_ctx* c = new (std::nothrow) lsc_euler_ctx();
    if (!c) return lsc_fail(LSC_ENOMEM, "host allocation failed");
    c->cfg = *cfg;
    if (c->cfg.nranks < 1) {
        c->cfg.nranks = 1;
        c->cfg.rank = 0;
    }
    int b0, b1;
    lsc::euler_slab_range(c->cfg, b0, b1);
    c->i0 = b0;
    c->nxl = b1 - b0;
    if (c->nxl < 4) {
        delete c;
        return lsc_fail(LSC_EINVAL, "slab of rank %d has %d columns (need >= 4)", cfg->rank, b1 - b0);
    }
    c->dx = (cfg->x1 - cfg->x0) / cfg->nx;  // S:546-547
    c->dy = (cfg->y1 - cfg->y0) / cfg->ny;
    c->g.nxl = c->nxl;
    c->g.ny = cfg->ny;
    c->g.pitch = ((cfg->ny + 4 + lsc::JOFF) + lsc::JBLK - 1) / lsc::JBLK * lsc::JBLK;
    c->g.plane = (size_t)(c->nxl + 4) * (size_t)c->g.pitch;
    const size_t state_bytes = c->g.plane * 4 * sizeof(double);
    const size_t col_bytes = (size_t)(cfg->ny + 4) * 4 * sizeof(double);
    long long cols = (long long)((256ull << 20) / col_bytes);
    if (cols < 2) cols = 2;  // two halves (lsc_euler_upload / lsc_euler_download)
    if (cols > c->nxl + 4) cols = c->nxl + 4;
    c->stage_cols = (int)cols;
    bool ok = true;
    for (int b = 0; b < 3 && ok; ++b) {
        ok = lsc_dev_malloc(&c->state[b], state_bytes) == cudaSuccess;
        if (ok) cudaMemset(c->state[b], 0, state_bytes);
    }
    c->slot = lsc_slot_acquire(cfg->device, 1);
    if (c->slot < 0) {
        lsc_euler_destroy(c);
        return lsc_fail(LSC_ENOMEM, "more than %d live Euler contexts on device %d", lsc::STEP_SLOTS, cfg->device);
    }
    ok = ok && lsc_dev_malloc(&c->d_stage, col_bytes * (size_t)cols) == cudaSuccess &&
         lsc_dev_malloc(&c->d_status, 32) == cudaSuccess && cudaMemset(c->d_status, 0, 32) == cudaSuccess &&
         lsc_dev_malloc(&c->d_ctl, sizeof(lsc::EulerStepCtl)) == cudaSuccess &&
         lsc_dev_malloc(&c->d_cmd, sizeof(lsc::StepCmd)) == cudaSuccess &&
         cudaMemset(c->d_cmd, 0, sizeof(lsc::StepCmd)) == cudaSuccess &&
         lsc_dev_malloc(&c->d_log, sizeof(double) * 3 * lsc::STEP_LOG_CAP) == cudaSuccess &&
         lsc_dev_malloc(&c->d_gather, sizeof(unsigned long long) * (size_t)(c->cfg.nranks + 1)) == cudaSuccess &&
         cudaHostAlloc(&c->h, sizeof(EulerHostBlock), cudaHostAllocPortable) == cudaSuccess &&
         cudaEventCreateWithFlags(&c->red_event, cudaEventDisableTiming) == cudaSuccess &&
         cudaEventCreateWithFlags(&c->max_event, cudaEventDisableTiming) == cudaSuccess &&
         lsc_dev_malloc(&c->d_fix, sizeof(unsigned) * (size_t)(lsc::euler_fix_list_entries(c->g) +
                                                           lsc::euler_edge_list_entries(c->g))) == cudaSuccess &&
         cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) == cudaSuccess &&
         cudaEventCreateWithFlags(&c->halo_event, cudaEventDisableTiming) == cudaSuccess &&
         cudaEventCreateWithFlags(&c->edges_done, cudaEventDisableTiming) == cudaSuccess &&
         cudaEventCreateWithFlags(&c->halo_done, cudaEventDisableTiming) == cudaSuccess &&
         lsc_create_comm_stream(&c->comm_stream) == cudaSuccess;
    if (!ok) {
        cudaGetLastError();
        lsc_euler_destroy(c);
        return lsc_fail(LSC_ENOMEM, "device allocation failed (%zu bytes per state)", state_bytes);
    }
    std::memset(c->h, 0, sizeof(EulerHostBlock));
    if (const char* e = std::getenv("LSC_NO_GRAPH")) c->graph_disabled = (*e && *e != '0');
    *out = c;
    return LSC_OK;
}

int lsc_euler_destroy(lsc_euler_ctx* c) {
    if (!c) return LSC_OK;
    cudaSetDevice(c->cfg.device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->comm_stream) cudaStreamSynchronize(c->comm_stream);
    for (int b = 0; b < 3; ++b)
        if (c->state[b]) lsc_dev_free(c->state[b]);
    if (c->copy_stream) {
        cudaStreamSynchronize(c->copy_stream);
        cudaStreamDestroy(c->copy_stream);
    }
    for (int b = 0; b < 2; ++b) {
        if (c->stage_filled[b]) cudaEventDestroy(c->stage_filled[b]);
        if (c->stage_free[b]) cudaEventDestroy(c->stage_free[b]);
    }
    if (c->d_stage) lsc_dev_free(c->d_stage);
    if (c->d_status) lsc_dev_free(c->d_status);
    if (c->d_fix) lsc_dev_free(c->d_fix);
    if (c->d_ctl) lsc_dev_free(c->d_ctl);
    if (c->d_cmd) lsc_dev_free(c->d_cmd);
    if (c->d_log) lsc_dev_free(c->d_log);
    if (c->d_gather) lsc_dev_free(c->d_gather);
    if (c->h) cudaFreeHost(c->h);
    if (c->red_event) cudaEventDestroy(c->red_event);
    if (c->max_event) cudaEventDestroy(c->max_event);
    for (auto& g : c->graphs) cudaGraphExecDestroy(g.exec);
    lsc_slot_release(c->cfg.device, 1, c->slot);
    if (c->d_table) lsc_dev_free(c->d_table);
    if (c->d_hints) lsc_dev_free(c->d_hints);
    if (c->comm) lsc::comm_destroy(c->comm);
    if (c->halo_event) cudaEventDestroy(c->halo_event);
    if (c->edges_done) cudaEventDestroy(c->edges_done);
    if (c->halo_done) cudaEventDestroy(c->halo_done);
    if (c->comm_stream) cudaStreamDestroy(c->comm_stream);
    for (auto e : c->prof_events) cudaEventDestroy(e);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
    return LSC_OK;
}

int lsc_euler_initial_condition(const lsc_euler_config* cfg, int problem, double* aos) {
    if (!cfg || !aos) return lsc_fail(LSC_EINVAL, "null argument");
    if (problem < LSC_EULER_IC_ORSZAG_TANG || problem > LSC_EULER_IC_CYLINDER)
        return lsc_fail(LSC_EINVAL, "unknown problem %d (the table IC goes through lsc_euler_tabeos_initial_condition)",
                        problem);
    lsc::euler_initial_condition(*cfg, problem, aos);
    return LSC_OK;
}

int lsc_eos_table_read(const char* path, double* densities, double* pressures, double* sound_speeds,
                       double* energies) {
    if (!path || !densities || !pressures || !sound_speeds || !energies) return lsc_fail(LSC_EINVAL, "null argument");
    const int bad = lsc::eos_table_read(path, densities, pressures, sound_speeds, energies);
    if (bad < 0) return lsc_fail(LSC_EINVAL, "cannot open %s", path);
    if (bad > 0) return lsc_fail(LSC_EINVAL, "%s: line %d does not hold the 500 values the table layout needs", path, bad);
    return LSC_OK;
}

int lsc_euler_set_eos_table(lsc_euler_ctx* c, const double* densities, const double* pressures,
                            const double* sound_speeds, const double* energies) {
    if (!c) return lsc_fail(LSC_EINVAL, "null argument");
    LSC_CUDA_TRY(cudaSetDevice(c->cfg.device));
    for (auto& g : c->graphs) cudaGraphExecDestroy(g.exec);  // captured steps hold the old table pointers
    c->graphs.clear();
    if (!densities) {  // back to the ideal gas of S / G
        if (c->d_table) lsc_dev_free(c->d_table);
        c->d_table = nullptr;
        c->h_table.clear();
        return LSC_OK;
    }
    if (!pressures || !sound_speeds || !energies) return lsc_fail(LSC_EINVAL, "null table");
    for (size_t k = 1; k < TAB_N; ++k)
        if (!(densities[k] > densities[k - 1]) || !(pressures[k] > pressures[k - 1]))
            return lsc_fail(LSC_EINVAL, "table axes must be strictly ascending (entry %zu)", k);
    try {
        c->h_table.resize(TAB_DOUBLES);
    } catch (...) {
        return lsc_fail(LSC_ENOMEM, "host allocation failed");
    }
    double* t = c->h_table.data();
    std::memcpy(t, densities, TAB_N * sizeof(double));
    std::memcpy(t + TAB_N, pressures, TAB_N * sizeof(double));
    std::memcpy(t + 2 * TAB_N, sound_speeds, TAB_N * TAB_N * sizeof(double));
    std::memcpy(t + 2 * TAB_N + TAB_N * TAB_N, energies, TAB_N * TAB_N * sizeof(double));
    if (!c->d_table && lsc_dev_malloc(&c->d_table, TAB_DOUBLES * sizeof(double)) != cudaSuccess) {
        cudaGetLastError();
        c->d_table = nullptr;
        return lsc_fail(LSC_ENOMEM, "device allocation of the EoS table failed");
    }
    LSC_CUDA_TRY(cudaMemcpy(c->d_table, t, TAB_DOUBLES * sizeof(double), cudaMemcpyHostToDevice));
    c->table_monotone = lsc::eos_table_is_monotone(energies);
    if (!c->d_hints) {  // any in-range bracket is a valid starting point: 0 will do
        const size_t bytes = 3 * sizeof(int) * (size_t)lsc::euler_y_block_count(c->g);
        if (lsc_dev_malloc(&c->d_hints, bytes) != cudaSuccess) {
            cudaGetLastError();
            c->d_hints = nullptr;
            return lsc_fail(LSC_ENOMEM, "device allocation of the search hints failed");
        }
        LSC_CUDA_TRY(cudaMemset(c->d_hints, 0, bytes));
    }
    return LSC_OK;
}

int lsc_euler_set_eos_search(lsc_euler_ctx* c, int warm_start) {
    if (!c) return lsc_fail(LSC_EINVAL, "null argument");
    c->warm_search = (warm_start != 0);
    return LSC_OK;
}

int lsc_euler_eos_search_is_warm(const lsc_euler_ctx* c) {
    if (!c) return 0;
    return (c->d_table && c->table_monotone && c->warm_search) ? 1 : 0;
}

int lsc_euler_tabeos_initial_condition(const lsc_euler_config* cfg, const double* densities, const double* pressures,
                                       const double* energies, double* aos) {
    if (!cfg || !densities || !pressures || !energies || !aos) return lsc_fail(LSC_EINVAL, "null argument");
    lsc::tabeos_initial_condition(*cfg, densities, pressures, energies, aos);
    return LSC_OK;
}

int lsc_euler_tabeos_write_dat(const lsc_euler_config* cfg, const double* aos, const char* path,
                               const double* densities, const double* pressures, const double* energies) {
    if (!cfg || !aos || !path || !densities || !pressures || !energies) return lsc_fail(LSC_EINVAL, "null argument");
    if (lsc::tabeos_write_dat(*cfg, aos, path, densities, pressures, energies))
        return lsc_fail(LSC_EINVAL, "cannot write %s", path);
    return LSC_OK;
}

static int e_transfer_init(lsc_euler_ctx* c) {
    if (c->copy_stream) return LSC_OK;
    LSC_CUDA_TRY(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
    for (int b = 0; b < 2; ++b) {
        LSC_CUDA_TRY(cudaEventCreateWithFlags(&c->stage_filled[b], cudaEventDisableTiming));
        LSC_CUDA_TRY(cudaEventCreateWithFlags(&c->stage_free[b], cudaEventDisableTiming));
    }
    return LSC_OK;
}

int lsc_euler_upload(lsc_euler_ctx* c, const double* aos) {
    if (!c || !aos) return lsc_fail(LSC_EINVAL, "null argument");
    LSC_CUDA_TRY(cudaSetDevice(c->cfg.device));
    if (int rc = e_transfer_init(c)) return rc;
    const size_t col = (size_t)(c->cfg.ny + 4) * 4;
    const int ncol = c->nxl + 4;
    const int half = std::max(1, c->stage_cols / 2);
    int chunk = 0;
    for (int il = 0; il < ncol; il += half, ++chunk) {
        const int ni = std::min(half, ncol - il);
        const int b = chunk & 1;
        double* buf = c->d_stage + (size_t)b * half * col;
        if (chunk >= 2) LSC_CUDA_TRY(cudaStreamWaitEvent(c->copy_stream, c->stage_free[b], 0));
        LSC_CUDA_TRY(cudaMemcpyAsync(buf, aos + col * il, col * ni * sizeof(double), cudaMemcpyHostToDevice,
                                     c->copy_stream));
        LSC_CUDA_TRY(cudaEventRecord(c->stage_filled[b], c->copy_stream));
        LSC_CUDA_TRY(cudaStreamWaitEvent(c->stream, c->stage_filled[b], 0));
        lsc::launch_euler_aos_to_soa(c->g, buf, il, ni, c->state[c->cur], c->stream);
        lsc_count_launches(1);
        LSC_CUDA_TRY(cudaGetLastError());
        LSC_CUDA_TRY(cudaEventRecord(c->stage_free[b], c->stream));
    }
    LSC_CUDA_TRY(cudaStreamSynchronize(c->stream));
    // ComputeTimeStep of the first step sees the array as given, over i,j in [2, n+2] (S:173-174); then the
    // loop's own boundary fill (S:604) -- on the device only the x ghost columns are ever read
    LSC_CUDA_TRY(cudaMemsetAsync(c->d_status, 0, 32, c->stream));
    // (a slab's share of that range: its own interior columns, plus global column nx+2 on the last rank)
    const int i_hi = c->nxl + ((c->cfg.rank == c->cfg.nranks - 1) ? 2 : 1);
    lsc::EulerParams ep{};
    fill_eos(c, ep);
    lsc::launch_euler_max_speed(c->g, ep, c->state[c->cur], 2, i_hi, 2, c->cfg.ny + 2, c->d_status, c->stream);
    lsc_count_launches(1);
    LSC_CUDA_TRY(cudaGetLastError());
    c->ghosts_current = false;
    if (c->cfg.nranks <= 1) {
        int rc = e_fill_x_ghosts(c, c->cur);
        if (rc) return rc;
        c->ghosts_current = true;
    }  // slabs: the first pass of the loop exchanges the halo columns / forms the images (e_enqueue_steps)
    // fresh loop state: t = 0, step = 0, the run's constants (S:537, S:166-169, S:546-547)
    lsc::EulerStepCtl& k = c->h->ctl;
    std::memset(&k, 0, sizeof k);
    k.cfl = c->cfg.cfl;
    k.dspace = c->dx;  // S:166-169
    if (c->dy < c->dx) k.dspace = c->dy;
    k.dx = c->dx;
    k.dy = c->dy;
    k.batch_seen = c->batch;
    LSC_CUDA_TRY(cudaMemcpyAsync(c->d_ctl, &k, sizeof k, cudaMemcpyHostToDevice, c->stream));
    LSC_CUDA_TRY(cudaStreamSynchronize(c->stream));
    c->t = c->dt = 0;
    c->step = 0;
    c->ieee_redos = 0;
    c->uploaded = true;
    return LSC_OK;
}

int lsc_euler_init_problem(lsc_euler_ctx* c, int problem) {
    if (!c) return lsc_fail(LSC_EINVAL, "null argument");
    std::vector<double> host;
    try {
        host.resize((size_t)(c->nxl + 4) * (c->cfg.ny + 4) * 4);
    } catch (...) {
        return lsc_fail(LSC_ENOMEM, "host allocation failed");
    }
    if (problem == LSC_EULER_IC_TABEOS) {
        if (c->h_table.empty()) return lsc_fail(LSC_ESTATE, "LSC_EULER_IC_TABEOS needs lsc_euler_set_eos_table first");
        const double* t = c->h_table.data();
        lsc::tabeos_initial_condition(c->cfg, t, t + TAB_N, t + 2 * TAB_N + TAB_N * TAB_N, host.data());
        return lsc_euler_upload(c, host.data());
    }
    int rc = lsc_euler_initial_condition(&c->cfg, problem, host.data());
    if (rc) return rc;
    return lsc_euler_upload(c, host.data());
}

int lsc_euler_step(lsc_euler_ctx* c, long nsteps) {
    int rc = e_check_plain(c, "lsc_euler_step");
    if (rc) return rc;
    ETeam T{&c, 1};
    return e_team_step(T, nsteps);
}

int lsc_euler_step_timed(lsc_euler_ctx* c, long nsteps, float* ms) {
    if (!ms) return lsc_fail(LSC_EINVAL, "null argument");
    int rc = e_check_plain(c, "lsc_euler_step_timed");
    if (rc) return rc;
    ETeam T{&c, 1};
    return e_team_step_timed(T, nsteps, ms);
}

int lsc_euler_run_until(lsc_euler_ctx* c, double t_stop, lsc_euler_log_cb log_cb, void* user, long* steps_out) {
    int rc = e_check_plain(c, "lsc_euler_run_until");
    if (rc) return rc;
    ETeam T{&c, 1};
    return e_team_run_until(T, t_stop, log_cb, user, steps_out);
}

int lsc_euler_loop_stats(const lsc_euler_ctx* c, long* host_syncs) {
    if (!c) return lsc_fail(LSC_EINVAL, "null argument");
    if (host_syncs) *host_syncs = c->batches;
    return LSC_OK;
}

int lsc_euler_get_scalars(const lsc_euler_ctx* c, double* t, double* dt, long* step) {
    if (!c) return lsc_fail(LSC_EINVAL, "null argument");
    if (t) *t = c->t;
    if (dt) *dt = c->dt;
    if (step) *step = c->step;
    return LSC_OK;
}

int lsc_euler_download(lsc_euler_ctx* c, double* aos) {
    if (!c || !aos) return lsc_fail(LSC_EINVAL, "null argument");
    if (!c->uploaded) return lsc_fail(LSC_ESTATE, "lsc_euler_download before lsc_euler_upload");
    LSC_CUDA_TRY(cudaSetDevice(c->cfg.device));
    if (int rc = e_transfer_init(c)) return rc;
    const size_t col = (size_t)(c->cfg.ny + 4) * 4;
    const int ncol = c->nxl + 4;
    const int half = std::max(1, c->stage_cols / 2);
    int chunk = 0;
    for (int il = 0; il < ncol; il += half, ++chunk) {
        const int ni = std::min(half, ncol - il);
        const int b = chunk & 1;
        double* buf = c->d_stage + (size_t)b * half * col;
        if (chunk >= 2) LSC_CUDA_TRY(cudaStreamWaitEvent(c->stream, c->stage_free[b], 0));
        lsc::launch_euler_soa_to_aos(c->g, c->state[c->cur], il, ni, buf, c->stream);
        lsc_count_launches(1);
        LSC_CUDA_TRY(cudaGetLastError());
        LSC_CUDA_TRY(cudaEventRecord(c->stage_filled[b], c->stream));
        LSC_CUDA_TRY(cudaStreamWaitEvent(c->copy_stream, c->stage_filled[b], 0));
        LSC_CUDA_TRY(cudaMemcpyAsync(aos + col * il, buf, col * ni * sizeof(double), cudaMemcpyDeviceToHost,
                                     c->copy_stream));
        LSC_CUDA_TRY(cudaEventRecord(c->stage_free[b], c->copy_stream));
    }
    LSC_CUDA_TRY(cudaStreamSynchronize(c->copy_stream));
    LSC_CUDA_TRY(cudaStreamSynchronize(c->stream));
    // y ghost rows are never stored on the device: rebuild the reference's ghost ring (S:616) on the host
    if (c->step > 0) lsc::euler_fill_ghost_ring(c->cfg, aos);
    return LSC_OK;
}

int lsc_euler_state_hash(lsc_euler_ctx* c, unsigned long long* hash) {
    if (!c || !hash) return lsc_fail(LSC_EINVAL, "null argument");
    if (!c->uploaded) return lsc_fail(LSC_ESTATE, "lsc_euler_state_hash before lsc_euler_upload");
    LSC_CUDA_TRY(cudaSetDevice(c->cfg.device));
    unsigned long long* d = nullptr;
    LSC_CUDA_TRY(cudaMalloc(&d, sizeof *d));
    cudaMemsetAsync(d, 0, sizeof *d, c->stream);
    lsc::launch_state_hash(c->g, 4, c->state[c->cur], c->i0, d, c->stream);
    lsc_count_launches(1);
    cudaError_t e = cudaMemcpyAsync(hash, d, sizeof *d, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(d);
    LSC_CUDA_TRY(e);
    return LSC_OK;
}

int lsc_euler_group_download(lsc_euler_ctx** g, int n, double* aos) {
    int rc = eg_check(g, n);
    if (rc) return rc;
    if (!aos) return lsc_fail(LSC_EINVAL, "null argument");
    const size_t col = (size_t)(g[0]->cfg.ny + 4) * 4;
    std::vector<double> slab;
    for (int r = 0; r < n; ++r) {
        lsc_euler_ctx* c = g[r];
        try {
            slab.resize(col * (size_t)(c->nxl + 4));
        } catch (...) {
            return lsc_fail(LSC_ENOMEM, "host allocation failed");
        }
        rc = lsc_euler_download(c, slab.data());
        if (rc) return rc;
        const int lo = (r == 0) ? 0 : 2, hi = (r == n - 1) ? c->nxl + 4 : c->nxl + 2;
        std::memcpy(aos + col * (size_t)(c->i0 + lo), slab.data() + col * lo, col * (size_t)(hi - lo) * sizeof(double));
    }
    if (g[0]->step > 0) {
        lsc_euler_config whole = g[0]->cfg;
        whole.rank = 0;
        whole.nranks = 1;
        lsc::euler_fill_ghost_ring(whole, aos);
    }
    return LSC_OK;
}

int lsc_euler_write_dat(const lsc_euler_config* cfg, const double* aos, const char* path, int blank_line_per_row) {
    if (!cfg || !aos || !path) return lsc_fail(LSC_EINVAL, "null argument");
    if (lsc::euler_write_dat(*cfg, aos, path, blank_line_per_row)) return lsc_fail(LSC_EINVAL, "cannot write %s", path);
    return LSC_OK;
}

long lsc_euler_ieee_fallback_count(const lsc_euler_ctx* c) { return c ? c->ieee_redos : -1; }

int lsc_euler_set_overlap(lsc_euler_ctx* c, int enabled) {
    if (!c) return lsc_fail(LSC_EINVAL, "null context");
    if (enabled < 0 || enabled > 2) return lsc_fail(LSC_EINVAL, "overlap mode %d", enabled);
    c->overlap = enabled;
    return LSC_OK;
}

int lsc_euler_set_profiling(lsc_euler_ctx* c, int enabled) {
    if (!c) return lsc_fail(LSC_EINVAL, "null argument");
    c->profiling = enabled != 0;
    return LSC_OK;
}

int lsc_euler_kernel_times(lsc_euler_ctx* c, double ms[2], long launches[2]) {
    if (!c || !ms || !launches) return lsc_fail(LSC_EINVAL, "null argument");
    LSC_CUDA_TRY(cudaSetDevice(c->cfg.device));
    LSC_CUDA_TRY(cudaStreamSynchronize(c->stream));
    ms[0] = ms[1] = 0;
    launches[0] = launches[1] = 0;
    for (size_t q = 0; q + 3 < c->prof_events.size(); q += 4) {
        float a = 0, b = 0;
        LSC_CUDA_TRY(cudaEventElapsedTime(&a, c->prof_events[q], c->prof_events[q + 1]));
        LSC_CUDA_TRY(cudaEventElapsedTime(&b, c->prof_events[q + 2], c->prof_events[q + 3]));
        ms[0] += a;
        ms[1] += b;
        launches[0] += 1;
        launches[1] += 1;
    }
    for (auto e : c->prof_events) cudaEventDestroy(e);
    c->prof_events.clear();
    return LSC_OK;
}

}  // extern "C"

extern "C" int lsc_euler_kat_cfl_bound(int device, long n, const double* U, double gamma, double run_max,
                                       double* exact_and_bound) {
    if (n <= 0 || !U || !exact_and_bound) return LSC_EINVAL;
    if (cudaSetDevice(device) != cudaSuccess) return LSC_ECUDA;
    double *du = nullptr, *dout = nullptr;
    int rc = LSC_OK;
    if (cudaMalloc(&du, (size_t)n * 32) != cudaSuccess || cudaMalloc(&dout, (size_t)n * 16) != cudaSuccess) {
        rc = LSC_ENOMEM;
    } else {
        cudaMemcpy(du, U, (size_t)n * 32, cudaMemcpyHostToDevice);
        lsc::launch_euler_kat_cfl_bound(n, du, gamma, run_max, dout);
        lsc_count_launches(1);
        if (cudaDeviceSynchronize() != cudaSuccess ||
            cudaMemcpy(exact_and_bound, dout, (size_t)n * 16, cudaMemcpyDeviceToHost) != cudaSuccess) {
            cudaGetLastError();
            rc = LSC_ECUDA;
        }
    }
    cudaFree(du);
    cudaFree(dout);
    return rc;
}

// lsc_comm.cu -- NCCL halo exchange + max-allreduce for the slab decomposition along x.
#include "lsc_comm.h"

#include <cstdio>
#include <cstring>
#include <new>

#include <nccl.h>

#include "euler_kernels.cuh"

void lsc_count_launches(unsigned long long n);

namespace lsc {

static thread_local char g_comm_err[256] = "";
const char* comm_error() { return g_comm_err; }

#define NCCL_TRY(expr, ret)                                                                      \
    do {                                                                                         \
        ncclResult_t r_ = (expr);                                                                \
        if (r_ != ncclSuccess) {                                                                 \
            snprintf(g_comm_err, sizeof g_comm_err, "%s: %s", #expr, ncclGetErrorString(r_));    \
            return ret;                                                                          \
        }                                                                                        \
    } while (0)

struct Comm {
    ncclComm_t nccl = nullptr;
    int rank = 0, nranks = 1, nvar = 0;
};

static_assert(sizeof(ncclUniqueId) == 128, "lsc_comm_unique_id hands out 128 bytes");

int comm_unique_id(void* id128) {
    ncclUniqueId id;
    NCCL_TRY(ncclGetUniqueId(&id), -1);
    std::memcpy(id128, &id, sizeof id);
    return 0;
}

Comm* comm_create(const void* id128, int rank, int nranks, const GridDims&, int nvar) {
    Comm* c = new (std::nothrow) Comm();
    if (!c) return nullptr;
    ncclUniqueId id;
    std::memcpy(&id, id128, sizeof id);
    c->rank = rank;
    c->nranks = nranks;
    c->nvar = nvar;
    ncclResult_t r = ncclCommInitRank(&c->nccl, nranks, id, rank);
    if (r != ncclSuccess) {
        snprintf(g_comm_err, sizeof g_comm_err, "ncclCommInitRank: %s", ncclGetErrorString(r));
        delete c;
        return nullptr;
    }
    return c;
}

void comm_destroy(Comm* c) {
    if (!c) return;
    if (c->nccl) ncclCommDestroy(c->nccl);
    delete c;
}

int comm_exchange_halos(Comm* c, const GridDims& g, int bc_x, double* u, cudaStream_t s) {
    const bool ring = (bc_x == BC_PERIODIC);
    const int left = (c->rank > 0) ? c->rank - 1 : (ring ? c->nranks - 1 : -1);
    const int right = (c->rank < c->nranks - 1) ? c->rank + 1 : (ring ? 0 : -1);
    // A column of the blocked layout (mhd_kernels.cuh) is one contiguous run of nvar*pitch doubles, so the two
    // halo columns a neighbour needs are ONE message per direction.
    const size_t col = column_doubles(g, c->nvar);
    const size_t n = 2 * col;
    NCCL_TRY(ncclGroupStart(), -1);
    // Posting order matters when left == right (two ranks on a periodic ring): NCCL matches the sends and
    // receives of a pair in order, so "my low columns" must meet the peer's "high ghost" receive.
    // A failing call must not leave the group open: remember the first error, close the group, then report.
    ncclResult_t first = ncclSuccess;
    auto post = [&](ncclResult_t r) {
        if (first == ncclSuccess && r != ncclSuccess) first = r;
    };
    if (left >= 0) post(ncclSend(u + 2 * col, n, ncclDouble, left, c->nccl, s));
    if (right >= 0) post(ncclRecv(u + (size_t)(g.nxl + 2) * col, n, ncclDouble, right, c->nccl, s));
    if (right >= 0) post(ncclSend(u + (size_t)g.nxl * col, n, ncclDouble, right, c->nccl, s));
    if (left >= 0) post(ncclRecv(u, n, ncclDouble, left, c->nccl, s));
    post(ncclGroupEnd());
    if (first != ncclSuccess) {
        snprintf(g_comm_err, sizeof g_comm_err, "halo exchange: %s", ncclGetErrorString(first));
        return -1;
    }
    // end ranks of a non-periodic decomposition: the boundary image of that side (MHD: M:617-622, 4 variables:
    // the Euler map G:203-208)
    for (int side = 0; side < 2; ++side) {
        if ((side == 0 ? left : right) >= 0) continue;
        if (c->nvar == NV) launch_x_ghosts_side(g, side, u, s);
        else launch_euler_x_ghosts_side(g, side, u, s);
        lsc_count_launches(1);
    }
    return 0;
}

int comm_allreduce_max_u64(Comm* c, unsigned long long* d_words, int count, cudaStream_t s) {
    NCCL_TRY(ncclAllReduce(d_words, d_words, count, ncclUint64, ncclMax, c->nccl, s), -1);
    return 0;
}

}  // namespace lsc

// lsc_comm.h -- the only cross-rank traffic of the path: two halo
// columns of u^{n+1} per x neighbour per step and one max-allreduce of the signal speed.
// One process per GPU; NCCL over NVLink 5 / NVSwitch.
#pragma once
#include <cuda_runtime.h>

#include "mhd_kernels.cuh"

namespace lsc {

struct Comm;

const char* comm_error();
int comm_unique_id(void* id128);
Comm* comm_create(const void* id128, int rank, int nranks, const GridDims& g, int nvar);
void comm_destroy(Comm* c);
// Fill the ghost columns 0,1 and nxl+2,nxl+3 of `u` (one contiguous run each in the blocked layout): neighbour ranks' edge columns
// (a ring when bc_x is periodic), boundary images on the end ranks otherwise.
int comm_exchange_halos(Comm* c, const GridDims& g, int bc_x, double* u, cudaStream_t s);
// In-place max over ranks of `count` unsigned 64-bit words on the device (bit pattern of a
// non-negative double: same ordering).
int comm_allreduce_max_u64(Comm* c, unsigned long long* d_words, int count, cudaStream_t s);

}  // namespace lsc

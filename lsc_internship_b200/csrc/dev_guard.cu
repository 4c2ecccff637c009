// dev_guard.cu -- guard zones around the library's device buffers (LSC_GUARD=1).
//
// compute-sanitizer is not available on the GPU pool this library is developed on, so out-of-bounds WRITES of the
// kernels are caught the old way: with LSC_GUARD=1 in the environment every device buffer of a context is allocated
// with a 64 KiB zone of 0xA5 bytes in front of it and behind it, and lsc_guard_check() counts the guard bytes that
// no longer hold the pattern -- in the buffers that are alive and in every buffer freed so far (checked when it is
// freed); tests/test_guards_gpu.py runs the parity subset under it.  Without the variable the
// allocator is plain cudaMalloc / cudaFree.
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include <cuda_runtime.h>

#include "lsc_common.h"

namespace {
constexpr size_t GUARD = 64 * 1024;
struct Block {
    char* base;     // what cudaMalloc returned
    size_t bytes;   // user bytes (rounded up to 256)
    int device;
};
std::mutex g_mutex;
std::vector<Block> g_blocks;
unsigned long long g_corrupted_freed = 0;  // guard bytes found overwritten when their buffer was freed

// guard bytes of one block that no longer hold the pattern (the block's device must be current)
unsigned long long corrupted_in(const Block& b) {
    static std::vector<unsigned char> host(GUARD + 256);
    const size_t user = (b.bytes + 255) / 256 * 256;
    const size_t tail = GUARD + (user - b.bytes);
    const char* zones[2] = {b.base, b.base + GUARD + b.bytes};
    const size_t lens[2] = {GUARD, tail};
    unsigned long long bad = 0;
    for (int z = 0; z < 2; ++z) {
        if (cudaMemcpy(host.data(), zones[z], lens[z], cudaMemcpyDeviceToHost) != cudaSuccess) return ~0ull >> 1;
        for (size_t q = 0; q < lens[z]; ++q) bad += (host[q] != 0xA5);
    }
    return bad;
}

bool guard_on() {
    const char* e = std::getenv("LSC_GUARD");
    return e && *e && *e != '0';
}
}  // namespace

cudaError_t lsc_dev_malloc(void** p, size_t bytes) {
    if (!guard_on()) return cudaMalloc(p, bytes);
    const size_t user = (bytes + 255) / 256 * 256;
    char* base = nullptr;
    cudaError_t e = cudaMalloc(&base, user + 2 * GUARD);
    if (e != cudaSuccess) return e;
    e = cudaMemset(base, 0xA5, GUARD);
    if (e == cudaSuccess) e = cudaMemset(base + GUARD + user, 0xA5, GUARD);
    // the slack between `bytes` and the rounded size belongs to the guard as well
    if (e == cudaSuccess && user > bytes) e = cudaMemset(base + GUARD + bytes, 0xA5, user - bytes);
    if (e != cudaSuccess) {
        cudaFree(base);
        return e;
    }
    int dev = 0;
    cudaGetDevice(&dev);
    {
        std::lock_guard<std::mutex> lock(g_mutex);
        g_blocks.push_back({base, bytes, dev});
    }
    *p = base + GUARD;
    return cudaSuccess;
}

cudaError_t lsc_dev_free(void* p) {
    if (!p) return cudaSuccess;
    {
        std::lock_guard<std::mutex> lock(g_mutex);
        for (size_t q = 0; q < g_blocks.size(); ++q)
            if (g_blocks[q].base + GUARD == p) {
                char* base = g_blocks[q].base;
                cudaDeviceSynchronize();
                g_corrupted_freed += corrupted_in(g_blocks[q]);  // the verdict outlives the buffer
                g_blocks.erase(g_blocks.begin() + q);
                return cudaFree(base);
            }
    }
    return cudaFree(p);
}

extern "C" int lsc_guard_check(unsigned long long* corrupted_bytes, unsigned long long* guarded_buffers) {
    if (!corrupted_bytes) return LSC_EINVAL;
    *corrupted_bytes = 0;
    std::lock_guard<std::mutex> lock(g_mutex);
    if (guarded_buffers) *guarded_buffers = g_blocks.size();
    *corrupted_bytes = g_corrupted_freed;
    int restore = 0;
    cudaGetDevice(&restore);
    for (const Block& b : g_blocks) {
        cudaSetDevice(b.device);
        if (cudaDeviceSynchronize() != cudaSuccess) return lsc_fail(LSC_ECUDA, "device %d failed before the guard check", b.device);
        *corrupted_bytes += corrupted_in(b);
    }
    cudaSetDevice(restore);
    return LSC_OK;
}

// Device allocations of the library, with an optional guard mode (environment ORB_B200_GUARD=1): every buffer gets a 4 KB canary
// in front and behind, and orb_debug_guard_check() reads all canaries back.  It is this repository's stand-in for
// `compute-sanitizer --tool memcheck`, which the GPU pool does not allow ("closed on this pool and stays closed"): an
// out-of-bounds WRITE within 4 KB of any library buffer flips canary bytes and is reported with the buffer's size and the
// side it happened on (tests/test_gpu_guard.py runs every kernel family under it).  Off by default: one pointer compare per
// allocation, no extra memory.
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <unordered_map>
#include <vector>
#include "orb_internal.h"

namespace {
const size_t GUARD = 4096;
const unsigned char PATTERN = 0xA5;
struct Block { unsigned char* base; size_t bytes; };
std::mutex g_mu;
std::unordered_map<void*, Block> g_blocks;   // user pointer -> block
bool guard_on() { static const bool on = [] { const char* v = getenv("ORB_B200_GUARD"); return v && atoi(v) != 0; }(); return on; }
}  // namespace

cudaError_t orb_dev_alloc(void** p, size_t bytes) {
    if (!guard_on()) return cudaMalloc(p, bytes);
    unsigned char* base = nullptr;
    const size_t padded = (bytes + 255) & ~(size_t)255;   // the rear canary starts at the next 256-byte boundary
    cudaError_t e = cudaMalloc((void**)&base, padded + 2 * GUARD);
    if (e != cudaSuccess) return e;
    e = cudaMemset(base, PATTERN, padded + 2 * GUARD);
    if (e != cudaSuccess) { cudaFree(base); return e; }
    *p = base + GUARD;
    std::lock_guard<std::mutex> l(g_mu);
    g_blocks[*p] = Block{base, padded};
    return cudaSuccess;
}

cudaError_t orb_dev_free(void* p) {
    if (!p) return cudaSuccess;
    if (guard_on()) {
        std::lock_guard<std::mutex> l(g_mu);
        auto it = g_blocks.find(p);
        if (it != g_blocks.end()) { unsigned char* base = it->second.base; g_blocks.erase(it); return cudaFree(base); }
    }
    return cudaFree(p);
}

extern "C" int orb_debug_guard_check(int32_t* n_buffers, int32_t* n_damaged) {
    if (!n_buffers || !n_damaged) return ORB_ERR_BAD_ARG;
    *n_buffers = 0; *n_damaged = 0;
    if (!guard_on()) return ORB_OK;
    if (cudaDeviceSynchronize() != cudaSuccess) { orb_set_error("guard check: device error pending"); return ORB_ERR_CUDA; }
    std::lock_guard<std::mutex> l(g_mu);
    std::vector<unsigned char> h(GUARD);
    std::string report;
    for (const auto& kv : g_blocks) {
        const Block& b = kv.second;
        ++*n_buffers;
        bool bad = false;
        for (int side = 0; side < 2; ++side) {
            const unsigned char* src = side == 0 ? b.base : b.base + GUARD + b.bytes;
            if (cudaMemcpy(h.data(), src, GUARD, cudaMemcpyDeviceToHost) != cudaSuccess) { orb_set_error("guard check: copy failed"); return ORB_ERR_CUDA; }
            for (size_t i = 0; i < GUARD; ++i)
                if (h[i] != PATTERN) {
                    bad = true;
                    report += "buffer of " + std::to_string(b.bytes) + " bytes: " + (side == 0 ? "front" : "rear") + " canary damaged at offset " +
                              std::to_string(side == 0 ? (long long)i - (long long)GUARD : (long long)i) + "; ";
                    break;
                }
        }
        if (bad) ++*n_damaged;
    }
    if (*n_damaged) { orb_set_error("guard check: " + report); return ORB_ERR_CUDA; }
    return ORB_OK;
}

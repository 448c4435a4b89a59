// Host-side helpers shared by the gforb translation units: a bump arena in device memory used to stage host inputs
// and hold per-call scratch (one cudaMalloc at create(), no allocation on the call path).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>
#include <map>
#include <mutex>
#include <utility>
#include <vector>

namespace gforb {

struct Arena {
    uint8_t* base = nullptr;
    size_t cap = 0, used = 0;
    cudaError_t init(size_t bytes) {
        cap = bytes; used = 0;
        return cudaMalloc((void**)&base, bytes);
    }
    void release() { if (base) cudaFree(base); base = nullptr; cap = used = 0; }
    void reset() { used = 0; }
    void* take(size_t bytes) {
        size_t a = (used + 255) & ~(size_t)255;
        if (a + bytes > cap) return nullptr;
        used = a + bytes;
        return base + a;
    }
};

// cudaFuncAttributeMaxDynamicSharedMemorySize is one value per (function, device) shared by every handle of the process:
// only ever raise it, so that creating a small handle after a large one cannot starve the large one's launches.
template <class F>
inline cudaError_t raise_dyn_smem(F func, int bytes, int device) {
    static std::mutex mu;
    static std::map<std::pair<const void*, int>, int> granted;
    std::lock_guard<std::mutex> g(mu);
    int& cur = granted[std::make_pair(reinterpret_cast<const void*>(func), device)];
    if (bytes <= cur) return cudaSuccess;
    const cudaError_t e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e == cudaSuccess) cur = bytes;
    return e;
}

// deferred device -> host copies of a call's outputs
struct OutCopy { void* host; const void* dev; size_t bytes; };

}  // namespace gforb

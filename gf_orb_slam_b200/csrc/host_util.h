// Host-side helpers shared by the gforb translation units: a bump arena in device memory used to stage host inputs
// and hold per-call scratch (one cudaMalloc at create(), no allocation on the call path).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>
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

// deferred device -> host copies of a call's outputs
struct OutCopy { void* host; const void* dev; size_t bytes; };

}  // namespace gforb

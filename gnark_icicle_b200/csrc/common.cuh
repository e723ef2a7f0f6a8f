// Shared host-side plumbing for libg16b200: error propagation, device buffers, the context object.
// No CPU fallback exists anywhere in this library: every entry point fails loudly when CUDA fails.
#pragma once
#include "curve_select.h"
#include <cuda_runtime.h>
#include <cstdint>
#include <atomic>
#include <cstdio>
#include <map>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include "curve.cuh"

namespace g16 {

struct CudaError : std::runtime_error {
    explicit CudaError(const std::string& s) : std::runtime_error(s) {}
};

#define G16_CUDA(expr)                                                                                   \
    do {                                                                                                 \
        cudaError_t _e = (expr);                                                                         \
        if (_e != cudaSuccess)                                                                           \
            throw ::g16::CudaError(std::string(#expr) + " failed: " + cudaGetErrorString(_e) + " at " +  \
                                   __FILE__ + ":" + std::to_string(__LINE__));                           \
    } while (0)

// every kernel launch of this library goes through this (bench.py reports the count as gpu_launches)
inline std::atomic<uint64_t>& launch_counter() { static std::atomic<uint64_t> c{0}; return c; }
#define G16_LAUNCH_CHECK()                 \
    do {                                   \
        ::g16::launch_counter()++;         \
        G16_CUDA(cudaGetLastError());      \
    } while (0)

// growable device buffer (never shrinks; freed with the context) -- replaces the per-proof
// cudaMalloc/cudaFree churn of the reference (icicle.go:356,415-419,492,505-507)
struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    void reserve(size_t bytes) {
        if (bytes <= cap) return;
        if (p) G16_CUDA(cudaFree(p));
        p = nullptr; cap = 0;
        G16_CUDA(cudaMalloc(&p, bytes));
        cap = bytes;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

struct PinnedBuf {
    void* p = nullptr;
    size_t cap = 0;
    void reserve(size_t bytes) {
        if (bytes <= cap) return;
        if (p) G16_CUDA(cudaFreeHost(p));
        p = nullptr; cap = 0;
        G16_CUDA(cudaMallocHost(&p, bytes));
        cap = bytes;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

inline int ilog2_ceil(uint64_t n) { int l = 0; while ((1ull << l) < n) l++; return l; }

}  // namespace g16

// Shared helpers for the cyten_b200 CUDA sources (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdarg>
#include <cstring>
#include "../../include/cyten_b200.h"

namespace cyb {

// ---- error plumbing -----------------------------------------------------------------------
void set_error(const char* fmt, ...);

#define CYB_CHECK_CUDA(expr)                                                              \
    do {                                                                                  \
        cudaError_t _e = (expr);                                                          \
        if (_e != cudaSuccess) {                                                          \
            cyb::set_error("%s failed at %s:%d: %s", #expr, __FILE__, __LINE__,           \
                           cudaGetErrorString(_e));                                       \
            return CYB_ERR_CUDA;                                                          \
        }                                                                                 \
    } while (0)

#define CYB_REQUIRE(cond, ...)                                                            \
    do {                                                                                  \
        if (!(cond)) {                                                                    \
            cyb::set_error(__VA_ARGS__);                                                  \
            return CYB_ERR_INVALID;                                                       \
        }                                                                                 \
    } while (0)

#define CYB_LAUNCH_CHECK()                                                                \
    do {                                                                                  \
        cudaError_t _e = cudaGetLastError();                                              \
        if (_e != cudaSuccess) {                                                          \
            cyb::set_error("kernel launch failed at %s:%d: %s", __FILE__, __LINE__,       \
                           cudaGetErrorString(_e));                                       \
            return CYB_ERR_CUDA;                                                          \
        }                                                                                 \
    } while (0)

int sm_count();

// Stream-ordered scratch memory owned by the library (cudaMallocAsync pool, kept cached).
struct Scratch {
    void* ptr = nullptr;
    cudaStream_t stream = nullptr;
    cudaError_t err = cudaSuccess;
    Scratch(size_t bytes, cudaStream_t s);
    ~Scratch();
    Scratch(const Scratch&) = delete;
    Scratch& operator=(const Scratch&) = delete;
    template <class T> T* as(size_t byte_offset = 0) const {
        return reinterpret_cast<T*>(static_cast<char*>(ptr) + byte_offset);
    }
};

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// One-time work PER DEVICE at a call site (cudaFuncSetAttribute is a per-device setting): ``static PerDeviceOnce once;
// if (once.first()) { ... }``.  Repeating the guarded work is harmless, so no lock.
struct PerDeviceOnce {
    unsigned long long done = 0ull;
    bool first()
    {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess) return true;
        const unsigned long long bit = 1ull << (dev & 63);
        if (done & bit) return false;
        done |= bit;
        return true;
    }
};

// ---- device helpers -----------------------------------------------------------------------
#ifdef __CUDACC__

// exclusive scan of one value per thread across a CTA of up to 1024 threads.
// `smem` must hold 33 T's. Returns the exclusive prefix; *total receives the CTA-wide sum.
template <class T>
__device__ __forceinline__ T block_exclusive_scan(T v, T* smem, T* total)
{
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    T incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        T o = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= (unsigned)d) incl += o;
    }
    __syncthreads();  // protect smem reuse between consecutive calls
    if (lane == 31) smem[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        const unsigned nw = (blockDim.x + 31u) >> 5;
        T w = lane < nw ? smem[lane] : T(0);
        T wi = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            T o = __shfl_up_sync(0xffffffffu, wi, d);
            if (lane >= (unsigned)d) wi += o;
        }
        smem[lane] = wi - w;  // exclusive per-warp offsets
        if (lane == 31) smem[32] = wi;
    }
    __syncthreads();
    T res = incl - v + smem[warp];
    *total = smem[32];
    return res;
}

__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    return v;
}

#endif  // __CUDACC__

}  // namespace cyb

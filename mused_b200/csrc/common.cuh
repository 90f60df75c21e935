// Shared helpers for the mused_b200 CUDA sources (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cooperative_groups.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <string>
#include <atomic>

#include "../../include/mused_b200.h"

namespace mused {

void set_error(const char* fmt, ...);

#define MUSED_CUDA(call)                                                              \
    do {                                                                              \
        cudaError_t _e = (call);                                                      \
        if (_e != cudaSuccess) {                                                      \
            mused::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(_e),  \
                             __FILE__, __LINE__);                                     \
            return MUSED_ERR_CUDA;                                                    \
        }                                                                             \
    } while (0)

#define MUSED_REQUIRE(cond, ...)            \
    do {                                    \
        if (!(cond)) {                      \
            mused::set_error(__VA_ARGS__);  \
            return MUSED_ERR_ARG;           \
        }                                   \
    } while (0)

// Opt a kernel in to more than 48 KB of dynamic shared memory, once per DEVICE (the attribute belongs to the
// device's context; a process may drive several GPUs and several host threads).
#define MUSED_SMEM_ATTR(func, bytes)                                                                       \
    do {                                                                                                   \
        static std::atomic<unsigned long long> _done{0ull};                                                \
        int _dev = 0;                                                                                      \
        MUSED_CUDA(cudaGetDevice(&_dev));                                                                  \
        if (!((_done.load(std::memory_order_acquire) >> (_dev & 63)) & 1ull)) {                            \
            MUSED_CUDA(cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes))); \
            _done.fetch_or(1ull << (_dev & 63), std::memory_order_release);                                \
        }                                                                                                  \
    } while (0)

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_prod(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v *= __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float warp_sum_f(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// 256-thread named barrier (ids 1..15; 0 is __syncthreads)
__device__ __forceinline__ void bar_sync_256(int id) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(256) : "memory");
}

// ---------------------------------------------------------------------------
// 64x64 output tile of a float64 GEMM computed by one 256-thread worker
// (16x16 threads, 4x4 register micro-tile, K staged 16 at a time in shared
// memory).  Operands are described by functors so that the same code serves
// the eigenvector update, the shrink-rotate, the Gram and the merge output:
//   fa(m, k) -> A[m][k]      fb(k, n) -> B[k][n]      fc(m, n, value)
// A_KFAST / B_NFAST say which index is contiguous in memory (coalescing).
// `sync` is the worker's barrier.
// ---------------------------------------------------------------------------
constexpr int GT = 64;   // tile edge
constexpr int GK = 16;   // k chunk
constexpr int GLD = GT + 2;
struct GemmSmem {
    double a[GK][GLD];
    double b[GK][GLD];
};

template <bool A_KFAST, bool B_NFAST, class FA, class FB, class FC, class SYNC>
__device__ __forceinline__ void gemm_tile_f64(GemmSmem& sm, int wt, int m0, int n0, int K, FA fa,
                                              FB fb, FC fc, SYNC sync) {
    double acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.0;
    const int ty = wt >> 4, tx = wt & 15;
    for (int k0 = 0; k0 < K; k0 += GK) {
#pragma unroll
        for (int e = wt; e < GT * GK; e += 256) {
            int mm, kk;
            if (A_KFAST) { kk = e & (GK - 1); mm = e >> 4; }
            else         { mm = e & (GT - 1); kk = e >> 6; }
            sm.a[kk][mm] = fa(m0 + mm, k0 + kk);
        }
#pragma unroll
        for (int e = wt; e < GT * GK; e += 256) {
            int nn, kk;
            if (B_NFAST) { nn = e & (GT - 1); kk = e >> 6; }
            else         { kk = e & (GK - 1); nn = e >> 4; }
            sm.b[kk][nn] = fb(k0 + kk, n0 + nn);
        }
        sync();
#pragma unroll
        for (int kk = 0; kk < GK; ++kk) {
            double av[4], bv[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) av[i] = sm.a[kk][ty * 4 + i];
#pragma unroll
            for (int j = 0; j < 4; ++j) bv[j] = sm.b[kk][tx * 4 + j];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fma(av[i], bv[j], acc[i][j]);
        }
        sync();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) fc(m0 + ty * 4 + i, n0 + tx * 4 + j, acc[i][j]);
}

}  // namespace mused

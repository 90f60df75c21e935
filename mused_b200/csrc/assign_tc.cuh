// K7: fused cosine-similarity (or Euclidean) + threshold + argmax event assignment on the
// tcgen05 tensor cores (north_star operator; nearest reference code
// matrix_operations.py:106 cosine_similarity and :286-289 nearest centroid).
//
// The scores of one 128-row tile of X against all centroids are a (128 x NP x d)
// GEMM; with k = 150 centroids that is 300 flop per 4 bytes of X, far above what
// fp32 SIMT can stream, so the products run on the tensor pipe and X is read from
// HBM exactly once:
//   warp 0      TMA producer: X tile(s) (128 rows x 32 floats, SWIZZLE_128B) and the
//               matching slices of the centroid hi / lo planes, per K step of 32
//   warps 2..   convert: split the X tile in place into TF32 hi (cvt.rna) and a lo
//               plane (fp32-exact 3xTF32 products), accumulate ||x||^2 of "their" row
//   warp 1      UMMA issue: acc += hi.hi^T + hi.lo^T + lo.hi^T (kind::tf32, fp32 in TMEM)
//   warps 2..   epilogue: tcgen05.ld the row of scores, scale by 1/||c_j||, running
//               argmax (ties -> lowest j), threshold against ||x||, write label/score
// MH = accumulators (128 rows each) per CTA, NST = pipeline stages; the launcher picks the
// deepest pipeline that fits the 227 KB of shared memory.
#pragma once
#include <float.h>
#include "umma.cuh"

namespace mused {
namespace atc {

constexpr int KB = 32;                 // floats of K per step (one 128-byte swizzle row)

struct Args {
    long long n;
    int k, NP, kblocks, metric;
    float tau;
    const float* cn2;                  // [NP] squared centroid norms (0 for padding)
    int* labels;
    float* score;                      // nullable
};

template <int MH, int NST>
struct Cfg {
    static constexpr int NTHREADS = 64 + MH * 128;
    static size_t stage_bytes(int NP) { return (size_t)MH * 2 * 128 * 128 + (size_t)2 * NP * 128; }
    static size_t smem_bytes(int NP) { return NST * stage_bytes(NP) + 1024 + 256 + (size_t)NP * 4; }
};

// centroids -> TF32 hi / lo planes padded to NP rows, squared norms
__global__ void k_assign_prep(const float* __restrict__ C, long long k, long long d, long long ldc, int NP,
                              float* __restrict__ Chi, float* __restrict__ Clo, float* __restrict__ cn2) {
    const int r = blockIdx.x;
    if (r >= NP) return;
    double acc = 0.0;
    for (long long c = threadIdx.x; c < d; c += blockDim.x) {
        const float v = (r < k) ? C[(size_t)r * ldc + c] : 0.f;
        float hi, lo;
        umma::split_tf32(v, hi, lo);
        Chi[(size_t)r * d + c] = hi;
        Clo[(size_t)r * d + c] = lo;
        acc += (double)v * (double)v;
    }
    __shared__ double red[8];
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int i = 0; i < (int)(blockDim.x >> 5); ++i) t += red[i];
        cn2[r] = (float)t;
    }
}

template <int MH, int NST>
__global__ void __launch_bounds__(Cfg<MH, NST>::NTHREADS, 1)
k_assign_tc(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_chi,
            const __grid_constant__ CUtensorMap tm_clo, Args A) {
    extern __shared__ unsigned char smem_dyn[];
    unsigned char* base = (unsigned char*)(((uintptr_t)smem_dyn + 1023) & ~(uintptr_t)1023);
    const int NP = A.NP;
    const uint32_t a_bytes = MH * 128 * 128;                   // one X plane (hi or lo) per stage
    const uint32_t c_bytes = (uint32_t)NP * 128;               // one centroid plane per stage
    const uint32_t stage_bytes = 2 * a_bytes + 2 * c_bytes;
    uint64_t* full = (uint64_t*)(base + NST * stage_bytes);
    uint64_t* conv = full + NST;
    uint64_t* empty = conv + NST;
    uint64_t* accum = empty + NST;
    uint32_t* tmem_slot = (uint32_t*)(accum + 1);
    float* cscale = (float*)(base + NST * stage_bytes + 256);   // [NP] 1/||c|| (cosine) or ||c||^2 (L2)

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long row0 = (long long)blockIdx.x * (MH * 128);

    if (warp == 0 && lane == 0) {
        umma::tma_prefetch_desc(&tm_x);
        umma::tma_prefetch_desc(&tm_chi);
        umma::tma_prefetch_desc(&tm_clo);
        for (int i = 0; i < NST; ++i) {
            umma::mbar_init(&full[i], 1);
            umma::mbar_init(&conv[i], MH * 128);
            umma::mbar_init(&empty[i], 1);
        }
        umma::mbar_init(accum, 1);
        umma::fence_mbar_init();
    }
    if (warp == 1) umma::tmem_alloc<512>(tmem_slot);
    for (int j = threadIdx.x; j < NP; j += blockDim.x) {
        const float c2 = A.cn2[j];
        cscale[j] = (A.metric == MUSED_METRIC_COSINE) ? rsqrtf(c2 > 0.f ? c2 : 1.f) : c2;
    }
    umma::tc_fence_before();
    __syncthreads();
    umma::tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            for (int kb = 0; kb < A.kblocks; ++kb) {
                const int st = kb % NST;
                if (kb >= NST) umma::mbar_wait(&empty[st], ((kb / NST) - 1) & 1);
                unsigned char* sb = base + st * stage_bytes;
                umma::mbar_expect_tx(&full[st], a_bytes + 2 * c_bytes);
#pragma unroll
                for (int h = 0; h < MH; ++h)
                    umma::tma_load_2d(sb + h * 128 * 128, &tm_x, &full[st], kb * KB, (int)(row0 + h * 128));
                umma::tma_load_2d(sb + 2 * a_bytes, &tm_chi, &full[st], kb * KB, 0);
                umma::tma_load_2d(sb + 2 * a_bytes + c_bytes, &tm_clo, &full[st], kb * KB, 0);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc = umma::idesc_tf32(128, NP, 0, 0);
            for (int kb = 0; kb < A.kblocks; ++kb) {
                const int st = kb % NST;
                umma::mbar_wait(&conv[st], (kb / NST) & 1);
                umma::tc_fence_after();
                const uint32_t sb = umma::smem_u32(base + st * stage_bytes);
                const uint32_t chi = sb + 2 * a_bytes, clo = chi + c_bytes;
#pragma unroll
                for (int h = 0; h < MH; ++h) {
                    const uint32_t ahi = sb + h * 128 * 128, alo = ahi + a_bytes;
                    const uint32_t d_tmem = tmem + h * 256;
#pragma unroll
                    for (int p = 0; p < 3; ++p) {
                        const uint32_t a_base = (p == 2) ? alo : ahi;
                        const uint32_t b_base = (p == 1) ? clo : chi;
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const uint64_t ad = umma::smem_desc_k_sw128(a_base + k * 32);
                            const uint64_t bd = umma::smem_desc_k_sw128(b_base + k * 32);
                            umma::mma_tf32(d_tmem, ad, bd, idesc, (kb | p | k) != 0);
                        }
                    }
                }
                umma::mma_commit(&empty[st]);
            }
            umma::mma_commit(accum);
        }
    } else {
        // convert + epilogue: this thread owns row (h, q, lane) of the CTA's tile, q = its TMEM lane quarter
        const int q = warp & 3, h = (warp - 2) >> 2;
        const int r_in = q * 32 + lane;
        float xn2 = 0.f;
        for (int kb = 0; kb < A.kblocks; ++kb) {
            const int st = kb % NST;
            umma::mbar_wait(&full[st], (kb / NST) & 1);
            unsigned char* hi_row = base + st * stage_bytes + h * 128 * 128 + r_in * 128;
            unsigned char* lo_row = hi_row + a_bytes;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int c = (lane + j) & 7;                  // rotated chunk order: conflict-free quarter warps
                float4 v = *reinterpret_cast<float4*>(hi_row + c * 16);
                xn2 = fmaf(v.x, v.x, xn2);
                xn2 = fmaf(v.y, v.y, xn2);
                xn2 = fmaf(v.z, v.z, xn2);
                xn2 = fmaf(v.w, v.w, xn2);
                float4 hv, lv;
                umma::split_tf32(v.x, hv.x, lv.x);
                umma::split_tf32(v.y, hv.y, lv.y);
                umma::split_tf32(v.z, hv.z, lv.z);
                umma::split_tf32(v.w, hv.w, lv.w);
                *reinterpret_cast<float4*>(hi_row + c * 16) = hv;
                *reinterpret_cast<float4*>(lo_row + c * 16) = lv;
            }
            umma::fence_proxy_async();                          // generic-proxy writes -> visible to the UMMA
            umma::mbar_arrive(&conv[st]);
        }
        umma::mbar_wait(accum, 0);
        umma::tc_fence_after();
        float best = -FLT_MAX;
        int bidx = 0x7fffffff;
        const uint32_t trow = tmem + ((uint32_t)(q * 32) << 16) + h * 256;
        for (int c0 = 0; c0 < NP; c0 += 32) {
            uint32_t v[32];
            umma::tmem_ld32(trow + c0, v);
            umma::tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                const int cj = c0 + j;
                if (cj < A.k) {
                    const float dot = __uint_as_float(v[j]);
                    const float key = (A.metric == MUSED_METRIC_COSINE) ? dot * cscale[cj]
                                                                        : -(cscale[cj] - 2.f * dot);
                    if (key > best) { best = key; bidx = cj; }
                }
            }
        }
        const long long r = row0 + h * 128 + r_in;
        if (r < A.n) {
            float sc;
            bool keep;
            if (A.metric == MUSED_METRIC_COSINE) {
                sc = best * rsqrtf(xn2 > 0.f ? xn2 : 1.f);
                keep = sc >= A.tau;
            } else {
                sc = sqrtf(fmaxf(xn2 - best, 0.f));
                keep = (A.tau < 0.f) || (sc <= A.tau);
            }
            A.labels[r] = keep ? bidx : -1;
            if (A.score) A.score[r] = sc;
        }
    }
    umma::tc_fence_before();
    __syncthreads();
    if (warp == 1) umma::tmem_dealloc<512>(tmem);
}

}  // namespace atc
}  // namespace mused

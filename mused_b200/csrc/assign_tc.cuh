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
//
// PASSES = 3 is the float32-exact product above (tensor-bound: three UMMAs per byte of X).  PASSES = 1 is the
// SCREENING kernel of the labels-only cosine call: one TF32 pass (hi.hi^T; the lo planes are neither written nor
// loaded), one persistent CTA per SM looping over 256-row tiles (two accumulators share every centroid slice: the
// L2 -> SM traffic of the centroid re-reads, 0.6 byte per byte of X, stays under the L2 throughput cap), the TMA
// producer running ahead into the next tile during the epilogue, so the kernel follows HBM instead of the tensor pipe.  Both operands are rounded to
// nearest TF32 (relative error <= 2^-11 each), hence |error of a cosine| <= 2^-10 for every (row, centroid) by
// Cauchy-Schwarz; SCREEN_EPS adds the accumulation slack.  The epilogue finds the keys within 2 eps of the row's best: a
// label is FINAL when the decision is outside that bound (best - second >= 2 eps and |cos - tau| >= eps, or
// cos + eps < tau whatever the argmax); otherwise the row's label slot receives a PENDING code naming what is
// undecided (the threshold only, two or three candidates, or more), and k_assign_rescore recomputes exactly those dots in
// float32 FMA from X (a second read of the undecided rows only) -- final labels are those of the exact kernel.
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
    uint32_t* cmask;                   // screening: [n][nw] bit masks of the keys within 2 eps of the row's best
    int nw;                            //            (read back by k_assign_rescore for rows with > 3 candidates)
};

#ifndef MUSED_ASSIGN_RAWX
#define MUSED_ASSIGN_RAWX 1
#endif
// RAWX: the screening UMMA reads the fp32 rows of X as TMA delivered them (kind::tf32 uses the upper 19 bits of each
// word: truncation, relative error < 2^-10) and the centroids rounded to nearest (<= 2^-11): |error of a cosine| <=
// 1.5 * 2^-10 = 1.465e-3 by Cauchy-Schwarz.  Otherwise X is rounded to nearest in shared memory first (2^-10 in all).
// SCREEN_EPS adds the fp32 accumulation slack (d <= 64k).
constexpr float SCREEN_EPS = MUSED_ASSIGN_RAWX ? 1.6e-3f : 1.1e-3f;

// pending codes left in labels[] by the screening kernel: label = -2 - (type << 24 | third << 16 | second << 8 | best)
constexpr int PEND_ONE = 0, PEND_PAIR = 1, PEND_TRIPLE = 2, PEND_FULL = 3;
__device__ __forceinline__ int pend_code(int type, int best, int second, int third) {
    return -2 - ((type << 24) | ((third & 255) << 16) | ((second & 255) << 8) | (best & 255));
}

template <int MH, int NST, int PASSES = 3>
struct Cfg {
    static constexpr int NTHREADS = 64 + MH * 128;
    static constexpr int PLANES = PASSES == 1 ? 1 : 2;       // hi (+ lo) planes of X and of the centroids per stage
    static constexpr int CTAS_PER_SM = 1;
    static constexpr int TMEM_COLS = MH == 1 ? 256 : 512;
    static size_t stage_bytes(int NP) { return (size_t)MH * PLANES * 128 * 128 + (size_t)PLANES * NP * 128; }
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

// Third pass of the screening epilogue for warps that hold a row with more than three candidate keys: the bit mask of
// the keys >= thr, one word per 32 centroids, written for the lanes that pass a destination.
__device__ __noinline__ void screen_write_masks(uint32_t trow, int NP, int k, const float* cscale, float thr, uint32_t* dst) {
    for (int c0 = 0; c0 < NP; c0 += 32) {
        uint32_t v[32];
        umma::tmem_ld32(trow + c0, v);
        umma::tmem_ld_wait();
        uint32_t m = 0;
#pragma unroll
        for (int j = 0; j < 32; ++j)
            if (c0 + j < k && __uint_as_float(v[j]) * cscale[c0 + j] >= thr) m |= 1u << j;
        if (dst) dst[c0 >> 5] = m;
    }
}

template <int MH, int NST, int PASSES = 3>
__global__ void __launch_bounds__(Cfg<MH, NST, PASSES>::NTHREADS, Cfg<MH, NST, PASSES>::CTAS_PER_SM)
k_assign_tc(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_chi,
            const __grid_constant__ CUtensorMap tm_clo, Args A) {
    using CF = Cfg<MH, NST, PASSES>;
    constexpr int PL = CF::PLANES;
    constexpr bool RAWX = PASSES == 1 && MUSED_ASSIGN_RAWX;     // no conversion of X in shared memory (see SCREEN_EPS)
    static_assert(PASSES == 1 || PASSES == 3, "one screening pass or the three passes of 3xTF32");
    extern __shared__ unsigned char smem_dyn[];
    unsigned char* base = (unsigned char*)(((uintptr_t)smem_dyn + 1023) & ~(uintptr_t)1023);
    const int NP = A.NP;
    const uint32_t a_bytes = MH * 128 * 128;                   // one X plane (hi or lo) per stage
    const uint32_t c_bytes = (uint32_t)NP * 128;               // one centroid plane per stage
    const uint32_t stage_bytes = PL * (a_bytes + c_bytes);
    uint64_t* full = (uint64_t*)(base + NST * stage_bytes);
    uint64_t* conv = full + NST;
    uint64_t* empty = conv + NST;
    uint64_t* accum = empty + NST;
    uint64_t* accfree = accum + 1;
    uint32_t* tmem_slot = (uint32_t*)(accfree + 1);
    float* cscale = (float*)(base + NST * stage_bytes + 256);   // [NP] 1/||c|| (cosine) or ||c||^2 (L2)

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // tiles of MH * 128 rows: blockIdx.x, blockIdx.x + gridDim.x, ...  (the exact kernel is launched with one tile
    // per CTA; the screening kernel with one CTA per SM: the producer runs ahead into the next tile while the
    // convert warps are in the epilogue, so HBM stays busy)
    const long long tiles = (A.n + MH * 128 - 1) / (MH * 128);

    if (warp == 0 && lane == 0) {
        umma::tma_prefetch_desc(&tm_x);
        umma::tma_prefetch_desc(&tm_chi);
        umma::tma_prefetch_desc(&tm_clo);
        for (int i = 0; i < NST; ++i) {
            umma::mbar_init(&full[i], 1);
            umma::mbar_init(&conv[i], MH * 128);
            umma::mbar_init(&empty[i], RAWX ? 1 + MH * 128 : 1);   // RAWX: the UMMA commit + every row's norm reader
        }
        umma::mbar_init(accum, 1);
        umma::mbar_init(accfree, MH * 128);
        umma::fence_mbar_init();
    }
    if (warp == 1) umma::tmem_alloc<CF::TMEM_COLS>(tmem_slot);
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
            uint32_t it = 0;                                    // K steps issued so far, over all tiles of this CTA
            for (long long tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
                const long long row0 = tile * (MH * 128);
                for (int kb = 0; kb < A.kblocks; ++kb, ++it) {
                    const uint32_t st = it % NST;
                    if (it >= NST) umma::mbar_wait(&empty[st], ((it / NST) - 1) & 1);
                    unsigned char* sb = base + st * stage_bytes;
                    umma::mbar_expect_tx(&full[st], a_bytes + PL * c_bytes);
#pragma unroll
                    for (int h = 0; h < MH; ++h)
                        umma::tma_load_2d(sb + h * 128 * 128, &tm_x, &full[st], kb * KB, (int)(row0 + h * 128));
                    umma::tma_load_2d(sb + PL * a_bytes, &tm_chi, &full[st], kb * KB, 0);
                    if constexpr (PL == 2) umma::tma_load_2d(sb + PL * a_bytes + c_bytes, &tm_clo, &full[st], kb * KB, 0);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc = umma::idesc_tf32(128, NP, 0, 0);
            uint32_t it = 0, t = 0;
            for (long long tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++t) {
                if (t > 0) {                                    // the epilogue of the previous tile has read the accumulators
                    umma::mbar_wait(accfree, (t - 1) & 1);
                    umma::tc_fence_after();
                }
                for (int kb = 0; kb < A.kblocks; ++kb, ++it) {
                    const uint32_t st = it % NST;
                    umma::mbar_wait(RAWX ? &full[st] : &conv[st], (it / NST) & 1);
                    umma::tc_fence_after();
                    const uint32_t sb = umma::smem_u32(base + st * stage_bytes);
                    const uint32_t chi = sb + PL * a_bytes, clo = chi + c_bytes;
#pragma unroll
                    for (int h = 0; h < MH; ++h) {
                        const uint32_t ahi = sb + h * 128 * 128, alo = ahi + a_bytes;
                        const uint32_t d_tmem = tmem + h * 256;
#pragma unroll
                        for (int p = 0; p < PASSES; ++p) {
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
        }
    } else {
        // convert + epilogue: this thread owns row (h, q, lane) of the CTA's tile, q = its TMEM lane quarter
        const int q = warp & 3, h = (warp - 2) >> 2;
        const int r_in = q * 32 + lane;
        const uint32_t trow = tmem + ((uint32_t)(q * 32) << 16) + h * 256;
        uint32_t it = 0, t = 0;
        for (long long tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++t) {
            float xn2 = 0.f;
            for (int kb = 0; kb < A.kblocks; ++kb, ++it) {
                const uint32_t st = it % NST;
                umma::mbar_wait(&full[st], (it / NST) & 1);
                unsigned char* hi_row = base + st * stage_bytes + h * 128 * 128 + r_in * 128;
                unsigned char* lo_row = hi_row + a_bytes;
                if constexpr (RAWX) {
                    // only the row's squared norm is needed: read the tile the UMMA is reading, release the stage
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const int c = (lane + j) & 7;
                        const float4 v = *reinterpret_cast<const float4*>(hi_row + c * 16);
                        xn2 = fmaf(v.x, v.x, xn2);
                        xn2 = fmaf(v.y, v.y, xn2);
                        xn2 = fmaf(v.z, v.z, xn2);
                        xn2 = fmaf(v.w, v.w, xn2);
                    }
                    umma::mbar_arrive(&empty[st]);
                } else {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int c = (lane + j) & 7;              // rotated chunk order: conflict-free quarter warps
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
                    if constexpr (PL == 2) *reinterpret_cast<float4*>(lo_row + c * 16) = lv;
                }
                umma::fence_proxy_async();                      // generic-proxy writes -> visible to the UMMA
                umma::mbar_arrive(&conv[st]);
                }
            }
            umma::mbar_wait(accum, t & 1);
            umma::tc_fence_after();
            const long long r = tile * (MH * 128) + h * 128 + r_in;
            if constexpr (PASSES == 1) {
                // screening (cosine), branch-free on purpose (data-dependent insertion sorts diverge in every warp):
                // pass 1 = the row's best key; pass 2 = bit mask of the keys within 2 eps of it, from which the
                // label is final (one bit, threshold clear) or a pending code (header comment)
                float b1 = -FLT_MAX;
                for (int c0 = 0; c0 < NP; c0 += 32) {
                    uint32_t v[32];
                    umma::tmem_ld32(trow + c0, v);
                    umma::tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (c0 + j < A.k) b1 = fmaxf(b1, __uint_as_float(v[j]) * cscale[c0 + j]);
                }
                const float xnorm = sqrtf(xn2 > 0.f ? xn2 : 1.f);
                const float sc = b1 / xnorm, thr = b1 - 2.f * SCREEN_EPS * xnorm;
                const bool below = sc + SCREEN_EPS < A.tau || r >= A.n;          // label -1 whichever centroid wins
                int cnt = 0, ca = 0, cb = 0, cc = 0;                              // first three candidates, ascending index
                if (!__all_sync(0xffffffffu, below)) {
                    for (int c0 = 0; c0 < NP; c0 += 32) {
                        uint32_t v[32];
                        umma::tmem_ld32(trow + c0, v);
                        umma::tmem_ld_wait();
                        uint32_t m = 0;
#pragma unroll
                        for (int j = 0; j < 32; ++j)
                            if (c0 + j < A.k && __uint_as_float(v[j]) * cscale[c0 + j] >= thr) m |= 1u << j;
                        while (m) {                                              // normally one bit in the whole row
                            const int j = __ffs(m) - 1;
                            m &= m - 1;
                            ca = cnt == 0 ? c0 + j : ca;
                            cb = cnt == 1 ? c0 + j : cb;
                            cc = cnt == 2 ? c0 + j : cc;
                            ++cnt;
                        }
                    }
                }
                // rare (more than three keys inside the bound): leave the rows' candidate masks for the re-score.  Kept
                // out of line on purpose: inlined, this never-taken loop cost the hot epilogue 10 % (measured).
                if (__any_sync(0xffffffffu, cnt > 3))
                    screen_write_masks(trow, NP, A.k, cscale, thr, cnt > 3 && r < A.n ? A.cmask + (size_t)r * A.nw : nullptr);
                umma::tc_fence_before();
                umma::mbar_arrive(accfree);
                if (r < A.n) {
                    int lab;
                    if (below) lab = -1;
                    else if (cnt > 3) lab = pend_code(PEND_FULL, ca, ca, ca);
                    else if (cnt == 3) lab = pend_code(PEND_TRIPLE, ca, cb, cc);
                    else if (cnt == 2) lab = pend_code(PEND_PAIR, ca, cb, cb);
                    else if (sc - SCREEN_EPS < A.tau) lab = pend_code(PEND_ONE, ca, ca, ca);
                    else lab = ca;
                    A.labels[r] = lab;
                }
            } else {
                float best = -FLT_MAX;
                int bidx = 0x7fffffff;
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
                umma::tc_fence_before();
                umma::mbar_arrive(accfree);
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
        }
    }
    umma::tc_fence_before();
    __syncthreads();
    if (warp == 1) umma::tmem_dealloc<CF::TMEM_COLS>(tmem);
}

// Second step of the screening path: every label slot holding a pending code is decided in float32 FMA arithmetic
// from the row itself (PEND_ONE: the threshold of a known winner; PEND_PAIR / PEND_TRIPLE: the candidates named in
// the code; PEND_FULL: the candidates of the row's bit mask).  A warp scans 32 label slots, then takes its pending rows one at a time, lanes
// striding the d columns; up to four centroids per pass over the row so their loads are in flight together.
__device__ __forceinline__ void rescore_dots(const float* __restrict__ x, const float* __restrict__ C, long long ldc,
                                             const int (&j)[4], int nj, long long d, int lane, float (&dot)[4], float& xn2) {
    float acc[4] = {0.f, 0.f, 0.f, 0.f}, xx = 0.f;
    for (long long i = 4 * lane; i < d; i += 128) {          // d % 32 == 0 and 16-byte aligned rows on this path
        const float4 xv = *reinterpret_cast<const float4*>(x + i);
        xx = fmaf(xv.x, xv.x, fmaf(xv.y, xv.y, fmaf(xv.z, xv.z, fmaf(xv.w, xv.w, xx))));
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            if (c < nj) {
                const float4 cv = *reinterpret_cast<const float4*>(C + (size_t)j[c] * ldc + i);
                acc[c] = fmaf(xv.x, cv.x, fmaf(xv.y, cv.y, fmaf(xv.z, cv.z, fmaf(xv.w, cv.w, acc[c]))));
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        xx += __shfl_xor_sync(0xffffffffu, xx, o);
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[c] += __shfl_xor_sync(0xffffffffu, acc[c], o);
    }
#pragma unroll
    for (int c = 0; c < 4; ++c) dot[c] = acc[c];
    xn2 = xx;
}

__global__ void __launch_bounds__(256) k_assign_rescore(const float* __restrict__ X, long long n, long long d, long long ldx,
                                                        const float* __restrict__ C, int k, long long ldc,
                                                        const float* __restrict__ cn2, float tau, int* __restrict__ labels,
                                                        const uint32_t* __restrict__ cmask, int nw,
                                                        unsigned long long* __restrict__ n_pending) {
    const int lane = threadIdx.x & 31;
    const long long warps = (long long)gridDim.x * (blockDim.x >> 5);
    unsigned long long mine = 0;
    for (long long base = ((long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * 32; base < n; base += warps * 32) {
        const int lab = (base + lane < n) ? labels[base + lane] : 0;
        unsigned pend = __ballot_sync(0xffffffffu, lab <= -2);
        mine += __popc(pend);
        while (pend) {
            const int src = __ffs(pend) - 1;
            pend &= pend - 1;
            const int code = -__shfl_sync(0xffffffffu, lab, src) - 2;
            const int type = code >> 24;
            const long long r = base + src;
            const float* x = X + (size_t)r * ldx;
            float best = -FLT_MAX, xn2 = 0.f, dot[4];
            int bidx = 0x7fffffff;
            int j[4] = {code & 255, (code >> 8) & 255, (code >> 16) & 255, 0};
            if (type != PEND_FULL) {
                const int nj = type + 1;
                rescore_dots(x, C, ldc, j, nj, d, lane, dot, xn2);
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    if (c < nj) {
                        const float c2 = cn2[j[c]];
                        const float key = dot[c] * rsqrtf(c2 > 0.f ? c2 : 1.f);
                        if (key > best || (key == best && j[c] < bidx)) { best = key; bidx = j[c]; }   // ties -> lowest index
                    }
                }
            } else {
                // more than three candidates: walk the row's candidate mask, four at a time, ascending index
                int w = 0;
                uint32_t m = cmask[(size_t)r * nw];
                for (;;) {
                    int nj = 0;
                    while (nj < 4) {
                        while (!m && w + 1 < nw) m = cmask[(size_t)r * nw + ++w];
                        if (!m) break;
                        j[nj++] = 32 * w + __ffs(m) - 1;
                        m &= m - 1;
                    }
                    if (nj == 0) break;
                    for (int c = nj; c < 4; ++c) j[c] = j[0];
                    rescore_dots(x, C, ldc, j, nj, d, lane, dot, xn2);
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        if (c < nj) {
                            const float c2 = cn2[j[c]];
                            const float key = dot[c] * rsqrtf(c2 > 0.f ? c2 : 1.f);
                            if (key > best) { best = key; bidx = j[c]; }                               // ascending index
                        }
                    }
                    if (nj < 4) break;
                }
            }
            const float sc = best * rsqrtf(xn2 > 0.f ? xn2 : 1.f);
            if (lane == 0) labels[r] = sc >= tau ? bidx : -1;
        }
    }
    if (n_pending && lane == 0 && mine) atomicAdd(n_pending, mine);
}

}  // namespace atc
}  // namespace mused

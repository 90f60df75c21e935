// K4: shrink + rotate of every FD instance on the tcgen05 tensor cores:
//     sketch'_s (l x d) = T_s (l x 2l) . R_s (2l x d),   R_s = [sketch_s ; chunk]
// (swfd/frequent_directions.py:58 `sketch = Vt * sqrt(sigma^2)` with the shrink of
// :86-93 folded into the coefficient rows T_s = sqrt(sigma^2_r / lambda_r) v_r^T that
// the eigensolve emits).
//
// One CTA = one instance x one tile of 128 columns of d.  D[r][c] (M = 128 rows r,
// N = 128 columns c, fp32 in 128 TMEM columns).  A = T is K-major (rows of 32
// coefficients, 128 B, SWIZZLE_128B) exactly like the Gram operands; B = R is
// MN-MAJOR in the SWIZZLE_128B_BASE32B layout (the only one the tensor core takes
// for MN-major TF32): a TMA box of {32 columns c, k rows} written with
// CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B lands as 128-byte rows "32 consecutive c of
// one k", 4 such rows form one 512-byte swizzle atom, atoms follow each other along
// K every 512 B (SBO) and along N every `rows x 128` B (LBO).  3xTF32 like the Gram: Thi.Rhi + Thi.Rlo + Tlo.Rhi.  The epilogue splits
// the result into the hi / lo planes of the instance's other ping-pong buffer; rows
// whose direction is dumped (fastfdwithdump.py:115-131) go to their snapshot-ring
// entry instead and leave a zero row in the sketch.
#pragma once
#include "bswfd.h"
#include "umma.cuh"

namespace mused {

struct RotCfg {
    static constexpr int KB = 32;                    // K elements per stage
    static constexpr int NT = 128;                   // columns of d per CTA
    static constexpr int A_BYTES = 128 * 128;        // one plane of the T tile (128 rows x 32 coefficients)
    static constexpr int B_BYTES = (NT / 32) * KB * 128;   // one plane of the R tile: 4 chunks x 32 k-rows x 128 B
    static constexpr int STAGE_BYTES = 2 * A_BYTES + 2 * B_BYTES;
    static constexpr int NST = 3;
    static constexpr int SMEM_BYTES = NST * STAGE_BYTES + 1024 + 256;
};

__global__ void __launch_bounds__(192, 1)
k_rotate_tc(const __grid_constant__ CUtensorMap tm_t_hi, const __grid_constant__ CUtensorMap tm_t_lo,
            const __grid_constant__ CUtensorMap tm_sk_hi, const __grid_constant__ CUtensorMap tm_sk_lo,
            const __grid_constant__ CUtensorMap tm_x_hi, const __grid_constant__ CUtensorMap tm_x_lo, BDev D) {
    using C = RotCfg;
    const int s = blockIdx.y;
    const BSlot& sl = D.slots[s];
    // raw-row snapshots of this chunk (sliding_windows_frequent_directions.py:56-80): this CTA's 128 columns
    // of every such row, staging planes -> ring entry chosen by K5.  `nw` warps, warp `wi` takes rows wi, wi+nw, ..
    const int nraw = sl.nraw;
    auto raw_copy = [&](int wi, int nw) {
        const int* rl = D.raw_list + (size_t)s * 2 * D.ell;
        const long long c = (long long)blockIdx.x * RotCfg::NT + 4 * (threadIdx.x & 31);
        if (c >= D.d) return;
        for (int k = wi; k < nraw; k += nw) {
            const size_t xr = ((size_t)(s / D.S1) * D.xrows + D.x0 + rl[k]) * D.d + c;
            const float4 h = *reinterpret_cast<const float4*>(D.xhi + xr);
            const float4 l = *reinterpret_cast<const float4*>(D.xlo + xr);
            *reinterpret_cast<float4*>(D.ringV + ((size_t)s * D.capn + rl[D.ell + k]) * D.d + c) =
                make_float4(h.x + l.x, h.y + l.y, h.z + l.z, h.w + l.w);
        }
    };
    const int mt = blockIdx.z;                        // rows [128 mt, +128) of the new sketch (l > 128: l / 128 CTAs per tile)
    if (sl.skip) {
        if (nraw > 0 && mt == 0) raw_copy(threadIdx.x >> 5, 6);
        return;
    }
    extern __shared__ unsigned char smem_dyn[];
    unsigned char* base = (unsigned char*)(((uintptr_t)smem_dyn + 1023) & ~(uintptr_t)1023);
    uint64_t* full = (uint64_t*)(base + C::NST * C::STAGE_BYTES);
    uint64_t* empty = full + C::NST;
    uint64_t* accum = empty + C::NST;
    uint32_t* tmem_slot = (uint32_t*)(accum + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ell = D.ell, n2 = D.n2;
    const int c0 = blockIdx.x * C::NT;
    const int src = sl.src;
    const int xbase = (s / D.S1) * (int)D.xrows + D.x0;
    const int kblocks = n2 / C::KB;                  // n2 = 32, 64, 128, 256
    const int brows = ell < C::KB ? ell : C::KB;     // rows of one R box (a box never mixes sketch and chunk rows)

    if (warp == 0 && lane == 0) {
        umma::tma_prefetch_desc(&tm_t_hi);
        umma::tma_prefetch_desc(&tm_t_lo);
        umma::tma_prefetch_desc(&tm_sk_hi);
        umma::tma_prefetch_desc(&tm_sk_lo);
        umma::tma_prefetch_desc(&tm_x_hi);
        umma::tma_prefetch_desc(&tm_x_lo);
        for (int i = 0; i < C::NST; ++i) {
            umma::mbar_init(&full[i], 1);
            umma::mbar_init(&empty[i], 1);
        }
        umma::mbar_init(accum, 1);
        umma::fence_mbar_init();
    }
    if (warp == 1) umma::tmem_alloc<128>(tmem_slot);
    umma::tc_fence_before();
    __syncthreads();
    umma::tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            const int t_rows = ell < 128 ? ell : 128;                 // rows of one T box
            const int row_t = s * ell + 128 * mt;
            const int row_sk = (s * 2 + src) * ell;
            for (int kb = 0; kb < kblocks; ++kb) {
                const int st = kb % C::NST;
                if (kb >= C::NST) umma::mbar_wait(&empty[st], ((kb / C::NST) - 1) & 1);
                unsigned char* a_hi = base + st * C::STAGE_BYTES;
                unsigned char* a_lo = a_hi + C::A_BYTES;
                unsigned char* b_hi = a_lo + C::A_BYTES;
                unsigned char* b_lo = b_hi + C::B_BYTES;
                umma::mbar_expect_tx(&full[st], 2 * t_rows * 128 + 2 * (C::NT / 32) * C::KB * 128);
                umma::tma_load_2d(a_hi, &tm_t_hi, &full[st], kb * C::KB, row_t);
                umma::tma_load_2d(a_lo, &tm_t_lo, &full[st], kb * C::KB, row_t);
                for (int ch = 0; ch < C::NT / 32; ++ch) {
                    for (int r0 = 0; r0 < C::KB; r0 += brows) {
                        const int k = kb * C::KB + r0;                // row of R
                        const int off = ch * C::KB * 128 + r0 * 128;
                        if (k < ell) {
                            umma::tma_load_2d(b_hi + off, &tm_sk_hi, &full[st], c0 + ch * 32, row_sk + k);
                            umma::tma_load_2d(b_lo + off, &tm_sk_lo, &full[st], c0 + ch * 32, row_sk + k);
                        } else {
                            umma::tma_load_2d(b_hi + off, &tm_x_hi, &full[st], c0 + ch * 32, xbase + k - ell);
                            umma::tma_load_2d(b_lo + off, &tm_x_lo, &full[st], c0 + ch * 32, xbase + k - ell);
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = umma::idesc_tf32(128, C::NT, 0, 1);     // A K-major, B MN-major
            for (int kb = 0; kb < kblocks; ++kb) {
                const int st = kb % C::NST;
                umma::mbar_wait(&full[st], (kb / C::NST) & 1);
                umma::tc_fence_after();
                const uint32_t a_hi = umma::smem_u32(base + st * C::STAGE_BYTES);
                const uint32_t a_lo = a_hi + C::A_BYTES;
                const uint32_t b_hi = a_lo + C::A_BYTES;
                const uint32_t b_lo = b_hi + C::B_BYTES;
#pragma unroll
                for (int p = 0; p < 3; ++p) {
                    const uint32_t a_base = (p == 2 ? a_lo : a_hi);
                    const uint32_t b_base = (p == 1 ? b_lo : b_hi);
#pragma unroll
                    for (int k = 0; k < C::KB / 8; ++k) {
                        const uint64_t ad = umma::smem_desc_k_sw128(a_base + k * 32);
                        const uint64_t bd = umma::smem_desc_mn_sw128_32b(b_base + k * 1024, C::KB * 128, 512);
                        umma::mma_tf32(tmem, ad, bd, idesc, (kb | p | k) != 0);
                    }
                }
                umma::mma_commit(&empty[st]);
            }
            umma::mma_commit(accum);
        }
    } else {
        const int q = warp & 3;
        if (nraw > 0 && mt == 0) raw_copy(warp - 2, 4);  // under the TMA / MMA pipeline of warps 0 and 1
        umma::mbar_wait(accum, 0);
        umma::tc_fence_after();
        const int r = 128 * mt + q * 32 + lane;
        const long long d = D.d;
        int ds = -1;
        if (r < ell) ds = D.dump_slot[(size_t)s * ell + r];
        float* ring = ds >= 0 ? D.ringV + ((size_t)s * D.capn + ds) * d : nullptr;
        float* dh = D.skhi + (((size_t)s * 2 + (src ^ 1)) * ell + r) * d;
        float* dl = D.sklo + (((size_t)s * 2 + (src ^ 1)) * ell + r) * d;
#pragma unroll 1
        for (int cc = 0; cc < C::NT; cc += 32) {
            uint32_t v[32];
            umma::tmem_ld32(tmem + ((uint32_t)(q * 32) << 16) + cc, v);
            umma::tmem_ld_wait();
            const int c = c0 + cc;
            if (r < ell && c < d) {                      // d % 32 == 0: a 32-column group is all in or all out
                float4 *ph = (float4*)(dh + c), *pl = (float4*)(dl + c);
                float4* pr = ring ? (float4*)(ring + c) : nullptr;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    float h[4], l[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float x = __uint_as_float(v[4 * j + e]);
                        if (ring) { h[e] = 0.f; l[e] = 0.f; }
                        else umma::split_tf32(x, h[e], l[e]);
                    }
                    if (pr) pr[j] = make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]),
                                                __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3]));
                    ph[j] = make_float4(h[0], h[1], h[2], h[3]);
                    pl[j] = make_float4(l[0], l[1], l[2], l[3]);
                }
            }
        }
    }
    umma::tc_fence_before();
    __syncthreads();
    if (warp == 1) umma::tmem_dealloc<128>(tmem);
}

}  // namespace mused

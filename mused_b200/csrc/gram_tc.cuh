// K2: Gram matrices G_s = R_s R_s^T of the stacked rows R_s = [sketch_s ; chunk]
// (2l x d) of every FD instance, on the tcgen05 tensor cores.
//
// Replaces the left half of the reference's thin SVD of [sketch; buffer]
// (swfd/frequent_directions.py:45-55): sigma^2 = eig(R R^T).
//
// Precision: operands are fp32 split into TF32 hi / lo planes (stored that way
// in HBM by the producers), D = hi.hi^T + hi.lo^T + lo.hi^T accumulated in fp32
// in TMEM ("3xTF32").  Split-K over d: CTA (ks, s) contracts the columns
// [ks*dk, (ks+1)*dk) and writes one fp32 partial; the eigensolve adds the
// partials in float64 in a fixed order (deterministic, no atomics).
//
// Structure per CTA (192 threads): warp 0 = TMA producer (one lane), warp 1 =
// TMEM allocation + UMMA issue (one lane), warps 2..5 = epilogue
// (tcgen05.ld -> global).  Operand tiles are K-major rows of 32 floats (128 B,
// SWIZZLE_128B): rows [0, l) come from the instance's sketch plane, rows [l, 2l)
// from the shared chunk plane, so A (M = 128 rows) and B (N = 2l rows) are two
// views of the SAME shared-memory tile.
#pragma once
#include "umma.cuh"
#include "eig_common.cuh"

namespace mused {

struct GramArgs {
    const BSlot* slots;   // cur plane per instance
    float* Gp;            // [S][KS][n2][n2]
    int ell, KS, kblocks; // kblocks = k-blocks of 32 floats per CTA
    int x0;               // first row of the open chunk in the staging planes
    int S1, xrows;        // instances per stream, plane rows per stream
};

template <int N2>
struct GramCfg {
    static constexpr int ROWS = (N2 < 128) ? 128 : N2;         // rows addressable by the A view
    static constexpr int MH = (N2 > 128) ? N2 / 128 : 1;       // M halves
    static constexpr int TILE_BYTES = ROWS * 128;
    static constexpr int STAGE_BYTES = 2 * TILE_BYTES;         // hi + lo
    static constexpr int NST = (N2 >= 256) ? 3 : 4;
    static constexpr int TMEM_COLS = (MH * N2 < 32) ? 32 : MH * N2;
    static constexpr int SMEM_BYTES = NST * STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/;
};

template <int N2>
__global__ void __launch_bounds__(192, 1)
k_gram_tc(const __grid_constant__ CUtensorMap tm_sk_hi, const __grid_constant__ CUtensorMap tm_sk_lo,
          const __grid_constant__ CUtensorMap tm_x_hi, const __grid_constant__ CUtensorMap tm_x_lo, GramArgs A) {
    using C = GramCfg<N2>;
    constexpr int ELL = N2 / 2;
    extern __shared__ unsigned char smem_dyn[];
    unsigned char* base = (unsigned char*)(((uintptr_t)smem_dyn + 1023) & ~(uintptr_t)1023);
    uint64_t* full = (uint64_t*)(base + C::NST * C::STAGE_BYTES);
    uint64_t* empty = full + C::NST;
    uint64_t* accum = empty + C::NST;
    uint32_t* tmem_slot = (uint32_t*)(accum + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ks = blockIdx.x, s = blockIdx.y;
    const int cur = A.slots[s].cur;
    const int xrow = (s / A.S1) * A.xrows + A.x0;

    if (warp == 0 && lane == 0) {
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
    if (warp == 1) umma::tmem_alloc<C::TMEM_COLS>(tmem_slot);
    umma::tc_fence_before();
    __syncthreads();
    umma::tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            const int row_sk = (s * 2 + cur) * ELL;
            for (int kb = 0; kb < A.kblocks; ++kb) {
                const int st = kb % C::NST;
                if (kb >= C::NST) umma::mbar_wait(&empty[st], ((kb / C::NST) - 1) & 1);
                unsigned char* hi = base + st * C::STAGE_BYTES;
                unsigned char* lo = hi + C::TILE_BYTES;
                const int col = (ks * A.kblocks + kb) * 32;
                umma::mbar_expect_tx(&full[st], 2 * N2 * 128);
                umma::tma_load_2d(hi, &tm_sk_hi, &full[st], col, row_sk);
                umma::tma_load_2d(hi + ELL * 128, &tm_x_hi, &full[st], col, xrow);
                umma::tma_load_2d(lo, &tm_sk_lo, &full[st], col, row_sk);
                umma::tma_load_2d(lo + ELL * 128, &tm_x_lo, &full[st], col, xrow);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = umma::idesc_tf32(128, N2, 0, 0);
            for (int kb = 0; kb < A.kblocks; ++kb) {
                const int st = kb % C::NST;
                umma::mbar_wait(&full[st], (kb / C::NST) & 1);
                umma::tc_fence_after();
                const uint32_t hi = umma::smem_u32(base + st * C::STAGE_BYTES);
                const uint32_t lo = hi + C::TILE_BYTES;
#pragma unroll
                for (int h = 0; h < C::MH; ++h) {
                    const uint32_t d_tmem = tmem + h * N2;
                    const uint32_t a_off = h * 128 * 128;
#pragma unroll
                    for (int p = 0; p < 3; ++p) {
                        const uint32_t a_base = (p == 2 ? lo : hi) + a_off;
                        const uint32_t b_base = (p == 1 ? lo : hi);
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
        // epilogue: warp q reads TMEM lanes [32q, 32q+32)
        const int q = warp & 3;
        umma::mbar_wait(accum, 0);
        umma::tc_fence_after();
        float* G = A.Gp + ((size_t)s * A.KS + ks) * N2 * N2;
#pragma unroll 1
        for (int h = 0; h < C::MH; ++h) {
            const int row = h * 128 + q * 32 + lane;
#pragma unroll 1
            for (int c0 = 0; c0 < N2; c0 += 32) {
                uint32_t v[32];
                umma::tmem_ld32(tmem + ((uint32_t)(q * 32) << 16) + h * N2 + c0, v);
                umma::tmem_ld_wait();
                if (row < N2) {
                    float4* dst = (float4*)(G + (size_t)row * N2 + c0);
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        dst[j] = make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]),
                                             __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3]));
                }
            }
        }
    }
    umma::tc_fence_before();
    __syncthreads();
    if (warp == 1) umma::tmem_dealloc<C::TMEM_COLS>(tmem);
}

// ---------------------------------------------------------------------------------------------
// n2 = 512 (l = 256): the Gram no longer fits one accumulator (512 TMEM columns, UMMA N <= 256), so the output
// is cut into 128 x 256 tiles; CTA (ks, s, z) computes tile (mt, nt) = (z / NTN, z % NTN) of the split-K
// partial: A = rows [128 mt, +128) (TMA box of 128 rows), B = rows [256 nt, +256) (box of 256 rows) of
// R = [sketch; chunk], each from the sketch plane (rows < l) or the shared chunk plane.  Same 3xTF32 passes,
// same warp roles; two stages of 96 KB.
// ---------------------------------------------------------------------------------------------
struct GramTiledCfg {
    static constexpr int TM = 128, TN = 256;
    static constexpr int A_BYTES = TM * 128, B_BYTES = TN * 128;            // one plane each
    static constexpr int STAGE_BYTES = 2 * (A_BYTES + B_BYTES);
    static constexpr int NST = 2;
    static constexpr int SMEM_BYTES = NST * STAGE_BYTES + 1024 + 256;
};

__global__ void __launch_bounds__(192, 1)
k_gram_tiled(const __grid_constant__ CUtensorMap ta_sk_hi, const __grid_constant__ CUtensorMap ta_sk_lo,
             const __grid_constant__ CUtensorMap ta_x_hi, const __grid_constant__ CUtensorMap ta_x_lo,
             const __grid_constant__ CUtensorMap tb_sk_hi, const __grid_constant__ CUtensorMap tb_sk_lo,
             const __grid_constant__ CUtensorMap tb_x_hi, const __grid_constant__ CUtensorMap tb_x_lo, GramArgs A) {
    using C = GramTiledCfg;
    extern __shared__ unsigned char smem_dyn[];
    unsigned char* base = (unsigned char*)(((uintptr_t)smem_dyn + 1023) & ~(uintptr_t)1023);
    uint64_t* full = (uint64_t*)(base + C::NST * C::STAGE_BYTES);
    uint64_t* empty = full + C::NST;
    uint64_t* accum = empty + C::NST;
    uint32_t* tmem_slot = (uint32_t*)(accum + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ks = blockIdx.x, s = blockIdx.y;
    const int ell = A.ell, n2 = 2 * ell;
    const int ntn = n2 / C::TN;
    const int mt = blockIdx.z / ntn, nt = blockIdx.z % ntn;
    const int cur = A.slots[s].cur;
    const int xrow = (s / A.S1) * A.xrows + A.x0;
    const int row_sk = (s * 2 + cur) * ell;
    // source of the A rows / B rows: the sketch plane or the chunk plane
    const bool a_sk = C::TM * mt < ell, b_sk = C::TN * nt < ell;
    const int a_row = a_sk ? row_sk + C::TM * mt : xrow + C::TM * mt - ell;
    const int b_row = b_sk ? row_sk + C::TN * nt : xrow + C::TN * nt - ell;

    if (warp == 0 && lane == 0) {
        umma::tma_prefetch_desc(a_sk ? &ta_sk_hi : &ta_x_hi);
        umma::tma_prefetch_desc(b_sk ? &tb_sk_hi : &tb_x_hi);
        for (int i = 0; i < C::NST; ++i) {
            umma::mbar_init(&full[i], 1);
            umma::mbar_init(&empty[i], 1);
        }
        umma::mbar_init(accum, 1);
        umma::fence_mbar_init();
    }
    if (warp == 1) umma::tmem_alloc<C::TN>(tmem_slot);
    umma::tc_fence_before();
    __syncthreads();
    umma::tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            for (int kb = 0; kb < A.kblocks; ++kb) {
                const int st = kb % C::NST;
                if (kb >= C::NST) umma::mbar_wait(&empty[st], ((kb / C::NST) - 1) & 1);
                unsigned char* a_hi = base + st * C::STAGE_BYTES;
                unsigned char* a_lo = a_hi + C::A_BYTES;
                unsigned char* b_hi = a_lo + C::A_BYTES;
                unsigned char* b_lo = b_hi + C::B_BYTES;
                const int col = (ks * A.kblocks + kb) * 32;
                umma::mbar_expect_tx(&full[st], C::STAGE_BYTES);
                umma::tma_load_2d(a_hi, a_sk ? &ta_sk_hi : &ta_x_hi, &full[st], col, a_row);
                umma::tma_load_2d(a_lo, a_sk ? &ta_sk_lo : &ta_x_lo, &full[st], col, a_row);
                umma::tma_load_2d(b_hi, b_sk ? &tb_sk_hi : &tb_x_hi, &full[st], col, b_row);
                umma::tma_load_2d(b_lo, b_sk ? &tb_sk_lo : &tb_x_lo, &full[st], col, b_row);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = umma::idesc_tf32(C::TM, C::TN, 0, 0);
            for (int kb = 0; kb < A.kblocks; ++kb) {
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
                    for (int k = 0; k < 4; ++k) {
                        const uint64_t ad = umma::smem_desc_k_sw128(a_base + k * 32);
                        const uint64_t bd = umma::smem_desc_k_sw128(b_base + k * 32);
                        umma::mma_tf32(tmem, ad, bd, idesc, (kb | p | k) != 0);
                    }
                }
                umma::mma_commit(&empty[st]);
            }
            umma::mma_commit(accum);
        }
    } else {
        const int q = warp & 3;
        umma::mbar_wait(accum, 0);
        umma::tc_fence_after();
        float* G = A.Gp + ((size_t)s * A.KS + ks) * n2 * n2;
        const int row = C::TM * mt + q * 32 + lane;
#pragma unroll 1
        for (int c0 = 0; c0 < C::TN; c0 += 32) {
            uint32_t v[32];
            umma::tmem_ld32(tmem + ((uint32_t)(q * 32) << 16) + c0, v);
            umma::tmem_ld_wait();
            float4* dst = (float4*)(G + (size_t)row * n2 + C::TN * nt + c0);
#pragma unroll
            for (int j = 0; j < 8; ++j)
                dst[j] = make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]),
                                     __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3]));
        }
    }
    umma::tc_fence_before();
    __syncthreads();
    if (warp == 1) umma::tmem_dealloc<C::TN>(tmem);
}

}  // namespace mused

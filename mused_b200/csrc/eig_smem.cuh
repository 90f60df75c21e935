// K3: batched symmetric eigensolve of the (2l x 2l) Gram matrices of the batched FD shrink.
// One thread-block cluster per FD instance, the whole matrix in SHARED memory, 2 + 2 columns per warp step.
//
// The thin SVD of [sketch; batch] that the reference takes with LAPACK
// (swfd/frequent_directions.py:55) is replaced by the eigen-decomposition of G = R R^T.  One-sided
// (Hestenes) Jacobi on the columns of W = G: plane rotations from the right orthogonalise the
// columns, W V -> V Lambda, so column norms converge to the eigenvalues and w_j / lambda_j to the
// eigenvectors; no separate V is carried.
//
// Layout: CTA c of the cluster (n2 / 32 CTAs, 256 threads = 8 warps, two CTAs per SM up to n2 = 256) holds 32
// columns, two blocks of 16, all n2 rows, column-major in shared memory.  In a ROUND warp w keeps columns 2w, 2w+1
// of the CTA's first block in registers and the column pairs of the second block pass by through shared memory:
// 8 STEPS per round, each a 2 x 2 block of rotations (four inner products, packed FFMA2 on float32, one 5-stage
// shuffle reduction, the angles, the rotations, 128-bit shared-memory accesses), then __syncthreads.  Squared column
// norms are cached and follow the rotations (only the cross products are reduced).  Between rounds the blocks move
// through distributed shared memory following the round-robin tournament on 2 NC blocks (block 0 fixed, 2 NC - 1
// rounds: every block pair meets once); pairs inside a block are rotated at the head of the sweep (7 steps on 8
// column pairs per block + the pair inside each 2-column unit).  A sweep at l = 128 is 15 rounds x 8 steps + the
// head, ~214 k cycles in float32.  Round 1 of this repo kept the matrix in the REGISTERS of 4 warps per CTA
// (8 columns per warp): one warp per scheduler; DESIGN.md 4b lists what else was measured.
//
// Mixed precision as before: float32 sweeps on G / max|G| until the largest cosine of a sweep is
// <= tol32, V32 = normalised columns, W = G V32 in float64 (G re-read from L2), float64 sweeps
// (normally one), check sum(lambda^2) = ||G||_F^2, epilogue.
//
// Epilogue (frequent_directions.py:86-93, :58 and the multi-direction dump of
// fastfdwithdump.py:115-131): rank the eigenvalues, cut = lambda[l], sigma^2_r = lambda_r - cut,
// delta += cut/2, coefficient rows T[r] = sqrt(sigma^2_r / lambda_r) * v_r^T for the rotate GEMM,
// and ring slots for every direction with sigma^2 >= threshold.
#pragma once
#include <float.h>
#include "common.cuh"
#include "eig_common.cuh"

namespace mused {

#ifdef MUSED_EIG_PROF
#define ES_PROF_DECL long long pf_t = clock64(), pf_rot = 0, pf_sync = 0, pf_exch = 0, pf_head = 0
#define ES_LAP(acc) do { const long long _n = clock64(); acc += _n - pf_t; pf_t = _n; } while (0)
#else
#define ES_PROF_DECL
#define ES_LAP(acc)
#endif

constexpr int ES_BC = 16;        // columns per block
constexpr int ES_CPC = 32;       // columns per CTA
constexpr int ES_WARPS = 8;
constexpr int ES_NT = ES_WARPS * 32;

template <int EPL>
struct ESLayout {
    static constexpr int N2 = 32 * EPL;
    static constexpr size_t W_BYTES = (size_t)ES_CPC * N2 * 8;    // one float64 copy = two float32 copies
    static constexpr size_t OFF_NRM = W_BYTES;                     // [2][32] float64 slots
    static constexpr size_t OFF_LAM = OFF_NRM + 2 * 32 * 8;        // [N2] eigenvalue per column position
    static constexpr size_t OFF_RED = OFF_LAM + (size_t)N2 * 8;    // [2 parity][2][16 ranks]
    static constexpr size_t OFF_CTA = OFF_RED + 2 * 2 * 16 * 8;    // [2][16 warps]
    static constexpr size_t OFF_RN = OFF_CTA + 2 * 16 * 8;         // [32] float64 column scales
    static constexpr size_t OFF_CUT = OFF_RN + 32 * 8;
    static constexpr size_t BYTES = OFF_CUT + 16;
};

// ---- 128-bit column accesses: lane l owns rows VW*l + 32*VW*j .. +VW-1 (VW = 16 / sizeof(T)); columns shorter
// than 32 vectors fall back to rows l + 32 e ------------------------------------------------------------------
template <class T, int EPL>
struct ColIO {
    static constexpr int VW = 16 / (int)sizeof(T);
    static constexpr bool VEC = (EPL % VW) == 0;
    static constexpr int NV = VEC ? EPL / VW : 0;
    __device__ static __forceinline__ int row(int lane, int e) {
        if constexpr (VEC) return VW * lane + 32 * VW * (e / VW) + (e % VW);
        else return lane + 32 * e;
    }
    __device__ static __forceinline__ void load(const T* col, int lane, T (&x)[EPL]) {
        if constexpr (VEC) {
#pragma unroll
            for (int j = 0; j < NV; ++j) {
                const int4 v = *reinterpret_cast<const int4*>(col + VW * lane + 32 * VW * j);
                memcpy(&x[VW * j], &v, 16);
            }
        } else {
#pragma unroll
            for (int e = 0; e < EPL; ++e) x[e] = col[lane + 32 * e];
        }
    }
    __device__ static __forceinline__ void store(T* col, int lane, const T (&x)[EPL]) {
        if constexpr (VEC) {
#pragma unroll
            for (int j = 0; j < NV; ++j) {
                int4 v;
                memcpy(&v, &x[VW * j], 16);
                *reinterpret_cast<int4*>(col + VW * lane + 32 * VW * j) = v;
            }
        } else {
#pragma unroll
            for (int e = 0; e < EPL; ++e) col[lane + 32 * e] = x[e];
        }
    }
    // the same column at the same place in CTA `rank` of the cluster
    __device__ static __forceinline__ void store_remote(T* col, int rank, int lane, const T (&x)[EPL]) {
        const uint32_t base = dsmem_addr(col, rank);
        if constexpr (VEC) {
#pragma unroll
            for (int j = 0; j < NV; ++j) {
                int4 v;
                memcpy(&v, &x[VW * j], 16);
                asm volatile("st.shared::cluster.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(
                                 base + (uint32_t)((VW * lane + 32 * VW * j) * sizeof(T))),
                             "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
                             : "memory");
            }
        } else {
#pragma unroll
            for (int e = 0; e < EPL; ++e) {
                if constexpr (sizeof(T) == 8) dsmem_st(base + (uint32_t)((lane + 32 * e) * 8), (double)x[e]);
                else asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(base + (uint32_t)((lane + 32 * e) * 4)), "f"((float)x[e]) : "memory");
            }
        }
    }
};

template <class T>
__device__ __forceinline__ T warp_sum_t(T v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

template <class T, int EPL>
__device__ __forceinline__ T col_dot(const T (&x)[EPL], const T (&y)[EPL]) {
    if constexpr (sizeof(T) == 4 && EPL % 2 == 0) {
        unsigned long long acc = 0ull;
#pragma unroll
        for (int e = 0; e < EPL; e += 2) acc = fma2(pk2(x[e], x[e + 1]), pk2(y[e], y[e + 1]), acc);
        float lo, hi;
        upk2(acc, lo, hi);
        return lo + hi;
    } else {
        T g0 = T(0), g1 = T(0);
#pragma unroll
        for (int e = 0; e < EPL; e += 2) {
            g0 = fma(x[e], y[e], g0);
            if (e + 1 < EPL) g1 = fma(x[e + 1], y[e + 1], g1);
        }
        return g0 + g1;
    }
}

// Rotation of a column pair with inner product g and squared norms (nx, ny); returns whether the pair is rotated.
// float32 (the sweeps that only have to converge): two MUFU on the dependent chain, c^2 = (1 + |a| / r) / 2,
// s = sgn(a) b / (2 r c) with a = ny - nx, b = 2 g, r = sqrt(a^2 + b^2) -- the small-angle root.
// float64: the tangent from 20-bit MUFU seeds, then (c, s) refined so that c^2 + s^2 = 1 to 1e-16 (W stays G times
// an orthogonal matrix).  Norms follow the rotation actually applied.
template <class T>
__device__ __forceinline__ bool es_angle(T g, T nx, T ny, T nmax, T& off, T& c, T& s, T& mx, T& my) {
    constexpr bool F64 = sizeof(T) == 8;
    const T prod = nx * ny;
    const T rel = fabs(g) * rsqrt_approx(prod);
    const bool live = prod > T(0);
    if (live && prod > (F64 ? T(1e-24) : T(1e-12)) * nmax * nmax) off = fmax(off, rel);
    const bool apply = live && rel > (F64 ? T(1e-10) : T(3e-8));
    const T a = ny - nx, b = g + g;
    const T q2 = fma(a, a, b * b);
    if constexpr (F64) {
        const T r = q2 * rsqrt_approx(q2);
        const T tb = b * rcp_approx(fabs(a) + r);
        const T t = sel(a < T(0), -tb, tb);
        const T u = fma(t, t, T(1));
        T c0 = rsqrt_approx(u);
        c0 = c0 * fma(T(-0.5) * u, c0 * c0, T(1.5));
        c0 = c0 * fma(T(-0.5) * u, c0 * c0, T(1.5));     // 20 -> 40 -> 80 bits
        c = c0;
        s = c0 * t;
    } else {
        const T rinv = rsqrt_approx(q2);
        const T cc = fma(T(0.5) * fabs(a), rinv, T(0.5));
        const T cinv = rsqrt_approx(cc);
        c = cc * cinv;
        const T sb = (T(0.5) * b) * (rinv * cinv);
        s = sel(a < T(0), -sb, sb);
    }
    const T cc2 = c * c, ss = s * s, cs2 = (c + c) * s * g;
    mx = fma(cc2, nx, fma(ss, ny, -cs2));     // |c x - s y|^2
    my = fma(ss, nx, fma(cc2, ny, cs2));      // |s x + c y|^2
    return apply;
}

// Two independent rotations of one warp, (xa, ya) and (xb, yb): the two inner products are reduced TOGETHER
// (after the xor-16 step lanes 0-15 own pair a, lanes 16-31 pair b: 5 shuffles instead of 10), each half
// computes its angle, (c, s, new norms) are broadcast back.  `dirty` reports whether anything moved.
template <class T, int EPL>
__device__ __forceinline__ void es_rot2(T (&xa)[EPL], T (&ya)[EPL], T& nxa, T& nya, T (&xb)[EPL], T (&yb)[EPL], T& nxb,
                                        T& nyb, T nmax, T& off, int lane, bool& dirty) {
    const T ga = col_dot<T, EPL>(xa, ya), gb = col_dot<T, EPL>(xb, yb);
    const bool hi = (lane & 16) != 0;
    T k = sel(hi, gb, ga) + __shfl_xor_sync(0xffffffffu, sel(hi, ga, gb), 16);
    k += __shfl_xor_sync(0xffffffffu, k, 8);
    k += __shfl_xor_sync(0xffffffffu, k, 4);
    k += __shfl_xor_sync(0xffffffffu, k, 2);
    k += __shfl_xor_sync(0xffffffffu, k, 1);
    const T nx = sel(hi, nxb, nxa), ny = sel(hi, nyb, nya);
    T c, s, mx, my;
    const bool apply = es_angle<T>(k, nx, ny, nmax, off, c, s, mx, my);
    if (__ballot_sync(0xffffffffu, apply) == 0u) return;      // warp-uniform
    dirty = true;
    c = sel(apply, c, T(1));
    s = sel(apply, s, T(0));
    mx = sel(apply, mx, nx);
    my = sel(apply, my, ny);
    const T ca = __shfl_sync(0xffffffffu, c, 0), sa = __shfl_sync(0xffffffffu, s, 0);
    const T cb = __shfl_sync(0xffffffffu, c, 16), sb = __shfl_sync(0xffffffffu, s, 16);
    nxa = __shfl_sync(0xffffffffu, mx, 0);
    nya = __shfl_sync(0xffffffffu, my, 0);
    nxb = __shfl_sync(0xffffffffu, mx, 16);
    nyb = __shfl_sync(0xffffffffu, my, 16);
    auto rot = [&](T (&x)[EPL], T (&y)[EPL], T cc, T ss) {
        if constexpr (sizeof(T) == 4 && EPL % 2 == 0) {
            const unsigned long long c2 = pk2(cc, cc), s2 = pk2(ss, ss), ns2 = pk2(-ss, -ss);
#pragma unroll
            for (int e = 0; e < EPL; e += 2) {
                const unsigned long long X = pk2(x[e], x[e + 1]);
                const unsigned long long Y = pk2(y[e], y[e + 1]);
                upk2(fma2(c2, X, mul2(ns2, Y)), x[e], x[e + 1]);
                upk2(fma2(s2, X, mul2(c2, Y)), y[e], y[e + 1]);
            }
        } else {
#pragma unroll
            for (int e = 0; e < EPL; ++e) {
                const T xv = x[e], yv = y[e];
                x[e] = fma(cc, xv, -ss * yv);
                y[e] = fma(ss, xv, cc * yv);
            }
        }
    };
    rot(xa, ya, ca, sa);
    rot(xb, yb, cb, sb);
}

// all four rotations between the column pair (x0, x1) and the column pair (y0, y1)
template <class T, int EPL>
__device__ __forceinline__ void es_cross2(T (&x0)[EPL], T (&x1)[EPL], T (&nx)[2], T (&y0)[EPL], T (&y1)[EPL], T (&ny)[2],
                                          T nmax, T& off, int lane, bool& dirty) {
    es_rot2<T, EPL>(x0, y0, nx[0], ny[0], x1, y1, nx[1], ny[1], nmax, off, lane, dirty);
    es_rot2<T, EPL>(x0, y1, nx[0], ny[1], x1, y0, nx[1], ny[0], nmax, off, lane, dirty);
}

template <class T>
__device__ __forceinline__ void ld_norm2(const T* nrm, int c, T (&n)[2]) {   // c even
    if constexpr (sizeof(T) == 4) {
        const float2 v = *reinterpret_cast<const float2*>(nrm + c);
        n[0] = v.x; n[1] = v.y;
    } else {
        const double2 v = *reinterpret_cast<const double2*>(nrm + c);
        n[0] = v.x; n[1] = v.y;
    }
}
template <class T>
__device__ __forceinline__ void st_norm2(T* nrm, int c, const T (&n)[2]) {
    if constexpr (sizeof(T) == 4) *reinterpret_cast<float2*>(nrm + c) = make_float2(n[0], n[1]);
    else *reinterpret_cast<double2*>(nrm + c) = make_double2(n[0], n[1]);
}

// Cluster-wide reduction of two values per thread: a by max, b by max or sum.  One cluster barrier; the
// receive arrays alternate (par) so that a CTA still reading the previous result is never overwritten.
template <bool BSUM>
__device__ __forceinline__ void es_cluster_red2(double* cta, double* red, int NC, int rank, uint32_t& par, double& a,
                                                double& b) {
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    a = warp_max(a);
    b = BSUM ? warp_sum(b) : warp_max(b);
    if (lane == 0) {
        cta[w] = a;
        cta[16 + w] = b;
    }
    __syncthreads();
    if (w == 0) {
        double x = lane < ES_WARPS ? cta[lane] : -DBL_MAX;
        double y = lane < ES_WARPS ? cta[16 + lane] : (BSUM ? 0.0 : -DBL_MAX);
        x = warp_max(x);
        y = BSUM ? warp_sum(y) : warp_max(y);
        if (lane < NC) {
            const uint32_t dst = dsmem_addr(red + par * 32, lane);
            dsmem_st(dst + (uint32_t)(rank * 8), x);
            dsmem_st(dst + (uint32_t)((16 + rank) * 8), y);
        }
    }
    cluster_barrier();
    const double* r = red + par * 32;
    a = -DBL_MAX;
    b = BSUM ? 0.0 : -DBL_MAX;
    for (int i = 0; i < NC; ++i) {
        a = fmax(a, r[i]);
        b = BSUM ? b + r[16 + i] : fmax(b, r[16 + i]);
    }
    par ^= 1u;
}

// Blocks change places after a round (round-robin tournament on the 2 NC block slots: slot "first of CTA 0"
// is fixed, the others move one place along the cycle first[1] -> .. -> first[NC-1] -> second[NC-1] -> .. ->
// second[0] -> first[1]).  Warp w carries columns 2w, 2w+1 of both blocks: the first block's straight from its
// registers, the second block's from shared memory.  DB: the blocks are written into the OTHER copy of the
// destination (one cluster barrier); otherwise everybody loads, barrier, stores, barrier.
template <class T, int EPL, bool DB>
__device__ __forceinline__ void es_exchange(const T (&xt0)[EPL], const T (&xt1)[EPL], const T (&nt)[2], T* Wsrc, T* nsrc,
                                            T* Wdst, T* ndst, int NC, int rank, int w, int lane) {
    constexpr int N2 = 32 * EPL;
    using IO = ColIO<T, EPL>;
    int fr, fs, sr, ss;                       // destination (CTA, slot) of my first / second block
    if (rank == 0) { fr = 0; fs = 0; sr = 1; ss = 0; }
    else {
        if (rank == NC - 1) { fr = rank; fs = 1; } else { fr = rank + 1; fs = 0; }
        sr = rank - 1; ss = 1;
    }
    T xs0[EPL], xs1[EPL], ns[2];
    IO::load(Wsrc + (ES_BC + 2 * w) * N2, lane, xs0);
    IO::load(Wsrc + (ES_BC + 2 * w + 1) * N2, lane, xs1);
    ld_norm2<T>(nsrc, ES_BC + 2 * w, ns);
    if (!DB) cluster_barrier();
    IO::store_remote(Wdst + (fs * ES_BC + 2 * w) * N2, fr, lane, xt0);
    IO::store_remote(Wdst + (fs * ES_BC + 2 * w + 1) * N2, fr, lane, xt1);
    IO::store_remote(Wdst + (ss * ES_BC + 2 * w) * N2, sr, lane, xs0);
    IO::store_remote(Wdst + (ss * ES_BC + 2 * w + 1) * N2, sr, lane, xs1);
    if (lane < 4) {
        const bool first = lane < 2;
        const T v = first ? nt[lane & 1] : ns[lane & 1];
        const uint32_t a = dsmem_addr(ndst + (first ? fs : ss) * ES_BC + 2 * w + (lane & 1), first ? fr : sr);
        if constexpr (sizeof(T) == 8) dsmem_st(a, (double)v);
        else asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(a), "f"((float)v) : "memory");
    }
    cluster_barrier();
}

// The sweeps of one phase.  Returns the sweeps done; `last` = the largest cosine the last started sweep saw at
// its entry (measured during the sweep before).  DB: two copies of W / norms in shared memory, `cur` says which.
template <class T, int EPL, bool DB>
__device__ __forceinline__ int es_sweeps(unsigned char* smem, int& cur, double tol, int max_sweeps, int NC, int rank,
                                         uint32_t& par, double& last) {
    constexpr int N2 = 32 * EPL;
    using L = ESLayout<EPL>;
    using IO = ColIO<T, EPL>;
    T* const Wbase = reinterpret_cast<T*>(smem);
    T* const nbase = reinterpret_cast<T*>(smem + L::OFF_NRM);
    double* const cta = reinterpret_cast<double*>(smem + L::OFF_CTA);
    double* const red = reinterpret_cast<double*>(smem + L::OFF_RED);
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int NR = NC > 1 ? 2 * NC - 1 : 1;
    T off = T(0);
    int done = 0;
    ES_PROF_DECL;
    for (int sweep = 0; sweep < max_sweeps; ++sweep) {
        ES_LAP(pf_head);
        T* W = Wbase + (DB ? cur * ES_CPC * N2 : 0);
        T* nrm = nbase + (DB ? cur * 64 : 0);
        T x0[EPL], x1[EPL], y0[EPL], y1[EPL], nx[2], ny[2];
        // exact norms of columns 4w .. 4w+3, cluster-wide largest norm and last sweep's largest cosine; the pairs
        // (4w, 4w+1) and (4w+2, 4w+3) are rotated on the way
        IO::load(W + (4 * w + 0) * N2, lane, x0);
        IO::load(W + (4 * w + 1) * N2, lane, x1);
        IO::load(W + (4 * w + 2) * N2, lane, y0);
        IO::load(W + (4 * w + 3) * N2, lane, y1);
        nx[0] = warp_sum_t(col_dot<T, EPL>(x0, x0));
        nx[1] = warp_sum_t(col_dot<T, EPL>(x1, x1));
        ny[0] = warp_sum_t(col_dot<T, EPL>(y0, y0));
        ny[1] = warp_sum_t(col_dot<T, EPL>(y1, y1));
        double a = (double)fmax(fmax(nx[0], nx[1]), fmax(ny[0], ny[1])), b = (double)off;
        es_cluster_red2<false>(cta, red, NC, rank, par, a, b);
        last = b;
        if (sweep > 0 && b <= tol) break;
        const T nmax = (T)a;
        off = T(0);
        bool dirty = false;
        es_rot2<T, EPL>(x0, x1, nx[0], nx[1], y0, y1, ny[0], ny[1], nmax, off, lane, dirty);
        IO::store(W + (4 * w + 0) * N2, lane, x0);
        IO::store(W + (4 * w + 1) * N2, lane, x1);
        IO::store(W + (4 * w + 2) * N2, lane, y0);
        IO::store(W + (4 * w + 3) * N2, lane, y1);
        st_norm2<T>(nrm, 4 * w, nx);
        st_norm2<T>(nrm, 4 * w + 2, ny);
        __syncthreads();
        // pairs inside each block: round-robin on its 8 column pairs, warps 0-3 the first block, 4-7 the second
        {
            const int base = (w < 4) ? 0 : ES_BC, i = w & 3;
#pragma unroll 1
            for (int t = 0; t < 7; ++t) {
                int ua, ub;
                if (i == 0) { ua = 7; ub = t; }
                else { ua = (t + i) % 7; ub = (t + 7 - i) % 7; }
                const int ca = base + 2 * ua, cb = base + 2 * ub;
                IO::load(W + ca * N2, lane, x0);
                IO::load(W + (ca + 1) * N2, lane, x1);
                IO::load(W + cb * N2, lane, y0);
                IO::load(W + (cb + 1) * N2, lane, y1);
                ld_norm2<T>(nrm, ca, nx);
                ld_norm2<T>(nrm, cb, ny);
                dirty = false;
                es_cross2<T, EPL>(x0, x1, nx, y0, y1, ny, nmax, off, lane, dirty);
                if (dirty) {
                    IO::store(W + ca * N2, lane, x0);
                    IO::store(W + (ca + 1) * N2, lane, x1);
                    IO::store(W + cb * N2, lane, y0);
                    IO::store(W + (cb + 1) * N2, lane, y1);
                    if (lane == 0) {
                        st_norm2<T>(nrm, ca, nx);
                        st_norm2<T>(nrm, cb, ny);
                    }
                }
                __syncthreads();
            }
        }
#pragma unroll 1
        for (int round = 0; round < NR; ++round) {
            // the warp keeps columns 2w, 2w+1 of the first block in registers for the whole round; the column
            // pairs of the second block pass by through shared memory
            ES_LAP(pf_head);
            IO::load(W + (2 * w) * N2, lane, x0);
            IO::load(W + (2 * w + 1) * N2, lane, x1);
            ld_norm2<T>(nrm, 2 * w, nx);
#pragma unroll 1
            for (int m = 0; m < 8; ++m) {
                const int cb = ES_BC + 2 * ((w + m) & 7);
                IO::load(W + cb * N2, lane, y0);
                IO::load(W + (cb + 1) * N2, lane, y1);
                ld_norm2<T>(nrm, cb, ny);
                dirty = false;
                es_cross2<T, EPL>(x0, x1, nx, y0, y1, ny, nmax, off, lane, dirty);
                if (dirty) {
                    IO::store(W + cb * N2, lane, y0);
                    IO::store(W + (cb + 1) * N2, lane, y1);
                    if (lane == 0) st_norm2<T>(nrm, cb, ny);
                }
                ES_LAP(pf_rot);
                __syncthreads();
                ES_LAP(pf_sync);
            }
            if (NC > 1) {
                if constexpr (DB) {
                    T* W2 = Wbase + (cur ^ 1) * ES_CPC * N2;
                    T* n2 = nbase + (cur ^ 1) * 64;
                    es_exchange<T, EPL, true>(x0, x1, nx, W, nrm, W2, n2, NC, rank, w, lane);
                    cur ^= 1;
                    W = W2;
                    nrm = n2;
                } else {
                    es_exchange<T, EPL, false>(x0, x1, nx, W, nrm, W, nrm, NC, rank, w, lane);
                }
            } else {
                IO::store(W + (2 * w) * N2, lane, x0);
                IO::store(W + (2 * w + 1) * N2, lane, x1);
                if (lane == 0) st_norm2<T>(nrm, 2 * w, nx);
                __syncthreads();
            }
            ES_LAP(pf_exch);
        }
        done = sweep + 1;
    }
#ifdef MUSED_EIG_PROF
    if (blockIdx.x == 17 && w == 3 && lane == 0 && done > 0)
        printf("K3 prof %s: %d sweeps; cycles per sweep: head+inblock %lld, cross rot %lld, syncthreads %lld, exchange %lld\n",
               sizeof(T) == 4 ? "f32" : "f64", done, pf_head / done, pf_rot / done, pf_sync / done, pf_exch / done);
#endif
    return done;
}

// Grid = S * EPL CTAs, cluster = EPL CTAs (n2 / 32), ES_NT threads, dynamic smem = ESLayout<EPL>::BYTES.
template <int EPL>
__global__ void __launch_bounds__(ES_NT, (EPL <= 8 ? 2 : 1)) k_eig_jacobi(EigArgs A) {
    constexpr int N2 = 32 * EPL;
    constexpr int ELL = 16 * EPL;
    constexpr int NC = EPL;
    constexpr int CW = ES_CPC / ES_WARPS;      // columns a warp looks after outside the sweeps (4w .. 4w+3)
    using L = ESLayout<EPL>;
    using IOD = ColIO<double, EPL>;
    using IOF = ColIO<float, EPL>;
    extern __shared__ __align__(16) unsigned char smem[];
    double* const Wd = reinterpret_cast<double*>(smem);
    float* const Wf = reinterpret_cast<float*>(smem);
    double* const lam = reinterpret_cast<double*>(smem + L::OFF_LAM);
    double* const red = reinterpret_cast<double*>(smem + L::OFF_RED);
    double* const cta = reinterpret_cast<double*>(smem + L::OFF_CTA);
    double* const rn = reinterpret_cast<double*>(smem + L::OFF_RN);
    double* const cutp = reinterpret_cast<double*>(smem + L::OFF_CUT);
    uint32_t rank_u;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank_u));
    const int rank = (int)rank_u;
    const int s = blockIdx.x / NC;
    const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
    BSlot& sl = A.slots[s];
    const int npend = sl.npend;
    if (npend == 0) {                          // nothing to shrink for this instance
        if (rank == 0 && tid == 0) sl.skip = 1;
        return;
    }
#ifdef MUSED_EIG_PROF
    long long kp[6];
    kp[0] = clock64();
#define ES_KSTAMP(i) kp[i] = clock64()
#else
#define ES_KSTAMP(i)
#endif
    const int q_head = sl.q_head, q_len = sl.q_len;
    const unsigned long long last_dump = sl.last_dump;
    const double thr = sl.error;
    uint32_t par = 0;

    // ---- load: my 32 columns of the symmetrised, masked Gram -> Gs (global, float64) and W32 = Gs / max|Gs| ----
    const unsigned char* mk = A.mask + (size_t)s * ELL;
    auto dead = [&](int i) -> bool {
        if (i < ELL) return false;
        const int r = i - ELL;
        return r >= A.n_rows || mk[r] != 0;
    };
    const double* G = A.G + (size_t)s * N2 * N2;
    double* Gs = A.Gs + (size_t)s * N2 * N2;
    double gmax = 0.0, fro = 0.0;
#pragma unroll 1
    for (int h = 0; h < CW; ++h) {
        const int j = rank * ES_CPC + CW * w + h;
        const bool dj = dead(j);
#pragma unroll
        for (int e = 0; e < EPL; ++e) {
            const int i = lane + 32 * e;
            double v = 0.0;
            if (!dj && !dead(i)) v = 0.5 * (G[(size_t)j * N2 + i] + G[(size_t)i * N2 + j]);
            Gs[(size_t)j * N2 + i] = v;
            gmax = fmax(gmax, fabs(v));
            fro = fma(v, v, fro);
        }
    }
    es_cluster_red2<true>(cta, red, NC, rank, par, gmax, fro);     // fro = ||Gs||_F^2 = sum of squared eigenvalues

    int sweeps32 = 0, cur = 0;
    double off = 0.0;
    bool from_gram = !A.mixed;
    if (A.mixed) {
        // ---- float32 phase -------------------------------------------------------------------------------
        const double inv = gmax > 0.0 ? 1.0 / gmax : 0.0;
#pragma unroll 1
        for (int h = 0; h < CW; ++h) {
            const int c = CW * w + h;
#pragma unroll
            for (int e = 0; e < EPL; ++e)
                Wf[c * N2 + lane + 32 * e] = (float)(Gs[(size_t)(rank * ES_CPC + c) * N2 + lane + 32 * e] * inv);
        }
        __syncthreads();
        ES_KSTAMP(1);
        sweeps32 = es_sweeps<float, EPL, true>(smem, cur, A.tol32, A.max_sweeps, NC, rank, par, off);
        ES_KSTAMP(2);
        // ---- is every column's direction trustworthy?  (see the note on rank-deficient Grams in DESIGN.md 5) --
        float* Wc = Wf + cur * ES_CPC * N2;
        double tmax = 0.0, tmin = DBL_MAX;
#pragma unroll 1
        for (int h = 0; h < CW; ++h) {
            float x[EPL];
            IOF::load(Wc + (CW * w + h) * N2, lane, x);
            double t = 0.0;
#pragma unroll
            for (int e = 0; e < EPL; ++e) t = fma((double)x[e], (double)x[e], t);
            t = warp_sum(t);
            if (lane == 0) rn[CW * w + h] = t > 0.0 ? rsqrt(t) : 0.0;
            tmax = fmax(tmax, t);
            if (t > 0.0) tmin = fmin(tmin, t);
        }
        double ntmin = -tmin;
        es_cluster_red2<false>(cta, red, NC, rank, par, tmax, ntmin);
        tmin = -ntmin;
        if (tmin < A.small_col * tmax) {
            sweeps32 += 50;                     // flag in BSlot::sweeps: numerically rank-deficient
            from_gram = true;
        } else {
            // ---- W = Gs V32 in float64: thread = RPT rows x CPT columns, V32 = rn[c] * W32[c] ---------
            constexpr int RS = N2 < 64 ? N2 : 64;
            constexpr int RPT = N2 / RS;
            constexpr int CPT = ES_CPC / (ES_NT / RS);
            const int r0 = tid % RS, c0 = (tid / RS) * CPT;
            double acc[RPT][CPT];
#pragma unroll
            for (int r = 0; r < RPT; ++r)
#pragma unroll
                for (int c = 0; c < CPT; ++c) acc[r][c] = 0.0;
#pragma unroll 2
            for (int k = 0; k < N2; ++k) {
                double gk[RPT];
#pragma unroll
                for (int r = 0; r < RPT; ++r) gk[r] = __ldcg(Gs + (size_t)k * N2 + r0 + RS * r);
#pragma unroll
                for (int c = 0; c < CPT; ++c) {
                    const double v = (double)Wc[(c0 + c) * N2 + k];
#pragma unroll
                    for (int r = 0; r < RPT; ++r) acc[r][c] = fma(gk[r], v, acc[r][c]);
                }
            }
            __syncthreads();                    // everyone has read W32 before W (float64) takes its place
#pragma unroll
            for (int c = 0; c < CPT; ++c) {
                const double sc = rn[c0 + c];
#pragma unroll
                for (int r = 0; r < RPT; ++r) Wd[(c0 + c) * N2 + r0 + RS * r] = acc[r][c] * sc;
            }
            __syncthreads();
        }
    }

    // ---- float64 sweeps; eigenvalues per column position, replicated into every CTA -------------------------
    // After a mixed-precision solve the result is CHECKED: sum of lambda^2 must equal ||G||_F^2 (1e-6 relative; a
    // restart basis V32 that was not orthonormal fails it by orders of magnitude); a failed check repeats the
    // float64 sweeps from G itself.
    ES_KSTAMP(3);
    int sweeps_done = 0;
    double lamv[CW];
#pragma unroll 1
    for (int attempt = 0; attempt < 2; ++attempt) {
        if (from_gram) {
            __syncthreads();
#pragma unroll 1
            for (int h = 0; h < CW; ++h)
#pragma unroll
                for (int e = 0; e < EPL; ++e)
                    Wd[(CW * w + h) * N2 + lane + 32 * e] =
                        __ldcg(Gs + (size_t)(rank * ES_CPC + CW * w + h) * N2 + lane + 32 * e);
            __syncthreads();
        }
        int cur64 = 0;
        sweeps_done += es_sweeps<double, EPL, false>(smem, cur64, A.tol, A.max_sweeps, NC, rank, par, off);
#pragma unroll
        for (int h = 0; h < CW; ++h) {
            double x[EPL];
            IOD::load(Wd + (CW * w + h) * N2, lane, x);
            lamv[h] = sqrt(warp_sum(col_dot<double, EPL>(x, x)));
        }
        if (lane < NC) {
            const uint32_t dst = dsmem_addr(lam, lane);
#pragma unroll
            for (int h = 0; h < CW; ++h) dsmem_st(dst + (uint32_t)((rank * ES_CPC + CW * w + h) * 8), lamv[h]);
        }
        cluster_barrier();
        if (from_gram) break;
        double l2 = 0.0;
        for (int i = lane; i < N2; i += 32) l2 = fma(lam[i], lam[i], l2);
        l2 = warp_sum(l2);
        if (fabs(l2 - fro) <= 1e-6 * fro) break;
        from_gram = true;
        sweeps32 += 70;                         // flag in BSlot::sweeps: the check failed
        cluster_barrier();                      // everyone has read lam
    }
    sweeps_done += 100 * sweeps32;
    ES_KSTAMP(4);

    // rank (descending, ties by column position) of my columns
    int rk[CW];
#pragma unroll
    for (int h = 0; h < CW; ++h) rk[h] = 0;
    const int pos0 = rank * ES_CPC + CW * w;
    for (int i = lane; i < N2; i += 32) {
        const double v = lam[i];
#pragma unroll
        for (int h = 0; h < CW; ++h) rk[h] += (v > lamv[h]) || (v == lamv[h] && i < pos0 + h);
    }
#pragma unroll
    for (int h = 0; h < CW; ++h) rk[h] = __reduce_add_sync(0xffffffffu, rk[h]);
    if (lane < NC) {
        const uint32_t dst = dsmem_addr(cutp, lane);
#pragma unroll
        for (int h = 0; h < CW; ++h)
            if (rk[h] == A.ell_l) dsmem_st(dst, lamv[h]);
    }
    cluster_barrier();
    const double cut = *cutp;
    // Directions below the float64 round-off of the largest eigenvalue have no usable eigenvector:
    // their rows are zero, as LAPACK's null singular vectors times sqrt(~0) are.
    double lam_max = 0.0;
    for (int i = lane; i < N2; i += 32) lam_max = fmax(lam_max, lam[i]);
    lam_max = warp_max(lam_max);
    const double lam_floor = 1e-10 * lam_max;

    // ---- outputs ----------------------------------------------------------------------------------------
    const int src_plane = sl.cur;
    float* Th = A.Thi + (size_t)s * ELL * N2;
    float* Tl = A.Tlo + (size_t)s * ELL * N2;
#pragma unroll
    for (int h = 0; h < CW; ++h) {
        const int r = rk[h];
        if (r >= ELL) continue;
        const double lm = lamv[h];
        const double s2v = (r < A.ell_l && lm > lam_floor) ? fmax(lm - cut, 0.0) : 0.0;
        const double coef = (lm > 0.0 && s2v > 0.0) ? sqrt(s2v / lm) / lm : 0.0;
        const double* col = Wd + (CW * w + h) * N2;
#pragma unroll
        for (int e = 0; e < EPL; ++e) {
            const float v = (float)(coef * col[lane + 32 * e]);
            uint32_t hb, lb;
            asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hb) : "f"(v));
            const float hi = __uint_as_float(hb);
            asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(lb) : "f"(v - hi));
            Th[(size_t)r * N2 + lane + 32 * e] = hi;
            Tl[(size_t)r * N2 + lane + 32 * e] = __uint_as_float(lb);
        }
        if (lane == 0) {
            const bool dumped = s2v >= thr;
            int slot = -1;
            if (dumped) {
                slot = (q_head + q_len + r) % A.capn;
                const unsigned long long s0 = last_dump < A.t_end ? last_dump : A.t_end;
                A.ringS[(size_t)s * A.capn + slot] = s0;
                A.ringT[(size_t)s * A.capn + slot] = A.t_end;
            }
            A.dump_slot[(size_t)s * ELL + r] = slot;
            A.s2[(size_t)s * ELL + r] = dumped ? 0.0 : s2v;
        }
    }
    if (rank == 0 && w == 0) {
        int nd = 0;
        for (int i = lane; i < N2; i += 32) nd += (fmax(lam[i] - cut, 0.0) >= thr) ? 1 : 0;
        nd = __reduce_add_sync(0xffffffffu, nd);
        if (lane == 0) {
            sl.delta += cut / 2.0;
            sl.q_len = q_len + nd;
            if (nd > 0) sl.last_dump = A.t_end + 1ull;
            sl.n_flush += 1;
            sl.n_dump += (unsigned long long)nd;
            sl.nd = nd;
            sl.src = src_plane;
            sl.cur = src_plane ^ 1;
            sl.skip = 0;
            sl.sweeps = sweeps_done;
        }
    }
    cluster_barrier();                          // no CTA leaves while its shared memory may still be written
#ifdef MUSED_EIG_PROF
    ES_KSTAMP(5);
    if (blockIdx.x == 17 && w == 3 && lane == 0)
        printf("K3 prof kernel: prologue %lld, float32 sweeps %lld, W = G V32 %lld, float64 sweeps + check %lld, epilogue %lld cycles\n",
               kp[1] - kp[0], kp[2] - kp[1], kp[3] - kp[2], kp[4] - kp[3], kp[5] - kp[4]);
#endif
}

}  // namespace mused

// Symmetric eigensolve of the query Gram (SlidingWindowFD.get,
// sliding_windows_frequent_directions.py:160: the thin SVD of [C; snapshots],
// up to 11 l rows) on the whole GPU, float64.
//
// One-sided (Hestenes) BLOCK Jacobi on the columns of W = G: W V -> V Lambda, so
// column norms converge to the eigenvalues and w_j / lambda_j to the eigenvectors.
// The matrix lives in SHARED MEMORY, distributed over P co-resident CTAs: CTA p
// keeps the two column blocks at positions (2p, 2p+1) of the odd-even
// transposition ordering (CB columns each, all n rows).  One round:
//   Gram of the 2 CB resident columns (thread per row, register accumulators,
//   butterfly reduce-scatter) -> cyclic two-sided Jacobi of that (2CB x 2CB) Gram
//   by one warp, accumulating the rotations in Q -> W <- W Q (thread per row).
// Blocks trade places after every round; a block that crosses a CTA boundary
// goes through an L2 mailbox, and the only synchronisation is a release/acquire
// flag between NEIGHBOURS (no grid barrier inside a sweep; one per sweep for the
// convergence test).  Launched cooperatively so that all P CTAs are resident.
#pragma once
#include "common.cuh"

namespace mused {
namespace big {
namespace cg = cooperative_groups;

constexpr int NT = 256;
constexpr int NW = NT / 32;
constexpr unsigned SPIN_LIMIT = 1u << 22;   // ~seconds; a lost neighbour poisons the output instead of hanging

struct Args {
    const double* G;                 // n x n symmetric (leading dimension lda)
    int lda, n, nr, P;               // nr = n rounded up to 32 (rows kept per column)
    double *mailL, *mailR;           // [P][CB * nr] blocks on their way to the left / right neighbour
    unsigned long long *flagL, *flagR;   // [P] iteration stamps, zero before the launch
    unsigned long long* red;         // [128]: max cosine per sweep | max norm^2 per sweep (bits), zero before the launch
    double* lam;                     // [2 P CB] eigenvalue of every column slot (0 for padding)
    double* VT;                      // [2 P CB][ldt] eigenvectors, one per ROW
    int ldt;
    double tol;
    int max_sweeps;
    int* info;                       // [0] sweeps done, [1] != 0: a neighbour wait timed out
    unsigned long long* prof;        // [8] clock cycles of CTA P/2 per phase: gram, small, apply, out, wait, in
};

__device__ __forceinline__ double rsqrt_a(double x) {
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    return y;
}
__device__ __forceinline__ double rcp_a(double x) {
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    return y;
}
__device__ __forceinline__ void flag_publish(unsigned long long* f, unsigned long long v) {
    __threadfence();
    asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(f), "l"(v) : "memory");
}
__device__ __forceinline__ bool flag_wait(const unsigned long long* f, unsigned long long v) {
    unsigned long long x;
    unsigned spins = 0;
    do {
        asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(x) : "l"(f) : "memory");
        if (x >= v) return true;
    } while (++spins < SPIN_LIMIT);
    return false;
}

// Rotation of the column pair with inner product g and squared norms (nx, ny):
// x' = c x - s y, y' = s x + c y.  20-bit seeds for the tangent, (c, s) refined to
// c^2 + s^2 = 1 in float64 (same scheme as eig_smem.cuh).
__device__ __forceinline__ void angle(double g, double nx, double ny, double nmax, double& off, double& c,
                                      double& s) {
    const double prod = nx * ny;
    const bool live = prod > 0.0;
    const double rel = live ? fabs(g) * rsqrt_a(prod) : 0.0;
    if (live && prod > 1e-24 * nmax * nmax) off = fmax(off, rel);
    const bool apply = live && rel > 1e-12;
    const double a = ny - nx, b = g + g;
    const double q2 = fma(a, a, b * b);
    const double r = (q2 > 0.0) ? q2 * rsqrt_a(q2) : 0.0;
    const double den = fabs(a) + r;
    const double tb = (den > 0.0) ? b * rcp_a(den) : 0.0;
    const double t = (a < 0.0) ? -tb : tb;
    const double u = fma(t, t, 1.0);
    double c0 = rsqrt_a(u);
    c0 = c0 * fma(-0.5 * u, c0 * c0, 1.5);
    c0 = c0 * fma(-0.5 * u, c0 * c0, 1.5);
    c = apply ? c0 : 1.0;
    s = apply ? c0 * t : 0.0;
}

template <int CB>
struct Cfg {
    static constexpr int M = 2 * CB;
    static constexpr int NACC = M * (M + 1) / 2;
    static constexpr int K = (NACC + 31) / 32 * 32;      // padded for the butterfly
    // doubles after the two column blocks
    static constexpr int SMALL = 2 * M * M + NW * K + 2 * CB + 8;
    static size_t smem_bytes(int nr) { return ((size_t)M * nr + SMALL) * sizeof(double) + 4 * CB * sizeof(int); }
};

// Gram of the M resident columns -> S (M x M, symmetric) in shared memory.
template <int CB>
__device__ __forceinline__ void block_gram(const double* __restrict__ W, int nr, double* S, double* part,
                                           int tid) {
    constexpr int M = Cfg<CB>::M, NACC = Cfg<CB>::NACC, K = Cfg<CB>::K;
    double acc[K];
#pragma unroll
    for (int e = 0; e < K; ++e) acc[e] = 0.0;
    for (int r = tid; r < nr; r += NT) {
        double x[M];
#pragma unroll
        for (int i = 0; i < M; ++i) x[i] = W[i * nr + r];
        int e = 0;
#pragma unroll
        for (int i = 0; i < M; ++i)
#pragma unroll
            for (int j = i; j < M; ++j, ++e) acc[e] = fma(x[i], x[j], acc[e]);
    }
    // butterfly reduce-scatter: K values per lane -> lane L ends with elements (K/32) L + j
    const int lane = tid & 31, warp = tid >> 5;
#pragma unroll
    for (int st = 0; st < 5; ++st) {
        const int o = 16 >> st;
        const int H = K >> (st + 1);
        const bool up = (lane & o) != 0;
#pragma unroll
        for (int j = 0; j < H; ++j) {
            const double keep = up ? acc[j + H] : acc[j];
            const double send = up ? acc[j] : acc[j + H];
            acc[j] = keep + __shfl_xor_sync(0xffffffffu, send, o);
        }
    }
    constexpr int PER = K / 32;
#pragma unroll
    for (int j = 0; j < PER; ++j) part[warp * K + PER * lane + j] = acc[j];
    __syncthreads();
    if (tid < NACC) {
        double v = 0.0;
#pragma unroll
        for (int w = 0; w < NW; ++w) v += part[w * K + tid];
        // element tid of the packed upper triangle -> (i, j)
        int i = 0, e = tid;
        while (e >= M - i) { e -= M - i; ++i; }
        const int j = i + e;
        S[i * M + j] = v;
        S[j * M + i] = v;
    }
    __syncthreads();
}

// pair k of step t: FULL = round-robin over all M columns (M - 1 steps);
// cross = column k of the first block with column (k + t) % CB of the second (CB steps)
template <int CB, bool FULL>
__device__ __forceinline__ constexpr int pair_p(int t, int k) {
    constexpr int M = 2 * CB, m = M - 1;
    if (!FULL) return k;
    const int a = (k == 0) ? m : (t + k) % m;
    const int b = (k == 0) ? t % m : (t - k + m) % m;
    return a < b ? a : b;
}
template <int CB, bool FULL>
__device__ __forceinline__ constexpr int pair_q(int t, int k) {
    constexpr int M = 2 * CB, m = M - 1;
    if (!FULL) return CB + (k + t) % CB;
    const int a = (k == 0) ? m : (t + k) % m;
    const int b = (k == 0) ? t % m : (t - k + m) % m;
    return a < b ? b : a;
}

// Cyclic two-sided Jacobi on the small Gram S (M x M) by ONE warp, in REGISTERS:
// lane i < M holds row i of S and row i of the accumulated rotation Q; every step
// is unrolled so that the pair indices are compile-time register names.  Per step:
// partner diagonals by shuffle -> angles on the lanes p_k -> (c, s) broadcast ->
// column update (local) -> row update (one row exchange with the partner lane).
template <int CB, bool FULL>
__device__ __forceinline__ void small_jacobi(const double* S, double* Q, double nmax, double& off, double& nmx,
                                             int lane) {
    constexpr int M = 2 * CB;
    constexpr int STEPS = FULL ? M - 1 : CB;
    const int li = lane < M ? lane : 0;
    double Sr[M], Qr[M];
#pragma unroll
    for (int j = 0; j < M; ++j) {
        Sr[j] = (lane < M) ? S[li * M + j] : 0.0;
        Qr[j] = (lane == j) ? 1.0 : 0.0;
    }
#pragma unroll
    for (int j = 0; j < M; ++j)
        if (lane == j) nmx = fmax(nmx, Sr[j]);
#pragma unroll
    for (int t = 0; t < STEPS; ++t) {
        double g = 0.0, nx = 0.0, ny = 0.0;
        bool is_p = false;
        int partner = lane;
#pragma unroll
        for (int k = 0; k < CB; ++k) {
            constexpr int dummy = 0;
            (void)dummy;
            const int pk = pair_p<CB, FULL>(t, k), qk = pair_q<CB, FULL>(t, k);
            const double dq = __shfl_sync(0xffffffffu, Sr[qk], qk);
            if (lane == pk) {
                g = Sr[qk];
                nx = Sr[pk];
                ny = dq;
                is_p = true;
                partner = qk;
            }
            if (lane == qk) partner = pk;
        }
        double c, sn, o2 = 0.0;
        angle(g, nx, ny, nmax, o2, c, sn);
        if (is_p) off = fmax(off, o2);
        double myc = 1.0, mys = 0.0;
#pragma unroll
        for (int k = 0; k < CB; ++k) {
            const int pk = pair_p<CB, FULL>(t, k), qk = pair_q<CB, FULL>(t, k);
            const double ck = __shfl_sync(0xffffffffu, c, pk);
            const double sk = __shfl_sync(0xffffffffu, sn, pk);
            if (lane == pk) { myc = ck; mys = -sk; }
            if (lane == qk) { myc = ck; mys = sk; }
            // columns pk, qk of my rows of S and Q
            const double a = Sr[pk], b = Sr[qk];
            Sr[pk] = fma(ck, a, -sk * b);
            Sr[qk] = fma(sk, a, ck * b);
            const double qa = Qr[pk], qb = Qr[qk];
            Qr[pk] = fma(ck, qa, -sk * qb);
            Qr[qk] = fma(sk, qa, ck * qb);
        }
        // rows: lane p gets c S_p - s S_q, lane q gets s S_p + c S_q
#pragma unroll
        for (int j = 0; j < M; ++j) {
            const double other = __shfl_sync(0xffffffffu, Sr[j], partner);
            Sr[j] = fma(myc, Sr[j], mys * other);
        }
    }
    if (lane < M) {
#pragma unroll
        for (int j = 0; j < M; ++j) Q[lane * M + j] = Qr[j];
    }
}

// W <- W Q, thread per row (two rows at a time share the loads of Q).  OUT = 1: the
// first block's new columns go straight to the mailbox `gout` instead of shared
// memory (the block is leaving); OUT = 2: the second block's; OUT = 0: all stay.
template <int CB, int OUT>
__device__ __forceinline__ void block_apply(double* W, int nr, const double* Q, double* __restrict__ gout,
                                            int tid) {
    constexpr int M = 2 * CB;
    for (int r0 = tid; r0 < nr; r0 += 2 * NT) {
        const int r1 = r0 + NT;
        const bool two = r1 < nr;
        const int r1c = two ? r1 : r0;
        double x0[M], x1[M], y0[M], y1[M];
#pragma unroll
        for (int i = 0; i < M; ++i) {
            x0[i] = W[i * nr + r0];
            x1[i] = W[i * nr + r1c];
            y0[i] = 0.0;
            y1[i] = 0.0;
        }
#pragma unroll
        for (int i = 0; i < M; ++i) {
#pragma unroll
            for (int j = 0; j < M; j += 2) {
                const double2 qv = *reinterpret_cast<const double2*>(Q + i * M + j);
                y0[j] = fma(x0[i], qv.x, y0[j]);
                y0[j + 1] = fma(x0[i], qv.y, y0[j + 1]);
                y1[j] = fma(x1[i], qv.x, y1[j]);
                y1[j + 1] = fma(x1[i], qv.y, y1[j + 1]);
            }
        }
#pragma unroll
        for (int i = 0; i < M; ++i) {
            const bool leaves = (OUT == 1 && i < CB) || (OUT == 2 && i >= CB);
            if (leaves) {
                const int c = (OUT == 1) ? i : i - CB;
                __stcg(gout + (size_t)c * nr + r0, y0[i]);
                if (two) __stcg(gout + (size_t)c * nr + r1, y1[i]);
            } else {
                W[i * nr + r0] = y0[i];
                if (two) W[i * nr + r1] = y1[i];
            }
        }
    }
}

__device__ __forceinline__ void copy_out(double* __restrict__ dst, const double* src, int cnt, int tid) {
    double2* d2 = reinterpret_cast<double2*>(dst);
    const double2* s2 = reinterpret_cast<const double2*>(src);
    for (int i = tid; i < cnt / 2; i += NT) __stcg(d2 + i, s2[i]);
}
__device__ __forceinline__ void copy_in(double* dst, const double* __restrict__ src, int cnt, int tid) {
    double2* d2 = reinterpret_cast<double2*>(dst);
    const double2* s2 = reinterpret_cast<const double2*>(src);
    for (int i = tid; i < cnt / 2; i += NT) d2[i] = __ldcg(s2 + i);
}

template <int CB>
__global__ void __launch_bounds__(NT, 1) k_eig_big(Args A) {
    constexpr int M = Cfg<CB>::M, K = Cfg<CB>::K;
    cg::grid_group grid = cg::this_grid();
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int nr = A.nr, P = A.P, n = A.n;
    const int p = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    double* W = reinterpret_cast<double*>(smem_raw);        // [M][nr]: block A = columns 0..CB-1, block B the rest
    double* blkA = W;
    double* blkB = W + (size_t)CB * nr;
    double* S = W + (size_t)M * nr;
    double* Q = S + M * M;
    double* part = Q + M * M;
    const int bcnt = CB * nr;                                // doubles per block

    // ---- load my 2 CB columns of G (column j of the symmetric G = row j) ----------------
    for (int i = 0; i < M; ++i) {
        const int col = (2 * p) * CB + i;
        for (int r = tid; r < nr; r += NT)
            W[i * nr + r] = (col < n && r < n) ? A.G[(size_t)col * A.lda + r] : 0.0;
    }
    __syncthreads();

    double off = 0.0, nmx = 0.0;        // live in warp 0
    long long pc[6] = {0, 0, 0, 0, 0, 0};
    long long t0 = clock64();
#define PROF(k) do { const long long t1_ = clock64(); pc[k] += t1_ - t0; t0 = t1_; } while (0)
    double nmax = 0.0;
    bool lost = false;
    unsigned long long seq = 0;
    int sweeps_done = 0;
    for (int sweep = 0; sweep < A.max_sweeps; ++sweep) {
        off = 0.0;
        nmx = 0.0;
        for (int it = 0; it < P; ++it) {
            ++seq;
            // even round: positions (2p, 2p+1), both resident.  B (afterwards at position 2p)
            // meets position 2p-1 in CTA p-1 next: its new columns go straight to the mailbox
            // (CTA 0 parks it in its own).
            t0 = clock64();
            block_gram<CB>(W, nr, S, part, tid);
            PROF(0);
            if (warp == 0) {
                if (it == 0) small_jacobi<CB, true>(S, Q, nmax, off, nmx, lane);
                else small_jacobi<CB, false>(S, Q, nmax, off, nmx, lane);
            }
            __syncthreads();
            PROF(1);
            if (P == 1) {
                block_apply<CB, 0>(W, nr, Q, nullptr, tid);
                __syncthreads();
                PROF(2);
                continue;
            }
            block_apply<CB, 2>(W, nr, Q, A.mailL + (size_t)p * bcnt, tid);
            __syncthreads();
            PROF(2);
            if (tid == 0 && p > 0) flag_publish(A.flagL + p, seq);
            PROF(3);
            if (p + 1 < P) {
                // odd round: my A (position 2p+1) with the right neighbour's B (position 2p+2)
                if (tid == 0 && !flag_wait(A.flagL + p + 1, seq)) A.info[1] = 1;
                __syncthreads();
                PROF(4);
                copy_in(blkB, A.mailL + (size_t)(p + 1) * bcnt, bcnt, tid);
                __syncthreads();
                PROF(5);
                block_gram<CB>(W, nr, S, part, tid);
                PROF(0);
                if (warp == 0) small_jacobi<CB, false>(S, Q, nmax, off, nmx, lane);
                __syncthreads();
                PROF(1);
                // the neighbour's block stays (position 2p+1), my A moves on to position 2p+2
                block_apply<CB, 1>(W, nr, Q, A.mailR + (size_t)p * bcnt, tid);
                __syncthreads();
                PROF(2);
                if (tid == 0) flag_publish(A.flagR + p, seq);
                PROF(3);
            } else {
                // the last position sits the odd round out: A stays, as the new B
                for (int i = tid; i < bcnt; i += NT) blkB[i] = blkA[i];
                __syncthreads();
            }
            if (p > 0) {
                if (tid == 0 && !flag_wait(A.flagR + p - 1, seq)) A.info[1] = 1;
                __syncthreads();
                PROF(4);
                copy_in(blkA, A.mailR + (size_t)(p - 1) * bcnt, bcnt, tid);
            } else {
                copy_in(blkA, A.mailL, bcnt, tid);           // position 0 sat the odd round out
            }
            __syncthreads();
            PROF(5);
        }
        sweeps_done = sweep + 1;
        // ---- convergence: largest cosine met in this sweep, over the whole grid -----------
        if (warp == 0) {
            off = warp_max(off);
            nmx = warp_max(nmx);
            if (lane == 0) {
                atomicMax(A.red + sweep, (unsigned long long)__double_as_longlong(off));
                atomicMax(A.red + 64 + sweep, (unsigned long long)__double_as_longlong(nmx));
            }
        }
        grid.sync();
        const double o = __longlong_as_double((long long)__ldcg(A.red + sweep));
        nmax = __longlong_as_double((long long)__ldcg(A.red + 64 + sweep));
        if (o <= A.tol || !(o == o)) break;
    }
    (void)lost;

    // ---- eigenvalues = column norms, eigenvectors = normalised columns -------------------
    block_gram<CB>(W, nr, S, part, tid);
    const bool bad = __ldcg(A.info + 1) != 0;
    // A column whose norm sank to round-off level (lambda <= 1e-10 lambda_max) is noise, not a
    // direction: its eigenvalue is kept, its eigenvector is returned as a ZERO row so that a caller
    // that projects onto it (vt = M^T v / sqrt(lambda)) gets nothing instead of amplified noise.
    const double lam_floor = 1e-10 * sqrt(fmax(nmax, 0.0));
    for (int i = 0; i < M; ++i) {
        const double lam = sqrt(fmax(S[i * M + i], 0.0));
        const int slot = 2 * p * CB + i;
        const double inv = (lam > lam_floor && lam > 0.0) ? 1.0 / lam : 0.0;
        if (tid == 0) A.lam[slot] = bad ? __longlong_as_double(0x7ff8000000000000ll) : lam;
        for (int r = tid; r < nr; r += NT) A.VT[(size_t)slot * A.ldt + r] = W[i * nr + r] * inv;
    }
    if (p == 0 && tid == 0) A.info[0] = sweeps_done;
    if (p == P / 2 && tid == 0)
        for (int k = 0; k < 6; ++k) A.prof[k] = (unsigned long long)pc[k];
#undef PROF
}

}  // namespace big
}  // namespace mused

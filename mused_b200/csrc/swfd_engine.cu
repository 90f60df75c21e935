// Host side of the float64 SWFD engine + the query (get) kernels.
//   mused_swfd_*  <->  swfd/seq_based_swfd.py:11-103 (SeqBasedSWFD)
//   mused_fd_*    <->  swfd/frequent_directions.py:79-153
// See swfd_device.cuh for the per-row device algorithm.
#include <math.h>
#include <stdarg.h>
#include <vector>

#include "swfd_device.cuh"
#include "bswfd.h"
#include "eig_big.cuh"

namespace mused {

static thread_local std::string g_err;
static unsigned long long g_launches = 0;
#define COUNT_LAUNCH(n) (mused::g_launches += (n))
void set_error(const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
}

// ---------------------------------------------------------------------------
// K1 staging: rows of any accepted dtype -> float64 staging rows + ||x||^2.
// One CTA per row, fixed-order reduction (the routing rule X@X.T >= error,
// seq_based_swfd.py:77, and the energy rule, frequent_directions.py:116, read it).
// ---------------------------------------------------------------------------
template <class Tin>
__global__ void __launch_bounds__(256) k_stage_rows(const Tin* __restrict__ in, long long ld,
                                                    long long d, double* __restrict__ out,
                                                    double* __restrict__ norm2) {
    __shared__ double red[8];
    const long long r = blockIdx.x;
    const Tin* src = in + (size_t)r * ld;
    double* dst = out + (size_t)r * d;
    double acc = 0.0;
    for (long long c = threadIdx.x; c < d; c += 256) {
        const double v = (double)src[c];
        dst[c] = v;
        acc = fma(v, v, acc);
    }
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int i = 0; i < 8; ++i) t += red[i];
        norm2[r] = t;
    }
}

// ---------------------------------------------------------------------------
// Query kernels (SlidingWindowFD.get, sliding_windows_frequent_directions.py:154-173):
// stack [C; live snapshots] -> Gram -> symmetric eigensolve -> shrink by the
// (l+1)-th eigenvalue -> l x d sketch and Vt.
// ---------------------------------------------------------------------------
__global__ void k_gather_stack(Dev D, int s, double* __restrict__ M) {
    const Slot& sl = D.slots[s];
    const int ell = D.ell;
    const int rows = ell + sl.q_len;
    const double* R = D.Rp(s, sl.rcur);
    for (int i = blockIdx.x; i < rows; i += gridDim.x) {
        const double* src = (i < ell) ? R + (size_t)i * D.d
                                      : D.ring_row(s, (sl.q_head + (i - ell)) % D.capn);
        double* dst = M + (size_t)i * D.d;
        for (long long c = threadIdx.x; c < D.d; c += blockDim.x) dst[c] = src[c];
    }
}

// G = M M^T (ng x ng, leading dimension lda), float64, one 64x64 tile per CTA.
__global__ void __launch_bounds__(256) k_gram_f64(const double* __restrict__ M, int ng, long long d,
                                                  double* __restrict__ G, int lda) {
    __shared__ GemmSmem sm;
    const int tiles = (ng + GT - 1) / GT;
    for (int tile = blockIdx.x; tile < tiles * tiles; tile += gridDim.x) {
        const int ti = tile / tiles, tj = tile % tiles;
        if (tj > ti) continue;  // lower triangle, mirrored on store
        const int m0 = ti * GT, n0 = tj * GT;
        gemm_tile_f64<true, false>(
            sm, threadIdx.x, m0, n0, (int)d,
            [&](int i, int c) -> double { return (i < ng && c < d) ? M[(size_t)i * d + c] : 0.0; },
            [&](int c, int j) -> double { return (j < ng && c < d) ? M[(size_t)j * d + c] : 0.0; },
            [&](int i, int j, double v) {
                if (i < ng && j < ng) {
                    G[(size_t)i * lda + j] = v;
                    G[(size_t)j * lda + i] = v;
                }
            },
            [&]() { __syncthreads(); });
    }
}

// round-robin pairing: n even, round r in [0, n-1), pair k in [0, n/2)
__device__ __forceinline__ void rr_pair(int n, int r, int k, int& p, int& q) {
    const int m = n - 1;
    int a, b;
    if (k == 0) { a = m; b = r % m; }
    else { a = (r + k) % m; b = (r - k + m) % m; }
    p = min(a, b);
    q = max(a, b);
}

// Two-sided parallel cyclic Jacobi on the symmetric A (n x n, n even), V = I at
// start; on exit diag(A) holds the eigenvalues and column i of V the vectors.
// Cooperative launch (grid-wide sync between the rotation and update phases).
__global__ void __launch_bounds__(256) k_jacobi(double* A, double* V, int n, int lda, double* cs,
                                                int max_sweeps, int* counters, double* scratch,
                                                double rel_thr) {
    cg::grid_group grid = cg::this_grid();
    const int gtid = blockIdx.x * blockDim.x + threadIdx.x;
    const int gth = gridDim.x * blockDim.x;
    const int half = n / 2;
    for (size_t e = gtid; e < (size_t)n * n; e += gth) {
        const int r = (int)(e / n), c = (int)(e % n);
        V[(size_t)r * lda + c] = (r == c) ? 1.0 : 0.0;
    }
    if (blockIdx.x == 0) {
        __shared__ double red[8];
        double t = 0.0;
        for (int i = threadIdx.x; i < n; i += 256) t += fabs(A[(size_t)i * lda + i]);
        t = warp_sum(t);
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = t;
        __syncthreads();
        if (threadIdx.x == 0) {
            double s = 0.0;
            for (int i = 0; i < 8; ++i) s += red[i];
            scratch[0] = s;
        }
    }
    grid.sync();
    const double thr = rel_thr * scratch[0] / (double)n;    // rotations below this are skipped
    for (int sweep = 0; sweep < max_sweeps; ++sweep) {
        for (int r = 0; r < n - 1; ++r) {
            for (int k = gtid; k < half; k += gth) {
                int p, q;
                rr_pair(n, r, k, p, q);
                const double apq = A[(size_t)p * lda + q];
                double c = 1.0, s = 0.0;
                if (fabs(apq) > thr) {
                    const double app = A[(size_t)p * lda + p], aqq = A[(size_t)q * lda + q];
                    const double tau = (aqq - app) / (2.0 * apq);
                    const double t = copysign(1.0, tau) / (fabs(tau) + sqrt(1.0 + tau * tau));
                    c = 1.0 / sqrt(1.0 + t * t);
                    s = t * c;
                    atomicAdd(&counters[sweep], 1);
                }
                cs[2 * k] = c;
                cs[2 * k + 1] = s;
            }
            grid.sync();
            for (size_t e = gtid; e < (size_t)half * half; e += gth) {
                const int k1 = (int)(e / half), k2 = (int)(e % half);
                int p1, q1, p2, q2;
                rr_pair(n, r, k1, p1, q1);
                rr_pair(n, r, k2, p2, q2);
                const double c1 = cs[2 * k1], s1 = cs[2 * k1 + 1];
                const double c2 = cs[2 * k2], s2 = cs[2 * k2 + 1];
                double* a_pp = A + (size_t)p1 * lda + p2;
                double* a_pq = A + (size_t)p1 * lda + q2;
                double* a_qp = A + (size_t)q1 * lda + p2;
                double* a_qq = A + (size_t)q1 * lda + q2;
                const double a = *a_pp, b = *a_pq, c = *a_qp, dd = *a_qq;
                // rows: J1^T * blk
                const double ra = c1 * a - s1 * c, rb = c1 * b - s1 * dd;
                const double rc = s1 * a + c1 * c, rd = s1 * b + c1 * dd;
                // cols: * J2
                double na = c2 * ra - s2 * rb, nb = s2 * ra + c2 * rb;
                double nc = c2 * rc - s2 * rd, ndv = s2 * rc + c2 * rd;
                if (k1 == k2) { nb = 0.0; nc = 0.0; }
                *a_pp = na; *a_pq = nb; *a_qp = nc; *a_qq = ndv;
            }
            for (size_t e = gtid; e < (size_t)n * half; e += gth) {
                const int row = (int)(e / half), k2 = (int)(e % half);
                int p2, q2;
                rr_pair(n, r, k2, p2, q2);
                const double c2 = cs[2 * k2], s2 = cs[2 * k2 + 1];
                double* vp = V + (size_t)row * lda + p2;
                double* vq = V + (size_t)row * lda + q2;
                const double a = *vp, b = *vq;
                *vp = c2 * a - s2 * b;
                *vq = s2 * a + c2 * b;
            }
            grid.sync();
        }
        if (counters[sweep] == 0) break;
    }
}

// descending rank of diag(A) -> perm[rank] = index, eval[rank] = value
__global__ void k_sort_eigs(const double* A, int n, int lda, int* perm, double* eval) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const double v = A[(size_t)i * lda + i];
        int r = 0;
        for (int j = 0; j < n; ++j) {
            const double vj = A[(size_t)j * lda + j];
            r += (vj > v) || (vj == v && j < i);
        }
        perm[r] = i;
        eval[r] = v;
    }
}

// sketch[i] = sqrt(s_i) * vt[i],  vt[i] = (1/sqrt(lam_i)) sum_j V[j][perm[i]] M[j],
// s_i = lam_i - lam_l (when ng > l).  One 64x64 tile per CTA.
template <bool VT_LAYOUT>   // VT_LAYOUT: eigenvector i is ROW perm[i] of V (eig_big.cuh); else column perm[i]
__global__ void __launch_bounds__(256) k_merge_out(const double* __restrict__ M, int ng, long long d,
                                                   const double* __restrict__ V, int lda,
                                                   const int* __restrict__ perm,
                                                   const double* __restrict__ eval, int ell,
                                                   double* sketch, double* vt, double* sigma2) {
    __shared__ GemmSmem sm;
    const double cut = (ng > ell) ? fmax(eval[ell], 0.0) : 0.0;
    const int tm = (ell + GT - 1) / GT, tn = (int)((d + GT - 1) / GT);
    for (int tile = blockIdx.x; tile < tm * tn; tile += gridDim.x) {
        const int m0 = (tile / tn) * GT, n0 = (tile % tn) * GT;
        gemm_tile_f64<VT_LAYOUT, true>(
            sm, threadIdx.x, m0, n0, ng,
            [&](int i, int j) -> double {
                if (!(i < ell && i < ng && j < ng)) return 0.0;
                return VT_LAYOUT ? V[(size_t)perm[i] * lda + j] : V[(size_t)j * lda + perm[i]];
            },
            [&](int j, int c) -> double { return (j < ng && c < d) ? M[(size_t)j * d + c] : 0.0; },
            [&](int i, int c, double v) {
                if (i < ell && c < d) {
                    const double li = (i < ng) ? fmax(eval[i], 0.0) : 0.0;
                    const double vv = (li > 0.0) ? v / sqrt(li) : 0.0;
                    if (vt) vt[(size_t)i * d + c] = vv;
                    sketch[(size_t)i * d + c] = vv * sqrt(fmax(li - cut, 0.0));
                }
            },
            [&]() { __syncthreads(); });
    }
    if (blockIdx.x == 0)
        for (int i = threadIdx.x; i < ell; i += blockDim.x)
            sigma2[i] = (i < ng) ? fmax(eval[i], 0.0) - cut : 0.0;
}

// FD state out (FrequentDirections.get, frequent_directions.py:60-61): sketch,
// sigma2, Vt = sketch / sqrt(sigma2) (rows with sigma2 == 0 -> 0), delta.
__global__ void k_fd_out(Dev D, int s, double* sketch, double* sigma2, double* vt, double* delta) {
    const Slot& sl = D.slots[s];
    const double* R = D.Rp(s, sl.rcur);
    const double* s2 = D.v(s, V_S2);
    const size_t tot = (size_t)D.ell * D.d;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < tot;
         e += (size_t)gridDim.x * blockDim.x) {
        const int i = (int)(e / D.d);
        const double v = R[e];
        sketch[e] = v;
        if (vt) vt[e] = (s2[i] > 0.0) ? v / sqrt(s2[i]) : 0.0;
    }
    if (blockIdx.x == 0) {
        for (int i = threadIdx.x; i < D.ell; i += blockDim.x) sigma2[i] = s2[i];
        if (threadIdx.x == 0) *delta = sl.delta;
    }
}


// C = M^T M (d x d, leading dimension lda), float64: the covariance side of the stack, used
// when it has more rows than columns (its eigenvectors ARE the right singular vectors).
__global__ void __launch_bounds__(256) k_cov_f64(const double* __restrict__ M, int ng, long long d,
                                                 double* __restrict__ Cm, int lda) {
    __shared__ GemmSmem sm;
    const int tiles = (int)((d + GT - 1) / GT);
    for (int tile = blockIdx.x; tile < tiles * tiles; tile += gridDim.x) {
        const int ti = tile / tiles, tj = tile % tiles;
        if (tj > ti) continue;
        const int m0 = ti * GT, n0 = tj * GT;
        gemm_tile_f64<false, true>(
            sm, threadIdx.x, m0, n0, ng,
            [&](int i, int r) -> double { return (i < d && r < ng) ? M[(size_t)r * d + i] : 0.0; },
            [&](int r, int j) -> double { return (j < d && r < ng) ? M[(size_t)r * d + j] : 0.0; },
            [&](int i, int j, double v) {
                if (i < d && j < d) {
                    Cm[(size_t)i * lda + j] = v;
                    Cm[(size_t)j * lda + i] = v;
                }
            },
            [&]() { __syncthreads(); });
    }
}

// covariance-side output: vt[i] = eigenvector i (row perm[i] of VT), sketch[i] = sqrt(lam_i - lam_l) vt[i]
__global__ void k_cov_out(const double* __restrict__ VT, int ldt, const int* __restrict__ perm,
                          const double* __restrict__ eval, int n, int ell, long long d, double* sketch, double* vt,
                          double* sigma2) {
    const double cut = (n > ell) ? fmax(eval[ell], 0.0) : 0.0;
    for (int i = blockIdx.x; i < ell; i += gridDim.x) {
        const double li = (i < n) ? fmax(eval[i], 0.0) : 0.0;
        const double sc = sqrt(fmax(li - cut, 0.0));
        const double* src = VT + (size_t)perm[i < n ? i : 0] * ldt;
        for (long long c = threadIdx.x; c < d; c += blockDim.x) {
            const double v = (i < n && li > 0.0) ? src[c] : 0.0;
            if (vt) vt[(size_t)i * d + c] = v;
            sketch[(size_t)i * d + c] = v * sc;
        }
        if (threadIdx.x == 0) sigma2[i] = li - cut;
    }
}

// rows of VT in descending eigenvalue order -> vecs (n x n, row i = eigenvector i)
__global__ void k_eig_permute_out(const double* __restrict__ VT, int ldt, const int* __restrict__ perm,
                                  const double* __restrict__ eval, int n, double* evals, double* vecs, long long ldv) {
    for (int i = blockIdx.x; i < n; i += gridDim.x) {
        const double* src = VT + (size_t)perm[i] * ldt;
        for (int c = threadIdx.x; c < n; c += blockDim.x) vecs[(size_t)i * ldv + c] = src[c];
        if (threadIdx.x == 0) evals[i] = eval[i];
    }
}

// Shared tail of both engines' get(): M (ng x d, float64) -> Gram -> Jacobi ->
// top-l rows shrunk by the (l+1)-th eigenvalue (sliding_windows_frequent_directions.py:160-171).
int merge_query_gram(const QueryScratch& Q, int ng, long long d, int sm_count, cudaStream_t st) {
    const int nj = ng + (ng & 1);
    MUSED_REQUIRE(nj <= Q.ng_max, "internal: stack of %d rows exceeds scratch %d", nj, Q.ng_max);
    MUSED_CUDA(cudaMemsetAsync(Q.G, 0, (size_t)Q.lda * Q.lda * sizeof(double), st));
    const int tiles = (ng + GT - 1) / GT;
    k_gram_f64<<<std::min(tiles * tiles, 8 * sm_count), 256, 0, st>>>(Q.M, ng, d, Q.G, Q.lda);
    MUSED_CUDA(cudaGetLastError());
    COUNT_LAUNCH(1);
    return 0;
}

// Whole-GPU block Jacobi (eig_big.cuh) when the matrix fits the co-resident CTAs'
// shared memory; *done = false -> the caller falls back to k_jacobi.
template <int CB>
static int launch_eig_big(const QueryScratch& Q, int n, int sm_count, cudaStream_t st, int* npad, bool* done) {
    *done = false;
    const int nr = (n + 31) / 32 * 32;
    const int P = (n + 2 * CB - 1) / (2 * CB);
    const size_t smem = big::Cfg<CB>::smem_bytes(nr);
    if (smem > 227u * 1024u || P > sm_count || P > 256 || nr > Q.ldt || 2 * P * CB > Q.npad_max ||
        (size_t)P * CB * nr > Q.mail_count)
        return 0;
    MUSED_SMEM_ATTR(big::k_eig_big<CB>, 227 * 1024);
    int nb = 0;
    MUSED_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, big::k_eig_big<CB>, big::NT, smem));
    if (nb < 1) return 0;
    MUSED_CUDA(cudaMemsetAsync(Q.sync, 0, (size_t)QSYNC_WORDS * sizeof(unsigned long long), st));
    big::Args a;
    a.G = Q.G; a.lda = Q.lda; a.n = n; a.nr = nr; a.P = P;
    a.mailL = Q.mailL; a.mailR = Q.mailR;
    a.flagL = Q.sync; a.flagR = Q.sync + 256; a.red = Q.sync + 512;
    a.lam = Q.lam; a.VT = Q.VT; a.ldt = Q.ldt;
    a.tol = Q.big_tol; a.max_sweeps = 40;
    a.info = reinterpret_cast<int*>(Q.sync + 640);
    a.prof = Q.sync + 700;
    void* args[] = {&a};
    MUSED_CUDA(cudaLaunchCooperativeKernel((void*)big::k_eig_big<CB>, dim3(P), dim3(big::NT), args, smem, st));
    *npad = 2 * P * CB;
    *done = true;
    return 0;
}

static int eig_big_cb_floor() {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("MUSED_EIG_CB");
        v = e ? atoi(e) : 4;
        if (v < 1) v = 1;
        if (v > 6) v = 6;
    }
    return v;
}

static int dispatch_eig_big(const QueryScratch& Q, int n, int sm_count, cudaStream_t st, int* npad, bool* done) {
    *done = false;
    int cb = (n + 2 * sm_count - 1) / (2 * sm_count);
    if (cb < eig_big_cb_floor()) cb = eig_big_cb_floor();
    switch (cb) {
        case 1: return launch_eig_big<1>(Q, n, sm_count, st, npad, done);
        case 2: return launch_eig_big<2>(Q, n, sm_count, st, npad, done);
        case 3: return launch_eig_big<3>(Q, n, sm_count, st, npad, done);
        case 4: return launch_eig_big<4>(Q, n, sm_count, st, npad, done);
        case 5: return launch_eig_big<5>(Q, n, sm_count, st, npad, done);
        case 6: return launch_eig_big<6>(Q, n, sm_count, st, npad, done);
        default: return 0;
    }
}

// Column-sharded callers add the other ranks' Grams into Q.G between the two halves.
int merge_query_solve(const QueryScratch& Q, int ng, long long d, int ell, int sm_count, double* o_sketch,
                      double* o_vt, double* o_s2, cudaStream_t st) {
    const int tm = (ell + GT - 1) / GT, tn = (int)((d + GT - 1) / GT);
    bool big_done = false;
    int npad = 0;
    if (Q.VT && !getenv("MUSED_EIG_LEGACY")) {
        int rc = dispatch_eig_big(Q, ng, sm_count, st, &npad, &big_done);
        if (rc) return rc;
    }
    if (big_done) {
        k_sort_eigs<<<(npad + 255) / 256, 256, 0, st>>>(Q.lam, npad, 0, Q.perm, Q.eval);
        MUSED_CUDA(cudaGetLastError());
        k_merge_out<true><<<std::min(tm * tn, 8 * sm_count), 256, 0, st>>>(Q.M, ng, d, Q.VT, Q.ldt, Q.perm, Q.eval, ell,
                                                                          o_sketch, o_vt, o_s2);
        MUSED_CUDA(cudaGetLastError());
        COUNT_LAUNCH(3);
        return 0;
    }
    const int nj = ng + (ng & 1);
    MUSED_CUDA(cudaMemsetAsync(Q.counters, 0, 64 * sizeof(int), st));
    {
        int nb = 0;
        MUSED_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_jacobi, 256, 0));
        if (nb < 1) nb = 1;
        if (nb > 4) nb = 4;
        int grid = nb * sm_count;
        const long long work = ((long long)nj / 2) * (nj / 2);
        const int need = (int)std::max(1ll, (work + 255) / 256);
        if (grid > need) grid = need;
        double *A_ = Q.G, *V_ = Q.Vq, *cs_ = Q.cs, *scr_ = Q.scr;
        int n_ = nj, lda_ = Q.lda, sweeps_ = 40;
        int* cnt_ = Q.counters;
        double thr_ = Q.rel_thr;
        void* args[] = {&A_, &V_, &n_, &lda_, &cs_, &sweeps_, &cnt_, &scr_, &thr_};
        MUSED_CUDA(cudaLaunchCooperativeKernel((void*)k_jacobi, dim3(grid), dim3(256), args, 0, st));
    }
    k_sort_eigs<<<(nj + 255) / 256, 256, 0, st>>>(Q.G, nj, Q.lda, Q.perm, Q.eval);
    MUSED_CUDA(cudaGetLastError());
    k_merge_out<false><<<std::min(tm * tn, 8 * sm_count), 256, 0, st>>>(Q.M, ng, d, Q.Vq, Q.lda, Q.perm, Q.eval, ell,
                                                                       o_sketch, o_vt, o_s2);
    MUSED_CUDA(cudaGetLastError());
    COUNT_LAUNCH(3);
    return 0;
}

int merge_query(const QueryScratch& Q, int ng, long long d, int ell, int sm_count, double* o_sketch,
                double* o_vt, double* o_s2, cudaStream_t st) {
    // more rows than columns: eigensolve the d x d covariance instead of the ng x ng Gram
    // (smaller, no structural null space, and its eigenvectors are the rows of Vt)
    if (Q.VT && d < ng && d <= Q.lda && !getenv("MUSED_EIG_LEGACY") && !getenv("MUSED_EIG_NOCOV")) {
        const int tiles = (int)((d + GT - 1) / GT);
        k_cov_f64<<<std::min(tiles * tiles, 8 * sm_count), 256, 0, st>>>(Q.M, ng, d, Q.G, Q.lda);
        MUSED_CUDA(cudaGetLastError());
        bool done = false;
        int npad = 0;
        int rc = dispatch_eig_big(Q, (int)d, sm_count, st, &npad, &done);
        if (rc) return rc;
        if (done) {
            k_sort_eigs<<<(npad + 255) / 256, 256, 0, st>>>(Q.lam, npad, 0, Q.perm, Q.eval);
            k_cov_out<<<std::min(ell, 4 * sm_count), 256, 0, st>>>(Q.VT, Q.ldt, Q.perm, Q.eval, (int)d, ell, d, o_sketch,
                                                                    o_vt, o_s2);
            MUSED_CUDA(cudaGetLastError());
            COUNT_LAUNCH(4);
            return 0;
        }
    }
    int rc = merge_query_gram(Q, ng, d, sm_count, st);
    if (rc) return rc;
    return merge_query_solve(Q, ng, d, ell, sm_count, o_sketch, o_vt, o_s2, st);
}

// ---------------------------------------------------------------------------
// Engine
// ---------------------------------------------------------------------------
struct Engine {
    Dev D{};
    int device = 0, P = 1, n_levels = 0;
    long long chunk_rows = 0;
    unsigned long long time = 0;
    bool fd_only = false;
    std::vector<double> thresholds;
    // staging
    unsigned char* raw = nullptr;  size_t raw_bytes = 0;
    double* stage = nullptr;
    double* norm2 = nullptr;
    unsigned long long* times_dev = nullptr;   // explicit arrival times of the staged chunk
    // query scratch
    QueryScratch Q;
    double *o_sketch = nullptr, *o_vt = nullptr, *o_s2 = nullptr, *o_delta = nullptr;
    int* o_flag = nullptr;
    std::vector<StateAlloc> allocs;   // every device allocation; .state = part of the checkpoint (mused_swfd_state_*)
    int sm_count = 148;
    bool timing = false;
    std::vector<cudaEvent_t> ev;   // start/stop pairs around k_swfd_rows (MODE_SEQ/FD)
    double timed_ms = 0.0;
    long long timed_launches = 0, timed_rows = 0;

    template <class T>
    int alloc(T** p, size_t count, bool state = true) {
        void* q = nullptr;
        cudaError_t e = cudaMalloc(&q, count * sizeof(T));
        if (e != cudaSuccess) {
            set_error("cudaMalloc of %zu bytes failed: %s", count * sizeof(T), cudaGetErrorString(e));
            return MUSED_ERR_NOMEM;
        }
        e = cudaMemset(q, 0, count * sizeof(T));
        if (e != cudaSuccess) { set_error("cudaMemset failed: %s", cudaGetErrorString(e)); return MUSED_ERR_CUDA; }
        allocs.push_back(StateAlloc{q, count * sizeof(T), state});
        *p = (T*)q;
        return 0;
    }
    ~Engine() {
        for (const StateAlloc& a : allocs) cudaFree(a.p);
    }
};

static int launch_rows(Engine& E, const RowArgs& A, int nslots, cudaStream_t st) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(nslots * E.P));
    cfg.blockDim = dim3(NT);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)E.P;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    switch (E.P) {
        case 1: MUSED_CUDA(cudaLaunchKernelEx(&cfg, k_swfd_rows<1>, E.D, A)); break;
        case 2: MUSED_CUDA(cudaLaunchKernelEx(&cfg, k_swfd_rows<2>, E.D, A)); break;
        case 4: MUSED_CUDA(cudaLaunchKernelEx(&cfg, k_swfd_rows<4>, E.D, A)); break;
        case 8: MUSED_CUDA(cudaLaunchKernelEx(&cfg, k_swfd_rows<8>, E.D, A)); break;
        default: set_error("cluster must be 1, 2, 4 or 8"); return MUSED_ERR_ARG;
    }
    COUNT_LAUNCH(1);
    return 0;
}

static int engine_create(Engine** out, long long N, long long d, long long ell, double beta,
                         int n_levels, const double* thr, int device, int cluster,
                         long long chunk_rows, bool fd_only, double fd_error) {
    int ndev = 0;
    cudaError_t ce = cudaGetDeviceCount(&ndev);
    if (ce != cudaSuccess || ndev == 0) {
        set_error("no CUDA device available (%s); mused_b200 has no CPU fallback",
                  ce != cudaSuccess ? cudaGetErrorString(ce) : "device count is 0");
        return MUSED_ERR_CUDA;
    }
    MUSED_REQUIRE(device >= 0 && device < ndev, "device %d out of range (have %d)", device, ndev);
    MUSED_REQUIRE(ell >= 1 && ell <= 512, "sketch_dim must be in [1, 512], got %lld", ell);
    MUSED_REQUIRE(d > ell, "the B200 engine requires d > sketch_dim (d=%lld, sketch_dim=%lld)", d, ell);
    MUSED_REQUIRE(N >= 1, "window N must be >= 1");
    MUSED_REQUIRE(beta > 0, "beta must be > 0");
    MUSED_REQUIRE(n_levels >= 0 && n_levels <= 64, "n_levels out of range");
    MUSED_CUDA(cudaSetDevice(device));
    Engine* E = new Engine();
    E->device = device;
    E->n_levels = n_levels;
    E->fd_only = fd_only;
    cudaDeviceProp prop;
    MUSED_CUDA(cudaGetDeviceProperties(&prop, device));
    E->sm_count = prop.multiProcessorCount;
    const int S = fd_only ? 1 : 2 * n_levels;
    if (cluster == 0) {
        // the team splits the O(n^3) eigenvector update and the d-wide GEMV /
        // rotate; small problems do not amortise the cluster barrier
        cluster = (ell >= 96) ? 8 : (ell >= 48 ? 4 : (ell >= 24 ? 2 : 1));
        while (cluster > 1 && (long long)S * cluster > 2LL * E->sm_count) cluster >>= 1;
    }
    MUSED_REQUIRE(cluster == 1 || cluster == 2 || cluster == 4 || cluster == 8,
                  "cluster must be 0 (auto), 1, 2, 4 or 8");
    E->P = cluster;
    Dev& D = E->D;
    D.S = S;
    D.ell = (int)ell;
    D.nmax = 2 * (int)ell;
    D.ldq = D.nmax + 2;
    D.d = d;
    D.N = (unsigned long long)N;
    D.qcap = (2.0 + 8.0 / beta) * (double)ell;
    D.capn = fd_only ? 1 : (int)floor(D.qcap) + 2;
    if (chunk_rows <= 0) {
        chunk_rows = (long long)((64ll << 20) / (d * 8));
        if (chunk_rows < 64) chunk_rows = 64;
        if (chunk_rows > 8192) chunk_rows = 8192;
    }
    E->chunk_rows = chunk_rows;
    int rc = 0;
#define A_(p, n) if ((rc = E->alloc(&(p), (n))) != 0) { delete E; return rc; }
#define S_(p, n) if ((rc = E->alloc(&(p), (n), false)) != 0) { delete E; return rc; }   /* scratch: not checkpointed */
    if (S > 0) {
        A_(D.slots, (size_t)S);
        A_(D.R, (size_t)S * 2 * D.nmax * d);
        A_(D.Q, (size_t)S * 2 * D.ldq * D.ldq);
        A_(D.UT, (size_t)S * D.ldq * D.ldq);
        A_(D.vec, (size_t)S * NV * D.ldq);
        A_(D.rotcs, (size_t)S * 2 * D.ldq);
        A_(D.ivec, (size_t)S * NI * D.ldq);
        A_(D.ringV, (size_t)S * D.capn * d);
        A_(D.ringS, (size_t)S * D.capn);
        A_(D.ringT, (size_t)S * D.capn);
    }
    E->raw_bytes = (size_t)chunk_rows * d * 8;
    S_(E->raw, E->raw_bytes);
    S_(E->stage, (size_t)chunk_rows * d);
    S_(E->norm2, (size_t)chunk_rows);
    S_(E->times_dev, (size_t)chunk_rows);
    QueryScratch& Q = E->Q;
    Q.ng_max = D.ell + D.capn + 1;
    Q.ng_max += Q.ng_max & 1;
    Q.lda = Q.ng_max;
    if (!fd_only) {
        S_(Q.M, (size_t)Q.ng_max * d);
        S_(Q.G, (size_t)Q.lda * Q.lda);
        S_(Q.Vq, (size_t)Q.lda * Q.lda);
        S_(Q.cs, (size_t)Q.ng_max);
        Q.plan_big();
        S_(Q.eval, (size_t)Q.npad_max);
        S_(Q.perm, (size_t)Q.npad_max);
        S_(Q.lam, (size_t)Q.npad_max);
        S_(Q.VT, (size_t)Q.npad_max * Q.ldt);
        S_(Q.mailL, Q.mail_count);
        S_(Q.mailR, Q.mail_count);
        S_(Q.sync, (size_t)QSYNC_WORDS);
        S_(Q.counters, 64);
        S_(Q.scr, 8);
    }
    S_(E->o_sketch, (size_t)ell * d);
    S_(E->o_vt, (size_t)ell * d);
    S_(E->o_s2, (size_t)ell);
    S_(E->o_delta, 1);
    S_(E->o_flag, 1);
#undef S_
#undef A_
    std::vector<Slot> init(S);
    memset(init.data(), 0, sizeof(Slot) * S);
    for (int s = 0; s < S; ++s) {
        Slot& sl = init[s];
        sl.error = fd_only ? fd_error : thr[s / 2];
        sl.n = (int)ell;
        sl.q_ident = 1;
        sl.last_dump = 1;  // DumpFDwithSnapshotQueue.last_dump_time, sliding_windows_...py:23
        sl.parity = s & 1;
    }
    if (S > 0)
        MUSED_CUDA(cudaMemcpy(D.slots, init.data(), sizeof(Slot) * S, cudaMemcpyHostToDevice));
    if (!fd_only) E->thresholds.assign(thr, thr + n_levels);
    *out = E;
    return 0;
}

static size_t dtype_size(int dt) {
    switch (dt) {
        case MUSED_F64: return 8;
        case MUSED_F32: return 4;
        case MUSED_I64: return 8;
        case MUSED_I32: return 4;
        case MUSED_U8: return 1;
    }
    return 0;
}

static int stage_chunk(Engine& E, const void* dev_rows, long long n, long long ld, int dtype,
                       cudaStream_t st) {
    const long long d = E.D.d;
    switch (dtype) {
        case MUSED_F64: k_stage_rows<double><<<(unsigned)n, 256, 0, st>>>((const double*)dev_rows, ld, d, E.stage, E.norm2); break;
        case MUSED_F32: k_stage_rows<float><<<(unsigned)n, 256, 0, st>>>((const float*)dev_rows, ld, d, E.stage, E.norm2); break;
        case MUSED_I64: k_stage_rows<long long><<<(unsigned)n, 256, 0, st>>>((const long long*)dev_rows, ld, d, E.stage, E.norm2); break;
        case MUSED_I32: k_stage_rows<int><<<(unsigned)n, 256, 0, st>>>((const int*)dev_rows, ld, d, E.stage, E.norm2); break;
        case MUSED_U8: k_stage_rows<unsigned char><<<(unsigned)n, 256, 0, st>>>((const unsigned char*)dev_rows, ld, d, E.stage, E.norm2); break;
    }
    MUSED_CUDA(cudaGetLastError());
    COUNT_LAUNCH(1);
    return 0;
}

// feed rows chunk by chunk; `mode` selects the per-row rule set
static int engine_feed(Engine& E, const void* rows, long long n_rows, long long ld, int dtype,
                       int where, int mode, long long batch, cudaStream_t st,
                       const unsigned long long* times_host = nullptr) {
    const size_t es = dtype_size(dtype);
    MUSED_REQUIRE(es != 0, "unknown dtype %d", dtype);
    MUSED_REQUIRE(n_rows >= 0 && ld >= E.D.d, "bad shape: n_rows=%lld ld=%lld d=%lld", n_rows, ld, E.D.d);
    MUSED_REQUIRE(rows != nullptr || n_rows == 0, "rows is NULL");
    MUSED_CUDA(cudaSetDevice(E.device));
    if (E.D.S == 0) {
        E.time = (times_host && n_rows > 0) ? times_host[n_rows - 1] : E.time + (unsigned long long)n_rows;
        return 0;
    }
    long long done = 0;
    while (done < n_rows) {
        const long long n = std::min(E.chunk_rows, n_rows - done);
        const char* src = (const char*)rows + (size_t)done * ld * es;
        const void* dev_rows = src;
        long long dev_ld = ld;
        if (where == MUSED_HOST) {
            // contiguous rows: one copy; otherwise a strided 2-D copy
            if (ld == E.D.d) {
                MUSED_CUDA(cudaMemcpyAsync(E.raw, src, (size_t)n * E.D.d * es, cudaMemcpyHostToDevice, st));
            } else {
                MUSED_CUDA(cudaMemcpy2DAsync(E.raw, (size_t)E.D.d * es, src, (size_t)ld * es,
                                             (size_t)E.D.d * es, (size_t)n, cudaMemcpyHostToDevice, st));
            }
            dev_rows = E.raw;
            dev_ld = E.D.d;
        }
        int rc = stage_chunk(E, dev_rows, n, dev_ld, dtype, st);
        if (rc) return rc;
        RowArgs A = {};
        A.rows = E.stage;
        A.norm2 = E.norm2;
        A.n_rows = n;
        A.time0 = E.time;
        if (times_host) {
            MUSED_CUDA(cudaMemcpyAsync(E.times_dev, times_host + done, (size_t)n * sizeof(unsigned long long),
                                       cudaMemcpyHostToDevice, st));
            A.times = E.times_dev;
        }
        A.mode = mode;
        A.slot0 = 0;
        A.batch = batch;
        A.last_call = (done + n == n_rows);
        cudaEvent_t e0 = nullptr, e1 = nullptr;
        if (E.timing) {
            MUSED_CUDA(cudaEventCreate(&e0));
            MUSED_CUDA(cudaEventCreate(&e1));
            MUSED_CUDA(cudaEventRecord(e0, st));
        }
        rc = launch_rows(E, A, E.D.S, st);
        if (rc) return rc;
        if (E.timing) {
            MUSED_CUDA(cudaEventRecord(e1, st));
            E.ev.push_back(e0);
            E.ev.push_back(e1);
            E.timed_launches += 1;
            E.timed_rows += n;
        }
        E.time = times_host ? times_host[done + n - 1] : E.time + (unsigned long long)n;
        done += n;
        if ((where == MUSED_HOST || times_host) && done < n_rows) MUSED_CUDA(cudaStreamSynchronize(st));
    }
    if (where == MUSED_HOST || times_host) MUSED_CUDA(cudaStreamSynchronize(st));
    return 0;
}

static int flush_slot(Engine& E, int s, cudaStream_t st) {
    RowArgs A = {};
    A.mode = MODE_FLUSH;
    A.slot0 = s;
    return launch_rows(E, A, 1, st);
}

static int copy_out(double* dst, const double* src, size_t n, int where, cudaStream_t st) {
    if (!dst) return 0;
    MUSED_CUDA(cudaMemcpyAsync(dst, src, n * sizeof(double),
                               where == MUSED_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice, st));
    return 0;
}

// the physical slot that currently plays the reference's `fd` of level j
static int fd_slot(const Engine& E, int level) {
    const unsigned long long p = E.time / E.D.N;
    return 2 * level + (int)(p & 1ull);
}

static int engine_get(Engine& E, int s, bool merge, double* sketch, double* sigma2, double* vt,
                      double* delta, int where, cudaStream_t st) {
    Dev& D = E.D;
    int rc = flush_slot(E, s, st);  // FrequentDirectionsWithDump.get_sketch flushes (:148-150)
    if (rc) return rc;
    Slot sl;
    MUSED_CUDA(cudaMemcpyAsync(&sl, D.slots + s, sizeof(Slot), cudaMemcpyDeviceToHost, st));
    MUSED_CUDA(cudaStreamSynchronize(st));
    const size_t ld_out = (size_t)D.ell * D.d;
    if (!merge || sl.q_len == 0) {
        k_fd_out<<<E.sm_count, 256, 0, st>>>(D, s, E.o_sketch, E.o_s2, vt ? E.o_vt : nullptr, E.o_delta);
        MUSED_CUDA(cudaGetLastError());
        COUNT_LAUNCH(1);
    } else {
        const int ng = D.ell + sl.q_len;
        k_gather_stack<<<std::min(ng, 4 * E.sm_count), 256, 0, st>>>(D, s, E.Q.M);
        MUSED_CUDA(cudaGetLastError());
        COUNT_LAUNCH(1);
        if ((rc = merge_query(E.Q, ng, D.d, D.ell, E.sm_count, E.o_sketch, vt ? E.o_vt : nullptr, E.o_s2, st))) return rc;
        MUSED_CUDA(cudaMemsetAsync(E.o_delta, 0, sizeof(double), st));  // returns 0.0 (:171)
    }
    if ((rc = copy_out(sketch, E.o_sketch, ld_out, where, st))) return rc;
    if ((rc = copy_out(sigma2, E.o_s2, (size_t)D.ell, where, st))) return rc;
    if ((rc = copy_out(vt, E.o_vt, ld_out, where, st))) return rc;
    if (delta) MUSED_CUDA(cudaMemcpyAsync(delta, E.o_delta, sizeof(double), cudaMemcpyDeviceToHost, st));
    MUSED_CUDA(cudaStreamSynchronize(st));
    return 0;
}

}  // namespace mused

using namespace mused;

struct mused_swfd { Engine* E; BEngine* B; };
struct mused_fd { Engine* E; };

extern "C" {

const char* mused_last_error(void) { return g_err.c_str(); }
uint64_t mused_launch_count(void) { return g_launches; }
void mused_count_launches(uint64_t n) { g_launches += n; }
int mused_version(void) { return 100; }
int mused_device_count(void) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) { set_error("cudaGetDeviceCount: %s", cudaGetErrorString(e)); return MUSED_ERR_CUDA; }
    return n;
}

int mused_swfd_create(mused_swfd_t** out, const mused_swfd_cfg* cfg) {
    MUSED_REQUIRE(out && cfg, "NULL argument");
    MUSED_REQUIRE(cfg->n_levels == 0 || cfg->thresholds, "thresholds is NULL");
    MUSED_REQUIRE(cfg->schedule == MUSED_SCHEDULE_REFERENCE || cfg->schedule == MUSED_SCHEDULE_BATCHED,
                  "unknown schedule %d", cfg->schedule);
    if (cfg->schedule == MUSED_SCHEDULE_BATCHED) {
        BEngine* B = nullptr;
        int rc = bengine_create(&B, cfg->N, cfg->d, cfg->ell, cfg->beta, cfg->n_levels, cfg->thresholds,
                                cfg->device, cfg->batch_rows, cfg->n_streams > 0 ? cfg->n_streams : 1);
        if (rc) return rc;
        *out = new mused_swfd{nullptr, B};
        return 0;
    }
    MUSED_REQUIRE(cfg->n_streams <= 1, "n_streams > 1 (independent streams batched per GPU) is served by schedule=batched");
    Engine* E = nullptr;
    int rc = engine_create(&E, cfg->N, cfg->d, cfg->ell, cfg->beta, cfg->n_levels, cfg->thresholds,
                           cfg->device, cfg->cluster, cfg->chunk_rows, false, 0.0);
    if (rc) return rc;
    *out = new mused_swfd{E, nullptr};
    return 0;
}
int mused_swfd_destroy(mused_swfd_t* h) {
    if (!h) return 0;
    if (h->B) {
        bengine_destroy(h->B);
    } else {
        cudaSetDevice(h->E->device);
        delete h->E;
    }
    delete h;
    return 0;
}
int mused_swfd_fit(mused_swfd_t* h, const void* rows, int64_t n_rows, int64_t ld, int dtype,
                   int where, void* stream) {
    MUSED_REQUIRE(h, "NULL handle");
    if (h->B) {
        MUSED_REQUIRE(bengine_streams(h->B) == 1, "this handle batches %d streams: use mused_swfd_fit_streams", bengine_streams(h->B));
        return bengine_fit(h->B, rows, n_rows, ld, 0, dtype, where, (cudaStream_t)stream);
    }
    return engine_feed(*h->E, rows, n_rows, ld, dtype, where, MODE_SEQ, 0, (cudaStream_t)stream);
}
int mused_swfd_fit_streams(mused_swfd_t* h, const void* rows, int64_t n_rows, int64_t ld, int64_t stream_stride,
                           int dtype, int where, void* stream) {
    MUSED_REQUIRE(h, "NULL handle");
    MUSED_REQUIRE(h->B, "independent streams batched per GPU are served by schedule=batched");
    return bengine_fit(h->B, rows, n_rows, ld, stream_stride, dtype, where, (cudaStream_t)stream);
}
int mused_swfd_get_stream(mused_swfd_t* h, int32_t k, double* sketch, double* sigma2, double* vt, double* delta,
                          int32_t* level, int where, void* stream) {
    MUSED_REQUIRE(h && sketch && sigma2, "NULL argument");
    MUSED_REQUIRE(h->B, "independent streams batched per GPU are served by schedule=batched");
    return bengine_get(h->B, k, sketch, sigma2, vt, delta, level, where, (cudaStream_t)stream);
}
int mused_swfd_size_stream(mused_swfd_t* h, int32_t k, int64_t* n) {
    MUSED_REQUIRE(h && n && h->B, "NULL argument / not a batched handle");
    long long v = 0;
    int rc = bengine_size(h->B, k, &v);
    *n = v;
    return rc;
}
int mused_swfd_debug_stream(mused_swfd_t* h, int32_t k, double* scal, double* sigma2, uint64_t* q_s, uint64_t* q_t,
                            int64_t q_cap) {
    MUSED_REQUIRE(h && h->B, "NULL argument / not a batched handle");
    return bengine_debug(h->B, k, scal, sigma2, (unsigned long long*)q_s, (unsigned long long*)q_t, q_cap);
}
int mused_comm_unique_id(void* id_out, int64_t cap, int64_t* n_bytes, const char* nccl_path) {
    long long n = 0;
    int rc = nccl_unique_id(id_out, cap, &n, nccl_path);
    if (n_bytes) *n_bytes = n;
    return rc;
}
int mused_swfd_attach_comm(mused_swfd_t* h, const void* unique_id, int64_t n_bytes, int32_t rank, int32_t nranks,
                           const char* nccl_path) {
    MUSED_REQUIRE(h, "NULL handle");
    MUSED_REQUIRE(h->B, "the column-sharded mode is served by schedule=batched");
    return bengine_attach_comm(h->B, unique_id, n_bytes, rank, nranks, nccl_path);
}
int mused_swfd_comm_stats(mused_swfd_t* h, uint64_t* n_allreduce, uint64_t* bytes) {
    MUSED_REQUIRE(h && h->B, "NULL argument / not a batched handle");
    unsigned long long a = 0, b = 0;
    bengine_comm_stats(h->B, &a, &b);
    if (n_allreduce) *n_allreduce = a;
    if (bytes) *bytes = b;
    return 0;
}
int mused_swfd_fit_timed(mused_swfd_t* h, const void* rows, int64_t n_rows, int64_t ld, int dtype,
                         int where, const uint64_t* times, void* stream) {
    MUSED_REQUIRE(h && (times || n_rows == 0), "NULL argument");
    MUSED_REQUIRE(!h->B, "explicit arrival times (TimeBasedSWFD) are served by schedule=reference");
    return engine_feed(*h->E, rows, n_rows, ld, dtype, where, MODE_SEQ, 0, (cudaStream_t)stream,
                       reinterpret_cast<const unsigned long long*>(times));
}
int mused_swfd_time(mused_swfd_t* h, uint64_t* t) {
    MUSED_REQUIRE(h && t, "NULL argument");
    *t = h->B ? bengine_time(h->B) : h->E->time;
    return 0;
}
// Checkpoint / resume (SURVEY 5 "state_dict"; the reference keeps its state in Python objects only and has no
// counterpart): the handle's persistent device buffers + host counters as one host blob.  A blob loads only into
// a handle created with the same configuration.
int mused_swfd_state_size(mused_swfd_t* h, int64_t* bytes) {
    MUSED_REQUIRE(h && bytes, "NULL argument");
    if (h->B) { long long b = 0; int rc = bengine_state(h->B, 0, nullptr, 0, &b); *bytes = b; return rc; }
    *bytes = state_blob_bytes(h->E->allocs);
    return 0;
}
int mused_swfd_state_save(mused_swfd_t* h, void* buf, int64_t cap) {
    MUSED_REQUIRE(h && buf, "NULL argument");
    if (h->B) return bengine_state(h->B, 1, buf, cap, nullptr);
    Engine& E = *h->E;
    MUSED_CUDA(cudaSetDevice(E.device));
    const unsigned long long host[6] = {E.time, 0ull, 0ull, 0ull, 0ull, 0ull};
    return state_blob_save(0u, host, E.allocs, buf, cap);
}
int mused_swfd_state_load(mused_swfd_t* h, const void* buf, int64_t bytes) {
    MUSED_REQUIRE(h && buf, "NULL argument");
    if (h->B) return bengine_state(h->B, 2, const_cast<void*>(buf), bytes, nullptr);
    Engine& E = *h->E;
    MUSED_CUDA(cudaSetDevice(E.device));
    unsigned long long host[6] = {0, 0, 0, 0, 0, 0};
    if (int rc = state_blob_load(0u, host, E.allocs, buf, bytes)) return rc;
    E.time = host[0];
    return 0;
}
int mused_swfd_profile(mused_swfd_t* h, uint64_t* cycles, int64_t n_inst_cap) {
    MUSED_REQUIRE(h && cycles, "NULL argument");
    MUSED_REQUIRE(!h->B, "mused_swfd_profile: per-phase cycles exist for schedule=reference only");
    std::vector<Slot> sl;
    Engine& E = *h->E;
    sl.resize(E.D.S);
    if (E.D.S) {
        MUSED_CUDA(cudaSetDevice(E.device));
        MUSED_CUDA(cudaMemcpy(sl.data(), E.D.slots, sizeof(Slot) * E.D.S, cudaMemcpyDeviceToHost));
    }
    MUSED_REQUIRE(n_inst_cap >= E.D.S, "n_inst_cap too small");
    for (int s = 0; s < E.D.S; ++s)
        for (int k = 0; k < 16; ++k) cycles[(size_t)s * 16 + k] = sl[s].prof[k];
    return 0;
}
int mused_swfd_timing(mused_swfd_t* h, int enable, double* ms, int64_t* launches, int64_t* rows) {
    MUSED_REQUIRE(h, "NULL handle");
    if (h->B) {
        double k[5];
        long long c = 0, r = 0;
        int rc = bengine_timing(h->B, enable, k, &c, &r);
        if (rc) return rc;
        if (ms) *ms = k[0] + k[1] + k[2] + k[3] + k[4];
        if (launches) *launches = c;
        if (rows) *rows = r;
        return 0;
    }
    Engine& E = *h->E;
    MUSED_CUDA(cudaSetDevice(E.device));
    for (size_t i = 0; i + 1 < E.ev.size(); i += 2) {
        float t = 0.f;
        MUSED_CUDA(cudaEventSynchronize(E.ev[i + 1]));
        MUSED_CUDA(cudaEventElapsedTime(&t, E.ev[i], E.ev[i + 1]));
        E.timed_ms += t;
        cudaEventDestroy(E.ev[i]);
        cudaEventDestroy(E.ev[i + 1]);
    }
    E.ev.clear();
    if (ms) *ms = E.timed_ms;
    if (launches) *launches = E.timed_launches;
    if (rows) *rows = E.timed_rows;
    if (enable >= 0) {
        E.timing = enable != 0;
        E.timed_ms = 0.0;
        E.timed_launches = 0;
        E.timed_rows = 0;
    }
    return 0;
}
int mused_swfd_kernel_gaps(mused_swfd_t* h, double* gaps2) {
    MUSED_REQUIRE(h && h->B && gaps2, "mused_swfd_kernel_gaps needs a schedule=batched handle");
    return bengine_gaps(h->B, gaps2);
}
int mused_swfd_kernel_ms(mused_swfd_t* h, int enable, double* ms5, int64_t* chunks, int64_t* rows) {
    MUSED_REQUIRE(h && h->B, "mused_swfd_kernel_ms needs a schedule=batched handle");
    long long c = 0, r = 0;
    int rc = bengine_timing(h->B, enable, ms5, &c, &r);
    if (chunks) *chunks = c;
    if (rows) *rows = r;
    return rc;
}
// ---- phase API of the column-sharded mode (batched schedule only) -----------------
#define NEED_B(h) MUSED_REQUIRE((h) && (h)->B, "this call needs a schedule=batched handle")
int mused_swfd_chunk_room(mused_swfd_t* h, int64_t* room, int64_t* pending) {
    NEED_B(h);
    if (room) *room = bengine_room(h->B);
    if (pending) *pending = bengine_pending(h->B);
    return 0;
}
int mused_swfd_stage(mused_swfd_t* h, const void* rows, int64_t n_rows, int64_t ld, int dtype, int where,
                     void* stream) {
    NEED_B(h);
    return bengine_stage(h->B, rows, n_rows, ld, dtype, where, (cudaStream_t)stream);
}
int mused_swfd_chunk_gram(mused_swfd_t* h, void* stream) {
    NEED_B(h);
    return bengine_chunk_gram(h->B, (cudaStream_t)stream);
}
int mused_swfd_chunk_finish(mused_swfd_t* h, void* stream) {
    NEED_B(h);
    return bengine_chunk_finish(h->B, (cudaStream_t)stream);
}
int mused_swfd_exchange_buffers(mused_swfd_t* h, double** norm2, int64_t* norm2_count, double** gram,
                                int64_t* gram_count, double** query_gram, int64_t* query_gram_count) {
    NEED_B(h);
    long long gc = 0, qc = 0;
    int rc = bengine_buffers(h->B, norm2, gram, &gc, query_gram, &qc);
    if (norm2_count) *norm2_count = bengine_ell(h->B);
    if (gram_count) *gram_count = gc;
    if (query_gram_count) *query_gram_count = qc;
    return rc;
}
int mused_swfd_get_begin(mused_swfd_t* h, int32_t* level, int32_t* n_stack, void* stream) {
    NEED_B(h);
    int lv = 0, ng = 0;
    int rc = bengine_get_begin(h->B, &lv, &ng, (cudaStream_t)stream);
    if (level) *level = lv;
    if (n_stack) *n_stack = ng;
    return rc;
}
int mused_swfd_get_finish(mused_swfd_t* h, double* sketch, double* sigma2, double* vt, double* delta,
                          int where, void* stream) {
    NEED_B(h);
    MUSED_REQUIRE(sketch && sigma2, "NULL argument");
    return bengine_get_finish(h->B, sketch, sigma2, vt, delta, where, (cudaStream_t)stream);
}
#undef NEED_B

int mused_swfd_queue_capacity(mused_swfd_t* h, int64_t* q) {
    MUSED_REQUIRE(h && q, "NULL argument");
    *q = h->B ? bengine_qcap(h->B) : h->E->D.capn;
    return 0;
}

static int read_slots(Engine& E, std::vector<Slot>& sl) {
    sl.resize(E.D.S);
    if (E.D.S == 0) return 0;
    MUSED_CUDA(cudaSetDevice(E.device));
    MUSED_CUDA(cudaMemcpy(sl.data(), E.D.slots, sizeof(Slot) * E.D.S, cudaMemcpyDeviceToHost));
    return 0;
}

int mused_swfd_get(mused_swfd_t* h, double* sketch, double* sigma2, double* vt, double* delta,
                   int32_t* level, int where, void* stream) {
    MUSED_REQUIRE(h && sketch && sigma2, "NULL argument");
    if (h->B) return bengine_get(h->B, 0, sketch, sigma2, vt, delta, level, where, (cudaStream_t)stream);
    Engine& E = *h->E;
    if (E.n_levels == 0) {  // the reference raises KeyError: 0 here
        set_error("SeqBasedSWFD has zero levels (R <= 1): get() is undefined");
        return MUSED_ERR_STATE;
    }
    MUSED_CUDA(cudaSetDevice(E.device));
    cudaStream_t st = (cudaStream_t)stream;
    MUSED_CUDA(cudaStreamSynchronize(st));
    std::vector<Slot> sl;
    int rc = read_slots(E, sl);
    if (rc) return rc;
    // level choice, seq_based_swfd.py:83-98 (uint64 arithmetic as np.uint64)
    int rj = 0;
    const unsigned long long time = E.time, N = E.D.N;
    for (int j = 0; j < E.n_levels; ++j) {
        const int s = fd_slot(E, j);
        if (sl[s].q_len == 0) break;
        unsigned long long head_s = 0;
        MUSED_CUDA(cudaMemcpy(&head_s, E.D.ringS + (size_t)s * E.D.capn + sl[s].q_head, 8,
                              cudaMemcpyDeviceToHost));
        const unsigned long long lhs = time - head_s;
        const unsigned long long rhs = std::min(N - 1ull, time - 1ull);
        if (lhs >= rhs) { rj = j; break; }
    }
    if (level) *level = rj;
    return engine_get(E, fd_slot(E, rj), true, sketch, sigma2, vt, delta, where, st);
}

int mused_swfd_size(mused_swfd_t* h, int64_t* n) {
    MUSED_REQUIRE(h && n, "NULL argument");
    if (h->B) { long long v = 0; int rc = bengine_size(h->B, 0, &v); *n = v; return rc; }
    std::vector<Slot> sl;
    int rc = read_slots(*h->E, sl);
    if (rc) return rc;
    int64_t tot = 0;
    for (int j = 0; j < h->E->n_levels; ++j)
        tot += sl[2 * j].q_len + sl[2 * j + 1].q_len + 2 * (int64_t)h->E->D.ell;
    *n = tot;
    return 0;
}

int mused_swfd_debug(mused_swfd_t* h, double* scal, double* sigma2, uint64_t* q_s, uint64_t* q_t,
                     int64_t q_cap) {
    MUSED_REQUIRE(h, "NULL handle");
    if (h->B) return bengine_debug(h->B, 0, scal, sigma2, (unsigned long long*)q_s, (unsigned long long*)q_t, q_cap);
    Engine& E = *h->E;
    std::vector<Slot> sl;
    int rc = read_slots(E, sl);
    if (rc) return rc;
    const int capn = E.D.capn, ell = E.D.ell;
    std::vector<unsigned long long> rs(capn), rt(capn);
    for (int j = 0; j < E.n_levels; ++j) {
        for (int side = 0; side < 2; ++side) {
            const int fd = fd_slot(E, j);
            const int s = side == 0 ? fd : (fd ^ 1);
            const int o = 2 * j + side;
            if (scal) {
                double* r = scal + (size_t)o * 8;
                r[0] = sl[s].energy; r[1] = sl[s].delta; r[2] = sl[s].b;
                r[3] = (double)sl[s].last_dump; r[4] = sl[s].q_len;
                r[5] = (double)sl[s].n_flush; r[6] = (double)sl[s].n_dump; r[7] = (double)sl[s].n_raw;
            }
            if (sigma2)
                MUSED_CUDA(cudaMemcpy(sigma2 + (size_t)o * ell, E.D.vec + ((size_t)s * NV + V_S2) * E.D.ldq,
                                      sizeof(double) * ell, cudaMemcpyDeviceToHost));
            if (q_s || q_t) {
                MUSED_REQUIRE(q_cap >= sl[s].q_len, "q_cap too small");
                MUSED_CUDA(cudaMemcpy(rs.data(), E.D.ringS + (size_t)s * capn, 8 * (size_t)capn, cudaMemcpyDeviceToHost));
                MUSED_CUDA(cudaMemcpy(rt.data(), E.D.ringT + (size_t)s * capn, 8 * (size_t)capn, cudaMemcpyDeviceToHost));
                for (int e = 0; e < sl[s].q_len; ++e) {
                    const int idx = (sl[s].q_head + e) % capn;
                    if (q_s) q_s[(size_t)o * q_cap + e] = rs[idx];
                    if (q_t) q_t[(size_t)o * q_cap + e] = rt[idx];
                }
            }
        }
    }
    return 0;
}

// ---- single FD instance ------------------------------------------------------
int mused_fd_create(mused_fd_t** out, int64_t d, int64_t ell, double error, int device, int cluster) {
    MUSED_REQUIRE(out, "NULL argument");
    Engine* E = nullptr;
    int rc = engine_create(&E, 1, d, ell, 1.0, 0, nullptr, device, cluster, 0, true, error);
    if (rc) return rc;
    *out = new mused_fd{E};
    return 0;
}
int mused_fd_destroy(mused_fd_t* h) {
    if (!h) return 0;
    cudaSetDevice(h->E->device);
    delete h->E;
    delete h;
    return 0;
}
int mused_fd_fit(mused_fd_t* h, const void* rows, int64_t n_rows, int64_t ld, int dtype, int where,
                 void* stream) {
    MUSED_REQUIRE(h, "NULL handle");
    return engine_feed(*h->E, rows, n_rows, ld, dtype, where, MODE_FD, 0, (cudaStream_t)stream);
}
int mused_fd_fit_batches(mused_fd_t* h, const void* rows, int64_t n_rows, int64_t ld, int dtype,
                         int where, int64_t batch_size, void* stream) {
    MUSED_REQUIRE(h, "NULL handle");
    MUSED_REQUIRE(batch_size >= 1 && batch_size <= h->E->D.ell,
                  "batch_size must be in [1, sketch_dim] for the B200 engine, got %lld", (long long)batch_size);
    return engine_feed(*h->E, rows, n_rows, ld, dtype, where, MODE_FD_BATCH, batch_size, (cudaStream_t)stream);
}
int mused_fd_flush(mused_fd_t* h, void* stream) {
    MUSED_REQUIRE(h, "NULL handle");
    MUSED_CUDA(cudaSetDevice(h->E->device));
    int rc = flush_slot(*h->E, 0, (cudaStream_t)stream);
    if (rc) return rc;
    MUSED_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    return 0;
}
int mused_fd_dump(mused_fd_t* h, double* v, int32_t* dumped, int where, void* stream) {
    MUSED_REQUIRE(h && v && dumped, "NULL argument");
    Engine& E = *h->E;
    cudaStream_t st = (cudaStream_t)stream;
    MUSED_CUDA(cudaSetDevice(E.device));
    RowArgs A = {};
    A.mode = MODE_DUMP;
    A.slot0 = 0;
    A.out = E.o_vt;
    A.out_flag = E.o_flag;
    int rc = launch_rows(E, A, 1, st);
    if (rc) return rc;
    int flag = 0;
    MUSED_CUDA(cudaMemcpyAsync(&flag, E.o_flag, sizeof(int), cudaMemcpyDeviceToHost, st));
    MUSED_CUDA(cudaStreamSynchronize(st));
    *dumped = flag;
    if (flag) {
        rc = copy_out(v, E.o_vt, (size_t)E.D.d, where, st);
        if (rc) return rc;
        MUSED_CUDA(cudaStreamSynchronize(st));
    }
    return 0;
}
int mused_fd_get(mused_fd_t* h, double* sketch, double* sigma2, double* vt, double* delta, int where,
                 void* stream) {
    MUSED_REQUIRE(h && sketch && sigma2, "NULL argument");
    MUSED_CUDA(cudaSetDevice(h->E->device));
    return engine_get(*h->E, 0, false, sketch, sigma2, vt, delta, where, (cudaStream_t)stream);
}

// FD merge / shrink of an arbitrary stack of rows (the thin SVD + shrink + top-l cut that LMFD.__merge, lmfd.py:27-47,
// DIFD, difd.py:117-127, FastFrequentDirectionsWithDump.fit, fastfdwithdump.py:98-108, and SlidingWindowFD.get,
// sliding_windows_frequent_directions.py:160-171, all spell out with a LAPACK thin SVD): rows (ng x d float64, device,
// leading dimension ld) -> sketch (l x d: sqrt(s_i^2 - s_l^2) v_i^T, zero rows beyond the rank), sigma2 (l), vt (l x d,
// nullable); s_l^2 is subtracted only when ng > l, as in the reference.  Scratch is allocated per call.
int mused_fd_merge(const double* rows, int64_t ng, int64_t d, int64_t ld, int64_t ell, double* sketch, double* sigma2,
                   double* vt, void* stream) {
    MUSED_REQUIRE(rows && sketch && sigma2, "NULL argument");
    MUSED_REQUIRE(ng >= 1 && d >= 1 && ld >= d && ell >= 1, "bad shape ng=%lld d=%lld l=%lld", (long long)ng, (long long)d, (long long)ell);
    cudaStream_t st = (cudaStream_t)stream;
    int dev = 0;
    MUSED_CUDA(cudaGetDevice(&dev));
    int sm = 148;
    MUSED_CUDA(cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, dev));
    QueryScratch Q;
    Q.ng_max = (int)std::max<int64_t>(ng, d) + 2;
    Q.ng_max += Q.ng_max & 1;
    Q.lda = Q.ng_max;
    Q.rel_thr = 1e-11;
    Q.plan_big();
    std::vector<void*> allocs;
    auto fail = [&](int code) { for (void* p : allocs) cudaFree(p); return code; };
    auto alloc = [&](void** p, size_t bytes) -> int {
        if (cudaMalloc(p, bytes) != cudaSuccess) { set_error("mused_fd_merge: cudaMalloc of %zu bytes failed", bytes); return MUSED_ERR_NOMEM; }
        allocs.push_back(*p);
        cudaMemsetAsync(*p, 0, bytes, st);
        return 0;
    };
    int rc = 0;
    double* out_s = nullptr; double* out_v = nullptr; double* out_2 = nullptr;
    if ((rc = alloc((void**)&Q.M, sizeof(double) * (size_t)Q.ng_max * d)) || (rc = alloc((void**)&Q.G, sizeof(double) * (size_t)Q.lda * Q.lda)) ||
        (rc = alloc((void**)&Q.Vq, sizeof(double) * (size_t)Q.lda * Q.lda)) || (rc = alloc((void**)&Q.cs, sizeof(double) * Q.ng_max)) ||
        (rc = alloc((void**)&Q.eval, sizeof(double) * Q.npad_max)) || (rc = alloc((void**)&Q.perm, sizeof(int) * Q.npad_max)) ||
        (rc = alloc((void**)&Q.lam, sizeof(double) * Q.npad_max)) || (rc = alloc((void**)&Q.VT, sizeof(double) * (size_t)Q.npad_max * Q.ldt)) ||
        (rc = alloc((void**)&Q.mailL, sizeof(double) * Q.mail_count)) || (rc = alloc((void**)&Q.mailR, sizeof(double) * Q.mail_count)) ||
        (rc = alloc((void**)&Q.sync, sizeof(unsigned long long) * QSYNC_WORDS)) || (rc = alloc((void**)&Q.counters, sizeof(int) * 64)) ||
        (rc = alloc((void**)&Q.scr, sizeof(double) * 8)) || (rc = alloc((void**)&out_s, sizeof(double) * (size_t)ell * d)) ||
        (rc = alloc((void**)&out_v, sizeof(double) * (size_t)ell * d)) || (rc = alloc((void**)&out_2, sizeof(double) * (size_t)ell)))
        return fail(rc);
    if (cudaMemcpy2DAsync(Q.M, (size_t)d * 8, rows, (size_t)ld * 8, (size_t)d * 8, (size_t)ng, cudaMemcpyDeviceToDevice, st) != cudaSuccess) {
        set_error("mused_fd_merge: copying the stack failed");
        return fail(MUSED_ERR_CUDA);
    }
    rc = merge_query(Q, (int)ng, d, (int)ell, sm, out_s, out_v, out_2, st);
    if (rc) return fail(rc);
    cudaMemcpyAsync(sketch, out_s, sizeof(double) * (size_t)ell * d, cudaMemcpyDeviceToDevice, st);
    cudaMemcpyAsync(sigma2, out_2, sizeof(double) * (size_t)ell, cudaMemcpyDeviceToDevice, st);
    if (vt) cudaMemcpyAsync(vt, out_v, sizeof(double) * (size_t)ell * d, cudaMemcpyDeviceToDevice, st);
    cudaError_t e = cudaStreamSynchronize(st);
    for (void* p : allocs) cudaFree(p);
    if (e != cudaSuccess) { set_error("mused_fd_merge failed: %s", cudaGetErrorString(e)); return MUSED_ERR_CUDA; }
    return 0;
}

int mused_sym_eig(const double* G, int64_t n, int64_t ld, double* evals, double* vecs, int64_t ld_vecs,
                  int32_t* sweeps, void* stream) {
    MUSED_REQUIRE(G && evals && vecs, "NULL argument");
    MUSED_REQUIRE(n >= 1 && ld >= n && ld_vecs >= n, "bad shape n=%lld ld=%lld", (long long)n, (long long)ld);
    cudaStream_t st = (cudaStream_t)stream;
    int dev = 0;
    MUSED_CUDA(cudaGetDevice(&dev));
    cudaDeviceProp prop;
    MUSED_CUDA(cudaGetDeviceProperties(&prop, dev));
    QueryScratch Q;
    Q.ng_max = (int)n + ((int)n & 1);
    Q.lda = (int)ld;
    Q.plan_big();
    std::vector<void*> allocs;
    auto alloc = [&](void** p, size_t bytes) -> int {
        MUSED_CUDA(cudaMalloc(p, bytes));
        allocs.push_back(*p);
        return 0;
    };
    int rc = 0;
    auto fail = [&](int code) { for (void* p : allocs) cudaFree(p); return code; };
    if ((rc = alloc((void**)&Q.eval, sizeof(double) * Q.npad_max))) return fail(rc);
    if ((rc = alloc((void**)&Q.perm, sizeof(int) * Q.npad_max))) return fail(rc);
    if ((rc = alloc((void**)&Q.lam, sizeof(double) * Q.npad_max))) return fail(rc);
    if ((rc = alloc((void**)&Q.VT, sizeof(double) * (size_t)Q.npad_max * Q.ldt))) return fail(rc);
    if ((rc = alloc((void**)&Q.mailL, sizeof(double) * Q.mail_count))) return fail(rc);
    if ((rc = alloc((void**)&Q.mailR, sizeof(double) * Q.mail_count))) return fail(rc);
    if ((rc = alloc((void**)&Q.sync, sizeof(unsigned long long) * QSYNC_WORDS))) return fail(rc);
    Q.G = const_cast<double*>(G);
    bool done = false;
    int npad = 0;
    const int sm = prop.multiProcessorCount;
    rc = dispatch_eig_big(Q, (int)n, sm, st, &npad, &done);
    if (rc) return fail(rc);
    if (!done) {
        set_error("mused_sym_eig: n=%lld does not fit the shared memory of %d SMs", (long long)n, sm);
        return fail(MUSED_ERR_ARG);
    }
    k_sort_eigs<<<(npad + 255) / 256, 256, 0, st>>>(Q.lam, npad, 0, Q.perm, Q.eval);
    k_eig_permute_out<<<std::min((int)n, 4 * sm), 256, 0, st>>>(Q.VT, Q.ldt, Q.perm, Q.eval, (int)n, evals, vecs,
                                                               (long long)ld_vecs);
    COUNT_LAUNCH(3);
    int info[2] = {0, 0};
    cudaError_t e = cudaMemcpyAsync(info, Q.sync + 640, sizeof(info), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) {
        set_error("mused_sym_eig failed: %s", cudaGetErrorString(e));
        return fail(MUSED_ERR_CUDA);
    }
    if (sweeps) *sweeps = info[0];
    if (getenv("MUSED_EIG_DEBUG")) {
        unsigned long long h[QSYNC_WORDS];
        cudaMemcpy(h, Q.sync, sizeof(h), cudaMemcpyDeviceToHost);
        fprintf(stderr, "[eig_big] n=%lld npad=%d sweeps=%d  cosines:", (long long)n, npad, info[0]);
        for (int k = 0; k < info[0] && k < 64; ++k) {
            double v;
            memcpy(&v, &h[512 + k], 8);
            fprintf(stderr, " %.1e", v);
        }
        fprintf(stderr, "\n[eig_big] Mcycles gram %.2f small %.2f apply %.2f out %.2f wait %.2f in %.2f\n", h[700] * 1e-6,
                h[701] * 1e-6, h[702] * 1e-6, h[703] * 1e-6, h[704] * 1e-6, h[705] * 1e-6);
    }
    for (void* p : allocs) cudaFree(p);
    if (info[1]) {
        set_error("mused_sym_eig: a neighbour exchange timed out");
        return MUSED_ERR_STATE;
    }
    return 0;
}

}  // extern "C"

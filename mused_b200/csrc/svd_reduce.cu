// Randomized truncated SVD of the fused adjacency matrix on the device: perform_svd_reduction
// (matrix_operations.py:143-147 = sklearn TruncatedSVD(n_components, random_state=seed).fit_transform), the reducer of
// the reference's sSVDMC approaches (main.py:79, :146).
//
// sklearn's algorithm (randomized_svd with n_iter power iterations, n_oversamples extra columns): start from a
// Gaussian test matrix Omega (n x q, q = n_components + n_oversamples; drawn by the CALLER with numpy's generator so
// that the same seed gives the same Omega as sklearn), Q <- M Q, Q <- M^T Q (n_iter times, re-conditioned after every
// product), Q <- orth(M Q), B = Q^T M, SVD of the small B, V^T = its right singular vectors with the sign of the
// largest |entry| of each row made positive (svd_flip, u_based_decision=False), result M V.  The range of Q after the
// iterations does not depend on HOW the basis is re-conditioned between products (sklearn: LU; here: Cholesky-QR,
// twice at the end), so the result is sklearn's up to rounding.
//
// Everything is GEMM-shaped on a 0/1 uint8 (or float64) matrix in float64; W = 2 000, q = 60: 11 products of 0.5
// GFLOP -- a tiled SIMT float64 kernel, one generic routine for every product (transposes, uint8 operand), the small
// Cholesky in one CTA, the q x q eigenproblem of B B^T through mused_sym_eig.
#include <stdlib.h>
#include "common.cuh"

namespace mused {

constexpr int SGT = 64;    // C tile
constexpr int SGK = 16;    // k chunk

template <class TA>
__device__ __forceinline__ double ld_a(const TA* A, long long lda, bool ta, long long i, long long kk) {
    return (double)(ta ? A[(size_t)kk * lda + i] : A[(size_t)i * lda + kk]);
}

// C (m x n, ldc) = op(A) (m x k) * op(B) (k x n) [* diag(colscale)], float64 accumulate; A may be uint8 (the adjacency)
template <class TA>
__global__ void __launch_bounds__(256) k_dgemm(const TA* __restrict__ A, long long lda, int ta, const double* __restrict__ B,
                                               long long ldb, int tb, double* __restrict__ C, long long ldc, long long m,
                                               long long n, long long k, const double* __restrict__ colscale) {
    __shared__ double sa[SGK][SGT + 1];
    __shared__ double sb[SGK][SGT + 1];
    const int t = threadIdx.x, ty = t >> 4, tx = t & 15;
    const long long r0 = (long long)blockIdx.y * SGT, c0 = (long long)blockIdx.x * SGT;
    double acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.0;
    for (long long k0 = 0; k0 < k; k0 += SGK) {
        for (int e = t; e < SGT * SGK; e += 256) {
            // A tile: consecutive threads along the contiguous dimension of the operand
            int ii, kk;
            if (ta) { ii = e % SGT; kk = e / SGT; } else { kk = e % SGK; ii = e / SGK; }
            const long long gi = r0 + ii, gk = k0 + kk;
            sa[kk][ii] = (gi < m && gk < k) ? ld_a<TA>(A, lda, ta != 0, gi, gk) : 0.0;
            int jj, kb;
            if (tb) { kb = e % SGK; jj = e / SGK; } else { jj = e % SGT; kb = e / SGT; }
            const long long gj = c0 + jj, gkb = k0 + kb;
            sb[kb][jj] = (gj < n && gkb < k) ? (tb ? B[(size_t)gj * ldb + gkb] : B[(size_t)gkb * ldb + gj]) : 0.0;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < SGK; ++kk) {
            double av[4], bv[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) av[i] = sa[kk][ty * 4 + i];
#pragma unroll
            for (int j = 0; j < 4; ++j) bv[j] = sb[kk][tx * 4 + j];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fma(av[i], bv[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const long long r = r0 + ty * 4 + i;
        if (r >= m) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const long long c = c0 + tx * 4 + j;
            if (c < n) C[(size_t)r * ldc + c] = colscale ? acc[i][j] * colscale[c] : acc[i][j];
        }
    }
}

// G (q x q, symmetric positive definite) -> Rinv = inverse of its upper Cholesky factor (G = R^T R), one CTA.
// A pivot at round-off level (rank-deficient sample) gets a zero column: that direction drops out of the basis.
__global__ void __launch_bounds__(256) k_chol_inv(const double* __restrict__ G, int q, double* __restrict__ Rinv, int* flag) {
    extern __shared__ double sm[];
    double* R = sm;              // q x q, upper
    __shared__ double piv;
    const int t = threadIdx.x;
    for (int e = t; e < q * q; e += 256) { R[e] = G[e]; Rinv[e] = 0.0; }
    __syncthreads();
    double dmax = 0.0;
    for (int i = 0; i < q; ++i) dmax = fmax(dmax, G[i * q + i]);
    for (int j = 0; j < q; ++j) {
        // row j of R: R[j][c] = (G[j][c] - sum_{p<j} R[p][j] R[p][c]) / R[j][j]
        if (t == 0) {
            double s = R[j * q + j];
            for (int p = 0; p < j; ++p) s -= R[p * q + j] * R[p * q + j];
            if (s <= 1e-14 * dmax) { s = 0.0; if (flag) atomicAdd(flag, 1); }
            piv = sqrt(s);
        }
        __syncthreads();
        const double pv = piv;
        for (int c = j + 1 + t; c < q; c += 256) {
            double s = R[j * q + c];
            for (int p = 0; p < j; ++p) s -= R[p * q + j] * R[p * q + c];
            R[j * q + c] = pv > 0.0 ? s / pv : 0.0;
        }
        __syncthreads();
        if (t == 0) R[j * q + j] = pv;
        __syncthreads();
    }
    // Rinv = R^-1 (upper): column c by back substitution, one thread per column (it re-reads only its own column)
    for (int c = t; c < q; c += 256) {
        for (int i = c; i >= 0; --i) {
            double s = (i == c) ? 1.0 : 0.0;
            for (int p = i + 1; p <= c; ++p) s -= R[i * q + p] * Rinv[p * q + c];
            const double d = R[i * q + i];
            Rinv[i * q + c] = d > 0.0 ? s / d : 0.0;
        }
    }
}

// per row of VT (k x n): sign of its largest |entry| (first on ties, like numpy argmax) and 1 / sigma
__global__ void __launch_bounds__(256) k_flip_scale(const double* __restrict__ VT, long long ld, long long n, const double* __restrict__ ev,
                                                    double* __restrict__ rowscale, double* __restrict__ sigma) {
    const int r = blockIdx.x, t = threadIdx.x;
    double best = -1.0, val = 0.0;
    long long bi = 0x7fffffffffffffffll;
    for (long long c = t; c < n; c += 256) {
        const double v = VT[(size_t)r * ld + c], a = fabs(v);
        if (a > best) { best = a; val = v; bi = c; }
    }
    __shared__ double sb[256], sv[256];
    __shared__ long long si[256];
    sb[t] = best; sv[t] = val; si[t] = bi;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (t < o && (sb[t + o] > sb[t] || (sb[t + o] == sb[t] && si[t + o] < si[t]))) {
            sb[t] = sb[t + o]; sv[t] = sv[t + o]; si[t] = si[t + o];
        }
        __syncthreads();
    }
    if (t == 0) {
        const double s = sqrt(fmax(ev[r], 0.0));
        sigma[r] = s;
        rowscale[r] = (s > 0.0 ? 1.0 / s : 0.0) * (sv[0] < 0.0 ? -1.0 : 1.0);
    }
}

__global__ void k_scale_rows(double* __restrict__ VT, long long ld, long long n, const double* __restrict__ rowscale) {
    const long long c = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (c < n) VT[(size_t)blockIdx.y * ld + c] *= rowscale[blockIdx.y];
}

template <class TA>
static int dgemm(const TA* A, long long lda, int ta, const double* B, long long ldb, int tb, double* C, long long ldc,
                 long long m, long long n, long long k, const double* colscale, cudaStream_t st) {
    dim3 grid((unsigned)((n + SGT - 1) / SGT), (unsigned)((m + SGT - 1) / SGT));
    k_dgemm<TA><<<grid, 256, 0, st>>>(A, lda, ta, B, ldb, tb, C, ldc, m, n, k, colscale);
    MUSED_CUDA(cudaGetLastError());
    return 0;
}

struct SvdScratch {
    double *A1 = nullptr, *A2 = nullptr, *A3 = nullptr, *G = nullptr, *Rinv = nullptr, *ev = nullptr, *vec = nullptr,
           *VT = nullptr, *rs = nullptr;
    int* flag = nullptr;
    cudaStream_t st = nullptr;
    ~SvdScratch() {
        for (void* p : {(void*)A1, (void*)A2, (void*)A3, (void*)G, (void*)Rinv, (void*)ev, (void*)vec, (void*)VT, (void*)rs,
                        (void*)flag})
            if (p) cudaFreeAsync(p, st);
    }
};

}  // namespace mused

using namespace mused;
extern "C" void mused_count_launches(uint64_t n);

// out (rows x q) = in * Rinv with in^T in = R^T R: orthonormal columns (one Cholesky-QR pass)
static int chol_qr(const double* in, double* out, long long rows, int q, SvdScratch& S, cudaStream_t st) {
    if (int rc = dgemm<double>(in, q, 1, in, q, 0, S.G, q, q, q, rows, nullptr, st)) return rc;     // G = in^T in
    const size_t smem = sizeof(double) * (size_t)q * q;
    MUSED_SMEM_ATTR(k_chol_inv, sizeof(double) * 160 * 160);        // once per device: the largest q (mused_svd_reduce)
    k_chol_inv<<<1, 256, smem, st>>>(S.G, q, S.Rinv, S.flag);
    MUSED_CUDA(cudaGetLastError());
    if (int rc = dgemm<double>(in, q, 0, S.Rinv, q, 0, out, q, rows, q, q, nullptr, st)) return rc;
    mused_count_launches(3);
    return 0;
}

template <class TM>
static int svd_reduce(const TM* M, int64_t m, int64_t n, int64_t ld, const double* omega, int q, int k, int n_iter,
                      double* Xt, int64_t ld_xt, double* sigma, double* vt_out, int64_t ld_vt, cudaStream_t st) {
    SvdScratch S;
    S.st = st;
    const size_t big = (size_t)(m > n ? m : n) * q;
    MUSED_CUDA(cudaMallocAsync((void**)&S.A1, sizeof(double) * big, st));
    MUSED_CUDA(cudaMallocAsync((void**)&S.A2, sizeof(double) * big, st));
    MUSED_CUDA(cudaMallocAsync((void**)&S.A3, sizeof(double) * big, st));
    MUSED_CUDA(cudaMallocAsync((void**)&S.G, sizeof(double) * (size_t)q * q, st));
    MUSED_CUDA(cudaMallocAsync((void**)&S.Rinv, sizeof(double) * (size_t)q * q, st));
    MUSED_CUDA(cudaMallocAsync((void**)&S.ev, sizeof(double) * (size_t)q, st));
    MUSED_CUDA(cudaMallocAsync((void**)&S.vec, sizeof(double) * (size_t)q * q, st));
    MUSED_CUDA(cudaMallocAsync((void**)&S.VT, sizeof(double) * (size_t)k * n, st));
    MUSED_CUDA(cudaMallocAsync((void**)&S.rs, sizeof(double) * (size_t)q, st));
    MUSED_CUDA(cudaMallocAsync((void**)&S.flag, sizeof(int), st));
    MUSED_CUDA(cudaMemsetAsync(S.flag, 0, sizeof(int), st));
    // range finder (sklearn.utils.extmath.randomized_range_finder): Q = Omega (n x q)
    const double* Q = omega;
    for (int it = 0; it < n_iter; ++it) {
        if (int rc = dgemm<TM>(M, ld, 0, Q, q, 0, S.A1, q, m, q, n, nullptr, st)) return rc;        // M Q      (m x q)
        if (int rc = chol_qr(S.A1, S.A2, m, q, S, st)) return rc;
        if (int rc = dgemm<TM>(M, ld, 1, S.A2, q, 0, S.A1, q, n, q, m, nullptr, st)) return rc;     // M^T (.)  (n x q)
        if (int rc = chol_qr(S.A1, S.A3, n, q, S, st)) return rc;
        Q = S.A3;
        mused_count_launches(2);
    }
    if (int rc = dgemm<TM>(M, ld, 0, Q, q, 0, S.A1, q, m, q, n, nullptr, st)) return rc;            // M Q
    if (int rc = chol_qr(S.A1, S.A2, m, q, S, st)) return rc;
    if (int rc = chol_qr(S.A2, S.A1, m, q, S, st)) return rc;                                       // twice: orthonormal to round-off
    // B = Q^T M (q x n), kept as B^T = M^T Q (n x q);  B B^T = Uh diag(s^2) Uh^T;  V^T = diag(1/s) Uh^T B
    if (int rc = dgemm<TM>(M, ld, 1, S.A1, q, 0, S.A2, q, n, q, m, nullptr, st)) return rc;
    if (int rc = dgemm<double>(S.A2, q, 1, S.A2, q, 0, S.G, q, q, q, n, nullptr, st)) return rc;
    if (int rc = mused_sym_eig(S.G, q, q, S.ev, S.vec, q, nullptr, (void*)st)) return rc;           // rows of vec = Uh^T
    if (int rc = dgemm<double>(S.vec, q, 0, S.A2, q, 1, S.VT, n, k, n, q, nullptr, st)) return rc;  // Uh^T B   (k x n)
    k_flip_scale<<<k, 256, 0, st>>>(S.VT, n, n, S.ev, S.rs, sigma ? sigma : S.ev + 0);
    MUSED_CUDA(cudaGetLastError());
    k_scale_rows<<<dim3((unsigned)((n + 255) / 256), (unsigned)k), 256, 0, st>>>(S.VT, n, n, S.rs);
    MUSED_CUDA(cudaGetLastError());
    // the reduced matrix: M V (TruncatedSVD.fit_transform returns X @ components_.T)
    if (int rc = dgemm<TM>(M, ld, 0, S.VT, n, 1, Xt, ld_xt, m, k, n, nullptr, st)) return rc;
    if (vt_out) MUSED_CUDA(cudaMemcpy2DAsync(vt_out, sizeof(double) * ld_vt, S.VT, sizeof(double) * n, sizeof(double) * n, k,
                                             cudaMemcpyDeviceToDevice, st));
    mused_count_launches(7);
    return 0;
}

extern "C" int mused_svd_reduce(const void* M, int dtype, int64_t m, int64_t n, int64_t ld, const double* omega, int32_t q,
                                int32_t k, int32_t n_iter, double* Xt, int64_t ld_xt, double* sigma, double* vt, int64_t ld_vt,
                                void* stream) {
    MUSED_REQUIRE(M && omega && Xt, "NULL argument");
    MUSED_REQUIRE(m >= 1 && n >= 2 && ld >= n, "bad shape");
    MUSED_REQUIRE(m >= n, "mused_svd_reduce takes matrices with at least as many rows as columns (the fused adjacency is square)");
    MUSED_REQUIRE(k >= 1 && k <= q && q <= n && q <= 160, "need 1 <= k <= q <= min(n, 160) (q = n_components + n_oversamples)");
    MUSED_REQUIRE(n_iter >= 0 && ld_xt >= k && (!vt || ld_vt >= n), "bad argument");
    MUSED_REQUIRE(dtype == MUSED_U8 || dtype == MUSED_F64, "matrix dtype must be MUSED_U8 or MUSED_F64");
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == MUSED_U8)
        return svd_reduce<unsigned char>((const unsigned char*)M, m, n, ld, omega, q, k, n_iter, Xt, ld_xt, sigma, vt, ld_vt, st);
    return svd_reduce<double>((const double*)M, m, n, ld, omega, q, k, n_iter, Xt, ld_xt, sigma, vt, ld_vt, st);
}

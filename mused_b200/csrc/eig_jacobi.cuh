// Batched symmetric eigensolve of the (2l x 2l) Gram matrices of the batched FD
// shrink, one thread-block cluster per FD instance, float64.
//
// The thin SVD of [sketch; batch] that the reference takes with LAPACK
// (swfd/frequent_directions.py:55) is replaced by the eigen-decomposition of
// G = R R^T.  One-sided (Hestenes) Jacobi on the columns of W = G: plane
// rotations from the right orthogonalise the columns, W V -> V Lambda, so column
// norms converge to the eigenvalues and w_j / lambda_j to the eigenvectors; no
// separate V is carried.  The whole matrix lives in REGISTERS: warp w owns the
// two columns at positions (2w, 2w+1), 2l/32 elements per lane.  Pairs follow the
// odd-even transposition ordering (every pair meets once in 2l rounds); a round
// moves one column per warp through (distributed) shared memory and costs one
// cluster barrier.
//
// Epilogue (frequent_directions.py:86-93, :58 and the multi-direction dump of
// fastfdwithdump.py:115-131): rank the eigenvalues, cut = lambda[l],
// sigma^2_r = lambda_r - cut, delta += cut/2, coefficient rows
// T[r] = sqrt(sigma^2_r / lambda_r) * v_r^T for the rotate GEMM, and ring slots
// for every direction with sigma^2 >= threshold.
#pragma once
#include <float.h>
#include "common.cuh"

namespace mused {
namespace cg = cooperative_groups;

constexpr int EIG_WARPS = 16;               // warps per CTA
constexpr int EIG_NT = EIG_WARPS * 32;

struct BSlot {                               // one FD instance of the batched engine
    double delta, error;
    unsigned long long last_dump, period;
    unsigned long long n_flush, n_dump, n_raw, overflow;
    int q_head, q_len;
    int cur, src;                            // ping-pong plane holding the sketch / source of the pending rotate
    int parity, skip, npend, nd;
    int sweeps, pad;
};

struct EigArgs {
    BSlot* slots;
    const float* Gp;         // [S][KS][n2][n2] split-K partial Grams
    int KS, ell, n2, capn;
    const unsigned char* mask;  // [S][ell] 1 = row of the chunk is NOT part of this instance's batch
    int n_rows;              // rows in the chunk
    unsigned long long t_end;
    float* T;                // [S][ell][n2] rotate coefficients
    double* s2;              // [S][ell]
    int* dump_slot;          // [S][ell] ring entry or -1
    unsigned long long* ringS;  // [S][capn]
    unsigned long long* ringT;
    double tol;              // convergence: max |x.y| / (|x||y|)
    int max_sweeps;
};

template <int EPL>
struct EigSmem {
    static constexpr int N2 = 32 * EPL;
    double slotA[EIG_WARPS][N2];
    double slotB[EIG_WARPS][N2];
    double lam[N2];           // eigenvalue per position (replicated in every CTA)
    double red_off[8 * EIG_WARPS];   // per (cta, warp) partial maxima, written into CTA 0
    double red_nmax[8 * EIG_WARPS];
    double normA[EIG_WARPS];         // squared norms travelling with slotA / slotB
    double normB[EIG_WARPS];
    double cut;
};

// One Hestenes rotation of the column pair (x, y) with squared norms (nx, ny);
// afterwards the two columns have traded places (odd-even transposition
// ordering), which is folded into the rotation so no register moves are spent.
// FP64 issue is the bottleneck of this kernel (64 DFMA/clk/SM), so the rotation
// ANGLE is taken in fp32 (MUFU rsqrt / sqrt, tangent good to ~1e-7) while the
// pair (c, s) is refined to float64 (two Newton steps on rsqrt) so that
// c^2 + s^2 = 1 to 1e-16 and W stays G times an orthogonal matrix.  The norms
// follow the rotation that was actually applied.
template <int EPL>
__device__ __forceinline__ void jacobi_pair(double (&x)[EPL], double (&y)[EPL], double& nx, double& ny,
                                            double nmax, double& off) {
    double g = 0.0;
#pragma unroll
    for (int e = 0; e < EPL; ++e) g = fma(x[e], y[e], g);
    g = warp_sum(g);
    const float gf = (float)g, nxf = (float)nx, nyf = (float)ny;
    float relf = 0.f;
    if (nxf > 0.f && nyf > 0.f) relf = fabsf(gf) * rsqrtf(nxf) * rsqrtf(nyf);
    if (nx * ny > 1e-24 * nmax * nmax) off = fmax(off, (double)relf);
    double c = 0.0, s = 1.0;            // pure swap: x_new = y, y_new = x
    double mx = ny, my = nx;
    if (relf > 1e-10f) {
        const float zf = (float)(ny - nx) / (2.f * gf);
        const float tf = copysignf(1.f, zf) / (fabsf(zf) + sqrtf(fmaf(zf, zf, 1.f)));
        const double t = (double)tf;
        const double u = fma(t, t, 1.0);
        double c0 = (double)rsqrtf((float)u);
        c0 = c0 * fma(-0.5 * u, c0 * c0, 1.5);
        c0 = c0 * fma(-0.5 * u, c0 * c0, 1.5);
        const double s0 = c0 * t;
        // rotated pair (c0 x - s0 y, s0 x + c0 y), stored swapped
        c = s0;
        s = c0;
        const double cc = c0 * c0, ss = s0 * s0, cs2 = 2.0 * c0 * s0 * g;
        mx = fma(ss, nx, fma(cc, ny, cs2));      // |s0 x + c0 y|^2
        my = fma(cc, nx, fma(ss, ny, -cs2));     // |c0 x - s0 y|^2
    }
    // x_new = c x + s y,  y_new = s x - c y   (swap: c = 0, s = 1 gives x_new = y, y_new = x)
    const bool rotated = relf > 1e-10f;
#pragma unroll
    for (int e = 0; e < EPL; ++e) {
        const double xv = x[e], yv = y[e];
        x[e] = fma(c, xv, s * yv);
        y[e] = rotated ? fma(s, xv, -c * yv) : xv;
    }
    nx = mx;
    ny = my;
}

// Grid = S * EPL CTAs, cluster = EPL CTAs, EIG_NT threads, dynamic smem = sizeof(EigSmem<EPL>).
template <int EPL>
__global__ void __launch_bounds__(EIG_NT, 1) k_eig_jacobi(EigArgs A) {
    constexpr int N2 = 32 * EPL;
    constexpr int ELL = 16 * EPL;
    constexpr int M = N2 / 2;                 // warps in the cluster
    extern __shared__ __align__(16) unsigned char smem_raw[];
    EigSmem<EPL>& sm = *reinterpret_cast<EigSmem<EPL>*>(smem_raw);
    cg::cluster_group cluster = cg::this_cluster();
    const int rank = (EPL == 1) ? 0 : (int)cluster.block_rank();
    const int s = blockIdx.x / EPL;
    const int lw = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int w = rank * EIG_WARPS + lw;
    BSlot& sl = A.slots[s];
    const int npend = sl.npend;
    if (npend == 0) {                          // nothing to shrink for this instance
        if (w == 0 && lane == 0) sl.skip = 1;
        return;
    }
    const int q_head = sl.q_head, q_len = sl.q_len;
    const unsigned long long last_dump = sl.last_dump;
    const double thr = sl.error;
    auto sync_all = [&]() {
        if (EPL == 1) __syncthreads();
        else cluster.sync();
    };

    // ---- load: columns 2w, 2w+1 of the symmetrised, masked Gram ------------------
    const unsigned char* mk = A.mask + (size_t)s * ELL;
    auto dead = [&](int i) -> bool {
        if (i < ELL) return false;
        const int r = i - ELL;
        return r >= A.n_rows || mk[r] != 0;
    };
    double a[EPL], b[EPL];
    {
        const float* G0 = A.Gp + (size_t)s * A.KS * N2 * N2;
        const int ja = 2 * w, jb = 2 * w + 1;
        const bool da = dead(ja), db = dead(jb);
#pragma unroll
        for (int e = 0; e < EPL; ++e) {
            const int i = lane + 32 * e;
            double va = 0.0, vb = 0.0;
            if (!dead(i)) {
                for (int k = 0; k < A.KS; ++k) {
                    const float* G = G0 + (size_t)k * N2 * N2;
                    va += (double)G[(size_t)ja * N2 + i] + (double)G[(size_t)i * N2 + ja];
                    vb += (double)G[(size_t)jb * N2 + i] + (double)G[(size_t)i * N2 + jb];
                }
            }
            a[e] = da ? 0.0 : 0.5 * va;
            b[e] = db ? 0.0 : 0.5 * vb;
        }
    }

    // ---- sweeps -------------------------------------------------------------------------
    double na = 0.0, nb = 0.0;
    int sweeps_done = 0;
    double off = 0.0;
    for (int sweep = 0; sweep < A.max_sweeps; ++sweep) {
        na = 0.0;
        nb = 0.0;
#pragma unroll
        for (int e = 0; e < EPL; ++e) {
            na = fma(a[e], a[e], na);
            nb = fma(b[e], b[e], nb);
        }
        na = warp_sum(na);
        nb = warp_sum(nb);
        // cluster-wide: largest column norm^2 and last sweep's largest off-diagonal cosine
        {
            double* ro = (EPL == 1) ? sm.red_off : cluster.map_shared_rank(sm.red_off, 0);
            double* rn = (EPL == 1) ? sm.red_nmax : cluster.map_shared_rank(sm.red_nmax, 0);
            if (lane == 0) {
                ro[w] = off;
                rn[w] = fmax(na, nb);
            }
            sync_all();
            double o = 0.0, nm = 0.0;
            for (int i = lane; i < M; i += 32) {
                o = fmax(o, ro[i]);
                nm = fmax(nm, rn[i]);
            }
            o = warp_max(o);
            nm = warp_max(nm);
            sync_all();                         // everyone has read before the arrays are rewritten
            if (sweep > 0 && o <= A.tol) break;
            off = 0.0;
            const double nmax = nm;
            // ---- one sweep = N2 rounds, two per iteration -----------------------------
            for (int it = 0; it < M; ++it) {
                // positions (2w, 2w+1)
                jacobi_pair<EPL>(a, b, na, nb, nmax, off);
                // positions (2w+1, 2w+2): fetch the right neighbour's first column
#pragma unroll
                for (int e = 0; e < EPL; ++e) sm.slotA[lw][lane + 32 * e] = a[e];
                if (lane == 0) sm.normA[lw] = na;
                sync_all();
                double c2[EPL];
                double nc = 0.0;
                const bool has_right = (w + 1 < M);
                if (has_right) {
                    const int nlw = (lw + 1) % EIG_WARPS;
                    const int nrk = (lw + 1 == EIG_WARPS) ? rank + 1 : rank;
                    const double* src = (EPL == 1 || nrk == rank) ? &sm.slotA[nlw][0]
                                                                    : cluster.map_shared_rank(&sm.slotA[nlw][0], nrk);
                    const double* nsrc = (EPL == 1 || nrk == rank) ? &sm.normA[nlw]
                                                                     : cluster.map_shared_rank(&sm.normA[nlw], nrk);
#pragma unroll
                    for (int e = 0; e < EPL; ++e) c2[e] = src[lane + 32 * e];
                    nc = *nsrc;
                    jacobi_pair<EPL>(b, c2, nb, nc, nmax, off);
                    // now b = position 2w+1, c2 = position 2w+2 (goes back to the right neighbour)
#pragma unroll
                    for (int e = 0; e < EPL; ++e) sm.slotB[lw][lane + 32 * e] = c2[e];
                    if (lane == 0) sm.normB[lw] = nc;
                }
                sync_all();
                if (w > 0) {
                    const int plw = (lw + EIG_WARPS - 1) % EIG_WARPS;
                    const int prk = (lw == 0) ? rank - 1 : rank;
                    const double* src = (EPL == 1 || prk == rank) ? &sm.slotB[plw][0]
                                                                    : cluster.map_shared_rank(&sm.slotB[plw][0], prk);
                    const double* nsrc = (EPL == 1 || prk == rank) ? &sm.normB[plw]
                                                                     : cluster.map_shared_rank(&sm.normB[plw], prk);
#pragma unroll
                    for (int e = 0; e < EPL; ++e) a[e] = src[lane + 32 * e];
                    na = *nsrc;
                }
                // warp 0 keeps its own first column (position 0 sits this round out)
            }
            sweeps_done = sweep + 1;
        }
    }

    // ---- eigenvalues per position, replicated into every CTA ---------------------------
    na = 0.0;
    nb = 0.0;
#pragma unroll
    for (int e = 0; e < EPL; ++e) {
        na = fma(a[e], a[e], na);
        nb = fma(b[e], b[e], nb);
    }
    na = sqrt(warp_sum(na));
    nb = sqrt(warp_sum(nb));
    sync_all();
    if (lane < EPL) {
        double* dst = (EPL == 1) ? sm.lam : cluster.map_shared_rank(sm.lam, lane);
        dst[2 * w] = na;
        dst[2 * w + 1] = nb;
    }
    sync_all();
    // rank (descending, ties by position) of my two columns
    int ra = 0, rb = 0;
    for (int i = lane; i < N2; i += 32) {
        const double v = sm.lam[i];
        ra += (v > na) || (v == na && i < 2 * w);
        rb += (v > nb) || (v == nb && i < 2 * w + 1);
    }
    ra = __reduce_add_sync(0xffffffffu, ra);
    rb = __reduce_add_sync(0xffffffffu, rb);
    if (lane < EPL) {
        double* dst = (EPL == 1) ? &sm.cut : cluster.map_shared_rank(&sm.cut, lane);
        if (ra == ELL) *dst = na;
        if (rb == ELL) *dst = nb;
    }
    sync_all();
    const double cut = sm.cut;

    // ---- outputs -------------------------------------------------------------------------
    const int src_plane = sl.cur;
    float* Ts = A.T + (size_t)s * ELL * N2;
    auto emit = [&](const double (&col)[EPL], double lam, int r) {
        if (r >= ELL) return;
        const double s2v = fmax(lam - cut, 0.0);
        const double coef = (lam > 0.0 && s2v > 0.0) ? sqrt(s2v / lam) / lam : 0.0;
#pragma unroll
        for (int e = 0; e < EPL; ++e) Ts[(size_t)r * N2 + lane + 32 * e] = (float)(coef * col[e]);
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
    };
    emit(a, na, ra);
    emit(b, nb, rb);
    if (w == 0) {
        int nd = 0;
        for (int i = lane; i < N2; i += 32) nd += (fmax(sm.lam[i] - cut, 0.0) >= thr) ? 1 : 0;
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
}

}  // namespace mused

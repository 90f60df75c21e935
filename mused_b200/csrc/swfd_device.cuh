// Device side of the float64 SWFD engine: one FD instance ("slot") is driven by a
// team of P CTAs (a thread-block cluster) that walks the staged rows in arrival
// order and keeps every data-dependent decision of the reference on the device:
//   * FrequentDirectionsWithDump.fit / flush / dump   swfd/frequent_directions.py:96-153
//   * SlidingWindowFD.fit / append period + snapshots swfd/sliding_windows_frequent_directions.py:56-138
//   * SeqBasedSWFD.fit queue trim + routing           swfd/seq_based_swfd.py:47-80
// The reference's thin SVD of [sketch; buffer] (frequent_directions.py:55) is
// replaced by an incrementally maintained eigen-decomposition of the Gram matrix
// of the stored rows R (n x d):  R R^T = Q diag(lam) Q^T.  Each arriving row
// borders the diagonalised Gram with one row/column -- a symmetric arrowhead
// eigenproblem solved through its secular equation (roots kept as pole+offset,
// coupling vector recomputed from the roots, Gu-Eisenstat) -- and the shrink
// (sigma^2 - sigma^2[l], :86-93) plus rotation Vt*sqrt(sigma^2) (:58) is one GEMM
// over the d-wide rows.  tests/proto_arrowhead.py states the same algorithm in
// NumPy and is checked against the reference on the CPU.
#pragma once
#include <float.h>
#include "common.cuh"

namespace mused {
namespace cg = cooperative_groups;

constexpr int NT = 512;          // threads per CTA
constexpr int NWORK = NT / 256;  // GEMM workers per CTA

struct Slot {
    double energy, delta, error;
    double tol, alpha, wnorm;
    unsigned long long last_dump, period;
    unsigned long long n_flush, n_dump, n_raw, overflow;
    int n, b, rcur, qcur, q_ident, q_head, q_len;
    int f_reset, f_raw, f_flush, tail;
    int m, rot_n, parity, pad;
    unsigned long long prof[16];  // cycles per phase, accumulated by the leader thread
};

enum { V_LAM = 0, V_LAMN, V_S2, V_G, V_W, V_DD, V_WW, V_WHAT, V_MU, NV };
enum { I_ORG = 0, I_SIDX, I_KIDX, I_ND, I_POS, I_ROTI, I_ROTJ, NI };

struct Dev {
    int S, ell, nmax, ldq, capn, pad;
    long long d;
    unsigned long long N;
    double qcap;
    Slot* slots;
    double* R;      // [S][2][nmax][d]
    double* Q;      // [S][2][ldq][ldq]
    double* UT;     // [S][ldq][ldq]
    double* vec;    // [S][NV][ldq]
    double* rotcs;  // [S][2][ldq]
    int* ivec;      // [S][NI][ldq]
    double* ringV;  // [S][capn][d]
    unsigned long long* ringS;  // [S][capn]
    unsigned long long* ringT;

    __device__ __forceinline__ double* Rp(int s, int which) const {
        return R + ((size_t)s * 2 + which) * (size_t)nmax * d;
    }
    __device__ __forceinline__ double* Qp(int s, int which) const {
        return Q + ((size_t)s * 2 + which) * (size_t)ldq * ldq;
    }
    __device__ __forceinline__ double* UTp(int s) const { return UT + (size_t)s * ldq * ldq; }
    __device__ __forceinline__ double* v(int s, int which) const {
        return vec + ((size_t)s * NV + which) * ldq;
    }
    __device__ __forceinline__ int* iv(int s, int which) const {
        return ivec + ((size_t)s * NI + which) * ldq;
    }
    __device__ __forceinline__ double* ring_row(int s, int e) const {
        return ringV + ((size_t)s * capn + e) * d;
    }
};

template <int P>
struct Team {
    int tid, nth, wid, nw, lane, rank;
    __device__ Team() {
        rank = (P == 1) ? 0 : (int)cg::this_cluster().block_rank();
        tid = rank * NT + threadIdx.x;
        nth = P * NT;
        wid = tid >> 5;
        nw = nth >> 5;
        lane = threadIdx.x & 31;
    }
    __device__ __forceinline__ void sync() const {
        if (P == 1) __syncthreads();
        else cg::this_cluster().sync();
    }
    __device__ __forceinline__ bool leader() const { return tid == 0; }
};

enum { PH_A = 0, PH_B, PH_C, PH_D, PH_E, PH_F, PH_G, PH_H, PH_FLUSH, PH_DUMP, PH_RAW, PH_ROWS, PH_N };
#define PROF_MARK(ph)                                          \
    do {                                                       \
        if (T.leader()) {                                      \
            const unsigned long long _now = clock64();         \
            sl.prof[ph] += _now - t_prof;                      \
            t_prof = _now;                                     \
        }                                                      \
    } while (0)

struct TeamSmem {
    GemmSmem g[NWORK];
    double red[32];
    int ia, ib;
};

__device__ __forceinline__ double block_reduce_max(double v, double* red) {
    v = warp_max(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    double r = red[0];
    for (int i = 1; i < NT / 32; ++i) r = fmax(r, red[i]);
    return r;
}
__device__ __forceinline__ double block_reduce_sum(double v, double* red) {
    v = warp_sum(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    double r = 0.0;
    for (int i = 0; i < NT / 32; ++i) r += red[i];
    return r;
}

// ---------------------------------------------------------------------------
// secular root k of   f(x) = alpha - x - sum_i w_i^2 / (d_i - x),  d sorted
// descending and distinct, w_i != 0: roots interlace, root 0 above d_0, root m
// below d_{m-1}.  Returned as origin pole `org` + offset `mu`, so that every
// difference root - d_i is available without cancellation.  One warp per root.
// ---------------------------------------------------------------------------
__device__ __forceinline__ void secular_root(int k, int m, const double* dd,
                                             const double* ww, double alpha,
                                             double wnorm, int lane, double& mu_out, int& org_out) {
    const double eps = DBL_EPSILON;
    int o;
    double a, b;
    bool lower_pole;  // origin pole is the lower end of the bracket (root above it)
    if (k == 0) {
        o = 0;
        a = 0.0;
        b = fmax(dd[0], alpha) + wnorm - dd[0];
        lower_pole = true;
    } else if (k == m) {
        o = m - 1;
        a = fmin(dd[m - 1], alpha) - wnorm - dd[m - 1];
        b = 0.0;
        lower_pole = false;
    } else {
        const double dk = dd[k];
        const double half = 0.5 * (dd[k - 1] - dk);
        double s = 0.0;
        for (int i = lane; i < m; i += 32) {
            double w = ww[i];
            s += (w * w) / ((dd[i] - dk) - half);
        }
        s = warp_sum(s);
        const double hm = (alpha - dk) - half - s;
        if (hm == 0.0) {
            mu_out = half;
            org_out = k;
            return;
        }
        if (hm > 0.0) { o = k - 1; a = -half; b = 0.0; lower_pole = false; }
        else          { o = k;     a = 0.0;   b = half; lower_pole = true; }
    }
    const double dorg = dd[o];
    const double c0 = alpha - dorg;
    double x = 0.5 * (a + b);
    for (int it = 0; it < 100; ++it) {
        double s0 = 0.0, s1 = 0.0, s2 = 0.0;
        for (int i = lane; i < m; i += 32) {
            double w = ww[i];
            double r = 1.0 / ((dd[i] - dorg) - x);
            double q = (w * w) * r;
            s0 += q;
            s1 = fma(q, r, s1);
            s2 += fabs(q);
        }
        s0 = warp_sum(s0);
        s1 = warp_sum(s1);
        s2 = warp_sum(s2);
        const double h = c0 - x - s0;
        const double hp = -1.0 - s1;
        const double hs = fabs(c0) + fabs(x) + s2;
        if (h == 0.0 || fabs(h) <= 8.0 * eps * hs) break;
        if (h > 0.0) a = x; else b = x;
        const double p = lower_pole ? x : -x;
        const double pp = lower_pole ? 1.0 : -1.0;
        const double F = p * h;
        const double Fp = pp * h + p * hp;
        double xn = (Fp != 0.0) ? x - F / Fp : 0.5 * (a + b);
        if (!(xn > a && xn < b)) xn = 0.5 * (a + b);
        const bool stop = (xn == x) || ((b - a) <= 2.0 * eps * fmax(fabs(a), fabs(b)));
        x = xn;
        if (stop) break;
    }
    mu_out = x;
    org_out = o;
}

// rank-sort the non-deflated indices by (value desc, index asc); CTA-local.
__device__ __forceinline__ int rank_sort_nd(int n, const double* val, const int* nd, int* sidx,
                                            int* counter) {
    if (threadIdx.x == 0) *counter = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += NT) {
        if (!nd[i]) continue;
        const double vi = val[i];
        int r = 0;
        for (int j = 0; j < n; ++j) {
            if (!nd[j]) continue;
            const double vj = val[j];
            r += (vj > vi) || (vj == vi && j < i);
        }
        sidx[r] = i;
        atomicAdd(counter, 1);
    }
    __syncthreads();
    return *counter;
}

// ---------------------------------------------------------------------------
// Border the slot's Gram eigen-decomposition with the row x (||x||^2 = alpha).
// ---------------------------------------------------------------------------
template <int P>
__device__ void fd_border(const Dev& D, int s, const Team<P>& T, TeamSmem& sm,
                          const double* __restrict__ x, double alpha, unsigned long long& t_prof) {
    Slot& sl = D.slots[s];
    const int n = sl.n, rcur = sl.rcur, qcur = sl.qcur;
    int q_ident = sl.q_ident;
    const long long d = D.d;
    const int ldq = D.ldq;
    double* R = D.Rp(s, rcur);
    double* Qc = D.Qp(s, qcur);
    double* Qn = D.Qp(s, qcur ^ 1);
    double* UT = D.UTp(s);
    double* lam = D.v(s, V_LAM);
    double* lamn = D.v(s, V_LAMN);
    double* g = D.v(s, V_G);
    double* w = D.v(s, V_W);
    double* dd = D.v(s, V_DD);
    double* ww = D.v(s, V_WW);
    double* what = D.v(s, V_WHAT);
    double* mu = D.v(s, V_MU);
    int* org = D.iv(s, I_ORG);
    int* sidx = D.iv(s, I_SIDX);
    int* kidx = D.iv(s, I_KIDX);
    int* nd = D.iv(s, I_ND);
    int* pos = D.iv(s, I_POS);
    int* roti = D.iv(s, I_ROTI);
    int* rotj = D.iv(s, I_ROTJ);
    double* rotc = D.rotcs + (size_t)s * 2 * ldq;
    double* rots = rotc + ldq;

    // ---- B: g = R x ; store x as row n -------------------------------------
    for (int i = T.wid; i < n; i += T.nw) {
        const double* r = R + (size_t)i * d;
        double acc = 0.0;
        for (long long c = T.lane; c < d; c += 32) acc = fma(r[c], x[c], acc);
        acc = warp_sum(acc);
        if (T.lane == 0) g[i] = acc;
    }
    {
        double* dst = R + (size_t)n * d;
        for (long long c = T.tid; c < d; c += T.nth) dst[c] = x[c];
    }
    T.sync();
    PROF_MARK(PH_B);

    // ---- C: w = Q^T g --------------------------------------------------------
    if (q_ident) {
        for (int k = T.tid; k < n; k += T.nth) w[k] = g[k];
    } else {
        // 32 columns x 16 row-slices per CTA pass
        const int ki = threadIdx.x & 31, si = threadIdx.x >> 5;
        for (int c0 = T.rank * 32; c0 < n; c0 += P * 32) {
            const int k = c0 + ki;
            double acc = 0.0;
            if (k < n)
                for (int i = si; i < n; i += NT / 32) acc = fma(Qc[(size_t)i * ldq + k], g[i], acc);
            __syncthreads();
            double* part = &sm.g[0].b[0][0];
            part[si * 32 + ki] = acc;
            __syncthreads();
            if (si == 0 && k < n) {
                double t = 0.0;
                for (int q = 0; q < NT / 32; ++q) t += part[q * 32 + ki];
                w[k] = t;
            }
        }
    }
    T.sync();
    PROF_MARK(PH_C);

    // ---- D: deflation + ordering (CTA 0) -------------------------------------
    if (T.rank == 0) {
        double mx = 0.0;
        for (int i = threadIdx.x; i < n; i += NT) mx = fmax(mx, fmax(fabs(lam[i]), fabs(w[i])));
        mx = block_reduce_max(mx, sm.red);
        const double tol = 8.0 * DBL_EPSILON * fmax(mx, fabs(alpha));
        for (int i = threadIdx.x; i < n; i += NT) {
            nd[i] = fabs(w[i]) > tol;
            lamn[i] = lam[i];
        }
        if (threadIdx.x == 0) sm.ib = 0;
        __syncthreads();
        int m = rank_sort_nd(n, lamn, nd, sidx, &sm.ia);
        // type-2 screen: adjacent (sorted) diagonals whose rotation is negligible
        for (int r = threadIdx.x; r + 1 < m; r += NT) {
            const int i = sidx[r], j = sidx[r + 1];
            const double rr = hypot(w[i], w[j]);
            const double c = w[j] / rr, sn = w[i] / rr;
            if (fabs((lamn[i] - lamn[j]) * c * sn) <= tol) atomicOr(&sm.ib, 1);
        }
        __syncthreads();
        int nrot = 0;
        if (sm.ib) {
            if (threadIdx.x == 0) {
                int cur = sidx[0];
                for (int r = 1; r < m; ++r) {
                    const int nxt = sidx[r];
                    const double rr = hypot(w[cur], w[nxt]);
                    const double c = w[nxt] / rr, sn = w[cur] / rr;
                    const double t = lamn[cur] - lamn[nxt];
                    if (fabs(t * c * sn) <= tol) {
                        roti[nrot] = cur; rotj[nrot] = nxt; rotc[nrot] = c; rots[nrot] = sn;
                        ++nrot;
                        const double dc = lamn[cur], dn = lamn[nxt];
                        lamn[cur] = dc * c * c + dn * sn * sn;
                        lamn[nxt] = dc * sn * sn + dn * c * c;
                        w[cur] = 0.0;
                        w[nxt] = rr;
                        nd[cur] = 0;
                    }
                    cur = nxt;
                }
                sm.ib = nrot;
            }
            __syncthreads();
            nrot = sm.ib;
            m = rank_sort_nd(n, lamn, nd, sidx, &sm.ia);
        }
        for (int i = threadIdx.x; i < n; i += NT) kidx[i] = -1;
        __syncthreads();
        double wsq = 0.0;
        for (int r = threadIdx.x; r < m; r += NT) {
            const int i = sidx[r];
            dd[r] = lamn[i];
            ww[r] = w[i];
            kidx[i] = r;
            wsq += w[i] * w[i];
        }
        wsq = block_reduce_sum(wsq, sm.red);
        if (threadIdx.x == 0) {
            kidx[n] = m;
            sl.m = m;
            sl.rot_n = nrot;
            sl.tol = tol;
            sl.alpha = alpha;
            sl.wnorm = sqrt(wsq);
        }
    }
    T.sync();
    PROF_MARK(PH_D);
    const int m = sl.m;
    const int rot_n = sl.rot_n;
    const double wnorm = sl.wnorm;

    // ---- D2: rare -- apply type-2 rotations to the eigenvector columns -------
    if (rot_n > 0) {
        if (q_ident) {
            for (int e = T.tid; e < n * n; e += T.nth) {
                const int r = e / n, c = e - r * n;
                Qc[(size_t)r * ldq + c] = (r == c) ? 1.0 : 0.0;
            }
            q_ident = 0;
            T.sync();
        }
        for (int r = T.tid; r < n; r += T.nth) {
            double* row = Qc + (size_t)r * ldq;
            for (int q = 0; q < rot_n; ++q) {
                const int i = roti[q], j = rotj[q];
                const double c = rotc[q], sn = rots[q];
                const double gi = row[i], gj = row[j];
                row[i] = c * gi - sn * gj;
                row[j] = sn * gi + c * gj;
            }
        }
        T.sync();
    }

    // ---- E: secular roots, one warp each ---------------------------------------
    if (m == 0) {
        if (T.leader()) lamn[n] = alpha;
    } else {
        for (int k = T.wid; k <= m; k += T.nw) {
            double muk;
            int ok;
            secular_root(k, m, dd, ww, alpha, wnorm, T.lane, muk, ok);
            if (T.lane == 0) {
                mu[k] = muk;
                org[k] = ok;
                lamn[(k < m) ? sidx[k] : n] = dd[ok] + muk;
            }
        }
    }
    T.sync();
    PROF_MARK(PH_E);

    // ---- F: coupling vector for which the roots are exact + sorted positions ---
    for (int i = T.wid; i < m; i += T.nw) {
        const double di = dd[i];
        double prod = 1.0;
        for (int j = T.lane; j < m; j += 32) {
            if (j == i) continue;
            const int kk = (j < i) ? j : j + 1;
            const double num = (dd[org[kk]] - di) + mu[kk];
            prod *= num / (dd[j] - di);
        }
        prod = warp_prod(prod);
        if (T.lane == 0) {
            const double e0 = (dd[org[i]] - di) + mu[i];
            const double e1 = (dd[org[i + 1]] - di) + mu[i + 1];
            prod *= -(e0 * e1);
            what[i] = copysign(sqrt(fmax(prod, 0.0)), ww[i]);
        }
    }
    for (int slot = T.tid; slot <= n; slot += T.nth) {
        const double v = lamn[slot];
        int cnt = 0;
        for (int j = 0; j <= n; ++j) {
            const double vj = lamn[j];
            cnt += (vj > v) || (vj == v && j < slot);
        }
        pos[slot] = cnt;
    }
    T.sync();
    PROF_MARK(PH_F);

    // ---- G: eigenvector rows (transposed, already in sorted position) ----------
    for (int slot = T.wid; slot <= n; slot += T.nw) {
        const int k = kidx[slot];
        const int p = pos[slot];
        double* row = UT + (size_t)p * ldq;
        for (int c = T.lane; c <= n; c += 32) row[c] = 0.0;
        __syncwarp();
        if (k < 0) {
            if (T.lane == 0) row[slot] = 1.0;
        } else {
            const double dk = (m > 0) ? dd[org[k]] : 0.0;
            const double muk = (m > 0) ? mu[k] : 0.0;
            double ss = 0.0;
            for (int i = T.lane; i < m; i += 32) {
                const double u = what[i] / -((dk - dd[i]) + muk);
                ss = fma(u, u, ss);
            }
            ss = warp_sum(ss);
            const double inv = 1.0 / sqrt(1.0 + ss);
            for (int i = T.lane; i < m; i += 32) {
                const double u = what[i] / -((dk - dd[i]) + muk);
                row[sidx[i]] = u * inv;
            }
            if (T.lane == 0) row[n] = -inv;
        }
        if (T.lane == 0) lam[p] = lamn[slot];
    }
    T.sync();
    PROF_MARK(PH_G);

    // ---- H: Q_new = diag(Q,1) * U  (rows of UT are eigenvectors) ---------------
    const int n1 = n + 1;
    if (q_ident) {
        for (int e = T.tid; e < n1 * n1; e += T.nth) {
            const int p = e / n1, r = e - p * n1;
            Qn[(size_t)r * ldq + p] = UT[(size_t)p * ldq + r];
        }
    } else {
        const int tiles = (n1 + GT - 1) / GT;
        const int worker = T.rank * NWORK + (threadIdx.x >> 8);
        const int wt = threadIdx.x & 255;
        const int bar = 1 + (threadIdx.x >> 8);
        GemmSmem& gs = sm.g[threadIdx.x >> 8];
        for (int tile = worker; tile < tiles * tiles; tile += P * NWORK) {
            const int m0 = (tile / tiles) * GT, n0 = (tile % tiles) * GT;
            gemm_tile_f64<true, false>(
                gs, wt, m0, n0, n1,
                [&](int r, int c) -> double {
                    if (r < n && c < n) return Qc[(size_t)r * ldq + c];
                    return (r == n && c == n) ? 1.0 : 0.0;
                },
                [&](int c, int p) -> double {
                    return (c < n1 && p < n1) ? UT[(size_t)p * ldq + c] : 0.0;
                },
                [&](int r, int p, double v) {
                    if (r < n1 && p < n1) Qn[(size_t)r * ldq + p] = v;
                },
                [&]() { bar_sync_256(bar); });
        }
    }
    if (T.leader()) {
        sl.n = n1;
        sl.b = sl.b + 1;
        sl.q_ident = 0;
        sl.qcur = qcur ^ 1;
    }
    T.sync();
    PROF_MARK(PH_H);
}

// ---------------------------------------------------------------------------
// Shrink + rotate (frequent_directions.py:86-93 and :58) of a slot with pending
// rows: R_new[i] = sqrt((lam_i - lam_l)/lam_i) * sum_j Q[j][i] R[j],  i < l.
// ---------------------------------------------------------------------------
template <int P>
__device__ void fd_flush(const Dev& D, int s, const Team<P>& T, TeamSmem& sm) {
    Slot& sl = D.slots[s];
    const int n = sl.n, b = sl.b, rcur = sl.rcur, qcur = sl.qcur;
    if (b == 0) return;
    const int ell = D.ell, ldq = D.ldq;
    const long long d = D.d;
    const double* R = D.Rp(s, rcur);
    double* Rn = D.Rp(s, rcur ^ 1);
    const double* Qc = D.Qp(s, qcur);
    double* lam = D.v(s, V_LAM);
    double* s2 = D.v(s, V_S2);
    double* scl = D.v(s, V_G);
    const double cut = fmax(lam[ell], 0.0);
    for (int i = T.tid; i < ell; i += T.nth) {
        const double li = fmax(lam[i], 0.0);
        scl[i] = (li > 0.0) ? sqrt((li - cut) / li) : 0.0;
    }
    T.sync();
    {
        const int tm = (ell + GT - 1) / GT;
        const int tn = (int)((d + GT - 1) / GT);
        const int worker = T.rank * NWORK + (threadIdx.x >> 8);
        const int wt = threadIdx.x & 255;
        const int bar = 1 + (threadIdx.x >> 8);
        GemmSmem& gs = sm.g[threadIdx.x >> 8];
        for (int tile = worker; tile < tm * tn; tile += P * NWORK) {
            const int m0 = (tile / tn) * GT, n0 = (tile % tn) * GT;
            gemm_tile_f64<false, true>(
                gs, wt, m0, n0, n,
                [&](int i, int j) -> double {
                    return (i < ell && j < n) ? Qc[(size_t)j * ldq + i] : 0.0;
                },
                [&](int j, int c) -> double {
                    return (j < n && c < d) ? R[(size_t)j * d + c] : 0.0;
                },
                [&](int i, int c, double v) {
                    if (i < ell && c < d) Rn[(size_t)i * d + c] = v * scl[i];
                },
                [&]() { bar_sync_256(bar); });
        }
    }
    T.sync();
    double v0 = 0.0;
    for (int i = T.tid; i < ell; i += T.nth) {
        const double v = fmax(lam[i], 0.0) - cut;
        lam[i] = v;
        s2[i] = v;
        if (i == 0) v0 = v;
    }
    if (T.leader()) {
        sl.delta += cut / 2.0;
        sl.energy = v0;
        sl.n = ell;
        sl.b = 0;
        sl.q_ident = 1;
        sl.rcur = rcur ^ 1;
        sl.n_flush += 1;
    }
    T.sync();
}

// dump (frequent_directions.py:127-142): move sketch row 0 out when
// sigma^2[0] >= error.  dst = snapshot ring entry (or a caller buffer).
template <int P>
__device__ bool fd_dump_row(const Dev& D, int s, const Team<P>& T, double* dst) {
    Slot& sl = D.slots[s];
    double* s2 = D.v(s, V_S2);
    const bool doit = (sl.b == 0) && (s2[0] >= sl.error);
    if (!doit) return false;
    double* row = D.Rp(s, sl.rcur);
    for (long long c = T.tid; c < D.d; c += T.nth) {
        dst[c] = row[c];
        row[c] = 0.0;
    }
    T.sync();
    if (T.leader()) {
        s2[0] = 0.0;
        D.v(s, V_LAM)[0] = 0.0;
        sl.energy = 0.0;
        sl.n_dump += 1;
    }
    return true;
}

// ---------------------------------------------------------------------------
// Kernels.  Grid = (#slots * P) CTAs, cluster = P CTAs, NT threads.
// ---------------------------------------------------------------------------
enum { MODE_SEQ = 0, MODE_FD = 1, MODE_FD_BATCH = 2, MODE_FLUSH = 3, MODE_DUMP = 4 };

struct RowArgs {
    const double* rows;   // staged float64 rows [n_rows][d]
    const double* norm2;  // ||row||^2
    long long n_rows;
    unsigned long long time0;  // rows consumed before this launch
    const unsigned long long* times;  // explicit arrival time per row (TimeBasedSWFD), or NULL: time0 + r + 1
    int mode;
    int slot0;            // first slot handled by this launch
    long long batch;      // MODE_FD_BATCH
    int last_call;        // MODE_FD_BATCH: shrink the ragged tail
    double* out;          // MODE_DUMP: destination row
    int* out_flag;        // MODE_DUMP
};

template <int P>
__global__ void __launch_bounds__(NT, 1) k_swfd_rows(Dev D, RowArgs A) {
    __shared__ TeamSmem sm;
    const Team<P> T;
    const int s = A.slot0 + blockIdx.x / P;
    Slot& sl = D.slots[s];

    if (A.mode == MODE_FLUSH) {
        fd_flush<P>(D, s, T, sm);
        return;
    }
    if (A.mode == MODE_DUMP) {
        const bool did = fd_dump_row<P>(D, s, T, A.out);
        if (T.leader()) *A.out_flag = did ? 1 : 0;
        return;
    }
    unsigned long long t_prof = clock64();
    for (long long r = 0; r < A.n_rows; ++r) {
        const double* x = A.rows + (size_t)r * D.d;
        const double nx = A.norm2[r];
        const unsigned long long t = A.times ? A.times[r] : A.time0 + (unsigned long long)r + 1ull;
        // ---- A: scalar bookkeeping by the leader ------------------------------
        if (T.leader()) {
            sl.f_reset = 0;
            sl.f_raw = 0;
            sl.f_flush = 0;
            if (A.mode == MODE_SEQ) {
                // SeqBasedSWFD.fit :59-74 -- cap + expiry trim
                unsigned long long* qt = D.ringT + (size_t)s * D.capn;
                while (sl.q_len > 0 &&
                       ((double)sl.q_len > D.qcap || qt[sl.q_head] + D.N <= t)) {
                    sl.q_head = (sl.q_head + 1) % D.capn;
                    sl.q_len -= 1;
                }
                // SlidingWindowFD period swap :99-107 -- the slot that becomes
                // fd_aux starts over with last_dump_time = time
                const unsigned long long p = t / D.N;
                if (p > sl.period) {
                    // explicit times may skip whole periods: after two or more swaps both sides are
                    // fresh (the `while self.period < self.time // self.N` loop, :66-75 / :99-107)
                    const bool both = p >= sl.period + 2ull;
                    sl.period = p;
                    if (both || (int)(p & 1ull) != sl.parity) {
                        sl.f_reset = 1;
                        sl.n = D.ell; sl.b = 0; sl.q_ident = 1;
                        sl.energy = 0.0; sl.delta = 0.0;
                        sl.q_head = 0; sl.q_len = 0;
                        sl.last_dump = t;
                    }
                }
                if (nx >= sl.error) {
                    // SlidingWindowFD.append :56-80 -- raw row becomes a snapshot
                    sl.f_raw = 1;
                    if (sl.q_len == D.capn) {
                        sl.q_head = (sl.q_head + 1) % D.capn; sl.q_len -= 1; sl.overflow += 1;
                    }
                    const int tail = (sl.q_head + sl.q_len) % D.capn;
                    sl.tail = tail;
                    D.ringS[(size_t)s * D.capn + tail] = sl.last_dump;
                    D.ringT[(size_t)s * D.capn + tail] = t;
                    sl.q_len += 1;
                    sl.last_dump = t + 1ull;
                    sl.n_raw += 1;
                } else {
                    sl.energy = sl.energy + nx;  // frequent_directions.py:116
                    sl.f_flush = (sl.b + 1 >= D.ell) || (sl.energy >= sl.error);  // :122-125
                }
            } else if (A.mode == MODE_FD) {
                sl.energy = sl.energy + nx;
                sl.f_flush = (sl.b + 1 >= D.ell) || (sl.energy >= sl.error);
            } else {  // MODE_FD_BATCH: FrequentDirections.fit(X, batch_size) :39-58
                sl.f_flush = (sl.b + 1 >= A.batch) || (A.last_call && r + 1 == A.n_rows);
            }
        }
        T.sync();
        PROF_MARK(PH_A);
        if (sl.f_reset) {
            double* R = D.Rp(s, sl.rcur);
            const size_t tot = (size_t)D.ell * D.d;
            for (size_t e = T.tid; e < tot; e += T.nth) R[e] = 0.0;
            double* lam = D.v(s, V_LAM);
            double* s2 = D.v(s, V_S2);
            for (int i = T.tid; i < D.ell; i += T.nth) { lam[i] = 0.0; s2[i] = 0.0; }
            T.sync();
        }
        if (sl.f_raw) {
            double* dst = D.ring_row(s, sl.tail);
            for (long long c = T.tid; c < D.d; c += T.nth) dst[c] = x[c];
            T.sync();
            PROF_MARK(PH_RAW);
            continue;
        }
        const bool flush = sl.f_flush != 0;
        fd_border<P>(D, s, T, sm, x, nx, t_prof);
        if (flush) {
            fd_flush<P>(D, s, T, sm);
            PROF_MARK(PH_FLUSH);
            if (A.mode == MODE_SEQ) {
                // SlidingWindowFD.fit :126-130 / :134-138
                const int tail = (sl.q_head + sl.q_len) % D.capn;
                const bool did = fd_dump_row<P>(D, s, T, D.ring_row(s, tail));
                if (did && T.leader()) {
                    D.ringS[(size_t)s * D.capn + tail] = sl.last_dump;
                    D.ringT[(size_t)s * D.capn + tail] = t;
                    sl.q_len += 1;
                    sl.last_dump = t + 1ull;
                }
                T.sync();
                PROF_MARK(PH_DUMP);
            }
        }
    }
    if (T.leader()) sl.prof[PH_ROWS] += (unsigned long long)A.n_rows;
}

}  // namespace mused

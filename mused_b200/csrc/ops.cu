// Similarity / assignment operators of matrix_operations.py as sm_100a kernels.
//   mused_assign          fused cosine|L2 + threshold + argmax|argmin (north_star operator;
//                         nearest reference code matrix_operations.py:106, :286-289)
//   mused_topk_adjacency  kNN -> 0/1 adjacency, OR-fused (matrix_operations.py:101-141)
//   mused_topk_adjacency_static  the same with file-scope device scratch (no allocation; W <= 1024)
//   mused_group_adjacency / mused_jaccard_topk_adjacency  the username / tags branches (:55-89) on integer ids
//   mused_contingency     overlap counts of match_clusters (matrix_operations.py:169-172)
#include <float.h>
#include <stdlib.h>
#include <atomic>
#include "common.cuh"
#include "assign_tc.cuh"

namespace mused {

constexpr int AT = 64;   // rows / centroids per tile
constexpr int AK = 32;   // k chunk (floats)

struct AssignSmem {
    float a[AK][AT + 4];
    float b[AK][AT + 4];
};

// squared row norms of a small matrix (centroids / features)
__global__ void k_row_norm2_f32(const float* __restrict__ X, long long n, long long d, long long ld,
                                float* __restrict__ out) {
    const long long r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (r >= n) return;
    const float* x = X + (size_t)r * ld;
    float acc = 0.f;
    for (long long c = threadIdx.x & 31; c < d; c += 32) acc = fmaf(x[c], x[c], acc);
    acc = warp_sum_f(acc);
    if ((threadIdx.x & 31) == 0) out[r] = acc;
}

// 64 x 64 tile of X C^T in fp32 (16x16 threads, 4x4 micro-tile); the squared
// norm of each X row falls out of the same pass (tx == 0 lanes accumulate it).
__device__ __forceinline__ void dot_tile_f32(AssignSmem& sm, const float* __restrict__ X,
                                             long long n, long long ldx, const float* __restrict__ C,
                                             long long k, long long ldc, long long d, long long r0,
                                             long long c0, float acc[4][4], float xn[4]) {
    const int t = threadIdx.x, ty = t >> 4, tx = t & 15;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        xn[i] = 0.f;
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    }
    for (long long k0 = 0; k0 < d; k0 += AK) {
#pragma unroll
        for (int e = t; e < AT * AK; e += 256) {
            const int kk = e & (AK - 1), mm = e >> 5;
            const long long r = r0 + mm, c = k0 + kk;
            sm.a[kk][mm] = (r < n && c < d) ? X[(size_t)r * ldx + c] : 0.f;
            const long long cr = c0 + mm;
            sm.b[kk][mm] = (cr < k && c < d) ? C[(size_t)cr * ldc + c] : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < AK; ++kk) {
            float av[4], bv[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) av[i] = sm.a[kk][ty * 4 + i];
#pragma unroll
            for (int j = 0; j < 4; ++j) bv[j] = sm.b[kk][tx * 4 + j];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                xn[i] = fmaf(av[i], av[i], xn[i]);
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
            }
        }
        __syncthreads();
    }
}

// One CTA per 64 rows of X; loops over centroid tiles keeping the running best.
__global__ void __launch_bounds__(256) k_assign(const float* __restrict__ X, long long n, long long d,
                                                long long ldx, const float* __restrict__ C, long long k,
                                                long long ldc, const float* __restrict__ cn2, int metric,
                                                float tau, int* __restrict__ labels,
                                                float* __restrict__ score) {
    __shared__ AssignSmem sm;
    const int t = threadIdx.x, ty = t >> 4, tx = t & 15;
    for (long long r0 = (long long)blockIdx.x * AT; r0 < n; r0 += (long long)gridDim.x * AT) {
        float best[4];
        int bidx[4];
        float xn_keep[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int i = 0; i < 4; ++i) { best[i] = -FLT_MAX; bidx[i] = 0x7fffffff; }
        for (long long c0 = 0; c0 < k; c0 += AT) {
            float acc[4][4], xn[4];
            dot_tile_f32(sm, X, n, ldx, C, k, ldc, d, r0, c0, acc, xn);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                xn_keep[i] = xn[i];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const long long cj = c0 + tx * 4 + j;
                    if (cj >= k) continue;
                    const float c2 = cn2[cj];
                    float key;  // larger is better for both metrics
                    if (metric == MUSED_METRIC_COSINE) {
                        // sklearn normalises with norm 0 -> 1 (cosine of a zero row is 0)
                        key = acc[i][j] * rsqrtf(c2 > 0.f ? c2 : 1.f);
                    } else {
                        key = -(c2 - 2.f * acc[i][j]);  // -(||c||^2 - 2 x.c); ||x||^2 is common
                    }
                    if (key > best[i] || (key == best[i] && (int)cj < bidx[i])) {
                        best[i] = key;
                        bidx[i] = (int)cj;
                    }
                }
            }
        }
        // reduce over the 16 tx lanes that share a row (same half-warp)
#pragma unroll
        for (int i = 0; i < 4; ++i) {
#pragma unroll
            for (int o = 8; o > 0; o >>= 1) {
                const float ob = __shfl_xor_sync(0xffffffffu, best[i], o);
                const int oi = __shfl_xor_sync(0xffffffffu, bidx[i], o);
                if (ob > best[i] || (ob == best[i] && oi < bidx[i])) { best[i] = ob; bidx[i] = oi; }
            }
            const long long r = r0 + ty * 4 + i;
            if (tx == 0 && r < n) {
                const float x2 = xn_keep[i];
                float sc;
                bool keep;
                if (metric == MUSED_METRIC_COSINE) {
                    sc = best[i] * rsqrtf(x2 > 0.f ? x2 : 1.f);
                    keep = sc >= tau;
                } else {
                    sc = sqrtf(fmaxf(x2 - best[i], 0.f));
                    keep = (tau < 0.f) || (sc <= tau);
                }
                labels[r] = keep ? bidx[i] : -1;
                if (score) score[r] = sc;
            }
        }
    }
}

// S = similarity key matrix (W x W, larger = closer) for the adjacency operator.
__global__ void __launch_bounds__(256) k_pair_keys(const float* __restrict__ F, long long W, long long f,
                                                   long long ld, const float* __restrict__ n2, int metric,
                                                   float* __restrict__ S) {
    __shared__ AssignSmem sm;
    const int t = threadIdx.x, ty = t >> 4, tx = t & 15;
    const long long tiles = (W + AT - 1) / AT;
    for (long long tile = blockIdx.x; tile < tiles * tiles; tile += gridDim.x) {
        const long long r0 = (tile / tiles) * AT, c0 = (tile % tiles) * AT;
        float acc[4][4], xn[4];
        dot_tile_f32(sm, F, W, ld, F, W, ld, f, r0, c0, acc, xn);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const long long r = r0 + ty * 4 + i;
            if (r >= W) continue;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const long long c = c0 + tx * 4 + j;
                if (c >= W) continue;
                float key;
                if (metric == MUSED_METRIC_COSINE) {
                    const float a2 = n2[r], b2 = n2[c];
                    key = acc[i][j] * rsqrtf(a2 > 0.f ? a2 : 1.f) * rsqrtf(b2 > 0.f ? b2 : 1.f);
                } else {
                    key = (r == c) ? 0.f : -fmaxf(n2[r] + n2[c] - 2.f * acc[i][j], 0.f);
                }
                S[(size_t)r * W + c] = key;
            }
        }
    }
}

// per row: k best keys (ties -> lowest index), adj[i][j] |= 1 for j != i.
template <typename T>
__global__ void __launch_bounds__(256) k_topk_rows(const T* __restrict__ S, long long W, long long k,
                                                   unsigned char* __restrict__ adj) {
    extern __shared__ __align__(8) unsigned char row_raw[];
    T* row = reinterpret_cast<T*>(row_raw);
    __shared__ T rb[8];
    __shared__ int ri[8];
    for (long long i = blockIdx.x; i < W; i += gridDim.x) {
        for (long long c = threadIdx.x; c < W; c += 256) row[c] = S[(size_t)i * W + c];
        __syncthreads();
        const long long kk = k < W ? k : W;
        const T lowest = (T)(-FLT_MAX);            // keys are >= -FLT_MAX in both precisions
        for (long long it = 0; it < kk; ++it) {
            T b = lowest;
            int bi = 0x7fffffff;
            for (long long c = threadIdx.x; c < W; c += 256) {
                const T v = row[c];
                if (v > b) { b = v; bi = (int)c; }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const T ob = __shfl_xor_sync(0xffffffffu, b, o);
                const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                if (ob > b || (ob == b && oi < bi)) { b = ob; bi = oi; }
            }
            if ((threadIdx.x & 31) == 0) { rb[threadIdx.x >> 5] = b; ri[threadIdx.x >> 5] = bi; }
            __syncthreads();
            if (threadIdx.x == 0) {
                for (int q = 1; q < 8; ++q)
                    if (rb[q] > b || (rb[q] == b && ri[q] < bi)) { b = rb[q]; bi = ri[q]; }
                if (bi != 0x7fffffff) {
                    row[bi] = lowest;  // taken (an all -FLT_MAX row simply stops matching)
                    if (bi != (int)i) adj[(size_t)i * W + bi] = 1;
                }
            }
            __syncthreads();
        }
    }
}

// float64 keys of the two low-dimensional modalities (larger = closer):
//   HAVERSINE  (matrix_operations.py:22-30, :250-263): rows are (latitude, longitude) in degrees;
//              key = -a with a = sin^2(dlat/2) + cos(lat1) cos(lat2) sin^2(dlon/2); the distance
//              2 r asin(sqrt(a)) is monotone in a, so the ranking is the reference's.
//   L1SUM      (:32-53): key = -(|x0 - y0| + |x1 - y1| + ...), the reference's taken + upload differences.
__global__ void __launch_bounds__(256) k_pair_keys_f64(const double* __restrict__ F, long long W, long long f,
                                                       long long ld, int metric, double* __restrict__ S) {
    const double rad = 0.017453292519943295;
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < W * W;
         e += (long long)gridDim.x * blockDim.x) {
        const long long r = e / W, c = e % W;
        const double* x = F + r * ld;
        const double* y = F + c * ld;
        double key;
        if (metric == MUSED_METRIC_HAVERSINE) {
            const double lat1 = x[0] * rad, lon1 = x[1] * rad, lat2 = y[0] * rad, lon2 = y[1] * rad;
            const double sa = sin((lat2 - lat1) / 2.0), sb = sin((lon2 - lon1) / 2.0);
            key = -(sa * sa + cos(lat1) * cos(lat2) * sb * sb);
        } else {
            double acc = 0.0;
            for (long long j = 0; j < f; ++j) acc += fabs(y[j] - x[j]);
            key = -acc;
        }
        S[e] = fmax(key, -(double)FLT_MAX);
    }
}

__global__ void k_contingency(const int* __restrict__ prev, const int* __restrict__ cur, long long n, int P,
                              int Q, int* __restrict__ counts) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
         i += (long long)gridDim.x * blockDim.x) {
        const int p = prev[i], q = cur[i];
        if (p >= 0 && p < P && q >= 0 && q < Q) atomicAdd(&counts[(size_t)p * Q + q], 1);
    }
}

// ---------------------------------------------------------------------------
// Device KMeans (perform_clustering, matrix_operations.py:149-153: KMeans(n_clusters, random_state).fit_predict).
// k-means++ seeding (D^2 sampling, one trial per centre, counter-based RNG) by ONE CTA; Lloyd iterations =
// the assignment operator (mused_assign, L2) + k_kmeans_update below.  Every reduction runs in a fixed order:
// the result is a pure function of (X, k, seed).
// ---------------------------------------------------------------------------
__device__ __forceinline__ double kmeans_uniform(unsigned long long seed, unsigned long long ctr) {
    unsigned long long z = seed + 0x9E3779B97F4A7C15ull * (ctr + 1ull);       // splitmix64
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    return (double)(z >> 11) * (1.0 / 9007199254740992.0);
}

__global__ void __launch_bounds__(1024) k_kmeans_seed(const float* __restrict__ X, long long n, long long d, long long ld,
                                                      int k, unsigned long long seed, float* __restrict__ C,
                                                      float* __restrict__ mind) {
    __shared__ double part[1024];
    __shared__ long long pick;
    const int t = threadIdx.x;
    if (t == 0) pick = (long long)(kmeans_uniform(seed, 0) * (double)n) % n;
    __syncthreads();
    for (int j = 0; j < k; ++j) {
        const long long p = pick;
        for (long long c = t; c < d; c += 1024) C[(size_t)j * d + c] = X[(size_t)p * ld + c];
        __syncthreads();
        if (j + 1 == k) break;
        // squared distance of my rows (a contiguous slice: the cumulative search below needs an order) to the new centre
        const long long per = (n + 1023) / 1024, lo = t * per, hi = lo + per < n ? lo + per : n;
        double mine = 0.0;
        for (long long i = lo; i < hi; ++i) {
            float acc = 0.f;
            const float* x = X + (size_t)i * ld;
            const float* cj = C + (size_t)j * d;
            for (long long c = 0; c < d; ++c) { const float df = x[c] - cj[c]; acc = fmaf(df, df, acc); }
            const float m = (j == 0) ? acc : fminf(mind[i], acc);
            mind[i] = m;
            mine += (double)m;
        }
        part[t] = mine;
        __syncthreads();
        if (t == 0) {
            double tot = 0.0;
            for (int q = 0; q < 1024; ++q) tot += part[q];
            const double u = kmeans_uniform(seed, (unsigned long long)(j + 1)) * tot;
            double run = 0.0;
            int q = 0;
            for (; q < 1023; ++q) { if (run + part[q] > u) break; run += part[q]; }
            const long long qlo = q * per, qhi = qlo + per < n ? qlo + per : n;
            long long i = qlo;
            for (; i + 1 < qhi; ++i) { if (run + (double)mind[i] > u) break; run += (double)mind[i]; }
            pick = i < n ? i : n - 1;
        }
        __syncthreads();
    }
}

// new centre j = mean of the rows labelled j (CTA j; thread = column, rows in index order: deterministic); an empty
// cluster keeps its centre.  shift2[j] = squared movement, inertia partial left to the assignment scores.
__global__ void __launch_bounds__(256) k_kmeans_update(const float* __restrict__ X, long long n, long long d, long long ld,
                                                       const int* __restrict__ labels, float* __restrict__ C,
                                                       double* __restrict__ shift2, int* __restrict__ counts) {
    __shared__ double red[8];
    const int j = blockIdx.x;
    double sh = 0.0;
    int cnt = 0;
    for (long long c0 = 0; c0 < d; c0 += 256) {
        const long long c = c0 + threadIdx.x;
        double acc = 0.0;
        cnt = 0;
        for (long long i = 0; i < n; ++i) {
            if (labels[i] == j) {
                ++cnt;
                if (c < d) acc += (double)X[(size_t)i * ld + c];
            }
        }
        if (c < d && cnt > 0) {
            const float nv = (float)(acc / (double)cnt);
            const double df = (double)nv - (double)C[(size_t)j * d + c];
            sh = fma(df, df, sh);
            C[(size_t)j * d + c] = nv;
        }
    }
    sh = warp_sum(sh);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = sh;
    __syncthreads();
    if (threadIdx.x == 0) {
        double tsum = 0.0;
        for (int q = 0; q < 8; ++q) tsum += red[q];
        shift2[j] = tsum;
        counts[j] = cnt;
    }
}

// stats[0] = sum of score^2 (inertia), stats[1] = sum over centres of shift2, stats[2] = labels changed
__global__ void __launch_bounds__(256) k_kmeans_stats(const float* __restrict__ score, const int* __restrict__ lab,
                                                      int* __restrict__ prev, long long n, const double* __restrict__ shift2,
                                                      int k, double* __restrict__ stats) {
    __shared__ double r0[8], r2[8];
    double a = 0.0, ch = 0.0;
    for (long long i = threadIdx.x; i < n; i += 256) {
        a = fma((double)score[i], (double)score[i], a);
        if (lab[i] != prev[i]) ch += 1.0;
        prev[i] = lab[i];
    }
    a = warp_sum(a);
    ch = warp_sum(ch);
    if ((threadIdx.x & 31) == 0) { r0[threadIdx.x >> 5] = a; r2[threadIdx.x >> 5] = ch; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double x = 0.0, z = 0.0, y = 0.0;
        for (int q = 0; q < 8; ++q) { x += r0[q]; z += r2[q]; }
        for (int j = 0; j < k; ++j) y += shift2[j];
        stats[0] = x; stats[1] = y; stats[2] = z;
    }
}

// "username" branch of create_adjacency_matrix (matrix_operations.py:55-72): rows with the same group id are
// neighbours of each other.  group[i] < 0 = row without a user name (no edges).
__global__ void __launch_bounds__(256) k_group_adjacency(const int* __restrict__ group, long long W,
                                                         unsigned char* __restrict__ adj) {
    for (long long e = (long long)blockIdx.x * 256 + threadIdx.x; e < W * W; e += (long long)gridDim.x * 256) {
        const long long i = e / W, j = e % W;
        const int gi = group[i];
        if (gi >= 0 && i != j && gi == group[j]) adj[e] = 1;
    }
}

// "tags" branch (:74-89, jaccard_similarity :245-248): row i's tag set = sorted unique token ids
// ids[offsets[i] .. offsets[i+1]).  S[i][j] = |A & B| / |A u B| as the float64 quotient of the two integers (the
// reference's Python division), 0 when either set is empty, -1 on the diagonal (:88).
__global__ void __launch_bounds__(256) k_jaccard_keys(const int* __restrict__ offsets, const int* __restrict__ ids,
                                                      long long W, double* __restrict__ S) {
    for (long long e = (long long)blockIdx.x * 256 + threadIdx.x; e < W * W; e += (long long)gridDim.x * 256) {
        const long long i = e / W, j = e % W;
        double key;
        if (i == j) key = -1.0;
        else {
            int a = offsets[i], b = offsets[j];
            const int ae = offsets[i + 1], be = offsets[j + 1];
            const int na = ae - a, nb = be - b;
            int common = 0;
            while (a < ae && b < be) {
                const int x = ids[a], y = ids[b];
                common += (x == y);
                a += (x <= y);
                b += (y <= x);
            }
            key = (na == 0 || nb == 0) ? 0.0 : (double)common / (double)(na + nb - common);
        }
        S[e] = key;
    }
}

// Scratch comes from the stream-ordered pool; by default that pool hands its memory back to the
// driver at every synchronisation, which makes the next cudaMallocAsync map fresh pages (slow and
// erratic).  Keep the pages: one attribute on the device's default pool, set once per device.
static void keep_pool_memory() {
    static bool done[64] = {false};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64 || done[dev]) return;
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
        unsigned long long keep = ~0ull;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    }
    done[dev] = true;
}

static int sm_count_of_current() {
    int dev = 0, n = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    return n;
}

}  // namespace mused

using namespace mused;
extern "C" void mused_count_launches(uint64_t n);

// tensor-core path of mused_assign (assign_tc.cuh)
template <int MH, int NST, int PASSES = 3>
static int launch_assign_tc(const float* X, int64_t n, int64_t d, int64_t ld_x, int NP, const float* Chi,
                            const float* Clo, atc::Args& A, cudaStream_t st, int max_ctas = 0) {
    using CF = atc::Cfg<MH, NST, PASSES>;
    CUtensorMap tx, thi, tlo;
    if (encode_tmap_f32_2d(&tx, X, (uint64_t)n, (uint64_t)d, (uint64_t)ld_x, 128) != CUDA_SUCCESS ||
        encode_tmap_f32_2d(&thi, Chi, (uint64_t)NP, (uint64_t)d, (uint64_t)d, (uint32_t)NP) != CUDA_SUCCESS ||
        encode_tmap_f32_2d(&tlo, Clo, (uint64_t)NP, (uint64_t)d, (uint64_t)d, (uint32_t)NP) != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed for the assignment operands");
        return MUSED_ERR_CUDA;
    }
    const size_t smem = CF::smem_bytes(NP);
    MUSED_CUDA(cudaFuncSetAttribute(atc::k_assign_tc<MH, NST, PASSES>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long tiles = (n + MH * 128 - 1) / (MH * 128);
    const long long grid = max_ctas > 0 ? std::min<long long>(tiles, max_ctas) : tiles;   // max_ctas: persistent tile loop
    atc::k_assign_tc<MH, NST, PASSES><<<(unsigned)grid, CF::NTHREADS, smem, st>>>(tx, thi, tlo, A);
    MUSED_CUDA(cudaGetLastError());
    return 0;
}

// rows the screening kernel left to k_assign_rescore, cumulative per device (mused_assign_pending)
static __device__ unsigned long long g_assign_pending;
static int assign_pending_addr(unsigned long long** out) {
    static std::atomic<unsigned long long*> addr[64];
    int dev = 0;
    MUSED_CUDA(cudaGetDevice(&dev));
    MUSED_REQUIRE(dev >= 0 && dev < 64, "device ordinal %d out of range", dev);
    unsigned long long* p = addr[dev].load(std::memory_order_acquire);
    if (!p) {
        MUSED_CUDA(cudaGetSymbolAddress((void**)&p, g_assign_pending));
        addr[dev].store(p, std::memory_order_release);
    }
    *out = p;
    return 0;
}

// ---- scratch of mused_topk_adjacency_static: lives in the module image, one copy on every GPU that loads the
// library, from load to exit.  [0, W*W) = the key matrix S (row stride W), [MAX_W^2, MAX_W^2 + W) = squared row norms.
// The NAME below is a host-side stand-in in host code: the device address comes from cudaGetSymbolAddress.
// MUSED_TOPK_STATIC_MAX_W (include/mused_b200.h) = 1024 rows -> 4 MiB + 4 KiB
static __device__ __align__(256) float g_topk_static_scratch[(size_t)MUSED_TOPK_STATIC_MAX_W * MUSED_TOPK_STATIC_MAX_W +
                                                             MUSED_TOPK_STATIC_MAX_W];

// device address of the scratch on the CURRENT device, resolved once per device
static int topk_static_scratch(float** out) {
    static std::atomic<float*> addr[64];
    int dev = 0;
    MUSED_CUDA(cudaGetDevice(&dev));
    MUSED_REQUIRE(dev >= 0 && dev < 64, "device ordinal %d out of range", dev);
    float* p = addr[dev].load(std::memory_order_acquire);
    if (!p) {
        MUSED_CUDA(cudaGetSymbolAddress((void**)&p, g_topk_static_scratch));
        MUSED_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned a NULL scratch address");
        addr[dev].store(p, std::memory_order_release);
    }
    *out = p;
    return 0;
}

extern "C" {

int mused_assign(const float* X, int64_t n, int64_t d, int64_t ld_x, const float* C, int64_t k,
                 int64_t ld_c, int metric, float tau, int32_t* labels, float* score, void* stream) {
    MUSED_REQUIRE(n >= 0 && d >= 1 && k >= 1 && ld_x >= d && ld_c >= d, "bad shape");
    MUSED_REQUIRE(X && C && labels, "NULL argument");
    MUSED_REQUIRE(metric == MUSED_METRIC_COSINE || metric == MUSED_METRIC_L2, "unknown metric %d", metric);
    if (n == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    keep_pool_memory();
    // tensor-core path: K in whole 128-byte rows, every centroid in one UMMA N, TMA-addressable X
    const int NP = (int)((k + 15) / 16 * 16);
    if (d % 32 == 0 && k <= 256 && ld_x % 4 == 0 && ((uintptr_t)X & 15) == 0 && n < (1ll << 31) &&
        !getenv("MUSED_ASSIGN_SIMT")) {
        float *Chi = nullptr, *Clo = nullptr, *c2 = nullptr;
        MUSED_CUDA(cudaMallocAsync((void**)&Chi, sizeof(float) * (size_t)NP * d, st));
        MUSED_CUDA(cudaMallocAsync((void**)&Clo, sizeof(float) * (size_t)NP * d, st));
        MUSED_CUDA(cudaMallocAsync((void**)&c2, sizeof(float) * (size_t)NP, st));
        atc::k_assign_prep<<<NP, 256, 0, st>>>(C, k, d, ld_c, NP, Chi, Clo, c2);
        MUSED_CUDA(cudaGetLastError());
        atc::Args A;
        A.n = n; A.k = (int)k; A.NP = NP; A.kblocks = (int)(d / 32); A.metric = metric; A.tau = tau;
        A.cn2 = c2; A.labels = labels; A.score = score; A.cmask = nullptr; A.nw = 0;
        const int sms = sm_count_of_current();
        const size_t cap = 227u * 1024u;
        int rc;
        // labels-only cosine call: one-pass TF32 screening (persistent, one CTA per SM) + float32 re-scoring of the rows
        // whose decision is inside the screening error bound (assign_tc.cuh); MUSED_ASSIGN_EXACT=1 forces 3xTF32
        static int exact = -1;
        if (exact < 0) exact = getenv("MUSED_ASSIGN_EXACT") ? atoi(getenv("MUSED_ASSIGN_EXACT")) : 0;
        if (metric == MUSED_METRIC_COSINE && score == nullptr && !exact && ld_c % 4 == 0 && ((uintptr_t)C & 15) == 0) {
            // 256-row tiles once every SM has one (centroid slices shared by two accumulators), else 128-row tiles;
            // grid = one persistent CTA per SM
            const bool big = (n + 255) / 256 >= sms;
            A.nw = NP / 32 + (NP % 32 != 0);
            MUSED_CUDA(cudaMallocAsync((void**)&A.cmask, sizeof(uint32_t) * (size_t)n * A.nw, st));
            if (big && atc::Cfg<2, 4, 1>::smem_bytes(NP) <= cap) rc = launch_assign_tc<2, 4, 1>(X, n, d, ld_x, NP, Chi, Chi, A, st, sms);
            else if (big && atc::Cfg<2, 3, 1>::smem_bytes(NP) <= cap) rc = launch_assign_tc<2, 3, 1>(X, n, d, ld_x, NP, Chi, Chi, A, st, sms);
            else if (atc::Cfg<1, 4, 1>::smem_bytes(NP) <= cap) rc = launch_assign_tc<1, 4, 1>(X, n, d, ld_x, NP, Chi, Chi, A, st, sms);
            else rc = launch_assign_tc<1, 3, 1>(X, n, d, ld_x, NP, Chi, Chi, A, st, sms);
            if (!rc) {
                unsigned long long* pend = nullptr;
                rc = assign_pending_addr(&pend);
                if (!rc) {
                    const int grid = (int)std::min<long long>((n + 255) / 256, 8ll * sms);
                    atc::k_assign_rescore<<<grid, 256, 0, st>>>(X, n, d, ld_x, C, (int)k, ld_c, c2, tau, labels, A.cmask, A.nw, pend);
                    if (cudaGetLastError() != cudaSuccess) { set_error("k_assign_rescore launch failed"); rc = MUSED_ERR_CUDA; }
                }
            }
            MUSED_CUDA(cudaFreeAsync(A.cmask, st));
            MUSED_CUDA(cudaFreeAsync(Chi, st));
            MUSED_CUDA(cudaFreeAsync(Clo, st));
            MUSED_CUDA(cudaFreeAsync(c2, st));
            if (rc) return rc;
            mused_count_launches(3);
            return 0;
        }
        // tile shape (measured on B200, k = 150, d = 1024): two 128-row accumulators per CTA halve the
        // centroid re-reads from L2 and win once every SM has a tile (2.43 TB/s at 1M rows against
        // 2.08 TB/s); a window-sized call (10k rows) wants more, smaller CTAs with a deeper pipeline.
        static int shape = -1;
        if (shape < 0) shape = getenv("MUSED_ASSIGN_SHAPE") ? atoi(getenv("MUSED_ASSIGN_SHAPE")) : 0;
        const bool fits22 = atc::Cfg<2, 2>::smem_bytes(NP) <= cap;
        const bool big = (n + 255) / 256 >= sms;
        if (fits22 && (shape == 22 || (shape == 0 && big))) rc = launch_assign_tc<2, 2>(X, n, d, ld_x, NP, Chi, Clo, A, st);
        else if (shape != 12 && atc::Cfg<1, 4>::smem_bytes(NP) <= cap) rc = launch_assign_tc<1, 4>(X, n, d, ld_x, NP, Chi, Clo, A, st);
        else if (shape != 12 && atc::Cfg<1, 3>::smem_bytes(NP) <= cap) rc = launch_assign_tc<1, 3>(X, n, d, ld_x, NP, Chi, Clo, A, st);
        else rc = launch_assign_tc<1, 2>(X, n, d, ld_x, NP, Chi, Clo, A, st);
        MUSED_CUDA(cudaFreeAsync(Chi, st));
        MUSED_CUDA(cudaFreeAsync(Clo, st));
        MUSED_CUDA(cudaFreeAsync(c2, st));
        if (rc) return rc;
        mused_count_launches(2);
        return 0;
    }
    float* cn2 = nullptr;
    MUSED_CUDA(cudaMallocAsync((void**)&cn2, sizeof(float) * (size_t)k, st));
    k_row_norm2_f32<<<(unsigned)((k + 7) / 8), 256, 0, st>>>(C, k, d, ld_c, cn2);
    MUSED_CUDA(cudaGetLastError());
    const long long tiles = (n + AT - 1) / AT;
    const int grid = (int)std::min<long long>(tiles, 16ll * sm_count_of_current());
    k_assign<<<grid, 256, 0, st>>>(X, n, d, ld_x, C, k, ld_c, cn2, metric, tau, labels, score);
    MUSED_CUDA(cudaGetLastError());
    MUSED_CUDA(cudaFreeAsync(cn2, st));
    mused_count_launches(2);
    return 0;
}

int mused_assign_pending(uint64_t* n_rows) {
    MUSED_REQUIRE(n_rows, "NULL argument");
    unsigned long long v = 0;
    MUSED_CUDA(cudaMemcpyFromSymbol(&v, g_assign_pending, sizeof(v)));
    *n_rows = (uint64_t)v;
    return 0;
}

int mused_topk_adjacency(const float* feats, int64_t W, int64_t f, int64_t ld, int metric, int64_t k,
                         uint8_t* adj, void* stream) {
    MUSED_REQUIRE(W >= 1 && f >= 1 && ld >= f && k >= 0, "bad shape");
    MUSED_REQUIRE(feats && adj, "NULL argument");
    MUSED_REQUIRE(metric == MUSED_METRIC_COSINE || metric == MUSED_METRIC_L2, "unknown metric %d", metric);
    MUSED_REQUIRE(W * 4 <= 200 * 1024, "window of %lld rows exceeds the per-row shared-memory select (max 51200)", (long long)W);
    if (k == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    keep_pool_memory();
    float *n2 = nullptr, *S = nullptr;
    MUSED_CUDA(cudaMallocAsync((void**)&n2, sizeof(float) * (size_t)W, st));
    MUSED_CUDA(cudaMallocAsync((void**)&S, sizeof(float) * (size_t)W * W, st));
    k_row_norm2_f32<<<(unsigned)((W + 7) / 8), 256, 0, st>>>(feats, W, f, ld, n2);
    MUSED_CUDA(cudaGetLastError());
    const long long tiles = ((W + AT - 1) / AT) * ((W + AT - 1) / AT);
    const int sms = sm_count_of_current();
    k_pair_keys<<<(int)std::min<long long>(tiles, 16ll * sms), 256, 0, st>>>(feats, W, f, ld, n2, metric, S);
    MUSED_CUDA(cudaGetLastError());
    const size_t smem = (size_t)W * sizeof(float);
    MUSED_CUDA(cudaFuncSetAttribute(k_topk_rows<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_topk_rows<float><<<(int)std::min<long long>(W, 8ll * sms), 256, smem, st>>>(S, W, k, adj);
    MUSED_CUDA(cudaGetLastError());
    MUSED_CUDA(cudaFreeAsync(S, st));
    MUSED_CUDA(cudaFreeAsync(n2, st));
    mused_count_launches(3);
    return 0;
}

// The same three launches as mused_topk_adjacency with the scratch in g_topk_static_scratch: no cudaMalloc /
// cudaMallocAsync / cudaFree on this path.  Both scratch regions are fully overwritten for rows and columns
// [0, W) before they are read, so nothing left by the previous call is ever consumed.
int mused_topk_adjacency_static(const float* feats, int64_t W, int64_t f, int64_t ld, int metric, int64_t k,
                                uint8_t* adj, void* stream) {
    MUSED_REQUIRE(W >= 1 && f >= 1 && ld >= f && k >= 0, "bad shape");
    MUSED_REQUIRE(feats && adj, "NULL argument");
    MUSED_REQUIRE(metric == MUSED_METRIC_COSINE || metric == MUSED_METRIC_L2, "unknown metric %d", metric);
    MUSED_REQUIRE(W <= MUSED_TOPK_STATIC_MAX_W,
                  "window of %lld rows exceeds the static scratch of mused_topk_adjacency_static (max %d rows); "
                  "use mused_topk_adjacency", (long long)W, MUSED_TOPK_STATIC_MAX_W);
    if (k == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    float* scratch = nullptr;
    if (int rc = topk_static_scratch(&scratch)) return rc;
    float* S = scratch;
    float* n2 = scratch + (size_t)MUSED_TOPK_STATIC_MAX_W * MUSED_TOPK_STATIC_MAX_W;
    k_row_norm2_f32<<<(unsigned)((W + 7) / 8), 256, 0, st>>>(feats, W, f, ld, n2);
    MUSED_CUDA(cudaGetLastError());
    const long long tiles = ((W + AT - 1) / AT) * ((W + AT - 1) / AT);
    const int sms = sm_count_of_current();
    k_pair_keys<<<(int)std::min<long long>(tiles, 16ll * sms), 256, 0, st>>>(feats, W, f, ld, n2, metric, S);
    MUSED_CUDA(cudaGetLastError());
    // mused_topk_adjacency sets this attribute to ITS window's size on every call (possibly smaller than this one's,
    // which would make the launch below invalid): set it here too.  An attribute of the function, not an allocation.
    const size_t smem = (size_t)W * sizeof(float);
    MUSED_CUDA(cudaFuncSetAttribute(k_topk_rows<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_topk_rows<float><<<(int)std::min<long long>(W, 8ll * sms), 256, smem, st>>>(S, W, k, adj);
    MUSED_CUDA(cudaGetLastError());
    mused_count_launches(3);
    return 0;
}

int mused_topk_adjacency_f64(const double* feats, int64_t W, int64_t f, int64_t ld, int metric, int64_t k,
                             uint8_t* adj, void* stream) {
    MUSED_REQUIRE(W >= 1 && f >= 1 && ld >= f && k >= 0, "bad shape");
    MUSED_REQUIRE(feats && adj, "NULL argument");
    MUSED_REQUIRE(metric == MUSED_METRIC_HAVERSINE || metric == MUSED_METRIC_L1SUM, "unknown float64 metric %d", metric);
    MUSED_REQUIRE(metric != MUSED_METRIC_HAVERSINE || f == 2, "haversine rows are (latitude, longitude)");
    MUSED_REQUIRE(W * 8 <= 200 * 1024, "window of %lld rows exceeds the per-row shared-memory select (max 25600)", (long long)W);
    if (k == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    keep_pool_memory();
    double* S = nullptr;
    MUSED_CUDA(cudaMallocAsync((void**)&S, sizeof(double) * (size_t)W * W, st));
    const int sms = sm_count_of_current();
    k_pair_keys_f64<<<(int)std::min<long long>((W * W + 255) / 256, 16ll * sms), 256, 0, st>>>(feats, W, f, ld, metric, S);
    MUSED_CUDA(cudaGetLastError());
    const size_t smem = (size_t)W * sizeof(double);
    MUSED_CUDA(cudaFuncSetAttribute(k_topk_rows<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_topk_rows<double><<<(int)std::min<long long>(W, 8ll * sms), 256, smem, st>>>(S, W, k, adj);
    MUSED_CUDA(cudaGetLastError());
    MUSED_CUDA(cudaFreeAsync(S, st));
    mused_count_launches(2);
    return 0;
}

int mused_group_adjacency(const int32_t* group, int64_t W, uint8_t* adj, void* stream) {
    MUSED_REQUIRE(W >= 1 && group && adj, "bad argument");
    cudaStream_t st = (cudaStream_t)stream;
    const int grid = (int)std::min<long long>((W * W + 255) / 256, 16ll * sm_count_of_current());
    k_group_adjacency<<<grid, 256, 0, st>>>(group, W, adj);
    MUSED_CUDA(cudaGetLastError());
    mused_count_launches(1);
    return 0;
}

int mused_jaccard_topk_adjacency(const int32_t* offsets, const int32_t* ids, int64_t W, int64_t k, uint8_t* adj,
                                 void* stream) {
    MUSED_REQUIRE(W >= 1 && k >= 0 && offsets && adj, "bad argument");
    MUSED_REQUIRE(W * 8 <= 200 * 1024, "window of %lld rows exceeds the per-row shared-memory select (max 25600)", (long long)W);
    if (k == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    keep_pool_memory();
    double* S = nullptr;
    MUSED_CUDA(cudaMallocAsync((void**)&S, sizeof(double) * (size_t)W * W, st));
    const int sms = sm_count_of_current();
    k_jaccard_keys<<<(int)std::min<long long>((W * W + 255) / 256, 16ll * sms), 256, 0, st>>>(offsets, ids, W, S);
    MUSED_CUDA(cudaGetLastError());
    const size_t smem = (size_t)W * sizeof(double);
    MUSED_CUDA(cudaFuncSetAttribute(k_topk_rows<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_topk_rows<double><<<(int)std::min<long long>(W, 8ll * sms), 256, smem, st>>>(S, W, k, adj);
    MUSED_CUDA(cudaGetLastError());
    MUSED_CUDA(cudaFreeAsync(S, st));
    mused_count_launches(2);
    return 0;
}

int mused_contingency(const int32_t* prev, const int32_t* cur, int64_t n, int32_t P, int32_t Q,
                      int32_t* counts, void* stream) {
    MUSED_REQUIRE(n >= 0 && P >= 1 && Q >= 1, "bad shape");
    MUSED_REQUIRE((prev && cur) || n == 0, "NULL argument");
    MUSED_REQUIRE(counts, "NULL argument");
    cudaStream_t st = (cudaStream_t)stream;
    MUSED_CUDA(cudaMemsetAsync(counts, 0, sizeof(int32_t) * (size_t)P * Q, st));
    if (n == 0) return 0;
    const int grid = (int)std::min<long long>((n + 255) / 256, 4ll * sm_count_of_current());
    k_contingency<<<grid, 256, 0, st>>>(prev, cur, n, P, Q, counts);
    MUSED_CUDA(cudaGetLastError());
    mused_count_launches(1);
    return 0;
}

int mused_kmeans(const float* X, int64_t n, int64_t d, int64_t ld, int32_t k, uint64_t seed, int32_t max_iter, double tol,
                 float* centroids, int32_t* labels, double* inertia, int32_t* n_iter, void* stream) {
    MUSED_REQUIRE(n >= 1 && d >= 1 && ld >= d && k >= 1 && k <= n, "bad shape (n=%lld d=%lld k=%d)", (long long)n, (long long)d, k);
    MUSED_REQUIRE(X && centroids && labels, "NULL argument");
    MUSED_REQUIRE(max_iter >= 1 && tol >= 0.0, "bad max_iter / tol");
    cudaStream_t st = (cudaStream_t)stream;
    keep_pool_memory();
    float *mind = nullptr, *score = nullptr;
    int *prev = nullptr, *counts = nullptr;
    double *shift2 = nullptr, *stats = nullptr;
    MUSED_CUDA(cudaMallocAsync((void**)&mind, sizeof(float) * (size_t)n, st));
    MUSED_CUDA(cudaMallocAsync((void**)&score, sizeof(float) * (size_t)n, st));
    MUSED_CUDA(cudaMallocAsync((void**)&prev, sizeof(int) * (size_t)n, st));
    MUSED_CUDA(cudaMallocAsync((void**)&counts, sizeof(int) * (size_t)k, st));
    MUSED_CUDA(cudaMallocAsync((void**)&shift2, sizeof(double) * (size_t)k, st));
    MUSED_CUDA(cudaMallocAsync((void**)&stats, sizeof(double) * 4, st));
    MUSED_CUDA(cudaMemsetAsync(prev, 0xff, sizeof(int) * (size_t)n, st));
    k_kmeans_seed<<<1, 1024, 0, st>>>(X, n, d, ld, k, seed, centroids, mind);
    MUSED_CUDA(cudaGetLastError());
    mused_count_launches(1);
    // tolerance on the centre movement, scaled like sklearn's: tol * mean feature variance is the caller's business
    double h[3] = {0, 0, 0};
    int it = 0, rc = 0;
    for (; it < max_iter; ++it) {
        rc = mused_assign(X, n, d, ld, centroids, k, d, MUSED_METRIC_L2, -1.f, labels, score, stream);
        if (rc) break;
        k_kmeans_update<<<k, 256, 0, st>>>(X, n, d, ld, labels, centroids, shift2, counts);
        k_kmeans_stats<<<1, 256, 0, st>>>(score, labels, prev, n, shift2, k, stats);
        if (cudaGetLastError() != cudaSuccess) { set_error("kmeans kernel launch failed"); rc = MUSED_ERR_CUDA; break; }
        mused_count_launches(2);
        if (cudaMemcpyAsync(h, stats, sizeof(h), cudaMemcpyDeviceToHost, st) != cudaSuccess ||
            cudaStreamSynchronize(st) != cudaSuccess) { set_error("kmeans: reading the convergence record failed"); rc = MUSED_ERR_CUDA; break; }
        if (h[2] == 0.0 || h[1] <= tol) { ++it; break; }
    }
    if (!rc) {
        // labels of the FINAL centres (what fit_predict returns) and their inertia
        rc = mused_assign(X, n, d, ld, centroids, k, d, MUSED_METRIC_L2, -1.f, labels, score, stream);
        if (!rc) {
            MUSED_CUDA(cudaMemsetAsync(shift2, 0, sizeof(double) * (size_t)k, st));
            k_kmeans_stats<<<1, 256, 0, st>>>(score, labels, prev, n, shift2, k, stats);
            MUSED_CUDA(cudaGetLastError());
            MUSED_CUDA(cudaMemcpyAsync(h, stats, sizeof(h), cudaMemcpyDeviceToHost, st));
            MUSED_CUDA(cudaStreamSynchronize(st));
            if (inertia) *inertia = h[0];
            if (n_iter) *n_iter = it;
        }
    }
    cudaFreeAsync(mind, st); cudaFreeAsync(score, st); cudaFreeAsync(prev, st);
    cudaFreeAsync(counts, st); cudaFreeAsync(shift2, st); cudaFreeAsync(stats, st);
    return rc;
}

}  // extern "C"

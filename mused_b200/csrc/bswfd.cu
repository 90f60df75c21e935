// Batched-schedule SWFD engine (north_star's form of the path): the stream is cut
// into chunks of <= l rows; per chunk every FD instance of the level ladder runs
//   K1  k_bstage    rows (any dtype) -> TF32 hi/lo planes of the shared chunk + ||x||^2
//   K5  k_bbook     period swap, snapshot expiry / cap trim, raw-row snapshots, masks
//   K2  k_gram_tc   G = [sketch; chunk][sketch; chunk]^T          (tcgen05, gram_tc.cuh)
//   K3  k_eig_jacobi  eig(G), shrink, dump decisions                (eig_smem.cuh)
//   K4  k_rotate_tc sketch' = T [sketch; chunk] (tcgen05, rotate_tc.cuh), dumped rows -> snapshot ring
// with no host synchronisation between chunks.  The schedule and its arithmetic
// are stated in oracle/swfd_batched_oracle.py; reference lines:
//   swfd/seq_based_swfd.py:47-103, swfd/sliding_windows_frequent_directions.py:56-173,
//   swfd/frequent_directions.py:39-58,86-93, swfd/fastfdwithdump.py:115-131.
#include <math.h>
#include <algorithm>
#include <vector>

#include "bswfd.h"
#include "gram_tc.cuh"
#include "eig_smem.cuh"
#include "rotate_tc.cuh"
#include "nccl_dyn.h"

namespace mused {

// ---------------------------------------------------------------------------
// K1: stage rows [row_off, row_off + m) of the chunk.  One CTA per row, fixed
// reduction order for ||x||^2 (it decides routing, seq_based_swfd.py:76-77).
// ---------------------------------------------------------------------------
template <class Tin>
__global__ void __launch_bounds__(256) k_bstage(const Tin* __restrict__ in, long long ld, long long stream_stride,
                                                long long xrows, long long d_in, long long d,
                                                int row_off, float* __restrict__ xhi,
                                                float* __restrict__ xlo, double* __restrict__ norm2) {
    __shared__ double red[8];
    const int r = blockIdx.x;
    in += (size_t)blockIdx.y * stream_stride;                 // grid.y = stream
    row_off += (int)(blockIdx.y * xrows);
    const Tin* src = in + (size_t)r * ld;
    float* hi = xhi + (size_t)(row_off + r) * d;
    float* lo = xlo + (size_t)(row_off + r) * d;
    double acc = 0.0;
    for (long long c = threadIdx.x; c < d_in; c += 256) {    // columns d_in .. d-1 of the planes stay zero
        const float v = (float)src[c];
        float h, l;
        umma::split_tf32(v, h, l);
        hi[c] = h;
        lo[c] = l;
        acc = fma((double)v, (double)v, acc);
    }
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int i = 0; i < 8; ++i) t += red[i];
        norm2[row_off + r] = t;
    }
}

// fp32 rows with 16-byte aligned, 4-divisible geometry: one WARP per row, 128-bit loads and stores
// (the append is pure streaming: 4 bytes read and 8 written per element).  Same contract as
// k_bstage: hi / lo planes and ||x||^2 in a fixed order (lane partials in element order, then the
// xor tree) that does not depend on how the caller slices the stream.
__global__ void __launch_bounds__(256) k_bstage_v4(const float* __restrict__ in, long long ld, long long stream_stride,
                                                   long long xrows, long long d_in, long long d, long long m,
                                                   int row_off, float* __restrict__ xhi, float* __restrict__ xlo,
                                                   double* __restrict__ norm2) {
    const int lane = threadIdx.x & 31;
    in += (size_t)blockIdx.y * stream_stride;                 // grid.y = stream
    row_off += (int)(blockIdx.y * xrows);
    const long long warps = (long long)gridDim.x * 8;
    for (long long r = (long long)blockIdx.x * 8 + (threadIdx.x >> 5); r < m; r += warps) {
        const float4* src = reinterpret_cast<const float4*>(in + (size_t)r * ld);
        float4* hi = reinterpret_cast<float4*>(xhi + (size_t)(row_off + r) * d);
        float4* lo = reinterpret_cast<float4*>(xlo + (size_t)(row_off + r) * d);
        double acc = 0.0;
        for (long long c = lane; c < d_in / 4; c += 32) {
            const float4 v = __ldcs(src + c);
            float4 h, l;
            umma::split_tf32(v.x, h.x, l.x);
            umma::split_tf32(v.y, h.y, l.y);
            umma::split_tf32(v.z, h.z, l.z);
            umma::split_tf32(v.w, h.w, l.w);
            hi[c] = h;
            lo[c] = l;
            acc = fma((double)v.x, (double)v.x, acc);
            acc = fma((double)v.y, (double)v.y, acc);
            acc = fma((double)v.z, (double)v.z, acc);
            acc = fma((double)v.w, (double)v.w, acc);
        }
        acc = warp_sum(acc);
        if (lane == 0) norm2[row_off + r] = acc;
    }
}

// The same append with the rows staged by TMA (bulk copies): fp32 rows that are contiguous in memory (ld == d) and
// fill the planes' rows (d_in == d).  CTAs of 4 warps (three per SM) loop over tiles of RT whole rows (<= 16 KB): a
// 2-stage ring of bulk loads global -> shared (mbarrier); ONE WARP PER ROW splits the row in shared memory -- hi in
// place, lo into a second buffer; lane-strided float4 in element order and the xor tree, i.e. exactly k_bstage_v4's
// ||x||^2 -- and two bulk stores shared -> global per tile write the planes.  4 bytes read + 8 written per element,
// all through the TMA engine; the SM only converts.
constexpr int STMA_TILE_BYTES = 16 * 1024;
constexpr int STMA_STAGES = 2;
constexpr int STMA_NT = 128;
__global__ void __launch_bounds__(STMA_NT, 3) k_bstage_tma(const float* __restrict__ in, long long stream_stride, long long xrows,
                                                           long long d, long long m, int RT, int row_off,
                                                           float* __restrict__ xhi, float* __restrict__ xlo,
                                                           double* __restrict__ norm2) {
    extern __shared__ __align__(128) unsigned char sraw[];
    float* s_hi = reinterpret_cast<float*>(sraw);                                            // [STAGES][TILE] rows in, hi out
    float* s_lo = reinterpret_cast<float*>(sraw + STMA_STAGES * STMA_TILE_BYTES);            // [STAGES][TILE]
    uint64_t* full = reinterpret_cast<uint64_t*>(sraw + 2 * STMA_STAGES * STMA_TILE_BYTES);
    const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
    // this CTA's tiles: blockIdx.x, blockIdx.x + gridDim.x, ... of ITS stream (grid.y = stream)
    const long long stride = gridDim.x;
    const long long t_begin = blockIdx.x, t_end = (m + RT - 1) / RT;
    const float* src0 = in + (size_t)blockIdx.y * stream_stride;
    const size_t dst_row0 = (size_t)blockIdx.y * xrows + row_off;
    if (tid == 0) {
        for (int i = 0; i < STMA_STAGES; ++i) umma::mbar_init(&full[i], 1);
        umma::fence_mbar_init();
    }
    __syncthreads();
    auto issue = [&](long long t, int st) {
        const long long r0 = t * RT;
        const uint32_t bytes = (uint32_t)(std::min<long long>(RT, m - r0) * d * 4);
        umma::mbar_expect_tx(&full[st], bytes);
        umma::bulk_load_1d(s_hi + (size_t)st * (STMA_TILE_BYTES / 4), src0 + (size_t)r0 * d, bytes, &full[st]);
    };
    if (tid == 0) {
        long long t = t_begin;
        for (int i = 0; i < STMA_STAGES && t < t_end; ++i, t += stride) issue(t, i);
    }
    int it = 0;
    for (long long t = t_begin; t < t_end; t += stride, ++it) {
        const int st = it % STMA_STAGES;
        const long long r0 = t * RT;
        const int rows = (int)std::min<long long>(RT, m - r0);
        umma::mbar_wait(&full[st], (it / STMA_STAGES) & 1);
        float* th = s_hi + (size_t)st * (STMA_TILE_BYTES / 4);
        float* tl = s_lo + (size_t)st * (STMA_TILE_BYTES / 4);
        for (int r = w; r < rows; r += STMA_NT / 32) {
            float4* hi = reinterpret_cast<float4*>(th + (size_t)r * d);
            float4* lo = reinterpret_cast<float4*>(tl + (size_t)r * d);
            double acc = 0.0;
            for (long long c = lane; c < d / 4; c += 32) {
                const float4 v = hi[c];
                float4 h, l;
                umma::split_tf32(v.x, h.x, l.x);
                umma::split_tf32(v.y, h.y, l.y);
                umma::split_tf32(v.z, h.z, l.z);
                umma::split_tf32(v.w, h.w, l.w);
                hi[c] = h;
                lo[c] = l;
                acc = fma((double)v.x, (double)v.x, acc);
                acc = fma((double)v.y, (double)v.y, acc);
                acc = fma((double)v.z, (double)v.z, acc);
                acc = fma((double)v.w, (double)v.w, acc);
            }
            acc = warp_sum(acc);
            if (lane == 0) norm2[dst_row0 + r0 + r] = acc;
        }
        umma::fence_proxy_async();                 // the planes written above -> visible to the bulk stores
        __syncthreads();
        if (tid == 0) {
            const uint32_t bytes = (uint32_t)((size_t)rows * d * 4);
            umma::bulk_store_1d(xhi + (dst_row0 + r0) * d, th, bytes);
            umma::bulk_store_1d(xlo + (dst_row0 + r0) * d, tl, bytes);
            umma::bulk_commit();
            const long long tn = t + (long long)STMA_STAGES * stride;
            if (tn < t_end) {
                umma::bulk_wait_read<0>();         // this stage's buffers have been read by the stores just issued
                issue(tn, st);
            }
        }
        // the next iteration uses the OTHER stage; this stage is touched again only after thread 0 re-armed it and
        // its mbarrier completed, so no further barrier is needed here
    }
    if (tid == 0) umma::bulk_wait<0>();            // the planes are in global memory before the kernel ends
}

// ---------------------------------------------------------------------------
// K5: per-instance bookkeeping of one chunk (times t_a .. t_a + n - 1).
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_bbook(BDev D, int n, unsigned long long t_a) {
    __shared__ int raw_row[BS_MAX_ELL], raw_slot[BS_MAX_ELL];
    __shared__ unsigned char is_raw[BS_MAX_ELL];
    __shared__ int n_raw_s, do_reset;
    const int s = blockIdx.x;
    BSlot& sl = D.slots[s];
    const unsigned long long t_b = t_a + (unsigned long long)n - 1ull;
    unsigned long long* qs = D.ringS + (size_t)s * D.capn;
    unsigned long long* qt = D.ringT + (size_t)s * D.capn;
    // routing test of every row in parallel (seq_based_swfd.py:76-77); the serial part below only walks flags
    const double* nrm2 = D.norm2 + (size_t)(s / D.S1) * D.xrows + D.x0;
    for (int r = threadIdx.x; r < n; r += 256) is_raw[r] = nrm2[r] >= sl.error ? 1 : 0;
    __syncthreads();
    if (threadIdx.x == 0) {
        int reset = 0;
        const unsigned long long p = t_a / D.N;
        if (p > sl.period) {                       // sliding_windows_frequent_directions.py:99-107
            sl.period = p;
            if ((int)(p & 1ull) != sl.parity) {    // this slot becomes the fresh fd_aux
                reset = 1;
                sl.delta = 0.0;
                sl.q_head = 0;
                sl.q_len = 0;
                sl.last_dump = t_a;
                sl.n_flush = 0;
                sl.n_dump = 0;
                sl.n_raw = 0;
            }
        }
        int head = sl.q_head, len = sl.q_len;
        while (len > 0 && ((double)len > D.qcap || qt[head] + D.N <= t_b)) {   // seq_based_swfd.py:59-74
            head = (head + 1) % D.capn;
            --len;
        }
        int nr = 0;
        unsigned long long last = sl.last_dump;
        for (int r = 0; r < n; ++r) {
            if (is_raw[r]) {                       // -> SlidingWindowFD.append (:56-80)
                if (len == D.capn) { head = (head + 1) % D.capn; --len; sl.overflow += 1; }
                const int tail = (head + len) % D.capn;
                qs[tail] = last;
                qt[tail] = t_a + (unsigned long long)r;
                last = t_a + (unsigned long long)r + 1ull;
                ++len;
                raw_row[nr] = r;
                raw_slot[nr] = tail;
                ++nr;
            }
        }
        sl.q_head = head;
        sl.q_len = len;
        sl.last_dump = last;
        sl.n_raw += (unsigned long long)nr;
        sl.npend = n - nr;
        n_raw_s = nr;
        do_reset = reset;
    }
    __syncthreads();
    unsigned char* mk = D.mask + (size_t)s * D.ell;
    for (int r = threadIdx.x; r < D.ell; r += 256) mk[r] = 0;
    __syncthreads();
    const int nr = n_raw_s;
    for (int k = threadIdx.x; k < nr; k += 256) mk[raw_row[k]] = 1;
    if (do_reset) {
        const size_t off = ((size_t)s * 2 + sl.cur) * D.ell * D.d;
        const size_t tot = (size_t)D.ell * D.d;
        for (size_t e = threadIdx.x; e < tot; e += 256) {
            D.skhi[off + e] = 0.f;
            D.sklo[off + e] = 0.f;
        }
        for (int i = threadIdx.x; i < D.ell; i += 256) D.s2[(size_t)s * D.ell + i] = 0.0;
    }
    // the rows themselves are copied into the ring by K4's CTAs (one column tile each), not by this one CTA
    int* rl = D.raw_list + (size_t)s * 2 * D.ell;
    for (int k = threadIdx.x; k < nr; k += 256) {
        rl[k] = raw_row[k];
        rl[D.ell + k] = raw_slot[k];
    }
    if (threadIdx.x == 0) sl.nraw = nr;
}

// split-K reduction of K2's partial Grams, float64, fixed order (deterministic)
__global__ void __launch_bounds__(256) k_gram_reduce(BDev D) {
    const size_t per = (size_t)D.n2 * D.n2;
    const size_t tot = (size_t)D.S * per;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < tot; e += (size_t)gridDim.x * blockDim.x) {
        const size_t s = e / per, ij = e - s * per;
        const float* p = D.Gp + s * D.KS * per + ij;
        double acc = 0.0;
        for (int k = 0; k < D.KS; ++k) acc += (double)p[(size_t)k * per];
        D.G[e] = acc;
    }
}

// ---------------------------------------------------------------------------
// Query support: float64 stack [sketch; live snapshots] of one instance, and the
// plain FD output (sketch, sigma^2, Vt = sketch / sqrt(sigma^2), delta).
// ---------------------------------------------------------------------------
__global__ void k_bgather(BDev D, int s, double* __restrict__ M) {
    const BSlot& sl = D.slots[s];
    const int rows = D.ell + sl.q_len;
    const float* skh = D.skhi + ((size_t)s * 2 + sl.cur) * D.ell * D.d;
    const float* skl = D.sklo + ((size_t)s * 2 + sl.cur) * D.ell * D.d;
    for (int i = blockIdx.x; i < rows; i += gridDim.x) {
        double* dst = M + (size_t)i * D.d;
        if (i < D.ell) {
            for (long long c = threadIdx.x; c < D.d; c += blockDim.x)
                dst[c] = (double)skh[(size_t)i * D.d + c] + (double)skl[(size_t)i * D.d + c];
        } else {
            const float* src = D.ringV + ((size_t)s * D.capn + (sl.q_head + (i - D.ell)) % D.capn) * D.d;
            for (long long c = threadIdx.x; c < D.d; c += blockDim.x) dst[c] = (double)src[c];
        }
    }
}

__global__ void k_bfd_out(BDev D, int s, double* sketch, double* sigma2, double* vt, double* delta) {
    const BSlot& sl = D.slots[s];
    const float* skh = D.skhi + ((size_t)s * 2 + sl.cur) * D.ell * D.d;
    const float* skl = D.sklo + ((size_t)s * 2 + sl.cur) * D.ell * D.d;
    const double* s2 = D.s2 + (size_t)s * D.ell;
    const size_t tot = (size_t)D.ell * D.d;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < tot; e += (size_t)gridDim.x * blockDim.x) {
        const int i = (int)(e / D.d);
        const double v = (double)skh[e] + (double)skl[e];
        sketch[e] = v;
        if (vt) vt[e] = (s2[i] > 0.0) ? v / sqrt(s2[i]) : 0.0;
    }
    if (blockIdx.x == 0) {
        for (int i = threadIdx.x; i < D.ell; i += blockDim.x) sigma2[i] = s2[i];
        if (threadIdx.x == 0) *delta = sl.delta;
    }
}

// ---------------------------------------------------------------------------
// Engine
// ---------------------------------------------------------------------------
struct BEngine {
    BDev D{};
    int device = 0, n_levels = 0, sm_count = 148;
    int K = 1;                        // independent streams advanced in lock-step (same rows per fit() call each)
    long long batch = 0;
    unsigned long long time = 0;      // rows handed to fit()
    unsigned long long done = 0;      // rows folded into the state
    int pend = 0;                     // rows staged in the chunk planes, not yet processed
    CUtensorMap tm_sk_hi, tm_sk_lo, tm_x_hi, tm_x_lo;                 // K2 boxes: l rows x 32 columns
    CUtensorMap tr_t_hi, tr_t_lo, tr_sk_hi, tr_sk_lo, tr_x_hi, tr_x_lo; // K4 boxes
    CUtensorMap ga_sk_hi, ga_sk_lo, ga_x_hi, ga_x_lo, gb_sk_hi, gb_sk_lo, gb_x_hi, gb_x_lo;   // tiled K2 (l = 256)
    unsigned char* raw = nullptr;     // H2D staging of caller rows: [K][raw_rows][d] of the widest dtype
    size_t raw_bytes = 0;
    long long raw_rows = 0;
    long long sb_rows = 0;            // rows of the staging planes (+ ell of slack for the last TMA box)
    long long ahead = 0;              // rows staged beyond the open chunk (inside one fit() call)
    QueryScratch Q;
    double *o_sketch = nullptr, *o_vt = nullptr, *o_s2 = nullptr, *o_delta = nullptr;
    std::vector<StateAlloc> allocs;   // every device allocation; .state = part of the checkpoint (mused_swfd_state_*)
    cudaStream_t last_stream = nullptr;   // stream of the last fit(): size / debug read the slots behind it
    // column-sharded mode: this engine holds the columns of ONE rank; the two exchanges of a chunk
    // (||x||^2 of the rows, the Gram) are NCCL allreduces posted on the engine's stream from here
    ncclComm_t comm = nullptr;
    int comm_rank = 0, comm_size = 1;
    unsigned long long n_allreduce = 0, allreduce_bytes = 0;
    int q_slot = 0, q_ng = 0;         // pending query (get_begin -> get_finish)
    bool q_gram = true;               // get_begin left the stack Gram in Q.G
    bool timing = false;
    std::vector<cudaEvent_t> ev[5];
    double timed_ms[5] = {0, 0, 0, 0, 0};
    double gap_last[2] = {0, 0};
    double gap_ms[2] = {0, 0};        // GPU idle between the kernels of a chunk / between chunks of one fit()
    long long timed_chunks = 0, timed_rows = 0;

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
    ~BEngine() {
        if (comm && nccl_api().ok()) nccl_api().CommDestroy(comm);
        for (int k = 0; k < 5; ++k)
            for (cudaEvent_t e : ev[k]) cudaEventDestroy(e);
        for (const StateAlloc& a : allocs) cudaFree(a.p);
    }
    void mark(int k, cudaStream_t st) {
        if (!timing) return;
        cudaEvent_t e;
        cudaEventCreate(&e);
        cudaEventRecord(e, st);
        ev[k].push_back(e);
    }
};

extern "C" void mused_count_launches(uint64_t n);

static bool stage_simt() {            // MUSED_STAGE_SIMT=1: the SIMT append kernels also for TMA-eligible calls (diagnostic)
    static int v = -1;
    if (v < 0) v = getenv("MUSED_STAGE_SIMT") ? 1 : 0;
    return v == 1;
}

static size_t b_dtype_size(int dt) {
    switch (dt) {
        case MUSED_F64: case MUSED_I64: return 8;
        case MUSED_F32: case MUSED_I32: return 4;
        case MUSED_U8: return 1;
    }
    return 0;
}

int bengine_create(BEngine** out, long long N, long long d, long long ell, double beta, int n_levels,
                   const double* thr, int device, long long batch, int n_streams) {
    int ndev = 0;
    cudaError_t ce = cudaGetDeviceCount(&ndev);
    if (ce != cudaSuccess || ndev == 0) {
        set_error("no CUDA device available (%s); mused_b200 has no CPU fallback",
                  ce != cudaSuccess ? cudaGetErrorString(ce) : "device count is 0");
        return MUSED_ERR_CUDA;
    }
    MUSED_REQUIRE(device >= 0 && device < ndev, "device %d out of range (have %d)", device, ndev);
    MUSED_REQUIRE(ell >= 2 && ell <= BS_MAX_ELL,
                  "schedule=batched supports sketch_dim 2 .. %d (got %lld); use schedule=reference", BS_MAX_ELL, ell);
    MUSED_REQUIRE(d > ell, "schedule=batched needs d > sketch_dim (d=%lld, sketch_dim=%lld)", d, ell);
    // The kernels run on sketch_dim in {16, 32, 64, 128, 256} and d %% 32 == 0; any other shape is embedded in the
    // next such one with zero rows / columns, which changes no Gram entry, eigenvalue or product.
    const long long ell_l = ell, d_l = d;
    ell = ell_l <= 16 ? 16 : (ell_l <= 32 ? 32 : (ell_l <= 64 ? 64 : (ell_l <= 128 ? 128 : 256)));
    d = (d_l + 31) / 32 * 32;
    MUSED_REQUIRE(N >= 1 && beta > 0, "bad N / beta");
    MUSED_REQUIRE(n_levels >= 0 && n_levels <= 64, "n_levels out of range");
    MUSED_REQUIRE(n_streams >= 1 && n_streams <= 1024 && (long long)n_streams * 2 * n_levels <= 65535,
                  "n_streams out of range (1 .. 1024, n_streams * 2 * levels <= 65535)");
    if (batch <= 0) batch = ell_l;
    MUSED_REQUIRE(batch <= ell_l, "batch_rows must be <= sketch_dim");
    MUSED_CUDA(cudaSetDevice(device));
    BEngine* E = new BEngine();
    E->device = device;
    E->n_levels = n_levels;
    E->K = n_streams;
    E->batch = batch;
    cudaDeviceProp prop;
    MUSED_CUDA(cudaGetDeviceProperties(&prop, device));
    E->sm_count = prop.multiProcessorCount;
    BDev& D = E->D;
    const int S1 = 2 * n_levels;
    const int S = S1 * n_streams;
    D.S = S;
    D.S1 = std::max(S1, 1);
    D.ell = (int)ell;
    D.ell_l = (int)ell_l;
    D.n2 = 2 * (int)ell;
    D.d = d;
    D.d_l = d_l;
    D.N = (unsigned long long)N;
    D.qcap = (2.0 + 8.0 / beta) * (double)ell_l;   // seq_based_swfd.py:61
    D.capn = (int)floor(D.qcap) + (int)batch + (int)ell + 2;
    // split-K: enough CTAs to cover the SMs, K slices of whole 32-float blocks
    int KS = 16;
    // (sized on ONE stream's instances so that a stream's arithmetic does not depend on how many share the handle)
    const int gram_tiles = ell > 128 ? (int)((2 * ell / GramTiledCfg::TM) * (2 * ell / GramTiledCfg::TN)) : 1;
    while (KS > 1 && (d % (32 * KS) != 0 || (long long)S1 * KS * gram_tiles > 1LL * E->sm_count)) KS >>= 1;
    if (getenv("MUSED_GRAM_KS")) {
        const int want = atoi(getenv("MUSED_GRAM_KS"));
        if (want >= 1 && want <= 16 && d % (32 * want) == 0) KS = want;
    }
    D.KS = KS;
    int rc = 0;
#define A_(p, n) if ((rc = E->alloc(&(p), (n))) != 0) { delete E; return rc; }
#define S_(p, n) if ((rc = E->alloc(&(p), (n), false)) != 0) { delete E; return rc; }   /* scratch: not checkpointed */
    const size_t Sa = (size_t)std::max(S, 1);
    A_(D.slots, Sa);
    A_(D.skhi, Sa * 2 * ell * d);
    A_(D.sklo, Sa * 2 * ell * d);
    E->sb_rows = std::max<long long>(3 * ell, std::min<long long>(8192, (64ll << 20) / (d * 8 * n_streams)));
    D.xrows = E->sb_rows + ell;
    A_(D.xhi, (size_t)n_streams * D.xrows * d);
    A_(D.xlo, (size_t)n_streams * D.xrows * d);
    A_(D.norm2, (size_t)n_streams * D.xrows);
    D.x0 = 0;
    A_(D.mask, Sa * ell);
    S_(D.Gp, Sa * KS * D.n2 * D.n2);
    S_(D.G, Sa * D.n2 * D.n2);
    S_(D.Gs, Sa * D.n2 * D.n2);
    S_(D.Thi, Sa * ell * D.n2 + 128 * D.n2);   // + slack: the 128-row T box of the last instance
    S_(D.Tlo, Sa * ell * D.n2 + 128 * D.n2);
    A_(D.s2, Sa * ell);
    A_(D.dump_slot, Sa * ell);
    A_(D.raw_list, Sa * 2 * ell);
    A_(D.ringV, Sa * D.capn * d);
    A_(D.ringS, Sa * D.capn);
    A_(D.ringT, Sa * D.capn);
    E->raw_rows = std::max<long long>(ell, std::min<long long>(8192, (64ll << 20) / (d * 8 * n_streams)));
    E->raw_bytes = (size_t)n_streams * E->raw_rows * d * 8;
    S_(E->raw, E->raw_bytes);
    QueryScratch& Q = E->Q;
    Q.ng_max = D.ell + D.capn + 1;
    Q.ng_max += Q.ng_max & 1;
    Q.lda = Q.ng_max;
    Q.rel_thr = 1e-11;     // the fp32-product Grams upstream carry ~1e-6; 1e-11 leaves eigenvalues exact to ~1e-20
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
    S_(E->o_sketch, (size_t)ell * d);
    S_(E->o_vt, (size_t)ell * d);
    S_(E->o_s2, (size_t)ell);
    S_(E->o_delta, 1);
#undef A_
#undef S_
    std::vector<BSlot> init(Sa);
    memset(init.data(), 0, sizeof(BSlot) * Sa);
    for (int s = 0; s < S; ++s) {
        init[s].error = thr[(s % S1) / 2];
        init[s].last_dump = 1;       // sliding_windows_frequent_directions.py:23
        init[s].parity = (s % S1) & 1;
        init[s].skip = 1;
    }
    MUSED_CUDA(cudaMemcpy(D.slots, init.data(), sizeof(BSlot) * Sa, cudaMemcpyHostToDevice));
    if (ell > 128) {
        // tiled Gram (n2 = 512): A boxes of 128 rows, B boxes of 256 rows
        CUresult q[8];
        const uint64_t xt = (uint64_t)n_streams * (uint64_t)D.xrows;
        q[0] = encode_tmap_f32_2d(&E->ga_sk_hi, D.skhi, (uint64_t)Sa * 2 * ell, (uint64_t)d, (uint64_t)d, 128);
        q[1] = encode_tmap_f32_2d(&E->ga_sk_lo, D.sklo, (uint64_t)Sa * 2 * ell, (uint64_t)d, (uint64_t)d, 128);
        q[2] = encode_tmap_f32_2d(&E->ga_x_hi, D.xhi, xt, (uint64_t)d, (uint64_t)d, 128);
        q[3] = encode_tmap_f32_2d(&E->ga_x_lo, D.xlo, xt, (uint64_t)d, (uint64_t)d, 128);
        q[4] = encode_tmap_f32_2d(&E->gb_sk_hi, D.skhi, (uint64_t)Sa * 2 * ell, (uint64_t)d, (uint64_t)d, 256);
        q[5] = encode_tmap_f32_2d(&E->gb_sk_lo, D.sklo, (uint64_t)Sa * 2 * ell, (uint64_t)d, (uint64_t)d, 256);
        q[6] = encode_tmap_f32_2d(&E->gb_x_hi, D.xhi, xt, (uint64_t)d, (uint64_t)d, 256);
        q[7] = encode_tmap_f32_2d(&E->gb_x_lo, D.xlo, xt, (uint64_t)d, (uint64_t)d, 256);
        for (int i = 0; i < 8; ++i)
            if (q[i] != CUDA_SUCCESS) {
                set_error("cuTensorMapEncodeTiled failed for the tiled Gram (%d)", (int)q[i]);
                delete E;
                return MUSED_ERR_CUDA;
            }
    }
    CUresult r1 = encode_tmap_f32_2d(&E->tm_sk_hi, D.skhi, (uint64_t)Sa * 2 * ell, (uint64_t)d, (uint64_t)d, (uint32_t)ell);
    CUresult r2 = encode_tmap_f32_2d(&E->tm_sk_lo, D.sklo, (uint64_t)Sa * 2 * ell, (uint64_t)d, (uint64_t)d, (uint32_t)ell);
    const uint64_t xtot = (uint64_t)n_streams * (uint64_t)D.xrows;
    CUresult r3 = encode_tmap_f32_2d(&E->tm_x_hi, D.xhi, xtot, (uint64_t)d, (uint64_t)d, (uint32_t)ell);
    CUresult r4 = encode_tmap_f32_2d(&E->tm_x_lo, D.xlo, xtot, (uint64_t)d, (uint64_t)d, (uint32_t)ell);
    {
        const uint32_t trows = (uint32_t)std::min<long long>(ell, 128), brows = (uint32_t)std::min<long long>(ell, 32);
        CUresult q[6];
        q[0] = encode_tmap_f32_2d(&E->tr_t_hi, D.Thi, (uint64_t)Sa * ell, (uint64_t)D.n2, (uint64_t)D.n2, trows);
        q[1] = encode_tmap_f32_2d(&E->tr_t_lo, D.Tlo, (uint64_t)Sa * ell, (uint64_t)D.n2, (uint64_t)D.n2, trows);
        q[2] = encode_tmap_f32_2d(&E->tr_sk_hi, D.skhi, (uint64_t)Sa * 2 * ell, (uint64_t)d, (uint64_t)d, brows, 32, true);
        q[3] = encode_tmap_f32_2d(&E->tr_sk_lo, D.sklo, (uint64_t)Sa * 2 * ell, (uint64_t)d, (uint64_t)d, brows, 32, true);
        q[4] = encode_tmap_f32_2d(&E->tr_x_hi, D.xhi, xtot, (uint64_t)d, (uint64_t)d, brows, 32, true);
        q[5] = encode_tmap_f32_2d(&E->tr_x_lo, D.xlo, xtot, (uint64_t)d, (uint64_t)d, brows, 32, true);
        for (int i = 0; i < 6; ++i)
            if (q[i] != CUDA_SUCCESS) r1 = q[i];
    }
    if (r1 != CUDA_SUCCESS || r2 != CUDA_SUCCESS || r3 != CUDA_SUCCESS || r4 != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed (%d %d %d %d)", (int)r1, (int)r2, (int)r3, (int)r4);
        delete E;
        return MUSED_ERR_CUDA;
    }
    *out = E;
    return 0;
}

void bengine_destroy(BEngine* E) {
    if (!E) return;
    cudaSetDevice(E->device);
    delete E;
}

template <int N2>
static int launch_gram(BEngine& E, cudaStream_t st) {
    using C = GramCfg<N2>;
    MUSED_SMEM_ATTR(k_gram_tc<N2>, C::SMEM_BYTES);
    GramArgs A;
    A.slots = E.D.slots;
    A.Gp = E.D.Gp;
    A.ell = E.D.ell;
    A.KS = E.D.KS;
    A.kblocks = (int)(E.D.d / (32 * E.D.KS));
    A.x0 = E.D.x0;
    A.S1 = E.D.S1;
    A.xrows = (int)E.D.xrows;
    k_gram_tc<N2><<<dim3(E.D.KS, E.D.S), 192, C::SMEM_BYTES, st>>>(E.tm_sk_hi, E.tm_sk_lo, E.tm_x_hi, E.tm_x_lo, A);
    MUSED_CUDA(cudaGetLastError());
    return 0;
}

static int launch_gram_tiled(BEngine& E, cudaStream_t st) {
    using C = GramTiledCfg;
    MUSED_SMEM_ATTR(k_gram_tiled, C::SMEM_BYTES);
    GramArgs A;
    A.slots = E.D.slots;
    A.Gp = E.D.Gp;
    A.ell = E.D.ell;
    A.KS = E.D.KS;
    A.kblocks = (int)(E.D.d / (32 * E.D.KS));
    A.x0 = E.D.x0;
    A.S1 = E.D.S1;
    A.xrows = (int)E.D.xrows;
    const int tiles = (E.D.n2 / C::TM) * (E.D.n2 / C::TN);
    k_gram_tiled<<<dim3(E.D.KS, E.D.S, tiles), 192, C::SMEM_BYTES, st>>>(E.ga_sk_hi, E.ga_sk_lo, E.ga_x_hi, E.ga_x_lo,
                                                                       E.gb_sk_hi, E.gb_sk_lo, E.gb_x_hi, E.gb_x_lo, A);
    MUSED_CUDA(cudaGetLastError());
    return 0;
}

template <int EPL>
static int launch_eig(BEngine& E, int n, unsigned long long t_end, cudaStream_t st) {
    static const int mixed = getenv("MUSED_EIG_MIXED") ? atoi(getenv("MUSED_EIG_MIXED")) : 1;   // 0 = float64 sweeps only
    const int smem = (int)ESLayout<EPL>::BYTES;
    MUSED_SMEM_ATTR(k_eig_jacobi<EPL>, smem);
    if (EPL > 8) {                             // clusters of 16 CTAs are outside the portable limit of 8
        static std::atomic<unsigned long long> np_done{0ull};
        int dev = 0;
        MUSED_CUDA(cudaGetDevice(&dev));
        if (!((np_done.load() >> (dev & 63)) & 1ull)) {
            MUSED_CUDA(cudaFuncSetAttribute(k_eig_jacobi<EPL>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
            np_done.fetch_or(1ull << (dev & 63));
        }
    }
    EigArgs A;
    A.slots = E.D.slots;
    A.G = E.D.G;
    A.ell = E.D.ell;
    A.ell_l = E.D.ell_l;
    A.n2 = E.D.n2;
    A.capn = E.D.capn;
    A.mask = E.D.mask;
    A.n_rows = n;
    A.t_end = t_end;
    A.Thi = E.D.Thi;
    A.Tlo = E.D.Tlo;
    A.s2 = E.D.s2;
    A.dump_slot = E.D.dump_slot;
    A.ringS = E.D.ringS;
    A.ringT = E.D.ringT;
    A.tol = 1e-3;          // quadratic phase: the sweep that measures <= 1e-3 leaves ~1e-6 (Gram accuracy)
    A.max_sweeps = 24;
    A.mixed = mixed;
    A.tol32 = getenv("MUSED_EIG_TOL32") ? atof(getenv("MUSED_EIG_TOL32")) : 4e-3;   // measured at configs[1]: 1e-3 costs one more float32 sweep (+4.6%), 1e-2 a second float64 one
    A.small_col = getenv("MUSED_EIG_SMALLCOL") ? atof(getenv("MUSED_EIG_SMALLCOL")) : 1e-14;
    A.Gs = E.D.Gs;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(E.D.S * EPL));
    cfg.blockDim = dim3(ES_NT);
    cfg.dynamicSmemBytes = (size_t)smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = EPL;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    MUSED_CUDA(cudaLaunchKernelEx(&cfg, k_eig_jacobi<EPL>, A));
    return 0;
}

// Fold the staged chunk (E.pend rows, times done+1 ..) into every FD instance, in two
// phases: everything up to the Gram (a column-sharded run adds the other shards'
// Grams in between), then eigensolve + rotate.
int bengine_chunk_gram(BEngine* Ep, cudaStream_t st) {
    BEngine& E = *Ep;
    const int n = E.pend;
    if (n == 0 || E.D.S == 0) return 0;
    BDev& D = E.D;
    const unsigned long long t_a = E.done + 1ull;
    int rc = 0;
    E.mark(1, st);
    k_bbook<<<D.S, 256, 0, st>>>(D, n, t_a);
    MUSED_CUDA(cudaGetLastError());
    E.mark(1, st);
    E.mark(2, st);
    switch (D.ell) {
        case 16: rc = launch_gram<32>(E, st); break;
        case 32: rc = launch_gram<64>(E, st); break;
        case 64: rc = launch_gram<128>(E, st); break;
        case 128: rc = launch_gram<256>(E, st); break;
        default: rc = launch_gram_tiled(E, st); break;
    }
    if (rc) return rc;
    k_gram_reduce<<<std::min<long long>(((long long)D.S * D.n2 * D.n2 + 255) / 256, 8ll * E.sm_count), 256, 0, st>>>(D);
    MUSED_CUDA(cudaGetLastError());
    E.mark(2, st);
    mused_count_launches(3);
    return 0;
}

int bengine_chunk_finish(BEngine* Ep, cudaStream_t st) {
    BEngine& E = *Ep;
    const int n = E.pend;
    if (n == 0) return 0;
    BDev& D = E.D;
    int rc = 0;
    if (D.S > 0) {
        const unsigned long long t_end = E.done + (unsigned long long)n;
        E.mark(3, st);
        switch (D.ell) {
            case 16: rc = launch_eig<1>(E, n, t_end, st); break;
            case 32: rc = launch_eig<2>(E, n, t_end, st); break;
            case 64: rc = launch_eig<4>(E, n, t_end, st); break;
            case 128: rc = launch_eig<8>(E, n, t_end, st); break;
            default: rc = launch_eig<16>(E, n, t_end, st); break;
        }
        if (rc) return rc;
        E.mark(3, st);
        E.mark(4, st);
        MUSED_SMEM_ATTR(k_rotate_tc, RotCfg::SMEM_BYTES);
        k_rotate_tc<<<dim3((unsigned)((D.d + RotCfg::NT - 1) / RotCfg::NT), (unsigned)D.S, (unsigned)std::max(1, D.ell / 128)), 192, RotCfg::SMEM_BYTES, st>>>(
            E.tr_t_hi, E.tr_t_lo, E.tr_sk_hi, E.tr_sk_lo, E.tr_x_hi, E.tr_x_lo, D);
        MUSED_CUDA(cudaGetLastError());
        E.mark(4, st);
        mused_count_launches(2);
    }
    if (E.timing) { E.timed_chunks += 1; E.timed_rows += n; }
    E.done += (unsigned long long)n;
    E.pend = 0;
    E.D.x0 += n;                       // the next chunk starts behind this one in the staging planes
    return 0;
}

static int allreduce_f64(BEngine& E, double* buf, size_t count, cudaStream_t st) {
    MUSED_NCCL(nccl_api().AllReduce(buf, buf, count, ncclDouble, ncclSum, E.comm, st));
    E.n_allreduce += 1;
    E.allreduce_bytes += count * sizeof(double);
    return 0;
}

static int process_chunk(BEngine& E, cudaStream_t st) {
    if (E.comm && E.pend > 0 && E.D.S > 0) {
        // exchange 1: global ||x||^2 of the chunk's rows (routing rule X@X.T >= threshold, seq_based_swfd.py:77)
        int rc = allreduce_f64(E, E.D.norm2 + E.D.x0, (size_t)E.pend, st);
        if (rc) return rc;
        if ((rc = bengine_chunk_gram(&E, st))) return rc;
        // exchange 2: G = sum over the column shards, the one real exchange per shrink; same stream, no host sync
        if ((rc = allreduce_f64(E, E.D.G, (size_t)E.D.S * E.D.n2 * E.D.n2, st))) return rc;
        return bengine_chunk_finish(&E, st);
    }
    int rc = bengine_chunk_gram(&E, st);
    if (rc) return rc;
    return bengine_chunk_finish(&E, st);
}

// rows the current chunk may still take before it must be processed
static long long chunk_room(const BEngine& E) {
    const unsigned long long t_a = E.done + 1ull;
    const unsigned long long N = E.D.N;
    const unsigned long long to_boundary = (t_a % N) ? N - (t_a % N) : N;   // rows until the next multiple of N
    const long long cap = (long long)std::min<unsigned long long>((unsigned long long)E.batch, to_boundary);
    return cap - E.pend;
}

// K1 launch: m rows behind everything already staged.  logical = the rows join the open chunk at
// once (phase API, one chunk at a time); otherwise they wait in the planes (fit() stages a whole
// block of caller rows with ONE launch and then walks it chunk by chunk).
static int stage_rows(BEngine& E, const void* dev_rows, long long m, long long ld, long long stream_stride, int dtype,
                      bool logical, cudaStream_t st) {
    BDev& D = E.D;
    if (E.pend == 0 && E.ahead == 0) D.x0 = 0;
    const int row_off = D.x0 + E.pend + (int)E.ahead;
    MUSED_REQUIRE(row_off + m <= E.sb_rows + D.ell, "internal: staging planes overflow (%d + %lld rows)", row_off, m);
    E.mark(0, st);
    const dim3 gr((unsigned)m, (unsigned)E.K);
#define STAGE_(T) k_bstage<T><<<gr, 256, 0, st>>>((const T*)dev_rows, ld, stream_stride, D.xrows, D.d_l, D.d, row_off, D.xhi, D.xlo, D.norm2)
    switch (dtype) {
        case MUSED_F64: STAGE_(double); break;
        case MUSED_F32:
            if (ld == D.d && D.d_l == D.d && D.d * 4 <= STMA_TILE_BYTES && stream_stride % 4 == 0 &&
                ((uintptr_t)dev_rows & 15) == 0 && m >= 64 && !stage_simt()) {
                // TMA-staged append (contiguous fp32 rows): persistent CTAs, whole-row tiles of <= 16 KB
                const int RT = (int)(STMA_TILE_BYTES / (D.d * 4));
                const size_t smem = 2 * STMA_STAGES * STMA_TILE_BYTES + 64;
                MUSED_SMEM_ATTR(k_bstage_tma, smem);
                const long long tiles = (m + RT - 1) / RT;
                const unsigned gx = (unsigned)std::min<long long>(tiles, std::max(1, 3 * E.sm_count / E.K));
                k_bstage_tma<<<dim3(gx, (unsigned)E.K), STMA_NT, smem, st>>>((const float*)dev_rows, stream_stride, D.xrows, D.d, m, RT,
                                                                         row_off, D.xhi, D.xlo, D.norm2);
            } else if (D.d_l % 4 == 0 && ld % 4 == 0 && stream_stride % 4 == 0 && ((uintptr_t)dev_rows & 15) == 0)
                k_bstage_v4<<<dim3((unsigned)std::min<long long>((m + 7) / 8, std::max(1ll, 16ll * E.sm_count / E.K)), (unsigned)E.K), 256, 0, st>>>(
                    (const float*)dev_rows, ld, stream_stride, D.xrows, D.d_l, D.d, m, row_off, D.xhi, D.xlo, D.norm2);
            else
                STAGE_(float);
            break;
        case MUSED_I64: STAGE_(long long); break;
        case MUSED_I32: STAGE_(int); break;
        case MUSED_U8: STAGE_(unsigned char); break;
    }
#undef STAGE_
    MUSED_CUDA(cudaGetLastError());
    E.mark(0, st);
    mused_count_launches(1);
    if (logical) E.pend += (int)m;
    else E.ahead += m;
    return 0;
}

// the open chunk back to row 0 of the staging planes (rare: only when a fit() call finds the planes
// nearly full behind a partial chunk)
static int compact_open_chunk(BEngine& E, cudaStream_t st) {
    BDev& D = E.D;
    if (D.x0 == 0) return 0;
    if (E.pend > 0) {
        // cudaMemcpyAsync needs disjoint ranges: move in pieces of at most x0 rows, front to back (dst < src, so a
        // piece never overwrites rows that are still to be moved)
        for (int k = 0; k < E.K; ++k) {
            const size_t base = (size_t)k * D.xrows;
            for (int done = 0; done < E.pend; done += D.x0) {
                const int r = std::min(D.x0, E.pend - done);
                const size_t bytes = (size_t)r * D.d * sizeof(float);
                MUSED_CUDA(cudaMemcpyAsync(D.xhi + (base + done) * D.d, D.xhi + (base + D.x0 + done) * D.d, bytes, cudaMemcpyDeviceToDevice, st));
                MUSED_CUDA(cudaMemcpyAsync(D.xlo + (base + done) * D.d, D.xlo + (base + D.x0 + done) * D.d, bytes, cudaMemcpyDeviceToDevice, st));
                MUSED_CUDA(cudaMemcpyAsync(D.norm2 + base + done, D.norm2 + base + D.x0 + done, (size_t)r * sizeof(double), cudaMemcpyDeviceToDevice, st));
            }
        }
    }
    D.x0 = 0;
    return 0;
}

int bengine_fit(BEngine* Ep, const void* rows, long long n_rows, long long ld, long long stream_stride, int dtype,
                int where, cudaStream_t st) {
    BEngine& E = *Ep;
    const size_t es = b_dtype_size(dtype);
    MUSED_REQUIRE(es != 0, "unknown dtype %d", dtype);
    MUSED_REQUIRE(n_rows >= 0 && ld >= E.D.d_l, "bad shape: n_rows=%lld ld=%lld d=%lld", n_rows, ld, E.D.d_l);
    MUSED_REQUIRE(rows != nullptr || n_rows == 0, "rows is NULL");
    MUSED_REQUIRE(E.K == 1 || stream_stride >= n_rows * ld, "stream_stride %lld overlaps the streams (n_rows %lld x ld %lld)",
                  stream_stride, n_rows, ld);
    MUSED_CUDA(cudaSetDevice(E.device));
    const long long d = E.D.d_l;
    const long long host_rows = E.raw_rows;
    long long fed = 0;
    while (fed < n_rows) {
        // a block of caller rows (of every stream) that is resident on the device
        long long blk = n_rows - fed;
        const char* src = (const char*)rows + (size_t)fed * ld * es;
        const void* dev = src;
        long long dev_ld = ld, dev_ss = stream_stride;
        if (where == MUSED_HOST) {
            blk = std::min(blk, host_rows);
            for (int k = 0; k < E.K; ++k)
                MUSED_CUDA(cudaMemcpy2DAsync(E.raw + (size_t)k * host_rows * d * es, (size_t)d * es,
                                             src + (size_t)k * stream_stride * es, (size_t)ld * es, (size_t)d * es,
                                             (size_t)blk, cudaMemcpyHostToDevice, st));
            dev = E.raw;
            dev_ld = d;
            dev_ss = host_rows * d;
        }
        long long off = 0;
        while (off < blk) {
            // stage as many of the block's rows as the planes take, in one launch ...
            long long space = E.sb_rows - (E.D.x0 + E.pend);
            if (space < std::min<long long>(blk - off, E.batch)) {
                int rc = compact_open_chunk(E, st);
                if (rc) return rc;
                space = E.sb_rows - E.pend;
            }
            const long long m = std::min(blk - off, space);
            int rc = stage_rows(E, (const char*)dev + (size_t)off * dev_ld * es, m, dev_ld, dev_ss, dtype, false, st);
            if (rc) return rc;
            off += m;
            // ... then fold them chunk by chunk
            while (E.ahead > 0) {
                const long long take = std::min(E.ahead, chunk_room(E));
                E.pend += (int)take;
                E.ahead -= take;
                E.time += (unsigned long long)take;
                if (chunk_room(E) == 0) {
                    rc = process_chunk(E, st);
                    if (rc) return rc;
                }
            }
        }
        fed += blk;
        if (where == MUSED_HOST) MUSED_CUDA(cudaStreamSynchronize(st));   // E.raw is reused / caller buffer is free
    }
    E.last_stream = st;
    return 0;
}

int bengine_streams(BEngine* E) { return E->K; }
long long state_blob_bytes(const std::vector<StateAlloc>& a) {
    long long n = (long long)sizeof(StateHeader);
    for (const StateAlloc& x : a)
        if (x.state) n += 8 + (long long)x.bytes;
    return n;
}

int state_blob_save(unsigned kind, const unsigned long long (&host)[6], const std::vector<StateAlloc>& a, void* buf, long long cap) {
    MUSED_REQUIRE(buf && cap >= state_blob_bytes(a), "state buffer of %lld bytes is too small (need %lld)", cap, state_blob_bytes(a));
    MUSED_CUDA(cudaDeviceSynchronize());
    unsigned char* o = (unsigned char*)buf;
    StateHeader h;
    h.magic = STATE_MAGIC;
    h.kind = kind;
    h.n_allocs = 0;
    for (const StateAlloc& x : a) h.n_allocs += x.state ? 1u : 0u;
    memcpy(h.host, host, sizeof(h.host));
    memcpy(o, &h, sizeof(h));
    o += sizeof(h);
    for (const StateAlloc& x : a) {
        if (!x.state) continue;
        const unsigned long long b = x.bytes;
        memcpy(o, &b, 8);
        o += 8;
        MUSED_CUDA(cudaMemcpy(o, x.p, x.bytes, cudaMemcpyDeviceToHost));
        o += x.bytes;
    }
    return 0;
}

int state_blob_load(unsigned kind, unsigned long long (&host)[6], const std::vector<StateAlloc>& a, const void* buf, long long bytes) {
    MUSED_REQUIRE(buf && bytes >= (long long)sizeof(StateHeader), "state blob too short");
    const unsigned char* o = (const unsigned char*)buf;
    StateHeader h;
    memcpy(&h, o, sizeof(h));
    MUSED_REQUIRE(h.magic == STATE_MAGIC, "not a mused_b200 state blob");
    MUSED_REQUIRE(h.kind == kind, "state blob of the other schedule (saved %u, this handle %u)", h.kind, kind);
    MUSED_REQUIRE(bytes == state_blob_bytes(a), "state blob of %lld bytes does not fit this handle's configuration (%lld bytes)",
                  bytes, state_blob_bytes(a));
    // first pass: every size must match before anything is overwritten
    const unsigned char* q = o + sizeof(h);
    unsigned n = 0;
    for (const StateAlloc& x : a) {
        if (!x.state) continue;
        unsigned long long b;
        memcpy(&b, q, 8);
        MUSED_REQUIRE(b == x.bytes, "state blob buffer %u has %llu bytes, this handle's has %zu: another configuration", n, b, x.bytes);
        q += 8 + x.bytes;
        ++n;
    }
    MUSED_REQUIRE(n == h.n_allocs, "state blob holds %u buffers, this handle %u", h.n_allocs, n);
    MUSED_CUDA(cudaDeviceSynchronize());
    q = o + sizeof(h);
    for (const StateAlloc& x : a) {
        if (!x.state) continue;
        q += 8;
        MUSED_CUDA(cudaMemcpy(x.p, q, x.bytes, cudaMemcpyHostToDevice));
        q += x.bytes;
    }
    memcpy(host, h.host, sizeof(h.host));
    return 0;
}

// checkpoint of the batched engine: the open chunk (staged, not yet folded rows) is part of the state
int bengine_state(BEngine* E, int op, void* buf, long long cap, long long* bytes) {
    MUSED_CUDA(cudaSetDevice(E->device));
    if (bytes) *bytes = state_blob_bytes(E->allocs);
    if (op == 0) return 0;
    unsigned long long host[6] = {E->time, E->done, (unsigned long long)E->pend, (unsigned long long)E->ahead,
                                  (unsigned long long)E->D.x0, 0ull};
    if (op == 1) return state_blob_save(1u, host, E->allocs, buf, cap);
    MUSED_REQUIRE(!E->comm, "load_state on a handle attached to a communicator: load before attaching");
    if (int rc = state_blob_load(1u, host, E->allocs, buf, cap)) return rc;
    E->time = host[0];
    E->done = host[1];
    E->pend = (int)host[2];
    E->ahead = (long long)host[3];
    E->D.x0 = (int)host[4];
    E->q_ng = 0;
    return 0;
}

unsigned long long bengine_time(BEngine* E) { return E->time; }
int bengine_qcap(BEngine* E) { return E->D.capn; }

static int b_read_slots(BEngine& E, std::vector<BSlot>& sl) {
    sl.resize(std::max(E.D.S, 1));
    MUSED_CUDA(cudaSetDevice(E.device));
    MUSED_CUDA(cudaStreamSynchronize(E.last_stream));      // the kernels of the last fit() may still be running
    MUSED_CUDA(cudaMemcpy(sl.data(), E.D.slots, sizeof(BSlot) * sl.size(), cudaMemcpyDeviceToHost));
    return 0;
}
static int b_fd_slot(const BEngine& E, int level, int stream_k = 0) {
    // the chunk that contains row `done` decided the period: period = t_a / N of the last chunk;
    // rows are only folded in whole chunks, so done / N is that period once a boundary row was folded
    const unsigned long long p = E.done / E.D.N;
    return stream_k * E.D.S1 + 2 * level + (int)(p & 1ull);
}

int bengine_size(BEngine* E, int stream_k, long long* n) {
    MUSED_REQUIRE(stream_k >= 0 && stream_k < E->K, "stream %d out of range (n_streams %d)", stream_k, E->K);
    std::vector<BSlot> sl;
    int rc = b_read_slots(*E, sl);
    if (rc) return rc;
    long long tot = 0;
    const int o = stream_k * E->D.S1;
    for (int j = 0; j < E->n_levels; ++j) tot += sl[o + 2 * j].q_len + sl[o + 2 * j + 1].q_len + 2ll * E->D.ell_l;
    *n = tot;
    return 0;
}

static int b_copy_out(double* dst, const double* src, size_t n, int where, cudaStream_t st) {
    if (!dst) return 0;
    MUSED_CUDA(cudaMemcpyAsync(dst, src, n * sizeof(double),
                               where == MUSED_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice, st));
    return 0;
}
// rows x cols (the caller's sketch_dim x d, packed) out of a matrix with the physical row length ld
static int b_copy_out_2d(double* dst, const double* src, size_t rows, size_t cols, size_t ld, int where, cudaStream_t st) {
    if (!dst) return 0;
    if (cols == ld) return b_copy_out(dst, src, rows * cols, where, st);
    MUSED_CUDA(cudaMemcpy2DAsync(dst, cols * sizeof(double), src, ld * sizeof(double), cols * sizeof(double), rows,
                                 where == MUSED_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice, st));
    return 0;
}

// SeqBasedSWFD.get, first half: level choice (seq_based_swfd.py:83-98), stack
// [C; snapshots] of the chosen level and its float64 Gram.  *ng = rows of the
// stack, 0 when the level has no snapshots (no Gram needed).
static int b_get_begin(BEngine* Ep, int stream_k, int* level, int* ng_out, bool with_gram, cudaStream_t st) {
    BEngine& E = *Ep;
    BDev& D = E.D;
    if (E.n_levels == 0) {
        set_error("SeqBasedSWFD has zero levels (R <= 1): get() is undefined");
        return MUSED_ERR_STATE;
    }
    MUSED_REQUIRE(E.pend == 0, "get_begin: the open chunk must be closed first");
    MUSED_REQUIRE(stream_k >= 0 && stream_k < E.K, "stream %d out of range (n_streams %d)", stream_k, E.K);
    MUSED_CUDA(cudaSetDevice(E.device));
    MUSED_CUDA(cudaStreamSynchronize(st));
    std::vector<BSlot> sl;
    int rc = 0;
    if ((rc = b_read_slots(E, sl))) return rc;
    int rj = 0;
    const unsigned long long time = E.time, N = D.N;
    for (int j = 0; j < E.n_levels; ++j) {
        const int s = b_fd_slot(E, j, stream_k);
        if (sl[s].q_len == 0) break;
        unsigned long long head_s = 0;
        MUSED_CUDA(cudaMemcpy(&head_s, D.ringS + (size_t)s * D.capn + sl[s].q_head, 8, cudaMemcpyDeviceToHost));
        if (time - head_s >= std::min(N - 1ull, time - 1ull)) { rj = j; break; }
    }
    if (level) *level = rj;
    const int s = b_fd_slot(E, rj, stream_k);
    E.q_slot = s;
    E.q_ng = 0;
    if (sl[s].q_len != 0) {
        const int ng = D.ell + sl[s].q_len;
        MUSED_REQUIRE(ng + (ng & 1) <= E.Q.ng_max, "internal: stack of %d rows exceeds scratch %d", ng, E.Q.ng_max);
        k_bgather<<<std::min(ng, 4 * E.sm_count), 256, 0, st>>>(D, s, E.Q.M);
        MUSED_CUDA(cudaGetLastError());
        mused_count_launches(1);
        if (with_gram && (rc = merge_query_gram(E.Q, ng, D.d, E.sm_count, st))) return rc;
        E.q_gram = with_gram;
        E.q_ng = ng;
    }
    if (ng_out) *ng_out = E.q_ng;
    return 0;
}

int bengine_get_begin(BEngine* Ep, int* level, int* ng_out, cudaStream_t st) {
    return b_get_begin(Ep, 0, level, ng_out, true, st);
}

int bengine_get_finish(BEngine* Ep, double* sketch, double* sigma2, double* vt, double* delta, int where,
                       cudaStream_t st) {
    BEngine& E = *Ep;
    BDev& D = E.D;
    int rc = 0;
    const int s = E.q_slot;
    if (E.q_ng == 0) {
        k_bfd_out<<<E.sm_count, 256, 0, st>>>(D, s, E.o_sketch, E.o_s2, vt ? E.o_vt : nullptr, E.o_delta);
        MUSED_CUDA(cudaGetLastError());
        mused_count_launches(1);
    } else {
        // single-GPU get(): Gram or covariance side, whichever is smaller; phase API: the Gram is in Q.G
        if (E.q_gram) rc = merge_query_solve(E.Q, E.q_ng, D.d, D.ell_l, E.sm_count, E.o_sketch, vt ? E.o_vt : nullptr, E.o_s2, st);
        else rc = merge_query(E.Q, E.q_ng, D.d, D.ell_l, E.sm_count, E.o_sketch, vt ? E.o_vt : nullptr, E.o_s2, st);
        if (rc) return rc;
        MUSED_CUDA(cudaMemsetAsync(E.o_delta, 0, sizeof(double), st));
    }
    if ((rc = b_copy_out_2d(sketch, E.o_sketch, (size_t)D.ell_l, (size_t)D.d_l, (size_t)D.d, where, st))) return rc;
    if ((rc = b_copy_out(sigma2, E.o_s2, (size_t)D.ell_l, where, st))) return rc;
    if ((rc = b_copy_out_2d(vt, E.o_vt, (size_t)D.ell_l, (size_t)D.d_l, (size_t)D.d, where, st))) return rc;
    if (delta) MUSED_CUDA(cudaMemcpyAsync(delta, E.o_delta, sizeof(double), cudaMemcpyDeviceToHost, st));
    MUSED_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int bengine_get(BEngine* Ep, int stream_k, double* sketch, double* sigma2, double* vt, double* delta, int* level,
                int where, cudaStream_t st) {
    BEngine& E = *Ep;
    if (E.n_levels == 0) {
        set_error("SeqBasedSWFD has zero levels (R <= 1): get() is undefined");
        return MUSED_ERR_STATE;
    }
    MUSED_CUDA(cudaSetDevice(E.device));
    int rc = process_chunk(E, st);      // get() closes the open chunk (the reference flushes on get, too)
    if (rc) return rc;
    int ng = 0;
    if (E.comm) {
        // sharded query: float64 Gram of the local columns of the stack, summed over the ranks (only the live
        // ng x ng block travels: the scratch is addressed with leading dimension ng for this query)
        const int lda_full = E.Q.lda;
        if ((rc = b_get_begin(Ep, stream_k, level, &ng, false, st))) return rc;
        if (ng > 0) {
            E.Q.lda = ng + (ng & 1);
            rc = merge_query_gram(E.Q, ng, E.D.d, E.sm_count, st);
            if (!rc) rc = allreduce_f64(E, E.Q.G, (size_t)E.Q.lda * E.Q.lda, st);
            E.q_gram = true;
        }
        if (!rc) rc = bengine_get_finish(Ep, sketch, sigma2, vt, delta, where, st);
        E.Q.lda = lda_full;
        return rc;
    }
    if ((rc = b_get_begin(Ep, stream_k, level, &ng, false, st))) return rc;
    return bengine_get_finish(Ep, sketch, sigma2, vt, delta, where, st);
}

int bengine_attach_comm(BEngine* Ep, const void* unique_id, long long id_bytes, int rank, int nranks, const char* nccl_path) {
    BEngine& E = *Ep;
    MUSED_REQUIRE(E.K == 1, "the column-sharded mode drives one stream per handle");
    MUSED_REQUIRE(!E.comm, "a communicator is already attached");
    MUSED_REQUIRE(unique_id && id_bytes == (long long)sizeof(ncclUniqueId), "unique id must be %zu bytes", sizeof(ncclUniqueId));
    MUSED_REQUIRE(nranks >= 1 && rank >= 0 && rank < nranks, "bad rank %d of %d", rank, nranks);
    MUSED_REQUIRE(E.time == 0, "attach the communicator before the first row");
    std::string err;
    if (!nccl_load(nccl_path, &err)) { set_error("%s", err.c_str()); return MUSED_ERR_STATE; }
    MUSED_CUDA(cudaSetDevice(E.device));
    ncclUniqueId id;
    memcpy(&id, unique_id, sizeof(id));
    MUSED_NCCL(nccl_api().CommInitRank(&E.comm, nranks, id, rank));
    E.comm_rank = rank;
    E.comm_size = nranks;
    return 0;
}

int bengine_comm_stats(BEngine* E, unsigned long long* n, unsigned long long* bytes) {
    if (n) *n = E->n_allreduce;
    if (bytes) *bytes = E->allreduce_bytes;
    return 0;
}

int nccl_unique_id(void* out, long long cap, long long* n_bytes, const char* nccl_path) {
    MUSED_REQUIRE(out && cap >= (long long)sizeof(ncclUniqueId), "unique id buffer must hold %zu bytes", sizeof(ncclUniqueId));
    std::string err;
    if (!nccl_load(nccl_path, &err)) { set_error("%s", err.c_str()); return MUSED_ERR_STATE; }
    ncclUniqueId id;
    MUSED_NCCL(nccl_api().GetUniqueId(&id));
    memcpy(out, &id, sizeof(id));
    if (n_bytes) *n_bytes = (long long)sizeof(id);
    return 0;
}

long long bengine_room(BEngine* E) { return chunk_room(*E); }
int bengine_pending(BEngine* E) { return E->pend; }
int bengine_ell(BEngine* E) { return E->D.ell; }

int bengine_buffers(BEngine* E, double** norm2, double** G, long long* g_count, double** QG, long long* qg_count) {
    if (norm2) *norm2 = E->D.norm2;
    if (G) *G = E->D.G;
    if (g_count) *g_count = (long long)E->D.S * E->D.n2 * E->D.n2;
    if (QG) *QG = E->Q.G;
    if (qg_count) *qg_count = (long long)E->Q.lda * E->Q.lda;
    return 0;
}

// K1 only: stage m (<= room) rows into the open chunk
int bengine_stage(BEngine* Ep, const void* rows, long long m, long long ld, int dtype, int where, cudaStream_t st) {
    BEngine& E = *Ep;
    const size_t es = b_dtype_size(dtype);
    MUSED_REQUIRE(es != 0, "unknown dtype %d", dtype);
    MUSED_REQUIRE(m >= 0 && m <= chunk_room(E), "stage: %lld rows do not fit the open chunk (room %lld)", m, chunk_room(E));
    MUSED_REQUIRE(ld >= E.D.d_l && (rows || m == 0), "bad rows / ld");
    if (m == 0) return 0;
    MUSED_CUDA(cudaSetDevice(E.device));
    const void* dev = rows;
    long long dev_ld = ld;
    if (where == MUSED_HOST) {
        MUSED_REQUIRE((size_t)m * E.D.d_l * es <= E.raw_bytes, "stage: block too large");
        MUSED_CUDA(cudaMemcpy2DAsync(E.raw, (size_t)E.D.d_l * es, rows, (size_t)ld * es, (size_t)E.D.d_l * es, (size_t)m, cudaMemcpyHostToDevice, st));
        dev = E.raw;
        dev_ld = E.D.d_l;
    }
    MUSED_REQUIRE(E.K == 1, "the phase API drives one stream");
    int rc = stage_rows(E, dev, m, dev_ld, 0, dtype, true, st);
    if (rc) return rc;
    E.time += (unsigned long long)m;
    if (where == MUSED_HOST) MUSED_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int bengine_debug(BEngine* Ep, int stream_k, double* scal, double* sigma2, unsigned long long* q_s,
                  unsigned long long* q_t, long long q_cap) {
    BEngine& E = *Ep;
    MUSED_REQUIRE(stream_k >= 0 && stream_k < E.K, "stream %d out of range (n_streams %d)", stream_k, E.K);
    std::vector<BSlot> sl;
    int rc = b_read_slots(E, sl);
    if (rc) return rc;
    const int capn = E.D.capn, ell = E.D.ell_l, ell_p = E.D.ell;
    std::vector<unsigned long long> rs(capn), rt(capn);
    for (int j = 0; j < E.n_levels; ++j) {
        for (int side = 0; side < 2; ++side) {
            const int fd = b_fd_slot(E, j, stream_k);
            const int s = side == 0 ? fd : (fd ^ 1);
            const int o = 2 * j + side;
            if (scal) {
                double* r = scal + (size_t)o * 8;
                r[0] = (double)sl[s].sweeps; r[1] = sl[s].delta; r[2] = 0.0;
                r[3] = (double)sl[s].last_dump; r[4] = sl[s].q_len;
                r[5] = (double)sl[s].n_flush; r[6] = (double)sl[s].n_dump; r[7] = (double)sl[s].n_raw;
            }
            if (sigma2)
                MUSED_CUDA(cudaMemcpy(sigma2 + (size_t)o * ell, E.D.s2 + (size_t)s * ell_p, sizeof(double) * ell, cudaMemcpyDeviceToHost));
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

int bengine_timing(BEngine* Ep, int enable, double* ms, long long* chunks, long long* rows) {
    BEngine& E = *Ep;
    MUSED_CUDA(cudaSetDevice(E.device));
    {
        // idle time on the stream: end mark of one kernel class -> start mark of the next (nothing
        // is enqueued in between), chunk by chunk; rotate -> next bookkeeping counts as the
        // inter-chunk gap unless a get() sits in between (> 5 ms)
        const size_t nc = std::min(std::min(E.ev[1].size(), E.ev[2].size()), std::min(E.ev[3].size(), E.ev[4].size())) / 2;
        for (size_t c = 0; c < nc; ++c) {
            float t = 0.f;
            MUSED_CUDA(cudaEventSynchronize(E.ev[4][2 * c + 1]));
            for (int k = 1; k < 4; ++k) {
                MUSED_CUDA(cudaEventElapsedTime(&t, E.ev[k][2 * c + 1], E.ev[k + 1][2 * c]));
                E.gap_ms[0] += t;
            }
            if (c + 1 < nc) {
                MUSED_CUDA(cudaEventElapsedTime(&t, E.ev[4][2 * c + 1], E.ev[1][2 * c + 2]));
                if (t < 5.f) E.gap_ms[1] += t;
            }
        }
    }
    for (int k = 0; k < 5; ++k) {
        for (size_t i = 0; i + 1 < E.ev[k].size(); i += 2) {
            float t = 0.f;
            MUSED_CUDA(cudaEventSynchronize(E.ev[k][i + 1]));
            MUSED_CUDA(cudaEventElapsedTime(&t, E.ev[k][i], E.ev[k][i + 1]));
            E.timed_ms[k] += t;
        }
        for (cudaEvent_t e : E.ev[k]) cudaEventDestroy(e);
        E.ev[k].clear();
    }
    if (ms) for (int k = 0; k < 5; ++k) ms[k] = E.timed_ms[k];
    if (chunks) *chunks = E.timed_chunks;
    if (rows) *rows = E.timed_rows;
    if (enable >= 0) {
        E.timing = enable != 0;
        for (int k = 0; k < 5; ++k) E.timed_ms[k] = 0.0;
        E.gap_last[0] = E.gap_ms[0];
        E.gap_last[1] = E.gap_ms[1];
        E.gap_ms[0] = E.gap_ms[1] = 0.0;
        E.timed_chunks = 0;
        E.timed_rows = 0;
    }
    return 0;
}

int bengine_gaps(BEngine* E, double* gaps2) {
    gaps2[0] = E->gap_last[0];
    gaps2[1] = E->gap_last[1];
    return 0;
}

}  // namespace mused

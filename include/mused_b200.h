/*
 * mused_b200 -- C ABI of the B200-native SWFD sketch-and-assign hot path.
 *
 * The reference (kelaendi/mused) is pure Python and has no FFI of its own; the
 * drop-in boundary is the Python surface that main.py calls.  Each entry point
 * below names the reference interface it stands behind (paths relative to the
 * reference tree).  INTEGRATION.md shows the ctypes binding a maintainer adds.
 *
 * Conventions: extern "C", plain pointers and sizes, no torch / CUDA types in
 * signatures (`stream` is a cudaStream_t passed as void*, NULL = default
 * stream).  Every function returns 0 on success and a negative code on error;
 * mused_last_error() returns the message of the calling thread's last error.
 * A handle is bound to one device and is not thread-safe; distinct handles are
 * independent.  There is no CPU fallback: without a CUDA device every call
 * fails with MUSED_ERR_CUDA.
 */
#ifndef MUSED_B200_H
#define MUSED_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MUSED_OK 0
#define MUSED_ERR_ARG (-1)
#define MUSED_ERR_CUDA (-2)
#define MUSED_ERR_STATE (-3)
#define MUSED_ERR_NOMEM (-4)

/* element types accepted for incoming rows */
#define MUSED_F64 0
#define MUSED_F32 1
#define MUSED_I64 2
#define MUSED_I32 3
#define MUSED_U8 4

/* where a buffer lives */
#define MUSED_HOST 0
#define MUSED_DEVICE 1

const char* mused_last_error(void);
int mused_version(void);
/* kernels launched by this library in this process so far (bench.py's gpu_launches) */
uint64_t mused_launch_count(void);
void mused_count_launches(uint64_t n);
/* number of CUDA devices visible, or a negative error code */
int mused_device_count(void);

/* ---------------------------------------------------------------------------
 * Sequence-based multi-level sliding-window Frequent Directions.
 * Replaces swfd/seq_based_swfd.py:11-103 (SeqBasedSWFD) together with the
 * per-level machinery of swfd/sliding_windows_frequent_directions.py:27-197
 * (SlidingWindowFD) and swfd/frequent_directions.py:96-153
 * (FrequentDirectionsWithDump).  Arithmetic is float64 like the reference.
 * ------------------------------------------------------------------------- */
typedef struct mused_swfd mused_swfd_t;

typedef struct {
    int64_t N;                /* window length            (seq_based_swfd.py:22)        */
    int64_t d;                /* row dimension            (:23)                         */
    int64_t ell;              /* sketch_dim               (:27)                         */
    double beta;              /* queue cap (2+8/beta)*ell (:61)                         */
    int32_t n_levels;         /* logR                     (:25 / :32)                   */
    const double* thresholds; /* n_levels dump thresholds (:36 / :42), host pointer     */
    int32_t device;           /* CUDA ordinal                                            */
    int32_t cluster;          /* CTAs cooperating on one FD instance: 1,2,4,8; 0 = auto */
    int64_t chunk_rows;       /* rows staged per launch; 0 = auto                        */
    int32_t schedule;         /* MUSED_SCHEDULE_REFERENCE | MUSED_SCHEDULE_BATCHED       */
    int32_t batch_rows;       /* batched: rows per shrink (<= ell); 0 = ell              */
    int32_t n_streams;        /* batched: independent streams advanced in lock-step by one
                                 launch sequence (north_star: "independent streams and windows
                                 are batched per GPU"); 0 / 1 = one stream                  */
} mused_swfd_cfg;

/* REFERENCE: every flush / dump decision of the reference, row by row
 * (frequent_directions.py:122-125) -- bit-faithful schedule, float64.
 * BATCHED: north_star's schedule -- rows are folded in chunks of batch_rows with
 * one tensor-core Gram + eigensolve + rotate per chunk and FD instance, every
 * direction with sigma^2 >= threshold dumped per chunk (fastfdwithdump.py:115-131);
 * stated exactly in oracle/swfd_batched_oracle.py.  fp32 rows, 3xTF32 products,
 * float64 eigensolve and scalars. */
#define MUSED_SCHEDULE_REFERENCE 0
#define MUSED_SCHEDULE_BATCHED 1

int mused_swfd_create(mused_swfd_t** out, const mused_swfd_cfg* cfg);
int mused_swfd_destroy(mused_swfd_t* h);

/* SeqBasedSWFD.fit (seq_based_swfd.py:47-80) applied to n_rows consecutive rows
 * in arrival order; n_rows = 1 is exactly one reference call.  `rows` is
 * row-major with leading dimension `ld` elements of `dtype`, on the host or on
 * the handle's device.  The rows are copied; the caller may reuse the buffer
 * when the call returns (host) / when `stream` reaches this point (device). */
int mused_swfd_fit(mused_swfd_t* h, const void* rows, int64_t n_rows, int64_t ld,
                   int dtype, int where, void* stream);

/* n_streams independent sketches behind one handle (cfg.n_streams = K, schedule = BATCHED): every stream
 * follows SeqBasedSWFD semantics (seq_based_swfd.py:11-103) on its own rows and shares nothing with the
 * others; one K1..K5 launch sequence per chunk covers the K x 2 x levels FD instances.  fit_streams feeds
 * n_rows rows to EVERY stream: stream k's rows start at rows + k * stream_stride elements (row-major, ld).
 * get_stream / size_stream / debug_stream address stream k (get closes the open chunk of all streams,
 * as SeqBasedSWFD.get flushes, then reads stream k without further mutation). */
int mused_swfd_fit_streams(mused_swfd_t* h, const void* rows, int64_t n_rows, int64_t ld, int64_t stream_stride,
                           int dtype, int where, void* stream);
int mused_swfd_get_stream(mused_swfd_t* h, int32_t k, double* sketch, double* sigma2, double* vt, double* delta,
                          int32_t* level, int where, void* stream);
int mused_swfd_size_stream(mused_swfd_t* h, int32_t k, int64_t* n);
int mused_swfd_debug_stream(mused_swfd_t* h, int32_t k, double* scal, double* sigma2, uint64_t* q_s, uint64_t* q_t,
                            int64_t q_cap);

/* TimeBasedSWFD.fit(X, t) (swfd/seq_based_swfd.py:150-189): the same per-row rules with an
 * explicit arrival time per row instead of the running counter; times[r] (HOST array of n_rows
 * values, non-decreasing) may repeat or skip whole windows (then both sides of a level start over,
 * sliding_windows_frequent_directions.py:66-75).  schedule = reference only. */
int mused_swfd_fit_timed(mused_swfd_t* h, const void* rows, int64_t n_rows, int64_t ld,
                         int dtype, int where, const uint64_t* times, void* stream);

/* SeqBasedSWFD.get (seq_based_swfd.py:83-100 -> sliding_windows_...py:154-173).
 * Like the reference it flushes the chosen level's pending rows (a state
 * mutation).  sketch: ell x d, sigma2: ell, vt: ell x d (may be NULL),
 * delta: 1, level: chosen level j.  Output buffers live on `where`. */
int mused_swfd_get(mused_swfd_t* h, double* sketch, double* sigma2, double* vt,
                   double* delta, int32_t* level, int where, void* stream);

/* SeqBasedSWFD.get_size (seq_based_swfd.py:102-103). */
int mused_swfd_size(mused_swfd_t* h, int64_t* n);

/* Rows consumed so far (SeqBasedSWFD.time). */
int mused_swfd_time(mused_swfd_t* h, uint64_t* t);

/* Checkpoint / resume of a sketch ("state_dict"; the reference keeps its state in Python objects and has no
 * counterpart, SURVEY.md section 5 lists it as an optional next item): the handle's persistent device buffers and
 * host counters as one HOST blob of mused_swfd_state_size bytes.  Both schedules; with schedule = BATCHED the rows
 * of the open chunk are part of the state.  A blob loads only into a handle created with the same configuration
 * (sizes are checked before anything is overwritten: MUSED_ERR_ARG otherwise).  save / load synchronise the device. */
int mused_swfd_state_size(mused_swfd_t* h, int64_t* bytes);
int mused_swfd_state_save(mused_swfd_t* h, void* buf, int64_t cap);
int mused_swfd_state_load(mused_swfd_t* h, const void* buf, int64_t bytes);

/* Introspection for the differential tests: per FD instance (2 per level,
 * index 2*level + side, side 0 = the reference's current `fd`, 1 = `fd_aux`)
 * energy, delta, pending rows, last_dump_time, queue length, flush/dump counts.
 * scal: n_inst x 8 doubles {energy, delta, pending, last_dump_time, q_len,
 * n_flush, n_dump, n_raw}; sigma2: n_inst x ell; q_s/q_t: n_inst x q_cap
 * (first q_len valid, oldest first).  Host buffers; any may be NULL. */
int mused_swfd_debug(mused_swfd_t* h, double* scal, double* sigma2, uint64_t* q_s,
                     uint64_t* q_t, int64_t q_cap);
int mused_swfd_queue_capacity(mused_swfd_t* h, int64_t* q_cap);
/* CUDA-event timing of the per-row kernel (the dominant kernel): returns the
 * milliseconds, launches and rows accumulated since timing was last (re)armed,
 * then enable = 1 arms and resets, 0 disarms and resets, -1 leaves the state alone. */
int mused_swfd_timing(mused_swfd_t* h, int enable, double* ms, int64_t* launches, int64_t* rows);
/* batched schedule: CUDA-event milliseconds per kernel class since timing was armed,
 * ms5 = {K1 stage, K5 bookkeeping, K2 Gram, K3 eigensolve, K4 rotate}; chunks and rows
 * folded meanwhile.  enable as in mused_swfd_timing. */
int mused_swfd_kernel_ms(mused_swfd_t* h, int enable, double* ms5, int64_t* chunks, int64_t* rows);
/* Stream idle time (ms) of the period closed by the last mused_swfd_kernel_ms(enable >= 0):
 * gaps2[0] between the kernels of a chunk, gaps2[1] between consecutive chunks of one fit. */
int mused_swfd_kernel_gaps(mused_swfd_t* h, double* gaps2);
/* SM cycles the per-row kernel spent per phase, per physical FD instance:
 * cycles[inst * 16 + k], k = {A bookkeeping, B gemv, C Q^T g, D deflate, E secular roots,
 * F coupling vector, G eigenvector rows, H eigenvector GEMM, flush, dump, raw, rows} */
int mused_swfd_profile(mused_swfd_t* h, uint64_t* cycles, int64_t n_inst_cap);

/* ---------------------------------------------------------------------------
 * Column-sharded mode (schedule = BATCHED): rank r holds columns
 * [lo_r, hi_r) of every row and of the sketch (cfg.d = hi_r - lo_r).  The caller
 * (mused_b200/sharded.py, torch.distributed / NCCL) drives one chunk in phases and
 * runs the two exchanges of the path on the device buffers the library exposes:
 *   stage (rows of the open chunk, <= room)
 *   -> allreduce-sum norm2[ell]            (routing rule X@X.T >= threshold, seq_based_swfd.py:77)
 *   -> chunk_gram                          (bookkeeping, tcgen05 Gram of the local columns)
 *   -> allreduce-sum gram[S][2l][2l]       (the one real exchange per shrink)
 *   -> chunk_finish                        (replicated eigensolve, local rotate)
 * and a query as  get_begin -> allreduce-sum query_gram (when n_stack > 0) -> get_finish.
 * Every decision depends on replicated scalars only, so all ranks stay in lock-step.
 * ------------------------------------------------------------------------- */
/* The same mode with the exchanges INSIDE the library: rank 0 draws a NCCL unique id
 * (mused_comm_unique_id), the host distributes its bytes (torch.distributed broadcast, MPI, a file ...),
 * every rank attaches it to its handle before the first row; from then on mused_swfd_fit / mused_swfd_get on
 * that handle are collective calls (every rank feeds its own columns of the same rows) and post
 * ncclAllReduce(norm2) and ncclAllReduce(gram) per chunk, ncclAllReduce(query gram) per get(), on the
 * caller's stream with no host synchronisation.  nccl_path: the libnccl.so.2 to dlopen (NULL = the one
 * the process already maps, e.g. torch's).  comm_stats: allreduces posted and their payload bytes. */
int mused_comm_unique_id(void* id_out, int64_t cap, int64_t* n_bytes, const char* nccl_path);
int mused_swfd_attach_comm(mused_swfd_t* h, const void* unique_id, int64_t n_bytes, int32_t rank, int32_t nranks,
                           const char* nccl_path);
int mused_swfd_comm_stats(mused_swfd_t* h, uint64_t* n_allreduce, uint64_t* bytes);

int mused_swfd_chunk_room(mused_swfd_t* h, int64_t* room, int64_t* pending);
int mused_swfd_stage(mused_swfd_t* h, const void* rows, int64_t n_rows, int64_t ld, int dtype, int where,
                     void* stream);
int mused_swfd_chunk_gram(mused_swfd_t* h, void* stream);
int mused_swfd_chunk_finish(mused_swfd_t* h, void* stream);
int mused_swfd_exchange_buffers(mused_swfd_t* h, double** norm2, int64_t* norm2_count, double** gram,
                                int64_t* gram_count, double** query_gram, int64_t* query_gram_count);
int mused_swfd_get_begin(mused_swfd_t* h, int32_t* level, int32_t* n_stack, void* stream);
int mused_swfd_get_finish(mused_swfd_t* h, double* sketch, double* sigma2, double* vt, double* delta,
                          int where, void* stream);

/* ---------------------------------------------------------------------------
 * One Frequent Directions instance with row buffer and dump.
 * Replaces swfd/frequent_directions.py:96-153 (FrequentDirectionsWithDump) and,
 * with error = +inf and explicit batches, :79-93 (RobustFrequentDirections).
 * ------------------------------------------------------------------------- */
typedef struct mused_fd mused_fd_t;

int mused_fd_create(mused_fd_t** out, int64_t d, int64_t ell, double error, int device,
                    int cluster);
int mused_fd_destroy(mused_fd_t* h);
/* FrequentDirectionsWithDump.fit row by row (:115-125); no dump is taken. */
int mused_fd_fit(mused_fd_t* h, const void* rows, int64_t n_rows, int64_t ld, int dtype,
                 int where, void* stream);
/* FrequentDirections.fit(X, batch_size) (:39-58): shrink after every
 * batch_size rows (and after the ragged tail), no energy rule. */
int mused_fd_fit_batches(mused_fd_t* h, const void* rows, int64_t n_rows, int64_t ld,
                         int dtype, int where, int64_t batch_size, void* stream);
/* FrequentDirectionsWithDump.dump (:127-142): *dumped = 1 and v (d doubles on
 * `where`) filled when sigma2[0] >= error, else *dumped = 0. */
int mused_fd_dump(mused_fd_t* h, double* v, int32_t* dumped, int where, void* stream);
int mused_fd_flush(mused_fd_t* h, void* stream);
/* get (:144-146): flushes, then returns (sketch, sigma2, vt, delta). */
int mused_fd_get(mused_fd_t* h, double* sketch, double* sigma2, double* vt, double* delta,
                 int where, void* stream);

/* ---------------------------------------------------------------------------
 * Similarity / assignment operators of matrix_operations.py.
 * ------------------------------------------------------------------------- */
#define MUSED_METRIC_COSINE 0 /* matrix_operations.py:106 cosine_similarity          */
#define MUSED_METRIC_L2 1     /* :118 NearestNeighbors (Euclidean), :286-289 argmin  */
#define MUSED_METRIC_HAVERSINE 2 /* :22-30 location kNN, haversine_distance :250-263 (float64 entry) */
#define MUSED_METRIC_L1SUM 3  /* :32-53 time: |d datetaken| + |d dateupload| (float64 entry)       */

/* Fused cosine-similarity + threshold + argmax event assignment (north_star
 * operator; nearest reference code matrix_operations.py:106 and :286-289).
 * X: n x d float32 rows (ld_x), C: k x d float32 centroids (ld_c), all device.
 * metric COSINE: label = argmax_j cos(x, c_j) if that cosine >= tau else -1;
 * metric L2:     label = argmin_j ||x - c_j||  if that distance <= tau else -1
 *                (tau < 0 disables the threshold).  Ties -> lowest j.
 * score (nullable): best cosine / best distance.
 * With score == NULL and metric COSINE the call takes the bandwidth path: one TF32 screening pass on the tensor
 * cores, then float32 re-scoring of exactly the rows whose decision (winner or threshold) lies inside the
 * screening error bound (|error of a cosine| <= 1.1e-3); the labels are those of the float32-exact path. */
int mused_assign(const float* X, int64_t n, int64_t d, int64_t ld_x, const float* C,
                 int64_t k, int64_t ld_c, int metric, float tau, int32_t* labels,
                 float* score, void* stream);
/* Diagnostics of the call above: rows re-scored after screening, cumulative on the current device (synchronises). */
int mused_assign_pending(uint64_t* n_rows);

/* kNN adjacency of one dense modality, OR-accumulated into a fused uint8
 * W x W matrix (create_adjacency_matrix numeric branches
 * matrix_operations.py:101-108 cosine top-k, :112-119 Euclidean kNN, writer
 * :123-130, and fuse_matrices :134-141).  feats: W x f float32 device; every
 * row i sets adj[i, j] = 1 for its k best j (self excluded after selection
 * exactly like the reference: the top-k list may contain i itself). */
int mused_topk_adjacency(const float* feats, int64_t W, int64_t f, int64_t ld, int metric,
                         int64_t k, uint8_t* adj, void* stream);

/* mused_topk_adjacency for callers that may not allocate (CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same
 * arguments, same three kernels and the same result as mused_topk_adjacency; the W x W key matrix and the row
 * norms live in a file-scope `static __device__` array of the library (4 MiB + 4 KiB, held by every GPU that loads
 * the library until the process exits) instead of stream-ordered pool memory, and the call makes no cudaMalloc /
 * cudaMallocAsync / cudaFree (the first call per device resolves the array's device address with
 * cudaGetSymbolAddress).  Limit: W <= MUSED_TOPK_STATIC_MAX_W rows; a larger window is MUSED_ERR_ARG with a message
 * naming the limit.  All calls on a device share the one array: calls that overlap on the GPU (different streams
 * without an ordering between them) race -- serialise them.  Nothing in the array has to start at zero (both
 * regions are rewritten for [0, W) before they are read). */
#define MUSED_TOPK_STATIC_MAX_W 1024
int mused_topk_adjacency_static(const float* feats, int64_t W, int64_t f, int64_t ld, int metric,
                                int64_t k, uint8_t* adj, void* stream);

/* Same operator for the two low-dimensional float64 modalities of
 * create_adjacency_matrix: metric HAVERSINE ("location", matrix_operations.py:22-30,
 * rows = (latitude, longitude) in degrees, k = k_basis + 1 there) and L1SUM
 * ("time", :32-53, rows = (datetaken, dateupload), k = 3 k_basis + 1 there).
 * Epoch seconds and degrees need float64: feats is W x f float64 device.
 * The caller drops invalid rows first (NaN / zero stamps, :24 and :35). */
int mused_topk_adjacency_f64(const double* feats, int64_t W, int64_t f, int64_t ld, int metric,
                             int64_t k, uint8_t* adj, void* stream);

/* The two string modalities of create_adjacency_matrix on integer ids (the strings themselves are mapped to ids by
 * the host wrapper -- dictionaries of user names / tag tokens are host string processing):
 *  - "username" (matrix_operations.py:55-72): group[i] = id of row i's user name, < 0 for an empty name; every
 *    pair of distinct rows with the same id gets an edge in both directions.
 *  - "tags" (:74-89 with jaccard_similarity :245-248): row i's tag set is the sorted unique ids
 *    ids[offsets[i] .. offsets[i+1]); similarity = |A & B| / |A u B| (float64 quotient, 0 if either set is empty,
 *    -1 on the diagonal), every row gets edges to its k most similar rows (ties -> lowest index; the reference's
 *    np.argsort leaves the order inside ties to numpy's introsort).
 * Device pointers; adj is the W x W uint8 accumulator (edges are OR-ed in). */
int mused_group_adjacency(const int32_t* group, int64_t W, uint8_t* adj, void* stream);
int mused_jaccard_topk_adjacency(const int32_t* offsets, const int32_t* ids, int64_t W, int64_t k, uint8_t* adj,
                                 void* stream);

/* Contingency table for match_clusters (matrix_operations.py:169-172):
 * counts[p * Q + q] = #{i : prev[i] == p and cur[i] == q}; labels outside
 * [0,P) x [0,Q) are ignored.  Device pointers. */
int mused_contingency(const int32_t* prev, const int32_t* cur, int64_t n, int32_t P,
                      int32_t Q, int32_t* counts, void* stream);

/* FD merge of an arbitrary stack of rows: thin SVD, top-l cut, shrink by the (l+1)-th squared singular value when
 * the stack has more than l rows -- LMFD.__merge (swfd/lmfd.py:27-47), DI-FD's level merge (swfd/difd.py:117-127),
 * FastFrequentDirectionsWithDump's full shrink (swfd/fastfdwithdump.py:98-108) and SlidingWindowFD.get
 * (swfd/sliding_windows_frequent_directions.py:160-171).  rows: ng x d float64 DEVICE (ld); sketch l x d, sigma2 l,
 * vt l x d (nullable): DEVICE.  Synchronises the stream. */
int mused_fd_merge(const double* rows, int64_t ng, int64_t d, int64_t ld, int64_t ell, double* sketch, double* sigma2,
                   double* vt, void* stream);

/* KMeans of the reduced window, matrix_operations.py:149-153 (KMeans(n_clusters, random_state=seed).fit_predict)
 * entirely on the device: k-means++ seeding (D^2 sampling, counter-based RNG from `seed`), Lloyd iterations with
 * the assignment operator above and a deterministic centre update, until no label changes, the summed squared
 * centre movement is <= tol, or max_iter.  X: n x d float32 device rows; centroids: k x d float32 device (out),
 * labels: n int32 device (out), inertia / n_iter: host (nullable).  The seeding is not sklearn's (its greedy
 * k-means++ draws from numpy's Mersenne twister): labels agree with sklearn's up to the usual KMeans ambiguity,
 * which is why the reference-parity loop keeps sklearn's fit and this entry is the throughput option. */
int mused_kmeans(const float* X, int64_t n, int64_t d, int64_t ld, int32_t k, uint64_t seed, int32_t max_iter, double tol,
                 float* centroids, int32_t* labels, double* inertia, int32_t* n_iter, void* stream);

/* perform_svd_reduction (matrix_operations.py:143-147: TruncatedSVD(n_components, random_state=seed).fit_transform),
 * the reducer of the reference's sSVDMC approaches (main.py:79, :146), as sklearn's randomized SVD on the device:
 * n_iter power iterations from the Gaussian test matrix omega (n x q float64 DEVICE, row-major, q = n_components +
 * n_oversamples; the caller draws it with numpy's RandomState(seed).normal so that it is sklearn's), re-conditioned
 * by Cholesky-QR, B = Q^T M, V^T = right singular vectors of B with svd_flip's sign rule, Xt = M V (m x k, ld_xt).
 * M: m x n (ld), dtype MUSED_U8 (the fused 0/1 adjacency) or MUSED_F64, DEVICE, m >= n.  sigma (k, nullable) and
 * vt (k x n, ld_vt, nullable): DEVICE.  1 <= k <= q <= min(n, 160).  Synchronises the stream. */
int mused_svd_reduce(const void* M, int dtype, int64_t m, int64_t n, int64_t ld, const double* omega, int32_t q,
                     int32_t k, int32_t n_iter, double* Xt, int64_t ld_xt, double* sigma, double* vt, int64_t ld_vt,
                     void* stream);

/* Symmetric eigen-decomposition of a positive semi-definite Gram matrix
 * (what replaces scipy.linalg.svd(..., lapack_driver="gesvd") of the stacked
 * rows in SlidingWindowFD.get, sliding_windows_frequent_directions.py:160, and
 * the query of SeqBasedSWFD.get, seq_based_swfd.py:83-100: singular values =
 * sqrt(evals), left singular vectors = vecs).  G: n x n float64 device, leading
 * dimension ld; evals (n, descending) and vecs (n x n, ROW i = eigenvector i,
 * leading dimension ld_vecs) are device buffers.  Whole-GPU one-sided block
 * Jacobi (csrc/eig_big.cuh); n up to ~1750 on a 148-SM B200.  Eigenvectors of
 * eigenvalues at round-off level (<= 1e-10 of the largest) are returned as
 * zero rows: their direction is numerical noise.  Synchronises
 * the stream.  *sweeps (nullable) = Jacobi sweeps taken. */
int mused_sym_eig(const double* G, int64_t n, int64_t ld, double* evals, double* vecs,
                  int64_t ld_vecs, int32_t* sweeps, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MUSED_B200_H */

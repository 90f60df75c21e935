// Batched-schedule SWFD engine: device state and the host entry points the C ABI
// (swfd_engine.cu) forwards to.
#pragma once
#include <vector>
#include "common.cuh"
#include "eig_common.cuh"

namespace mused {

constexpr int BS_MAX_ELL = 256;

struct BDev {
    int S, ell, n2, capn, KS;  // S = FD instances of ALL streams; ell = PHYSICAL sketch rows (16 / 32 / 64 / 128), n2 = 2 ell
    int S1;                    // FD instances per stream (2 per level); instance s belongs to stream s / S1
    long long xrows;           // rows of one stream's staging planes (stream k starts at row k * xrows)
    int ell_l;                 // the caller's sketch_dim <= ell; rows ell_l .. ell-1 of every sketch stay zero
    int x0;                    // row of the staging planes where the open chunk starts
    long long d;               // PHYSICAL row length (multiple of 32); columns d_l .. d-1 stay zero
    long long d_l;             // the caller's d
    unsigned long long N;
    double qcap;
    BSlot* slots;
    float *skhi, *sklo;        // [S][2][ell][d] sketch planes (TF32 hi / lo), ping-pong
    float *xhi, *xlo;          // [K][xrows = stage rows + ell][d] staged rows of a fit() call; the open chunk starts at row x0
    double* norm2;             // [K][xrows] ||x||^2 of the staged rows
    unsigned char* mask;       // [S][ell] 1 = chunk row is a raw snapshot at this instance's level
    float* Gp;                 // [S][KS][n2][n2] split-K partials of K2
    double* G;                 // [S][n2][n2] their float64 sum (the buffer a column-sharded run allreduces)
    double* Gs;                // [S][n2][n2] K3 scratch: symmetrised, masked Gram (mixed-precision eigensolve)
    float *Thi, *Tlo;          // [S][ell][n2] rotate coefficients (TF32 hi / lo planes)
    double* s2;                // [S][ell]
    int* dump_slot;            // [S][ell]
    int* raw_list;             // [S][2][ell] chunk rows that became raw snapshots | their ring entries (K5 -> K4)
    float* ringV;              // [S][capn][d] snapshot rows
    unsigned long long *ringS, *ringT;
};

struct BEngine;

int bengine_create(BEngine** out, long long N, long long d, long long ell, double beta, int n_levels,
                   const double* thr, int device, long long batch, int n_streams);
void bengine_destroy(BEngine* E);
// n_rows rows for EVERY stream: stream k's rows start at rows + k * stream_stride elements
int bengine_fit(BEngine* E, const void* rows, long long n_rows, long long ld, long long stream_stride, int dtype,
                int where, cudaStream_t st);
int bengine_get(BEngine* E, int stream_k, double* sketch, double* sigma2, double* vt, double* delta, int* level,
                int where, cudaStream_t st);
int bengine_size(BEngine* E, int stream_k, long long* n);
int bengine_streams(BEngine* E);
unsigned long long bengine_time(BEngine* E);
int bengine_qcap(BEngine* E);
int bengine_debug(BEngine* E, int stream_k, double* scal, double* sigma2, unsigned long long* q_s,
                  unsigned long long* q_t, long long q_cap);
// CUDA-event time per kernel class since the last reset: ms[5] = {stage, book, gram, eig, rotate},
// chunks processed.  enable: 1 arm+reset, 0 disarm+reset, -1 read only.
int bengine_timing(BEngine* E, int enable, double* ms, long long* chunks, long long* rows);
// GPU idle time of the period closed by the last bengine_timing(enable >= 0): ms between the kernels
// of a chunk, ms between consecutive chunks of one fit()
int bengine_gaps(BEngine* E, double* gaps2);

// Phase API for the column-sharded mode (mused_b200/sharded.py drives it; the two
// exchange points of SURVEY 8(e) sit between the phases):
//   stage rows (<= room) -> [allreduce norm2] -> chunk_gram -> [allreduce G] -> chunk_finish
//   get_begin -> [allreduce query Gram] -> get_finish
long long bengine_room(BEngine* E);                       // rows the open chunk still takes
int bengine_pending(BEngine* E);
int bengine_ell(BEngine* E);                          // rows staged in the open chunk
int bengine_stage(BEngine* E, const void* rows, long long m, long long ld, int dtype, int where, cudaStream_t st);
int bengine_chunk_gram(BEngine* E, cudaStream_t st);     // K5 + K2 + split-K reduction into G (f64)
int bengine_chunk_finish(BEngine* E, cudaStream_t st);   // K3 + K4
// device buffers the collectives run on: norm2 [ell] f64, G [S][n2][n2] f64, query Gram [lda][lda] f64
int bengine_buffers(BEngine* E, double** norm2, double** G, long long* g_count, double** QG, long long* qg_count);
int bengine_get_begin(BEngine* E, int* level, int* ng, cudaStream_t st);   // open chunk must be closed
int bengine_get_finish(BEngine* E, double* sketch, double* sigma2, double* vt, double* delta, int where,
                       cudaStream_t st);

// Column-sharded mode with the exchanges inside the library (NCCL allreduce on the engine's stream)
int nccl_unique_id(void* out, long long cap, long long* n_bytes, const char* nccl_path);
int bengine_attach_comm(BEngine* E, const void* unique_id, long long id_bytes, int rank, int nranks, const char* nccl_path);
int bengine_comm_stats(BEngine* E, unsigned long long* n, unsigned long long* bytes);

// ---- state_dict support (optional checkpoint / resume; the reference has none) -----------------------
// A handle's state = its persistent device allocations, byte for byte, + a few host scalars.  The blob can only be
// loaded into a handle created with the same configuration (same allocation list): sizes are checked.
struct StateAlloc { void* p; size_t bytes; bool state; };
struct StateHeader {
    unsigned long long magic;          // "MUSEDST1"
    unsigned int kind, n_allocs;       // 0 = reference schedule, 1 = batched
    unsigned long long host[6];
};
constexpr unsigned long long STATE_MAGIC = 0x315453444553554dull;
long long state_blob_bytes(const std::vector<StateAlloc>& a);
int state_blob_save(unsigned kind, const unsigned long long (&host)[6], const std::vector<StateAlloc>& a, void* buf, long long cap);
int state_blob_load(unsigned kind, unsigned long long (&host)[6], const std::vector<StateAlloc>& a, const void* buf, long long bytes);
int bengine_state(BEngine* E, int op, void* buf, long long cap, long long* bytes);   // op 0 size, 1 save, 2 load

// Shared query tail (swfd_engine.cu): top-l shrink of the float64 stack M (ng x d).
struct QueryScratch {
    double *M = nullptr, *G = nullptr, *Vq = nullptr, *cs = nullptr, *eval = nullptr, *scr = nullptr;
    int *perm = nullptr, *counters = nullptr;
    int ng_max = 0, lda = 0;
    double rel_thr = 0.1 * 2.220446049250313e-16;   // off-diagonal cut-off relative to mean |diag|
    // whole-GPU block Jacobi (eig_big.cuh): eigenvalues per column slot, eigenvectors one per row,
    // neighbour mailboxes, flags / reductions / info (zeroed per call)
    double *lam = nullptr, *VT = nullptr, *mailL = nullptr, *mailR = nullptr;
    unsigned long long* sync = nullptr;
    int npad_max = 0, ldt = 0;
    size_t mail_count = 0;
    double big_tol = 1e-4;                           // largest cosine measured in the last sweep
    void plan_big() {
        npad_max = ng_max + 12;
        ldt = (ng_max + 31) / 32 * 32;
        mail_count = (size_t)(ng_max / 2 + 8) * ldt;
    }
};
constexpr int QSYNC_WORDS = 1024;   // flagL[256] | flagR[256] | red[128] | info
int merge_query_gram(const QueryScratch& Q, int ng, long long d, int sm_count, cudaStream_t st);
int merge_query_solve(const QueryScratch& Q, int ng, long long d, int ell, int sm_count, double* o_sketch,
                      double* o_vt, double* o_s2, cudaStream_t st);
int merge_query(const QueryScratch& Q, int ng, long long d, int ell, int sm_count, double* o_sketch,
                double* o_vt, double* o_s2, cudaStream_t st);

}  // namespace mused

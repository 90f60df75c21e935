// Batched-schedule SWFD engine: device state and the host entry points the C ABI
// (swfd_engine.cu) forwards to.
#pragma once
#include "common.cuh"
#include "eig_jacobi.cuh"

namespace mused {

constexpr int BS_MAX_ELL = 128;

struct BDev {
    int S, ell, n2, capn, KS, pad;
    long long d;
    unsigned long long N;
    double qcap;
    BSlot* slots;
    float *skhi, *sklo;        // [S][2][ell][d] sketch planes (TF32 hi / lo), ping-pong
    float *xhi, *xlo;          // [ell][d] the chunk being assembled / processed
    double* norm2;             // [ell]
    unsigned char* mask;       // [S][ell] 1 = chunk row is a raw snapshot at this instance's level
    float* Gp;                 // [S][KS][n2][n2]
    float* T;                  // [S][ell][n2]
    double* s2;                // [S][ell]
    int* dump_slot;            // [S][ell]
    float* ringV;              // [S][capn][d] snapshot rows
    unsigned long long *ringS, *ringT;
};

struct BEngine;

int bengine_create(BEngine** out, long long N, long long d, long long ell, double beta, int n_levels,
                   const double* thr, int device, long long batch);
void bengine_destroy(BEngine* E);
int bengine_fit(BEngine* E, const void* rows, long long n_rows, long long ld, int dtype, int where,
                cudaStream_t st);
int bengine_get(BEngine* E, double* sketch, double* sigma2, double* vt, double* delta, int* level,
                int where, cudaStream_t st);
int bengine_size(BEngine* E, long long* n);
unsigned long long bengine_time(BEngine* E);
int bengine_qcap(BEngine* E);
int bengine_debug(BEngine* E, double* scal, double* sigma2, unsigned long long* q_s,
                  unsigned long long* q_t, long long q_cap);
// CUDA-event time per kernel class since the last reset: ms[5] = {stage, book, gram, eig, rotate},
// chunks processed.  enable: 1 arm+reset, 0 disarm+reset, -1 read only.
int bengine_timing(BEngine* E, int enable, double* ms, long long* chunks, long long* rows);

// Shared query tail (swfd_engine.cu): top-l shrink of the float64 stack M (ng x d).
struct QueryScratch {
    double *M = nullptr, *G = nullptr, *Vq = nullptr, *cs = nullptr, *eval = nullptr, *scr = nullptr;
    int *perm = nullptr, *counters = nullptr;
    int ng_max = 0, lda = 0;
};
int merge_query(const QueryScratch& Q, int ng, long long d, int ell, int sm_count, double* o_sketch,
                double* o_vt, double* o_s2, cudaStream_t st);

}  // namespace mused

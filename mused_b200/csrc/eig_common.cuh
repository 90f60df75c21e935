// Shared pieces of the batched eigensolve K3 (eig_smem.cuh): the per-instance slot record, the kernel
// arguments, distributed-shared-memory accessors and the MUFU / packed-float32 helpers.
#pragma once
#include <float.h>
#include "common.cuh"

namespace mused {
namespace cg = cooperative_groups;

struct BSlot {                               // one FD instance of the batched engine
    double delta, error;
    unsigned long long last_dump, period;
    unsigned long long n_flush, n_dump, n_raw, overflow;
    int q_head, q_len;
    int cur, src;                            // ping-pong plane holding the sketch / source of the pending rotate
    int parity, skip, npend, nd;
    int sweeps, nraw;                        // nraw: raw-row snapshots of the current chunk (K5 -> K4 copies the rows)
};

struct EigArgs {
    BSlot* slots;
    const double* G;         // [S][n2][n2] Gram (split-K partials already added, all column shards)
    int ell, n2, capn;
    int ell_l;               // logical sketch_dim (<= ell): shrink at rank ell_l, ranks >= ell_l give zero rows
    const unsigned char* mask;  // [S][ell] 1 = row of the chunk is NOT part of this instance's batch
    int n_rows;              // rows in the chunk
    unsigned long long t_end;
    float *Thi, *Tlo;        // [S][ell][n2] rotate coefficients, TF32 hi / lo planes
    double* s2;              // [S][ell]
    int* dump_slot;          // [S][ell] ring entry or -1
    unsigned long long* ringS;  // [S][capn]
    unsigned long long* ringT;
    double tol;              // convergence: max |x.y| / (|x||y|)
    int max_sweeps;
    double tol32;            // mixed: the float32 phase ends after the sweep that measured a largest cosine <= tol32
    double small_col;        // mixed: smallest trusted (column norm / largest column norm)^2 after the float32 phase
    int mixed;               // 1 = float32 sweeps first, then W = G V32 and float64 sweeps (see k_eig_jacobi)
    double* Gs;              // [S][n2][n2] scratch: the symmetrised, masked Gram (mixed only)
};

// ---- distributed shared memory by cluster address -----------------------------------------
__device__ __forceinline__ uint32_t dsmem_addr(const void* p, int rank) {
    uint32_t a = (uint32_t)__cvta_generic_to_shared(p), r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(rank));
    return r;
}
__device__ __forceinline__ double dsmem_ld(uint32_t addr) {
    double v;
    asm volatile("ld.shared::cluster.f64 %0, [%1];" : "=d"(v) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void dsmem_st(uint32_t addr, double v) {
    asm volatile("st.shared::cluster.f64 [%0], %1;" ::"r"(addr), "d"(v) : "memory");
}
__device__ __forceinline__ void dsmem_ld_t(uint32_t addr, double& v) { v = dsmem_ld(addr); }
__device__ __forceinline__ void dsmem_ld_t(uint32_t addr, float& v) {
    asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
}
__device__ __forceinline__ void cluster_barrier() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

__device__ __forceinline__ double rsqrt_approx(double x) {       // MUFU.RSQ64H: ~20 bits
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    return y;
}
__device__ __forceinline__ double rcp_approx(double x) {         // MUFU.RCP64H: ~20 bits
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    return y;
}
__device__ __forceinline__ double sel(bool p, double a, double b) { return p ? a : b; }
__device__ __forceinline__ float sel(bool p, float a, float b) { return p ? a : b; }
__device__ __forceinline__ float rsqrt_approx(float x) {         // MUFU.RSQ
    float y;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float rcp_approx(float x) {           // MUFU.RCP
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// packed float32 pairs (FFMA2 / FMUL2 on sm_100): the float32 phase is bound by issued instructions, the
// two halves of a register pair ride one instruction; element-wise results equal the scalar fmaf / fmul
__device__ __forceinline__ unsigned long long pk2(float lo, float hi) {
    unsigned long long r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void upk2(unsigned long long v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
    unsigned long long r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
__device__ __forceinline__ unsigned long long mul2(unsigned long long a, unsigned long long b) {
    unsigned long long r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}

}  // namespace mused

// Shared device/host helpers for the dvf_b200 kernels (sm_100a only).
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstdarg>
#include <cuda_runtime.h>
#include "../../include/dvf_b200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "dvf_b200 is written for sm_100a (B200) only"
#endif

namespace dvf {

// ---- error plumbing -------------------------------------------------------------------------
void set_error(const char* fmt, ...);
// optional CUDA-event timing of the Gram kernel inside dvf_fit_em (dvf_profile_enable)
void profile_gram_begin(cudaStream_t st);
void profile_gram_end(cudaStream_t st);
int cuda_fail(cudaError_t e, const char* what, int line);

#define DVF_CUDA(x)                                                        \
    do {                                                                   \
        cudaError_t e__ = (x);                                             \
        if (e__ != cudaSuccess) return dvf::cuda_fail(e__, #x, __LINE__);  \
    } while (0)
// every kernel launch of the library goes through this macro: it also feeds the launch counter that
// bench.py reports as gpu_launches (dvf_profile_get)
extern long long g_launch_count;
#define DVF_LAUNCH_CHECK()              \
    do {                                \
        ++dvf::g_launch_count;          \
        DVF_CUDA(cudaGetLastError());   \
    } while (0)
#define DVF_ARG(cond, ...)                                                 \
    do {                                                                   \
        if (!(cond)) { dvf::set_error(__VA_ARGS__); return DVF_ERR_ARG; }  \
    } while (0)

constexpr int kMaxD = 64;   // dimensions supported by the register-tiled kernels
inline int num_sms() {
    static int n = 0;
    if (!n) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev); }
    return n ? n : 148;
}
// D padded to the DMMA granularity (8)
__host__ __device__ inline int pad8(int d) { return (d + 7) & ~7; }
// leading dimension == 4 (mod 16) doubles: conflict-free DMMA fragment loads
__host__ __device__ inline int ld4mod16(int w) { int r = w & 15; return w + ((4 - r) & 15); }

// ---- DMMA m8n8k4 (the only native FP64 tensor shape on sm_100a; 37.1 TFLOP/s measured) -------
// fragment ownership, lane = 4*g + t (g = lane>>2, t = lane&3):
//   a = A[row g][k t]      b = B[k t][col g]      c0 = C[row g][col 2t], c1 = C[row g][col 2t+1]
__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// ---- cp.async (LDGSTS) 16-byte copies ---------------------------------------------------------
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async16_zfill(void* smem, const void* gmem, bool valid) {
    unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    int sz = valid ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(s), "l"(gmem), "r"(sz));
}
__device__ __forceinline__ void cp_async8(void* smem, const void* gmem) {
    unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N)); }

// ---- mbarrier + bulk async copy (TMA engine, 1-D) ----------------------------------------------
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "DVF_WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DVF_DONE_%=;\n"
        "bra DVF_WAIT_%=;\n"
        "DVF_DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// global -> shared bulk copy; bytes % 16 == 0, both addresses 16-byte aligned; completes on `bar` (tx bytes)
__device__ __forceinline__ void bulk_g2s(void* smem, const void* gmem, unsigned bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(smem)), "l"(gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ int ld_volatile_i32(const int* p) {
    int v;
    asm volatile("ld.volatile.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_volatile_i32(int* p, int v) {
    asm volatile("st.volatile.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// ---- reductions -------------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_min(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
// block-wide sum; result valid in thread 0.  `sh` needs 32 doubles.  Fixed order => deterministic.
__device__ __forceinline__ double block_sum(double v, double* sh) {
    v = warp_sum(v);
    int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = (blockDim.x + 31) >> 5;
    __syncthreads();
    if (l == 0) sh[w] = v;
    __syncthreads();
    if (w == 0) {
        v = (l < nw) ? sh[l] : 0.0;
        v = warp_sum(v);
    }
    return v;
}
__device__ __forceinline__ double block_min(double v, double* sh) {
    v = warp_min(v);
    int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = (blockDim.x + 31) >> 5;
    __syncthreads();
    if (l == 0) sh[w] = v;
    __syncthreads();
    if (w == 0) {
        v = (l < nw) ? sh[l] : __longlong_as_double(0x7ff0000000000000LL);
        v = warp_min(v);
    }
    return v;
}

}  // namespace dvf

// K8: gene-space transform of per-cell Jacobians, out[e][r][c] = float32(Qe[e,:] J[:,:,c] Qr[r,:]^T)
// (reference utils.py:544-554, :497-541, :430-453) and K16: V Q^T (utils.py:311-326).
// Factored as T_e[j][c] = sum_i Qe[e][i] J[i][j][c]  (D^2 flops per (e, c)), then
// out[e][r][c] = sum_j Qr[r][j] T_e[j][c]; one CTA = one effector x 32 cells, T_e lives in shared memory,
// every global access is a 32-cell (256 B / 128 B) run along the cell-fastest layouts of J and out.
#include <cstdlib>
#include "common.cuh"
#include "tma.cuh"

namespace dvf {

constexpr int GT_CELLS = 32;
constexpr int GT_RT = 4;   // regulators per thread in the second stage

__global__ void __launch_bounds__(256) k_gene_transform(const double* __restrict__ J, int64_t n, int D,
                                                        const double* __restrict__ Qe, int Ge,
                                                        const double* __restrict__ Qr, int Gr, int take_abs,
                                                        const int32_t* __restrict__ group, int n_groups,
                                                        float* __restrict__ out_full, double* __restrict__ out_group) {
    extern __shared__ double smem[];
    double* T = smem;                       // [D][GT_CELLS]
    double* qe = T + D * GT_CELLS;          // [D]
    const int e = blockIdx.y;
    const int64_t c0 = (int64_t)blockIdx.x * GT_CELLS;
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    const int64_t cell = c0 + lane;
    const bool live = cell < n;
    for (int i = threadIdx.x; i < D; i += 256) qe[i] = Qe[(int64_t)e * D + i];
    __syncthreads();
    // stage 1: warp w handles columns j = w, w+8, ...
    for (int j = wrp; j < D; j += 8) {
        double s = 0.0;
        if (live)
            for (int i = 0; i < D; ++i) s = fma(qe[i], J[((int64_t)i * D + j) * n + cell], s);
        T[j * GT_CELLS + lane] = s;
    }
    __syncthreads();
    // stage 2: warp w handles regulators in groups of GT_RT
    const int lab = (group && live) ? group[cell] : -1;
    for (int r0 = wrp * GT_RT; r0 < Gr; r0 += 8 * GT_RT) {
        double acc[GT_RT];
#pragma unroll
        for (int q = 0; q < GT_RT; ++q) acc[q] = 0.0;
        for (int j = 0; j < D; ++j) {
            const double tv = T[j * GT_CELLS + lane];
#pragma unroll
            for (int q = 0; q < GT_RT; ++q)
                if (r0 + q < Gr) acc[q] = fma(__ldg(Qr + (int64_t)(r0 + q) * D + j), tv, acc[q]);
        }
#pragma unroll
        for (int q = 0; q < GT_RT; ++q) {
            const int r = r0 + q;
            if (r >= Gr) continue;
            float f = (float)acc[q];                       // reference stores float32 (utils.py:546)
            if (take_abs) f = fabsf(f);
            if (out_full && live) out_full[((int64_t)e * Gr + r) * n + cell] = f;
            if (out_group) {
                // per-group sums of the float32-rounded values, accumulated in float64
                for (int gidx = 0; gidx < n_groups; ++gidx) {
                    const unsigned m = __ballot_sync(0xffffffffu, lab == gidx);
                    if (!m) continue;
                    double v = (lab == gidx) ? (double)f : 0.0;
                    v = warp_sum(v);
                    if (lane == 0) atomicAdd(out_group + ((int64_t)gidx * Ge + e) * Gr + r, v);
                }
            }
        }
    }
}

__global__ void __launch_bounds__(256) k_gene_elem(const double* __restrict__ J, int64_t n, int D,
                                                   const double* __restrict__ qe, const double* __restrict__ qr,
                                                   double* __restrict__ out) {
    const int64_t cell = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (cell >= n) return;
    double s = 0.0;
    for (int i = 0; i < D; ++i) {
        double row = 0.0;
        for (int j = 0; j < D; ++j) row = fma(J[((int64_t)i * D + j) * n + cell], qr[j], row);
        s = fma(qe[i], row, s);
    }
    out[cell] = s;
}

// out[c][gidx] = sum_d V[c][d] Q[gidx][d]; 64 cells x 64 genes per CTA, 4 x 4 per thread
__global__ void __launch_bounds__(256) k_vector_transform(const double* __restrict__ V, int64_t n, int D,
                                                          const double* __restrict__ Q, int G, double* __restrict__ out) {
    __shared__ double vs[32][64 + 4];
    __shared__ double qs[32][64 + 4];
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const int64_t n0 = (int64_t)blockIdx.x * 64;
    const int g0 = blockIdx.y * 64;
    double acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.0;
    for (int d0 = 0; d0 < D; d0 += 32) {
        __syncthreads();
        for (int e = threadIdx.x; e < 64 * 32; e += 256) {
            const int r = e / 32, d = e % 32;
            vs[d][r] = (n0 + r < n && d0 + d < D) ? V[(n0 + r) * D + d0 + d] : 0.0;
            qs[d][r] = (g0 + r < G && d0 + d < D) ? Q[(int64_t)(g0 + r) * D + d0 + d] : 0.0;
        }
        __syncthreads();
#pragma unroll 8
        for (int d = 0; d < 32; ++d) {
            double a[4], b[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) { a[i] = vs[d][ty * 4 + i]; b[i] = qs[d][tx * 4 + i]; }
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fma(a[i], b[j], acc[i][j]);
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int64_t c = n0 + ty * 4 + i;
            const int gg = g0 + tx * 4 + j;
            if (c < n && gg < G) out[c * G + gg] = acc[i][j];
        }
}

}  // namespace dvf

using namespace dvf;

extern "C" int dvf_gene_transform(const double* J, int64_t n, int D, const double* Qe, int Ge, const double* Qr, int Gr,
                                  int take_abs, const int32_t* group, int n_groups, float* out_full,
                                  double* out_group_sum, void* stream) {
    DVF_ARG(J && Qe && Qr && n >= 0 && D >= 1 && Ge >= 1 && Gr >= 1, "bad argument");
    DVF_ARG(out_full || out_group_sum, "no output requested");
    DVF_ARG(!out_group_sum || (group && n_groups >= 1), "group labels missing");
    DVF_ARG(Ge <= 65535, "Ge > 65535: split the effector set");
    if (n == 0) return DVF_OK;
    // large gene sets: two batched DGEMMs on the FP64 tensor cores (kern_gene_gemm.cu); its 128-row tiles would be
    // mostly empty for short gene lists, which stay on the per-effector kernel below.  DVF_GENE_GEMM=0/1 forces a path.
    {
        static const char* force = getenv("DVF_GENE_GEMM");
        const bool ok = (n % 2) == 0 && ((uintptr_t)J & 15) == 0 && D <= kMaxD;
        const bool want = force ? atoi(force) != 0 : (Ge >= 96 && Gr >= 96 && n >= 128);
        if (ok && want) {
            size_t bytes = 0;
            if (int rc = dvf_gene_gemm_workspace_bytes(n, D, Ge, Gr, out_full != nullptr, out_group_sum != nullptr, &bytes)) return rc;
            void* ws = nullptr;
            if (int rc = pool_alloc(&ws, bytes, (cudaStream_t)stream)) return rc;
            const int rc = dvf_gene_transform_gemm(J, n, D, Qe, Ge, Qr, Gr, take_abs, group, n_groups, out_full, out_group_sum,
                                                   ws, bytes, stream);
            cudaFreeAsync(ws, (cudaStream_t)stream);
            return rc;
        }
    }
    const size_t smem = sizeof(double) * ((size_t)D * GT_CELLS + D);
    DVF_ARG(smem <= 200 * 1024, "D too large for the gene transform");
    if (smem > 48 * 1024)
        DVF_CUDA(cudaFuncSetAttribute(k_gene_transform, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((unsigned)((n + GT_CELLS - 1) / GT_CELLS), Ge);
    k_gene_transform<<<grid, 256, smem, (cudaStream_t)stream>>>(J, n, D, Qe, Ge, Qr, Gr, take_abs, group, n_groups,
                                                              out_full, out_group_sum);
    DVF_LAUNCH_CHECK();
    return DVF_OK;
}

extern "C" int dvf_gene_transform_elem(const double* J, int64_t n, int D, const double* qe, const double* qr,
                                       double* out_elem, void* stream) {
    DVF_ARG(J && qe && qr && out_elem && n >= 0 && D >= 1, "bad argument");
    if (n == 0) return DVF_OK;
    k_gene_elem<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(J, n, D, qe, qr, out_elem);
    DVF_LAUNCH_CHECK();
    return DVF_OK;
}

extern "C" int dvf_vector_transform(const double* V, int64_t n, int D, const double* Q, int G, double* out, void* stream) {
    DVF_ARG(V && Q && out && n >= 0 && D >= 1 && G >= 1, "bad argument");
    if (n == 0) return DVF_OK;
    dim3 grid((unsigned)((n + 63) / 64), (unsigned)((G + 63) / 64));
    k_vector_transform<<<grid, 256, 0, (cudaStream_t)stream>>>(V, n, D, Q, G, out);
    DVF_LAUNCH_CHECK();
    return DVF_OK;
}

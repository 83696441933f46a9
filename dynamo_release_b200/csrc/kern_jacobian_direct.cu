// K6 (direct form, kept for cross-checks: DVF_JAC_DIRECT=1): per-cell analytical Jacobian of the Gaussian-RKHS field
// (reference utils.py:351-379)
//     J_c[i][j] = -2 beta sum_m k(x_c, Xc_m) C[m][i] (x_c[j] - Xc[m][j])
// computed as one small FP64 tensor-core GEMM per cell:  J_c = A_c^T B_c  with
//     A_c[m][i] = k_cm C[m][i]      B_c[m][j] = x_c[j] - Xc[m][j]
// Both operands are generated in registers straight into DMMA m8n8k4 fragments (one DMUL / DADD per fragment
// element, 2/D of the MMA work); the D x D accumulators of a cell live in the registers of one warp; K is
// never materialised.  Control points / coefficients are staged through shared memory in chunks of 64.
#include "common.cuh"

namespace dvf {

constexpr int JC_MC = 64;        // ctrl points per chunk
constexpr int JC_WARPS = 8;

template <int NT>
struct jac_cfg {
    static constexpr int DP = NT * 8;
    static constexpr int CPW = (NT <= 4) ? 2 : 1;            // cells per warp
    static constexpr int SPLIT = (NT >= 7) ? 2 : 1;          // warps sharing one cell (row halves of J)
    static constexpr int NTI = (SPLIT == 2) ? (NT + 1) / 2 : NT;   // row tiles owned by a warp
    static constexpr int CELLS = JC_WARPS * CPW / SPLIT;     // cells per CTA
    static constexpr int LDC = ((DP + 15) / 16) * 16 + 4;    // == 4 (mod 16): conflict-free fragment loads
    static constexpr int LDT = JC_MC + 1;                    // transposed ctrl tile, lane = ctrl point
    static constexpr size_t SMEM = sizeof(double) * (size_t)(2 * JC_MC * LDC + DP * LDT + CELLS * DP);
};

template <int NT>
__global__ void __launch_bounds__(JC_WARPS * 32) k_jacobian(const double* __restrict__ x, int64_t n, int D,
                                                            const double* __restrict__ Xc, int M, double beta,
                                                            const double* __restrict__ C, double* __restrict__ J) {
    using cfg = jac_cfg<NT>;
    constexpr int DP = cfg::DP, CPW = cfg::CPW, LDC = cfg::LDC, LDT = cfg::LDT, NTI = cfg::NTI;
    extern __shared__ double smem[];
    double* cs = smem;                     // [MC][LDC]  ctrl points, row-major (B fragments)
    double* Cs = cs + JC_MC * LDC;         // [MC][LDC]  coefficients, row-major (A fragments)
    double* cT = Cs + JC_MC * LDC;         // [DP][LDT]  ctrl points transposed (distance phase)
    double* xs = cT + DP * LDT;            // [CELLS][DP] this CTA's cells

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int64_t cell0 = (int64_t)blockIdx.x * cfg::CELLS;
    const int wcell = (cfg::SPLIT == 2) ? (warp >> 1) : warp * CPW;   // first local cell of this warp
    const int ib0 = (cfg::SPLIT == 2) ? (warp & 1) * NTI : 0;         // first row tile of this warp

    for (int e = threadIdx.x; e < cfg::CELLS * DP; e += blockDim.x) {
        int r = e / DP, d = e % DP;
        int64_t c = cell0 + r;
        xs[e] = (c < n && d < D) ? x[c * D + d] : 0.0;
    }
    __syncthreads();

    double acc[CPW][NTI][NT][2];
    double xg[CPW][NT];
#pragma unroll
    for (int c = 0; c < CPW; ++c) {
#pragma unroll
        for (int jb = 0; jb < NT; ++jb) xg[c][jb] = xs[(wcell + c) * DP + jb * 8 + g];
#pragma unroll
        for (int ib = 0; ib < NTI; ++ib)
#pragma unroll
            for (int jb = 0; jb < NT; ++jb) acc[c][ib][jb][0] = acc[c][ib][jb][1] = 0.0;
    }

    for (int m0 = 0; m0 < M; m0 += JC_MC) {
        __syncthreads();
        for (int e = threadIdx.x; e < JC_MC * DP; e += blockDim.x) {
            int r = e / DP, d = e % DP;
            bool ok = (m0 + r < M) && d < D;
            double cv = ok ? Xc[(int64_t)(m0 + r) * D + d] : 0.0;
            cs[r * LDC + d] = cv;
            cT[d * LDT + r] = cv;
            Cs[r * LDC + d] = ok ? C[(int64_t)(m0 + r) * D + d] : 0.0;
        }
        __syncthreads();

        // distance phase: lane = ctrl point, two rounds of 32
        double kk[CPW][JC_MC / 32];
#pragma unroll
        for (int c = 0; c < CPW; ++c) {
            const double* xc = xs + (wcell + c) * DP;
#pragma unroll
            for (int r = 0; r < JC_MC / 32; ++r) {
                double d2 = 0.0;
#pragma unroll 8
                for (int d = 0; d < DP; ++d) {
                    double tt = xc[d] - cT[d * LDT + r * 32 + lane];
                    d2 = fma(tt, tt, d2);
                }
                kk[c][r] = (m0 + r * 32 + lane < M) ? exp(-beta * d2) : 0.0;
            }
        }

        // MMA phase: k-steps of 4 ctrl points
#pragma unroll 2
        for (int s = 0; s < JC_MC / 4; ++s) {
            const int row = 4 * s + t;
            double cfrag[NT], Cfrag[NTI];
#pragma unroll
            for (int b = 0; b < NT; ++b) cfrag[b] = cs[row * LDC + b * 8 + g];
#pragma unroll
            for (int b = 0; b < NTI; ++b) Cfrag[b] = (ib0 + b < NT) ? Cs[row * LDC + (ib0 + b) * 8 + g] : 0.0;
#pragma unroll
            for (int c = 0; c < CPW; ++c) {
                const double kv = __shfl_sync(0xffffffffu, (s < 8) ? kk[c][0] : kk[c][1], ((4 * s) & 31) + t);
                double a[NTI], b[NT];
#pragma unroll
                for (int q = 0; q < NTI; ++q) a[q] = kv * Cfrag[q];
#pragma unroll
                for (int q = 0; q < NT; ++q) b[q] = xg[c][q] - cfrag[q];
#pragma unroll
                for (int ib = 0; ib < NTI; ++ib)
                    if (cfg::SPLIT == 1 || ib0 + ib < NT) {
#pragma unroll
                        for (int jb = 0; jb < NT; ++jb) dmma(acc[c][ib][jb][0], acc[c][ib][jb][1], a[ib], b[jb]);
                    }
            }
        }
    }

    // epilogue: J[(i*D + j)*n + cell] (cell index fastest, reference layout (D, D, n))
    const double sc = -2.0 * beta;
#pragma unroll
    for (int c = 0; c < CPW; ++c) {
        const int64_t cell = cell0 + wcell + c;
        if (cell >= n) continue;
#pragma unroll
        for (int ib = 0; ib < NTI; ++ib)
#pragma unroll
            for (int jb = 0; jb < NT; ++jb) {
                const int i = (ib0 + ib) * 8 + g, j = jb * 8 + 2 * t;
                if (i < D) {
                    if (j < D) J[((int64_t)i * D + j) * n + cell] = sc * acc[c][ib][jb][0];
                    if (j + 1 < D) J[((int64_t)i * D + j + 1) * n + cell] = sc * acc[c][ib][jb][1];
                }
            }
    }
}

template <int NT>
static int launch_jacobian(const double* x, int64_t n, int D, const double* Xc, int M, double beta, const double* C,
                           double* J, cudaStream_t st) {
    using cfg = jac_cfg<NT>;
    static bool attr_set = false;
    if (!attr_set) {
        DVF_CUDA(cudaFuncSetAttribute(k_jacobian<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg::SMEM));
        attr_set = true;
    }
    const unsigned grid = (unsigned)((n + cfg::CELLS - 1) / cfg::CELLS);
    k_jacobian<NT><<<grid, JC_WARPS * 32, cfg::SMEM, st>>>(x, n, D, Xc, M, beta, C, J);
    DVF_LAUNCH_CHECK();
    return DVF_OK;
}

}  // namespace dvf

using namespace dvf;

int dvf_jacobian_direct_impl(const double* x, int64_t n, int D, const double* Xc, int M, double beta,
                             const double* C, double* J, void* stream) {
    DVF_ARG(n >= 0 && M >= 1, "n=%lld M=%d: need n >= 0, M >= 1", (long long)n, M);
    DVF_ARG(D >= 1 && D <= kMaxD, "D=%d unsupported (1..%d)", D, kMaxD);
    DVF_ARG((x || n == 0) && Xc && C && (J || n == 0), "NULL pointer");
    if (n == 0) return DVF_OK;
    cudaStream_t st = (cudaStream_t)stream;
    switch (pad8(D) / 8) {
        case 1: return launch_jacobian<1>(x, n, D, Xc, M, beta, C, J, st);
        case 2: return launch_jacobian<2>(x, n, D, Xc, M, beta, C, J, st);
        case 3: return launch_jacobian<3>(x, n, D, Xc, M, beta, C, J, st);
        case 4: return launch_jacobian<4>(x, n, D, Xc, M, beta, C, J, st);
        case 5: return launch_jacobian<5>(x, n, D, Xc, M, beta, C, J, st);
        case 6: return launch_jacobian<6>(x, n, D, Xc, M, beta, C, J, st);
        case 7: return launch_jacobian<7>(x, n, D, Xc, M, beta, C, J, st);
        default: return launch_jacobian<8>(x, n, D, Xc, M, beta, C, J, st);
    }
}

// K4: control-point solve of the M-step,  (G[:M,:M] + lambda sigma2 Kc) C = rhs,  rhs[m][d] = G[M+d][m].
// Reference: scVectorField.py:475-481 -> tools/utils.py:2948-2981 forms the normal equations
// solve(lhs^T lhs, lhs^T rhs).  lhs is symmetric positive definite (PSD Gram + lambda sigma2 * PD kernel
// matrix), so this kernel factors lhs itself (blocked right-looking Cholesky, FP64) -- the same solution in
// exact arithmetic with an error bound of cond(lhs) eps instead of cond(lhs)^2 eps ("solver").
// One cooperative launch: the D right-hand-side columns ride along as D extra rows of the panel, so the
// forward substitution costs nothing extra; the backward substitution follows block by block.
#include <algorithm>
#include <cstdlib>
#include <cooperative_groups.h>
#include "common.cuh"

namespace cg = cooperative_groups;

namespace dvf {

constexpr int SV_NB = 32;      // panel width
constexpr int SV_T = 64;       // trailing-update tile

struct solve_args {
    const double* G; int64_t ldg; const double* Kc; int M; int D;
    double lam_sigma2; const double* lam_sigma2_ptr;   // device scalar wins if non-NULL
    double* L; int ldl; double* C; int* info; const int* done;
};

__global__ void __launch_bounds__(256) k_chol_solve(solve_args a) {
    cg::grid_group grid = cg::this_grid();
    if (a.done && *a.done) return;            // uniform over the grid: no sync is skipped by a subset
    __shared__ double Dg[SV_NB][SV_NB + 1];
    __shared__ double Pi[SV_T][SV_NB + 1];
    __shared__ double Pj[SV_T][SV_NB + 1];
    static_assert(SV_NB * (kMaxD + 1) <= SV_T * (SV_NB + 1), "Cb must fit in Pi");
    double (*Cb)[kMaxD + 1] = reinterpret_cast<double (*)[kMaxD + 1]>(&Pi[0][0]);   // backward phase only
    const int M = a.M, D = a.D, R = a.M + a.D, ldl = a.ldl;
    double* L = a.L;
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int64_t gtid = (int64_t)blockIdx.x * nthr + tid, gsz = (int64_t)gridDim.x * nthr;
    const double ls = a.lam_sigma2_ptr ? *a.lam_sigma2_ptr : a.lam_sigma2;

    // phase 0: working matrix, rows 0..M-1 = lower triangle of lhs, rows M..M+D-1 = rhs^T
    for (int64_t e = gtid; e < (int64_t)R * M; e += gsz) {
        const int i = (int)(e / M), j = (int)(e % M);
        double v = 0.0;
        if (i >= M) v = a.G[(int64_t)i * a.ldg + j];
        else if (j <= i) v = a.G[(int64_t)i * a.ldg + j] + ls * a.Kc[(int64_t)i * M + j];
        L[(int64_t)i * ldl + j] = v;
    }
    grid.sync();

    for (int k0 = 0; k0 < M; k0 += SV_NB) {
        const int nb = min(SV_NB, M - k0);
        // 1. every CTA factors the nb x nb diagonal block redundantly in shared memory
        for (int e = tid; e < SV_NB * SV_NB; e += nthr) {
            const int i = e / SV_NB, j = e % SV_NB;
            Dg[i][j] = (i < nb && j <= i) ? L[(int64_t)(k0 + i) * ldl + k0 + j] : (i == j ? 1.0 : 0.0);
        }
        __syncthreads();
        for (int j = 0; j < nb; ++j) {
            const double d = Dg[j][j];
            __syncthreads();
            if (!(d > 0.0) && blockIdx.x == 0 && tid == 0) atomicCAS(a.info, 0, k0 + j + 1);
            const double l = sqrt(d);
            if (tid > j && tid < nb) Dg[tid][j] /= l;
            if (tid == j) Dg[j][j] = l;
            __syncthreads();
            for (int e = tid; e < nb * nb; e += nthr) {
                const int i = e / nb, c = e % nb;
                if (c > j && i >= c) Dg[i][c] -= Dg[i][j] * Dg[c][j];
            }
            __syncthreads();
        }
        if (blockIdx.x == 0)
            for (int e = tid; e < nb * nb; e += nthr) {
                const int i = e / nb, j = e % nb;
                if (j <= i) L[(int64_t)(k0 + i) * ldl + k0 + j] = Dg[i][j];
            }
        // 2. panel rows below the block (including the D rhs rows): x Dg^T = a, one thread per row
        for (int64_t r = k0 + nb + (int64_t)tid * gridDim.x + blockIdx.x; r < R; r += gsz) {
            double* row = L + r * ldl + k0;
            double xv[SV_NB];
#pragma unroll
            for (int j = 0; j < SV_NB; ++j) xv[j] = (j < nb) ? row[j] : 0.0;
#pragma unroll
            for (int j = 0; j < SV_NB; ++j) {
                double sacc = xv[j];
#pragma unroll
                for (int c = 0; c < j; ++c) sacc = fma(-xv[c], Dg[j][c], sacc);
                xv[j] = sacc / Dg[j][j];
            }
#pragma unroll
            for (int j = 0; j < SV_NB; ++j)
                if (j < nb) row[j] = xv[j];
        }
        grid.sync();
        // 3. trailing update  L[i][j] -= sum_c L[i][k0+c] L[j][k0+c]   for i >= j >= k0+nb, j < M
        const int r0 = k0 + nb;
        if (r0 < M) {
            const int nti = (R - r0 + SV_T - 1) / SV_T, ntj = (M - r0 + SV_T - 1) / SV_T;
            for (int tile = blockIdx.x; tile < nti * ntj; tile += gridDim.x) {
                const int ti = tile / ntj, tj = tile % ntj;
                if (tj > ti) continue;
                const int i0 = r0 + ti * SV_T, j0 = r0 + tj * SV_T;
                __syncthreads();
                for (int e = tid; e < SV_T * SV_NB; e += nthr) {
                    const int r = e / SV_NB, c = e % SV_NB;
                    Pi[r][c] = (i0 + r < R && c < nb) ? L[(int64_t)(i0 + r) * ldl + k0 + c] : 0.0;
                    Pj[r][c] = (j0 + r < M && c < nb) ? L[(int64_t)(j0 + r) * ldl + k0 + c] : 0.0;
                }
                __syncthreads();
                const int ty = tid / 16, tx = tid % 16;
                double acc[4][4];
#pragma unroll
                for (int u = 0; u < 4; ++u)
#pragma unroll
                    for (int v = 0; v < 4; ++v) acc[u][v] = 0.0;
#pragma unroll 8
                for (int c = 0; c < SV_NB; ++c) {
                    double av[4], bv[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) { av[u] = Pi[ty + 16 * u][c]; bv[u] = Pj[tx + 16 * u][c]; }
#pragma unroll
                    for (int u = 0; u < 4; ++u)
#pragma unroll
                        for (int v = 0; v < 4; ++v) acc[u][v] = fma(av[u], bv[v], acc[u][v]);
                }
#pragma unroll
                for (int u = 0; u < 4; ++u)
#pragma unroll
                    for (int v = 0; v < 4; ++v) {
                        const int i = i0 + ty + 16 * u, j = j0 + tx + 16 * v;
                        if (i < R && j < M && j <= i) L[(int64_t)i * ldl + j] -= acc[u][v];
                    }
            }
        }
        grid.sync();
    }

    // backward substitution  L^T C = Z,  Z[m][d] = L[M+d][m]
    const int nblk = (M + SV_NB - 1) / SV_NB;
    for (int kb = nblk - 1; kb >= 0; --kb) {
        const int k0 = kb * SV_NB, nb = min(SV_NB, M - k0);
        __syncthreads();
        for (int e = tid; e < SV_NB * SV_NB; e += nthr) {
            const int i = e / SV_NB, j = e % SV_NB;
            Dg[i][j] = (i < nb && j <= i) ? L[(int64_t)(k0 + i) * ldl + k0 + j] : (i == j ? 1.0 : 0.0);
        }
        for (int e = tid; e < SV_NB * D; e += nthr) {
            const int i = e % SV_NB, d = e / SV_NB;
            Cb[i][d] = (i < nb) ? L[(int64_t)(M + d) * ldl + k0 + i] : 0.0;
        }
        __syncthreads();
        if (tid < D) {
            for (int i = nb - 1; i >= 0; --i) {
                double sacc = Cb[i][tid];
                for (int j = i + 1; j < nb; ++j) sacc = fma(-Dg[j][i], Cb[j][tid], sacc);
                Cb[i][tid] = sacc / Dg[i][i];
            }
        }
        __syncthreads();
        if (blockIdx.x == 0)
            for (int e = tid; e < nb * D; e += nthr) a.C[(int64_t)(k0 + e / D) * D + e % D] = Cb[e / D][e % D];
        // rows above the block: Z[i][d] -= sum_j L[k0+j][i] C[k0+j][d]
        for (int64_t e = gtid; e < (int64_t)k0 * D; e += gsz) {
            const int i = (int)(e % k0), d = (int)(e / k0);
            double sacc = L[(int64_t)(M + d) * ldl + i];
            for (int j = 0; j < nb; ++j) sacc = fma(-L[(int64_t)(k0 + j) * ldl + i], Cb[j][d], sacc);
            L[(int64_t)(M + d) * ldl + i] = sacc;
        }
        grid.sync();
    }
}

int solve_run(const double* G, int64_t ldg, const double* Kc, int M, int D, double lam_sigma2,
              const double* lam_sigma2_ptr, double* L, double* C, int* info, const int* done, cudaStream_t st) {
    solve_args a{G, ldg, Kc, M, D, lam_sigma2, lam_sigma2_ptr, L, M, C, info, done};
    static int max_grid = 0;
    if (!max_grid) {
        int per_sm = 0;
        DVF_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_chol_solve, 256, 0));
        max_grid = num_sms() * (per_sm > 0 ? 1 : 0);
        if (max_grid == 0) { set_error("k_chol_solve cannot be made resident"); return DVF_ERR_CUDA; }
    }
    int grid = max_grid;
    const int want = (M + D + 31) / 32 * 4;          // more CTAs than this only add grid-sync cost (tools/solve_tune.py)
    if (grid > want) grid = want < 8 ? 8 : want;
    if (grid > max_grid) grid = max_grid;
    static const char* force = getenv("DVF_SOLVE_GRID");   // developer knob (tools/solve_tune.py)
    if (force && atoi(force) > 0) grid = std::min(atoi(force), max_grid);
    void* params[] = {&a};
    ++g_launch_count;
    DVF_CUDA(cudaLaunchCooperativeKernel((void*)k_chol_solve, dim3(grid), dim3(256), params, 0, st));
    return DVF_OK;
}

}  // namespace dvf

using namespace dvf;

extern "C" int dvf_solve(const double* G, int64_t ldg, const double* Kc, int M, int D, double lam_sigma2,
                         double* L, double* C, int* info, void* stream) {
    DVF_ARG(G && Kc && L && C && info, "NULL pointer");
    DVF_ARG(M >= 1 && D >= 1 && D <= kMaxD && ldg >= M + D, "bad shape M=%d D=%d ldg=%lld", M, D, (long long)ldg);
    DVF_CUDA(cudaMemsetAsync(info, 0, sizeof(int), (cudaStream_t)stream));
    return solve_run(G, ldg, Kc, M, D, lam_sigma2, nullptr, L, C, info, nullptr, (cudaStream_t)stream);
}

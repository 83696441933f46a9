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

namespace dvf {

// =========================================================================================================
// K4': the reference's own solver form (lstsq_method="drouin"): tools/utils.py:2975-2976
//     i = dgemm(a.T, a.T, trans_b=True)          ->  N = lhs^T lhs
//     x = np.linalg.solve(i, dgemm(a.T, b))      ->  LAPACK gesv: LU with partial pivoting of N, then the two
//                                                    triangular solves applied to lhs^T rhs
// Squares cond(lhs) exactly as the reference does -- that is the point of this path.
// lhs is symmetric (lower triangle of the Gram mirrored + lambda sigma2 Kc), so lhs^T [lhs | rhs] = lhs [lhs | rhs]:
//   k_lu_prep  builds AB = [lhs | rhs]                               (grid-wide)
//   k_lu_ata   N = lhs AB, the augmented matrix [lhs^T lhs | lhs^T rhs] (M x (M + D)), tiles over the whole GPU
//   k_lu_factor  ONE thread-block cluster (16 CTAs; 8 where 16 cannot be scheduled): right-looking LU with partial
//              pivoting in panels of LU_NB columns, the right-hand sides riding along as extra columns (gesv's forward
//              substitution happens during the elimination), then the back substitution.  The three phases of a panel
//              are separated by cluster barriers (~0.2 us) instead of grid-wide ones; v1 of this kernel (cooperative
//              grid of 40 CTAs x 1024 threads, 4 block barriers per column, rows physically exchanged) took 2.28 ms at
//              M = 500, 46 % of it the other CTAs waiting for the panel.
// Panel (CTA 0, one matrix row per thread in registers): the pivot of column j is the first entry of largest magnitude
// (LAPACK idamax); rows are NOT moved -- every thread tracks the logical row number of the data it holds, a swap
// exchanges two numbers -- so a column costs one arg-max (warp shuffles + 16 shared words), one broadcast of the pivot
// row through shared memory and two block barriers.  The other columns of the swapped rows are gathered once per panel
// (composition of the panel's 16 swaps), fused with U12 = L11^-1 N12.
// =========================================================================================================
constexpr int LU_NB = 16;
constexpr int LU_THREADS = 512;
constexpr int LU_TR = 64, LU_TC = 128;   // tile of the product kernels: 64 rows x 128 columns, 4 x 4 outputs per thread
constexpr int LU_MAX_M = 4 * LU_THREADS;

struct lu_args {
    const double* G; int64_t ldg; const double* Kc; int M; int D;
    double lam_sigma2; const double* lam_sigma2_ptr;
    double* AB; double* N; int ldn; int* piv; double* C; int* info; const int* done;
};

// one 64 x 128 tile: Cd[i][j] = (acc ? Cd : 0) + sign * sum_{c<K} P[i][c] Q[c][j]
__device__ __forceinline__ void lu_tile(double* Cd, int ldc, const double* P, int ldp, const double* Q, int ldq, int rows,
                                        int cols, int K, double sign, bool acc, int i0, int j0,
                                        double (*Ps)[LU_NB + 1], double (*Qs)[LU_TC]) {
    const int tid = threadIdx.x;
    const int ty = tid >> 5, tx = tid & 31;      // 16 x 32 threads: rows ty + 16 u, columns tx + 32 v
    double s[4][4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int v = 0; v < 4; ++v) s[u][v] = 0.0;
    for (int c0 = 0; c0 < K; c0 += LU_NB) {
        __syncthreads();
        for (int e = tid; e < LU_TR * LU_NB; e += LU_THREADS) {
            const int r = e >> 4, c = e & 15;
            Ps[r][c] = (i0 + r < rows && c0 + c < K) ? P[(int64_t)(i0 + r) * ldp + c0 + c] : 0.0;
        }
        for (int e = tid; e < LU_NB * LU_TC; e += LU_THREADS) {
            const int c = e / LU_TC, j = e % LU_TC;
            Qs[c][j] = (c0 + c < K && j0 + j < cols) ? Q[(int64_t)(c0 + c) * ldq + j0 + j] : 0.0;
        }
        __syncthreads();
#pragma unroll
        for (int c = 0; c < LU_NB; ++c) {
            double pv[4], qv[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) pv[u] = Ps[ty + 16 * u][c];
#pragma unroll
            for (int v = 0; v < 4; ++v) qv[v] = Qs[c][tx + 32 * v];
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int v = 0; v < 4; ++v) s[u][v] = fma(pv[u], qv[v], s[u][v]);
        }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int v = 0; v < 4; ++v) {
            const int i = i0 + ty + 16 * u, j = j0 + tx + 32 * v;
            if (i < rows && j < cols) {
                double* dst = Cd + (int64_t)i * ldc + j;
                *dst = acc ? fma(sign, s[u][v], *dst) : sign * s[u][v];
            }
        }
}

// AB = [lhs | rhs], lhs symmetric from the lower triangle of G
__global__ void __launch_bounds__(256) k_lu_prep(lu_args a) {
    if (a.done && *a.done) return;
    const int M = a.M, W = a.M + a.D;
    const double ls = a.lam_sigma2_ptr ? *a.lam_sigma2_ptr : a.lam_sigma2;
    for (int64_t e = (int64_t)blockIdx.x * 256 + threadIdx.x; e < (int64_t)M * W; e += (int64_t)gridDim.x * 256) {
        const int i = (int)(e / W), j = (int)(e % W);
        double v;
        if (j >= M) v = a.G[(int64_t)j * a.ldg + i];                       // rhs[i][d] = G[M + d][i]
        else {
            const int hi = i > j ? i : j, lo = i > j ? j : i;
            v = a.G[(int64_t)hi * a.ldg + lo] + ls * a.Kc[(int64_t)i * M + j];
        }
        a.AB[(int64_t)i * a.ldn + j] = v;
    }
}

// N = lhs [lhs | rhs]: one 64 x 128 tile per CTA
__global__ void __launch_bounds__(LU_THREADS) k_lu_ata(lu_args a) {
    if (a.done && *a.done) return;
    __shared__ double Ps[LU_TR][LU_NB + 1];
    __shared__ double Qs[LU_NB][LU_TC];
    const int W = a.M + a.D, ntc = (W + LU_TC - 1) / LU_TC;
    lu_tile(a.N, a.ldn, a.AB, a.ldn, a.AB, a.ldn, a.M, W, a.M, 1.0, false, (blockIdx.x / ntc) * LU_TR,
            (blockIdx.x % ntc) * LU_TC, Ps, Qs);
}

__device__ __forceinline__ void lu_argmax_step(double& bv, int& bp, int& bo, int o) {
    const double ov = __shfl_xor_sync(0xffffffffu, bv, o);
    const int op = __shfl_xor_sync(0xffffffffu, bp, o);
    const int oo = __shfl_xor_sync(0xffffffffu, bo, o);
    if (ov > bv || (ov == bv && op < bp)) { bv = ov; bp = op; bo = oo; }
}

template <int R>
__global__ void __launch_bounds__(LU_THREADS) k_lu_factor(lu_args a) {
    cg::cluster_group cluster = cg::this_cluster();
    if (a.done && *a.done) return;               // uniform over the cluster
    __shared__ double Ps[LU_TR][LU_NB + 1];
    __shared__ double Qs[LU_NB][LU_TC];
    __shared__ double rowP[2][LU_NB];
    __shared__ double wval[2][LU_THREADS / 32];
    __shared__ int wpos[2][LU_THREADS / 32], wown[2][LU_THREADS / 32];
    __shared__ int s_piv[LU_NB];
    __shared__ int s_tp[2 * LU_NB], s_ts[2 * LU_NB], s_np;
    const int M = a.M, D = a.D, W = a.M + a.D, ldn = a.ldn;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int crank = (int)cluster.block_rank(), csize = (int)cluster.num_blocks();
    double* N = a.N;
    constexpr int NOROW = 0x7fffffff;

    for (int k0 = 0; k0 < M; k0 += LU_NB) {
        const int nbe = min(LU_NB, M - k0);
        // ---- (a) panel factorisation, CTA 0 ----
        if (crank == 0) {
            double av[R][LU_NB];
            int pos[R];
#pragma unroll
            for (int q = 0; q < R; ++q) {
                const int r = k0 + tid + LU_THREADS * q;
                pos[q] = r < M ? r : NOROW;
#pragma unroll
                for (int c = 0; c < LU_NB; ++c) av[q][c] = (r < M && c < nbe) ? N[(int64_t)r * ldn + k0 + c] : 0.0;
            }
#pragma unroll
            for (int j = 0; j < LU_NB; ++j) {
                if (j < nbe) {                     // uniform over the CTA
                    const int par = j & 1;
                    double bv = -1.0;
                    int bp = NOROW, bo = 0;
#pragma unroll
                    for (int q = 0; q < R; ++q) {
                        const double v = fabs(av[q][j]);
                        if (pos[q] != NOROW && pos[q] >= k0 + j && (v > bv || (v == bv && pos[q] < bp))) {
                            bv = v; bp = pos[q]; bo = tid * R + q;
                        }
                    }
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) lu_argmax_step(bv, bp, bo, o);
                    if (lane == 0) { wval[par][warp] = bv; wpos[par][warp] = bp; wown[par][warp] = bo; }
                    __syncthreads();
                    {   // every warp reduces the 16 warp candidates itself (no second barrier)
                        const int w = lane & (LU_THREADS / 32 - 1);
                        bv = wval[par][w]; bp = wpos[par][w]; bo = wown[par][w];
#pragma unroll
                        for (int o = 8; o > 0; o >>= 1) lu_argmax_step(bv, bp, bo, o);
                    }
                    if (tid == 0) {
                        s_piv[j] = bp;
                        if (!(bv > 0.0)) atomicCAS(a.info, 0, k0 + j + 1);      // exactly singular (or NaN)
                    }
#pragma unroll
                    for (int q = 0; q < R; ++q) {
                        if (tid * R + q == bo) {             // this thread holds the pivot row: it becomes row k0 + j
#pragma unroll
                            for (int c = 0; c < LU_NB; ++c) rowP[par][c] = av[q][c];
                            pos[q] = k0 + j;
                        } else if (pos[q] == k0 + j) {
                            pos[q] = bp;                     // the row that sat at k0 + j goes where the pivot row was
                        }
                    }
                    __syncthreads();
                    const double pv = rowP[par][j];
#pragma unroll
                    for (int q = 0; q < R; ++q) {
                        if (pos[q] != NOROW && pos[q] > k0 + j) {
                            const double l = av[q][j] / pv;
                            av[q][j] = l;
#pragma unroll
                            for (int c = 0; c < LU_NB; ++c)
                                if (c > j) av[q][c] = fma(-l, rowP[par][c], av[q][c]);
                        }
                    }
                }
            }
#pragma unroll
            for (int q = 0; q < R; ++q) {
                if (pos[q] != NOROW) {
#pragma unroll
                    for (int c = 0; c < LU_NB; ++c)
                        if (c < nbe) N[(int64_t)pos[q] * ldn + k0 + c] = av[q][c];
                }
            }
            __syncthreads();
            if (tid < nbe) a.piv[tid] = s_piv[tid];
        }
        cluster.sync();
        // ---- (b) the panel's row swaps applied to the columns right of it (incl. the right-hand sides), fused with
        //      U12 = L11^-1 N12.  The 16 swaps are composed first: position tp[i] receives the old row ts[i]. ----
        {
            double (*L11)[LU_NB + 1] = Ps;
            if (tid < LU_NB * LU_NB) {
                const int i = tid / LU_NB, j = tid % LU_NB;
                L11[i][j] = (i < nbe && j < i) ? N[(int64_t)(k0 + i) * ldn + k0 + j] : 0.0;
            }
            if (tid == 0) {
                int np = nbe;
                for (int i = 0; i < nbe; ++i) { s_tp[i] = k0 + i; s_ts[i] = k0 + i; }
                for (int j = 0; j < nbe; ++j) {
                    const int p = a.piv[j];
                    int b = -1;
                    for (int i = 0; i < np; ++i)
                        if (s_tp[i] == p) b = i;
                    if (b < 0) { b = np++; s_tp[b] = p; s_ts[b] = p; }
                    const int t = s_ts[j]; s_ts[j] = s_ts[b]; s_ts[b] = t;
                }
                s_np = np;
            }
            __syncthreads();
            const int np = s_np;
            for (int cc = k0 + nbe + crank * LU_THREADS + tid; cc < W; cc += csize * LU_THREADS) {
                double vals[2 * LU_NB];
#pragma unroll
                for (int i = 0; i < 2 * LU_NB; ++i) vals[i] = (i < np) ? N[(int64_t)s_ts[i] * ldn + cc] : 0.0;
#pragma unroll
                for (int j = 0; j < LU_NB; ++j) {
                    double s = vals[j];
#pragma unroll
                    for (int i = 0; i < LU_NB; ++i)
                        if (i < j) s = fma(-L11[j][i], vals[i], s);
                    vals[j] = s;
                }
#pragma unroll
                for (int i = 0; i < 2 * LU_NB; ++i)
                    if (i < np) N[(int64_t)s_tp[i] * ldn + cc] = vals[i];
            }
        }
        cluster.sync();
        // ---- (c) trailing update N22 -= L21 U12 ----
        const int r0 = k0 + nbe;
        if (r0 < M) {
            const int rows = M - r0, cols = W - r0;
            const int ntr = (rows + LU_TR - 1) / LU_TR, ntc = (cols + LU_TC - 1) / LU_TC;
            for (int tile = crank; tile < ntr * ntc; tile += csize)
                lu_tile(N + (int64_t)r0 * ldn + r0, ldn, N + (int64_t)r0 * ldn + k0, ldn, N + (int64_t)k0 * ldn + r0, ldn,
                        rows, cols, nbe, -1.0, true, (tile / ntc) * LU_TR, (tile % ntc) * LU_TC, Ps, Qs);
        }
        cluster.sync();
    }

    // ---- back substitution U X = Y (Y = columns M .. M+D-1 of N), CTA 0 ----
    if (crank != 0) return;
    double (*Ub)[LU_NB + 1] = Ps;                   // U block, 16 x 16
    double (*Xb)[LU_TC] = Qs;                       // solved block rows, 16 x D   (D <= 64 < LU_TC)
    const int nblk = (M + LU_NB - 1) / LU_NB;
    for (int kb = nblk - 1; kb >= 0; --kb) {
        const int k0 = kb * LU_NB, nbe = min(LU_NB, M - k0);
        __syncthreads();
        if (tid < LU_NB * LU_NB) {
            const int i = tid / LU_NB, j = tid % LU_NB;
            Ub[i][j] = (i < nbe && j < nbe && j >= i) ? N[(int64_t)(k0 + i) * ldn + k0 + j] : (i == j ? 1.0 : 0.0);
        }
        for (int e = tid; e < LU_NB * D; e += LU_THREADS) {
            const int i = e / D, d = e % D;
            Xb[i][d] = (i < nbe) ? N[(int64_t)(k0 + i) * ldn + M + d] : 0.0;
        }
        __syncthreads();
        if (tid < D) {
            for (int i = nbe - 1; i >= 0; --i) {
                double s = Xb[i][tid];
                for (int j = i + 1; j < nbe; ++j) s = fma(-Ub[i][j], Xb[j][tid], s);
                Xb[i][tid] = s / Ub[i][i];
            }
        }
        __syncthreads();
        for (int e = tid; e < nbe * D; e += LU_THREADS) a.C[(int64_t)(k0 + e / D) * D + e % D] = Xb[e / D][e % D];
        for (int e = tid; e < k0 * D; e += LU_THREADS) {
            const int i = e / D, d = e % D;
            double s = N[(int64_t)i * ldn + M + d];
            const double* urow = N + (int64_t)i * ldn + k0;
#pragma unroll 4
            for (int j = 0; j < nbe; ++j) s = fma(-urow[j], Xb[j][d], s);
            N[(int64_t)i * ldn + M + d] = s;
        }
    }
}

static inline int lu_ldn(int M, int D) { return (M + D + 1) & ~1; }

size_t solve_normal_ws_bytes(int M, int D) {
    return sizeof(double) * 2 * (size_t)M * lu_ldn(M, D) + 256;
}

template <int R>
static int lu_launch_factor(const lu_args& a, cudaStream_t st) {
    static int cluster_size = 0;
    if (!cluster_size) {
        DVF_CUDA(cudaFuncSetAttribute(k_lu_factor<R>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
        for (int cs : {16, 8, 4, 2, 1}) {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(cs); cfg.blockDim = dim3(LU_THREADS);
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension;
            at[0].val.clusterDim.x = cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            int n = 0;
            if (cudaOccupancyMaxActiveClusters(&n, k_lu_factor<R>, &cfg) == cudaSuccess && n >= 1) { cluster_size = cs; break; }
            cudaGetLastError();
        }
        if (!cluster_size) { set_error("k_lu_factor cannot be scheduled"); return DVF_ERR_CUDA; }
        static const char* force = getenv("DVF_LU_CLUSTER");   // developer knob
        if (force && atoi(force) > 0 && atoi(force) <= cluster_size) cluster_size = atoi(force);
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(cluster_size); cfg.blockDim = dim3(LU_THREADS); cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = cluster_size; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    ++g_launch_count;
    DVF_CUDA(cudaLaunchKernelEx(&cfg, k_lu_factor<R>, a));
    return DVF_OK;
}

int solve_normal_run(const double* G, int64_t ldg, const double* Kc, int M, int D, double lam_sigma2,
                     const double* lam_sigma2_ptr, void* ws, double* C, int* info, const int* done, cudaStream_t st) {
    if (M > LU_MAX_M) {
        set_error("lstsq_method='drouin' (normal equations + LU) supports M <= %d control points; M = %d (use "
                  "lstsq_method='cholesky')", LU_MAX_M, M);
        return DVF_ERR_ARG;
    }
    const int ldn = lu_ldn(M, D);
    double* AB = (double*)ws;
    double* N = AB + (size_t)M * ldn;
    int* piv = (int*)(N + (size_t)M * ldn);
    lu_args a{G, ldg, Kc, M, D, lam_sigma2, lam_sigma2_ptr, AB, N, ldn, piv, C, info, done};
    const int W = M + D;
    int nb = (int)(((int64_t)M * W + 255) / 256);
    if (nb > 4 * num_sms()) nb = 4 * num_sms();
    k_lu_prep<<<nb, 256, 0, st>>>(a);
    DVF_LAUNCH_CHECK();
    k_lu_ata<<<((M + LU_TR - 1) / LU_TR) * ((W + LU_TC - 1) / LU_TC), LU_THREADS, 0, st>>>(a);
    DVF_LAUNCH_CHECK();
    if (M <= LU_THREADS) return lu_launch_factor<1>(a, st);
    if (M <= 2 * LU_THREADS) return lu_launch_factor<2>(a, st);
    return lu_launch_factor<4>(a, st);
}

}  // namespace dvf

extern "C" int dvf_solve_normal_workspace_bytes(int M, int D, size_t* out) {
    DVF_ARG(out && M >= 1 && D >= 1 && D <= dvf::kMaxD, "bad argument");
    *out = dvf::solve_normal_ws_bytes(M, D);
    return DVF_OK;
}

extern "C" int dvf_solve_normal(const double* G, int64_t ldg, const double* Kc, int M, int D, double lam_sigma2,
                                void* ws, size_t ws_bytes, double* C, int* info, void* stream) {
    DVF_ARG(G && Kc && ws && C && info, "NULL pointer");
    DVF_ARG(M >= 1 && D >= 1 && D <= dvf::kMaxD && ldg >= M + D, "bad shape M=%d D=%d ldg=%lld", M, D, (long long)ldg);
    if (ws_bytes < dvf::solve_normal_ws_bytes(M, D)) {
        dvf::set_error("normal-equations solver workspace too small");
        return DVF_ERR_WORKSPACE;
    }
    DVF_CUDA(cudaMemsetAsync(info, 0, sizeof(int), (cudaStream_t)stream));
    return dvf::solve_normal_run(G, ldg, Kc, M, D, lam_sigma2, nullptr, ws, C, info, nullptr, (cudaStream_t)stream);
}

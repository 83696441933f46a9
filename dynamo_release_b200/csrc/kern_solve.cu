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
constexpr int LU_TR = 64, LU_TC = 128;   // tile of the trailing update: 64 rows x 128 columns, 4 x 4 outputs per thread
constexpr int LU_ATR = 32, LU_ATC = 64;  // tile of N = lhs [lhs | rhs]: small, so that M = 500 already fills the GPU (144 CTAs)
constexpr int LU_PWMAX = 16;             // at most all warps of CTA 0 factor the panel (named barrier 1), R rows per thread
constexpr int LU_MAX_M = 4 * LU_PWMAX * 32;

struct lu_args {
    const double* G; int64_t ldg; const double* Kc; int M; int D;
    double lam_sigma2; const double* lam_sigma2_ptr;
    double* AB; double* N; int ldn; int* piv; double* C; int* info; const int* done;
};

// one TR x TC tile (512 threads, 16 x 32; TR / 16 x TC / 32 outputs per thread):
//   Cd[i][j] = (acc ? Cd : 0) + sign * sum_{c<K} P[i][c] Q[c][j]
template <int TR, int TC>
__device__ __forceinline__ void lu_tile(double* Cd, int ldc, const double* P, int ldp, const double* Q, int ldq, int rows,
                                        int cols, int K, double sign, bool acc, int i0, int j0,
                                        double (*Ps)[LU_NB + 1], double (*Qs)[TC]) {
    constexpr int RU = TR / 16, CV = TC / 32;
    const int tid = threadIdx.x;
    const int ty = tid >> 5, tx = tid & 31;      // rows ty + 16 u, columns tx + 32 v
    double s[RU][CV];
#pragma unroll
    for (int u = 0; u < RU; ++u)
#pragma unroll
        for (int v = 0; v < CV; ++v) s[u][v] = 0.0;
    for (int c0 = 0; c0 < K; c0 += LU_NB) {
        __syncthreads();
        for (int e = tid; e < TR * LU_NB; e += LU_THREADS) {
            const int r = e >> 4, c = e & 15;
            Ps[r][c] = (i0 + r < rows && c0 + c < K) ? P[(int64_t)(i0 + r) * ldp + c0 + c] : 0.0;
        }
        for (int e = tid; e < LU_NB * TC; e += LU_THREADS) {
            const int c = e / TC, j = e % TC;
            Qs[c][j] = (c0 + c < K && j0 + j < cols) ? Q[(int64_t)(c0 + c) * ldq + j0 + j] : 0.0;
        }
        __syncthreads();
#pragma unroll
        for (int c = 0; c < LU_NB; ++c) {
            double pv[RU], qv[CV];
#pragma unroll
            for (int u = 0; u < RU; ++u) pv[u] = Ps[ty + 16 * u][c];
#pragma unroll
            for (int v = 0; v < CV; ++v) qv[v] = Qs[c][tx + 32 * v];
#pragma unroll
            for (int u = 0; u < RU; ++u)
#pragma unroll
                for (int v = 0; v < CV; ++v) s[u][v] = fma(pv[u], qv[v], s[u][v]);
        }
    }
#pragma unroll
    for (int u = 0; u < RU; ++u)
#pragma unroll
        for (int v = 0; v < CV; ++v) {
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

// N = lhs [lhs | rhs]: one 32 x 64 tile per CTA
__global__ void __launch_bounds__(LU_THREADS) k_lu_ata(lu_args a) {
    if (a.done && *a.done) return;
    __shared__ double Ps[LU_ATR][LU_NB + 1];
    __shared__ double Qs[LU_NB][LU_ATC];
    const int W = a.M + a.D, ntc = (W + LU_ATC - 1) / LU_ATC;
    lu_tile<LU_ATR, LU_ATC>(a.N, a.ldn, a.AB, a.ldn, a.AB, a.ldn, a.M, W, a.M, 1.0, false, (blockIdx.x / ntc) * LU_ATR,
                            (blockIdx.x % ntc) * LU_ATC, Ps, Qs);
}

__device__ __forceinline__ unsigned long long lu_now() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// phase timers of CTA 0 (ns, accumulated over the panels), kept behind the pivot array in the workspace: developer
// aid for tools/solve_tune.py, a handful of timer reads per panel
#define LU_MARK(slot)                                                    \
    do {                                                                 \
        if (tid == 0 && crank == 0) {                                    \
            const unsigned long long now__ = lu_now();                   \
            tacc[slot] += now__ - tlast;                                 \
            tlast = now__;                                               \
        }                                                                \
    } while (0)

// 1 / x to within an ulp without the division's dependent chain of ~10 FP64 instructions and its slow-path branch:
// hardware seed (MUFU.RCP64H, ~20 bits, full double range) + two Newton steps.  The panel computes it SPECULATIVELY for
// every row's entry of the next column, so that it overlaps the arg-max reductions and the barrier and the winner's
// reciprocal is simply picked (LAPACK getf2 also scales by the reciprocal of the pivot).
__device__ __forceinline__ double lu_rcp(double x) {
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    double e = fma(-x, r, 1.0);
    r = fma(r, e, r);
    e = fma(-x, r, 1.0);
    r = fma(r, e, r);
    return r;
}

// shared state of the panel factorisation
struct lu_panel_smem {
    double rowP[2][LU_NB];
    double wval[2][LU_PWMAX], wsgn[2][LU_PWMAX];
    int wpos[2][LU_PWMAX], wown[2][LU_PWMAX];
    int s_piv[LU_NB];
    double U12s[LU_NB][LU_NB + 1];
#ifdef DVF_LU_FINE
    long long fine[8];
#endif
};
#ifdef DVF_LU_FINE
#define LU_FINE(slot) do { if (tid == 0) { const long long c__ = clock64(); sh.fine[slot] += c__ - fl; fl = c__; } } while (0)
#else
#define LU_FINE(slot) do { } while (0)
#endif

// Panel factorisation of the columns k0 .. k0 + 15 (rows k0 .. M - 1) by the first LU_PW warps of a CTA, R rows per thread
// in registers, named barrier 1.  kprev >= 0: the trailing update of the PREVIOUS panel (columns kprev .. kprev + 15) has
// not been applied to these columns yet and is done here while they are loaded (look-ahead: the other CTAs of the
// cluster meanwhile update the rest of the trailing matrix).
template <int R, int LU_PW>
__device__ __forceinline__ void lu_panel(const lu_args& a, int k0, int kprev, lu_panel_smem& sh) {
    constexpr int NOROW = 0x7fffffff;
    const int M = a.M, ldn = a.ldn;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nbe = min(LU_NB, M - k0);
    double* N = a.N;
#ifdef DVF_LU_FINE
    long long fl = clock64();
#endif
    constexpr int PT = LU_PW * 32;
    double av[R][LU_NB];
    int pos[R];
    if (kprev >= 0) {       // look-ahead: U12 block of the previous panel above this panel's columns
        for (int e = tid; e < LU_NB * LU_NB; e += PT) {
            const int jj = e / LU_NB, c = e % LU_NB;
            sh.U12s[jj][c] = (c < nbe) ? N[(int64_t)(kprev + jj) * ldn + k0 + c] : 0.0;
        }
        asm volatile("bar.sync 1, %0;" ::"n"(PT) : "memory");
    }
#pragma unroll
    for (int q = 0; q < R; ++q) {
        const int r = k0 + tid + PT * q;
        pos[q] = r < M ? r : NOROW;
#pragma unroll
        for (int c = 0; c < LU_NB; ++c) av[q][c] = (r < M && c < nbe) ? N[(int64_t)r * ldn + k0 + c] : 0.0;
        if (kprev >= 0 && r < M) {
            // the previous panel's trailing update of these 16 columns, in registers: N22 -= L21 U12 (same arithmetic as
            // lu_tile: the sum first, then one fma into the entry)
            double lrow[LU_NB];
#pragma unroll
            for (int jj = 0; jj < LU_NB; ++jj) lrow[jj] = N[(int64_t)r * ldn + kprev + jj];
#pragma unroll
            for (int c = 0; c < LU_NB; ++c) {
                double sacc = 0.0;
#pragma unroll
                for (int jj = 0; jj < LU_NB; ++jj) sacc = fma(lrow[jj], sh.U12s[jj][c], sacc);
                av[q][c] = fma(-1.0, sacc, av[q][c]);
            }
        }
    }
#pragma unroll
    for (int j = 0; j < LU_NB; ++j) {
        if (j < nbe) {                     // uniform over the panel warps
            const int par = j & 1;
            LU_FINE(0);      // previous column's elimination (+ the panel load for j = 0)
            // arg-max of |column j| over the rows not yet eliminated.  |v| >= 0, so the bit pattern orders like the
            // value: two 32-bit warp reductions (REDUX.MAX: high word, then low word among the lanes that hold the
            // largest high word) instead of five 64-bit shuffle steps; ties in magnitude go to the smallest logical row
            // (LAPACK's idamax takes the first)
            double bv = -1.0, bs = 0.0;
            int bp = NOROW, bo = 0;
#pragma unroll
            for (int q = 0; q < R; ++q) {
                const double v = fabs(av[q][j]);
                if (pos[q] != NOROW && pos[q] >= k0 + j && (v > bv || (v == bv && pos[q] < bp))) {
                    bv = v; bs = av[q][j]; bp = pos[q]; bo = tid * R + q;
                }
            }
            bs = lu_rcp(bs);       // speculative: the reciprocal of this thread's candidate (used only if it wins)
            {
                const bool cand = bp != NOROW && bv == bv;           // has a row, not NaN
                const unsigned hi = cand ? (unsigned)__double2hiint(bv) : 0u, lo = (unsigned)__double2loint(bv);
                const unsigned mh = __reduce_max_sync(0xffffffffu, hi);
                const bool top = cand && hi == mh;
                const unsigned ml = __reduce_max_sync(0xffffffffu, top ? lo : 0u);
                const bool win = top && lo == ml;
                const unsigned mp = __reduce_min_sync(0xffffffffu, win ? (unsigned)bp : 0xffffffffu);
                const unsigned who = __ballot_sync(0xffffffffu, win && (unsigned)bp == mp);
                if (who == 0u) {           // no candidate in this warp (all its rows are eliminated / NaN)
                    if (lane == 0) { sh.wval[par][warp] = -1.0; sh.wsgn[par][warp] = 0.0; sh.wpos[par][warp] = NOROW; sh.wown[par][warp] = 0; }
                } else if (lane == __ffs(who) - 1) {
                    sh.wval[par][warp] = bv; sh.wsgn[par][warp] = bs; sh.wpos[par][warp] = bp; sh.wown[par][warp] = bo;
                }
            }
            LU_FINE(1);      // arg-max inside the warp
            asm volatile("bar.sync 1, %0;" ::"n"(PT) : "memory");
            LU_FINE(2);      // barrier 1
            {   // the LU_PW warp candidates: lane w < LU_PW takes candidate w, the same three reductions, then the winner's
                // fields are broadcast with shuffles (every warp does this redundantly; no second barrier for it)
                const int w = lane < LU_PW ? lane : 0;
                const double cv = sh.wval[par][w];
                const double cs_ = sh.wsgn[par][w];
                const int cp = sh.wpos[par][w], co = sh.wown[par][w];
                const bool cand = lane < LU_PW && cp != NOROW && cv >= 0.0;
                const unsigned hi = cand ? (unsigned)__double2hiint(cv) : 0u, lo = (unsigned)__double2loint(cv);
                const unsigned mh = __reduce_max_sync(0xffffffffu, hi);
                const bool top = cand && hi == mh;
                const unsigned ml = __reduce_max_sync(0xffffffffu, top ? lo : 0u);
                const bool win = top && lo == ml;
                const unsigned mp = __reduce_min_sync(0xffffffffu, win ? (unsigned)cp : 0xffffffffu);
                const unsigned who = __ballot_sync(0xffffffffu, win && (unsigned)cp == mp);
                const int src = who ? __ffs(who) - 1 : 0;
                bv = who ? __shfl_sync(0xffffffffu, cv, src) : -1.0;
                bs = __shfl_sync(0xffffffffu, cs_, src);
                bp = who ? __shfl_sync(0xffffffffu, cp, src) : NOROW;
                bo = __shfl_sync(0xffffffffu, co, src);
            }
            if (tid == 0) {
                sh.s_piv[j] = bp;
                if (!(bv > 0.0)) atomicCAS(a.info, 0, k0 + j + 1);      // exactly singular (or NaN)
            }
#pragma unroll
            for (int q = 0; q < R; ++q) {
                if (tid * R + q == bo) {             // this thread holds the pivot row: it becomes row k0 + j
#pragma unroll
                    for (int c = 0; c < LU_NB; ++c) sh.rowP[par][c] = av[q][c];
                    pos[q] = k0 + j;
                } else if (pos[q] == k0 + j) {
                    pos[q] = bp;                     // the row that sat at k0 + j goes where the pivot row was
                }
            }
            const double rpv = bs;     // the winner's reciprocal, computed before the reductions
            LU_FINE(3);      // scan of the warp candidates, reciprocal issued, pivot row to shared memory
            asm volatile("bar.sync 1, %0;" ::"n"(PT) : "memory");
            LU_FINE(4);      // barrier 2
#pragma unroll
            for (int q = 0; q < R; ++q) {
                if (pos[q] != NOROW && pos[q] > k0 + j) {
                    const double l = av[q][j] * rpv;
                    av[q][j] = l;
#pragma unroll
                    for (int c = 0; c < LU_NB; ++c)
                        if (c > j) av[q][c] = fma(-l, sh.rowP[par][c], av[q][c]);
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
    asm volatile("bar.sync 1, %0;" ::"n"(PT) : "memory");
    if (tid < nbe) a.piv[(k0 / LU_NB & 1) * LU_NB + tid] = sh.s_piv[tid];
}

template <int R, int LU_PW>
__global__ void __launch_bounds__(LU_THREADS) k_lu_factor(lu_args a) {
    cg::cluster_group cluster = cg::this_cluster();
    if (a.done && *a.done) return;               // uniform over the cluster
    unsigned long long tacc[8] = {0, 0, 0, 0, 0, 0, 0, 0}, tlast = lu_now();
    __shared__ double Ps[LU_TR][LU_NB + 1];
    __shared__ double Qs[LU_NB][LU_TC];
    __shared__ lu_panel_smem sh;
#ifdef DVF_LU_FINE
    if (threadIdx.x < 8) sh.fine[threadIdx.x] = 0;
    __syncthreads();
#endif
    __shared__ int s_tp[2 * LU_NB], s_ts[2 * LU_NB], s_np;
    const int M = a.M, D = a.D, W = a.M + a.D, ldn = a.ldn;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int crank = (int)cluster.block_rank(), csize = (int)cluster.num_blocks();
    double* N = a.N;
    // panel 0; every later panel is factored by CTA 0 during the trailing update of the panel before it (look-ahead)
    if (crank == 0 && warp < LU_PW) lu_panel<R, LU_PW>(a, 0, -1, sh);
    LU_MARK(0);
    cluster.sync();
    LU_MARK(1);
    for (int k0 = 0; k0 < M; k0 += LU_NB) {
        const int nbe = min(LU_NB, M - k0);
        // ---- (b) the panel's row swaps applied to the columns right of it (incl. the right-hand sides), fused with
        //      U12 = L11^-1 N12.  The 16 swaps are composed first: position tp[i] receives the old row ts[i]. ----
        {
            double (*L11)[LU_NB + 1] = Ps;
            if (tid < LU_NB * LU_NB) {
                const int i = tid / LU_NB, j = tid % LU_NB;
                L11[i][j] = (i < nbe && j < i) ? N[(int64_t)(k0 + i) * ldn + k0 + j] : 0.0;
            }
            if (warp == 0) {       // lane i holds touched position tp and the old row ts that ends up there
                int np = nbe;
                int tp = lane < nbe ? k0 + lane : -1, ts = tp;
                const int mypiv = lane < nbe ? a.piv[(k0 / LU_NB & 1) * LU_NB + lane] : 0;
                for (int j = 0; j < nbe; ++j) {
                    const int p = __shfl_sync(0xffffffffu, mypiv, j);
                    const unsigned hit = __ballot_sync(0xffffffffu, lane < np && tp == p);
                    int b;
                    if (hit) b = __ffs(hit) - 1;
                    else { b = np++; if (lane == b) { tp = p; ts = p; } }
                    const int tj = __shfl_sync(0xffffffffu, ts, j), tb = __shfl_sync(0xffffffffu, ts, b);
                    if (lane == j) ts = tb;
                    if (lane == b) ts = tj;
                }
                s_tp[lane] = tp; s_ts[lane] = ts;
                if (lane == 0) s_np = np;
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
        LU_MARK(2);
        cluster.sync();
        LU_MARK(3);
        // ---- (c) trailing update N22 -= L21 U12.  CTA 0 applies it to the NEXT panel's 16 columns only and factors that
        //      panel right away; the other CTAs take the rest of the trailing matrix meanwhile ----
        const int r0 = k0 + nbe;
        if (r0 < M) {
            const int nb2 = min(LU_NB, M - r0), c0 = r0 + nb2;
            if (crank == 0 && warp < LU_PW) lu_panel<R, LU_PW>(a, r0, k0, sh);
            if (csize == 1 || crank != 0) {
                const int workers = csize == 1 ? 1 : csize - 1, me = csize == 1 ? 0 : crank - 1;
                const int rows = M - r0, cols = W - c0;
                const int ntr = (rows + LU_TR - 1) / LU_TR, ntc = (cols + LU_TC - 1) / LU_TC;
                for (int tile = me; tile < ntr * ntc; tile += workers)
                    lu_tile<LU_TR, LU_TC>(N + (int64_t)r0 * ldn + c0, ldn, N + (int64_t)r0 * ldn + k0, ldn,
                                          N + (int64_t)k0 * ldn + c0, ldn, rows, cols, nbe, -1.0, true,
                                          (tile / ntc) * LU_TR, (tile % ntc) * LU_TC, Ps, Qs);
            }
        }
        LU_MARK(4);
        cluster.sync();
        LU_MARK(5);
    }

    // ---- back substitution U X = Y (Y = columns M .. M+D-1 of N).  The right-hand sides are independent: CTA c takes the
    // columns c, c + csize, ... (no exchange between CTAs).  Thread t keeps the entries of its rows t, t + 512, ... in
    // registers.  Per block of 16 rows (bottom up): every thread whose row lies in or above the block loads its 16
    // entries of U; the warp that owns the block's rows solves the 16 x 16 triangle with shuffles (v1 went through
    // global memory and one thread per column: 585 us of 1.9 ms at M = 500); the solved block goes through shared
    // memory to the rows above. ----
    {
        constexpr int RB = (LU_MAX_M + LU_THREADS - 1) / LU_THREADS;      // rows per thread, worst case
        constexpr int NCMAX = 4;                                          // columns per pass
        double (*Xs)[LU_NB][NCMAX] = reinterpret_cast<double (*)[LU_NB][NCMAX]>(&Qs[0][0]);   // [2][16][NCMAX]
        const int nblk = (M + LU_NB - 1) / LU_NB;
        for (int cbase = crank; cbase < D; cbase += csize * NCMAX) {
            // this CTA's columns of the pass: cbase, cbase + csize, ... (up to NCMAX)
            int ncol = 0;
            int col[NCMAX];
#pragma unroll
            for (int c = 0; c < NCMAX; ++c) {
                col[c] = cbase + c * csize;
                if (col[c] < D) ncol = c + 1;
            }
            double y[RB][NCMAX];
#pragma unroll
            for (int q = 0; q < RB; ++q)
#pragma unroll
                for (int c = 0; c < NCMAX; ++c) {
                    const int r = tid + LU_THREADS * q;
                    y[q][c] = (r < M && c < ncol) ? N[(int64_t)r * ldn + M + col[c]] : 0.0;
                }
            for (int kb = nblk - 1; kb >= 0; --kb) {
                const int k0 = kb * LU_NB, nbe = min(LU_NB, M - k0), par = kb & 1;
                const int qb = k0 / LU_THREADS, t0 = k0 % LU_THREADS;        // the block's rows: threads t0 .. t0 + nbe - 1, slot qb
                if ((tid >> 5) == (t0 >> 5)) {
                    // the warp that holds the block's rows: triangular solve with shuffles (lanes l0 .. l0 + nbe - 1)
                    const int l0 = t0 & 31, li = lane - l0;
                    const bool in_blk = li >= 0 && li < nbe;
                    double useg[LU_NB], yb[NCMAX];
#pragma unroll
                    for (int jj = 0; jj < LU_NB; ++jj)
                        useg[jj] = (in_blk && jj < nbe) ? N[(int64_t)(k0 + li) * ldn + k0 + jj] : 1.0;
#pragma unroll
                    for (int c = 0; c < NCMAX; ++c) {
                        yb[c] = 0.0;
#pragma unroll
                        for (int q = 0; q < RB; ++q)
                            if (q == qb) yb[c] = y[q][c];
                    }
                    // reciprocal of the own diagonal entry, once per row and off the 16-step chain (optimised BLAS trsm
                    // kernels, OpenBLAS among them, keep the inverted diagonal too)
                    double rdiag = 1.0;
#pragma unroll
                    for (int jj = 0; jj < LU_NB; ++jj)
                        if (jj == li) rdiag = lu_rcp(useg[jj]);
#pragma unroll
                    for (int i = LU_NB - 1; i >= 0; --i) {
                        if (i < nbe) {                          // uniform over the warp
#pragma unroll
                            for (int c = 0; c < NCMAX; ++c) {
                                const double cand = yb[c] * rdiag;                    // meaningful on the lane of row k0 + i
                                const double xi = __shfl_sync(0xffffffffu, cand, l0 + i);
                                if (li == i) yb[c] = xi;
                                else if (in_blk && li < i) yb[c] = fma(-useg[i], xi, yb[c]);
                            }
                        }
                    }
                    if (in_blk) {
#pragma unroll
                        for (int c = 0; c < NCMAX; ++c) {
                            Xs[par][li][c] = yb[c];
                            if (c < ncol) a.C[(int64_t)(k0 + li) * D + col[c]] = yb[c];
                        }
                    }
                }
                __syncthreads();
                // rows above the block: y_r -= U[r][block] x_block
#pragma unroll
                for (int q = 0; q < RB; ++q) {
                    const int r = tid + LU_THREADS * q;
                    if (r < k0) {
                        double useg[LU_NB];
#pragma unroll
                        for (int jj = 0; jj < LU_NB; ++jj) useg[jj] = (jj < nbe) ? N[(int64_t)r * ldn + k0 + jj] : 0.0;
#pragma unroll
                        for (int jj = 0; jj < LU_NB; ++jj)
#pragma unroll
                            for (int c = 0; c < NCMAX; ++c) y[q][c] = fma(-useg[jj], Xs[par][jj][c], y[q][c]);
                    }
                }
            }
            __syncthreads();
        }
    }
    __syncthreads();
    LU_MARK(6);
    if (tid == 0 && crank == 0) {
        unsigned long long* dbg = reinterpret_cast<unsigned long long*>(a.piv + 32);
        for (int i = 0; i < 8; ++i) dbg[i] = tacc[i];
#ifdef DVF_LU_FINE
        for (int i = 0; i < 8; ++i) dbg[8 + i] = (unsigned long long)sh.fine[i];
#endif
    }
}

static inline int lu_ldn(int M, int D) { return (M + D + 1) & ~1; }

size_t solve_normal_ws_bytes(int M, int D) {
    return sizeof(double) * 2 * (size_t)M * lu_ldn(M, D) + 256;   // + pivots (16 ints) + 8 phase timers at byte 128
}

template <int R, int PW>
static int lu_launch_factor(const lu_args& a, cudaStream_t st) {
    static int cluster_size = 0;
    if (!cluster_size) {
        DVF_CUDA(cudaFuncSetAttribute(k_lu_factor<R, PW>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
        for (int cs : {16, 8, 4, 2, 1}) {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(cs); cfg.blockDim = dim3(LU_THREADS);
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension;
            at[0].val.clusterDim.x = cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            int n = 0;
            if (cudaOccupancyMaxActiveClusters(&n, k_lu_factor<R, PW>, &cfg) == cudaSuccess && n >= 1) { cluster_size = cs; break; }
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
    DVF_CUDA(cudaLaunchKernelEx(&cfg, k_lu_factor<R, PW>, a));
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
    k_lu_ata<<<((M + LU_ATR - 1) / LU_ATR) * ((W + LU_ATC - 1) / LU_ATC), LU_THREADS, 0, st>>>(a);
    DVF_LAUNCH_CHECK();
    // rows per thread x warps of the panel: fewer warps make the two barriers per column cheaper, more rows per thread
    // cost registers (4 x 16 doubles spill)
    static const char* pw = getenv("DVF_LU_PW");           // developer knob (tools/solve_tune.py)
    if (pw && atoi(pw) == 16 && M <= 512) return lu_launch_factor<1, 16>(a, st);
    if (pw && atoi(pw) == 4 && M <= 512) return lu_launch_factor<4, 4>(a, st);
    if (M <= 256) return lu_launch_factor<1, 8>(a, st);
    if (M <= 512) return lu_launch_factor<2, 8>(a, st);
    if (M <= 1024) return lu_launch_factor<2, 16>(a, st);
    return lu_launch_factor<4, 16>(a, st);
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

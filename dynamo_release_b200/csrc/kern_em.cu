// K2: E-step (posterior P, energy terms) and the V = U C / sigma2 update of the SparseVFC EM loop.
// Reference: scVectorField.py:197-227 (get_P), :484-495.  All reductions are two-stage with a fixed block
// count and a fixed summation order, so results are bitwise reproducible run to run.
#include "em.h"

namespace dvf {

// ---- E-step, pass 1: t1 = exp(-r2 / (2 sigma2)); per-block min over the non-zero t1 and zero count -------
__global__ void __launch_bounds__(ES_THREADS) k_estep1(const double* __restrict__ r2, int64_t N,
                                                       const double* __restrict__ sigma2_ptr, double sigma2,
                                                       double* __restrict__ P, double* __restrict__ part,
                                                       const int* __restrict__ done) {
    if (done && *done) return;
    __shared__ double sh[32];
    const double s2 = sigma2_ptr ? *sigma2_ptr : sigma2;
    const double inf = __longlong_as_double(0x7ff0000000000000LL);
    double mn = inf, nz = 0.0;
    for (int64_t i = (int64_t)blockIdx.x * ES_THREADS + threadIdx.x; i < N; i += (int64_t)gridDim.x * ES_THREADS) {
        const double t1 = exp(-r2[i] / (2.0 * s2));
        P[i] = t1;
        if (t1 == 0.0) nz += 1.0; else mn = fmin(mn, t1);
    }
    mn = block_min(mn, sh);
    nz = block_sum(nz, sh);
    if (threadIdx.x == 0) { part[2 * blockIdx.x] = mn; part[2 * blockIdx.x + 1] = nz; }
}

// minbuf[0] = min over blocks, minbuf[1] = number of exact zeros (one block)
__global__ void __launch_bounds__(256) k_estep_min(const double* __restrict__ part, int nblk, double* __restrict__ minbuf,
                                                   const int* __restrict__ done) {
    if (done && *done) return;
    __shared__ double sh[32];
    const double inf = __longlong_as_double(0x7ff0000000000000LL);
    double mn = inf, nz = 0.0;
    for (int b = threadIdx.x; b < nblk; b += 256) { mn = fmin(mn, part[2 * b]); nz += part[2 * b + 1]; }
    mn = block_min(mn, sh);
    nz = block_sum(nz, sh);
    if (threadIdx.x == 0) { minbuf[0] = mn; minbuf[1] = nz; }
}

// ---- E-step, pass 2: P = t1 / (t1 + t2), zeros of t1 replaced by the global min non-zero value ---------
// part[4b..4b+3] = { sum P r2, sum P, sum max(P, minP), #(max(P, minP) > theta) } of block b
__global__ void __launch_bounds__(ES_THREADS) k_estep2(const double* __restrict__ r2, int64_t N, int D,
                                                       const em_ctl* __restrict__ ctl, double sigma2, double gamma,
                                                       double a, double minP, double theta,
                                                       const double* __restrict__ minbuf, double* __restrict__ P,
                                                       double* __restrict__ part, const int* __restrict__ done) {
    if (done && *done) return;
    __shared__ double sh[32];
    const double s2 = ctl ? ctl->sigma2 : sigma2;
    const double gm = ctl ? ctl->gamma : gamma;
    const double t2 = pow(2.0 * 3.141592653589793 * s2, 0.5 * D) * (1.0 - gm) / (gm * a);
    const double tmin = minbuf[0];
    double s_pr2 = 0.0, s_p = 0.0, s_pc = 0.0, s_nc = 0.0;
    for (int64_t i = (int64_t)blockIdx.x * ES_THREADS + threadIdx.x; i < N; i += (int64_t)gridDim.x * ES_THREADS) {
        double t1 = P[i];
        if (t1 == 0.0) t1 = tmin;
        const double p = t1 / (t1 + t2);
        s_pr2 = fma(p, r2[i], s_pr2);
        s_p += p;
        const double pc = fmax(p, minP);
        P[i] = pc;   // the M-step and everything after it use the clamped posterior (scVectorField.py:468)
        s_pc += pc;
        if (pc > theta) s_nc += 1.0;
    }
    s_pr2 = block_sum(s_pr2, sh);
    s_p = block_sum(s_p, sh);
    s_pc = block_sum(s_pc, sh);
    s_nc = block_sum(s_nc, sh);
    if (threadIdx.x == 0) {
        part[4 * blockIdx.x + 0] = s_pr2;
        part[4 * blockIdx.x + 1] = s_p;
        part[4 * blockIdx.x + 2] = s_pc;
        part[4 * blockIdx.x + 3] = s_nc;
    }
}

// out[v] = sum_b part[b * nvals + v]   (one block, fixed order)
__global__ void __launch_bounds__(256) k_sum_partials(const double* __restrict__ part, int nblk, int nvals,
                                                      double* __restrict__ out, const int* __restrict__ done) {
    if (done && *done) return;
    __shared__ double sh[32];
    for (int v = 0; v < nvals; ++v) {
        double s = 0.0;
        for (int b = threadIdx.x; b < nblk; b += 256) s += part[(int64_t)b * nvals + v];
        s = block_sum(s, sh);
        if (threadIdx.x == 0) out[v] = s;
    }
}

// ---- tr(C^T Kc C) partials (scVectorField.py:454) ----------------------------------------------------
// Block b takes the rows i = b, b + gridDim.x, ...; for a row the 256 threads are 8 groups x 32 dimensions: group g sums
// Kc[i][j] C[j][d] over j = g, g + 8, ... (Kc[i][j] is a broadcast read, C[j][:] is read coalesced), the 8 partial sums
// meet in shared memory in a fixed order.  (First version: one thread per (i, d) walking all M control points, 94 us per
// EM iteration at M = 500 -- pure load latency.)
__global__ void __launch_bounds__(256) k_trace(const double* __restrict__ C, const double* __restrict__ Kc, int M, int D,
                                               double* __restrict__ part, const int* __restrict__ done) {
    if (done && *done) return;
    __shared__ double red[8][33];
    __shared__ double sh[32];
    const int d = threadIdx.x & 31, jg = threadIdx.x >> 5;
    double s = 0.0;
    for (int i = blockIdx.x; i < M; i += gridDim.x) {
        const double* krow = Kc + (int64_t)i * M;
        for (int d0 = 0; d0 < D; d0 += 32) {
            const int dd = d0 + d;
            double acc = 0.0;
            if (dd < D)
                for (int j = jg; j < M; j += 8) acc = fma(krow[j], C[(int64_t)j * D + dd], acc);
            red[jg][d] = acc;
            __syncthreads();
            if (jg == 0 && dd < D) {
                double kc = 0.0;
#pragma unroll
                for (int g = 0; g < 8; ++g) kc += red[g][d];
                s = fma(C[(int64_t)i * D + dd], kc, s);
            }
            __syncthreads();
        }
    }
    s = block_sum(s, sh);
    if (threadIdx.x == 0) part[blockIdx.x] = s;
}

// ---- Cp[Mp][DP] = zero-padded copy of C (16-byte aligned rows for cp.async) ---------------------------
__global__ void k_pad_c(const double* __restrict__ C, int M, int D, double* __restrict__ Cp, int Mp, int DP,
                        const int* __restrict__ done) {
    if (done && *done) return;
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < Mp * DP; e += gridDim.x * blockDim.x) {
        const int m = e / DP, d = e % DP;
        Cp[e] = (m < M && d < D) ? C[(int64_t)m * D + d] : 0.0;
    }
}

// ---- V = U C on FP64 tensor cores; r2 = |Y - V|^2; partial sums of max(P, minP) r2 ---------------------
// CTA = 128 cells (16 per warp), contraction over ctrl points in chunks of 32 through a cp.async ring.
template <int NT>
struct uv_cfg {
    static constexpr int DP = NT * 8, KC = 32, STAGES = 2;   // 92 KB per CTA: two CTAs (16 warps) per SM
    static constexpr int LDU = KC + 4;                             // == 4 (mod 16)
    static constexpr int LDC = ((DP + 15) / 16) * 16 + 4;          // == 4 (mod 16)
    static constexpr size_t SMEM = sizeof(double) * (size_t)STAGES * (UV_CELLS * LDU + KC * LDC);
};

template <int NT>
__global__ void __launch_bounds__(256, 2) k_update_v(const double* __restrict__ UY, const double* __restrict__ Cp,
                                                  const double* __restrict__ P, double minP, int64_t N, int D, int M,
                                                  int W, double* __restrict__ V, double* __restrict__ r2,
                                                  double* __restrict__ part, const int* __restrict__ done) {
    using cfg = uv_cfg<NT>;
    constexpr int DP = cfg::DP, KC = cfg::KC, STAGES = cfg::STAGES, LDU = cfg::LDU, LDC = cfg::LDC;
    if (done && *done) return;
    extern __shared__ double smem[];
    __shared__ double sh[32];
    double* Us = smem;                               // [STAGES][128][LDU]
    double* Cs = Us + STAGES * UV_CELLS * LDU;       // [STAGES][KC][LDC]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int64_t cell0 = (int64_t)blockIdx.x * UV_CELLS;
    const int n_steps = (M + KC - 1) / KC;

    auto load_stage = [&](int step, int stage) {
        const int m0 = step * KC;
        double* u = Us + stage * UV_CELLS * LDU;
        double* c = Cs + stage * KC * LDC;
        for (int e = threadIdx.x; e < UV_CELLS * (KC / 2); e += 256) {
            const int r = e / (KC / 2), ch = e % (KC / 2);
            const int64_t row = cell0 + r;
            const bool ok = row < N;   // columns m0.. < W always (W is a multiple of 64 >= M); cols >= M hold Y/0 and
                                       // meet zero rows of Cp
            cp_async16_zfill(u + r * LDU + 2 * ch, UY + (ok ? row : 0) * (int64_t)W + m0 + 2 * ch, ok);
        }
        for (int e = threadIdx.x; e < KC * (DP / 2); e += 256) {
            const int r = e / (DP / 2), ch = e % (DP / 2);
            cp_async16(c + r * LDC + 2 * ch, Cp + (int64_t)(m0 + r) * DP + 2 * ch);
        }
    };

    double acc[2][NT][2];
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < NT; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

#pragma unroll
    for (int st = 0; st < STAGES - 1; ++st) {
        if (st < n_steps) load_stage(st, st);
        cp_async_commit();
    }
    for (int step = 0; step < n_steps; ++step) {
        cp_async_wait<STAGES - 2>();
        __syncthreads();
        const int nxt = step + STAGES - 1;
        if (nxt < n_steps) load_stage(nxt, nxt % STAGES);
        cp_async_commit();
        const int stage = step % STAGES;
        const double* u = Us + stage * UV_CELLS * LDU + (warp * 16 + g) * LDU + t;
        const double* c = Cs + stage * KC * LDC + t * LDC + g;
#pragma unroll
        for (int k4 = 0; k4 < KC / 4; ++k4) {
            const double a0 = u[4 * k4], a1 = u[8 * LDU + 4 * k4];
            double bf[NT];
#pragma unroll
            for (int j = 0; j < NT; ++j) bf[j] = c[4 * k4 * LDC + 8 * j];
#pragma unroll
            for (int j = 0; j < NT; ++j) {
                dmma(acc[0][j][0], acc[0][j][1], a0, bf[j]);
                dmma(acc[1][j][0], acc[1][j][1], a1, bf[j]);
            }
        }
    }
    cp_async_wait<0>();

    // epilogue: lane owns V[cell][8j + 2t], V[cell][8j + 2t + 1] for its two cells
    double wsum = 0.0;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int64_t cell = cell0 + warp * 16 + i * 8 + g;
        const bool live = cell < N;
        double rr = 0.0;
        if (live) {
            const double* y = UY + cell * (int64_t)W + M;
#pragma unroll
            for (int j = 0; j < NT; ++j) {
                const int d = 8 * j + 2 * t;
                if (d < D) {
                    const double e0 = y[d] - acc[i][j][0];
                    rr = fma(e0, e0, rr);
                    if (V) V[cell * D + d] = acc[i][j][0];
                }
                if (d + 1 < D) {
                    const double e1 = y[d + 1] - acc[i][j][1];
                    rr = fma(e1, e1, rr);
                    if (V) V[cell * D + d + 1] = acc[i][j][1];
                }
            }
        }
        rr += __shfl_xor_sync(0xffffffffu, rr, 1);
        rr += __shfl_xor_sync(0xffffffffu, rr, 2);
        if (live && t == 0) {
            r2[cell] = rr;
            wsum = fma(fmax(P[cell], minP), rr, wsum);
        }
    }
    wsum = block_sum(wsum, sh);
    if (threadIdx.x == 0) part[blockIdx.x] = wsum;
}

template <int NT>
static int launch_update_v(const double* UY, const double* Cp, const double* P, double minP, int64_t N, int D, int M,
                           double* V, double* r2, double* part, const int* done, cudaStream_t st) {
    using cfg = uv_cfg<NT>;
    static bool attr_set = false;
    if (!attr_set) {
        DVF_CUDA(cudaFuncSetAttribute(k_update_v<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg::SMEM));
        attr_set = true;
    }
    k_update_v<NT><<<(unsigned)updatev_blocks(N), 256, cfg::SMEM, st>>>(UY, Cp, P, minP, N, D, M,
                                                                      (int)dvf_fit_row_width(M, D), V, r2, part, done);
    DVF_LAUNCH_CHECK();
    return DVF_OK;
}

int update_v_run(const double* UY, const double* Cp, const double* P, double minP, int64_t N, int D, int M, double* V,
                 double* r2, double* part, const int* done, cudaStream_t st) {
    switch (pad8(D) / 8) {
        case 1: return launch_update_v<1>(UY, Cp, P, minP, N, D, M, V, r2, part, done, st);
        case 2: return launch_update_v<2>(UY, Cp, P, minP, N, D, M, V, r2, part, done, st);
        case 3: return launch_update_v<3>(UY, Cp, P, minP, N, D, M, V, r2, part, done, st);
        case 4: return launch_update_v<4>(UY, Cp, P, minP, N, D, M, V, r2, part, done, st);
        case 5: return launch_update_v<5>(UY, Cp, P, minP, N, D, M, V, r2, part, done, st);
        case 6: return launch_update_v<6>(UY, Cp, P, minP, N, D, M, V, r2, part, done, st);
        case 7: return launch_update_v<7>(UY, Cp, P, minP, N, D, M, V, r2, part, done, st);
        default: return launch_update_v<8>(UY, Cp, P, minP, N, D, M, V, r2, part, done, st);
    }
}

int estep1_run(const double* r2, int64_t N, const double* sigma2_ptr, double sigma2, double* P, double* part,
               const int* done, cudaStream_t st) {
    k_estep1<<<estep_blocks(N), ES_THREADS, 0, st>>>(r2, N, sigma2_ptr, sigma2, P, part, done);
    DVF_LAUNCH_CHECK();
    return DVF_OK;
}
int estep_min_run(const double* part, int nblk, double* minbuf, const int* done, cudaStream_t st) {
    k_estep_min<<<1, 256, 0, st>>>(part, nblk, minbuf, done);
    DVF_LAUNCH_CHECK();
    return DVF_OK;
}
int estep2_run(const double* r2, int64_t N, int D, const em_ctl* ctl, double sigma2, double gamma, double a, double minP,
               double theta, const double* minbuf, double* P, double* part, const int* done, cudaStream_t st) {
    k_estep2<<<estep_blocks(N), ES_THREADS, 0, st>>>(r2, N, D, ctl, sigma2, gamma, a, minP, theta, minbuf, P, part, done);
    DVF_LAUNCH_CHECK();
    return DVF_OK;
}
int sum_partials_run(const double* part, int nblk, int nvals, double* out, const int* done, cudaStream_t st) {
    k_sum_partials<<<1, 256, 0, st>>>(part, nblk, nvals, out, done);
    DVF_LAUNCH_CHECK();
    return DVF_OK;
}
int pad_c_run(const double* C, int M, int D, double* Cp, int Mp, int DP, const int* done, cudaStream_t st) {
    k_pad_c<<<(Mp * DP + 255) / 256, 256, 0, st>>>(C, M, D, Cp, Mp, DP, done);
    DVF_LAUNCH_CHECK();
    return DVF_OK;
}
int trace_run(const double* C, const double* Kc, int M, int D, double* part, const int* done, cudaStream_t st) {
    k_trace<<<TR_BLOCKS, 256, 0, st>>>(C, Kc, M, D, part, done);
    DVF_LAUNCH_CHECK();
    return DVF_OK;
}

}  // namespace dvf

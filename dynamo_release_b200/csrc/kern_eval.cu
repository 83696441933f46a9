// K1 (Gaussian kernel matrix / fused [U | Y | 0] build), K5 (field evaluation), K7 (fused divergence /
// acceleration / curvature).  FP64 throughout; control points and coefficients are staged in shared
// memory in chunks and broadcast to the threads, cells stream from HBM.
#include "common.cuh"
#include "tma.cuh"

namespace dvf {

// ------------------------------------------------------------------------------------------------
// K1: 64 cells x 64 ctrl points per CTA pass, 4 x 4 register tile per thread, runtime D in chunks of 32.
// Direct form |x - c|^2 = sum_d (x_d - c_d)^2 as scipy cdist does (reference utils.py:195) -- no
// |x|^2 + |c|^2 - 2 x.c cancellation.
// ------------------------------------------------------------------------------------------------
constexpr int K1_T = 64, K1_DC = 32;

__global__ void __launch_bounds__(256) k_kernel_matrix(const double* __restrict__ x, int64_t n, int D,
                                                       const double* __restrict__ Xc, int M, double beta,
                                                       double* __restrict__ K, int64_t ldk,
                                                       const double* __restrict__ Y, int ycols, int padcols) {
    __shared__ double xs[K1_DC][K1_T + 4];
    __shared__ double cs[K1_DC][K1_T + 4];
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const int64_t n0 = (int64_t)blockIdx.x * K1_T;
    if (D <= K1_DC) {
        // one chunk of dimensions: the cell tile is staged once, and the next control-point tile is fetched into
        // registers while the current one is being used
        for (int e = threadIdx.x; e < K1_T * K1_DC; e += 256) {
            const int r = e / K1_DC, d = e % K1_DC;
            const int64_t gn = n0 + r;
            xs[d][r] = (gn < n && d < D) ? x[gn * D + d] : 0.0;
        }
        double pre[K1_T * K1_DC / 256];
#pragma unroll
        for (int q = 0; q < K1_T * K1_DC / 256; ++q) {
            const int e = threadIdx.x + q * 256, r = e / K1_DC, d = e % K1_DC;
            pre[q] = (r < M && d < D) ? Xc[(int64_t)r * D + d] : 0.0;
        }
        for (int m0 = 0; m0 < M; m0 += K1_T) {
            __syncthreads();
#pragma unroll
            for (int q = 0; q < K1_T * K1_DC / 256; ++q) {
                const int e = threadIdx.x + q * 256;
                cs[e % K1_DC][e / K1_DC] = pre[q];
            }
            __syncthreads();
            if (m0 + K1_T < M) {
#pragma unroll
                for (int q = 0; q < K1_T * K1_DC / 256; ++q) {
                    const int e = threadIdx.x + q * 256, r = m0 + K1_T + e / K1_DC, d = e % K1_DC;
                    pre[q] = (r < M && d < D) ? Xc[(int64_t)r * D + d] : 0.0;
                }
            }
            double acc[4][4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = 0.0;
            for (int d = 0; d < D; ++d) {
                double xv[4], cv[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) { xv[i] = xs[d][ty * 4 + i]; cv[i] = cs[d][tx * 4 + i]; }
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) { double t = xv[i] - cv[j]; acc[i][j] = fma(t, t, acc[i][j]); }
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int64_t gn = n0 + ty * 4 + i;
                if (gn >= n) continue;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int gm = m0 + tx * 4 + j;
                    if (gm < M) K[gn * ldk + gm] = exp(-beta * acc[i][j]);
                }
            }
        }
    } else
    for (int m0 = 0; m0 < M; m0 += K1_T) {
        double acc[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = 0.0;
        for (int d0 = 0; d0 < D; d0 += K1_DC) {
            __syncthreads();
            for (int e = threadIdx.x; e < K1_T * K1_DC; e += 256) {
                int r = e / K1_DC, d = e % K1_DC;
                int64_t gn = n0 + r;
                int gm = m0 + r;
                xs[d][r] = (gn < n && d0 + d < D) ? x[gn * D + d0 + d] : 0.0;
                cs[d][r] = (gm < M && d0 + d < D) ? Xc[(int64_t)gm * D + d0 + d] : 0.0;
            }
            __syncthreads();
            const int dl = min(K1_DC, D - d0);
            for (int d = 0; d < dl; ++d) {
                double xv[4], cv[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) { xv[i] = xs[d][ty * 4 + i]; cv[i] = cs[d][tx * 4 + i]; }
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) { double t = xv[i] - cv[j]; acc[i][j] = fma(t, t, acc[i][j]); }
            }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            int64_t gn = n0 + ty * 4 + i;
            if (gn >= n) continue;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                int gm = m0 + tx * 4 + j;
                if (gm < M) K[gn * ldk + gm] = exp(-beta * acc[i][j]);
            }
        }
    }
    // fused [U | Y | 0]: copy Y (if any) and zero the padding columns
    if (ycols + padcols > 0) {
        const int extra = ycols + padcols;
        for (int e = threadIdx.x; e < K1_T * extra; e += 256) {
            int r = e / extra, c = e % extra;
            int64_t gn = n0 + r;
            if (gn < n) K[gn * ldk + M + c] = (c < ycols) ? Y[gn * ycols + c] : 0.0;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// K5 / K7: one thread per cell; x, v (and a) live in registers (DP = D padded, compile time); control
// points and coefficients are read from shared memory with warp-wide broadcast loads.
// ------------------------------------------------------------------------------------------------
// ctrl points per shared-memory chunk: 2 x 16 KB of static shared memory whatever DP is
template <int DP> struct ev_chunk { static constexpr int MC = 2048 / DP; };

template <int DP>
__device__ __forceinline__ void load_chunk(double* cs, double* Cs, const double* Xc, const double* C, int D,
                                           int m0, int mc) {
    for (int e = threadIdx.x; e < mc * DP; e += blockDim.x) {
        int r = e / DP, d = e % DP;
        bool ok = d < D;
        cs[e] = ok ? Xc[(int64_t)(m0 + r) * D + d] : 0.0;
        Cs[e] = ok ? C[(int64_t)(m0 + r) * D + d] : 0.0;
    }
}

template <int DP>
__global__ void __launch_bounds__(128) k_eval(const double* __restrict__ x, int64_t n, int D,
                                              const double* __restrict__ Xc, int M, double beta,
                                              const double* __restrict__ C, double* __restrict__ v) {
    __shared__ double cs[ev_chunk<DP>::MC * DP];
    __shared__ double Cs[ev_chunk<DP>::MC * DP];
    const int64_t cell = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = cell < n;
    double xr[DP], vr[DP];
#pragma unroll
    for (int d = 0; d < DP; ++d) { xr[d] = (live && d < D) ? x[cell * D + d] : 0.0; vr[d] = 0.0; }
    for (int m0 = 0; m0 < M; m0 += ev_chunk<DP>::MC) {
        const int mc = min(ev_chunk<DP>::MC, M - m0);
        __syncthreads();
        load_chunk<DP>(cs, Cs, Xc, C, D, m0, mc);
        __syncthreads();
        for (int m = 0; m < mc; ++m) {
            // two interleaved partial sums: the FP64 pipe needs independent work behind every dependent DFMA
            double d2a = 0.0, d2b = 0.0;
#pragma unroll
            for (int d = 0; d < DP; d += 2) {
                double ta = xr[d] - cs[m * DP + d], tb = xr[d + 1] - cs[m * DP + d + 1];
                d2a = fma(ta, ta, d2a);
                d2b = fma(tb, tb, d2b);
            }
            const double k = exp(-beta * (d2a + d2b));
#pragma unroll
            for (int d = 0; d < DP; ++d) vr[d] = fma(k, Cs[m * DP + d], vr[d]);
        }
    }
    if (live) {
#pragma unroll
        for (int d = 0; d < DP; ++d)
            if (d < D) v[cell * D + d] = vr[d];
    }
}

template <int DP>
__global__ void __launch_bounds__(128) k_dac(const double* __restrict__ x, int64_t n, int D,
                                             const double* __restrict__ Xc, int M, double beta,
                                             const double* __restrict__ C, int formula, double* __restrict__ v_out,
                                             double* __restrict__ div_out, double* __restrict__ acc_out,
                                             double* __restrict__ accn_out, double* __restrict__ curv_out,
                                             double* __restrict__ curvn_out) {
    __shared__ double cs[ev_chunk<DP>::MC * DP];
    __shared__ double Cs[ev_chunk<DP>::MC * DP];
    const int64_t cell = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = cell < n;
    double xr[DP], vr[DP];
#pragma unroll
    for (int d = 0; d < DP; ++d) { xr[d] = (live && d < D) ? x[cell * D + d] : 0.0; vr[d] = 0.0; }
    double div = 0.0;
    // pass 1: v = sum_m k_m C_m ; div = -2 beta sum_m k_m C_m . (x - c_m)
    for (int m0 = 0; m0 < M; m0 += ev_chunk<DP>::MC) {
        const int mc = min(ev_chunk<DP>::MC, M - m0);
        __syncthreads();
        load_chunk<DP>(cs, Cs, Xc, C, D, m0, mc);
        __syncthreads();
        for (int m = 0; m < mc; ++m) {
            double d2a = 0.0, d2b = 0.0, t2a = 0.0, t2b = 0.0;
#pragma unroll
            for (int d = 0; d < DP; d += 2) {
                double ta = xr[d] - cs[m * DP + d], tb = xr[d + 1] - cs[m * DP + d + 1];
                d2a = fma(ta, ta, d2a);
                d2b = fma(tb, tb, d2b);
                t2a = fma(ta, Cs[m * DP + d], t2a);
                t2b = fma(tb, Cs[m * DP + d + 1], t2b);
            }
            const double k = exp(-beta * (d2a + d2b));
            div = fma(k, t2a + t2b, div);
#pragma unroll
            for (int d = 0; d < DP; ++d) vr[d] = fma(k, Cs[m * DP + d], vr[d]);
        }
    }
    div *= -2.0 * beta;
    const bool need_acc = acc_out || accn_out || curv_out || curvn_out;
    double ar[DP];
#pragma unroll
    for (int d = 0; d < DP; ++d) ar[d] = 0.0;
    if (need_acc) {
        // pass 2: a = J v = -2 beta sum_m k_m ((x - c_m) . v) C_m
        for (int m0 = 0; m0 < M; m0 += ev_chunk<DP>::MC) {
            const int mc = min(ev_chunk<DP>::MC, M - m0);
            __syncthreads();
            load_chunk<DP>(cs, Cs, Xc, C, D, m0, mc);
            __syncthreads();
            for (int m = 0; m < mc; ++m) {
                double d2a = 0.0, d2b = 0.0, sa = 0.0, sb = 0.0;
#pragma unroll
                for (int d = 0; d < DP; d += 2) {
                    double ta = xr[d] - cs[m * DP + d], tb = xr[d + 1] - cs[m * DP + d + 1];
                    d2a = fma(ta, ta, d2a);
                    d2b = fma(tb, tb, d2b);
                    sa = fma(ta, vr[d], sa);
                    sb = fma(tb, vr[d + 1], sb);
                }
                const double ks = exp(-beta * (d2a + d2b)) * (sa + sb);
#pragma unroll
                for (int d = 0; d < DP; ++d) ar[d] = fma(ks, Cs[m * DP + d], ar[d]);
            }
        }
#pragma unroll
        for (int d = 0; d < DP; ++d) ar[d] *= -2.0 * beta;
    }
    if (!live) return;
    if (v_out) {
#pragma unroll
        for (int d = 0; d < DP; ++d)
            if (d < D) v_out[cell * D + d] = vr[d];
    }
    if (div_out) div_out[cell] = div;
    if (!need_acc) return;
    double vv = 0.0, va = 0.0, aa = 0.0;
#pragma unroll
    for (int d = 0; d < DP; ++d) { vv = fma(vr[d], vr[d], vv); va = fma(vr[d], ar[d], va); aa = fma(ar[d], ar[d], aa); }
    if (acc_out) {
#pragma unroll
        for (int d = 0; d < DP; ++d)
            if (d < D) acc_out[cell * D + d] = ar[d];
    }
    if (accn_out) accn_out[cell] = sqrt(aa);
    if (formula == 2) {
        // reference utils.py:752-757: (a (v.v) - v (v.a)) / |v|^4
        const double nv = sqrt(vv);
        const double den = (nv * nv) * (nv * nv);
        double kk = 0.0;
#pragma unroll
        for (int d = 0; d < DP; ++d) {
            double c = (ar[d] * vv - vr[d] * va) / den;   // padded dims: a = v = 0 -> c = 0
            if (curv_out && d < D) curv_out[cell * D + d] = c;
            kk = fma(c, c, kk);
        }
        if (curvn_out) curvn_out[cell] = sqrt(kk);
    } else if (curvn_out) {
        // reference utils.py:743-749: |outer(v, a)|_F / |v|^3 = |v| |a| / |v|^3
        const double nv = sqrt(vv);
        curvn_out[cell] = sqrt(vv * aa) / (nv * nv * nv);
    }
}

// K7 v3 (D <= 32): the kernel values of pass 1 stay ON CHIP for pass 2.  k_dac recomputes |x - c|^2 and exp in its second
// pass (118 of its 300 FP64 instructions per cell and control point; parking k in HBM was slower).  Here a CTA owns CC
// cells and Q threads share a cell, each taking every Q-th control point: the CC x M kernel values fit in shared memory
// (32 cells x 512 control points = 128 KB), x / partial v / partial a stay in registers, and the partial sums are
// combined with log2(Q) shuffles per dimension.  Pass 2 needs only  s_m = (x - c_m) . v = (x - s) . v - (c_m - s) . v
// (s = the first control point: any shift is algebraically exact and keeps both terms at the scale of the data
// spread), i.e. ONE fma per dimension, then a += (k_m s_m) C_m: 2 D instead of 4 D + exp.
// Per cell and control point: pass 1 4 D + exp, pass 2 2 D  ->  ~6 D + 20 = 212 FP64 instructions at D = 32 (was 300).
// Xcp = control points, Xcsp = control points minus the first one, Cpad = coefficients: each [M][DP], zero padded
__global__ void k_pad_shift(const double* __restrict__ Xc, const double* __restrict__ C, int M, int D, int DP,
                            double* __restrict__ Xcp, double* __restrict__ Xcsp, double* __restrict__ Cpad) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= M * DP) return;
    const int m = e / DP, d = e % DP;
    const bool ok = d < D;
    const double c = ok ? Xc[(int64_t)m * D + d] : 0.0;
    Xcp[e] = c;
    Xcsp[e] = ok ? c - Xc[d] : 0.0;
    Cpad[e] = ok ? C[(int64_t)m * D + d] : 0.0;
}

template <int DP, int CC, int Q>
__global__ void __launch_bounds__(CC * Q) k_dac3(const double* __restrict__ x, int64_t n, int D,
                                                 const double* __restrict__ Xcp, const double* __restrict__ Xcsp,
                                                 const double* __restrict__ Cpad, int M, int Mp, double beta,
                                                 int formula, double* __restrict__ v_out,
                                                 double* __restrict__ div_out, double* __restrict__ acc_out,
                                                 double* __restrict__ accn_out, double* __restrict__ curv_out,
                                                 double* __restrict__ curvn_out) {
    // Thread layout: warp q (0..Q-1) x lane = cell (CC = 32).  All lanes of a warp read the SAME control point at the same
    // time, so every shared-memory load is a broadcast (one wavefront); with Q adjacent lanes per cell (first version)
    // each quarter-warp of an LDS.128 touched 8 different rows = 4 wavefronts per load and the kernel was bound by
    // shared-memory bandwidth (23 ms).  The Q partial sums of v (and a) meet through shared memory.
    static_assert(CC == 32, "one warp = 32 cells");
    constexpr int LDR = DP;              // broadcast reads: no padding needed
    constexpr int MC = 64;               // control points per staged chunk
    constexpr int STG = 4 * MC * LDR;    // [2 stages][c | C][MC][LDR] doubles; also the Q x DP x CC reduction buffer
    static_assert(STG >= Q * DP * CC, "reduction buffer must fit the staging area");
    extern __shared__ __align__(128) double sm3[];
    double* stg = sm3;
    double* ks = stg + STG;              // [Mp][CC]   kernel values, cell fastest
    const int tid = threadIdx.x, q = tid >> 5, cl = tid & 31;
    const int64_t cell = (int64_t)blockIdx.x * CC + cl;
    const bool live = cell < n;
    double xr[DP], vr[DP], ar[DP];
#pragma unroll
    for (int d = 0; d < DP; ++d) { xr[d] = (live && d < D) ? x[cell * D + d] : 0.0; vr[d] = 0.0; ar[d] = 0.0; }
    double div = 0.0, xv = 0.0;
    const bool need_acc = acc_out || accn_out || curv_out || curvn_out;
    const int nch = (M + MC - 1) / MC;
    // sum of the Q partial vectors held in registers (w) -> every thread gets the complete vector back
    auto reduce_q = [&](double (&w)[DP]) {
        __syncthreads();                 // staging area free (pipeline drained by the caller)
#pragma unroll
        for (int d = 0; d < DP; ++d) stg[(q * DP + d) * CC + cl] = w[d];
        __syncthreads();
#pragma unroll
        for (int d = 0; d < DP; ++d) {
            double acc = 0.0;
#pragma unroll
            for (int qq = 0; qq < Q; ++qq) acc += stg[(qq * DP + d) * CC + cl];
            w[d] = acc;
        }
        fence_proxy_async_smem();
        __syncthreads();
    };
    // Staging: the control points, the shifted control points and the coefficients come PRE-PADDED to DP columns from a
    // global scratch (k_pad_shift), so a chunk of MC rows is one contiguous block and ONE thread stages it with two TMA
    // bulk copies that complete on an mbarrier; the compute warps never execute copy code.  Global chunk counter g:
    // stage g & 1, barrier phase (g >> 1) & 1.  A stage is refilled only after the block barrier that ends its use, and
    // generic-proxy accesses are fenced before the async-proxy write (see kern_gram.cu).
    __shared__ __align__(8) uint64_t full[2];
    if (tid == 0) { mbar_init(&full[0], 1); mbar_init(&full[1], 1); mbar_fence_init(); }
    __syncthreads();
    auto issue = [&](int g, const double* cbase, int t) {           // chunk t of the pass whose control points are cbase
        if (tid == 0) {
            const int m0 = t * MC, mc = min(MC, M - m0);
            const unsigned bytes = (unsigned)(mc * DP * sizeof(double));
            double* cdst = stg + (g & 1) * (2 * MC * LDR);
            mbar_arrive_expect_tx(&full[g & 1], 2 * bytes);
            bulk_g2s(cdst, cbase + (int64_t)m0 * DP, bytes, &full[g & 1]);
            bulk_g2s(cdst + MC * LDR, Cpad + (int64_t)m0 * DP, bytes, &full[g & 1]);
        }
    };
    int g = 0;
    // ---- pass 1: k_m (kept on chip), v = sum_m k_m C_m, div = -2 beta sum_m k_m C_m . (x - c_m) ----
    issue(g, Xcp, 0);
    for (int t = 0; t < nch; ++t, ++g) {
        if (t + 1 < nch) issue(g + 1, Xcp, t + 1);                  // the next chunk travels while this one is consumed
        mbar_wait(&full[g & 1], (g >> 1) & 1);
        const double* cs = stg + (g & 1) * (2 * MC * LDR);
        const double* Cs = cs + MC * LDR;
        const int m0 = t * MC, mc = min(MC, M - m0);
        for (int mm = q; mm < mc; mm += Q) {
            const double* cr = cs + mm * LDR;
            const double* Cr = Cs + mm * LDR;
            double d2a = 0.0, d2b = 0.0, t2a = 0.0, t2b = 0.0;
#pragma unroll
            for (int d = 0; d < DP; d += 2) {
                const double ta = xr[d] - cr[d], tb = xr[d + 1] - cr[d + 1];
                d2a = fma(ta, ta, d2a);
                d2b = fma(tb, tb, d2b);
                t2a = fma(ta, Cr[d], t2a);
                t2b = fma(tb, Cr[d + 1], t2b);
            }
            const double k = exp(-beta * (d2a + d2b));
            ks[(m0 + mm) * CC + cl] = k;
            div = fma(k, t2a + t2b, div);
#pragma unroll
            for (int d = 0; d < DP; ++d) vr[d] = fma(k, Cr[d], vr[d]);
        }
        fence_proxy_async_smem();
        __syncthreads();                 // this stage may be refilled by the copy issued in the next iteration
    }
    reduce_q(vr);
    {   // div: Q partials through the same buffer
        stg[q * CC + cl] = div;
        __syncthreads();
        div = 0.0;
#pragma unroll
        for (int qq = 0; qq < Q; ++qq) div += stg[qq * CC + cl];
        fence_proxy_async_smem();
        __syncthreads();
    }
    div *= -2.0 * beta;
    if (need_acc) {
        // ---- pass 2: a = J v = -2 beta sum_m k_m s_m C_m,  s_m = (x - s) . v - (c_m - s) . v  (Xcsp = c - s) ----
#pragma unroll
        for (int d = 0; d < DP; ++d) xv = fma((live && d < D) ? xr[d] - Xcp[d] : 0.0, vr[d], xv);
        issue(g, Xcsp, 0);
        for (int t = 0; t < nch; ++t, ++g) {
            if (t + 1 < nch) issue(g + 1, Xcsp, t + 1);
            mbar_wait(&full[g & 1], (g >> 1) & 1);
            const double* cs = stg + (g & 1) * (2 * MC * LDR);
            const double* Cs = cs + MC * LDR;
            const int m0 = t * MC, mc = min(MC, M - m0);
            for (int mm = q; mm < mc; mm += Q) {
                const double* cr = cs + mm * LDR;
                const double* Cr = Cs + mm * LDR;
                double sa = 0.0, sb = 0.0;
#pragma unroll
                for (int d = 0; d < DP; d += 2) {
                    sa = fma(cr[d], vr[d], sa);
                    sb = fma(cr[d + 1], vr[d + 1], sb);
                }
                const double ksm = ks[(m0 + mm) * CC + cl] * (xv - (sa + sb));
#pragma unroll
                for (int d = 0; d < DP; ++d) ar[d] = fma(ksm, Cr[d], ar[d]);
            }
            fence_proxy_async_smem();
            __syncthreads();
        }
        reduce_q(ar);
#pragma unroll
        for (int d = 0; d < DP; ++d) ar[d] *= -2.0 * beta;
    }
    if (!live) return;
    // every one of the Q threads of a cell now holds the complete v (and a); thread q writes the dimensions d = q (mod Q)
    if (v_out) {
#pragma unroll
        for (int d = 0; d < DP; ++d)
            if (d < D && (d % Q) == q) v_out[cell * D + d] = vr[d];
    }
    if (div_out && q == 0) div_out[cell] = div;
    if (!need_acc) return;
    double vv = 0.0, va = 0.0, aa = 0.0;
#pragma unroll
    for (int d = 0; d < DP; ++d) { vv = fma(vr[d], vr[d], vv); va = fma(vr[d], ar[d], va); aa = fma(ar[d], ar[d], aa); }
    if (acc_out) {
#pragma unroll
        for (int d = 0; d < DP; ++d)
            if (d < D && (d % Q) == q) acc_out[cell * D + d] = ar[d];
    }
    if (accn_out && q == 0) accn_out[cell] = sqrt(aa);
    if (formula == 2) {
        const double nv = sqrt(vv);
        const double den = (nv * nv) * (nv * nv);
        double kk = 0.0;
#pragma unroll
        for (int d = 0; d < DP; ++d) {
            const double c = (ar[d] * vv - vr[d] * va) / den;   // padded dims: a = v = 0 -> c = 0
            if (curv_out && d < D && (d % Q) == q) curv_out[cell * D + d] = c;
            kk = fma(c, c, kk);
        }
        if (curvn_out && q == 0) curvn_out[cell] = sqrt(kk);
    } else if (curvn_out && q == 0) {
        const double nv = sqrt(vv);
        curvn_out[cell] = sqrt(vv * aa) / (nv * nv * nv);
    }
}

template <int DP, int CC, int Q>
static int launch_dac3(const double* x, int64_t n, int D, const double* Xc, int M, double beta, const double* C,
                       int formula, double* v, double* div, double* acc, double* acc_norm, double* curv,
                       double* curv_norm, cudaStream_t st) {
    const int Mp = M;                             // ks is [M][CC], cell fastest
    const size_t smem = sizeof(double) * (4 * (size_t)64 * DP + (size_t)CC * Mp);
    static size_t attr = 0;
    if (smem > attr) {
        DVF_CUDA(cudaFuncSetAttribute(k_dac3<DP, CC, Q>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr = smem;
    }
    double* pad = nullptr;                        // [Xcp | Xcsp | Cpad], each M x DP, stream-ordered scratch
    const size_t one = (size_t)M * DP;
    DVF_CUDA(cudaMallocAsync(&pad, sizeof(double) * 3 * one, st));
    k_pad_shift<<<(M * DP + 255) / 256, 256, 0, st>>>(Xc, C, M, D, DP, pad, pad + one, pad + 2 * one);
    DVF_LAUNCH_CHECK();
    k_dac3<DP, CC, Q><<<(unsigned)((n + CC - 1) / CC), CC * Q, smem, st>>>(x, n, D, pad, pad + one, pad + 2 * one, M, Mp, beta,
                                                                        formula, v, div, acc, acc_norm, curv, curv_norm);
    DVF_LAUNCH_CHECK();
    DVF_CUDA(cudaFreeAsync(pad, st));
    return DVF_OK;
}

// Least-action-path terms (reference vectorfield/least_action_path.py:7-57): per path segment with centre x and
// finite-difference velocity vp,   dv = vp - f(x)   and   z = dv J(x)   (the row-vector / Jacobian product action_grad needs),
// without materialising J:  z_j = -2 beta sum_m k_m (dv . C_m) (x_j - c_mj).  Same skeleton as k_dac (one thread per point,
// two passes over the control points, the second one with the complete dv).  D <= 32.
template <int DP>
__global__ void __launch_bounds__(128) k_lap_terms(const double* __restrict__ x, const double* __restrict__ vp, int64_t n,
                                                   int D, const double* __restrict__ Xc, int M, double beta,
                                                   const double* __restrict__ C, double* __restrict__ dv_out,
                                                   double* __restrict__ z_out) {
    __shared__ double cs[ev_chunk<DP>::MC * DP];
    __shared__ double Cs[ev_chunk<DP>::MC * DP];
    const int64_t cell = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = cell < n;
    double xr[DP], wr[DP];
#pragma unroll
    for (int d = 0; d < DP; ++d) { xr[d] = (live && d < D) ? x[cell * D + d] : 0.0; wr[d] = 0.0; }
    for (int m0 = 0; m0 < M; m0 += ev_chunk<DP>::MC) {
        const int mc = min(ev_chunk<DP>::MC, M - m0);
        __syncthreads();
        load_chunk<DP>(cs, Cs, Xc, C, D, m0, mc);
        __syncthreads();
        for (int m = 0; m < mc; ++m) {
            double d2a = 0.0, d2b = 0.0;
#pragma unroll
            for (int d = 0; d < DP; d += 2) {
                double ta = xr[d] - cs[m * DP + d], tb = xr[d + 1] - cs[m * DP + d + 1];
                d2a = fma(ta, ta, d2a);
                d2b = fma(tb, tb, d2b);
            }
            const double k = exp(-beta * (d2a + d2b));
#pragma unroll
            for (int d = 0; d < DP; ++d) wr[d] = fma(k, Cs[m * DP + d], wr[d]);
        }
    }
#pragma unroll
    for (int d = 0; d < DP; ++d) wr[d] = (live && d < D) ? vp[cell * D + d] - wr[d] : 0.0;     // dv = vp - f(x)
    double zr[DP];
#pragma unroll
    for (int d = 0; d < DP; ++d) zr[d] = 0.0;
    if (z_out) {
        for (int m0 = 0; m0 < M; m0 += ev_chunk<DP>::MC) {
            const int mc = min(ev_chunk<DP>::MC, M - m0);
            __syncthreads();
            load_chunk<DP>(cs, Cs, Xc, C, D, m0, mc);
            __syncthreads();
            for (int m = 0; m < mc; ++m) {
                double d2a = 0.0, d2b = 0.0, sa = 0.0, sb = 0.0;
#pragma unroll
                for (int d = 0; d < DP; d += 2) {
                    double ta = xr[d] - cs[m * DP + d], tb = xr[d + 1] - cs[m * DP + d + 1];
                    d2a = fma(ta, ta, d2a);
                    d2b = fma(tb, tb, d2b);
                    sa = fma(wr[d], Cs[m * DP + d], sa);
                    sb = fma(wr[d + 1], Cs[m * DP + d + 1], sb);
                }
                const double ks = exp(-beta * (d2a + d2b)) * (sa + sb);
#pragma unroll
                for (int d = 0; d < DP; ++d) zr[d] = fma(ks, xr[d] - cs[m * DP + d], zr[d]);
            }
        }
    }
    if (!live) return;
#pragma unroll
    for (int d = 0; d < DP; ++d)
        if (d < D) {
            if (dv_out) dv_out[cell * D + d] = wr[d];
            if (z_out) z_out[cell * D + d] = -2.0 * beta * zr[d];
        }
}

// K7 for D > 32: x, v and a of one cell no longer fit the register file of one thread (D = 50: 696 bytes of spills in
// the kernel above), so TWO adjacent lanes share a cell, each holding HD = DP / 2 of its dimensions; the partial
// distances and dot products are completed with one shuffle each.  Same arithmetic per dimension as k_dac; the two
// halves of a sum are added lower half first.
template <int HD>
__global__ void __launch_bounds__(128) k_dac2(const double* __restrict__ x, int64_t n, int D,
                                              const double* __restrict__ Xc, int M, double beta,
                                              const double* __restrict__ C, int formula, double* __restrict__ v_out,
                                              double* __restrict__ div_out, double* __restrict__ acc_out,
                                              double* __restrict__ accn_out, double* __restrict__ curv_out,
                                              double* __restrict__ curvn_out) {
    constexpr int DP = 2 * HD;
    __shared__ double cs[ev_chunk<DP>::MC * DP];
    __shared__ double Cs[ev_chunk<DP>::MC * DP];
    const int64_t cell = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 1;
    const int h = threadIdx.x & 1, d0 = h * HD;
    const bool live = cell < n;
    double xr[HD], vr[HD];
#pragma unroll
    for (int d = 0; d < HD; ++d) { xr[d] = (live && d0 + d < D) ? x[cell * D + d0 + d] : 0.0; vr[d] = 0.0; }
    double div = 0.0;
    for (int m0 = 0; m0 < M; m0 += ev_chunk<DP>::MC) {
        const int mc = min(ev_chunk<DP>::MC, M - m0);
        __syncthreads();
        load_chunk<DP>(cs, Cs, Xc, C, D, m0, mc);
        __syncthreads();
        for (int m = 0; m < mc; ++m) {
            const double* cm = cs + m * DP + d0;
            const double* Cm = Cs + m * DP + d0;
            double d2a = 0.0, d2b = 0.0, t2a = 0.0, t2b = 0.0;
#pragma unroll
            for (int d = 0; d < HD; d += 2) {
                double ta = xr[d] - cm[d], tb = xr[d + 1] - cm[d + 1];
                d2a = fma(ta, ta, d2a);
                d2b = fma(tb, tb, d2b);
                t2a = fma(ta, Cm[d], t2a);
                t2b = fma(tb, Cm[d + 1], t2b);
            }
            double d2 = d2a + d2b, t2 = t2a + t2b;
            const double d2o = __shfl_xor_sync(0xffffffffu, d2, 1), t2o = __shfl_xor_sync(0xffffffffu, t2, 1);
            d2 = h ? d2o + d2 : d2 + d2o;          // lower half first on both lanes: identical k
            t2 = h ? t2o + t2 : t2 + t2o;
            const double k = exp(-beta * d2);
            div = fma(k, t2, div);
#pragma unroll
            for (int d = 0; d < HD; ++d) vr[d] = fma(k, Cm[d], vr[d]);
        }
    }
    div *= -2.0 * beta;
    const bool need_acc = acc_out || accn_out || curv_out || curvn_out;
    double ar[HD];
#pragma unroll
    for (int d = 0; d < HD; ++d) ar[d] = 0.0;
    if (need_acc) {
        for (int m0 = 0; m0 < M; m0 += ev_chunk<DP>::MC) {
            const int mc = min(ev_chunk<DP>::MC, M - m0);
            __syncthreads();
            load_chunk<DP>(cs, Cs, Xc, C, D, m0, mc);
            __syncthreads();
            for (int m = 0; m < mc; ++m) {
                const double* cm = cs + m * DP + d0;
                const double* Cm = Cs + m * DP + d0;
                double d2a = 0.0, d2b = 0.0, sa = 0.0, sb = 0.0;
#pragma unroll
                for (int d = 0; d < HD; d += 2) {
                    double ta = xr[d] - cm[d], tb = xr[d + 1] - cm[d + 1];
                    d2a = fma(ta, ta, d2a);
                    d2b = fma(tb, tb, d2b);
                    sa = fma(ta, vr[d], sa);
                    sb = fma(tb, vr[d + 1], sb);
                }
                double d2 = d2a + d2b, sv = sa + sb;
                const double d2o = __shfl_xor_sync(0xffffffffu, d2, 1), svo = __shfl_xor_sync(0xffffffffu, sv, 1);
                d2 = h ? d2o + d2 : d2 + d2o;
                sv = h ? svo + sv : sv + svo;
                const double ks = exp(-beta * d2) * sv;
#pragma unroll
                for (int d = 0; d < HD; ++d) ar[d] = fma(ks, Cm[d], ar[d]);
            }
        }
#pragma unroll
        for (int d = 0; d < HD; ++d) ar[d] *= -2.0 * beta;
    }
    // per-cell scalars need both halves (every lane of the warp takes part in the shuffles)
    double vv = 0.0, va = 0.0, aa = 0.0;
#pragma unroll
    for (int d = 0; d < HD; ++d) { vv = fma(vr[d], vr[d], vv); va = fma(vr[d], ar[d], va); aa = fma(ar[d], ar[d], aa); }
    {
        const double vvo = __shfl_xor_sync(0xffffffffu, vv, 1), vao = __shfl_xor_sync(0xffffffffu, va, 1),
                     aao = __shfl_xor_sync(0xffffffffu, aa, 1);
        vv = h ? vvo + vv : vv + vvo;
        va = h ? vao + va : va + vao;
        aa = h ? aao + aa : aa + aao;
    }
    double kk = 0.0;
    double cr[HD];
    const double nv = sqrt(vv);
    if (formula == 2) {
        const double den = (nv * nv) * (nv * nv);
#pragma unroll
        for (int d = 0; d < HD; ++d) { cr[d] = (ar[d] * vv - vr[d] * va) / den; kk = fma(cr[d], cr[d], kk); }
    }
    {
        const double kko = __shfl_xor_sync(0xffffffffu, kk, 1);
        kk = h ? kko + kk : kk + kko;
    }
    if (!live) return;
    if (v_out) {
#pragma unroll
        for (int d = 0; d < HD; ++d)
            if (d0 + d < D) v_out[cell * D + d0 + d] = vr[d];
    }
    if (div_out && h == 0) div_out[cell] = div;
    if (!need_acc) return;
    if (acc_out) {
#pragma unroll
        for (int d = 0; d < HD; ++d)
            if (d0 + d < D) acc_out[cell * D + d0 + d] = ar[d];
    }
    if (accn_out && h == 0) accn_out[cell] = sqrt(aa);
    if (formula == 2) {
        if (curv_out) {
#pragma unroll
            for (int d = 0; d < HD; ++d)
                if (d0 + d < D) curv_out[cell * D + d0 + d] = cr[d];
        }
        if (curvn_out && h == 0) curvn_out[cell] = sqrt(kk);
    } else if (curvn_out && h == 0) {
        curvn_out[cell] = sqrt(vv * aa) / (nv * nv * nv);
    }
}

template <typename F>
static int dispatch_dp(int D, F&& f) {
    if (D <= 2) return f(std::integral_constant<int, 2>());
    if (D <= 4) return f(std::integral_constant<int, 4>());
    if (D <= 8) return f(std::integral_constant<int, 8>());
    if (D <= 16) return f(std::integral_constant<int, 16>());
    if (D <= 24) return f(std::integral_constant<int, 24>());
    if (D <= 32) return f(std::integral_constant<int, 32>());
    if (D <= 40) return f(std::integral_constant<int, 40>());
    if (D <= 48) return f(std::integral_constant<int, 48>());
    return f(std::integral_constant<int, 64>());
}

}  // namespace dvf

using namespace dvf;

static int check_field_args(const void* x, int64_t n, int D, const void* Xc, int M) {
    DVF_ARG(n >= 0 && M >= 1, "n=%lld M=%d: need n >= 0, M >= 1", (long long)n, M);
    DVF_ARG(D >= 1 && D <= kMaxD, "D=%d unsupported (1..%d)", D, kMaxD);
    DVF_ARG((x || n == 0) && Xc, "NULL pointer");
    return DVF_OK;
}

extern "C" int dvf_kernel_matrix(const double* x, int64_t n, int D, const double* Xc, int M, double beta,
                                 double* K, int64_t ldk, void* stream) {
    DVF_ARG(n >= 0 && M >= 1 && D >= 1, "bad shape n=%lld D=%d M=%d", (long long)n, D, M);
    DVF_ARG(ldk >= M && K && Xc, "ldk < M or NULL pointer");
    if (n == 0) return DVF_OK;
    k_kernel_matrix<<<(unsigned)((n + K1_T - 1) / K1_T), 256, 0, (cudaStream_t)stream>>>(x, n, D, Xc, M, beta, K, ldk,
                                                                                       nullptr, 0, 0);
    DVF_LAUNCH_CHECK();
    return DVF_OK;
}

// K with the columns M .. ldk-1 zero-filled (operand of the Jacobian GEMM; internal)
extern "C" int dvf_kernel_matrix_padded(const double* x, int64_t n, int D, const double* Xc, int M, double beta, double* K,
                                        int64_t ldk, void* stream) {
    DVF_ARG(n >= 0 && M >= 1 && D >= 1 && ldk >= M && K && Xc, "bad argument");
    if (n == 0) return DVF_OK;
    k_kernel_matrix<<<(unsigned)((n + K1_T - 1) / K1_T), 256, 0, (cudaStream_t)stream>>>(x, n, D, Xc, M, beta, K, ldk,
                                                                                       nullptr, 0, (int)(ldk - M));
    DVF_LAUNCH_CHECK();
    return DVF_OK;
}

extern "C" int dvf_build_uy(const double* X, const double* Y, int64_t N, int D, const double* Xc, int M,
                            double beta, double* UY, void* stream) {
    if (int rc = check_field_args(X, N, D, Xc, M)) return rc;
    DVF_ARG(Y && UY, "NULL pointer");
    if (N == 0) return DVF_OK;
    const int64_t W = dvf_fit_row_width(M, D);
    k_kernel_matrix<<<(unsigned)((N + K1_T - 1) / K1_T), 256, 0, (cudaStream_t)stream>>>(X, N, D, Xc, M, beta, UY, W, Y,
                                                                                       D, (int)(W - M - D));
    DVF_LAUNCH_CHECK();
    return DVF_OK;
}

extern "C" int dvf_eval(const double* x, int64_t n, int D, const double* Xc, int M, double beta,
                        const double* C, double* v, void* stream) {
    if (int rc = check_field_args(x, n, D, Xc, M)) return rc;
    DVF_ARG(C && (v || n == 0), "NULL pointer");
    if (n == 0) return DVF_OK;
    return dispatch_dp(D, [&](auto dp) {
        constexpr int DP = decltype(dp)::value;
        k_eval<DP><<<(unsigned)((n + 127) / 128), 128, 0, (cudaStream_t)stream>>>(x, n, D, Xc, M, beta, C, v);
        DVF_LAUNCH_CHECK();
        return DVF_OK;
    });
}

extern "C" int dvf_div_acc_curv(const double* x, int64_t n, int D, const double* Xc, int M, double beta,
                                const double* C, int formula, double* v, double* div, double* acc,
                                double* acc_norm, double* curv, double* curv_norm, void* stream) {
    if (int rc = check_field_args(x, n, D, Xc, M)) return rc;
    DVF_ARG(C, "NULL pointer");
    DVF_ARG(formula == 1 || formula == 2, "formula must be 1 or 2");
    if (n == 0) return DVF_OK;
    if (D > 32) {   // two lanes per cell (k_dac2): HD = half of D rounded up to the next multiple of 4
        const unsigned blocks = (unsigned)((2 * n + 127) / 128);
        auto launch = [&](auto hd) {
            constexpr int HD = decltype(hd)::value;
            k_dac2<HD><<<blocks, 128, 0, (cudaStream_t)stream>>>(x, n, D, Xc, M, beta, C, formula, v, div, acc, acc_norm, curv,
                                                               curv_norm);
        };
        if (D <= 40) launch(std::integral_constant<int, 20>());
        else if (D <= 48) launch(std::integral_constant<int, 24>());
        else if (D <= 56) launch(std::integral_constant<int, 28>());
        else launch(std::integral_constant<int, 32>());
        DVF_LAUNCH_CHECK();
        return DVF_OK;
    }
    // v3 (k_dac3: kernel values kept on chip between the passes, 212 instead of 300 FP64 instructions per cell and control
    // point) is correct -- it agrees with k_dac to 5e-15 and passes the parity tests -- but MEASURED SLOWER on B200 in all
    // three forms tried (1M x 30, M = 500; k_dac 12.0 ms): Q adjacent lanes per cell 26 ms -> with cp.async double
    // buffering 23 ms (each quarter-warp of an LDS.128 touched 8 rows: 4 wavefronts per load) -> warp per q, lanes = cells
    // (broadcast loads) 16.4 ms -> staging by one thread with TMA bulk copies from pre-padded scratch 15.8 ms.  What is
    // left is the shared-memory -> register path: both passes load the same 2 x DP doubles per cell and control point, one
    // use per loaded value, so halving the FP64 work of pass 2 does not shorten it.  It stays behind DVF_DAC_V3=1.
    static const bool v3 = getenv("DVF_DAC_V3") && atoi(getenv("DVF_DAC_V3")) != 0;
    const bool want_acc = acc || acc_norm || curv || curv_norm;
    if (v3 && want_acc && D >= 8 && M >= 64 && M <= 600) {       // DP >= 8: the reduction buffer fits the staging area
        return dispatch_dp(D, [&](auto dp) {
            constexpr int DP = decltype(dp)::value;
            return launch_dac3<DP, 32, 8>(x, n, D, Xc, M, beta, C, formula, v, div, acc, acc_norm, curv, curv_norm,
                                          (cudaStream_t)stream);
        });
    }
    return dispatch_dp(D, [&](auto dp) {
        constexpr int DP = decltype(dp)::value;
        k_dac<DP><<<(unsigned)((n + 127) / 128), 128, 0, (cudaStream_t)stream>>>(x, n, D, Xc, M, beta, C, formula, v, div,
                                                                              acc, acc_norm, curv, curv_norm);
        DVF_LAUNCH_CHECK();
        return DVF_OK;
    });
}

extern "C" int dvf_lap_terms(const double* x, const double* vp, int64_t n, int D, const double* Xc, int M, double beta,
                             const double* C, double* dv, double* z, void* stream) {
    if (int rc = check_field_args(x, n, D, Xc, M)) return rc;
    DVF_ARG(C && vp, "NULL pointer");
    DVF_ARG(D <= 32, "dvf_lap_terms serves D <= 32 (D = %d)", D);
    if (n == 0) return DVF_OK;
    return dispatch_dp(D, [&](auto dp) {
        constexpr int DP = decltype(dp)::value;
        k_lap_terms<DP><<<(unsigned)((n + 127) / 128), 128, 0, (cudaStream_t)stream>>>(x, vp, n, D, Xc, M, beta, C, dv, z);
        DVF_LAUNCH_CHECK();
        return DVF_OK;
    });
}

// K9: per-cell post-processing that the reference does in Python loops over cells
//   dvf_jacobian_det : det(J_c) for every cell      (vector_calculus.py:329, np.linalg.det per cell)
//   dvf_cell_angles  : signed angle between the observed and the fitted velocity of every cell
//                      (VectorField.py:255-259 -> vectorfield/utils.py:1148-1173)
//   dvf_row_norms    : |v_c|                        (vector_calculus.py:192, speed)
// All three are bound by reading their input once from HBM; the determinant additionally does a D^3/3 LU per cell.
#include "common.cuh"

namespace dvf {

// ---- batched determinant -----------------------------------------------------------------------------------
// J is (D, D, n) with the cell index fastest (the reference layout), so the block's `cpb` consecutive cells are
// read as contiguous runs of cpb doubles per matrix element; each warp then factorises one cell's matrix in shared
// memory: LU with partial pivoting exactly as LAPACK dgetf2 does it (first largest |a_ik| as the pivot, the column
// scaled by the reciprocal of the pivot), det = (+-1) * prod u_kk, accumulated as mantissa x 2^exponent so that a
// product whose partial products leave the double range still comes out right when the result is representable.
__device__ __forceinline__ int det_ld(int D) { return D | 1; }
__host__ __device__ inline int det_mat_stride(int D) { int s = D * (D | 1); return s | 1; }

__global__ void __launch_bounds__(512)
k_jacobian_det(const double* __restrict__ J, int64_t n, int D, int cpb, double* __restrict__ det) {
    extern __shared__ __align__(16) double det_smem[];
    const int S = det_mat_stride(D), LD = det_ld(D);
    const int64_t c0 = (int64_t)blockIdx.x * cpb;
    // coalesced load: consecutive threads -> consecutive cells of one matrix element
    {
        const int c = threadIdx.x % cpb;                  // blockDim.x = 32 * cpb: 32 matrix elements per pass
        for (int e = threadIdx.x / cpb; e < D * D; e += 32) {
            const int i = e / D, j = e - i * D;
            if (c0 + c < n) det_smem[c * S + i * LD + j] = J[(int64_t)e * n + c0 + c];
        }
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp >= cpb || c0 + warp >= n) return;
    double* A = det_smem + warp * S;
    double mant = 1.0;
    int expo = 0;
    bool zero = false;
    for (int k = 0; k < D; ++k) {
        // pivot: first row with the largest |a_ik|, i >= k  (idamax)
        double best = -1.0;
        int bi = k;
        for (int i = k + lane; i < D; i += 32) {
            const double v = fabs(A[i * LD + k]);
            if (v > best) { best = v; bi = i; }           // strided scan keeps the first maximum of this lane
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
        }
        if (!(best > 0.0)) {
            // every lane holds the same `best`: 0 = a zero column (singular), -1 = only NaNs were seen
            zero = true;
            mant = best == 0.0 ? 0.0 : __longlong_as_double(0x7ff8000000000000LL);
            break;
        }
        if (bi != k) {
            for (int j = lane; j < D; j += 32) {
                const double t = A[k * LD + j];
                A[k * LD + j] = A[bi * LD + j];
                A[bi * LD + j] = t;
            }
            mant = -mant;
        }
        __syncwarp();
        const double piv = A[k * LD + k];
        {
            int e2;
            const double m2 = frexp(piv, &e2);
            mant *= m2;
            expo += e2;
            int e3;
            mant = frexp(mant, &e3);
            expo += e3;
        }
        // multipliers l_i = a_ik * (1 / pivot)  (dgetf2: DSCAL by the reciprocal when |pivot| >= sfmin, else divide)
        const bool recip_ok = fabs(piv) >= 2.2250738585072014e-308;
        const double rp = 1.0 / piv;
        for (int i = k + 1 + lane; i < D; i += 32) A[i * LD + k] = recip_ok ? A[i * LD + k] * rp : A[i * LD + k] / piv;
        __syncwarp();
        // trailing update, lanes over columns
        const int w = D - k - 1;
        if (w <= 16) {
            // two rows per pass: lanes 0-15 take row i, lanes 16-31 row i + 1
            const int jj = k + 1 + (lane & 15), half = lane >> 4;
            const double u = (lane & 15) < w ? A[k * LD + jj] : 0.0;
            for (int i = k + 1 + half; i < D; i += 2)
                if ((lane & 15) < w) A[i * LD + jj] = fma(-A[i * LD + k], u, A[i * LD + jj]);
        } else {
            for (int jb = k + 1; jb < D; jb += 32) {
                const int j = jb + lane;
                const double u = j < D ? A[k * LD + j] : 0.0;
                for (int i = k + 1; i < D; ++i)
                    if (j < D) A[i * LD + j] = fma(-A[i * LD + k], u, A[i * LD + j]);
            }
        }
        __syncwarp();
    }
    if (lane == 0) det[c0 + warp] = zero ? mant : ldexp(mant, expo);
}

// ---- angle between two row vectors (vectorfield/utils.py:1148-1173) ----------------------------------------------
// one warp per cell: norms, dot product of the unit vectors, sign from the 2 x 2 determinant of their last two
// components (np.linalg.det = LU with partial pivoting; only its sign is used)
__global__ void __launch_bounds__(256)
k_cell_angles(const double* __restrict__ U, const double* __restrict__ V, int64_t n, int D, double* __restrict__ out) {
    const int64_t c = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (c >= n) return;
    const double* u = U + c * D;
    const double* v = V + c * D;
    double su = 0.0, sv = 0.0;
    for (int j = lane; j < D; j += 32) { su = fma(u[j], u[j], su); sv = fma(v[j], v[j], sv); }
    su = warp_sum(su);
    sv = warp_sum(sv);
    const double nu = sqrt(su), nv = sqrt(sv);
    double dp = 0.0;
    for (int j = lane; j < D; j += 32) dp = fma(u[j] / nu, v[j] / nv, dp);
    dp = warp_sum(dp);
    if (lane != 0) return;
    if (nu == 0.0 || nv == 0.0 || nu != nu || nv != nv) { out[c] = __longlong_as_double(0x7ff8000000000000LL); return; }
    const double a = u[D - 2] / nu, b = u[D - 1] / nu, cc = v[D - 2] / nv, d = v[D - 1] / nv;
    double minor;   // sign-exact restatement of the pivoted LU of [[a, b], [cc, d]]
    if (fabs(a) >= fabs(cc)) {
        if (a == 0.0) minor = 0.0;
        else { const double l = cc * (1.0 / a); minor = a * (d - l * b); }
    } else {
        const double l = a * (1.0 / cc);
        minor = -(cc * (b - l * d));
    }
    const double sign = minor == 0.0 ? 1.0 : (minor > 0.0 ? -1.0 : 1.0);
    dp = fmin(fmax(dp, -1.0), 1.0);
    out[c] = sign * acos(dp);
}

__global__ void __launch_bounds__(256)
k_row_norms(const double* __restrict__ V, int64_t n, int D, double* __restrict__ out) {
    const int64_t c = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (c >= n) return;
    double s = 0.0;
    for (int j = lane; j < D; j += 32) { const double x = V[c * D + j]; s = fma(x, x, s); }
    s = warp_sum(s);
    if (lane == 0) out[c] = sqrt(s);
}

// ---- Hessian of the Gaussian-RKHS field (vectorfield/utils.py:589-614) -------------------------------------------
//   H[i][l][m] = d^2 f_i / dx_l dx_m = -2 beta ( delta_lm v_i - 2 beta sum_j k_j C_ji d_jl d_jm ),  d_j = x - c_j
// One CTA per cell.  Control points stream through shared memory in chunks of HS_JC (differences d_j and the rows
// k_j C_j); a thread owns up to two (l <= m, block of 8 outputs i) units and keeps their 16 sums in registers; if
// D (D+1) / 2 * ceil(D / 8) exceeds 1024 units the cell is done in several rounds.  Output: H[cell][i][l][m].
constexpr int HS_JC = 32, HS_THREADS = 512, HS_IB = 8;

__global__ void __launch_bounds__(HS_THREADS)
k_hessian(const double* __restrict__ x, int64_t n, int D, const double* __restrict__ Xc, int M, double beta,
          const double* __restrict__ C, double* __restrict__ H) {
    extern __shared__ __align__(16) double hs_smem[];
    const int DP = D + 1, DI = (D + HS_IB - 1) / HS_IB * HS_IB;
    double* dj = hs_smem;                       // [HS_JC][DP]
    double* kc = dj + HS_JC * DP;               // [HS_JC][DI]
    double* xs = kc + HS_JC * DI;               // [D]
    double* vs = xs + D;                        // [D]
    double* kj = vs + D;                        // [HS_JC]
    const int64_t cell = blockIdx.x;
    const int tid = threadIdx.x;
    for (int d = tid; d < D; d += HS_THREADS) { xs[d] = x[cell * D + d]; vs[d] = 0.0; }
    const int n_pairs = D * (D + 1) / 2, n_ib = DI / HS_IB, n_units = n_pairs * n_ib;
    double* Hc = H + cell * (int64_t)D * D * D;
    for (int u0 = 0; u0 < n_units; u0 += 2 * HS_THREADS) {
        int ul[2], um[2], ui[2];
        bool live[2];
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            const int u = u0 + tid + q * HS_THREADS;
            live[q] = u < n_units;
            int pr = live[q] ? u / n_ib : 0, l = 0;
            ui[q] = live[q] ? (u % n_ib) * HS_IB : 0;
            while (pr >= D - l) { pr -= D - l; ++l; }
            ul[q] = l;
            um[q] = l + pr;
        }
        double acc[2][HS_IB];
#pragma unroll
        for (int q = 0; q < 2; ++q)
#pragma unroll
            for (int i = 0; i < HS_IB; ++i) acc[q][i] = 0.0;
        for (int j0 = 0; j0 < M; j0 += HS_JC) {
            __syncthreads();
            for (int e = tid; e < HS_JC * D; e += HS_THREADS) {
                const int j = e / D, d = e - j * D;
                dj[j * DP + d] = (j0 + j < M) ? xs[d] - Xc[(int64_t)(j0 + j) * D + d] : 0.0;
            }
            __syncthreads();
            if (tid < HS_JC) {
                double r2 = 0.0;
                for (int d = 0; d < D; ++d) r2 = fma(dj[tid * DP + d], dj[tid * DP + d], r2);
                kj[tid] = (j0 + tid < M) ? exp(-beta * r2) : 0.0;
            }
            __syncthreads();
            for (int e = tid; e < HS_JC * DI; e += HS_THREADS) {
                const int j = e / DI, i = e - j * DI;
                kc[e] = (i < D && j0 + j < M) ? kj[j] * C[(int64_t)(j0 + j) * D + i] : 0.0;
            }
            __syncthreads();
            if (u0 == 0 && tid < D) {          // v_i rides along in the first round
                double v = vs[tid];
                for (int j = 0; j < HS_JC; ++j) v += kc[j * DI + tid];
                vs[tid] = v;
            }
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                if (!live[q]) continue;
                for (int j = 0; j < HS_JC; ++j) {
                    const double pj = dj[j * DP + ul[q]] * dj[j * DP + um[q]];
                    const double* kr = kc + j * DI + ui[q];
#pragma unroll
                    for (int i = 0; i < HS_IB; i += 2) {
                        const double2 k2 = *reinterpret_cast<const double2*>(kr + i);
                        acc[q][i] = fma(k2.x, pj, acc[q][i]);
                        acc[q][i + 1] = fma(k2.y, pj, acc[q][i + 1]);
                    }
                }
            }
        }
        __syncthreads();                        // vs complete (first round) before it is read below
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            if (!live[q]) continue;
            const int l = ul[q], m = um[q];
#pragma unroll
            for (int i = 0; i < HS_IB; ++i) {
                const int io = ui[q] + i;
                if (io >= D) continue;
                const double h = -2.0 * beta * ((l == m ? vs[io] : 0.0) - 2.0 * beta * acc[q][i]);
                Hc[((int64_t)io * D + l) * D + m] = h;
                Hc[((int64_t)io * D + m) * D + l] = h;
            }
        }
    }
}

// ---- "Laplacian" as the reference computes it from its Hessian array (vectorfield/utils.py:665-687 applied to the
// Hs[i, l, m, cell] of vector_calculus.py:413-417): L[k] = sum_a Hs[a, a, k] = sum_a d^2 f_a / dx_a dx_k
//   = -2 beta sum_j k_j ( C_jk - 2 beta (C_j . d_j) d_jk ).   One warp per cell, lane = dimension (two if D > 32).
__global__ void __launch_bounds__(256)
k_laplacian(const double* __restrict__ x, int64_t n, int D, const double* __restrict__ Xc, int M, double beta,
            const double* __restrict__ C, double* __restrict__ L) {
    const int64_t cell = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (cell >= n) return;
    const int d0 = lane, d1 = lane + 32;
    const double x0 = d0 < D ? x[cell * D + d0] : 0.0, x1 = d1 < D ? x[cell * D + d1] : 0.0;
    double a0 = 0.0, a1 = 0.0;
    for (int j = 0; j < M; ++j) {
        const double* c = Xc + (int64_t)j * D;
        const double* cj = C + (int64_t)j * D;
        const double e0 = d0 < D ? x0 - c[d0] : 0.0, e1 = d1 < D ? x1 - c[d1] : 0.0;
        const double c0 = d0 < D ? cj[d0] : 0.0, c1 = d1 < D ? cj[d1] : 0.0;
        double r2 = fma(e1, e1, e0 * e0), q = fma(c1, e1, c0 * e0);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            r2 += __shfl_xor_sync(0xffffffffu, r2, o);
            q += __shfl_xor_sync(0xffffffffu, q, o);
        }
        const double k = exp(-beta * r2), w = 2.0 * beta * q;
        a0 = fma(k, c0 - w * e0, a0);
        a1 = fma(k, c1 - w * e1, a1);
    }
    if (d0 < D) L[cell * D + d0] = -2.0 * beta * a0;
    if (d1 < D) L[cell * D + d1] = -2.0 * beta * a1;
}

}  // namespace dvf

using namespace dvf;

extern "C" int dvf_jacobian_det(const double* J, int64_t n, int D, double* det, void* stream) {
    DVF_ARG(n >= 0 && D >= 1 && D <= kMaxD, "bad argument");
    if (n == 0) return DVF_OK;
    DVF_ARG(J && det, "null pointer");
    const size_t mat = (size_t)det_mat_stride(D) * 8;
    int cpb = (int)((100 * 1024) / mat) & ~3;     // two blocks per SM: one loads while the other factorises
    if (cpb < 4) cpb = 4;
    if (cpb > 16) cpb = 16;
    const size_t smem = mat * cpb;
    static size_t set_smem = 0;
    if (smem > set_smem) {
        DVF_CUDA(cudaFuncSetAttribute(k_jacobian_det, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        set_smem = smem;
    }
    const int64_t blocks = (n + cpb - 1) / cpb;
    DVF_ARG(blocks <= 0x7fffffff, "too many cells for one launch");
    k_jacobian_det<<<(unsigned)blocks, cpb * 32, smem, (cudaStream_t)stream>>>(J, n, D, cpb, det);
    DVF_LAUNCH_CHECK();
    return DVF_OK;
}

extern "C" int dvf_cell_angles(const double* U, const double* V, int64_t n, int D, double* out, void* stream) {
    DVF_ARG(n >= 0 && D >= 2, "bad argument (the reference's angle needs at least two components)");
    if (n == 0) return DVF_OK;
    DVF_ARG(U && V && out, "null pointer");
    const int64_t blocks = (n + 7) / 8;
    DVF_ARG(blocks <= 0x7fffffff, "too many cells for one launch");
    k_cell_angles<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(U, V, n, D, out);
    DVF_LAUNCH_CHECK();
    return DVF_OK;
}

extern "C" int dvf_row_norms(const double* V, int64_t n, int D, double* out, void* stream) {
    DVF_ARG(n >= 0 && D >= 1, "bad argument");
    if (n == 0) return DVF_OK;
    DVF_ARG(V && out, "null pointer");
    const int64_t blocks = (n + 7) / 8;
    DVF_ARG(blocks <= 0x7fffffff, "too many cells for one launch");
    k_row_norms<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(V, n, D, out);
    DVF_LAUNCH_CHECK();
    return DVF_OK;
}

extern "C" int dvf_hessian(const double* x, int64_t n, int D, const double* Xc, int M, double beta, const double* C,
                           double* H, void* stream) {
    DVF_ARG(n >= 0 && D >= 1 && D <= kMaxD && M >= 1, "bad argument");
    if (n == 0) return DVF_OK;
    DVF_ARG(x && Xc && C && H, "null pointer");
    DVF_ARG(n <= 0x7fffffff, "too many cells for one launch");
    const int DP = D + 1, DI = (D + HS_IB - 1) / HS_IB * HS_IB;
    const size_t smem = sizeof(double) * ((size_t)HS_JC * DP + (size_t)HS_JC * DI + 2 * D + HS_JC);
    k_hessian<<<(unsigned)n, HS_THREADS, smem, (cudaStream_t)stream>>>(x, n, D, Xc, M, beta, C, H);
    DVF_LAUNCH_CHECK();
    return DVF_OK;
}

extern "C" int dvf_laplacian(const double* x, int64_t n, int D, const double* Xc, int M, double beta, const double* C,
                             double* L, void* stream) {
    DVF_ARG(n >= 0 && D >= 1 && D <= kMaxD && M >= 1, "bad argument");
    if (n == 0) return DVF_OK;
    DVF_ARG(x && Xc && C && L, "null pointer");
    const int64_t blocks = (n + 7) / 8;
    DVF_ARG(blocks <= 0x7fffffff, "too many cells for one launch");
    k_laplacian<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(x, n, D, Xc, M, beta, C, L);
    DVF_LAUNCH_CHECK();
    return DVF_OK;
}

// ---------------------------------------------------------------------------------------------------------
// bandwidth_selector (reference scVectorField.py:121-135): distances of every control point to its k nearest control
// points (itself included), ascending.  One CTA per control point: the M squared distances -- accumulated dimension by
// dimension with separate multiply and add, the order and rounding sklearn's trees use -- go to shared memory, a
// bitonic network sorts them, the first k + 1 leave as square roots.  Replaces 3 D eager launches + a library top-k.
// ---------------------------------------------------------------------------------------------------------
namespace dvf {
__global__ void __launch_bounds__(256) k_knn_dists(const double* __restrict__ Xc, int M, int D, int kk, int P2,
                                                   double* __restrict__ out) {
    extern __shared__ double d2[];
    const int i = blockIdx.x;
    const double inf = __longlong_as_double(0x7ff0000000000000LL);
    for (int j = threadIdx.x; j < P2; j += blockDim.x) {
        double s = inf;
        if (j < M) {
            s = 0.0;
            for (int d = 0; d < D; ++d) {
                const double t = __dsub_rn(Xc[(int64_t)i * D + d], Xc[(int64_t)j * D + d]);
                s = __dadd_rn(s, __dmul_rn(t, t));
            }
        }
        d2[j] = s;
    }
    __syncthreads();
    for (int k = 2; k <= P2; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int t = threadIdx.x; t < P2; t += blockDim.x) {
                const int p = t ^ j;
                if (p > t) {
                    const double a = d2[t], b = d2[p];
                    const bool up = (t & k) == 0;
                    if ((a > b) == up) { d2[t] = b; d2[p] = a; }
                }
            }
            __syncthreads();
        }
    for (int j = threadIdx.x; j < kk; j += blockDim.x) out[(int64_t)i * kk + j] = sqrt(d2[j]);
}
}  // namespace dvf

extern "C" int dvf_knn_dists(const double* Xc, int M, int D, int kk, double* out, void* stream) {
    DVF_ARG(Xc && out && M >= 1 && D >= 1 && kk >= 1 && kk <= M, "bad argument");
    int P2 = 1;
    while (P2 < M) P2 <<= 1;
    DVF_ARG(P2 <= 16384, "dvf_knn_dists serves up to 16384 control points (M = %d)", M);
    const size_t smem = sizeof(double) * (size_t)P2;
    if (smem > 48 * 1024) {
        static bool attr = false;
        if (!attr) {
            DVF_CUDA(cudaFuncSetAttribute(dvf::k_knn_dists, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 1024));
            attr = true;
        }
    }
    dvf::k_knn_dists<<<M, 256, smem, (cudaStream_t)stream>>>(Xc, M, D, kk, P2, out);
    DVF_LAUNCH_CHECK();
    return DVF_OK;
}

// ---------------------------------------------------------------------------------------------------------
// Cell reductions of the MATERIALISED gene-space Jacobian (Ge, Gr, n) float32 in NumPy's own float32 orders, so that
// the means rank_jacobian_genes sorts (reference rank_vf.py:483-508) come out bit for bit -- north_star: "top-k
// regulator rankings must match exactly".
//   dvf_f32_row_means_pairwise : np.mean(J, axis=2) of a C-contiguous float32 array = (0 + pairwise_sum(row)) / n in
//       float32.  pairwise_sum (numpy/_core/src/umath/loops_utils.h): n < 8 straight loop; n <= 128 eight interleaved
//       accumulators combined ((r0+r1)+(r2+r3))+((r4+r5)+(r6+r7)) plus a sequential tail; else halves split at
//       n/2 rounded down to a multiple of 8, recursively.  The tree only depends on n: the host lists its leaves and,
//       per leaf, how many pending additions close after it; a warp sums 32 leaves at a time (one per lane) and lane 0
//       replays the additions on a small stack.
//   dvf_f32_group_means_seq : average_jacobian_by_group (vectorfield/utils.py:557-582): per group a running float32
//       sum over the cells in order, divided by the count.  One thread per (effector, regulator) row walks the cells
//       sequentially; rows go through shared memory in 128 x 32 tiles so that the global loads are coalesced.
// ---------------------------------------------------------------------------------------------------------
namespace dvf {

__device__ __forceinline__ float f32_leaf_sum(const float* __restrict__ a, int n) {
    if (n < 8) {
        float res = 0.0f;
        for (int i = 0; i < n; ++i) res = __fadd_rn(res, a[i]);
        return res;
    }
    float r[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) r[j] = a[j];
    int i = 8;
    for (; i < n - (n % 8); i += 8) {
#pragma unroll
        for (int j = 0; j < 8; ++j) r[j] = __fadd_rn(r[j], a[i + j]);
    }
    float res = __fadd_rn(__fadd_rn(__fadd_rn(r[0], r[1]), __fadd_rn(r[2], r[3])),
                          __fadd_rn(__fadd_rn(r[4], r[5]), __fadd_rn(r[6], r[7])));
    for (; i < n; ++i) res = __fadd_rn(res, a[i]);
    return res;
}

// leaves: (lo, len, adds_after) triples; one warp per row
__global__ void __launch_bounds__(256) k_f32_row_means_pairwise(const float* __restrict__ J, int64_t R, int64_t n,
                                                                const int* __restrict__ leaves, int n_leaves,
                                                                float* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= R) return;
    const float* a = J + row * n;
    float stack[40];
    int sp = 0;
    for (int l0 = 0; l0 < n_leaves; l0 += 32) {
        const int k = l0 + lane;
        float s = 0.0f;
        int adds = 0;
        if (k < n_leaves) {
            s = f32_leaf_sum(a + leaves[3 * k], leaves[3 * k + 1]);
            adds = leaves[3 * k + 2];
        }
        const int cnt = min(32, n_leaves - l0);
        for (int q = 0; q < cnt; ++q) {
            const float v = __shfl_sync(0xffffffffu, s, q);
            const int na = __shfl_sync(0xffffffffu, adds, q);
            if (lane == 0) {
                stack[sp++] = v;
                for (int t = 0; t < na; ++t) { const float b = stack[--sp], c = stack[--sp]; stack[sp++] = __fadd_rn(c, b); }
            }
        }
    }
    if (lane == 0) out[row] = __fdiv_rn(__fadd_rn(0.0f, stack[0]), (float)n);
}

constexpr int GS_ROWS = 128, GS_CELLS = 32;
__global__ void __launch_bounds__(GS_ROWS) k_f32_group_means_seq(const float* __restrict__ J, int64_t R, int64_t n,
                                                                 const int* __restrict__ label, int G,
                                                                 const float* __restrict__ count, float* __restrict__ out) {
    extern __shared__ float gs_smem[];
    float (*tile)[GS_CELLS + 1] = reinterpret_cast<float (*)[GS_CELLS + 1]>(gs_smem);       // [GS_ROWS][33]
    float* acc = gs_smem + GS_ROWS * (GS_CELLS + 1);                                         // [G][GS_ROWS]
    int* lab = reinterpret_cast<int*>(acc + (size_t)G * GS_ROWS);                            // [GS_CELLS]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t row0 = (int64_t)blockIdx.x * GS_ROWS;
    for (int g = 0; g < G; ++g) acc[g * GS_ROWS + tid] = 0.0f;
    for (int64_t c0 = 0; c0 < n; c0 += GS_CELLS) {
        __syncthreads();
        // coalesced: a warp loads 32 consecutive cells of one row per request
        for (int r = warp; r < GS_ROWS; r += GS_ROWS / 32) {
            const int64_t row = row0 + r, c = c0 + lane;
            tile[r][lane] = (row < R && c < n) ? J[row * n + c] : 0.0f;
        }
        if (tid < GS_CELLS) lab[tid] = (c0 + tid < n) ? label[c0 + tid] : -1;
        __syncthreads();
#pragma unroll 8
        for (int c = 0; c < GS_CELLS; ++c) {
            const int g = lab[c];
            if (g >= 0) acc[g * GS_ROWS + tid] = __fadd_rn(acc[g * GS_ROWS + tid], tile[tid][c]);
        }
    }
    if (row0 + tid < R)
        for (int g = 0; g < G; ++g) out[(int64_t)g * R + row0 + tid] = __fdiv_rn(acc[g * GS_ROWS + tid], count[g]);
}

}  // namespace dvf

extern "C" int dvf_f32_row_means_pairwise(const float* J, int64_t R, int64_t n, const int* leaves, int n_leaves,
                                          float* out, void* stream) {
    DVF_ARG(R >= 0 && n >= 1 && n_leaves >= 1, "bad argument");
    if (R == 0) return DVF_OK;
    DVF_ARG(J && leaves && out, "null pointer");
    const int64_t blocks = (R + 7) / 8;
    DVF_ARG(blocks <= 0x7fffffff, "too many rows for one launch");
    dvf::k_f32_row_means_pairwise<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(J, R, n, leaves, n_leaves, out);
    DVF_LAUNCH_CHECK();
    return DVF_OK;
}

extern "C" int dvf_f32_group_means_seq(const float* J, int64_t R, int64_t n, const int* label, int G, const float* count,
                                       float* out, void* stream) {
    DVF_ARG(R >= 0 && n >= 1 && G >= 1, "bad argument");
    if (R == 0) return DVF_OK;
    DVF_ARG(J && label && count && out, "null pointer");
    const size_t smem = sizeof(float) * ((size_t)dvf::GS_ROWS * (dvf::GS_CELLS + 1) + (size_t)G * dvf::GS_ROWS) + sizeof(int) * dvf::GS_CELLS;
    DVF_ARG(smem <= 200 * 1024, "too many groups for dvf_f32_group_means_seq (%d)", G);
    if (smem > 48 * 1024) DVF_CUDA(cudaFuncSetAttribute(dvf::k_f32_group_means_seq, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t blocks = (R + dvf::GS_ROWS - 1) / dvf::GS_ROWS;
    DVF_ARG(blocks <= 0x7fffffff, "too many rows for one launch");
    dvf::k_f32_group_means_seq<<<(unsigned)blocks, dvf::GS_ROWS, smem, (cudaStream_t)stream>>>(J, R, n, label, G, count, out);
    DVF_LAUNCH_CHECK();
    return DVF_OK;
}

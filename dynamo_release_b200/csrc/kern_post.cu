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

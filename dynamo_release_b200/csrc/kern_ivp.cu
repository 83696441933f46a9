// K11: batched initial-value integration through the RKHS field -- one CTA per initial state runs the whole
// adaptive Dormand-Prince 5(4) integration on the device (step control, terminal "at rest" event, dense-output
// coefficients), replacing the per-cell scipy.integrate.solve_ivp loop of reference prediction/utils.py:103-165
// (integrate_vf_ivp) whose right-hand side is vector_field_function (vectorfield/utils.py:73-142).
//
// Trajectories are independent and each one is a strictly sequential chain of field evaluations
// f(y) = scale * sum_m exp(-beta |y - c_m|^2) C[m, :], so the parallelism is (a) across trajectories -- one CTA
// each, several resident per SM -- and (b) inside one evaluation: the control points are split over the CTA's
// threads for the M kernel values, then over IVP_PARTS row groups x D columns for the product with C.  The control
// points are read through a transposed copy (coalesced over m) and C row-wise (coalesced over d); both stay in L1/L2
// for the whole integration.  All sums run in a fixed order, so a trajectory does not depend on the launch shape.
#include "common.cuh"

namespace dvf {

constexpr int IVP_THREADS = 128;
constexpr int IVP_PARTS = 4;      // row groups of the k^T C product (IVP_THREADS / 32)

// Dormand-Prince tableau, written as the same quotients scipy evaluates (scipy/integrate/_ivp/rk.py RK45)
__constant__ double c_A[6][5] = {
    {0, 0, 0, 0, 0},
    {1.0 / 5, 0, 0, 0, 0},
    {3.0 / 40, 9.0 / 40, 0, 0, 0},
    {44.0 / 45, -56.0 / 15, 32.0 / 9, 0, 0},
    {19372.0 / 6561, -25360.0 / 2187, 64448.0 / 6561, -212.0 / 729, 0},
    {9017.0 / 3168, -355.0 / 33, 46732.0 / 5247, 49.0 / 176, -5103.0 / 18656}};
__constant__ double c_B[6] = {35.0 / 384, 0, 500.0 / 1113, 125.0 / 192, -2187.0 / 6784, 11.0 / 84};
__constant__ double c_E[7] = {-71.0 / 57600, 0, 71.0 / 16695, -71.0 / 1920, 17253.0 / 339200, -22.0 / 525, 1.0 / 40};
__constant__ double c_P[7][4] = {
    {1, -8048581381.0 / 2820520608.0, 8663915743.0 / 2820520608.0, -12715105075.0 / 11282082432.0},
    {0, 0, 0, 0},
    {0, 131558114200.0 / 32700410799.0, -68118460800.0 / 10900136933.0, 87487479700.0 / 32700410799.0},
    {0, -1754552775.0 / 470086768.0, 14199869525.0 / 1410260304.0, -10690763975.0 / 1880347072.0},
    {0, 127303824393.0 / 49829197408.0, -318862633887.0 / 49829197408.0, 701980252875.0 / 199316789632.0},
    {0, -282668133.0 / 205662961.0, 2019193451.0 / 616988883.0, -1453857185.0 / 822651844.0},
    {0, 40617522.0 / 29380423.0, -110615467.0 / 29380423.0, 69997945.0 / 29380423.0}};

struct IvpShared {
    double y[kMaxD], yt[kMaxD], ynew[kMaxD], buf[kMaxD];
    double K[7][kMaxD];
    double part[IVP_PARTS][kMaxD];
};

// f(xin) -> fout (both in shared memory).  xin must be visible to the CTA on entry; fout is visible on return.
__device__ __forceinline__ void ivp_field(const double* xin, double* fout, IvpShared& s, double* kv, int D, int M,
                                          const double* __restrict__ XcT, const double* __restrict__ Cm, double beta,
                                          double scale) {
    const int tid = threadIdx.x;
    for (int mb = tid; mb < M; mb += 4 * IVP_THREADS) {     // four control points in flight per thread
        int mi[4];
        double d2[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
        for (int u = 0; u < 4; ++u) mi[u] = min(mb + u * IVP_THREADS, M - 1);
        for (int d = 0; d < D; ++d) {
            const double xv = xin[d];
            const double* row = XcT + (int64_t)d * M;
#pragma unroll
            for (int u = 0; u < 4; ++u) { const double t = xv - row[mi[u]]; d2[u] = fma(t, t, d2[u]); }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (mb + u * IVP_THREADS < M) kv[mb + u * IVP_THREADS] = exp(-beta * d2[u]);
    }
    __syncthreads();
    const int lane = tid & 31, p = tid >> 5;
    const int per = (M + IVP_PARTS - 1) / IVP_PARTS;
    const int m0 = p * per, m1 = min(M, m0 + per);
    const bool two = lane + 32 < D;
    if (lane < D) {
        // four interleaved partial sums per column (m mod 4), combined in a fixed order
        double a[4] = {0.0, 0.0, 0.0, 0.0}, b[4] = {0.0, 0.0, 0.0, 0.0};
        int m = m0;
        for (; m + 4 <= m1; m += 4) {
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const double k = kv[m + u];
                a[u] = fma(k, Cm[(int64_t)(m + u) * D + lane], a[u]);
                if (two) b[u] = fma(k, Cm[(int64_t)(m + u) * D + lane + 32], b[u]);
            }
        }
        for (; m < m1; ++m) {
            const double k = kv[m];
            a[0] = fma(k, Cm[(int64_t)m * D + lane], a[0]);
            if (two) b[0] = fma(k, Cm[(int64_t)m * D + lane + 32], b[0]);
        }
        s.part[p][lane] = (a[0] + a[1]) + (a[2] + a[3]);
        if (two) s.part[p][lane + 32] = (b[0] + b[1]) + (b[2] + b[3]);
    }
    __syncthreads();
    if (tid < D) {
        double v = s.part[0][tid];
#pragma unroll
        for (int q = 1; q < IVP_PARTS; ++q) v += s.part[q][tid];
        fout[tid] = scale * v;
    }
    __syncthreads();
}

// RMS norm of s.buf[0..D): every thread runs the same ordered sum over shared memory, so all of them hold the
// identical value and take the same branch afterwards.  s.buf must be visible on entry.
__device__ __forceinline__ double ivp_rms(const IvpShared& s, int D) {
    double a = 0.0;
    for (int d = 0; d < D; ++d) a = fma(s.buf[d], s.buf[d], a);
    return sqrt(a) / sqrt((double)D);
}
__device__ __forceinline__ bool ivp_at_rest(const double* f, int D, double tol) {
    bool r = true;
    for (int d = 0; d < D; ++d) r = r && (fabs(f[d]) < tol);
    return r;
}

__global__ void __launch_bounds__(IVP_THREADS) k_ivp_rk45(const double* __restrict__ y0s, int64_t n, int D,
                                                          const double* __restrict__ XcT, int M, double beta,
                                                          const double* __restrict__ Cm, double scale, double t0,
                                                          double t_bound, double max_step, double rtol, double atol,
                                                          double rest_tol, int cap, double* __restrict__ ts,
                                                          double* __restrict__ ys, double* __restrict__ Qs,
                                                          int* __restrict__ n_steps, int* __restrict__ status) {
    extern __shared__ double kv[];       // M kernel values
    __shared__ IvpShared s;
    const int tid = threadIdx.x;
    const int64_t traj = blockIdx.x;
    double* ts_o = ts + traj * (int64_t)(cap + 2);
    double* ys_o = ys + traj * (int64_t)(cap + 1) * D;
    double* Q_o = Qs + traj * (int64_t)cap * D * 4;

    const double dir = (t_bound != t0) ? ((t_bound > t0) ? 1.0 : -1.0) : 1.0;
    double t = t0;
    if (tid < D) {
        s.y[tid] = y0s[traj * D + tid];
        ys_o[tid] = s.y[tid];
    }
    if (tid == 0) ts_o[0] = t0;
    __syncthreads();
    ivp_field(s.y, s.K[0], s, kv, D, M, XcT, Cm, beta, scale);
    bool rest_old = ivp_at_rest(s.K[0], D, rest_tol);

    // ---- initial step (Hairer, Norsett, Wanner II.4 as in scipy's select_initial_step) ----
    const double span = fabs(t_bound - t0);
    double h_abs;
    {
        double sc = 0.0;
        if (tid < D) { sc = atol + fabs(s.y[tid]) * rtol; s.buf[tid] = s.y[tid] / sc; }
        __syncthreads();
        const double d0 = ivp_rms(s, D);
        __syncthreads();
        if (tid < D) s.buf[tid] = s.K[0][tid] / sc;
        __syncthreads();
        const double d1 = ivp_rms(s, D);
        double h0 = (d0 < 1e-5 || d1 < 1e-5) ? 1e-6 : 0.01 * d0 / d1;
        h0 = fmin(h0, span);
        if (tid < D) s.yt[tid] = s.y[tid] + (h0 * dir) * s.K[0][tid];
        __syncthreads();
        ivp_field(s.yt, s.K[1], s, kv, D, M, XcT, Cm, beta, scale);
        if (tid < D) s.buf[tid] = (s.K[1][tid] - s.K[0][tid]) / sc;
        __syncthreads();
        const double d2 = ivp_rms(s, D) / h0;
        const double h1 = (d1 <= 1e-15 && d2 <= 1e-15) ? fmax(1e-6, h0 * 1e-3) : pow(0.01 / fmax(d1, d2), 0.2);
        h_abs = fmin(fmin(100.0 * h0, h1), fmin(span, max_step));
        __syncthreads();
    }

    int nst = 0, st = 0;
    bool done = (span == 0.0);
    while (!done) {
        const double min_step = 10.0 * fabs(nextafter(t, dir * __longlong_as_double(0x7ff0000000000000LL)) - t);
        if (h_abs > max_step) h_abs = max_step;
        else if (h_abs < min_step) h_abs = min_step;
        bool rejected = false, failed = false;
        double h = 0.0, t_new = t;
        for (;;) {
            if (h_abs < min_step) { failed = true; break; }
            h = h_abs * dir;
            t_new = t + h;
            if (dir * (t_new - t_bound) > 0.0) t_new = t_bound;
            h = t_new - t;
            h_abs = fabs(h);
            for (int sg = 1; sg < 6; ++sg) {
                if (tid < D) {
                    double dy = 0.0;
                    for (int j = 0; j < sg; ++j) dy += s.K[j][tid] * c_A[sg][j];
                    s.yt[tid] = s.y[tid] + dy * h;
                }
                __syncthreads();
                ivp_field(s.yt, s.K[sg], s, kv, D, M, XcT, Cm, beta, scale);
            }
            if (tid < D) {
                double dy = 0.0;
#pragma unroll
                for (int j = 0; j < 6; ++j) dy += s.K[j][tid] * c_B[j];
                s.ynew[tid] = s.y[tid] + h * dy;
            }
            __syncthreads();
            ivp_field(s.ynew, s.K[6], s, kv, D, M, XcT, Cm, beta, scale);
            if (tid < D) {
                double e = 0.0;
#pragma unroll
                for (int j = 0; j < 7; ++j) e += s.K[j][tid] * c_E[j];
                const double sc = atol + fmax(fabs(s.y[tid]), fabs(s.ynew[tid])) * rtol;
                s.buf[tid] = e * h / sc;
            }
            __syncthreads();
            const double err = ivp_rms(s, D);
            __syncthreads();   // buf is rewritten by the next attempt
            if (err < 1.0) {
                double factor = (err == 0.0) ? 10.0 : fmin(10.0, 0.9 * pow(err, -0.2));
                if (rejected) factor = fmin(1.0, factor);
                h_abs *= factor;
                break;
            }
            h_abs *= fmax(0.2, 0.9 * pow(err, -0.2));
            rejected = true;
        }
        if (failed) { st = -1; break; }

        // accepted: dense-output coefficients Q = K^T P of this step, then advance
        double q[4] = {0.0, 0.0, 0.0, 0.0};
        if (tid < D) {
#pragma unroll
            for (int j = 0; j < 7; ++j) {
                const double k = s.K[j][tid];
#pragma unroll
                for (int c = 0; c < 4; ++c) q[c] += k * c_P[j][c];
            }
            double* qo = Q_o + ((int64_t)nst * D + tid) * 4;
            qo[0] = q[0]; qo[1] = q[1]; qo[2] = q[2]; qo[3] = q[3];
        }
        const double t_old = t;
        t = t_new;
        if (dir * (t - t_bound) >= 0.0) done = true;
        const bool rest_new = ivp_at_rest(s.K[6], D, rest_tol);   // K[6] = f(y_new), visible since ivp_field returned
        if (rest_new) {
            // terminal event: the two-valued event function makes the bracketing root finder return an end point
            const double root = rest_old ? t_old : t;
            const double x = (root - t_old) / h;
            if (tid < D) {
                const double p1 = x, p2 = p1 * x, p3 = p2 * x, p4 = p3 * x;
                const double yo = s.y[tid];
                s.ynew[tid] = h * (((q[0] * p1 + q[1] * p2) + q[2] * p3) + q[3] * p4) + yo;
            }
            t = root;
            st = rest_old ? -3 : 1;
            done = true;
        }
        __syncthreads();
        if (tid < D) {
            s.y[tid] = s.ynew[tid];
            s.K[0][tid] = s.K[6][tid];
            ys_o[(int64_t)(nst + 1) * D + tid] = s.ynew[tid];
        }
        if (tid == 0) {
            ts_o[nst + 1] = t;
            if (st == -3) ts_o[nst + 2] = t_new;    // the end of the step the only interpolant was built on
        }
        ++nst;
        rest_old = rest_new;
        __syncthreads();
        if (!done && nst == cap) { st = -2; break; }
    }
    if (tid == 0) { n_steps[traj] = nst; status[traj] = st; }
}

__global__ void k_ivp_transpose(const double* __restrict__ Xc, int M, int D, double* __restrict__ XcT) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < M * D) { const int m = i / D, d = i % D; XcT[(int64_t)d * M + m] = Xc[i]; }
}

// Dense-output evaluation of finished trajectories on the device (scipy OdeSolution.__call__ over RkDenseOutput pieces):
// row r of t_eval belongs to trajectory traj[r]; the segment is chosen with numpy.searchsorted's rules (side "left" on
// ascending times, "right" on the reversed array for descending ones, clipped to the first / last segment), then
// y = ys[seg] + h (Q[seg] [x, x^2, x^3, x^4]).  One CTA per row; out is (m, D, S), sample index fastest.
__global__ void __launch_bounds__(128) k_ivp_dense_eval(const double* __restrict__ ts, const double* __restrict__ ys,
                                                        const double* __restrict__ Qs, const int* __restrict__ n_steps,
                                                        const int* __restrict__ status, int D, int cap,
                                                        const int* __restrict__ traj, const double* __restrict__ t_eval,
                                                        const int* __restrict__ n_eval, int S, double* __restrict__ out) {
    const int r = blockIdx.x;
    const int64_t i = traj[r];
    const int k = n_steps[i];
    const int ne = n_eval ? n_eval[r] : S;
    const double* t_i = ts + i * (int64_t)(cap + 2);
    const double* y_i = ys + i * (int64_t)(cap + 1) * D;
    const double* q_i = Qs + i * (int64_t)cap * D * 4;
    const bool asc = t_i[k] >= t_i[0];
    const bool at_start = status[i] == -3;
    for (int e = threadIdx.x; e < ne * D; e += blockDim.x) {
        const int j = e / D, d = e % D;
        const double t = t_eval[(int64_t)r * S + j];
        // cnt = #{ts < t} (ascending, side left)  or  #{ts <= t} (descending: side right on the reversed array)
        int lo = 0, hi = k + 1;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            const double v = asc ? t_i[mid] : t_i[k - mid];
            if (asc ? (v < t) : (v <= t)) lo = mid + 1; else hi = mid;
        }
        int seg = min(max(lo - 1, 0), k - 1);
        if (!asc) seg = k - 1 - seg;
        const double h = at_start ? (t_i[2] - t_i[0]) : (t_i[seg + 1] - t_i[seg]);
        const double x = (t - t_i[seg]) / h;
        const double p1 = x, p2 = p1 * x, p3 = p2 * x, p4 = p3 * x;
        const double* q = q_i + ((int64_t)seg * D + d) * 4;
        out[((int64_t)r * D + d) * S + j] = h * (((q[0] * p1 + q[1] * p2) + q[2] * p3) + q[3] * p4) + y_i[(int64_t)seg * D + d];
    }
}

}  // namespace dvf

using namespace dvf;

extern "C" int dvf_ivp_workspace_bytes(int D, int M, size_t* out) {
    DVF_ARG(out && D >= 1 && D <= kMaxD && M >= 1, "dvf_ivp_workspace_bytes: bad arguments");
    *out = (size_t)D * M * sizeof(double);
    return DVF_OK;
}

extern "C" int dvf_integrate_rk45(const double* y0, int64_t n, int D, const double* Xc, int M, double beta,
                                  const double* C, double scale, double t0, double t_bound, double max_step,
                                  double rtol, double atol, double rest_tol, int cap, double* ts, double* ys,
                                  double* Q, int* n_steps, int* status, void* ws, size_t ws_bytes, void* stream) {
    DVF_ARG(D >= 1 && D <= kMaxD, "dvf_integrate_rk45: D = %d outside 1..%d", D, kMaxD);
    DVF_ARG(M >= 1 && M <= 5000, "dvf_integrate_rk45: M = %d outside 1..5000", M);
    DVF_ARG(n >= 0 && cap >= 1, "dvf_integrate_rk45: bad n / cap");
    DVF_ARG(max_step > 0.0 && rtol > 0.0 && atol > 0.0, "dvf_integrate_rk45: max_step, rtol, atol must be positive");
    if (n == 0) return DVF_OK;
    DVF_ARG(y0 && Xc && C && ts && ys && Q && n_steps && status && ws, "dvf_integrate_rk45: NULL pointer");
    if (ws_bytes < (size_t)D * M * sizeof(double)) {
        set_error("dvf_integrate_rk45: workspace %zu < %zu bytes", ws_bytes, (size_t)D * M * sizeof(double));
        return DVF_ERR_WORKSPACE;
    }
    DVF_ARG(n <= 0x7fffffff, "dvf_integrate_rk45: at most 2^31 - 1 initial states per call");
    cudaStream_t st = (cudaStream_t)stream;
    double* XcT = (double*)ws;
    k_ivp_transpose<<<(M * D + 255) / 256, 256, 0, st>>>(Xc, M, D, XcT);
    DVF_LAUNCH_CHECK();
    const size_t smem = (size_t)M * sizeof(double);
    k_ivp_rk45<<<(unsigned)n, IVP_THREADS, smem, st>>>(y0, n, D, XcT, M, beta, C, scale, t0, t_bound, max_step, rtol,
                                                       atol, rest_tol, cap, ts, ys, Q, n_steps, status);
    DVF_LAUNCH_CHECK();
    return DVF_OK;
}

extern "C" int dvf_ivp_dense_eval(const double* ts, const double* ys, const double* Q, const int* n_steps,
                                  const int* status, int D, int cap, const int* traj, const double* t_eval,
                                  const int* n_eval, int64_t m, int S, double* out, void* stream) {
    DVF_ARG(D >= 1 && D <= kMaxD && cap >= 1 && S >= 0 && m >= 0, "dvf_ivp_dense_eval: bad sizes");
    if (m == 0 || S == 0) return DVF_OK;
    DVF_ARG(ts && ys && Q && n_steps && status && traj && t_eval && out, "dvf_ivp_dense_eval: NULL pointer");
    DVF_ARG(m <= 0x7fffffff, "dvf_ivp_dense_eval: at most 2^31 - 1 rows per call");
    k_ivp_dense_eval<<<(unsigned)m, 128, 0, (cudaStream_t)stream>>>(ts, ys, Q, n_steps, status, D, cap, traj, t_eval,
                                                                    n_eval, S, out);
    DVF_LAUNCH_CHECK();
    return DVF_OK;
}

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
#include <cstdlib>

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
    extern __shared__ __align__(16) double kv[];       // M kernel values
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
        if (rest_new || rest_old) {
            // terminal event: the two-valued event function (0 at rest, -1 otherwise) makes the bracketing root finder
            // return an end point.  scipy's find_active_events fires on -1 -> 0 (root = t), 0 -> 0 and 0 -> -1 (a start at
            // rest: root = t_old, so the result is t = [t0, t0]); the last two can only happen on the first step.
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

// ------------------------------------------------------------------------------------------------------------------
// T trajectories per CTA (k_ivp_rk45_t).  The one-trajectory kernel above streams Xc^T and C through L1 once per field
// evaluation with one memory instruction per FMA (ncu: 64 % of the stall samples wait on shared / L1 operands, FP64 pipe
// 18 % busy).  Here a CTA advances T trajectories in lock step -- one attempted step of every unfinished trajectory per
// round, each with its own t, h and accept / reject decision -- so that every control-point coordinate and every
// coefficient fetched from L1 feeds T FMAs, the trajectory states and kernel values being read from shared memory as
// 16-byte vectors ([element][trajectory] layout).  Every sum is formed in the SAME order as in k_ivp_rk45 (sequential
// over d per kernel value; per quarter of the control points four interleaved partial sums combined as
// (a0 + a1) + (a2 + a3), quarters added in turn), so both kernels return identical trajectories.
// ------------------------------------------------------------------------------------------------------------------
constexpr int IVP_T = 4;

struct __align__(16) IvpSharedT {
    double y[kMaxD * IVP_T], yt[kMaxD * IVP_T], ynew[kMaxD * IVP_T], buf[kMaxD * IVP_T];
    double K[7][kMaxD * IVP_T];
    double part[8][kMaxD * IVP_T];
};

template <typename X>
__device__ __forceinline__ X ivp_pick(const X (&a)[IVP_T], int i) {
    X v = a[0];
#pragma unroll
    for (int k = 1; k < IVP_T; ++k) v = (i == k) ? a[k] : v;
    return v;
}

// rows [m0, m1) of k^T C for the column pair (2cp, 2cp+1) and the interleave classes c0 .. c0+NC-1 (class = (m - m0) mod 4;
// the rows left over after the last full group of four belong to class 0)
template <int NC>
__device__ __forceinline__ void ivp_phase2_t(const double* __restrict__ Cp, int Dp, const double* kv, int m0, int m1,
                                             int c0, int cp, double* out /* part[g] */) {
    constexpr int T = IVP_T;
    double acc[NC][2][T];
#pragma unroll
    for (int u = 0; u < NC; ++u)
#pragma unroll
        for (int tr = 0; tr < T; ++tr) { acc[u][0][tr] = 0.0; acc[u][1][tr] = 0.0; }
    const int main_end = (m1 > m0) ? m0 + ((m1 - m0) / 4) * 4 : m0;
    for (int m = m0; m < main_end; m += 4) {
#pragma unroll
        for (int u = 0; u < NC; ++u) {
            const int row = m + c0 + u;
            const double2 c = *reinterpret_cast<const double2*>(Cp + (int64_t)row * Dp + 2 * cp);
            const double2 k01 = *reinterpret_cast<const double2*>(kv + row * T);
            const double2 k23 = *reinterpret_cast<const double2*>(kv + row * T + 2);
            const double k[T] = {k01.x, k01.y, k23.x, k23.y};
#pragma unroll
            for (int tr = 0; tr < T; ++tr) {
                acc[u][0][tr] = fma(k[tr], c.x, acc[u][0][tr]);
                acc[u][1][tr] = fma(k[tr], c.y, acc[u][1][tr]);
            }
        }
    }
    if (c0 == 0) {
        for (int row = main_end; row < m1; ++row) {
            const double2 c = *reinterpret_cast<const double2*>(Cp + (int64_t)row * Dp + 2 * cp);
            const double2 k01 = *reinterpret_cast<const double2*>(kv + row * T);
            const double2 k23 = *reinterpret_cast<const double2*>(kv + row * T + 2);
            const double k[T] = {k01.x, k01.y, k23.x, k23.y};
#pragma unroll
            for (int tr = 0; tr < T; ++tr) {
                acc[0][0][tr] = fma(k[tr], c.x, acc[0][0][tr]);
                acc[0][1][tr] = fma(k[tr], c.y, acc[0][1][tr]);
            }
        }
    }
#pragma unroll
    for (int col = 0; col < 2; ++col)
#pragma unroll
        for (int tr = 0; tr < T; ++tr) {
            double v;
            if (NC == 2) v = acc[0][col][tr] + acc[1][col][tr];
            else v = (acc[0][col][tr] + acc[1][col][tr]) + (acc[NC - 2][col][tr] + acc[NC - 1][col][tr]);
            out[(2 * cp + col) * T + tr] = v;
        }
}

// f for the T states xin[d * T + tr] -> fout[d * T + tr]; xin visible on entry, fout visible on return
__device__ __forceinline__ void ivp_field_t(const double* xin, double* fout, IvpSharedT& s, double* kv, int D, int Dp,
                                            int M, const double* __restrict__ XcT, const double* __restrict__ Cp,
                                            double beta, double scale) {
    constexpr int T = IVP_T;
    const int tid = threadIdx.x;
    for (int mb = tid; mb < M; mb += 2 * IVP_THREADS) {      // two control points x T trajectories in flight per thread
        const int ma = mb, mbb = min(mb + IVP_THREADS, M - 1);
        double da[T], db[T];
#pragma unroll
        for (int tr = 0; tr < T; ++tr) { da[tr] = 0.0; db[tr] = 0.0; }
        for (int d = 0; d < D; ++d) {
            const double ca = XcT[(int64_t)d * M + ma], cb = XcT[(int64_t)d * M + mbb];
            const double2 x01 = *reinterpret_cast<const double2*>(xin + d * T);
            const double2 x23 = *reinterpret_cast<const double2*>(xin + d * T + 2);
            const double xv[T] = {x01.x, x01.y, x23.x, x23.y};
#pragma unroll
            for (int tr = 0; tr < T; ++tr) {
                const double ta = xv[tr] - ca, tb = xv[tr] - cb;
                da[tr] = fma(ta, ta, da[tr]);
                db[tr] = fma(tb, tb, db[tr]);
            }
        }
#pragma unroll
        for (int tr = 0; tr < T; ++tr) kv[ma * T + tr] = exp(-beta * da[tr]);
        if (mb + IVP_THREADS < M) {
#pragma unroll
            for (int tr = 0; tr < T; ++tr) kv[(mb + IVP_THREADS) * T + tr] = exp(-beta * db[tr]);
        }
    }
    __syncthreads();
    const int per = (M + IVP_PARTS - 1) / IVP_PARTS;
    if (D <= 32) {        // 16 column pairs x 8 groups: group g = quarter g / 2, classes {0, 1} or {2, 3}
        const int cp = tid & 15, g = tid >> 4, p = g >> 1;
        const int m0 = p * per, m1 = min(M, m0 + per);
        if (2 * cp < Dp) ivp_phase2_t<2>(Cp, Dp, kv, m0, m1, 2 * (g & 1), cp, s.part[g]);
    } else {              // 32 column pairs x 4 groups: group = quarter, all four classes
        const int cp = tid & 31, g = tid >> 5;
        const int m0 = g * per, m1 = min(M, m0 + per);
        if (2 * cp < Dp) ivp_phase2_t<4>(Cp, Dp, kv, m0, m1, 0, cp, s.part[g]);
    }
    __syncthreads();
    for (int e = tid; e < D * T; e += IVP_THREADS) {
        double v;
        if (D <= 32) {
            v = s.part[0][e] + s.part[1][e];
#pragma unroll
            for (int q = 1; q < IVP_PARTS; ++q) v += s.part[2 * q][e] + s.part[2 * q + 1][e];
        } else {
            v = s.part[0][e];
#pragma unroll
            for (int q = 1; q < IVP_PARTS; ++q) v += s.part[q][e];
        }
        fout[e] = scale * v;
    }
    __syncthreads();
}

__device__ __forceinline__ void ivp_rms_t(const double* buf, int D, double (&out)[IVP_T]) {
    constexpr int T = IVP_T;
#pragma unroll
    for (int tr = 0; tr < T; ++tr) out[tr] = 0.0;
    for (int d = 0; d < D; ++d) {
        const double2 b01 = *reinterpret_cast<const double2*>(buf + d * T);
        const double2 b23 = *reinterpret_cast<const double2*>(buf + d * T + 2);
        out[0] = fma(b01.x, b01.x, out[0]); out[1] = fma(b01.y, b01.y, out[1]);
        out[2] = fma(b23.x, b23.x, out[2]); out[3] = fma(b23.y, b23.y, out[3]);
    }
    const double rd = sqrt((double)D);
#pragma unroll
    for (int tr = 0; tr < T; ++tr) out[tr] = sqrt(out[tr]) / rd;
}
__device__ __forceinline__ void ivp_at_rest_t(const double* f, int D, double tol, bool (&out)[IVP_T]) {
    constexpr int T = IVP_T;
#pragma unroll
    for (int tr = 0; tr < T; ++tr) out[tr] = true;
    for (int d = 0; d < D; ++d)
#pragma unroll
        for (int tr = 0; tr < T; ++tr) out[tr] = out[tr] && (fabs(f[d * T + tr]) < tol);
}

template <int MINB>
__global__ void __launch_bounds__(IVP_THREADS, MINB) k_ivp_rk45_t(const double* __restrict__ y0s, int64_t n, int D, int Dp,
                                                            const double* __restrict__ XcT, int M, double beta,
                                                            const double* __restrict__ Cp, double scale, double t0,
                                                            double t_bound, double max_step, double rtol, double atol,
                                                            double rest_tol, int cap, double* __restrict__ ts,
                                                            double* __restrict__ ys, double* __restrict__ Qs,
                                                            int* __restrict__ n_steps, int* __restrict__ status) {
    constexpr int T = IVP_T;
    extern __shared__ __align__(16) double kv[];      // M x T kernel values
    __shared__ IvpSharedT s;
    const int tid = threadIdx.x;
    const int my = tid & (T - 1);                      // every element this thread owns belongs to trajectory slot `my`
    const int64_t base = (int64_t)blockIdx.x * T;
    const int64_t my_traj = base + my;
    const bool my_valid = my_traj < n;
    const int E = D * T;
    double* ts_o = ts + my_traj * (int64_t)(cap + 2);
    double* ys_o = ys + my_traj * (int64_t)(cap + 1) * D;
    double* Q_o = Qs + my_traj * (int64_t)cap * D * 4;
    const double dir = (t_bound != t0) ? ((t_bound > t0) ? 1.0 : -1.0) : 1.0;
    const double inf = __longlong_as_double(0x7ff0000000000000LL);

    for (int e = tid; e < E; e += IVP_THREADS) {
        const int d = e / T;
        const double v = my_valid ? y0s[my_traj * D + d] : 0.0;
        s.y[e] = v;
        if (my_valid) ys_o[d] = v;
    }
    if (tid < T && my_valid) ts_o[0] = t0;
    __syncthreads();
    ivp_field_t(s.y, s.K[0], s, kv, D, Dp, M, XcT, Cp, beta, scale);
    bool rest_old[T];
    ivp_at_rest_t(s.K[0], D, rest_tol, rest_old);

    const double span = fabs(t_bound - t0);
    double h_abs[T];
    {
        double d0[T], d1[T], d2[T], h0[T];
        for (int e = tid; e < E; e += IVP_THREADS) s.buf[e] = s.y[e] / (atol + fabs(s.y[e]) * rtol);
        __syncthreads();
        ivp_rms_t(s.buf, D, d0);
        __syncthreads();
        for (int e = tid; e < E; e += IVP_THREADS) s.buf[e] = s.K[0][e] / (atol + fabs(s.y[e]) * rtol);
        __syncthreads();
        ivp_rms_t(s.buf, D, d1);
#pragma unroll
        for (int tr = 0; tr < T; ++tr) {
            h0[tr] = (d0[tr] < 1e-5 || d1[tr] < 1e-5) ? 1e-6 : 0.01 * d0[tr] / d1[tr];
            h0[tr] = fmin(h0[tr], span);
        }
        const double my_h0 = ivp_pick(h0, my);
        for (int e = tid; e < E; e += IVP_THREADS) s.yt[e] = s.y[e] + (my_h0 * dir) * s.K[0][e];
        __syncthreads();
        ivp_field_t(s.yt, s.K[1], s, kv, D, Dp, M, XcT, Cp, beta, scale);
        for (int e = tid; e < E; e += IVP_THREADS)
            s.buf[e] = (s.K[1][e] - s.K[0][e]) / (atol + fabs(s.y[e]) * rtol);
        __syncthreads();
        ivp_rms_t(s.buf, D, d2);
#pragma unroll
        for (int tr = 0; tr < T; ++tr) {
            d2[tr] = d2[tr] / h0[tr];
            const double h1 = (d1[tr] <= 1e-15 && d2[tr] <= 1e-15) ? fmax(1e-6, h0[tr] * 1e-3)
                                                                   : pow(0.01 / fmax(d1[tr], d2[tr]), 0.2);
            h_abs[tr] = fmin(fmin(100.0 * h0[tr], h1), fmin(span, max_step));
        }
        __syncthreads();
    }

    double t[T], h[T], t_new[T], min_step[T];
    int nst[T], st[T];
    bool active[T], fresh[T], rejected[T];
#pragma unroll
    for (int tr = 0; tr < T; ++tr) {
        t[tr] = t0; h[tr] = 0.0; t_new[tr] = t0; min_step[tr] = 0.0;
        nst[tr] = 0; st[tr] = 0;
        active[tr] = (base + tr < n) && (span != 0.0);
        fresh[tr] = true; rejected[tr] = false;
    }

    // every round accepts or rejects one step of each unfinished trajectory; a rejection shrinks the step by at least
    // 0.2 x towards 10 ulp(t), so 64 rounds per stored step is far beyond what any trajectory can use: the bound only
    // guarantees that the kernel returns (leftovers are reported as failed, status -1)
    const int64_t max_rounds = 64 * ((int64_t)cap + 1);
    for (int64_t round = 0;; ++round) {
        if (round == max_rounds) {
#pragma unroll
            for (int tr = 0; tr < T; ++tr) if (active[tr]) { st[tr] = -1; active[tr] = false; }
            break;
        }
        // ---- one attempted step of every unfinished trajectory ----
        bool any = false;
#pragma unroll
        for (int tr = 0; tr < T; ++tr) {
            if (!active[tr]) continue;
            if (fresh[tr]) {
                min_step[tr] = 10.0 * fabs(nextafter(t[tr], dir * inf) - t[tr]);
                if (h_abs[tr] > max_step) h_abs[tr] = max_step;
                else if (h_abs[tr] < min_step[tr]) h_abs[tr] = min_step[tr];
                rejected[tr] = false;
                fresh[tr] = false;
            }
            if (h_abs[tr] < min_step[tr]) { st[tr] = -1; active[tr] = false; continue; }
            double hh = h_abs[tr] * dir;
            double tn = t[tr] + hh;
            if (dir * (tn - t_bound) > 0.0) tn = t_bound;
            hh = tn - t[tr];
            h[tr] = hh; h_abs[tr] = fabs(hh); t_new[tr] = tn;
            any = true;
        }
        if (!any) break;
        const double my_h = ivp_pick(h, my);
        for (int sg = 1; sg < 6; ++sg) {
            for (int e = tid; e < E; e += IVP_THREADS) {
                double dy = 0.0;
                for (int j = 0; j < sg; ++j) dy += s.K[j][e] * c_A[sg][j];
                s.yt[e] = s.y[e] + dy * my_h;
            }
            __syncthreads();
            ivp_field_t(s.yt, s.K[sg], s, kv, D, Dp, M, XcT, Cp, beta, scale);
        }
        for (int e = tid; e < E; e += IVP_THREADS) {
            double dy = 0.0;
#pragma unroll
            for (int j = 0; j < 6; ++j) dy += s.K[j][e] * c_B[j];
            s.ynew[e] = s.y[e] + my_h * dy;
        }
        __syncthreads();
        ivp_field_t(s.ynew, s.K[6], s, kv, D, Dp, M, XcT, Cp, beta, scale);
        for (int e = tid; e < E; e += IVP_THREADS) {
            double er = 0.0;
#pragma unroll
            for (int j = 0; j < 7; ++j) er += s.K[j][e] * c_E[j];
            const double sc = atol + fmax(fabs(s.y[e]), fabs(s.ynew[e])) * rtol;
            s.buf[e] = er * my_h / sc;
        }
        __syncthreads();
        double err[T];
        ivp_rms_t(s.buf, D, err);
        bool rest_new[T];
        ivp_at_rest_t(s.K[6], D, rest_tol, rest_new);

        // ---- decisions (identical in every thread) ----
        bool accept[T], stop[T];
        double t_old[T], xroot[T];
#pragma unroll
        for (int tr = 0; tr < T; ++tr) {
            accept[tr] = false; stop[tr] = false; t_old[tr] = t[tr]; xroot[tr] = 0.0;
            if (!active[tr]) continue;
            if (err[tr] < 1.0) {
                double factor = (err[tr] == 0.0) ? 10.0 : fmin(10.0, 0.9 * pow(err[tr], -0.2));
                if (rejected[tr]) factor = fmin(1.0, factor);
                h_abs[tr] *= factor;
                accept[tr] = true;
                t[tr] = t_new[tr];
                if (dir * (t[tr] - t_bound) >= 0.0) stop[tr] = true;
                if (rest_new[tr] || rest_old[tr]) {      // incl. a start at rest that leaves it: root = t_old
                    const double root = rest_old[tr] ? t_old[tr] : t[tr];
                    xroot[tr] = (root - t_old[tr]) / h[tr];
                    t[tr] = root;
                    st[tr] = rest_old[tr] ? -3 : 1;
                    stop[tr] = true;
                }
            } else {
                h_abs[tr] *= fmax(0.2, 0.9 * pow(err[tr], -0.2));
                rejected[tr] = true;
            }
        }
        const bool my_accept = ivp_pick(accept, my);
        const bool my_rest = my_accept && (ivp_pick(rest_new, my) || ivp_pick(rest_old, my));
        const int my_nst = ivp_pick(nst, my);
        if (my_accept) {
            const double x = ivp_pick(xroot, my);
            for (int e = tid; e < E; e += IVP_THREADS) {
                double q[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
                for (int j = 0; j < 7; ++j) {
                    const double k = s.K[j][e];
#pragma unroll
                    for (int c = 0; c < 4; ++c) q[c] += k * c_P[j][c];
                }
                if (my_valid) {
                    double* qo = Q_o + ((int64_t)my_nst * D + e / T) * 4;
                    qo[0] = q[0]; qo[1] = q[1]; qo[2] = q[2]; qo[3] = q[3];
                }
                if (my_rest) {
                    const double p1 = x, p2 = p1 * x, p3 = p2 * x, p4 = p3 * x;
                    const double yo = s.y[e];
                    s.ynew[e] = my_h * (((q[0] * p1 + q[1] * p2) + q[2] * p3) + q[3] * p4) + yo;
                }
            }
        }
        __syncthreads();
        if (my_accept) {
            for (int e = tid; e < E; e += IVP_THREADS) {
                const double v = s.ynew[e];
                s.y[e] = v;
                s.K[0][e] = s.K[6][e];
                if (my_valid) ys_o[(int64_t)(my_nst + 1) * D + e / T] = v;
            }
            if (tid < T && my_valid) {
                ts_o[my_nst + 1] = ivp_pick(t, my);
                if (ivp_pick(st, my) == -3) ts_o[my_nst + 2] = ivp_pick(t_new, my);
            }
        }
#pragma unroll
        for (int tr = 0; tr < T; ++tr) {
            if (!accept[tr]) continue;
            ++nst[tr];
            rest_old[tr] = rest_new[tr];
            fresh[tr] = true;
            if (stop[tr]) active[tr] = false;
            else if (nst[tr] == cap) { st[tr] = -2; active[tr] = false; }
        }
        __syncthreads();
    }
    if (tid < T && my_valid) { n_steps[my_traj] = ivp_pick(nst, my); status[my_traj] = ivp_pick(st, my); }
}

// C (M x D) -> Cp (M x Dp), Dp = D rounded up to even, zero padded: aligned 16-byte loads of column pairs
__global__ void k_ivp_pad(const double* __restrict__ Cm, int M, int D, int Dp, double* __restrict__ Cp) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < M * Dp) { const int m = i / Dp, d = i % Dp; Cp[i] = (d < D) ? Cm[(int64_t)m * D + d] : 0.0; }
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

static size_t ivp_xct_doubles(int D, int M) { return ((size_t)D * M + 1) & ~(size_t)1; }    // keeps Cp 16-byte aligned
static size_t ivp_ws_doubles(int D, int M) { return ivp_xct_doubles(D, M) + (size_t)M * ((D + 1) & ~1); }

extern "C" int dvf_ivp_workspace_bytes(int D, int M, size_t* out) {
    DVF_ARG(out && D >= 1 && D <= kMaxD && M >= 1, "dvf_ivp_workspace_bytes: bad arguments");
    *out = ivp_ws_doubles(D, M) * sizeof(double);
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
    if (ws_bytes < ivp_ws_doubles(D, M) * sizeof(double)) {
        set_error("dvf_integrate_rk45: workspace %zu < %zu bytes", ws_bytes, ivp_ws_doubles(D, M) * sizeof(double));
        return DVF_ERR_WORKSPACE;
    }
    DVF_ARG(n <= 0x7fffffff, "dvf_integrate_rk45: at most 2^31 - 1 initial states per call");
    cudaStream_t st = (cudaStream_t)stream;
    double* XcT = (double*)ws;
    k_ivp_transpose<<<(M * D + 255) / 256, 256, 0, st>>>(Xc, M, D, XcT);
    DVF_LAUNCH_CHECK();
    // one trajectory per CTA for a handful of initial states, IVP_T per CTA (built for 4 CTAs per SM: 59 ms against
    // 72 ms with 3 and 163 ms with one trajectory per CTA for 4096 states, D = 30, M = 500) otherwise (same results;
    // DVF_IVP_VARIANT =
    // 1 forces the former, 2 / 3 the latter compiled for 3 / 4 CTAs per SM: developer experiments and the test that
    // compares them)
    int variant = (n >= 2 * IVP_T) ? 3 : 1;
    if (const char* e = getenv("DVF_IVP_VARIANT")) { const int v = atoi(e); if (v >= 1 && v <= 3) variant = v; }
    if (variant == 1) {
        const size_t smem = (size_t)M * sizeof(double);
        k_ivp_rk45<<<(unsigned)n, IVP_THREADS, smem, st>>>(y0, n, D, XcT, M, beta, C, scale, t0, t_bound, max_step,
                                                           rtol, atol, rest_tol, cap, ts, ys, Q, n_steps, status);
        DVF_LAUNCH_CHECK();
        return DVF_OK;
    }
    const int Dp = (D + 1) & ~1;
    double* Cp = XcT + ivp_xct_doubles(D, M);
    DVF_ARG((((uintptr_t)Cp) & 15) == 0, "dvf_integrate_rk45: workspace must be 16-byte aligned");
    k_ivp_pad<<<(M * Dp + 255) / 256, 256, 0, st>>>(C, M, D, Dp, Cp);
    DVF_LAUNCH_CHECK();
    const size_t smem = (size_t)M * IVP_T * sizeof(double);
    auto kern = (variant == 3) ? k_ivp_rk45_t<4> : k_ivp_rk45_t<3>;      // 128 / 168 registers: 4 / 3 CTAs per SM
    if (smem + sizeof(IvpSharedT) > 48 * 1024)     // beyond the default limit: opt in (cheap, so not cached per device)
        DVF_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t ctas = (n + IVP_T - 1) / IVP_T;
    kern<<<(unsigned)ctas, IVP_THREADS, smem, st>>>(y0, n, D, Dp, XcT, M, beta, Cp, scale, t0, t_bound, max_step, rtol,
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

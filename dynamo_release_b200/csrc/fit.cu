// The SparseVFC EM loop (reference scVectorField.py:417-497) as a stream of kernels with device-resident
// control state: no host round trip inside an iteration, the stop flag is read back every `check_every`
// iterations and kernels launched after it is raised exit at once.
#include "em.h"
#include "comm.h"

namespace dvf {

static inline size_t al(size_t x) { return (x + 255) & ~(size_t)255; }

struct fit_layout {
    size_t uy, r2, gbuf, gscratch, L, Kc, Cp, part, scal, ctl, total;
    int W, ldg, Mp, DP, nblk_es;
    int64_t nblk_uv;
    gram_plan gp;
};

static fit_layout make_layout(int64_t N, int D, int M) {
    fit_layout l;
    l.W = (int)dvf_fit_row_width(M, D);
    l.gp = make_gram_plan(N, D, M);
    l.ldg = l.gp.ldg;
    l.Mp = (M + 31) / 32 * 32;
    l.DP = pad8(D);
    l.nblk_es = estep_blocks(N);
    l.nblk_uv = updatev_blocks(N);
    size_t o = 0;
    l.uy = o;       o += al(sizeof(double) * (size_t)N * l.W);
    l.r2 = o;       o += al(sizeof(double) * (size_t)(N + 2));
    l.gbuf = o;     o += al(sizeof(double) * ((size_t)l.ldg * l.ldg + RED_COUNT));
    l.gscratch = o; o += al(l.gp.scratch_bytes);
    {
        size_t chol = sizeof(double) * (size_t)(M + D) * M, lu = solve_normal_ws_bytes(M, D);
        l.L = o;    o += al(chol > lu ? chol : lu);
    }
    l.Kc = o;       o += al(sizeof(double) * (size_t)M * M);
    l.Cp = o;       o += al(sizeof(double) * (size_t)l.Mp * l.DP);
    size_t np = (size_t)l.nblk_es * 4;
    if ((size_t)l.nblk_uv > np) np = (size_t)l.nblk_uv;
    if ((size_t)TR_BLOCKS > np) np = TR_BLOCKS;
    l.part = o;     o += al(sizeof(double) * np);
    l.scal = o;     o += al(sizeof(double) * 16);
    l.ctl = o;      o += al(sizeof(em_ctl));
    l.total = o;
    return l;
}

// r2[n] = |y_n|^2 (V = 0 at the start, scVectorField.py:430-444) + block partials of the sum
__global__ void __launch_bounds__(ES_THREADS) k_init_r2(const double* __restrict__ UY, int64_t N, int W, int M, int D,
                                                        double* __restrict__ r2, double* __restrict__ part) {
    __shared__ double sh[32];
    double s = 0.0;
    for (int64_t i = (int64_t)blockIdx.x * ES_THREADS + threadIdx.x; i < N; i += (int64_t)gridDim.x * ES_THREADS) {
        const double* y = UY + i * (int64_t)W + M;
        double rr = 0.0;
        for (int d = 0; d < D; ++d) rr = fma(y[d], y[d], rr);
        r2[i] = rr;
        s += rr;
    }
    s = block_sum(s, sh);
    if (threadIdx.x == 0) part[blockIdx.x] = s;
}

__global__ void k_ctl_init(em_ctl* ctl, const double* sum_y2, int64_t N_total, int D, double gamma, int max_iter,
                           double ecr) {
    double s2 = sum_y2[0] / ((double)N_total * D);
    if (s2 < 1e-8) s2 = 1e-7;                               // scVectorField.py:445
    ctl->sigma2 = s2;
    ctl->gamma = gamma;
    ctl->E = 1.0;
    ctl->tecr = 1.0;
    ctl->lam_sigma2 = 0.0;
    ctl->sumP_clamped = 0.0;
    ctl->n_inlier = 0.0;
    ctl->iter = 0;
    ctl->info = 0;
    ctl->done = !(0 < max_iter && 1.0 > ecr && s2 > 1e-8);  // loop condition, :449
}

// after the allreduce of [G | red]: energy, change rate, lambda sigma2 for the solve   (:452-457)
__global__ void k_ctl_energy(em_ctl* ctl, const double* red, const double* trpart, double lambda, int D,
                             double* E_traj, double* tecr_traj) {
    if (ctl->done) return;
    double tr = 0.0;
    for (int b = 0; b < TR_BLOCKS; ++b) tr += trpart[b];
    const double s2 = ctl->sigma2;
    double E = red[RED_SUM_PR2] / (2.0 * s2) + red[RED_SUM_P] * log(s2) * D / 2.0;
    E = E + lambda / 2.0 * tr;
    const double tecr = fabs((E - ctl->E) / E);
    E_traj[ctl->iter] = E;
    tecr_traj[ctl->iter] = tecr;
    ctl->E = E;
    ctl->tecr = tecr;
    ctl->lam_sigma2 = lambda * s2;
    ctl->sumP_clamped = red[RED_SUM_PC];
    ctl->n_inlier = red[RED_NCORR];
}

// after V = U C: sigma2, gamma, loop condition   (:485-495, :449)
__global__ void k_ctl_finish(em_ctl* ctl, const double* vsum, int64_t N_total, int D, int max_iter, double ecr) {
    if (ctl->done) return;
    const double s2 = vsum[0] / (ctl->sumP_clamped * D);
    double gm = ctl->n_inlier / (double)N_total;
    if (gm > 0.95) gm = 0.95;
    else if (gm < 0.05) gm = 0.05;
    ctl->sigma2 = s2;
    ctl->gamma = gm;
    ctl->iter += 1;
    if (ctl->info != 0) ctl->done = 1;
    if (!(ctl->iter < max_iter && ctl->tecr > ecr && s2 > 1e-8)) ctl->done = 1;
}

__global__ void k_clamp(double* P, int64_t N, double minP) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += (int64_t)gridDim.x * blockDim.x)
        P[i] = fmax(P[i], minP);
}

}  // namespace dvf

using namespace dvf;

extern "C" int dvf_fit_workspace_bytes(int64_t N_local, int D, int M, size_t* out) {
    DVF_ARG(out && N_local >= 1 && D >= 1 && D <= kMaxD && M >= 1, "bad argument");
    *out = make_layout(N_local, D, M).total;
    return DVF_OK;
}

extern "C" int dvf_fit_em(const double* X, const double* Y, int64_t N, int64_t N_total, int D, const double* Xc, int M,
                          double beta, const dvf_em_params* prm, double* C, double* P, double* V, double* E_traj,
                          double* tecr_traj, dvf_em_result* res, void* ws, size_t ws_bytes, void* comm,
                          void* stream) {
    DVF_ARG(X && Y && Xc && prm && C && P && E_traj && tecr_traj && res && ws, "NULL pointer");
    DVF_ARG(N >= 1 && N_total >= N && M >= 1, "bad shape N=%lld N_total=%lld M=%d", (long long)N, (long long)N_total, M);
    DVF_ARG(D >= 1 && D <= kMaxD, "D=%d unsupported (1..%d)", D, kMaxD);
    DVF_ARG(prm->max_iter >= 0, "max_iter < 0");
    DVF_ARG(prm->solver == DVF_SOLVER_CHOLESKY || prm->solver == DVF_SOLVER_NORMAL_LU, "unknown solver %d", prm->solver);
    const fit_layout l = make_layout(N, D, M);
    if (ws_bytes < l.total) {
        set_error("fit workspace too small: %zu < %zu", ws_bytes, l.total);
        return DVF_ERR_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    char* base = (char*)ws;
    double* UY = (double*)(base + l.uy);
    double* r2 = (double*)(base + l.r2);
    double* G = (double*)(base + l.gbuf);
    double* red = G + (size_t)l.ldg * l.ldg;
    double* gscratch = (double*)(base + l.gscratch);
    double* Lw = (double*)(base + l.L);
    double* Kc = (double*)(base + l.Kc);
    double* Cp = (double*)(base + l.Cp);
    double* part = (double*)(base + l.part);
    double* scal = (double*)(base + l.scal);     // [0..1] min/zero count, [2] sum P r2 (new V), [3] sum y^2
    em_ctl* ctl = (em_ctl*)(base + l.ctl);
    const int* done = &ctl->done;

    // ---- prologue: U, Kc, r2 = |y|^2, sigma2 (scVectorField.py:417-445) ----
    if (int rc = dvf_build_uy(X, Y, N, D, Xc, M, beta, UY, stream)) return rc;
    if (int rc = dvf_kernel_matrix(Xc, M, D, Xc, M, beta, Kc, M, stream)) return rc;
    DVF_CUDA(cudaMemsetAsync(G, 0, sizeof(double) * ((size_t)l.ldg * l.ldg + RED_COUNT), st));
    DVF_CUDA(cudaMemsetAsync(C, 0, sizeof(double) * (size_t)M * D, st));
    if (int rc = gram_prepare(l.gp, N, D, M, gscratch, st)) return rc;
    k_init_r2<<<l.nblk_es, ES_THREADS, 0, st>>>(UY, N, l.W, M, D, r2, part);
    DVF_LAUNCH_CHECK();
    if (int rc = sum_partials_run(part, l.nblk_es, 1, scal + 3, nullptr, st)) return rc;
    if (int rc = comm_allreduce(comm, scal + 3, 1, false, st)) return rc;
    k_ctl_init<<<1, 1, 0, st>>>(ctl, scal + 3, N_total, D, prm->gamma, prm->max_iter, prm->ecr);
    DVF_LAUNCH_CHECK();

    const int check_every = prm->check_every > 0 ? prm->check_every : 8;
    em_ctl h;
    memset(&h, 0, sizeof(h));
    int launched = 0;
    bool stop = false;
    while (!stop) {
        if (launched >= prm->max_iter) break;
        // ---- E-step (:452 -> :197-227) ----
        if (int rc = estep1_run(r2, N, &ctl->sigma2, 0.0, P, part, done, st)) return rc;
        if (int rc = estep_min_run(part, l.nblk_es, scal, done, st)) return rc;
        if (int rc = comm_allreduce(comm, scal, 1, true, st)) return rc;
        if (int rc = estep2_run(r2, N, D, ctl, 0.0, 0.0, prm->a, prm->minP, prm->theta, scal, P, part, done, st)) return rc;
        if (int rc = sum_partials_run(part, l.nblk_es, 4, red, done, st)) return rc;
        if (int rc = trace_run(C, Kc, M, D, part, done, st)) return rc;
        // ---- M-step assembly (:468-476), one allreduce of [G | scalars] ----
        profile_gram_begin(st);
        if (int rc = gram_run(l.gp, UY, P, N, l.W, G, gscratch, done, st)) return rc;
        profile_gram_end(st);
        if (int rc = comm_allreduce(comm, G, (int64_t)l.ldg * l.ldg + RED_COUNT, false, st)) return rc;
        k_ctl_energy<<<1, 1, 0, st>>>(ctl, red, part, prm->lambda_, D, E_traj, tecr_traj);
        DVF_LAUNCH_CHECK();
        // ---- solve (:481), V = U C and sigma2 / gamma (:484-495) ----
        if (prm->solver == DVF_SOLVER_NORMAL_LU) {
            if (int rc = solve_normal_run(G, l.ldg, Kc, M, D, 0.0, &ctl->lam_sigma2, Lw, C, &ctl->info, done, st)) return rc;
        } else {
            if (int rc = solve_run(G, l.ldg, Kc, M, D, 0.0, &ctl->lam_sigma2, Lw, C, &ctl->info, done, st)) return rc;
        }
        if (int rc = pad_c_run(C, M, D, Cp, l.Mp, l.DP, done, st)) return rc;
        if (int rc = update_v_run(UY, Cp, P, prm->minP, N, D, M, V, r2, part, done, st)) return rc;
        if (int rc = sum_partials_run(part, (int)l.nblk_uv, 1, scal + 2, done, st)) return rc;
        if (int rc = comm_allreduce(comm, scal + 2, 1, false, st)) return rc;
        k_ctl_finish<<<1, 1, 0, st>>>(ctl, scal + 2, N_total, D, prm->max_iter, prm->ecr);
        DVF_LAUNCH_CHECK();
        ++launched;
        if (launched % check_every == 0 || launched >= prm->max_iter) {
            DVF_CUDA(cudaMemcpyAsync(&h, ctl, sizeof(h), cudaMemcpyDeviceToHost, st));
            DVF_CUDA(cudaStreamSynchronize(st));
            if (h.done) stop = true;
        }
    }
    if (launched == 0) {
        DVF_CUDA(cudaMemcpyAsync(&h, ctl, sizeof(h), cudaMemcpyDeviceToHost, st));
        DVF_CUDA(cudaStreamSynchronize(st));
    }
    k_clamp<<<estep_blocks(N), ES_THREADS, 0, st>>>(P, N, prm->minP);
    DVF_LAUNCH_CHECK();
    DVF_CUDA(cudaStreamSynchronize(st));
    res->n_iter = h.iter;
    res->sigma2 = h.sigma2;
    res->gamma = h.gamma;
    res->n_inlier = h.n_inlier;
    res->status = 0;
    if (h.info != 0) {
        res->status = DVF_ERR_NUMERIC;
        if (prm->solver == DVF_SOLVER_NORMAL_LU)
            set_error("Singular matrix: LU of lhs^T lhs met a zero pivot at column %d in EM iteration %d", h.info - 1, h.iter);
        else
            set_error("Cholesky met a non-positive pivot at column %d in EM iteration %d (lhs not positive definite)",
                      h.info - 1, h.iter);
        return DVF_ERR_NUMERIC;
    }
    return DVF_OK;
}

// ---- step-level entry points -------------------------------------------------------------------------
extern "C" int dvf_estep(const double* r2, int64_t N, int D, double sigma2, double gamma, double a, double* P,
                         double* stats, void* stream) {
    DVF_ARG(r2 && P && stats && N >= 1 && D >= 1, "bad argument");
    cudaStream_t st = (cudaStream_t)stream;
    const int nb = estep_blocks(N);
    double* part = nullptr;
    DVF_CUDA(cudaMallocAsync(&part, sizeof(double) * ((size_t)nb * 4 + 8), st));
    double* minbuf = part + (size_t)nb * 4;
    double* sums = minbuf + 2;
    int rc = estep1_run(r2, N, nullptr, sigma2, P, part, nullptr, st);
    if (!rc) rc = estep_min_run(part, nb, minbuf, nullptr, st);
    if (!rc) rc = estep2_run(r2, N, D, nullptr, sigma2, gamma, a, 0.0, 2.0, minbuf, P, part, nullptr, st);
    if (!rc) rc = sum_partials_run(part, nb, 4, sums, nullptr, st);
    double h[6];
    if (!rc) {
        DVF_CUDA(cudaMemcpyAsync(h, minbuf, sizeof(double) * 6, cudaMemcpyDeviceToHost, st));
        DVF_CUDA(cudaStreamSynchronize(st));
        double out[4] = {h[2] / (2.0 * sigma2) + h[3] * log(sigma2) * D / 2.0, h[3], h[0], h[1]};
        DVF_CUDA(cudaMemcpyAsync(stats, out, sizeof(out), cudaMemcpyHostToDevice, st));
        DVF_CUDA(cudaStreamSynchronize(st));
    }
    cudaFreeAsync(part, st);
    return rc;
}

extern "C" int dvf_update_v(const double* UY, const double* C, const double* P, double minP, int64_t N, int D, int M,
                            double* V, double* r2, double* stats, void* stream) {
    DVF_ARG(UY && C && P && r2 && stats && N >= 1 && M >= 1, "bad argument");
    DVF_ARG(D >= 1 && D <= kMaxD, "D=%d unsupported (1..%d)", D, kMaxD);
    cudaStream_t st = (cudaStream_t)stream;
    const int Mp = (M + 31) / 32 * 32, DP = pad8(D);
    const int64_t nb = updatev_blocks(N);
    double* tmp = nullptr;
    DVF_CUDA(cudaMallocAsync(&tmp, sizeof(double) * ((size_t)Mp * DP + nb), st));
    double* part = tmp + (size_t)Mp * DP;
    int rc = pad_c_run(C, M, D, tmp, Mp, DP, nullptr, st);
    if (!rc) rc = update_v_run(UY, tmp, P, minP, N, D, M, V, r2, part, nullptr, st);
    if (!rc) rc = sum_partials_run(part, (int)nb, 1, stats, nullptr, st);
    cudaFreeAsync(tmp, st);
    return rc;
}

// ---- allocation-free variant of dvf_update_v -----------------------------------------------------------------
// No counterpart in the reference: added for callers that cannot allocate
// (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  dvf_update_v takes its scratch (the
// zero-padded copy of C and the per-CTA partial sums) from the stream-ordered pool; here the same scratch lives in a
// file-scope device array of a fixed, modest size that every GPU loading the library holds until exit:
//   padded C : Mp x DP doubles, Mp = roundup(M, 32), DP = roundup(D, 8)  ->  M <= 1024, D <= 64  =  64 Ki doubles
//   partials : one double per 128 cells                                   ->  N <= 4 194 304      =  32 Ki doubles
// 768 KB in total.  Nothing in it has to start at zero: every word that is read was written earlier in the same call
// (k_pad_c writes all Mp x DP entries, every CTA of k_update_v writes its partial), so the previous call's contents are
// harmless.  All calls share the one array: calls that overlap on the GPU (different streams, or the same entry point
// from two host threads) race -- serialise them.
namespace dvf {
constexpr int NA_MAX_M = 1024, NA_MAX_D = 64;
constexpr int64_t NA_MAX_N = 4194304;
constexpr size_t NA_C_DOUBLES = (size_t)NA_MAX_M * NA_MAX_D;                 // padded C
constexpr size_t NA_PART_DOUBLES = (size_t)(NA_MAX_N / UV_CELLS);            // partial sums
static __device__ __align__(256) double g_update_v_scratch[NA_C_DOUBLES + NA_PART_DOUBLES];

// device address of the array on the CURRENT device, looked up once per device (in host code the array's name is only
// a host-side stand-in, not a device address)
static int update_v_scratch_ptr(double** out) {
    constexpr int kMaxDev = 64;
    static double* cache[kMaxDev] = {nullptr};
    int dev = 0;
    DVF_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDev) { set_error("device ordinal %d out of range", dev); return DVF_ERR_ARG; }
    if (!cache[dev]) {
        double* p = nullptr;
        DVF_CUDA(cudaGetSymbolAddress((void**)&p, g_update_v_scratch));
        cache[dev] = p;
    }
    *out = cache[dev];
    return DVF_OK;
}
}  // namespace dvf

extern "C" int dvf_update_v_noalloc(const double* UY, const double* C, const double* P, double minP, int64_t N, int D,
                                    int M, double* V, double* r2, double* stats, void* stream) {
    DVF_ARG(UY && C && P && r2 && stats && N >= 1 && M >= 1, "bad argument");
    DVF_ARG(D >= 1 && D <= kMaxD, "D=%d unsupported (1..%d)", D, kMaxD);
    if (M > NA_MAX_M || D > NA_MAX_D || N > NA_MAX_N) {
        set_error("dvf_update_v_noalloc: fixed scratch serves M <= %d, D <= %d, N <= %lld (got M=%d D=%d N=%lld); use "
                  "dvf_update_v", NA_MAX_M, NA_MAX_D, (long long)NA_MAX_N, M, D, (long long)N);
        return DVF_ERR_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const int Mp = (M + 31) / 32 * 32, DP = pad8(D);
    const int64_t nb = updatev_blocks(N);
    double* scratch = nullptr;
    if (int rc = update_v_scratch_ptr(&scratch)) return rc;
    double* Cp = scratch;                      // Mp * DP <= NA_C_DOUBLES
    double* part = scratch + NA_C_DOUBLES;     // nb <= NA_PART_DOUBLES
    int rc = pad_c_run(C, M, D, Cp, Mp, DP, nullptr, st);
    if (!rc) rc = update_v_run(UY, Cp, P, minP, N, D, M, V, r2, part, nullptr, st);
    if (!rc) rc = sum_partials_run(part, (int)nb, 1, stats, nullptr, st);
    return rc;
}

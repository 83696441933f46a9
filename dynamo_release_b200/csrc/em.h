// internal: pieces of the EM loop shared between kern_em.cu, kern_gram.cu, kern_solve.cu and fit.cu
#pragma once
#include "common.cuh"

namespace dvf {

// device-resident loop state (one per fit); every EM kernel starts with `if (ctl->done) return;`
struct em_ctl {
    double sigma2, gamma, E, tecr, lam_sigma2;
    double sumP_clamped, n_inlier;
    int iter, done, info, pad;
};

// slots of the scalar block that is allreduced together with G (appended after ldg*ldg doubles)
enum { RED_SUM_PR2 = 0, RED_SUM_P = 1, RED_SUM_PC = 2, RED_NCORR = 3, RED_COUNT = 8 };

struct gram_plan {
    int ldg, n_ctas, n_tiles, n_bundles, n_groups, sb, lag;
    int use_parts, n_parts, sb_parts, n_counters;     // "parts" schedule (kern_gram.cu): one tile part per CTA
    size_t off_tiles, off_boff, off_btiles, off_progress, off_parts, off_poff, off_pclamp, off_slots, scratch_bytes;
};
gram_plan make_gram_plan(int64_t N, int D, int M);
int gram_prepare(const gram_plan& pl, int64_t N, int D, int M, void* scratch, cudaStream_t st);
// P: already clamped to >= minP
int gram_run(const gram_plan& pl, const double* UY, const double* P, int64_t N, int W, double* G, double* scratch,
             const int* done, cudaStream_t st);

int solve_run(const double* G, int64_t ldg, const double* Kc, int M, int D, double lam_sigma2,
              const double* lam_sigma2_ptr, double* L, double* C, int* info, const int* done, cudaStream_t st);

// the reference's solver form: normal equations + LU with partial pivoting (kern_solve.cu, k_lu_solve)
size_t solve_normal_ws_bytes(int M, int D);
int solve_normal_run(const double* G, int64_t ldg, const double* Kc, int M, int D, double lam_sigma2,
                     const double* lam_sigma2_ptr, void* ws, double* C, int* info, const int* done, cudaStream_t st);

constexpr int ES_THREADS = 256;
constexpr int ES_MAX_BLOCKS = 1184;   // 8 per SM on 148 SMs; fixed => partial-sum order is reproducible
inline int estep_blocks(int64_t N) {
    int64_t b = (N + ES_THREADS - 1) / ES_THREADS;
    return (int)(b < 1 ? 1 : (b > ES_MAX_BLOCKS ? ES_MAX_BLOCKS : b));
}
constexpr int UV_CELLS = 128;         // cells per CTA of the V-update kernel
inline int64_t updatev_blocks(int64_t N) { return (N + UV_CELLS - 1) / UV_CELLS; }

// E-step in two passes around the global min-nonzero rule (scVectorField.py:223)
int estep1_run(const double* r2, int64_t N, const double* sigma2_ptr, double sigma2, double* P, double* part,
               const int* done, cudaStream_t st);
int estep_min_run(const double* part, int nblk, double* minbuf, const int* done, cudaStream_t st);
int estep2_run(const double* r2, int64_t N, int D, const em_ctl* ctl, double sigma2, double gamma, double a,
               double minP, double theta, const double* minbuf, double* P, double* part, const int* done,
               cudaStream_t st);
int sum_partials_run(const double* part, int nblk, int nvals, double* out, const int* done, cudaStream_t st);

int pad_c_run(const double* C, int M, int D, double* Cp, int Mp, int DP, const int* done, cudaStream_t st);
int update_v_run(const double* UY, const double* Cp, const double* P, double minP, int64_t N, int D, int M,
                 double* V, double* r2, double* part, const int* done, cudaStream_t st);
int trace_run(const double* C, const double* Kc, int M, int D, double* part, const int* done, cudaStream_t st);
constexpr int TR_BLOCKS = 64;

}  // namespace dvf

/* dvf_b200.h -- C ABI of the B200-native SparseVFC vector-field hot path.
 *
 * dynamo (aristoteleo/dynamo-release) has no FFI: its boundary for this path is a set of module-level
 * Python functions.  Each entry point below replaces the reference function cited
 * beside it (paths relative to /root/reference/dynamo); the Python host layer in dynamo_release_b200/
 * binds them with ctypes and mirrors the reference signatures.
 *
 * Conventions
 *   - every pointer named x, Xc, C, Y, U, P, V, J, ... is a DEVICE pointer (float64 unless stated),
 *     C-contiguous, caller-owned; nothing is allocated behind the caller's back except inside
 *     dvf_fit_* where the workspace is queried first (dvf_fit_workspace_bytes) and passed in;
 *   - `stream` is a cudaStream_t passed as void* (0 = legacy default stream);
 *   - all functions return 0 on success, <0 on error; dvf_last_error() gives a thread-local message;
 *   - sizes: n / N cells, D dimensions (1 <= D <= 64), M control points.
 */
#ifndef DVF_B200_H
#define DVF_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DVF_OK 0
#define DVF_ERR_ARG (-1)       /* bad argument (shape, NULL pointer, unsupported D) */
#define DVF_ERR_CUDA (-2)      /* CUDA runtime error, see dvf_last_error() */
#define DVF_ERR_NUMERIC (-3)   /* solver met a non-positive pivot / EM could not start */
#define DVF_ERR_COMM (-4)      /* NCCL error */
#define DVF_ERR_WORKSPACE (-5) /* workspace too small */

int dvf_abi_version(void);
const char* dvf_last_error(void);

/* Measurement hooks (bench.py): CUDA events around every Gram launch inside dvf_fit_em and a counter of all
 * kernel launches of the library.  dvf_profile_enable(1) resets both. */
int dvf_profile_enable(int on);
int dvf_profile_get(double* gram_ms, int* gram_launches, long long* kernel_launches);

/* K1: K[i, j] = exp(-beta |x_i - Xc_j|^2), row-major n x M with leading dimension ldk (>= M).
 * replaces vectorfield/utils.py:180-212 (con_K, values path). */
int dvf_kernel_matrix(const double* x, int64_t n, int D, const double* Xc, int M, double beta,
                      double* K, int64_t ldk, void* stream);

/* K5: v = K(x, Xc) C  (n x D).   replaces vectorfield/utils.py:73-142 (vector_field_function;
 * ctrl-point masking and dim slicing are done by the host layer on C / the result). */
int dvf_eval(const double* x, int64_t n, int D, const double* Xc, int M, double beta,
             const double* C, double* v, void* stream);

/* K6: J[i, j, c] = -2 beta sum_m C[m, i] k(x_c, Xc_m) (x_c[j] - Xc[m, j]); layout (D, D, n), cell index
 * fastest.  replaces vectorfield/utils.py:351-379 (Jacobian_rkhs_gaussian). */
int dvf_jacobian(const double* x, int64_t n, int D, const double* Xc, int M, double beta,
                 const double* C, double* J, void* stream);
/* The same with a caller-owned workspace (query, then pass; 256-byte aligned; 0 bytes when the direct per-cell
 * kernel is used).  dvf_jacobian() takes its scratch from the device's stream-ordered memory pool instead. */
int dvf_jacobian_workspace_bytes(int64_t n, int D, int M, size_t* out);
int dvf_jacobian_ws(const double* x, int64_t n, int D, const double* Xc, int M, double beta, const double* C,
                    double* J, void* ws, size_t ws_bytes, void* stream);

/* Least-action-path terms (vectorfield/least_action_path.py:7-57, `action` / `action_grad`): for n path segments with
 * centres x (n x D) and finite-difference velocities vp (n x D):  dv = vp - f(x)  and  z = dv J(x)  (row vector times
 * Jacobian, the loop of :50-52) without materialising J.  Either output may be NULL.  D <= 32. */
int dvf_lap_terms(const double* x, const double* vp, int64_t n, int D, const double* Xc, int M, double beta,
                  const double* C, double* dv, double* z, void* stream);

/* K7: fused divergence / acceleration / curvature without materialising J.  Any output may be NULL.
 *   v (n x D) = f(x);  div (n) = tr J;  acc (n x D) = J v;  acc_norm (n) = |acc|;
 *   formula 2: curv (n x D) = (a (v.v) - v (v.a)) / |v|^4, curv_norm = |curv|;
 *   formula 1: curv_norm = |v a^T|_F / |v|^3, curv untouched.
 * replaces vectorfield/utils.py:699-725, 728-793, 743-757, 797-821. */
int dvf_div_acc_curv(const double* x, int64_t n, int D, const double* Xc, int M, double beta,
                     const double* C, int formula, double* v, double* div, double* acc,
                     double* acc_norm, double* curv, double* curv_norm, void* stream);

/* K8: out[e, r, c] = float32( Qe[e, :] J[:, :, c] Qr[r, :]^T ), J laid out (D, D, n), out (Ge, Gr, n).
 * If take_abs != 0 the absolute value is taken before any reduction.  If group != NULL (int32 label per
 * cell, 0 <= label < n_groups) the per-group SUM over cells of the float32-rounded values is accumulated in
 * float64 into out_group_sum (n_groups x Ge x Gr, must be zeroed by the caller).  out_full may be NULL.
 * replaces vectorfield/utils.py:544-554 (transform_jacobian), :497-541, and for Ge=Gr=1 with
 * out_elem (float64, n) :430-453 (elementwise_jacobian_transformation). */
int dvf_gene_transform(const double* J, int64_t n, int D, const double* Qe, int Ge, const double* Qr,
                       int Gr, int take_abs, const int32_t* group, int n_groups, float* out_full,
                       double* out_group_sum, void* stream);
/* Tensor-core path of K8 with a caller-owned workspace (1024-byte aligned); dvf_gene_transform() selects it for
 * large gene sets (Ge, Gr >= 96) and takes the workspace from the stream-ordered pool.  Needs an even n. */
int dvf_gene_gemm_workspace_bytes(int64_t n, int D, int Ge, int Gr, int full, int grouped, size_t* out);
int dvf_gene_transform_gemm(const double* J, int64_t n, int D, const double* Qe, int Ge, const double* Qr, int Gr,
                            int take_abs, const int32_t* group, int n_groups, float* out_full,
                            double* out_group_sum, void* ws, size_t ws_bytes, void* stream);
/* Reductions over cells of the float32 gene-space Jacobian that rank_jacobian_genes consumes
 * (vectorfield/rank_vf.py:483-508), without materialising the (Ge, Gr, n) array.  Per-cell value, in the reference's
 * order: |.| if take_abs (:484-485); divided in float32 by jmax[e][r] if jmax != NULL (normalize, :487-490);
 *   op 0: out[g][e][r] += value over the cells with group[c] == g        (means: :498-508, utils.py:557-582)
 *   op 1: out[g][e][r] += value * value^T where both are negative        (mode='switch', :492-496; needs Ge == Gr)
 *   op 2: out[e][r] = max(out[e][r], max over cells |value|)             (the Jmax of :488; group ignored)
 * `out` is accumulated into (zero it first); workspace = dvf_gene_gemm_workspace_bytes(n, D, Ge, Gr, 0, 1).
 * n must be even and J 16-byte aligned (TMA row pitch). */
int dvf_gene_transform_reduce(const double* J, int64_t n, int D, const double* Qe, int Ge, const double* Qr, int Gr,
                              int take_abs, int op, const float* jmax, const int32_t* group, int n_groups,
                              double* out, void* ws, size_t ws_bytes, void* stream);
int dvf_gene_transform_elem(const double* J, int64_t n, int D, const double* qe, const double* qr,
                            double* out_elem, void* stream);

/* K16: out (n x G) = V (n x D) Q^T, Q is (G x D).   replaces vectorfield/utils.py:311-326. */
int dvf_vector_transform(const double* V, int64_t n, int D, const double* Q, int G, double* out,
                         void* stream);

/* ---- K9: per-cell post-processing (Python loops over cells in the reference) ---------------
 * dvf_jacobian_det: det[c] = det(J[:, :, c]), J is (D, D, n) cell-fastest.  replaces the per-cell np.linalg.det of
 *   vectorfield/vector_calculus.py:329 (LU with partial pivoting, pivot rule and reciprocal scaling of LAPACK
 *   dgetf2; the product of the pivots is accumulated with a separate exponent).
 * dvf_cell_angles: out[c] = angle(U[c], V[c]) of vectorfield/utils.py:1148-1163 (called per cell at
 *   vectorfield/VectorField.py:255-259): sign * arccos(clip(u.v / |u||v|)), sign = -sign of the 2 x 2 determinant
 *   of the last two unit-vector components (1 if it is zero), NaN if either norm is zero.  U, V are n x D.
 * dvf_row_norms: out[c] = |V[c]|  (vector_calculus.py:192, `speed`). */
int dvf_jacobian_det(const double* J, int64_t n, int D, double* det, void* stream);
int dvf_cell_angles(const double* U, const double* V, int64_t n, int D, double* out, void* stream);
int dvf_row_norms(const double* V, int64_t n, int D, double* out, void* stream);
/* Cell reductions of the materialised gene-space Jacobian J (R = Ge*Gr rows of n float32 cells, the reference's
 * `jacobian_gene` layout) in NumPy's own float32 summation orders, so that the means rank_jacobian_genes sorts
 * (vectorfield/rank_vf.py:483-508) are bit-identical:
 *   dvf_f32_row_means_pairwise: out[r] = np.mean(J[r]) -- pairwise float32 sum (leaves: n_leaves triples
 *     {first cell, length, additions that close after this leaf} of NumPy's pairwise tree for n, built by the host);
 *   dvf_f32_group_means_seq: out[g][r] = running float32 sum of J[r][c] over the cells with label[c] == g, in cell
 *     order, divided by count[g] (vectorfield/utils.py:557-582, average_jacobian_by_group); label < 0 skips a cell. */
int dvf_f32_row_means_pairwise(const float* J, int64_t R, int64_t n, const int* leaves, int n_leaves,
                               float* out, void* stream);
int dvf_f32_group_means_seq(const float* J, int64_t R, int64_t n, const int* label, int G, const float* count,
                            float* out, void* stream);
/* bandwidth_selector's neighbour search (scVectorField.py:121-135 -> tools/connectivity.py:565-577, exact kNN): out
 * (M x kk, row-major) = Euclidean distances from control point i to its kk nearest control points, itself included,
 * ascending.  Squared distances are summed dimension by dimension with separate multiply / add (sklearn's order). */
int dvf_knn_dists(const double* Xc, int M, int D, int kk, double* out, void* stream);

/* ---- K10: second derivatives of the field ----------------------------------------------------
 * dvf_hessian: H[c][i][l][m] = d^2 f_i / dx_l dx_m at x_c  (n x D x D x D, cell-major).  replaces
 *   Hessian_rkhs_gaussian, vectorfield/utils.py:589-614, and the per-cell loop of vector_calculus.py:413-417.
 * dvf_laplacian: L[c][k] = sum_a H[c][a][a][k], what the reference's Laplacian() (utils.py:665-687) returns for the
 *   Hs[i, l, m, cell] array of vector_calculus.py:413 -- computed directly from the field, no Hessian array. */
int dvf_hessian(const double* x, int64_t n, int D, const double* Xc, int M, double beta, const double* C,
                double* H, void* stream);
int dvf_laplacian(const double* x, int64_t n, int D, const double* Xc, int M, double beta, const double* C,
                  double* L, void* stream);

/* ---- K11: batched initial-value integration through the field ------------------------------
 * One adaptive Dormand-Prince 5(4) integration per initial state y0[i, :] of  y' = scale * f(y)  from t0 to t_bound,
 * with the step control, the terminal "at rest" event all(|scale f| < rest_tol) and the dense-output coefficients of
 * the scipy.integrate.solve_ivp call the reference makes per cell (prediction/utils.py:103-165, integrate_vf_ivp;
 * event :107; max_step :96; right-hand side vectorfield/utils.py:73-142).  Per trajectory i (cap = step capacity):
 *   ts[i, 0..k]      accepted times (k = n_steps[i] <= cap), ts row length cap + 2
 *   ys[i, 0..k, :]   states at those times, (cap + 1) x D per trajectory
 *   Q[i, s, :, 0..3] interpolant of step s:  y(t) = ys[s] + h (Q [x, x^2, x^3, x^4]),  x = (t - ts[s]) / h,
 *                    h = ts[s+1] - ts[s]
 *   status[i]        0 reached t_bound, 1 stopped by the event, -1 step size below 10 ulp(t) (scipy status -1),
 *                    -2 more than cap steps (call again with a larger cap), -3 stopped by the event at the START of
 *                    the first step (the initial state and the end of the first step are both at rest): k = 1,
 *                    ts[i, 1] = t0, ys[i, 1] = y0, and ts[i, 2] holds the end of the step Q[i, 0] was built on
 * workspace: dvf_ivp_workspace_bytes(D, M) (a transposed copy of Xc and a padded copy of C; 16-byte aligned);
 * M <= 5000.  A handful of initial states run one per CTA, larger batches four per CTA in lock step; both give the
 * same trajectories (every sum is formed in the same order). */
int dvf_ivp_workspace_bytes(int D, int M, size_t* out);
int dvf_integrate_rk45(const double* y0, int64_t n, int D, const double* Xc, int M, double beta, const double* C,
                       double scale, double t0, double t_bound, double max_step, double rtol, double atol,
                       double rest_tol, int cap, double* ts, double* ys, double* Q, int* n_steps, int* status,
                       void* ws, size_t ws_bytes, void* stream);
/* The continuous solution of trajectories integrated by dvf_integrate_rk45 (same ts / ys / Q / n_steps / status / cap),
 * evaluated on the device: row r of t_eval (m x S, n_eval[r] <= S valid entries, NULL = all S) belongs to trajectory
 * traj[r]; out (m, D, S) = what scipy's OdeSolution.__call__ returns for it (segment choice of numpy.searchsorted as in
 * scipy/integrate/_ivp/common.py, clipped extrapolation from the first / last segment).  replaces `sol(self.t)` of
 * prediction/trajectory.py:167-176,213-224 and `SOL[i](valid_t_trans)` of prediction/utils.py:231-242. */
int dvf_ivp_dense_eval(const double* ts, const double* ys, const double* Q, const int* n_steps, const int* status,
                       int D, int cap, const int* traj, const double* t_eval, const int* n_eval, int64_t m, int S,
                       double* out, void* stream);

/* ---------------------------------------------------------------------------------------------
 * SparseVFC EM fit (vectorfield/scVectorField.py:417-497; Gaussian kernel branch)
 * ------------------------------------------------------------------------------------------- */
typedef struct dvf_em_params {
    double a;        /* outlier uniform-distribution volume, default 5            (:315) */
    double gamma;    /* initial inlier fraction, default 0.9                       (:318) */
    double lambda_;  /* regularisation weight, default 3                           (:319) */
    double minP;     /* posterior floor, default 1e-5                              (:320) */
    double ecr;      /* energy change-rate stop criterion, default 1e-5            (:317) */
    double theta;    /* inlier threshold on P, default 0.75                        (:322) */
    int max_iter;    /* MaxIter, default 500 (30 through VectorField)              (:321) */
    int check_every; /* iterations launched between host reads of the stop flag; 0 = default (8) */
    int solver;      /* DVF_SOLVER_*: how the M x M system of the M-step is solved (lstsq_method, :332)  */
    int reserved;
} dvf_em_params;

/* lstsq_method="drouin" (the reference's default, tools/utils.py:2948-2981): normal equations
 * solve(lhs^T lhs, lhs^T rhs) by LU with partial pivoting -- squares cond(lhs) exactly as the reference does. */
#define DVF_SOLVER_NORMAL_LU 1
/* Cholesky factorisation of the symmetric positive definite lhs itself: same solution in exact arithmetic,
 * error cond(lhs) eps instead of cond(lhs)^2 eps (lstsq_method="cholesky", an addition of this backend). */
#define DVF_SOLVER_CHOLESKY 0

/* scalars the EM loop leaves on the host when it returns */
typedef struct dvf_em_result {
    int n_iter;          /* number of loop bodies executed (reference `i` at exit)             */
    int status;          /* 0 ok, DVF_ERR_NUMERIC if the solver failed                          */
    double sigma2;       /* final sigma2                                               (:486)   */
    double gamma;        /* final gamma                                                (:489)   */
    double n_inlier;     /* #(P > theta) of the last iteration, global over ranks               */
} dvf_em_result;

/* width (in doubles) of one row of the fused [U | Y | 0-pad] matrix for given M, D */
int64_t dvf_fit_row_width(int M, int D);
/* bytes of workspace dvf_fit_em needs for N_local cells (includes the N_local x width U matrix) */
int dvf_fit_workspace_bytes(int64_t N_local, int D, int M, size_t* out);

/* Whole EM loop on the device.  X, Y: this rank's cells (N_local x D); N_total: cells over all ranks;
 * Xc: control points (M x D); on return C (M x D), P (N_local, clamped >= minP), V (N_local x D) = U C,
 * E_traj / tecr_traj (max_iter doubles each, first n_iter valid).  comm: handle from dvf_comm_init or NULL
 * (single GPU).  The stop flag is read back every `check_every` iterations; kernels launched after the
 * flag is raised exit immediately.  replaces scVectorField.py:417-497 and tools/utils.py:2948-2981. */
int dvf_fit_em(const double* X, const double* Y, int64_t N_local, int64_t N_total, int D,
               const double* Xc, int M, double beta, const dvf_em_params* prm, double* C, double* P,
               double* V, double* E_traj, double* tecr_traj, dvf_em_result* res, void* ws,
               size_t ws_bytes, void* comm, void* stream);

/* Step-level entry points (used by the teacher-forced parity tests; the loop above is built from them).
 *   UY is the N x W row-major matrix [U | Y | 0], W = dvf_fit_row_width(M, D).                       */
int dvf_build_uy(const double* X, const double* Y, int64_t N, int D, const double* Xc, int M,
                 double beta, double* UY, void* stream);
/* E-step, scVectorField.py:197-227: r2 (N) -> P (N, unclamped) ; stats[0..3] = {sum P r2/(2 sigma2) +
 * sum P log(sigma2) D/2 (= E of get_P), sum P, min nonzero t1, number of exact zeros}. */
int dvf_estep(const double* r2, int64_t N, int D, double sigma2, double gamma, double a, double* P,
              double* stats, void* stream);
/* M-step assembly, scVectorField.py:468-476: G (W x W, lower triangle of the columns < M valid) = UY^T diag(max(P, minP)) UY.
 * scratch: dvf_gram_scratch_bytes(). */
/* G is ldg x ldg with ldg = dvf_gram_ld(D, M) >= W (row-major; only lower-triangular tiles are written).
 * Valid on return: G[i][j] for j <= i, j < M -- i.e. lhs' Gram part G[:M,:M] (lower triangle) and rhs^T = G[M:M+D,:M].
 * The Y^T diag(P) Y block (i, j >= M) is read by nothing on the path and is left unspecified.
 * P must be 16-byte aligned and readable up to an even element count (allocate N + 1 doubles if N is odd). */
int dvf_gram_ld(int D, int M);
int dvf_gram_scratch_bytes(int64_t N, int D, int M, size_t* out);
int dvf_gram(const double* UY, const double* P, double minP, int64_t N, int D, int M, double* G,
             void* scratch, size_t scratch_bytes, void* stream);
/* Solve (G[:M,:M] + lam_sigma2 * Kc) C = G[M:M+D, :M]^T.  scVectorField.py:475-481; see DESIGN.md for the
 * choice of a Cholesky solve of the M x M system instead of the reference's normal equations.
 * L: (M + D) x M scratch.  info (device int): 0 ok, k>0 = non-positive pivot at column k-1. */
int dvf_solve(const double* G, int64_t ldg, const double* Kc, int M, int D, double lam_sigma2,
              double* L, double* C, int* info, void* stream);
/* Allocation-free variant of dvf_update_v.  It has NO counterpart in the reference: it was added
 * for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  Same arguments and results as dvf_update_v; the scratch (padded C + per-CTA partial sums, 768 KB) is a
 * file-scope device array that every GPU loading the library holds until exit, so the call performs no cudaMalloc /
 * cudaFree / pool allocation.  Limits: M <= 1024, D <= 64, N <= 4 194 304 -- beyond them DVF_ERR_WORKSPACE with a
 * message.  All calls share the one array: calls that overlap on the GPU race, serialise them. */
int dvf_update_v_noalloc(const double* UY, const double* C, const double* P, double minP, int64_t N, int D,
                         int M, double* V, double* r2, double* stats, void* stream);
/* The reference's solver form (DVF_SOLVER_NORMAL_LU) as a step: C = solve(lhs^T lhs, lhs^T rhs), LU with partial
 * pivoting (tools/utils.py:2975-2976 -> LAPACK gesv).  ws: dvf_solve_normal_workspace_bytes(); info: 0 ok, k>0 =
 * exactly zero pivot at column k-1 (numpy raises LinAlgError("Singular matrix") there). */
int dvf_solve_normal_workspace_bytes(int M, int D, size_t* out);
int dvf_solve_normal(const double* G, int64_t ldg, const double* Kc, int M, int D, double lam_sigma2,
                     void* ws, size_t ws_bytes, double* C, int* info, void* stream);
/* V = U C, r2 = |Y - V|^2 per cell, stats[0] = sum_n max(P,minP)_n r2_n.   scVectorField.py:484-486 */
int dvf_update_v(const double* UY, const double* C, const double* P, double minP, int64_t N, int D,
                 int M, double* V, double* r2, double* stats, void* stream);

/* ---------------------------------------------------------------------------------------------
 * multi-GPU: one NCCL communicator owned by the library (cells sharded over ranks; the only traffic is
 * the allreduce of the Gram/RHS block and a few scalars per EM iteration)
 * ------------------------------------------------------------------------------------------- */
int dvf_comm_unique_id(void* out128);                          /* rank 0: 128-byte ncclUniqueId */
int dvf_comm_init(const void* id128, int rank, int world, void** comm);
int dvf_comm_destroy(void* comm);
int dvf_comm_allreduce_sum(void* comm, double* buf, int64_t count, void* stream);

/* Strided (pitched) asynchronous copy between device and pinned host memory: `height` rows of `width_bytes`.
 * Used to stream the (D, D, n) Jacobian to the host in cell chunks while the next chunk is computed.
 * to_host != 0: device -> host, else host -> device. */
int dvf_memcpy2d_async(void* dst, size_t dpitch, const void* src, size_t spitch, size_t width_bytes, size_t height,
                       int to_host, void* stream);

/* ---- host helper of the prelude -------------------------------------------------------------------------
 * Control-point sampling (reference tools/sampling.py:242-245 -> numpy Generator.choice(p=..., replace=False)):
 * cdf[i] = (p[0] + ... + p[i]) / total, accumulated strictly left to right in double precision exactly as
 * numpy.cumsum does, so that searchsorted() picks the same cells as the reference.  HOST pointers, no CUDA. */
int dvf_host_cdf(const double* p, int64_t n, double* cdf);
/* p[i] = v[i] / total (tools/sampling.py:244, p = |v| / sum|v|); element-wise, split over a few host threads. */
int dvf_host_scale(const double* v, int64_t n, double total, double* p);
/* numpy.sum(a) of a contiguous float64 vector bit for bit (NumPy's pairwise summation tree; the top levels of the
 * recursion run on threads).  tools/sampling.py:244 (np.sum(tmp_V)). */
int dvf_host_pairwise_sum(const double* a, int64_t n, double* out);
/* raw[i] = sequential cumulative sum of p_k = v[k] / total, with p_k = 0 for the ascending indices in `skip`
 * (the cells already drawn): Generator.choice(replace=False)'s cumsum(p) without storing p.  Every partial sum has the
 * bits of numpy.cumsum's sequential recurrence; the work is spread over host threads (csrc/host_cumsum.cpp). */
int dvf_host_choice_cumsum(const double* v, int64_t n, double total, const int64_t* skip, int64_t n_skip, double* raw);
/* out[q] = searchsorted(raw / raw[n-1], x[q], side='right') by bisection on the fly (no quotient array). */
int dvf_host_searchsorted_div(const double* raw, int64_t n, const double* x, int64_t m, int64_t* out);

#ifdef __cplusplus
}
#endif
#endif

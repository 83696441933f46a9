// K6: per-cell analytical Jacobian of the Gaussian-RKHS field (reference utils.py:351-379)
//     J_c[i][j] = -2 beta sum_m k(x_c, Xc_m) C[m][i] (x_c[j] - Xc[m][j])
// written as ONE dense FP64 tensor-core GEMM over all cells.  With any fixed shift s (the centroid of the
// control points, which keeps the two terms small),
//     sum_m k_cm C_mi (x_cj - c_mj) = v_ci (x_cj - s_j) - sum_m k_cm T[m][(i, j)],   T[m][(i, j)] = C_mi (c_mj - s_j),
//     v_ci = sum_m k_cm C_mi,
// so the (D*D) x n matrix J is (T_ext)^T (M x rows) times K^T (M x n) plus a rank-one correction per cell; the rows
// that produce v_i are appended to the row tile that needs them, so one pass over K gives both.
// The output layout (D, D, n), cell index fastest, is the reference's (utils.py:368) and is exactly the row-major
// GEMM result: every row of 128 cells is a contiguous 1 KB run.
//   pre-pass : K = exp(-beta |x - c|^2) for a chunk of cells (k_kernel_matrix, direct distance form like cdist)
//   GEMM     : persistent CTAs, 16 consumer warps (32 x 32 block each, 128 rows x 128 cells per item) + 1 TMA producer
//              thread, 3-stage ring of 32 control points, 128-byte swizzled boxes; the k-step -> control-point mapping
//              m(s, t) = 8 (s/2) + 2 (s&1) + (t&1) + 4 (t/2) together with the in-tile row permutation (g0, g2, g1)
//              makes both the A loads (control point = box row) and the B loads (control point = box column)
//              bank-conflict free (ncu: 4 -> 2 shared wavefronts per LDS.64).
// The direct per-cell form (kern_jacobian_direct.cu) is kept for cross-checks: DVF_JAC_DIRECT=1.
#include <cstdlib>
#include "common.cuh"
#include "tma.cuh"

int dvf_jacobian_direct_impl(const double* x, int64_t n, int D, const double* Xc, int M, double beta, const double* C,
                             double* J, void* stream);
extern "C" int dvf_kernel_matrix_padded(const double* x, int64_t n, int D, const double* Xc, int M, double beta, double* K,
                                        int64_t ldk, void* stream);

namespace dvf {

constexpr int JG_STAGES = 3;
constexpr int JG_CWARPS = 16;
constexpr int JG_THREADS = (JG_CWARPS + 1) * 32;
constexpr int JG_KC = 32;                                   // control points per stage
constexpr int JG_STAGE_B = 16 * TMA_BOX_BYTES;              // A panel: 8 boxes (128 rows), B panel: 8 boxes (128 cells)
constexpr int JG_VROWS = 12;                                // v rows a tile can hold (n_i <= 11 for D >= 10, <= D below)
constexpr size_t JG_SMEM = (size_t)JG_STAGES * JG_STAGE_B + JG_VROWS * 128 * 8 + 1024 + 64;
constexpr int64_t JG_CHUNK = 131072;                        // cells per K chunk (0.5 GB at M = 500)

struct jg_ring { int stage; unsigned phase; };
__device__ __forceinline__ void jg_advance(jg_ring& r) { if (++r.stage == JG_STAGES) { r.stage = 0; r.phase ^= 1u; } }

// rows of tile t: group gl = r / (D+1) -> i = t * n_i + gl, e = r % (D+1): e < D is the (i, j=e) entry, e == D is v_i
__host__ __device__ inline int jg_ni(int D) { int n = 128 / (D + 1); return n > D ? D : n; }

// s[d] = mean_m Xc[m][d]  (any shift is algebraically exact; the centroid keeps |c - s| small)
__global__ void k_jg_centroid(const double* __restrict__ Xc, int M, int D, double* __restrict__ s) {
    const int d = blockIdx.x * blockDim.x + threadIdx.x;
    if (d >= D) return;
    double a = 0.0;
    for (int m = 0; m < M; ++m) a += Xc[(int64_t)m * D + d];
    s[d] = a / M;
}

// T_ext[m][t * 128 + r] for m < Mp (zero rows past M), zero for the padding rows of a tile
__global__ void k_jg_build_t(const double* __restrict__ Xc, const double* __restrict__ C, const double* __restrict__ s,
                             int M, int Mp, int D, int n_tiles, double* __restrict__ T) {
    const int ncols = n_tiles * 128, n_i = jg_ni(D);
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < (int64_t)Mp * ncols;
         e += (int64_t)gridDim.x * blockDim.x) {
        const int m = (int)(e / ncols), col = (int)(e % ncols);
        const int t = col >> 7, r = col & 127, gl = r / (D + 1), q = r % (D + 1), i = t * n_i + gl;
        double v = 0.0;
        if (m < M && gl < n_i && i < D) v = (q < D) ? C[(int64_t)m * D + i] * (Xc[(int64_t)m * D + q] - s[q]) : C[(int64_t)m * D + i];
        T[e] = v;
    }
}

__global__ void __launch_bounds__(JG_THREADS, 1)
k_jacobian_gemm(const __grid_constant__ CUtensorMap mapT, const __grid_constant__ CUtensorMap mapK,
                const double* __restrict__ x, const double* __restrict__ s, int64_t n_chunk, int64_t cell_base,
                int64_t n_total, int D, int n_tiles, int k_stages, double scale, double* __restrict__ J) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    unsigned char* ring = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    double* vbuf = reinterpret_cast<double*>(ring + JG_STAGES * JG_STAGE_B);          // [JG_VROWS][128]
    uint64_t* full = reinterpret_cast<uint64_t*>(vbuf + JG_VROWS * 128);
    uint64_t* empty = full + JG_STAGES;
    if (threadIdx.x == 0) {
        for (int q = 0; q < JG_STAGES; ++q) { mbar_init(&full[q], 1); mbar_init(&empty[q], JG_CWARPS); }
        mbar_fence_init();
    }
    __syncthreads();

    const int64_t n_cb = (n_chunk + 127) / 128, n_items = n_cb * n_tiles;
    jg_ring rs{0, 0u};
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (warp == JG_CWARPS) {
        // ---------------- producer: one thread ----------------
        if (lane != 0) return;
        for (int64_t item = blockIdx.x; item < n_items; item += gridDim.x) {
            const int cb = (int)(item / n_tiles);
            const int t = (int)((item % n_tiles + cb) % n_tiles);     // rotated: see the consumer side
            for (int ks = 0; ks < k_stages; ++ks) {
                mbar_wait(&empty[rs.stage], rs.phase ^ 1u);
                const unsigned sb = smem_u32(ring) + rs.stage * JG_STAGE_B, bar = smem_u32(&full[rs.stage]);
                mbar_arrive_expect_tx(&full[rs.stage], JG_STAGE_B);
#pragma unroll
                for (int c = 0; c < 8; ++c) tma_load_2d(sb + c * TMA_BOX_BYTES, &mapT, t * 128 + 16 * c, ks * JG_KC, bar);
#pragma unroll
                for (int c = 0; c < 8; ++c)   // box (mc, cg) = (c / 4, c % 4): 16 control points x 32 cells
                    tma_load_2d(sb + (8 + c) * TMA_BOX_BYTES, &mapK, ks * JG_KC + 16 * (c >> 2), cb * 128 + 32 * (c & 3), bar);
                jg_advance(rs);
            }
        }
        return;
    }

    // ---------------- consumers: warp (wi, wj) owns rows wi*32.., cells wj*32.. ----------------
    const int g = lane >> 2, tq = lane & 3, wi = warp >> 2, wj = warp & 3;
    // control point of k-step s, lane t:  m = 8 (s/2) + 2 (s&1) + (t&1) + 4 (t/2)
    // DMMA row/column g of a tile is row / cell pg = (g0, g2, g1) of the 8-row group (bits 1 and 2 swapped): with the
    // 128-byte swizzle this puts the 16 lanes of each half-warp on 16 distinct 8-byte bank pairs for A and for B
    const int pg = (g & 1) | (((g >> 1) & 1) << 2) | ((g >> 2) << 1);
    const int pc = 4 * (tq & 1) + 2 * (tq >> 1);             // cells pc, pc + 1 of a tile are this lane's accumulators
    unsigned toffA[2][2], toffB[2][2];
#pragma unroll
    for (int sp = 0; sp < 2; ++sp) {
        const int f = 2 * sp + (tq & 1) + 4 * (tq >> 1);      // (m & 7)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            // A box: row = control point, column = 8 h + pg           (h = DMMA tile parity inside the 16-column box)
            toffA[sp][h] = f * 128 + ((((h << 2) | (pg >> 1)) ^ f) << 4) + (pg & 1) * 8;
            // B box: row = cell pg (of an 8-cell tile), column = 8 h + 2 sp + (t&1) + 4 (t/2)   (h = (s/2) & 1)
            toffB[sp][h] = pg * 128 + (((4 * h + sp + 2 * (tq >> 1)) ^ pg) << 4) + (tq & 1) * 8;
        }
    }
    const unsigned wbaseA = (wi * 2) * TMA_BOX_BYTES;
    const unsigned wbaseB = (8 + wj) * TMA_BOX_BYTES;
    const int n_i = jg_ni(D);

    for (int64_t item = blockIdx.x; item < n_items; item += gridDim.x) {
        // the row tile is rotated by the cell block so that every CTA meets all tiles equally often (the grid stride
        // and n_tiles share factors, and the last tile is cheaper than the others)
        const int64_t cb = item / n_tiles;
        const int t = (int)((item % n_tiles + cb) % n_tiles);
        double acc[4][4][2];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int q = 0; q < 4; ++q) acc[i][q][0] = acc[i][q][1] = 0.0;

        // rows of this tile that exist: the last tile usually holds fewer groups (D = 30: 2 of 4 -> 62 rows), and a
        // warp whose 32 rows are all padding only keeps the ring protocol going; warp % 4 = wj, so every scheduler
        // keeps the same number of working warps and the tile costs the DMMA pipe only its real rows
        const int rows_here = min(n_i, D - t * n_i) * (D + 1);
        const bool active = wi * 32 < rows_here;
        for (int ks = 0; ks < k_stages; ++ks) {
            mbar_wait(&full[rs.stage], rs.phase);
            const unsigned char* sb = ring + rs.stage * JG_STAGE_B;
            if (active)
#pragma unroll
            for (int sk = 0; sk < 8; ++sk) {
                const int sp = sk & 1;
                double af[4], bf[4];
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    af[i] = *reinterpret_cast<const double*>(sb + wbaseA + (i >> 1) * TMA_BOX_BYTES + (sk >> 1) * 1024 +
                                                             toffA[sp][i & 1]);
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    bf[q] = *reinterpret_cast<const double*>(sb + wbaseB + (sk >> 2) * (4 * TMA_BOX_BYTES) + q * 1024 +
                                                             toffB[sp][(sk >> 1) & 1]);
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int q = 0; q < 4; ++q) dmma(acc[i][q][0], acc[i][q][1], af[i], bf[q]);
            }
            __syncwarp();
            fence_proxy_async_smem();
            if (lane == 0) mbar_arrive(&empty[rs.stage]);
            jg_advance(rs);
        }

        // ---- epilogue: v rows -> shared, then J = scale * (v_i (x_j - s_j) - acc) ----
        int rgl[4], re[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int r = wi * 32 + i * 8 + pg;
            rgl[i] = r / (D + 1);
            re[i] = r % (D + 1);
            if (rgl[i] >= n_i || t * n_i + rgl[i] >= D) rgl[i] = -1;     // padding row
        }
#pragma unroll
        for (int i = 0; i < 4; ++i)
            if (rgl[i] >= 0 && re[i] == D) {
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    *reinterpret_cast<double2*>(vbuf + rgl[i] * 128 + wj * 32 + q * 8 + pc) =
                        make_double2(acc[i][q][0], acc[i][q][1]);
            }
        named_barrier_sync(1, JG_CWARPS * 32);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int cl = wj * 32 + q * 8 + pc;
            const int64_t c0 = cb * 128 + cl;                          // cell inside the chunk
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                if (rgl[i] < 0 || re[i] == D) continue;
                const int gi = t * n_i + rgl[i], j = re[i];
                const double2 v = *reinterpret_cast<const double2*>(vbuf + rgl[i] * 128 + cl);
                const double sj = s[j];
                double* out = J + ((int64_t)gi * D + j) * n_total + cell_base + c0;
                if (c0 < n_chunk) out[0] = scale * (v.x * (x[(cell_base + c0) * D + j] - sj) - acc[i][q][0]);
                if (c0 + 1 < n_chunk) out[1] = scale * (v.y * (x[(cell_base + c0 + 1) * D + j] - sj) - acc[i][q][1]);
            }
        }
        named_barrier_sync(1, JG_CWARPS * 32);
    }
}

}  // namespace dvf

using namespace dvf;

static bool jacobian_uses_direct(int64_t n, int D, int M) {
    // The GEMM form pays a K pre-pass and a 128-row tile granularity: it wins once the contraction is long and the
    // tiles are full (measured on B200: 1M x 30, M=500: 39.9 vs 49.9 ms; 200k x 50, M=200: 11.6 vs 19.5 ms; but
    // 100k x 30, M=100: 3.3 vs 1.7 ms).  DVF_JAC_DIRECT=1 / 0 forces one form (cross-checks, tools/jac_tune.py).
    static const char* force = getenv("DVF_JAC_DIRECT");
    return force ? atoi(force) != 0 : !(M >= 192 && D >= 12 && n >= 16384);
}

struct jg_layout { int Mp, n_tiles, ncols; int64_t chunk; size_t t_bytes, s_bytes, total; };
static jg_layout jg_make_layout(int64_t n, int D, int M) {
    jg_layout l;
    l.Mp = (M + JG_KC - 1) / JG_KC * JG_KC;
    const int n_i = jg_ni(D);
    l.n_tiles = (D + n_i - 1) / n_i;
    l.ncols = l.n_tiles * 128;
    l.chunk = n < JG_CHUNK ? n : JG_CHUNK;
    l.t_bytes = ((size_t)l.Mp * l.ncols * 8 + 255) & ~(size_t)255;
    l.s_bytes = 1024;
    l.total = l.t_bytes + l.s_bytes + (size_t)l.chunk * l.Mp * 8;
    return l;
}

// scratch: T_ext (Mp x ncols) | centroid | K chunk (chunk x Mp)
extern "C" int dvf_jacobian_workspace_bytes(int64_t n, int D, int M, size_t* out) {
    DVF_ARG(out && n >= 0 && D >= 1 && D <= kMaxD && M >= 1, "bad argument");
    *out = (n == 0 || jacobian_uses_direct(n, D, M)) ? 0 : jg_make_layout(n, D, M).total;
    return DVF_OK;
}

extern "C" int dvf_jacobian_ws(const double* x, int64_t n, int D, const double* Xc, int M, double beta, const double* C,
                               double* J, void* ws, size_t ws_bytes, void* stream) {
    DVF_ARG(n >= 0 && M >= 1, "n=%lld M=%d: need n >= 0, M >= 1", (long long)n, M);
    DVF_ARG(D >= 1 && D <= kMaxD, "D=%d unsupported (1..%d)", D, kMaxD);
    DVF_ARG((x || n == 0) && Xc && C && (J || n == 0), "NULL pointer");
    if (n == 0) return DVF_OK;
    if (jacobian_uses_direct(n, D, M)) return dvf_jacobian_direct_impl(x, n, D, Xc, M, beta, C, J, stream);
    const jg_layout l = jg_make_layout(n, D, M);
    if (!ws || ws_bytes < l.total) {
        set_error("jacobian workspace too small: %zu < %zu", ws_bytes, l.total);
        return DVF_ERR_WORKSPACE;
    }
    DVF_ARG(((uintptr_t)ws & 255) == 0, "workspace must be 256-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    static bool attr_set = false;
    if (!attr_set) {
        DVF_CUDA(cudaFuncSetAttribute(k_jacobian_gemm, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)JG_SMEM));
        attr_set = true;
    }
    char* scratch = (char*)ws;
    double* T = (double*)scratch;
    double* sc = (double*)(scratch + l.t_bytes);
    double* K = (double*)(scratch + l.t_bytes + l.s_bytes);
    k_jg_centroid<<<(D + 63) / 64, 64, 0, st>>>(Xc, M, D, sc);
    DVF_LAUNCH_CHECK();
    k_jg_build_t<<<num_sms() * 4, 256, 0, st>>>(Xc, C, sc, M, l.Mp, D, l.n_tiles, T);
    DVF_LAUNCH_CHECK();
    CUtensorMap mapT, mapK;
    if (int rc = make_tmap_2d(&mapT, T, (uint64_t)l.ncols, (uint64_t)l.Mp, (uint64_t)l.ncols * 8)) return rc;
    for (int64_t c0 = 0; c0 < n; c0 += l.chunk) {
        const int64_t nc = (n - c0 < l.chunk) ? (n - c0) : l.chunk;
        if (int rc = dvf_kernel_matrix_padded(x + c0 * D, nc, D, Xc, M, beta, K, l.Mp, stream)) return rc;
        if (int rc = make_tmap_2d(&mapK, K, (uint64_t)l.Mp, (uint64_t)nc, (uint64_t)l.Mp * 8)) return rc;
        const int64_t n_items = (nc + 127) / 128 * l.n_tiles;
        const int grid = (int)(n_items < num_sms() ? n_items : num_sms());
        k_jacobian_gemm<<<grid, JG_THREADS, JG_SMEM, st>>>(mapT, mapK, x, sc, nc, c0, n, D, l.n_tiles, l.Mp / JG_KC,
                                                           -2.0 * beta, J);
        DVF_LAUNCH_CHECK();
    }
    return DVF_OK;
}

// convenience form without a caller workspace: scratch comes from the device's stream-ordered pool, whose release
// threshold is raised once so that the 0.5 GB chunk buffer is not returned to the driver at every synchronisation
extern "C" int dvf_jacobian(const double* x, int64_t n, int D, const double* Xc, int M, double beta,
                            const double* C, double* J, void* stream) {
    size_t bytes = 0;
    if (int rc = dvf_jacobian_workspace_bytes(n, D, M, &bytes)) return rc;
    if (bytes == 0) return dvf_jacobian_ws(x, n, D, Xc, M, beta, C, J, nullptr, 0, stream);
    void* ws = nullptr;
    if (int rc = pool_alloc(&ws, bytes, (cudaStream_t)stream)) return rc;
    const int rc = dvf_jacobian_ws(x, n, D, Xc, M, beta, C, J, ws, bytes, stream);
    cudaFreeAsync(ws, (cudaStream_t)stream);
    return rc;
}

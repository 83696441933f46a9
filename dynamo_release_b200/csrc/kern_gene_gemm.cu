// K8 on FP64 tensor cores: gene-space transform of per-cell Jacobians (reference utils.py:544-554, :497-541)
//     out[e][r][c] = float32( sum_ij Qe[e][i] J[i][j][c] Qr[r][j] )            J (D, D, n), out (Ge, Gr, n), cell fastest
// factored into two batched tall-skinny DGEMMs whose results are already in the layouts the next step / the reference
// want (every output row is a contiguous run of cells):
//     stage 1, batch j :  T[(e, j)][c]  = sum_i QeT[i][e] J[i][j][c]        rows e, contraction i   (2 Ge D^2 flops / cell)
//     stage 2, batch e :  out[(e, r)][c] = sum_j QrT[j][r] T[(e, j)][c]      rows r, contraction j   (2 Ge Gr D flops / cell)
// Both are  C[row][cell] = sum_k A[k][row] B[k][cell]  with K = D <= 64: the Gram pipeline (16 consumer warps x 32x32,
// TMA producer thread, 128-byte swizzled boxes, conflict-free k-step -> row mapping r = 8 (s/2) + 2 t + (s&1), proxy
// fence on stage release) with a ring that keeps running across the (cell tile, batch, row tile) items, because one
// item has only ceil(D/32) <= 2 stages.  Cells are processed in chunks so that T (Ge D doubles per cell) stays small.
// Group sums (rank_jacobian_genes with groups) are reduced from the float32 values by k_gene_group_sum, one warp per
// (e, r) row in a fixed order: no atomics, bitwise reproducible.
#include <algorithm>
#include <cstdlib>
#include "common.cuh"
#include "tma.cuh"

namespace dvf {

constexpr int TS_STAGES = 3;
constexpr int TS_CWARPS = 16;
constexpr int TS_THREADS = (TS_CWARPS + 1) * 32;
constexpr int TS_STAGE_B = 16 * TMA_BOX_BYTES;     // A panel 8 boxes (32 k x 128 rows) + B panel 8 boxes (32 k x 128 cells)
constexpr size_t TS_SMEM = (size_t)TS_STAGES * TS_STAGE_B + 1024 + 64;

struct ts_params {
    int K;                  // contraction length
    int nb, nrt;            // batches, row tiles (cell tiles follow from `cells`)
    int rows;               // valid rows of A / the output (Ge or Gr)
    int64_t cells;          // valid cells of this chunk
    int64_t b_cell0;        // B, rank 3: first cell of the chunk inside J
    int b_rows_per_batch;   // B, rank 2: first k-row of batch b is b * b_rows_per_batch
    void* out;              // double (stage 1) or float (stage 2)
    int64_t ld;             // output row pitch in elements
    int64_t out_cell0;      // first output column of this chunk
    int64_t row_mul, b_mul; // output row index = (row tile * 128 + row) * row_mul + b * b_mul
    int take_abs;
};

struct ts_ring { int stage; unsigned phase; };
__device__ __forceinline__ void ts_advance(ts_ring& r) { if (++r.stage == TS_STAGES) { r.stage = 0; r.phase ^= 1u; } }

template <bool OUT_F32, bool B3D>
__global__ void __launch_bounds__(TS_THREADS, 1)
k_tsgemm(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB, const ts_params p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    unsigned char* ring = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint64_t* full = reinterpret_cast<uint64_t*>(ring + TS_STAGES * TS_STAGE_B);
    uint64_t* empty = full + TS_STAGES;
    if (threadIdx.x == 0) {
        for (int q = 0; q < TS_STAGES; ++q) { mbar_init(&full[q], 1); mbar_init(&empty[q], TS_CWARPS); }
        mbar_fence_init();
    }
    __syncthreads();

    const int nct = (int)((p.cells + 127) / 128);
    const int64_t per_ct = (int64_t)p.nb * p.nrt, n_items = per_ct * nct;
    const int k_stages = (p.K + 31) / 32;
    ts_ring rs{0, 0u};
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (warp == TS_CWARPS) {
        // ---------------- producer: one thread ----------------
        if (lane != 0) return;
        for (int64_t item = blockIdx.x; item < n_items; item += gridDim.x) {
            const int ct = (int)(item / per_ct), rem = (int)(item % per_ct), b = rem / p.nrt, rt = rem % p.nrt;
            for (int ks = 0; ks < k_stages; ++ks) {
                mbar_wait(&empty[rs.stage], rs.phase ^ 1u);
                const unsigned sb = smem_u32(ring) + rs.stage * TS_STAGE_B, bar = smem_u32(&full[rs.stage]);
                mbar_arrive_expect_tx(&full[rs.stage], TS_STAGE_B);
#pragma unroll
                for (int c = 0; c < 8; ++c) tma_load_2d(sb + c * TMA_BOX_BYTES, &mapA, rt * 128 + 16 * c, ks * 32, bar);
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    if (B3D) tma_load_3d(sb + (8 + c) * TMA_BOX_BYTES, &mapB, (int)(p.b_cell0 + ct * 128 + 16 * c), b, ks * 32, bar);
                    else tma_load_2d(sb + (8 + c) * TMA_BOX_BYTES, &mapB, ct * 128 + 16 * c, b * p.b_rows_per_batch + ks * 32, bar);
                }
                ts_advance(rs);
            }
        }
        return;
    }

    // ---------------- consumers: warp (wi, wj) owns rows wi*32.., cells wj*32.. of the 128 x 128 item ----------------
    const int g = lane >> 2, t = lane & 3, wi = warp >> 2, wj = warp & 3;
    unsigned toff[2][2];
#pragma unroll
    for (int sp = 0; sp < 2; ++sp)
#pragma unroll
        for (int h = 0; h < 2; ++h)
            toff[sp][h] = (2 * t + sp) * 128 + ((((h << 2) | (g >> 1)) ^ (2 * t + sp)) << 4) + (g & 1) * 8;
    const unsigned wbaseA = (wi * 2) * TMA_BOX_BYTES, wbaseB = (8 + wj * 2) * TMA_BOX_BYTES;

    for (int64_t item = blockIdx.x; item < n_items; item += gridDim.x) {
        const int ct = (int)(item / per_ct), rem = (int)(item % per_ct), b = rem / p.nrt, rt = rem % p.nrt;
        double acc[4][4][2];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int q = 0; q < 4; ++q) acc[i][q][0] = acc[i][q][1] = 0.0;

        for (int ks = 0; ks < k_stages; ++ks) {
            mbar_wait(&full[rs.stage], rs.phase);
            const unsigned char* sb = ring + rs.stage * TS_STAGE_B;
            // k-steps 2u and 2u+1 together cover the stage's rows 8u .. 8u+7 (rows past K are zero-filled)
            const int n_sk = min(8, 2 * ((p.K - 32 * ks + 7) / 8));
#pragma unroll
            for (int sk = 0; sk < 8; ++sk) {
                if (sk < n_sk) {
                    const int sp = sk & 1;
                    double af[4], bf[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        af[i] = *reinterpret_cast<const double*>(sb + wbaseA + (i >> 1) * TMA_BOX_BYTES + (sk >> 1) * 1024 + toff[sp][i & 1]);
#pragma unroll
                    for (int q = 0; q < 4; ++q)
                        bf[q] = *reinterpret_cast<const double*>(sb + wbaseB + (q >> 1) * TMA_BOX_BYTES + (sk >> 1) * 1024 + toff[sp][q & 1]);
#pragma unroll
                    for (int i = 0; i < 4; ++i)
#pragma unroll
                        for (int q = 0; q < 4; ++q) dmma(acc[i][q][0], acc[i][q][1], af[i], bf[q]);
                }
            }
            __syncwarp();
            fence_proxy_async_smem();
            if (lane == 0) mbar_arrive(&empty[rs.stage]);
            ts_advance(rs);
        }

        // ---- epilogue: lane holds rows wi*32 + i*8 + g, cells wj*32 + q*8 + 2t (+1) ----
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int row = rt * 128 + wi * 32 + i * 8 + g;
            if (row >= p.rows) continue;
            const int64_t orow = (int64_t)row * p.row_mul + (int64_t)b * p.b_mul;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int64_t cell = (int64_t)ct * 128 + wj * 32 + q * 8 + 2 * t;
                if (OUT_F32) {
                    float f0 = (float)acc[i][q][0], f1 = (float)acc[i][q][1];     // reference stores float32 (utils.py:546)
                    if (p.take_abs) { f0 = fabsf(f0); f1 = fabsf(f1); }
                    float* o = static_cast<float*>(p.out) + orow * p.ld + p.out_cell0 + cell;
                    if (cell < p.cells) o[0] = f0;
                    if (cell + 1 < p.cells) o[1] = f1;
                } else {
                    // T chunk: pitch is a multiple of 128 cells, columns past `cells` hold zeros (TMA zero fill)
                    double* o = static_cast<double*>(p.out) + orow * p.ld + p.out_cell0 + cell;
                    *reinterpret_cast<double2*>(o) = make_double2(acc[i][q][0], acc[i][q][1]);
                }
            }
        }
    }
}

// QT[k][r] = Q[r][k]   (pitch ldq doubles)
__global__ void k_transpose_q(const double* __restrict__ Q, int G, int D, double* __restrict__ QT, int ldq) {
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < G * D; e += gridDim.x * blockDim.x) {
        const int r = e / D, k = e % D;
        QT[(int64_t)k * ldq + r] = Q[e];
    }
}

// out_group[g][row] += sum over the chunk's cells with label g of f(vals[row][cell]); one warp per row, one pass over the
// row, per-lane per-group partials in shared memory, fixed reduction order (bitwise reproducible, no atomics).
// The per-cell value f follows rank_jacobian_genes (reference vectorfield/rank_vf.py:483-497) on the float32 entries:
//   jmax != NULL : a = v / jmax[row]   in float32      (normalize=True, :487-490)
//   op 0         : f = a                                (mean of J or |J|)
//   op 1         : f = a * a^T if both are negative, else 0, product rounded to float32   (mode='switch', :492-496;
//                  a^T is the (r, e) entry, so the matrix must be square: G = Ge = Gr)
//   op 2         : no groups; out[row] = max(out[row], max over cells of |v|)             (the Jmax of :488)
__global__ void __launch_bounds__(256) k_gene_group_sum(const float* __restrict__ vals, int64_t ld, int64_t col0, int64_t cells,
                                                        int64_t n_rows, const int32_t* __restrict__ group, int n_groups,
                                                        double* __restrict__ out_group, int op, int G,
                                                        const float* __restrict__ jmax) {
    extern __shared__ double part[];                      // [8 warps][n_groups][32 lanes]
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    double* mine = part + (size_t)w * n_groups * 32;
    const int64_t warp0 = (int64_t)blockIdx.x * 8 + w, nwarps = (int64_t)gridDim.x * 8;
    for (int64_t row = warp0; row < n_rows; row += nwarps) {
        const float* v = vals + row * ld + col0;
        if (op == 2) {
            float m = 0.f;
            for (int64_t c = lane; c < cells; c += 32) m = fmaxf(m, fabsf(v[c]));
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
            if (lane == 0 && (double)m > out_group[row]) out_group[row] = (double)m;
            continue;
        }
        const int64_t trow = op == 1 ? (row % G) * G + row / G : row;
        const float* vt = vals + trow * ld + col0;
        const float sc = jmax ? jmax[row] : 1.f, sct = jmax ? jmax[trow] : 1.f;
        for (int gi = 0; gi < n_groups; ++gi) mine[gi * 32 + lane] = 0.0;
        for (int64_t c = lane; c < cells; c += 32) {
            const int gi = group[c];
            if (gi < 0 || gi >= n_groups) continue;
            float a = v[c];
            if (jmax) a = __fdiv_rn(a, sc);
            if (op == 1) {
                float bt = vt[c];
                if (jmax) bt = __fdiv_rn(bt, sct);
                a = (a < 0.f && bt < 0.f) ? __fmul_rn(a, bt) : 0.f;
            }
            mine[gi * 32 + lane] += (double)a;
        }
        // lane gi ends up with group gi's sum, then the read-modify-writes of up to 32 groups go out together
        for (int g0 = 0; g0 < n_groups; g0 += 32) {
            double keep = 0.0;
            for (int gi = g0; gi < min(g0 + 32, n_groups); ++gi) {
                const double s = warp_sum(mine[gi * 32 + lane]);
                if (lane == gi - g0) keep = s;
            }
            if (g0 + lane < n_groups && keep != 0.0) out_group[(int64_t)(g0 + lane) * n_rows + row] += keep;
        }
        __syncwarp();
    }
}

}  // namespace dvf

using namespace dvf;

namespace {
struct gg_layout { int ldqe, ldqr; int64_t chunk; size_t qet, qrt, T, tmp, total; bool need_tmp; };

gg_layout gg_make_layout(int64_t n, int D, int Ge, int Gr, bool full, bool grouped) {
    gg_layout l;
    l.ldqe = (Ge + 1) & ~1;
    l.ldqr = (Gr + 1) & ~1;
    l.need_tmp = grouped && !full;
    // cells per chunk: T = Ge D doubles per cell; the float32 scratch of the group-only mode = Ge Gr floats per cell
    int64_t chunk = 2048;
    if (l.need_tmp) {
        const int64_t by_tmp = (int64_t)(8.0e9 / ((double)Ge * Gr * 4.0));
        chunk = std::max<int64_t>(128, std::min<int64_t>(2048, by_tmp / 128 * 128));
    }
    const int64_t n_up = (n + 127) / 128 * 128;
    l.chunk = std::min(chunk, n_up);
    auto al = [](size_t x) { return (x + 1023) & ~(size_t)1023; };
    size_t o = 0;
    l.qet = o; o += al((size_t)D * l.ldqe * 8);
    l.qrt = o; o += al((size_t)D * l.ldqr * 8);
    l.T = o;   o += al((size_t)Ge * D * l.chunk * 8);
    l.tmp = o; o += l.need_tmp ? al((size_t)Ge * Gr * l.chunk * 4) : 0;
    l.total = o;
    return l;
}
}  // namespace

extern "C" int dvf_gene_gemm_workspace_bytes(int64_t n, int D, int Ge, int Gr, int full, int grouped, size_t* out) {
    DVF_ARG(out && n >= 0 && D >= 1 && D <= kMaxD && Ge >= 1 && Gr >= 1, "bad argument");
    *out = n == 0 ? 0 : gg_make_layout(n, D, Ge, Gr, full != 0, grouped != 0).total;
    return DVF_OK;
}

static int gene_gemm_run(const double* J, int64_t n, int D, const double* Qe, int Ge, const double* Qr, int Gr,
                         int take_abs, int op, const float* jmax, const int32_t* group, int n_groups, float* out_full,
                         double* out_group_sum, void* ws, size_t ws_bytes, void* stream) {
    DVF_ARG(J && Qe && Qr && n >= 0 && D >= 1 && D <= kMaxD && Ge >= 1 && Gr >= 1, "bad argument");
    DVF_ARG(out_full || out_group_sum, "no output requested");
    DVF_ARG(!out_group_sum || op == 2 || (group && n_groups >= 1), "group labels missing");
    DVF_ARG(op != 1 || Ge == Gr, "the switch product needs a square gene matrix (same effectors and regulators)");
    if (op == 2) n_groups = 1;
    if (n == 0) return DVF_OK;
    DVF_ARG(((uintptr_t)J & 15) == 0 && (n % 2) == 0, "the tensor-core gene transform needs a 16-byte aligned J and an even number of cells");
    const gg_layout l = gg_make_layout(n, D, Ge, Gr, out_full != nullptr, out_group_sum != nullptr);
    if (!ws || ws_bytes < l.total) {
        set_error("gene transform workspace too small: %zu < %zu", ws_bytes, l.total);
        return DVF_ERR_WORKSPACE;
    }
    DVF_ARG(((uintptr_t)ws & 1023) == 0, "workspace must be 1024-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    static bool attr_set = false;
    if (!attr_set) {
        DVF_CUDA(cudaFuncSetAttribute(k_tsgemm<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TS_SMEM));
        DVF_CUDA(cudaFuncSetAttribute(k_tsgemm<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TS_SMEM));
        attr_set = true;
    }
    char* base = (char*)ws;
    double* QeT = (double*)(base + l.qet);
    double* QrT = (double*)(base + l.qrt);
    double* T = (double*)(base + l.T);
    float* tmp = (float*)(base + l.tmp);
    DVF_CUDA(cudaMemsetAsync(QeT, 0, (size_t)D * l.ldqe * 8, st));
    DVF_CUDA(cudaMemsetAsync(QrT, 0, (size_t)D * l.ldqr * 8, st));
    k_transpose_q<<<64, 256, 0, st>>>(Qe, Ge, D, QeT, l.ldqe);
    DVF_LAUNCH_CHECK();
    k_transpose_q<<<64, 256, 0, st>>>(Qr, Gr, D, QrT, l.ldqr);
    DVF_LAUNCH_CHECK();
    CUtensorMap mapQe, mapQr, mapJ, mapT;
    if (int rc = make_tmap_2d(&mapQe, QeT, (uint64_t)Ge, (uint64_t)D, (uint64_t)l.ldqe * 8)) return rc;
    if (int rc = make_tmap_2d(&mapQr, QrT, (uint64_t)Gr, (uint64_t)D, (uint64_t)l.ldqr * 8)) return rc;
    // J as (cell, j, i): strides n (j) and D n (i) doubles
    if (int rc = make_tmap_3d(&mapJ, J, (uint64_t)n, (uint64_t)D, (uint64_t)D, (uint64_t)n * 8, (uint64_t)D * n * 8)) return rc;
    if (int rc = make_tmap_2d(&mapT, T, (uint64_t)l.chunk, (uint64_t)Ge * D, (uint64_t)l.chunk * 8)) return rc;
    const int sms = num_sms();
    for (int64_t c0 = 0; c0 < n; c0 += l.chunk) {
        const int64_t nc = std::min<int64_t>(l.chunk, n - c0);
        ts_params p1{};
        p1.K = D; p1.nb = D; p1.nrt = (Ge + 127) / 128; p1.rows = Ge; p1.cells = nc; p1.b_cell0 = c0; p1.b_rows_per_batch = 0;
        p1.out = T; p1.ld = l.chunk; p1.out_cell0 = 0; p1.row_mul = D; p1.b_mul = 1; p1.take_abs = 0;
        int64_t items = (int64_t)p1.nb * p1.nrt * ((nc + 127) / 128);
        k_tsgemm<false, true><<<(int)std::min<int64_t>(items, sms), TS_THREADS, TS_SMEM, st>>>(mapQe, mapJ, p1);
        DVF_LAUNCH_CHECK();
        ts_params p2{};
        p2.K = D; p2.nb = Ge; p2.nrt = (Gr + 127) / 128; p2.rows = Gr; p2.cells = nc; p2.b_cell0 = 0; p2.b_rows_per_batch = D;
        p2.row_mul = 1; p2.b_mul = Gr; p2.take_abs = take_abs;
        if (out_full) { p2.out = out_full; p2.ld = n; p2.out_cell0 = c0; }
        else { p2.out = tmp; p2.ld = l.chunk; p2.out_cell0 = 0; }
        items = (int64_t)p2.nb * p2.nrt * ((nc + 127) / 128);
        k_tsgemm<true, false><<<(int)std::min<int64_t>(items, sms), TS_THREADS, TS_SMEM, st>>>(mapQr, mapT, p2);
        DVF_LAUNCH_CHECK();
        if (out_group_sum) {
            const size_t gs_smem = (size_t)8 * 32 * n_groups * sizeof(double);
            DVF_ARG(gs_smem <= 200 * 1024, "n_groups=%d too large for the group-sum kernel (max 100)", n_groups);
            if (gs_smem > 48 * 1024)
                DVF_CUDA(cudaFuncSetAttribute(k_gene_group_sum, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gs_smem));
            k_gene_group_sum<<<sms * 4, 256, gs_smem, st>>>((const float*)p2.out, p2.ld, p2.out_cell0, nc, (int64_t)Ge * Gr,
                                                      group ? group + c0 : nullptr, n_groups, out_group_sum, op, Ge, jmax);
            DVF_LAUNCH_CHECK();
        }
    }
    return DVF_OK;
}

// tensor-core path of dvf_gene_transform (same arguments + workspace)
extern "C" int dvf_gene_transform_gemm(const double* J, int64_t n, int D, const double* Qe, int Ge, const double* Qr, int Gr,
                                       int take_abs, const int32_t* group, int n_groups, float* out_full,
                                       double* out_group_sum, void* ws, size_t ws_bytes, void* stream) {
    return gene_gemm_run(J, n, D, Qe, Ge, Qr, Gr, take_abs, 0, nullptr, group, n_groups, out_full, out_group_sum, ws, ws_bytes,
                         stream);
}

// reductions over cells of the gene-space Jacobian that rank_jacobian_genes consumes, without materialising it
// (workspace: dvf_gene_gemm_workspace_bytes(n, D, Ge, Gr, full = 0, grouped = 1))
extern "C" int dvf_gene_transform_reduce(const double* J, int64_t n, int D, const double* Qe, int Ge, const double* Qr, int Gr,
                                         int take_abs, int op, const float* jmax, const int32_t* group, int n_groups,
                                         double* out, void* ws, size_t ws_bytes, void* stream) {
    DVF_ARG(op >= 0 && op <= 2 && out, "bad argument");
    return gene_gemm_run(J, n, D, Qe, Ge, Qr, Gr, take_abs, op, jmax, group, n_groups, nullptr, out, ws, ws_bytes, stream);
}

// K3: weighted Gram + right-hand side of the SparseVFC M-step (reference scVectorField.py:468-476)
//        G = [U | Y]^T diag(P) [U | Y]      (lower triangle only; P already clamped to >= minP)
// so that  lhs = G[:M, :M] + lambda sigma2 Kc   and   rhs[m][d] = G[M + d][m].
// The reference materialises the M x N temporary U^T * repmat(P) every iteration; here the N x W matrix
// UY = [U | Y | 0] is built once (K1) and streamed through FP64 tensor cores (DMMA m8n8k4, the contraction
// index is the cell).  The lower triangle is covered by output tiles of a few shapes (128x128 off the diagonal,
// 64x64 pieces on it, 8r x 128 strips for the tail rows) so that little more than the needed triangle is computed.
//
// CTA = 16 consumer warps + 1 producer warp.  Four consumer warps per scheduler: a warp can issue a DMMA only
// every ~17 cycles while the pipe accepts one per 16, so two warps per scheduler cannot also hide their LDS /
// DMUL issue slots (measured 67% pipe utilisation with 8 warps).  Every consumer warp owns a 32 x 32 block
// (4 x 4 DMMA tiles, 0.5 LDS per DMMA); tiles smaller than 128 x 128 split the 32 cells of a stage between
// warps instead of shrinking the warp block (k-split: partials of the k-groups are summed by k_gram_reduce).
// The producer thread streams 32-cell row blocks of the tile's two column panels into a 3-stage shared-memory
// ring with TMA (2-D tensor map over UY, box 16 columns x 32 cells, 128-byte swizzle: with the k-step -> row
// mapping r = 8 (s/2) + 2 t + (s & 1) every fragment load is bank-conflict free), `full` / `empty` mbarriers,
// no __syncthreads in the loop, and the ring keeps running across tile boundaries.  Rows past N and columns
// past W are zero-filled by the TMA unit.
// The CTAs of a group walk the same cells; producers keep within `lag` blocks of the slowest peer (progress
// words in global memory) so that a row block is fetched from HBM once and served from L2 to the others.
// Every CTA writes private partial tiles; k_gram_reduce sums them in a fixed order: bitwise reproducible.
#include <algorithm>
#include <vector>
#include "em.h"
#include "tma.cuh"

namespace dvf {

constexpr int GR_KC = 32;            // cells per pipeline stage
constexpr int GR_STAGES = 3;
constexpr int GR_CWARPS = 16;        // consumer warps
constexpr int GR_THREADS = (GR_CWARPS + 1) * 32;
constexpr int GR_SLOT = 128 * 128;   // doubles per partial-tile slot (all k-groups of a tile)
constexpr int GR_CHUNK_B = 16 * GR_KC * 8;                // bytes of one TMA box: 16 columns x 32 cells
constexpr int GR_STAGE_B = 16 * GR_CHUNK_B + 1024;        // A panel (8 boxes) + B panel (8 boxes) + P (padded)
constexpr size_t GR_SMEM = (size_t)GR_STAGES * GR_STAGE_B + 1024 /* alignment slack */ + 64 /* barriers */;

// tile variants: every warp owns RT x CT DMMA tiles; 16 warps = WI x WJ positions x KS k-groups
enum { GV_FULL = 0, GV_HALF = 1, GV_S8 = 2, GV_S4 = 3, GV_S3 = 4, GV_S2 = 5, GV_S1 = 6, GV_COUNT = 7 };
template <int V> struct gvar;
template <> struct gvar<GV_FULL> { static constexpr int RT = 4, CT = 4, WI = 4, WJ = 4, KS = 1; };   // 128 x 128
template <> struct gvar<GV_HALF> { static constexpr int RT = 4, CT = 4, WI = 2, WJ = 2, KS = 4; };   //  64 x  64
template <> struct gvar<GV_S8>   { static constexpr int RT = 4, CT = 4, WI = 2, WJ = 4, KS = 2; };   //  64 x 128
template <> struct gvar<GV_S4>   { static constexpr int RT = 4, CT = 4, WI = 1, WJ = 4, KS = 4; };   //  32 x 128
template <> struct gvar<GV_S3>   { static constexpr int RT = 3, CT = 4, WI = 1, WJ = 4, KS = 4; };   //  24 x 128
template <> struct gvar<GV_S2>   { static constexpr int RT = 2, CT = 4, WI = 1, WJ = 4, KS = 4; };   //  16 x 128
template <> struct gvar<GV_S1>   { static constexpr int RT = 1, CT = 4, WI = 1, WJ = 4, KS = 4; };   //   8 x 128
__host__ __device__ inline int gv_rows(int v) {
    return v == GV_FULL ? 128 : (v == GV_HALF || v == GV_S8) ? 64 : v == GV_S4 ? 32 : v == GV_S3 ? 24 : v == GV_S2 ? 16 : 8;
}
__host__ __device__ inline int gv_cols(int v) { return v == GV_HALF ? 64 : 128; }
__host__ __device__ inline int gv_ks(int v) { return v == GV_FULL ? 1 : v == GV_S8 ? 2 : 4; }

// relative cost of one 32-cell step (DMMA count x a penalty for the smaller shapes, calibrated on B200 with
// DVF_GRAM_ONLY, tools/gram_tune.py)
static double kVarCost[GV_COUNT] = {256.0, 64.0 * 1.1, 128.0 * 1.05, 64.0 * 1.1, 48.0 * 1.15, 32.0 * 1.25, 16.0 * 1.6};
// parts schedule: plain DMMA counts down to the 24-row strips -- without accumulator swaps those shapes cost what they
// compute (swept on B200 with tools/gram_cost_sweep.sh: 9.80-9.83 ms on a plateau around these values, 10.4 ms with
// the bundle table); the two thinnest strips keep their penalty (M = 100: 1.36 ms with it, 1.69 ms without)
static double kPartCost[GV_COUNT] = {256.0, 64.0, 128.0, 64.0, 48.0, 32.0 * 1.25, 16.0 * 1.6};
static double g_diag_factor = 0.75;
// g_parts: -1 planner decides, 0 bundles only, 1 parts whenever possible.  g_sb_parts: block length in stages; 0 = the
// planner picks a length near 512 that the part counts divide (see build_host_plan)
static int g_parts = -1, g_sb_parts = 0;
static int g_force_G = 0, g_sb = 64, g_lag = 0, g_only = -1, g_debug = 0;   // sb, lag: measured on B200
static void read_tuning_env() {   // developer knobs (tools/gram_tune.py); not part of the ABI
    static bool done = false;
    if (done) return;
    done = true;
    if (const char* e = getenv("DVF_GRAM_COST"))
        sscanf(e, "%lf,%lf,%lf,%lf,%lf,%lf,%lf", &kVarCost[0], &kVarCost[1], &kVarCost[2], &kVarCost[3], &kVarCost[4],
               &kVarCost[5], &kVarCost[6]);
    if (const char* e = getenv("DVF_GRAM_PCOST"))
        sscanf(e, "%lf,%lf,%lf,%lf,%lf,%lf,%lf", &kPartCost[0], &kPartCost[1], &kPartCost[2], &kPartCost[3], &kPartCost[4],
               &kPartCost[5], &kPartCost[6]);
    if (const char* e = getenv("DVF_GRAM_G")) g_force_G = atoi(e);
    if (const char* e = getenv("DVF_GRAM_DIAG")) g_diag_factor = atof(e);
    if (const char* e = getenv("DVF_GRAM_PARTS")) g_parts = atoi(e);
    if (const char* e = getenv("DVF_GRAM_SBP")) g_sb_parts = std::max(0, atoi(e));
    if (const char* e = getenv("DVF_GRAM_SB")) g_sb = std::max(1, atoi(e));
    if (const char* e = getenv("DVF_GRAM_LAG")) g_lag = atoi(e);       // blocks a producer may run ahead of its group
    if (const char* e = getenv("DVF_GRAM_ONLY")) g_only = atoi(e);     // timing experiments: keep one tile variant
    if (const char* e = getenv("DVF_GRAM_DEBUG")) g_debug = atoi(e);
}

struct gram_tile { int variant, row0, col0, diag, pad1; };   // diag: 64 x 64 tile on the diagonal (upper-right 32 x 32 unused)
struct gr_ring { int stage; unsigned phase; };
__device__ __forceinline__ void gr_advance(gr_ring& r) { if (++r.stage == GR_STAGES) { r.stage = 0; r.phase ^= 1u; } }

// ---- consumer: one tile over `n_steps` ring stages ------------------------------------------------------
// Swizzled box layout (CU_TENSOR_MAP_SWIZZLE_128B): element (cell r, column cc) of a 16-column box sits at byte
//   r * 128 + ((cc / 2) ^ (r & 7)) * 16 + (cc & 1) * 8.
// k-step s of a stage uses the cells r(s, t) = 8 (s / 2) + 2 t + (s & 1), t = 0..3: the 32 lanes of a fragment load
// (8 columns x 4 cells) then touch every bank exactly twice (2 wavefronts, the minimum for 256 bytes).
template <int V>
__device__ __forceinline__ void gram_consume(int n_steps, double* __restrict__ out, bool accumulate, bool diag,
                                             const unsigned char* ring, uint64_t* full, uint64_t* empty, gr_ring& rs) {
    using gv = gvar<V>;
    constexpr int RT = gv::RT, CT = gv::CT, WJ = gv::WJ, KS = gv::KS;
    constexpr int TR = 8 * RT * gv::WI, TC = 8 * CT * WJ;
    constexpr int SPW = (GR_KC / 4) / KS;                 // k-steps per warp and stage (8, 4 or 2)
    constexpr int POS = gv::WI * WJ;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    // 64 x 64 tiles rotate the positions through the k-groups, so that the four warps of one position sit on four
    // different schedulers (warp % 4): a diagonal tile then idles one warp per scheduler, not one whole scheduler
    const int kg = warp / POS, pos = (V == GV_HALF) ? (warp + kg) % POS : warp % POS, wi = pos / WJ, wj = pos % WJ;
    out += kg * (TR * TC);
    if (V == GV_HALF && diag && wi < wj) {
        // the 32 x 32 block above the diagonal is never read: only keep the ring protocol going (a warp may arrive
        // on `empty` for a stage once per fill, so it still waits for every fill)
        for (int step = 0; step < n_steps; ++step) {
            mbar_wait(&full[rs.stage], rs.phase);
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[rs.stage]);
            gr_advance(rs);
        }
        return;
    }

    double acc[RT][CT][2];
#pragma unroll
    for (int i = 0; i < RT; ++i)
#pragma unroll
        for (int j = 0; j < CT; ++j) {
            acc[i][j][0] = acc[i][j][1] = 0.0;
            if (accumulate) {   // continue this CTA's private partial of the tile (same thread wrote it)
                const int r = wi * (8 * RT) + i * 8 + g, c = wj * (8 * CT) + j * 8 + 2 * t;
                const double2 v = *reinterpret_cast<const double2*>(out + r * TC + c);
                acc[i][j][0] = v.x;
                acc[i][j][1] = v.y;
            }
        }

    // per-thread byte offsets inside a box for (s & 1, 8-column half h); RT is even or wi == 0, CT == 4
    unsigned toff[2][2];
#pragma unroll
    for (int sp = 0; sp < 2; ++sp)
#pragma unroll
        for (int h = 0; h < 2; ++h)
            toff[sp][h] = (2 * t + sp) * 128 + ((((h << 2) | (g >> 1)) ^ (2 * t + sp)) << 4) + (g & 1) * 8;
    // warp bases: A panel box (wi * RT / 2), B panel box (wj * 2), first cell 8 * kg * SPW / 2
    const unsigned wbase_a = (wi * RT / 2) * GR_CHUNK_B + kg * (SPW / 2) * (8 * 128);
    const unsigned wbase_b = 8 * GR_CHUNK_B + (wj * 2) * GR_CHUNK_B + kg * (SPW / 2) * (8 * 128);
    const unsigned wbase_p = 16 * GR_CHUNK_B + (kg * (SPW / 2) * 8 + 2 * t) * 8;

    for (int step = 0; step < n_steps; ++step) {
        mbar_wait(&full[rs.stage], rs.phase);
#ifdef DVF_DBG_CONS_SLEEP
        __nanosleep(DVF_DBG_CONS_SLEEP);
#endif
        const unsigned char* sb = ring + rs.stage * GR_STAGE_B;
#pragma unroll
        for (int j = 0; j < SPW; ++j) {
            const int sp = j & 1;
            const unsigned so = (j >> 1) * (8 * 128);
            const double p = *reinterpret_cast<const double*>(sb + wbase_p + (j >> 1) * 64 + sp * 8);
            double af[RT], bf[CT];
#pragma unroll
            for (int i = 0; i < RT; ++i) af[i] = *reinterpret_cast<const double*>(sb + wbase_a + (i >> 1) * GR_CHUNK_B + so + toff[sp][i & 1]);
#pragma unroll
            for (int q = 0; q < CT; ++q) bf[q] = *reinterpret_cast<const double*>(sb + wbase_b + (q >> 1) * GR_CHUNK_B + so + toff[sp][q & 1]);
#pragma unroll
            for (int i = 0; i < RT; ++i) af[i] *= p;
#pragma unroll
            for (int i = 0; i < RT; ++i)
#pragma unroll
                for (int q = 0; q < CT; ++q) dmma(acc[i][q][0], acc[i][q][1], af[i], bf[q]);
        }
        // generic-proxy reads of this stage must be ordered before the producer's next TMA (async-proxy) write to
        // it: without the proxy fence the B200 shows stale / torn stages a few times per thousand steps
        __syncwarp();
        fence_proxy_async_smem();
        if (lane == 0) mbar_arrive(&empty[rs.stage]);
        gr_advance(rs);
    }
#pragma unroll
    for (int i = 0; i < RT; ++i)
#pragma unroll
        for (int j = 0; j < CT; ++j) {
            const int r = wi * (8 * RT) + i * 8 + g, c = wj * (8 * CT) + j * 8 + 2 * t;
            *reinterpret_cast<double2*>(out + r * TC + c) = make_double2(acc[i][j][0], acc[i][j][1]);
        }
}

// ---- producer (one thread): cells [step*32, step*32+32) of the tile's A (row0..) and B (col0..) panels ------
__device__ __forceinline__ void gram_produce(const CUtensorMap* mapUY, const CUtensorMap* mapP, const gram_tile& tl, int lo,
                                             int hi, unsigned char* ring, uint64_t* full, uint64_t* empty, gr_ring& rs) {
    const int na = (gv_rows(tl.variant) + 15) / 16, nb = gv_cols(tl.variant) / 16;
    const unsigned bytes = (na + nb) * GR_CHUNK_B + GR_KC * 8;
    for (int step = lo; step < hi; ++step) {
        mbar_wait(&empty[rs.stage], rs.phase ^ 1u);
#ifdef DVF_DBG_PROD_SLEEP
        __nanosleep(DVF_DBG_PROD_SLEEP);
#endif
        const unsigned sb = smem_u32(ring) + rs.stage * GR_STAGE_B, bar = smem_u32(&full[rs.stage]);
        const int n0 = step * GR_KC;
        mbar_arrive_expect_tx(&full[rs.stage], bytes);
        for (int c = 0; c < na; ++c) tma_load_2d(sb + c * GR_CHUNK_B, mapUY, tl.row0 + 16 * c, n0, bar);
        for (int c = 0; c < nb; ++c) tma_load_2d(sb + (8 + c) * GR_CHUNK_B, mapUY, tl.col0 + 16 * c, n0, bar);
        tma_load_2d(sb + 16 * GR_CHUNK_B, mapP, n0, 0, bar);
        gr_advance(rs);
    }
}

// One CTA = one bundle of tiles (equal total cost across bundles) x one group's cell range, walked in blocks of
// `sb` steps; a multi-tile bundle round-robins over its tiles inside a block, carrying each tile's partial in its
// private scratch slot.
__global__ void __launch_bounds__(GR_THREADS, 1) k_gram(const __grid_constant__ CUtensorMap mapUY,
                                                        const __grid_constant__ CUtensorMap mapP, int64_t N,
                                                        const gram_tile* __restrict__ tiles, int n_tiles,
                                                        const int* __restrict__ bundle_off, const int* __restrict__ bundle_tiles,
                                                        int n_bundles, int n_groups, int sb, int lag, double* __restrict__ slots,
                                                        int* __restrict__ progress, const int* __restrict__ done) {
    if (done && *done) return;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    // swizzled TMA boxes need 1024-byte alignment
    unsigned char* ring = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint64_t* full = reinterpret_cast<uint64_t*>(ring + GR_STAGES * GR_STAGE_B);
    uint64_t* empty = full + GR_STAGES;
    if (threadIdx.x == 0) {
        for (int s = 0; s < GR_STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], GR_CWARPS); }
        mbar_fence_init();
    }
    __syncthreads();

    const int grp = blockIdx.x / n_bundles, bun = blockIdx.x % n_bundles;
    const int64_t n_steps = (N + GR_KC - 1) / GR_KC;
    const int s_lo = (int)(n_steps * grp / n_groups), s_hi = (int)(n_steps * (grp + 1) / n_groups);
    const int t_lo = bundle_off[bun], t_hi = bundle_off[bun + 1];
    const int n_blocks = max(1, (s_hi - s_lo + sb - 1) / sb);
    const bool single = (t_hi - t_lo == 1);
    gr_ring rs{0, 0u};
    const int warp = threadIdx.x >> 5;

    if (warp == GR_CWARPS) {
        // ---------------- producer: one thread ----------------
        if ((threadIdx.x & 31) != 0) return;
        int* mine = progress + blockIdx.x;
        const int* peers = progress + grp * n_bundles;
        for (int b = 0; b < n_blocks; ++b) {
            if (n_bundles > 1) {   // stay within `lag` blocks of the slowest CTA of the group (L2 sharing)
                st_volatile_i32(mine, b + 1);
                for (int spin = 0; spin < (1 << 16); ++spin) {
                    int m = 0x7fffffff;
                    for (int q = 0; q < n_bundles; ++q) m = min(m, ld_volatile_i32(peers + q));
                    if (m >= b + 1 - lag) break;
                }
            }
            const int b_lo = s_lo + b * sb, b_hi = min(b_lo + sb, s_hi);
            for (int q = t_lo; q < t_hi; ++q) {
                const gram_tile tl = tiles[bundle_tiles[q]];
                gram_produce(&mapUY, &mapP, tl, b_lo, b_hi, ring, full, empty, rs);
            }
        }
        if (n_bundles > 1) st_volatile_i32(mine, 0x7ffffff0);   // finished: never hold anyone back
        return;
    }

    // ---------------- consumers ----------------
    auto run = [&](int ti, int n, bool acc) {
        const gram_tile tl = tiles[ti];
        double* out = slots + ((size_t)grp * n_tiles + ti) * GR_SLOT;
        switch (tl.variant) {
            case GV_FULL: gram_consume<GV_FULL>(n, out, acc, tl.diag != 0, ring, full, empty, rs); break;
            case GV_HALF: gram_consume<GV_HALF>(n, out, acc, tl.diag != 0, ring, full, empty, rs); break;
            case GV_S8:   gram_consume<GV_S8>(n, out, acc, tl.diag != 0, ring, full, empty, rs); break;
            case GV_S4:   gram_consume<GV_S4>(n, out, acc, tl.diag != 0, ring, full, empty, rs); break;
            case GV_S3:   gram_consume<GV_S3>(n, out, acc, tl.diag != 0, ring, full, empty, rs); break;
            case GV_S2:   gram_consume<GV_S2>(n, out, acc, tl.diag != 0, ring, full, empty, rs); break;
            default:      gram_consume<GV_S1>(n, out, acc, tl.diag != 0, ring, full, empty, rs); break;
        }
    };
    if (single) {
        run(bundle_tiles[t_lo], max(0, s_hi - s_lo), false);
    } else {
        for (int b = 0; b < n_blocks; ++b) {
            const int b_lo = s_lo + b * sb, b_hi = min(b_lo + sb, s_hi);
            for (int q = t_lo; q < t_hi; ++q) run(bundle_tiles[q], max(0, b_hi - b_lo), b > 0);
        }
    }
}

// G[row0 + r][col0 + c] = sum over groups and k-groups of the tile's partials (fixed order); one CTA column per
// tile.  Also rearms the progress words for the next launch.
__global__ void __launch_bounds__(256) k_gram_reduce(const gram_tile* __restrict__ tiles, int n_tiles, int n_groups,
                                                     const double* __restrict__ slots, double* __restrict__ G, int ldg,
                                                     int* __restrict__ progress, int n_ctas, const int* __restrict__ done) {
    if (done && *done) return;
    if (blockIdx.x == 0 && blockIdx.y == 0)
        for (int i = threadIdx.x; i < n_ctas; i += 256) progress[i] = 0;
    const gram_tile tl = tiles[blockIdx.x];
    const int TR = gv_rows(tl.variant), TC = gv_cols(tl.variant), KS = gv_ks(tl.variant);
    const double* src = slots + (size_t)blockIdx.x * GR_SLOT;
    for (int e = blockIdx.y * 256 + threadIdx.x; e < TR * TC; e += gridDim.y * 256) {
        if (tl.diag && e / TC < 32 && e % TC >= 32) continue;        // not computed (above the diagonal)
        double sum = 0.0;
        for (int gi = 0; gi < n_groups; ++gi)
            for (int k = 0; k < KS; ++k) sum += src[(size_t)gi * n_tiles * GR_SLOT + k * TR * TC + e];
        const int r = tl.row0 + e / TC, c = tl.col0 + e % TC;
        if (r < ldg && c < ldg) G[(int64_t)r * ldg + c] = sum;
    }
}

// ---- "parts" schedule: one CTA = one part of one tile for the whole launch ---------------------------------------
// Tile t is cut into p_t parts (p_t ~ cost_t, sum p_t <= #SMs): inside every block of `sb` steps, part k takes the
// steps [len k / p_t, len (k+1) / p_t).  All CTAs sweep the blocks together (a counter per block, `lag` blocks of
// slack), so a block of rows is fetched from HBM once and served from L2 to every tile that needs it, and no CTA ever
// swaps accumulators: its partial tile lives in registers from the first step to the last.
struct gram_part { int tile, k, p, pad; };

__device__ __forceinline__ void gp_range(int b, int sb, int n_steps, int k, int p, int& lo, int& hi) {
    const int b_lo = b * sb, len = min(sb, n_steps - b_lo);
    lo = b_lo + (int)((int64_t)len * k / p);
    hi = b_lo + (int)((int64_t)len * (k + 1) / p);
}

__global__ void __launch_bounds__(GR_THREADS, 1) k_gram_parts(const __grid_constant__ CUtensorMap mapUY,
                                                              const __grid_constant__ CUtensorMap mapP, int64_t N,
                                                              const gram_tile* __restrict__ tiles,
                                                              const gram_part* __restrict__ parts, int sb, int lag,
                                                              double* __restrict__ slots, int* __restrict__ counters,
                                                              const int* __restrict__ done) {
    if (done && *done) return;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    unsigned char* ring = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint64_t* full = reinterpret_cast<uint64_t*>(ring + GR_STAGES * GR_STAGE_B);
    uint64_t* empty = full + GR_STAGES;
    if (threadIdx.x == 0) {
        for (int q = 0; q < GR_STAGES; ++q) { mbar_init(&full[q], 1); mbar_init(&empty[q], GR_CWARPS); }
        mbar_fence_init();
    }
    __syncthreads();
    const gram_part pt = parts[blockIdx.x];
    const gram_tile tl = tiles[pt.tile];
    const int n_steps = (int)((N + GR_KC - 1) / GR_KC), n_blocks = (n_steps + sb - 1) / sb;
    gr_ring rs{0, 0u};
    const int warp = threadIdx.x >> 5;

    if (warp == GR_CWARPS) {
        if ((threadIdx.x & 31) != 0) return;
        const int n_ctas = (int)gridDim.x;
        for (int b = 0; b < n_blocks; ++b) {
            atomicAdd(counters + b, 1);                       // this CTA starts block b
            if (b - lag >= 0)
                for (int spin = 0; spin < (1 << 20) && ld_volatile_i32(counters + b - lag) < n_ctas; ++spin) {}
            int lo, hi;
            gp_range(b, sb, n_steps, pt.k, pt.p, lo, hi);
            gram_produce(&mapUY, &mapP, tl, lo, hi, ring, full, empty, rs);
        }
        return;
    }
    int total = 0;
    for (int b = 0; b < n_blocks; ++b) {
        int lo, hi;
        gp_range(b, sb, n_steps, pt.k, pt.p, lo, hi);
        total += hi - lo;
    }
    double* out = slots + (size_t)blockIdx.x * GR_SLOT;
    switch (tl.variant) {
        case GV_FULL: gram_consume<GV_FULL>(total, out, false, tl.diag != 0, ring, full, empty, rs); break;
        case GV_HALF: gram_consume<GV_HALF>(total, out, false, tl.diag != 0, ring, full, empty, rs); break;
        case GV_S8:   gram_consume<GV_S8>(total, out, false, tl.diag != 0, ring, full, empty, rs); break;
        case GV_S4:   gram_consume<GV_S4>(total, out, false, tl.diag != 0, ring, full, empty, rs); break;
        case GV_S3:   gram_consume<GV_S3>(total, out, false, tl.diag != 0, ring, full, empty, rs); break;
        case GV_S2:   gram_consume<GV_S2>(total, out, false, tl.diag != 0, ring, full, empty, rs); break;
        default:      gram_consume<GV_S1>(total, out, false, tl.diag != 0, ring, full, empty, rs); break;
    }
}

// G tile = sum over the tile's parts (contiguous CTAs part_off[t] .. part_off[t+1]) and k-groups, fixed order
__global__ void __launch_bounds__(256) k_gram_reduce_parts(const gram_tile* __restrict__ tiles, const int* __restrict__ part_off,
                                                           const double* __restrict__ slots, double* __restrict__ G, int ldg,
                                                           int* __restrict__ counters, int n_counters,
                                                           const int* __restrict__ done) {
    if (done && *done) return;
    if (blockIdx.x == 0 && blockIdx.y == 0)
        for (int i = threadIdx.x; i < n_counters; i += 256) counters[i] = 0;
    const gram_tile tl = tiles[blockIdx.x];
    const int TR = gv_rows(tl.variant), TC = gv_cols(tl.variant), KS = gv_ks(tl.variant);
    const int p_lo = part_off[blockIdx.x], p_hi = part_off[blockIdx.x + 1];
    for (int e = blockIdx.y * 256 + threadIdx.x; e < TR * TC; e += gridDim.y * 256) {
        if (tl.diag && e / TC < 32 && e % TC >= 32) continue;
        double sum = 0.0;
        for (int q = p_lo; q < p_hi; ++q)
            for (int k = 0; k < KS; ++k) sum += slots[(size_t)q * GR_SLOT + k * TR * TC + e];
        const int r = tl.row0 + e / TC, c = tl.col0 + e % TC;
        if (r < ldg && c < ldg) G[(int64_t)r * ldg + c] = sum;
    }
}

// P clamped to >= minP (step-level entry point; inside dvf_fit_em the E-step stores P already clamped)
__global__ void k_clamp_copy(const double* __restrict__ P, int64_t N, double minP, double* __restrict__ out) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += (int64_t)gridDim.x * blockDim.x)
        out[i] = fmax(P[i], minP);
}

// ---- launch plan ---------------------------------------------------------------------------------------
struct gram_host_plan {
    std::vector<gram_tile> tiles;
    std::vector<int> bundle_off, bundle_tiles;
    int n_groups = 1;
    // parts schedule (used when it is predicted to be faster)
    bool use_parts = false;
    int sb_parts = 512;
    std::vector<gram_part> parts;
    std::vector<int> part_off;
};

// a diagonal 64 x 64 tile runs 12 of its 16 warps
static double tile_cost(const gram_tile& tl) { return kVarCost[tl.variant] * (tl.diag ? g_diag_factor : 1.0); }
static double part_cost(const gram_tile& tl) { return kPartCost[tl.variant] * (tl.diag ? g_diag_factor : 1.0); }

static void build_host_plan(int64_t N, int D, int M, int n_sms, gram_host_plan& hp) {
    read_tuning_env();
    const int units = (M + D + 7) / 8, nfull = units / 16, rt = units % 16;
    // A tile whose rows AND columns all lie at or beyond M is part of the Y^T diag(P) Y block, which nothing reads (lhs is
    // G[:M,:M], rhs is G[M:M+D,:M]): it is not computed.  At M = 500, D = 30 that is the last piece of the tail strip,
    // 3072 of 154.6 k FMA per cell.
    auto add = [&](int v, int r0, int c0) {
        if (r0 >= M && c0 >= M) return;
        hp.tiles.push_back(gram_tile{v, r0, c0, (v == GV_HALF && r0 == c0) ? 1 : 0, 0});
    };
    for (int s = 0; s < nfull; ++s) {
        for (int jb = 0; jb < s; ++jb) add(GV_FULL, 128 * s, 128 * jb);
        add(GV_HALF, 128 * s, 128 * s);
        add(GV_HALF, 128 * s + 64, 128 * s);
        add(GV_HALF, 128 * s + 64, 128 * s + 64);
    }
    int row = 128 * nfull, rem = rt;
    const int chunk_units[5] = {8, 4, 3, 2, 1}, chunk_var[5] = {GV_S8, GV_S4, GV_S3, GV_S2, GV_S1};
    for (int q = 0; q < 5; ++q)
        while (rem >= chunk_units[q]) {
            for (int jb = 0; jb <= nfull; ++jb) add(chunk_var[q], row, 128 * jb);
            row += 8 * chunk_units[q];
            rem -= chunk_units[q];
        }
    if (g_only >= 0) {
        std::vector<gram_tile> keep;
        for (auto& tl : hp.tiles) if (tl.variant == g_only) keep.push_back(tl);
        if (!keep.empty()) hp.tiles = keep;
    }
    const int nt = (int)hp.tiles.size();
    const int64_t n_steps = (N + GR_KC - 1) / GR_KC;
    std::vector<int> by_cost(nt);
    for (int i = 0; i < nt; ++i) by_cost[i] = i;
    std::stable_sort(by_cost.begin(), by_cost.end(), [&](int a, int b) {
        return tile_cost(hp.tiles[a]) > tile_cost(hp.tiles[b]);
    });
    double total = 0.0;
    for (auto& tl : hp.tiles) total += tile_cost(tl);
    // choose the number of bundles G (LPT bin packing) that maximises
    //   total / (G * max_bin)  x  (G * groups) / n_sms ,  groups = min(n_sms / G, n_steps)
    double best_eff = -1.0;
    int best_g = 1;
    std::vector<int> best_assign;
    // bundles no heavier than ~1.25 full tiles: a CTA that owns many tiles re-reads its cells once per tile from
    // HBM and swaps accumulators at every block, and the group-wise L2 sharing needs several CTAs per group
    int g_min = (int)(total / (1.25 * kVarCost[GV_FULL]) + 0.5);
    g_min = std::max(1, std::min(g_min, std::min(nt, n_sms)));
    for (int G = g_min; G <= nt && G <= n_sms; ++G) {
        std::vector<double> bin(G, 0.0);
        std::vector<int> assign(nt, 0);
        for (int id : by_cost) {
            int k = (int)(std::min_element(bin.begin(), bin.end()) - bin.begin());
            bin[k] += tile_cost(hp.tiles[id]);
            assign[id] = k;
        }
        const double mx = *std::max_element(bin.begin(), bin.end());
        int64_t groups = n_sms / G;
        if (groups > n_steps) groups = n_steps;
        if (groups < 1) groups = 1;
        const double eff = total / (G * mx) * (double)(G * groups) / n_sms;
        if (g_force_G ? (G == g_force_G) : (eff > best_eff + 1e-12)) { best_eff = eff; best_g = G; best_assign = assign; }
    }
    hp.n_groups = (int)std::max<int64_t>(1, std::min<int64_t>(n_sms / best_g, n_steps));
    hp.bundle_off.assign(best_g + 1, 0);
    for (int k = 0; k < best_g; ++k) {
        hp.bundle_off[k] = (int)hp.bundle_tiles.size();
        for (int i = 0; i < nt; ++i)
            if (best_assign[i] == k) hp.bundle_tiles.push_back(i);
    }
    hp.bundle_off[best_g] = (int)hp.bundle_tiles.size();
    // ---- parts schedule: p_t parts per tile, greedily lowering max(cost_t / p_t) until the SMs are used up ----
    double parts_eff = 0.0;
    if (nt <= n_sms && n_steps >= 4 * (int64_t)n_sms && g_parts != 0) {
        std::vector<int> p(nt, 1);
        double ptotal = 0.0;
        for (int i = 0; i < nt; ++i) ptotal += part_cost(hp.tiles[i]);
        for (int used = nt; used < n_sms; ++used) {
            int arg = 0;
            double worst = -1.0;
            for (int i = 0; i < nt; ++i) {
                const double c = part_cost(hp.tiles[i]) / p[i];
                if (c > worst) { worst = c; arg = i; }
            }
            p[arg]++;
        }
        double worst = 0.0;
        int used = 0;
        for (int i = 0; i < nt; ++i) { worst = std::max(worst, part_cost(hp.tiles[i]) / p[i]); used += p[i]; }
        parts_eff = ptotal / (n_sms * worst);
        // measured: parts wins clearly where the bundles are unbalanced (M = 500: 0.957 vs 0.881 predicted, 9.8 vs
        // 11.0 ms) and loses by ~3 % where both balance well (M = 1000: 0.953 vs 0.972, 37.0 vs 35.9 ms)
        // (the bundle prediction ignores the accumulator swaps of multi-tile CTAs: with few tiles parts always won,
        // 100k x (100+30): 0.35 vs 1.28 ms, 500k x (300+30): 2.43 vs 2.63 ms)
        if (g_parts > 0 || nt <= 32 || parts_eff > best_eff + 0.01) {
            hp.use_parts = true;
            hp.part_off.assign(nt + 1, 0);
            for (int i = 0; i < nt; ++i) {
                hp.part_off[i] = (int)hp.parts.size();
                for (int k = 0; k < p[i]; ++k) hp.parts.push_back(gram_part{i, k, p[i], 0});
            }
            hp.part_off[nt] = (int)hp.parts.size();
            // Block length.  Part k of a tile takes the stages [len k / p, len (k+1) / p) of every block, and all CTAs
            // meet at every block start, so a block costs its LARGEST part: with 512 stages and p = 15 that is 35
            // stages where 34.1 are due (2.5 %, paid 61 times; measured 9.81 ms against 9.54 ms with 2048-stage
            // blocks, whose 300 MB no longer fit the L2: 12.3 GB of DRAM traffic per launch instead of 5.0 GB).
            // Pick the length in [448, 704] that the part counts divide best: same speed, blocks stay L2-sized.
            hp.sb_parts = 512;
            if (g_sb_parts > 0) hp.sb_parts = g_sb_parts;
            else {
                double best_loss = 1e30;
                for (int sb = 448; sb <= 704; sb += 4) {
                    double loss = 0.0;
                    for (int i = 0; i < nt; ++i) loss = std::max(loss, (double)((sb + p[i] - 1) / p[i]) * p[i] / sb);
                    if (loss < best_loss - 1e-9) { best_loss = loss; hp.sb_parts = sb; }
                }
            }
            (void)used;
        }
    }
    if (g_debug) {
        fprintf(stderr, "[gram plan] parts schedule: eff=%.3f %s, block of %d stages\n", parts_eff,
                hp.use_parts ? "(used)" : "(not used)", hp.sb_parts);
        int cnt[GV_COUNT] = {0};
        for (auto& tl : hp.tiles) cnt[tl.variant]++;
        fprintf(stderr, "[gram plan] N=%lld M=%d D=%d tiles=%d (full %d half %d s8 %d s4 %d s3 %d s2 %d s1 %d) bundles=%d groups=%d eff=%.3f\n",
                (long long)N, M, D, nt, cnt[0], cnt[1], cnt[2], cnt[3], cnt[4], cnt[5], cnt[6], best_g, hp.n_groups, best_eff);
        g_debug = 0;
    }
}

static inline size_t al256(size_t x) { return (x + 255) & ~(size_t)255; }

gram_plan make_gram_plan(int64_t N, int D, int M) {
    gram_plan pl;
    gram_host_plan hp;
    build_host_plan(N, D, M, num_sms(), hp);
    pl.ldg = (int)dvf_fit_row_width(M, D);
    pl.n_tiles = (int)hp.tiles.size();
    pl.n_bundles = (int)hp.bundle_off.size() - 1;
    pl.n_groups = hp.n_groups;
    pl.n_ctas = pl.n_bundles * pl.n_groups;
    pl.use_parts = hp.use_parts ? 1 : 0;
    pl.n_parts = (int)hp.parts.size();
    pl.sb_parts = hp.sb_parts;
    pl.n_counters = (int)(((N + GR_KC - 1) / GR_KC + hp.sb_parts - 1) / hp.sb_parts) + 1;
    size_t o = 0;
    pl.off_tiles = o;  o += al256(sizeof(gram_tile) * hp.tiles.size());
    pl.off_boff = o;   o += al256(sizeof(int) * hp.bundle_off.size());
    pl.off_btiles = o; o += al256(sizeof(int) * hp.bundle_tiles.size());
    pl.off_progress = o; o += al256(sizeof(int) * (size_t)std::max(pl.n_ctas, pl.n_counters));
    pl.off_parts = o;  o += al256(sizeof(gram_part) * hp.parts.size());
    pl.off_poff = o;   o += al256(sizeof(int) * hp.part_off.size());
    pl.off_pclamp = o; o += al256(sizeof(double) * (size_t)(N + 2));
    pl.off_slots = o;
    o += sizeof(double) * (size_t)GR_SLOT * (pl.use_parts ? (size_t)pl.n_parts : (size_t)pl.n_tiles * pl.n_groups);
    pl.scratch_bytes = o;
    pl.sb = g_sb;
    pl.lag = g_lag;
    return pl;
}

// upload the schedule tables into the head of the scratch buffer (once per fit)
int gram_prepare(const gram_plan& pl, int64_t N, int D, int M, void* scratch, cudaStream_t st) {
    gram_host_plan hp;
    build_host_plan(N, D, M, num_sms(), hp);
    char* base = (char*)scratch;
    // pageable sources: cudaMemcpyAsync stages them before returning, so the vectors may die right after
    DVF_CUDA(cudaMemcpyAsync(base + pl.off_tiles, hp.tiles.data(), sizeof(gram_tile) * hp.tiles.size(), cudaMemcpyHostToDevice, st));
    DVF_CUDA(cudaMemcpyAsync(base + pl.off_boff, hp.bundle_off.data(), sizeof(int) * hp.bundle_off.size(), cudaMemcpyHostToDevice, st));
    DVF_CUDA(cudaMemcpyAsync(base + pl.off_btiles, hp.bundle_tiles.data(), sizeof(int) * hp.bundle_tiles.size(), cudaMemcpyHostToDevice, st));
    DVF_CUDA(cudaMemsetAsync(base + pl.off_progress, 0, sizeof(int) * (size_t)std::max(pl.n_ctas, pl.n_counters), st));
    if (pl.use_parts) {
        if ((int)hp.parts.size() != pl.n_parts) { set_error("gram plan changed between planning and preparation"); return DVF_ERR_ARG; }
        DVF_CUDA(cudaMemcpyAsync(base + pl.off_parts, hp.parts.data(), sizeof(gram_part) * hp.parts.size(), cudaMemcpyHostToDevice, st));
        DVF_CUDA(cudaMemcpyAsync(base + pl.off_poff, hp.part_off.data(), sizeof(int) * hp.part_off.size(), cudaMemcpyHostToDevice, st));
    }
    return DVF_OK;
}

// P must already be clamped to >= minP (the E-step stores it that way; dvf_gram clamps a copy).
// UY and P need 16-byte alignment (TMA).
int gram_run(const gram_plan& pl, const double* UY, const double* P, int64_t N, int W, double* G, double* scratch,
             const int* done, cudaStream_t st) {
    static bool attr_set = false;
    if (!attr_set) {
        DVF_CUDA(cudaFuncSetAttribute(k_gram, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GR_SMEM));
        DVF_CUDA(cudaFuncSetAttribute(k_gram_parts, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GR_SMEM));
        attr_set = true;
    }
    DVF_ARG(((uintptr_t)UY & 15) == 0 && ((uintptr_t)P & 15) == 0, "UY and P must be 16-byte aligned");
    CUtensorMap mapUY, mapP;
    if (int rc = make_tmap_2d(&mapUY, UY, (uint64_t)W, (uint64_t)N, (uint64_t)W * 8)) return rc;
    if (int rc = make_tmap_vec(&mapP, P, (uint64_t)N)) return rc;
    char* base = (char*)scratch;
    const gram_tile* tiles = (const gram_tile*)(base + pl.off_tiles);
    double* slots = (double*)(base + pl.off_slots);
    int* progress = (int*)(base + pl.off_progress);
    if (pl.use_parts) {
        k_gram_parts<<<pl.n_parts, GR_THREADS, GR_SMEM, st>>>(mapUY, mapP, N, tiles, (const gram_part*)(base + pl.off_parts),
                                                            pl.sb_parts, pl.lag, slots, progress, done);
        DVF_LAUNCH_CHECK();
        // 16 blocks per tile: each element's parts are summed in the same fixed order, only more of them are in flight
        // (4 blocks per tile: 125 us of dependent loads per EM iteration at M = 500)
        k_gram_reduce_parts<<<dim3(pl.n_tiles, 16), 256, 0, st>>>(tiles, (const int*)(base + pl.off_poff), slots, G, pl.ldg,
                                                                progress, pl.n_counters, done);
        DVF_LAUNCH_CHECK();
        return DVF_OK;
    }
    k_gram<<<pl.n_ctas, GR_THREADS, GR_SMEM, st>>>(mapUY, mapP, N, tiles, pl.n_tiles, (const int*)(base + pl.off_boff),
                                                  (const int*)(base + pl.off_btiles), pl.n_bundles, pl.n_groups, pl.sb,
                                                  pl.lag, slots, progress, done);
    DVF_LAUNCH_CHECK();
    k_gram_reduce<<<dim3(pl.n_tiles, 4), 256, 0, st>>>(tiles, pl.n_tiles, pl.n_groups, slots, G, pl.ldg, progress,
                                                      pl.n_ctas, done);
    DVF_LAUNCH_CHECK();
    return DVF_OK;
}

}  // namespace dvf

using namespace dvf;

extern "C" int64_t dvf_fit_row_width(int M, int D) { return ((int64_t)M + D + 63) / 64 * 64; }

extern "C" int dvf_gram_scratch_bytes(int64_t N, int D, int M, size_t* out) {
    DVF_ARG(out && N >= 0 && D >= 1 && M >= 1, "bad argument");
    *out = make_gram_plan(N, D, M).scratch_bytes;
    return DVF_OK;
}

extern "C" int dvf_gram_ld(int D, int M) { return (int)dvf_fit_row_width(M, D); }

extern "C" int dvf_gram(const double* UY, const double* P, double minP, int64_t N, int D, int M, double* G,
                        void* scratch, size_t scratch_bytes, void* stream) {
    DVF_ARG(UY && P && G && scratch, "NULL pointer");
    DVF_ARG(N >= 1 && D >= 1 && M >= 1, "bad shape");
    gram_plan pl = make_gram_plan(N, D, M);
    if (scratch_bytes < pl.scratch_bytes) {
        set_error("gram scratch too small: %zu < %zu", scratch_bytes, pl.scratch_bytes);
        return DVF_ERR_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    if (int rc = gram_prepare(pl, N, D, M, scratch, st)) return rc;
    double* Pc = (double*)((char*)scratch + pl.off_pclamp);
    k_clamp_copy<<<(unsigned)std::min<int64_t>((N + 255) / 256, 1184), 256, 0, st>>>(P, N, minP, Pc);
    DVF_LAUNCH_CHECK();
    return gram_run(pl, UY, Pc, N, (int)dvf_fit_row_width(M, D), G, (double*)scratch, nullptr, st);
}

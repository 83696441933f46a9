// K3: weighted Gram + right-hand side of the SparseVFC M-step (reference scVectorField.py:468-476)
//        G = [U | Y]^T diag(max(P, minP)) [U | Y]      (lower triangle only)
// so that  lhs = G[:M, :M] + lambda sigma2 Kc   and   rhs[m][d] = G[M + d][m].
// The reference materialises the M x N temporary U^T * repmat(P) every iteration; here the N x W matrix
// UY = [U | Y | 0] is built once (K1) and streamed through FP64 tensor cores (DMMA m8n8k4, the contraction
// index is the cell).  The lower triangle is covered by output tiles of a few shapes (128x128 off the diagonal,
// 64x64 pieces on it, thin strips for the tail rows) so that little more than the needed triangle is computed;
// the (tile, cell-range) work space is cut into equal-cost contiguous segments, one run of segments per SM
// (persistent CTAs, static schedule).  Every segment writes a partial tile; k_gram_reduce sums the partials of a
// tile in a fixed order, so the result is bitwise reproducible (no atomics).
#include <algorithm>
#include <vector>
#include "em.h"

namespace dvf {

constexpr int GR_KC = 32;        // cells per pipeline stage
constexpr int GR_STAGES = 3;
constexpr int GR_SLOT = 128 * 128;   // doubles per partial-tile slot

// tile variants: RT x CT DMMA tiles per warp, 8 warps as WI x WJ
enum { GV_FULL = 0, GV_HALF = 1, GV_STRIP8 = 2, GV_STRIP4 = 3, GV_STRIP2 = 4, GV_STRIP1 = 5, GV_COUNT = 6 };
template <int V> struct gvar;
template <> struct gvar<GV_FULL>   { static constexpr int RT = 4, CT = 8, WI = 4, WJ = 2; };   // 128 x 128
template <> struct gvar<GV_HALF>   { static constexpr int RT = 2, CT = 4, WI = 4, WJ = 2; };   //  64 x  64
template <> struct gvar<GV_STRIP8> { static constexpr int RT = 8, CT = 2, WI = 1, WJ = 8; };   //  64 x 128
template <> struct gvar<GV_STRIP4> { static constexpr int RT = 4, CT = 2, WI = 1, WJ = 8; };   //  32 x 128
template <> struct gvar<GV_STRIP2> { static constexpr int RT = 2, CT = 2, WI = 1, WJ = 8; };   //  16 x 128
template <> struct gvar<GV_STRIP1> { static constexpr int RT = 1, CT = 2, WI = 1, WJ = 8; };   //   8 x 128
// relative cost of one 32-cell step (DMMA count x a penalty for the thinner shapes, measured on B200)
static double kVarCost[GV_COUNT] = {256.0, 64.0 * 1.12, 128.0 * 1.15, 64.0 * 1.25, 32.0 * 1.5, 16.0 * 2.2};
static int g_force_G = 0;
static void read_tuning_env() {   // developer knobs (tools/gram_tune.py); not part of the ABI
    static bool done = false;
    if (done) return;
    done = true;
    if (const char* e = getenv("DVF_GRAM_COST")) sscanf(e, "%lf,%lf,%lf,%lf,%lf,%lf", &kVarCost[0], &kVarCost[1], &kVarCost[2], &kVarCost[3], &kVarCost[4], &kVarCost[5]);
    if (const char* e = getenv("DVF_GRAM_G")) g_force_G = atoi(e);
}

struct gram_tile { int variant, row0, col0, seg0, nseg; };

constexpr size_t GR_SMEM = sizeof(double) * (size_t)GR_STAGES * (GR_KC * (132 + 132) + GR_KC);

template <int V>
__device__ __forceinline__ void gram_segment(const double* __restrict__ UY, const double* __restrict__ P, double minP,
                                             int64_t N, int W, int row0, int col0, int step_lo, int step_hi,
                                             double* __restrict__ out, bool accumulate, double* smem) {
    using gv = gvar<V>;
    constexpr int RT = gv::RT, CT = gv::CT, WI = gv::WI, WJ = gv::WJ;
    constexpr int TR = 8 * RT * WI, TC = 8 * CT * WJ;
    constexpr int LDA = TR + ((4 - (TR & 15)) & 15), LDB = TC + ((4 - (TC & 15)) & 15);   // == 4 (mod 16)
    constexpr int KC = GR_KC, STAGES = GR_STAGES;
    double* As = smem;                         // [STAGES][KC][LDA]
    double* Bs = As + STAGES * KC * LDA;       // [STAGES][KC][LDB]
    double* Ps = Bs + STAGES * KC * LDB;       // [STAGES][KC]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int wi = warp / WJ, wj = warp % WJ;
    const int n_steps = step_hi - step_lo;
    const int64_t n_beg = (int64_t)step_lo * KC;

    auto load_stage = [&](int step, int stage) {
        const int64_t n0 = n_beg + (int64_t)step * KC;
        double* a = As + stage * KC * LDA;
        double* b = Bs + stage * KC * LDB;
        for (int e = threadIdx.x; e < KC * (TR / 2); e += 256) {
            const int r = e / (TR / 2), c = 2 * (e % (TR / 2));
            const bool ok = (n0 + r < N) && (row0 + c < W);
            cp_async16_zfill(a + r * LDA + c, UY + (ok ? (n0 + r) * (int64_t)W + row0 + c : 0), ok);
        }
        for (int e = threadIdx.x; e < KC * (TC / 2); e += 256) {
            const int r = e / (TC / 2), c = 2 * (e % (TC / 2));
            const bool ok = (n0 + r < N) && (col0 + c < W);
            cp_async16_zfill(b + r * LDB + c, UY + (ok ? (n0 + r) * (int64_t)W + col0 + c : 0), ok);
        }
        if (threadIdx.x < KC / 2) {
            const int64_t row = n0 + 2 * threadIdx.x;   // even row, P readable up to an even count (padded)
            cp_async16_zfill(Ps + stage * KC + 2 * threadIdx.x, P + (row < N ? row : 0), row < N);
        }
    };

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

    __syncthreads();   // the previous segment's readers are done with the ring
#pragma unroll
    for (int st = 0; st < STAGES - 1; ++st) {
        if (st < n_steps) load_stage(st, st);
        cp_async_commit();
    }
    for (int step = 0; step < n_steps; ++step) {
        cp_async_wait<STAGES - 2>();
        __syncthreads();
        {
            const int nxt = step + STAGES - 1;
            if (nxt < n_steps) load_stage(nxt, nxt % STAGES);
            cp_async_commit();
        }
        const int stage = step % STAGES;
        const double* a = As + stage * KC * LDA + wi * (8 * RT) + g;
        const double* b = Bs + stage * KC * LDB + wj * (8 * CT) + g;
        const double* ps = Ps + stage * KC;
        const int64_t n0 = n_beg + (int64_t)step * KC;
#pragma unroll (RT * CT >= 32 ? 2 : 8)
        for (int k4 = 0; k4 < KC / 4; ++k4) {
            const int row = 4 * k4 + t;
            double p = fmax(ps[row], minP);
            if (n0 + row >= N) p = 0.0;   // rows past the end (zero-filled data, unspecified P)
#ifdef DVF_EXP_NOSCALE
            p = 1.0;
#endif
            double af[RT], bf[CT];
#pragma unroll
            for (int i = 0; i < RT; ++i) af[i] = a[row * LDA + i * 8];
#pragma unroll
            for (int j = 0; j < CT; ++j) bf[j] = b[row * LDB + j * 8];
#ifdef DVF_EXP_NOSCALE
            if (false) {
#else
            if (RT <= CT) {
#endif
#pragma unroll
                for (int i = 0; i < RT; ++i) af[i] *= p;
            } else {
#ifndef DVF_EXP_NOSCALE
#pragma unroll
                for (int j = 0; j < CT; ++j) bf[j] *= p;
#endif
            }
#pragma unroll
            for (int i = 0; i < RT; ++i)
#pragma unroll
                for (int j = 0; j < CT; ++j) dmma(acc[i][j][0], acc[i][j][1], af[i], bf[j]);
        }
    }
    cp_async_wait<0>();
#pragma unroll
    for (int i = 0; i < RT; ++i)
#pragma unroll
        for (int j = 0; j < CT; ++j) {
            const int r = wi * (8 * RT) + i * 8 + g, c = wj * (8 * CT) + j * 8 + 2 * t;
            *reinterpret_cast<double2*>(out + r * TC + c) = make_double2(acc[i][j][0], acc[i][j][1]);
        }
}

// One CTA = one bundle of tiles (equal total cost across bundles) x one group's cell range.  The G CTAs of a
// group walk the same cells in the same order at the same pace, so a row block of UY is fetched from HBM once
// and served to the rest of the group from L2.  A multi-tile bundle round-robins over its tiles in blocks of
// GR_SB steps, carrying each tile's partial in its private scratch slot.
constexpr int GR_SB = 32;

template <int V>
__device__ __forceinline__ void gram_tile_range(const double* UY, const double* P, double minP, int64_t N, int W,
                                                const gram_tile& tl, int lo, int hi, double* out, bool acc, double* smem) {
    gram_segment<V>(UY, P, minP, N, W, tl.row0, tl.col0, lo, hi, out, acc, smem);
}

__global__ void __launch_bounds__(256, 1) k_gram(const double* __restrict__ UY, const double* __restrict__ P, double minP,
                                                 int64_t N, int W, const gram_tile* __restrict__ tiles, int n_tiles,
                                                 const int* __restrict__ bundle_off, const int* __restrict__ bundle_tiles,
                                                 int n_bundles, int n_groups, double* __restrict__ slots,
                                                 const int* __restrict__ done) {
    if (done && *done) return;
    extern __shared__ double smem[];
    const int grp = blockIdx.x / n_bundles, bun = blockIdx.x % n_bundles;
    const int64_t n_steps = (N + GR_KC - 1) / GR_KC;
    const int s_lo = (int)(n_steps * grp / n_groups), s_hi = (int)(n_steps * (grp + 1) / n_groups);
    const int t_lo = bundle_off[bun], t_hi = bundle_off[bun + 1];
    const int sb = (t_hi - t_lo == 1) ? (s_hi - s_lo > 0 ? s_hi - s_lo : 1) : GR_SB;
    bool first = true;
    for (int b_lo = s_lo; first || b_lo < s_hi; b_lo += sb) {
        const int b_hi = min(b_lo + sb, s_hi);
        for (int q = t_lo; q < t_hi; ++q) {
            const int ti = bundle_tiles[q];
            const gram_tile tl = tiles[ti];
            double* out = slots + ((size_t)grp * n_tiles + ti) * GR_SLOT;
            switch (tl.variant) {
                case GV_FULL:   gram_tile_range<GV_FULL>(UY, P, minP, N, W, tl, b_lo, b_hi, out, !first, smem); break;
                case GV_HALF:   gram_tile_range<GV_HALF>(UY, P, minP, N, W, tl, b_lo, b_hi, out, !first, smem); break;
                case GV_STRIP8: gram_tile_range<GV_STRIP8>(UY, P, minP, N, W, tl, b_lo, b_hi, out, !first, smem); break;
                case GV_STRIP4: gram_tile_range<GV_STRIP4>(UY, P, minP, N, W, tl, b_lo, b_hi, out, !first, smem); break;
                case GV_STRIP2: gram_tile_range<GV_STRIP2>(UY, P, minP, N, W, tl, b_lo, b_hi, out, !first, smem); break;
                default:        gram_tile_range<GV_STRIP1>(UY, P, minP, N, W, tl, b_lo, b_hi, out, !first, smem); break;
            }
        }
        first = false;
    }
}

// G[row0 + r][col0 + c] = sum over groups of the tile's partials (fixed order); one CTA column per tile
__global__ void __launch_bounds__(256) k_gram_reduce(const gram_tile* __restrict__ tiles, int n_tiles, int n_groups,
                                                     const double* __restrict__ slots, double* __restrict__ G, int ldg,
                                                     const int* __restrict__ done) {
    if (done && *done) return;
    const gram_tile tl = tiles[blockIdx.x];
    const int TR = (tl.variant == GV_FULL) ? 128 : (tl.variant == GV_HALF || tl.variant == GV_STRIP8) ? 64
                   : (tl.variant == GV_STRIP4) ? 32 : (tl.variant == GV_STRIP2) ? 16 : 8;
    const int TC = (tl.variant == GV_HALF) ? 64 : 128;
    const double* src = slots + (size_t)blockIdx.x * GR_SLOT;
    for (int e = blockIdx.y * 256 + threadIdx.x; e < TR * TC; e += gridDim.y * 256) {
        double sum = 0.0;
        for (int gi = 0; gi < n_groups; ++gi) sum += src[(size_t)gi * n_tiles * GR_SLOT + e];
        const int r = tl.row0 + e / TC, c = tl.col0 + e % TC;
        if (r < ldg && c < ldg) G[(int64_t)r * ldg + c] = sum;
    }
}

// ---- launch plan ---------------------------------------------------------------------------------------
struct gram_host_plan {
    std::vector<gram_tile> tiles;
    std::vector<int> bundle_off, bundle_tiles;
    int n_groups = 1;
};

static void build_host_plan(int64_t N, int D, int M, int n_sms, gram_host_plan& hp) {
    read_tuning_env();
    const int units = (M + D + 7) / 8, nfull = units / 16, rt = units % 16;
    auto add = [&](int v, int r0, int c0) { hp.tiles.push_back(gram_tile{v, r0, c0, 0, 0}); };
    for (int s = 0; s < nfull; ++s) {
        for (int jb = 0; jb < s; ++jb) add(GV_FULL, 128 * s, 128 * jb);
        add(GV_HALF, 128 * s, 128 * s);
        add(GV_HALF, 128 * s + 64, 128 * s);
        add(GV_HALF, 128 * s + 64, 128 * s + 64);
    }
    int row = 128 * nfull, rem = rt;
    const int chunk_units[4] = {8, 4, 2, 1}, chunk_var[4] = {GV_STRIP8, GV_STRIP4, GV_STRIP2, GV_STRIP1};
    for (int q = 0; q < 4; ++q)
        while (rem >= chunk_units[q]) {
            for (int jb = 0; jb <= nfull; ++jb) add(chunk_var[q], row, 128 * jb);
            row += 8 * chunk_units[q];
            rem -= chunk_units[q];
        }
    const int nt = (int)hp.tiles.size();
    const int64_t n_steps = (N + GR_KC - 1) / GR_KC;
    std::vector<int> by_cost(nt);
    for (int i = 0; i < nt; ++i) by_cost[i] = i;
    std::stable_sort(by_cost.begin(), by_cost.end(), [&](int a, int b) {
        return kVarCost[hp.tiles[a].variant] > kVarCost[hp.tiles[b].variant];
    });
    double total = 0.0;
    for (auto& tl : hp.tiles) total += kVarCost[tl.variant];
    // choose the number of bundles G (LPT bin packing) that maximises
    //   total / (G * max_bin)  x  (G * groups) / n_sms ,  groups = min(n_sms / G, n_steps)
    double best_eff = -1.0;
    int best_g = 1;
    std::vector<int> best_assign;
    for (int G = 1; G <= nt && G <= n_sms; ++G) {
        std::vector<double> bin(G, 0.0);
        std::vector<int> assign(nt, 0);
        for (int id : by_cost) {
            int k = (int)(std::min_element(bin.begin(), bin.end()) - bin.begin());
            bin[k] += kVarCost[hp.tiles[id].variant];
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
    size_t o = 0;
    pl.off_tiles = o;  o += al256(sizeof(gram_tile) * hp.tiles.size());
    pl.off_boff = o;   o += al256(sizeof(int) * hp.bundle_off.size());
    pl.off_btiles = o; o += al256(sizeof(int) * hp.bundle_tiles.size());
    pl.off_slots = o;  o += sizeof(double) * (size_t)GR_SLOT * pl.n_tiles * pl.n_groups;
    pl.scratch_bytes = o;
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
    return DVF_OK;
}

int gram_run(const gram_plan& pl, const double* UY, const double* P, double minP, int64_t N, int W, double* G,
             double* scratch, const int* done, cudaStream_t st) {
    static bool attr_set = false;
    if (!attr_set) {
        DVF_CUDA(cudaFuncSetAttribute(k_gram, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GR_SMEM));
        attr_set = true;
    }
    char* base = (char*)scratch;
    const gram_tile* tiles = (const gram_tile*)(base + pl.off_tiles);
    double* slots = (double*)(base + pl.off_slots);
    k_gram<<<pl.n_ctas, 256, GR_SMEM, st>>>(UY, P, minP, N, W, tiles, pl.n_tiles, (const int*)(base + pl.off_boff),
                                            (const int*)(base + pl.off_btiles), pl.n_bundles, pl.n_groups, slots, done);
    DVF_LAUNCH_CHECK();
    k_gram_reduce<<<dim3(pl.n_tiles, 4), 256, 0, st>>>(tiles, pl.n_tiles, pl.n_groups, slots, G, pl.ldg, done);
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
    if (int rc = gram_prepare(pl, N, D, M, scratch, (cudaStream_t)stream)) return rc;
    return gram_run(pl, UY, P, minP, N, (int)dvf_fit_row_width(M, D), G, (double*)scratch, nullptr, (cudaStream_t)stream);
}

// K3: weighted Gram + right-hand side of the SparseVFC M-step (reference scVectorField.py:468-476)
//        G = [U | Y]^T diag(max(P, minP)) [U | Y]      (lower triangle, in B x B tiles)
// so that  lhs = G[:M, :M] + lambda sigma2 Kc   and   rhs[m][d] = G[M + d][m].
// The reference materialises the M x N temporary U^T * repmat(P) every iteration; here the N x W matrix
// UY = [U | Y | 0] is built once (K1) and streamed: a CTA owns one lower-triangular B x B output tile and one
// contiguous slice of cells, stages 32-cell row tiles through a cp.async ring in shared memory and feeds
// FP64 tensor cores (DMMA m8n8k4, the contraction index is the cell).  The weight p_n is folded into the A
// fragment on the fly.  Partial tiles are written to scratch and summed in a fixed order by k_gram_reduce,
// so the result is bitwise reproducible (no atomics).
#include "em.h"

namespace dvf {

template <int B>
struct gram_cfg {
    static constexpr int KC = 32;                       // cells per stage
    static constexpr int STAGES = 3;
    static constexpr int LD = B + 4;                    // == 4 (mod 16)
    static constexpr int WTI = B / 4, WTJ = B / 2;      // warp tile (8 warps as 4 x 2)
    static constexpr int TI = WTI / 8, TJ = WTJ / 8;    // DMMA tiles per warp
    static constexpr size_t SMEM = sizeof(double) * (size_t)STAGES * (2 * KC * LD + KC);
};

__host__ __device__ inline void tri_decode(int p, int& bi, int& bj) {
    int r = (int)((sqrt(8.0 * p + 1.0) - 1.0) * 0.5);
    while ((r + 1) * (r + 2) / 2 <= p) ++r;
    while (r * (r + 1) / 2 > p) --r;
    bi = r;
    bj = p - r * (r + 1) / 2;
}

template <int B>
__global__ void __launch_bounds__(256) k_gram(const double* __restrict__ UY, const double* __restrict__ P,
                                              double minP, int64_t N, int W, int64_t chunk,
                                              double* __restrict__ scratch, const int* __restrict__ done) {
    using cfg = gram_cfg<B>;
    constexpr int KC = cfg::KC, STAGES = cfg::STAGES, LD = cfg::LD, TI = cfg::TI, TJ = cfg::TJ;
    if (done && *done) return;
    extern __shared__ double smem[];
    double* As = smem;                               // [STAGES][KC][LD]
    double* Bs = As + STAGES * KC * LD;              // [STAGES][KC][LD]
    double* Ps = Bs + STAGES * KC * LD;              // [STAGES][KC]

    int bi, bj;
    tri_decode(blockIdx.x, bi, bj);
    const bool diag = (bi == bj);
    const int S = gridDim.y, s = blockIdx.y;
    const int64_t n_beg = (int64_t)s * chunk;
    const int64_t n_end = min(N, n_beg + chunk);
    const int n_steps = (n_end > n_beg) ? (int)((n_end - n_beg + KC - 1) / KC) : 0;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int wi = warp >> 1, wj = warp & 1;

    auto load_stage = [&](int step, int stage) {
        const int64_t n0 = n_beg + (int64_t)step * KC;
        double* a = As + stage * KC * LD;
        double* b = Bs + stage * KC * LD;
        constexpr int CH = B / 2;   // 16-byte chunks per row
        for (int e = threadIdx.x; e < KC * CH; e += 256) {
            const int r = e / CH, c = e % CH;
            const int64_t row = n0 + r;
            const bool rv = row < n_end;
            const int ca = bi * B + 2 * c, cb = bj * B + 2 * c;
            const double* src = UY + (rv ? row : 0) * (int64_t)W;
            cp_async16_zfill(a + r * LD + 2 * c, src + (ca < W ? ca : 0), rv && ca < W);
            if (!diag) cp_async16_zfill(b + r * LD + 2 * c, src + (cb < W ? cb : 0), rv && cb < W);
        }
        if (threadIdx.x < KC / 2) {
            const int64_t row = n0 + 2 * threadIdx.x;   // N-slices start on even rows; P is padded by the caller
            cp_async16_zfill(Ps + stage * KC + 2 * threadIdx.x, P + (row < n_end ? row : 0), row < n_end);
        }
    };

    double acc[TI][TJ][2];
#pragma unroll
    for (int i = 0; i < TI; ++i)
#pragma unroll
        for (int j = 0; j < TJ; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

#pragma unroll
    for (int st = 0; st < STAGES - 1; ++st) {
        if (st < n_steps) load_stage(st, st);
        cp_async_commit();
    }
    for (int step = 0; step < n_steps; ++step) {
        cp_async_wait<STAGES - 2>();
        __syncthreads();
        {   // prefetch the stage that was consumed one iteration ago
            const int nxt = step + STAGES - 1;
            if (nxt < n_steps) load_stage(nxt, nxt % STAGES);
            cp_async_commit();
        }
        const int stage = step % STAGES;
        const double* a = As + stage * KC * LD + wi * cfg::WTI + g;
        const double* b = (diag ? As : Bs) + stage * KC * LD + wj * cfg::WTJ + g;
        const double* ps = Ps + stage * KC;
        const int64_t n0 = n_beg + (int64_t)step * KC;
#pragma unroll 2
        for (int k4 = 0; k4 < KC / 4; ++k4) {
            const int row = 4 * k4 + t;
            double p = fmax(ps[row], minP);
            if (n0 + row >= n_end) p = 0.0;   // the odd tail cell of a 16-byte P copy
            double af[TI], bf[TJ];
#pragma unroll
            for (int i = 0; i < TI; ++i) af[i] = p * a[row * LD + i * 8];
#pragma unroll
            for (int j = 0; j < TJ; ++j) bf[j] = b[row * LD + j * 8];
#pragma unroll
            for (int i = 0; i < TI; ++i)
#pragma unroll
                for (int j = 0; j < TJ; ++j) dmma(acc[i][j][0], acc[i][j][1], af[i], bf[j]);
        }
    }
    cp_async_wait<0>();

    double* out = scratch + ((int64_t)blockIdx.x * S + s) * (B * B);
#pragma unroll
    for (int i = 0; i < TI; ++i)
#pragma unroll
        for (int j = 0; j < TJ; ++j) {
            const int r = wi * cfg::WTI + i * 8 + g, c = wj * cfg::WTJ + j * 8 + 2 * t;
            *reinterpret_cast<double2*>(out + r * B + c) = make_double2(acc[i][j][0], acc[i][j][1]);
        }
}

// G[bi*B + i][bj*B + j] = sum_s scratch[pair][s][i][j]  (fixed order)
template <int B>
__global__ void __launch_bounds__(256) k_gram_reduce(const double* __restrict__ scratch, int S, double* __restrict__ G,
                                                     int ldg, const int* __restrict__ done) {
    if (done && *done) return;
    int bi, bj;
    tri_decode(blockIdx.x, bi, bj);
    const double* src = scratch + (int64_t)blockIdx.x * S * (B * B);
    for (int e = blockIdx.y * 256 + threadIdx.x; e < B * B; e += gridDim.y * 256) {
        double sum = 0.0;
        for (int s = 0; s < S; ++s) sum += src[(int64_t)s * (B * B) + e];
        G[(int64_t)(bi * B + e / B) * ldg + bj * B + e % B] = sum;
    }
}

// ---- launch plan --------------------------------------------------------------------------------
gram_plan make_gram_plan(int64_t N, int D, int M) {
    gram_plan pl;
    const int W = (int)dvf_fit_row_width(M, D);
    const char* force = getenv("DVF_GRAM_TILE");
    pl.B = (W >= 384) ? 128 : 64;
    if (force && atoi(force) == 64) pl.B = 64;
    if (force && atoi(force) == 128) pl.B = 128;
    pl.nblk = (W + pl.B - 1) / pl.B;
    pl.pairs = pl.nblk * (pl.nblk + 1) / 2;
    pl.ldg = pl.nblk * pl.B;
    // one wave of CTAs: the largest S with pairs * S <= #SMs * ctas_per_sm, at least 1
    const int ctas_per_sm = (pl.B == 128) ? 1 : 2;
    const int slots = num_sms() * ctas_per_sm;
    int S = slots / pl.pairs;
    if (S < 1) S = 1;
    const int64_t steps = (N + 31) / 32;
    if (S > steps) S = (int)(steps > 0 ? steps : 1);
    pl.S = S;
    int64_t chunk = (steps + S - 1) / S * 32;
    pl.chunk = chunk;
    pl.scratch_bytes = sizeof(double) * (size_t)pl.pairs * S * pl.B * pl.B;
    return pl;
}

template <int B>
static int launch_gram(const gram_plan& pl, const double* UY, const double* P, double minP, int64_t N, int W, double* G,
                       double* scratch, const int* done, cudaStream_t st) {
    using cfg = gram_cfg<B>;
    static bool attr_set = false;
    if (!attr_set) {
        DVF_CUDA(cudaFuncSetAttribute(k_gram<B>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg::SMEM));
        attr_set = true;
    }
    k_gram<B><<<dim3(pl.pairs, pl.S), 256, cfg::SMEM, st>>>(UY, P, minP, N, W, pl.chunk, scratch, done);
    DVF_LAUNCH_CHECK();
    k_gram_reduce<B><<<dim3(pl.pairs, B * B / 1024), 256, 0, st>>>(scratch, pl.S, G, pl.ldg, done);
    DVF_LAUNCH_CHECK();
    return DVF_OK;
}

int gram_run(const gram_plan& pl, const double* UY, const double* P, double minP, int64_t N, int W, double* G,
             double* scratch, const int* done, cudaStream_t st) {
    if (pl.B == 128) return launch_gram<128>(pl, UY, P, minP, N, W, G, scratch, done, st);
    return launch_gram<64>(pl, UY, P, minP, N, W, G, scratch, done, st);
}

}  // namespace dvf

using namespace dvf;

extern "C" int64_t dvf_fit_row_width(int M, int D) { return ((int64_t)M + D + 63) / 64 * 64; }

extern "C" int dvf_gram_scratch_bytes(int64_t N, int D, int M, size_t* out) {
    DVF_ARG(out && N >= 0 && D >= 1 && M >= 1, "bad argument");
    *out = make_gram_plan(N, D, M).scratch_bytes;
    return DVF_OK;
}

extern "C" int dvf_gram_ld(int D, int M) { return make_gram_plan(1, D, M).ldg; }

extern "C" int dvf_gram(const double* UY, const double* P, double minP, int64_t N, int D, int M, double* G,
                        void* scratch, size_t scratch_bytes, void* stream) {
    DVF_ARG(UY && P && G && scratch, "NULL pointer");
    DVF_ARG(N >= 1 && D >= 1 && M >= 1, "bad shape");
    gram_plan pl = make_gram_plan(N, D, M);
    if (scratch_bytes < pl.scratch_bytes) {
        set_error("gram scratch too small: %zu < %zu", scratch_bytes, pl.scratch_bytes);
        return DVF_ERR_WORKSPACE;
    }
    return gram_run(pl, UY, P, minP, N, (int)dvf_fit_row_width(M, D), G, (double*)scratch, nullptr, (cudaStream_t)stream);
}

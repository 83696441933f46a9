// Error plumbing and ABI version of libdvf_b200.
#include "common.cuh"
#include "tma.cuh"
#include "host_cumsum.h"
#include <algorithm>
#include <thread>
#include <vector>

namespace dvf {
static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
int cuda_fail(cudaError_t e, const char* what, int line) {
    set_error("CUDA error '%s' (%d) in %s at line %d", cudaGetErrorString(e), (int)e, what, line);
    return DVF_ERR_CUDA;
}
long long g_launch_count = 0;

// ---- tensor maps ---------------------------------------------------------------------------------------
typedef CUresult (*encode_tiled_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                    const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static encode_tiled_fn get_encode_tiled() {
    static encode_tiled_fn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = (encode_tiled_fn)p;
    }
    return fn;
}
static int encode_2d(CUtensorMap* map, const void* base, uint64_t cols, uint64_t rows, uint64_t pitch, uint32_t bc,
                     uint32_t br, CUtensorMapSwizzle sw, CUtensorMapL2promotion l2) {
    encode_tiled_fn enc = get_encode_tiled();
    if (!enc) { set_error("cuTensorMapEncodeTiled is not available from this driver"); return DVF_ERR_CUDA; }
    if (((uintptr_t)base & 15) != 0 || (pitch & 15) != 0) {
        set_error("TMA source needs a 16-byte aligned base and row pitch");
        return DVF_ERR_ARG;
    }
    const cuuint64_t dims[2] = {cols, rows}, strides[1] = {pitch};
    const cuuint32_t box[2] = {bc, br}, estr[2] = {1, 1};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, const_cast<void*>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, sw, l2, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed: %d", (int)r); return DVF_ERR_CUDA; }
    return DVF_OK;
}
int make_tmap_2d(CUtensorMap* map, const void* base, uint64_t cols, uint64_t rows, uint64_t pitch_bytes) {
    return encode_2d(map, base, cols, rows, pitch_bytes, TMA_BOX_COLS, TMA_BOX_ROWS, CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_128B);
}
int make_tmap_3d(CUtensorMap* map, const void* base, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t stride1_bytes,
                 uint64_t stride2_bytes) {
    encode_tiled_fn enc = get_encode_tiled();
    if (!enc) { set_error("cuTensorMapEncodeTiled is not available from this driver"); return DVF_ERR_CUDA; }
    if (((uintptr_t)base & 15) != 0 || (stride1_bytes & 15) != 0 || (stride2_bytes & 15) != 0) {
        set_error("TMA source needs a 16-byte aligned base and strides");
        return DVF_ERR_ARG;
    }
    const cuuint64_t dims[3] = {d0, d1, d2}, strides[2] = {stride1_bytes, stride2_bytes};
    const cuuint32_t box[3] = {TMA_BOX_COLS, 1, TMA_BOX_ROWS}, estr[3] = {1, 1, 1};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, const_cast<void*>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled (rank 3) failed: %d", (int)r); return DVF_ERR_CUDA; }
    return DVF_OK;
}
int pool_alloc(void** p, size_t bytes, cudaStream_t st) {
    static bool pool_set = false;
    if (!pool_set) {
        int dev = 0;
        cudaMemPool_t pool;
        DVF_CUDA(cudaGetDevice(&dev));
        DVF_CUDA(cudaDeviceGetDefaultMemPool(&pool, dev));
        uint64_t thr = ~0ull;
        DVF_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr));
        pool_set = true;
    }
    DVF_CUDA(cudaMallocAsync(p, bytes, st));
    return DVF_OK;
}
int make_tmap_vec(CUtensorMap* map, const void* base, uint64_t n) {
    // rank 2 with a single row (a rank-1 map is rejected without a stride array)
    return encode_2d(map, base, n, 1, (n * 8 + 15) & ~(uint64_t)15, 32, 1, CU_TENSOR_MAP_SWIZZLE_NONE,
                     CU_TENSOR_MAP_L2_PROMOTION_NONE);
}

// ---- CUDA-event profile of the Gram launches (the dominant kernel) -----------------------------------
static bool g_prof_on = false;
static const int kMaxProf = 1024;
static cudaEvent_t g_ev0[kMaxProf], g_ev1[kMaxProf];
static int g_ev_made = 0, g_ev_used = 0;

void profile_gram_begin(cudaStream_t st) {
    if (!g_prof_on || g_ev_used >= kMaxProf) return;
    if (g_ev_used >= g_ev_made) {
        cudaEventCreate(&g_ev0[g_ev_made]);
        cudaEventCreate(&g_ev1[g_ev_made]);
        ++g_ev_made;
    }
    cudaEventRecord(g_ev0[g_ev_used], st);
}
void profile_gram_end(cudaStream_t st) {
    if (!g_prof_on || g_ev_used >= kMaxProf) return;
    cudaEventRecord(g_ev1[g_ev_used], st);
    ++g_ev_used;
}
}  // namespace dvf

extern "C" int dvf_profile_enable(int on) {
    dvf::g_prof_on = on != 0;
    dvf::g_ev_used = 0;
    dvf::g_launch_count = 0;
    return DVF_OK;
}

// gram_ms: summed duration of the recorded Gram launches (kernel + fixed-order reduce); call after the stream
// has been synchronised.  Launches of EM iterations that ran after the stop flag was raised return at once and
// are included (they cost microseconds).
extern "C" int dvf_profile_get(double* gram_ms, int* gram_launches, long long* kernel_launches) {
    double total = 0.0;
    for (int i = 0; i < dvf::g_ev_used; ++i) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, dvf::g_ev0[i], dvf::g_ev1[i]) == cudaSuccess) total += ms;
    }
    if (gram_ms) *gram_ms = total;
    if (gram_launches) *gram_launches = dvf::g_ev_used;
    if (kernel_launches) *kernel_launches = dvf::g_launch_count;
    return DVF_OK;
}

// element-wise host passes of the sampling helper, split over a few threads once they are long enough to matter
// (8M cells on rank 0 of an 8-GPU job); every element is computed exactly as the scalar loop would
template <typename F>
static void host_parallel_for(int64_t n, F body, int64_t grain = 1 << 19) {
    unsigned hw = std::thread::hardware_concurrency();
    int nt = (int)std::min<int64_t>(std::min<int64_t>(hw ? hw : 1, 8), (n + grain - 1) / grain);
    if (nt <= 1) { body((int64_t)0, n); return; }
    std::vector<std::thread> th;
    const int64_t per = (n + nt - 1) / nt;
    for (int t = 1; t < nt; ++t) th.emplace_back([=] { body(std::min(n, t * per), std::min(n, (t + 1) * per)); });
    body((int64_t)0, std::min(n, per));
    for (auto& x : th) x.join();
}

// p = v / total (tools/sampling.py:244)
extern "C" int dvf_host_scale(const double* v, int64_t n, double total, double* p) {
    DVF_ARG(v && p && n >= 0, "bad argument");
    host_parallel_for(n, [=](int64_t a, int64_t b) { for (int64_t i = a; i < b; ++i) p[i] = v[i] / total; });
    return DVF_OK;
}

// numpy.cumsum(p) followed by cdf /= cdf[-1] (Generator.choice, replace=False): one strictly sequential pass (the
// rounding of every partial sum matters to searchsorted) + one element-wise pass
extern "C" int dvf_host_cdf(const double* p, int64_t n, double* cdf) {
    DVF_ARG(p && cdf && n >= 1, "bad argument");
    dvf::host_scaled_cumsum(p, false, 1.0, nullptr, 0, cdf, n);
    const double tot = cdf[n - 1];
    host_parallel_for(n, [=](int64_t a, int64_t b) { for (int64_t i = a; i < b; ++i) cdf[i] /= tot; });
    return DVF_OK;
}

// ---- numpy.sum of a contiguous float64 vector, bit for bit: NumPy's pairwise summation (8 interleaved partial sums
// on blocks of <= 128 elements, halves split at a multiple of 8 above that), with the top levels of the recursion
// spread over threads -- the tree, and therefore every rounding, is the same as in the serial recursion.
static double pw_sum(const double* a, int64_t n) {
    if (n < 8) {
        double res = 0.0;
        for (int64_t i = 0; i < n; ++i) res += a[i];
        return res;
    }
    if (n <= 128) {
        double r[8];
        for (int j = 0; j < 8; ++j) r[j] = a[j];
        int64_t i = 8;
        for (; i < n - (n % 8); i += 8)
            for (int j = 0; j < 8; ++j) r[j] += a[i + j];
        double res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
        for (; i < n; ++i) res += a[i];
        return res;
    }
    int64_t n2 = n / 2;
    n2 -= n2 % 8;
    return pw_sum(a, n2) + pw_sum(a + n2, n - n2);
}
static double pw_sum_par(const double* a, int64_t n, int depth) {
    if (depth <= 0 || n <= (1 << 16)) return pw_sum(a, n);
    int64_t n2 = n / 2;
    n2 -= n2 % 8;
    double left = 0.0;
    std::thread th([&] { left = pw_sum_par(a, n2, depth - 1); });
    const double right = pw_sum_par(a + n2, n - n2, depth - 1);
    th.join();
    return left + right;
}
extern "C" int dvf_host_pairwise_sum(const double* a, int64_t n, double* out) {
    DVF_ARG(out && (a || n == 0) && n >= 0, "bad argument");
    *out = 0.0 + pw_sum_par(a, n, 3);
    return DVF_OK;
}

// raw[i] = sum_{k <= i} p_k, p_k = v[k] / total except p_k = 0 for the (ascending) indices in `skip`: the strictly
// sequential cumulative sum of Generator.choice's p (numpy/random/_generator.pyx, replace=False), formed in place.
extern "C" int dvf_host_choice_cumsum(const double* v, int64_t n, double total, const int64_t* skip, int64_t n_skip,
                                      double* raw) {
    DVF_ARG(v && raw && n >= 1 && (skip || n_skip == 0), "bad argument");
    // host_cumsum.cpp: quotients, zeroed entries and the additions, with the rounding of the sequential recurrence
    dvf::host_scaled_cumsum(v, true, total, skip, n_skip, raw, n);
    return DVF_OK;
}

// out[q] = numpy.searchsorted(raw / raw[n-1], x[q], side='right') without forming the quotient array: the first i
// with raw[i] / last > x (division by a positive constant keeps the order, so bisection on the fly is exact)
extern "C" int dvf_host_searchsorted_div(const double* raw, int64_t n, const double* x, int64_t m, int64_t* out) {
    DVF_ARG(raw && x && out && n >= 1 && m >= 0, "bad argument");
    const double last = raw[n - 1];
    for (int64_t q = 0; q < m; ++q) {
        int64_t lo = 0, hi = n;
        while (lo < hi) {
            const int64_t mid = lo + (hi - lo) / 2;
            if (raw[mid] / last > x[q]) hi = mid; else lo = mid + 1;
        }
        out[q] = lo;
    }
    return DVF_OK;
}

extern "C" int dvf_memcpy2d_async(void* dst, size_t dpitch, const void* src, size_t spitch, size_t width_bytes,
                                  size_t height, int to_host, void* stream) {
    DVF_ARG(dst && src && dpitch >= width_bytes && spitch >= width_bytes, "bad argument");
    if (width_bytes == 0 || height == 0) return DVF_OK;
    DVF_CUDA(cudaMemcpy2DAsync(dst, dpitch, src, spitch, width_bytes, height,
                               to_host ? cudaMemcpyDeviceToHost : cudaMemcpyHostToDevice, (cudaStream_t)stream));
    return DVF_OK;
}

extern "C" int dvf_abi_version(void) { return 2; }
extern "C" const char* dvf_last_error(void) { return dvf::g_err; }

// Host side of the control-point sampling (tools/sampling.py:242-245 -> numpy Generator.choice, replace=False):
//     raw = numpy.cumsum(speed / total)        with the cells already drawn set to zero
// bit for bit, on several host threads.  Rank 0 of an 8-GPU job does this over all 8M cells in every fit.
//
// The recurrence acc <- fl(acc + p) rounds at every step, so a tree of partial sums gives other bits.  But while acc
// stays inside one binade [2^e, 2^(e+1)) every partial sum is a multiple of u = 2^(e-52), and fl(acc + p) = acc + r(p)
// with r(p) = p rounded to the nearest multiple of u -- the same r whatever acc is, unless p sits exactly half way
// between two multiples (then the parity of acc decides).  r(p) = (B + p) - B for B = 2^e and p < B / 2, and sums of
// multiples of u below 2^(e+1) are exact in any order.  So for a block of elements that starts and ends in the same
// binade, with no half-way element, no element >= B / 2 and no negative / non-finite element, the value of acc after
// the block is acc + sum r(p): computable per block, on any thread, vectorised.  The blocks' exact start values then
// follow from a short serial scan, and every block runs the plain recurrence from its own start value.  Blocks that
// break a condition (a binade crossing, a tie, zeros at the head of the array, ...) run the plain recurrence inside
// the serial scan.  Binades are predicted from approximate block sums taken while dividing; a wrong prediction only
// sends the block to the serial path.  tests/test_prelude_cpu.py compares with numpy.cumsum on adversarial inputs.
#include "host_cumsum.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstring>
#include <exception>
#include <thread>
#include <vector>
#if defined(__x86_64__)
#include <immintrin.h>
#endif

namespace dvf {
namespace {

constexpr int64_t BLK = 1 << 14;   // 128 KB of doubles: stays in the core's L2 between the passes over a block

struct SpinBarrier {
    explicit SpinBarrier(int n) : n_(n) {}
    void wait() {
        const int g = gen_.load(std::memory_order_acquire);
        if (count_.fetch_add(1, std::memory_order_acq_rel) + 1 == n_) {
            count_.store(0, std::memory_order_relaxed);
            gen_.store(g + 1, std::memory_order_release);
            return;
        }
        for (int spins = 0; gen_.load(std::memory_order_acquire) == g; ++spins)
            if (spins > 2000) std::this_thread::yield();
    }
    std::atomic<int> count_{0}, gen_{0};
    const int n_;
};

void recurrence(double* a, int64_t lo, int64_t hi, double& acc) {
    double s = acc;
    for (int64_t i = lo; i < hi; ++i) { s += a[i]; a[i] = s; }
    acc = s;
}

// raw[i] = v[i] / total; returns an approximate sum of the quotients (any order)
double divide_block_scalar(const double* v, double total, double* raw, int64_t n) {
    double s0 = 0, s1 = 0;
    int64_t i = 0;
    for (; i + 2 <= n; i += 2) {
        const double q0 = v[i] / total, q1 = v[i + 1] / total;
        raw[i] = q0;
        raw[i + 1] = q1;
        s0 += q0;
        s1 += q1;
    }
    for (; i < n; ++i) { raw[i] = v[i] / total; s0 += raw[i]; }
    return s0 + s1;
}

double sum_block(const double* a, int64_t n) {
    double s0 = 0, s1 = 0, s2 = 0, s3 = 0;
    int64_t i = 0;
    for (; i + 4 <= n; i += 4) { s0 += a[i]; s1 += a[i + 1]; s2 += a[i + 2]; s3 += a[i + 3]; }
    for (; i < n; ++i) s0 += a[i];
    return (s0 + s1) + (s2 + s3);
}

// sum of r(p) over the block (exact while it stays below 2B); *bad is set when the block must take the serial path
double rounded_sum_scalar(const double* a, int64_t n, double B, bool* bad_out) {
    const double half_u = B * 1.1102230246251565e-16 /* 2^-53 */, cap = 0.5 * B;
    double s0 = 0, s1 = 0;
    bool bad = false;
    int64_t i = 0;
    for (; i + 2 <= n; i += 2) {
        const double p0 = a[i], p1 = a[i + 1];
        const double t0 = B + p0, t1 = B + p1;   // the rounding of B + p is the point: never build with -ffast-math
        const double r0 = t0 - B, r1 = t1 - B;
        bad |= !(p0 >= 0.0) | !(p0 < cap) | (std::fabs(p0 - r0) == half_u);
        bad |= !(p1 >= 0.0) | !(p1 < cap) | (std::fabs(p1 - r1) == half_u);
        s0 += r0;
        s1 += r1;
    }
    for (; i < n; ++i) {
        const double p0 = a[i], t0 = B + p0, r0 = t0 - B;
        bad |= !(p0 >= 0.0) | !(p0 < cap) | (std::fabs(p0 - r0) == half_u);
        s0 += r0;
    }
    *bad_out = bad;
    return s0 + s1;
}

#if defined(__x86_64__)
__attribute__((target("avx2"))) double divide_block_avx2(const double* v, double total, double* raw, int64_t n) {
    const __m256d vt = _mm256_set1_pd(total);
    __m256d s0 = _mm256_setzero_pd(), s1 = _mm256_setzero_pd();
    int64_t i = 0;
    for (; i + 8 <= n; i += 8) {
        const __m256d q0 = _mm256_div_pd(_mm256_loadu_pd(v + i), vt), q1 = _mm256_div_pd(_mm256_loadu_pd(v + i + 4), vt);
        _mm256_storeu_pd(raw + i, q0);
        _mm256_storeu_pd(raw + i + 4, q1);
        s0 = _mm256_add_pd(s0, q0);
        s1 = _mm256_add_pd(s1, q1);
    }
    double lanes[4];
    _mm256_storeu_pd(lanes, _mm256_add_pd(s0, s1));
    double s = (lanes[0] + lanes[1]) + (lanes[2] + lanes[3]);
    for (; i < n; ++i) { raw[i] = v[i] / total; s += raw[i]; }
    return s;
}

__attribute__((target("avx2"))) double rounded_sum_avx2(const double* a, int64_t n, double B, bool* bad_out) {
    const double half_u = B * 1.1102230246251565e-16, cap = 0.5 * B;
    const __m256d vB = _mm256_set1_pd(B), vh = _mm256_set1_pd(half_u), vcap = _mm256_set1_pd(cap);
    const __m256d vzero = _mm256_setzero_pd();
    const __m256d absmask = _mm256_castsi256_pd(_mm256_set1_epi64x(0x7fffffffffffffffLL));
    __m256d s0 = vzero, s1 = vzero, vbad = vzero;
    int64_t i = 0;
    for (; i + 8 <= n; i += 8) {
        const __m256d p0 = _mm256_loadu_pd(a + i), p1 = _mm256_loadu_pd(a + i + 4);
        const __m256d r0 = _mm256_sub_pd(_mm256_add_pd(vB, p0), vB), r1 = _mm256_sub_pd(_mm256_add_pd(vB, p1), vB);
        const __m256d d0 = _mm256_and_pd(_mm256_sub_pd(p0, r0), absmask), d1 = _mm256_and_pd(_mm256_sub_pd(p1, r1), absmask);
        __m256d b0 = _mm256_cmp_pd(d0, vh, _CMP_EQ_OQ), b1 = _mm256_cmp_pd(d1, vh, _CMP_EQ_OQ);
        b0 = _mm256_or_pd(b0, _mm256_cmp_pd(p0, vzero, _CMP_NGE_UQ));   // negative or NaN
        b1 = _mm256_or_pd(b1, _mm256_cmp_pd(p1, vzero, _CMP_NGE_UQ));
        b0 = _mm256_or_pd(b0, _mm256_cmp_pd(p0, vcap, _CMP_NLT_UQ));    // >= B / 2 (or NaN)
        b1 = _mm256_or_pd(b1, _mm256_cmp_pd(p1, vcap, _CMP_NLT_UQ));
        vbad = _mm256_or_pd(vbad, _mm256_or_pd(b0, b1));
        s0 = _mm256_add_pd(s0, r0);
        s1 = _mm256_add_pd(s1, r1);
    }
    double lanes[4];
    _mm256_storeu_pd(lanes, _mm256_add_pd(s0, s1));
    double s = (lanes[0] + lanes[1]) + (lanes[2] + lanes[3]);
    bool bad = _mm256_movemask_pd(vbad) != 0;
    if (i < n) {
        bool tail_bad = false;
        s += rounded_sum_scalar(a + i, n - i, B, &tail_bad);
        bad |= tail_bad;
    }
    *bad_out = bad;
    return s;
}

bool have_avx2() {
    static const bool yes = __builtin_cpu_supports("avx2");
    return yes;
}
#else
bool have_avx2() { return false; }
#endif

double divide_block(const double* v, double total, double* raw, int64_t n) {
#if defined(__x86_64__)
    if (have_avx2()) return divide_block_avx2(v, total, raw, n);
#endif
    return divide_block_scalar(v, total, raw, n);
}

double rounded_sum(const double* a, int64_t n, double B, bool* bad) {
#if defined(__x86_64__)
    if (have_avx2()) return rounded_sum_avx2(a, n, B, bad);
#endif
    return rounded_sum_scalar(a, n, B, bad);
}

}  // namespace

void host_scaled_cumsum(const double* v, bool divide, double total, const int64_t* skip, int64_t n_skip, double* raw,
                        int64_t n) {
    const int64_t nb = (n + BLK - 1) / BLK;
    const unsigned hw = std::thread::hardware_concurrency();
    const int nt = (int)std::min<int64_t>(std::min<int64_t>(hw ? hw : 1, 16), nb / 8);
    auto blk_lo = [](int64_t b) { return b * BLK; };
    auto blk_hi = [n](int64_t b) { return std::min(n, (b + 1) * BLK); };
    // what the scalar loop does to one block before the additions: quotients, drawn cells zeroed
    auto prepare = [&](int64_t b) -> double {
        const int64_t lo = blk_lo(b), hi = blk_hi(b);
        double approx;
        if (divide) approx = divide_block(v + lo, total, raw + lo, hi - lo);
        else {
            if (raw != v) std::memcpy(raw + lo, v + lo, sizeof(double) * (size_t)(hi - lo));
            approx = 0.0;
        }
        bool zeroed = false;
        for (const int64_t* s = std::lower_bound(skip, skip + n_skip, lo); s != skip + n_skip && *s < hi; ++s) {
            raw[*s] = 0.0;
            zeroed = true;
        }
        if (zeroed || !divide) approx = sum_block(raw + lo, hi - lo);
        return approx;
    };
    if (nt <= 1) {
        double acc = 0.0;
        for (int64_t b = 0; b < nb; ++b) { prepare(b); recurrence(raw, blk_lo(b), blk_hi(b), acc); }
        return;
    }

    std::vector<double> approx((size_t)nb), base((size_t)nb), rsum((size_t)nb), start((size_t)nb);
    std::vector<char> clean((size_t)nb);
    SpinBarrier bar(nt);
    auto worker = [&](int t) {
        const int64_t b0 = nb * t / nt, b1 = nb * (t + 1) / nt;
        for (int64_t b = b0; b < b1; ++b) approx[(size_t)b] = prepare(b);
        bar.wait();
        if (t == 0) {   // predicted binade per block: the approximate running sum before and after must agree on it
            double run = 0.0;
            for (int64_t b = 0; b < nb; ++b) {
                const double before = run;
                run += approx[(size_t)b];
                const bool ok = std::isfinite(run) && before >= 2.2250738585072014e-308 * 9007199254740992.0 &&
                                std::ilogb(before) == std::ilogb(run);
                clean[(size_t)b] = ok ? 1 : 0;
                base[(size_t)b] = ok ? std::ldexp(1.0, std::ilogb(before)) : 0.0;
            }
        }
        bar.wait();
        for (int64_t b = b0; b < b1; ++b) {
            if (!clean[(size_t)b]) continue;
            bool bad = false;
            rsum[(size_t)b] = rounded_sum(raw + blk_lo(b), blk_hi(b) - blk_lo(b), base[(size_t)b], &bad);
            if (bad) clean[(size_t)b] = 0;
        }
        bar.wait();
        if (t == 0) {   // exact start value of every block; the others run the recurrence here
            double acc = 0.0;
            for (int64_t b = 0; b < nb; ++b) {
                bool ok = clean[(size_t)b] != 0;
                if (ok) {
                    const double B = base[(size_t)b], end = acc + rsum[(size_t)b];   // multiples of u: exact below 2B
                    ok = acc >= B && acc < 2.0 * B && end < 2.0 * B;
                    if (ok) { start[(size_t)b] = acc; acc = end; }
                }
                if (!ok) { clean[(size_t)b] = 0; recurrence(raw, blk_lo(b), blk_hi(b), acc); }
            }
        }
        bar.wait();
        for (int64_t b = b0; b < b1; ++b) {
            if (!clean[(size_t)b]) continue;
            double s = start[(size_t)b];
            recurrence(raw, blk_lo(b), blk_hi(b), s);
        }
    };
    // helpers wait for `go`: if the host refuses a thread, the ones already started are released unused and the
    // whole array takes the serial path (nothing has been written yet)
    std::atomic<int> go{0};
    std::vector<std::thread> th;
    bool spawned = true;
    try {
        for (int t = 1; t < nt; ++t)
            th.emplace_back([&, t] {
                int g;
                while ((g = go.load(std::memory_order_acquire)) == 0) std::this_thread::yield();
                if (g == 1) worker(t);
            });
    } catch (const std::exception&) {
        spawned = false;
    }
    go.store(spawned ? 1 : 2, std::memory_order_release);
    if (spawned) worker(0);
    for (auto& x : th) x.join();
    if (!spawned) {
        double acc = 0.0;
        for (int64_t b = 0; b < nb; ++b) { prepare(b); recurrence(raw, blk_lo(b), blk_hi(b), acc); }
    }
}

}  // namespace dvf

// numerics.cuh — fp64 numerics of the scoring path, usable from device code and (for host-side unit tests of
// the expression order / table layout) from plain C++.
//
// Restates, operation for operation and left-associated exactly as written there:
//   stats.rs:41-53   log_beta_binomial_pmf / log_beta_calc
//   stats.rs:35-39   logsumexp
//   stats.rs:8-33    expected_log_beta_binomial_pmf (expectation only; the variance is never observable)
//   statrs 0.16.0    function::gamma::ln_gamma (Lanczos g=10.900511, n=11), function::factorial::ln_binomial
// Compile with -fmad=false (device) / -ffp-contract=off (host): Rust never fuses a*b+c.
#pragma once
#include <math.h>
#include <stdint.h>

#ifdef __CUDACC__
#define CLX_HD __host__ __device__ __forceinline__
#else
#define CLX_HD inline
#endif

namespace clx {

// Totals 1..K_LIGHT are memoised per locus; deeper entries go to the heavy list (direct Lanczos).
constexpr int K_LIGHT = 3;
// Heavy entries deeper than this are evaluated by one CTA each (k_heavy_giant) instead of the batched kernel.
constexpr uint32_t HEAVY_GIANT_TOTAL = 1000;
constexpr uint32_t HEAVY_CHAIN_TOTAL = 32;    // heavy entries up to this total: chain evaluated once per (locus, total) group
// One 128-byte record per locus, laid out so that everything an entry of total n needs sits in one
// 32-byte sector:   [0..1] T(a=0..1 | n=1)  [2] E(1)  [3] -
//                   [4..6] T(a=0..2 | n=2)  [7] E(2)
//                   [8..11] T(a=0..3 | n=3) [12] E(3)  [13] lnB(alpha,beta)  [14] alpha  [15] beta
// T(a | n) = log_beta_binomial_pmf(alt=a, ref=n-a), E(n) = expected_log_likelihood(total=n).
constexpr int TAB_STRIDE = 16;
constexpr int TAB_DEN = 13, TAB_ALPHA = 14, TAB_BETA = 15;
CLX_HD int tab_idx_T(int a, int n) { return ((n - 1) << 2) + a; }
CLX_HD int tab_idx_E(int n) { return n == 1 ? 2 : (n == 2 ? 7 : 12); }
// index of ln C(n, a) in the 10-entry constant block for n <= 3
CLX_HD int lbc_idx(int a, int n) { return (n * (n + 1)) / 2 + a; }

// Staged table tile of the run-stream scoring kernel (tiled.cu): per locus TAB_SLOTS 16-byte slots {T(a | n), E(n)},
// slot p(a, n) = n(n+1)/2 - 1 + a — one LDS.128 per nonzero fetches both the log-pmf and the expected value of its
// total.  Every slot of a switched-off locus is {0.0, 0.0} (loci_used_per_cell is kept separately); slot TILE_SLOTS
// of a staged tile is the all-zero slot padding entries point at.  In the posterior pass the same slots carry
// {T_minority, T_majority} (pass 0) and {T_doublet, 0} (pass 1).
constexpr int TILE_L = 512;                          // loci per table tile
constexpr int CLS_STRIDE = 12;                        // per-locus entry-class counters: 9 light (a | n) classes,
constexpr int CLS_OTHER = 9;                          // [9] = entries without a memoised value (total 0 or heavy), 2 pads
constexpr int TAB_SLOTS = 9;                         // (a | n) pairs with n <= K_LIGHT
constexpr int TILE_SLOTS = TILE_L * TAB_SLOTS;       // 4608 slots = 73,728 bytes of shared memory per tile
CLX_HD int tab_slot(int a, int n) { return (n * (n + 1)) / 2 - 1 + a; }

struct Lanczos {
    // statrs GAMMA_DK / GAMMA_R / constants
    static constexpr double R = 10.900511;
    static constexpr double LN_2_SQRT_E_OVER_PI = 0.6207822376352452223455184457816472122518527279025978;
    static constexpr double LN_PI = 1.1447298858494001741434273513530587116472948129153;
    static constexpr double PI = 3.14159265358979323846264338327950288;
    static constexpr double E = 2.71828182845904523536028747135266250;
};

CLX_HD double lanczos_dk(int i) {
    switch (i) {
        case 0: return 2.48574089138753565546e-5;
        case 1: return 1.05142378581721974210;
        case 2: return -3.45687097222016235469;
        case 3: return 4.51227709466894823700;
        case 4: return -2.98285225323576655721;
        case 5: return 1.05639711577126713077;
        case 6: return -1.95428773191645869583e-1;
        case 7: return 1.70970543404441224307e-2;
        case 8: return -5.71926117404305781283e-4;
        case 9: return 4.63399473359905636708e-6;
        default: return -2.71994908488607703910e-9;
    }
}

CLX_HD double ln_gamma(double x) {
    if (x < 0.5) {  // reflection branch; unreachable on this path (all arguments >= 1) but kept for completeness
        double s = lanczos_dk(0);
#pragma unroll
        for (int i = 1; i < 11; ++i) s = s + lanczos_dk(i) / ((double)i - x);
        return Lanczos::LN_PI - log(sin(Lanczos::PI * x)) - log(s) - Lanczos::LN_2_SQRT_E_OVER_PI -
               (0.5 - x) * log((0.5 - x + Lanczos::R) / Lanczos::E);
    }
    double s = lanczos_dk(0);
#pragma unroll
    for (int i = 1; i < 11; ++i) s = s + lanczos_dk(i) / (x + (double)i - 1.0);
    return log(s) + Lanczos::LN_2_SQRT_E_OVER_PI + (x - 0.5) * log((x - 0.5 + Lanczos::R) / Lanczos::E);
}

// ln_gamma of an integer-valued argument through the memo lng[i] = ln_gamma((double)i): the same double, bit for
// bit (ln_gamma is a pure function of its argument); arguments beyond the memo are evaluated directly.
#ifdef __CUDACC__
// out of line: the Lanczos series is ~700 instructions and is the cold path here; inlined at every call site it
// made the lookup kernels instruction-fetch bound
static __device__ __noinline__ double ln_gamma_outofline(double x) { return ln_gamma(x); }
#else
inline double ln_gamma_outofline(double x) { return ln_gamma(x); }
#endif
CLX_HD double ln_gamma_int(double x, const double* __restrict__ lng, uint32_t lng_n) {
    if (x < (double)lng_n) return lng[(uint32_t)x];
#ifdef __CUDA_ARCH__
    return ln_gamma_outofline(x);
#else
    return ln_gamma(x);
#endif
}

// lnfact[i] = ln(i!) for i <= 170, computed on the host exactly like statrs (ln of the f64 product table).
CLX_HD double ln_factorial(uint64_t n, const double* __restrict__ lnfact) {
    if (n <= 170) return lnfact[n];
    return ln_gamma((double)n + 1.0);
}
CLX_HD double ln_binomial(uint64_t n, uint64_t k, const double* __restrict__ lnfact) {
    return ln_factorial(n, lnfact) - ln_factorial(k, lnfact) - ln_factorial(n - k, lnfact);
}

CLX_HD double log_beta_calc(double alpha, double beta) {
    const double lga = ln_gamma(alpha);
    const double lgb = ln_gamma(beta);
    const double lgab = ln_gamma(alpha + beta);
    return lga + lgb - lgab;
}
CLX_HD double log_beta_binomial_pmf(double alt, double ref, double alpha, double beta, double ln_coefficient) {
    const double ln_numerator = log_beta_calc(alt + alpha, ref + beta);
    const double ln_denominator = log_beta_calc(alpha, beta);
    return ln_coefficient + ln_numerator - ln_denominator;
}
// same value as log_beta_binomial_pmf when `den` == log_beta_calc(alpha, beta) (it is a pure function)
CLX_HD double log_beta_binomial_pmf_den(double alt, double ref, double alpha, double beta, double ln_coefficient,
                                        double den) {
    const double ln_numerator = log_beta_calc(alt + alpha, ref + beta);
    return ln_coefficient + ln_numerator - den;
}
// stats.rs:35-39: max + ln(exp(v1 - max) + exp(v2 - max)).  One of the two differences is exactly 0 and
// exp(0) == 1 exactly, so the sum is 1 + exp(other difference) bit for bit: (v1 - mx) + (v2 - mx) adds an exact
// zero to the other difference (and still propagates NaN), which saves one of the two exp evaluations.
CLX_HD double logsumexp2(double v1, double v2) {
    const double mx = fmax(v1, v2);
    const double sum = 1.0 + exp((v1 - mx) + (v2 - mx));
    return mx + log(sum);
}

// Per-locus memo record (see layout above).  lbc10[lbc_idx(a, n)] = ln C(n, a) for n <= 3.
CLX_HD void build_locus_table(double alpha, double beta, const double* __restrict__ lbc10, bool with_expected,
                              double* __restrict__ out) {
    const double den = log_beta_calc(alpha, beta);
    // ln_gamma(a + alpha) and ln_gamma(r + beta) are shared between the (a, r) combinations (same argument bits)
    double lga[K_LIGHT + 1], lgb[K_LIGHT + 1];
#pragma unroll
    for (int i = 0; i <= K_LIGHT; ++i) {
        lga[i] = ln_gamma((double)i + alpha);
        lgb[i] = ln_gamma((double)i + beta);
    }
    out[3] = 0.0;
#pragma unroll
    for (int n = 1; n <= K_LIGHT; ++n) {
        double pm[K_LIGHT + 1];
#pragma unroll
        for (int a = 0; a <= n; ++a) {
            const int r = n - a;
            // log_beta_calc(alt + alpha, ref + beta): third term is ln_gamma((alt+alpha) + (ref+beta))
            const double num = lga[a] + lgb[r] - ln_gamma(((double)a + alpha) + ((double)r + beta));
            pm[a] = lbc10[lbc_idx(a, n)] + num - den;
            out[tab_idx_T(a, n)] = pm[a];
        }
        double e = 0.0;
        if (with_expected) {
            e = 2.0 * pm[0];  // k = 0 -> alt = 0
#pragma unroll
            for (int k = 1; k <= n; ++k) e = logsumexp2(e, 2.0 * pm[k]);
        }
        out[tab_idx_E(n)] = e;
    }
    out[TAB_DEN] = den;
    out[TAB_ALPHA] = alpha;
    out[TAB_BETA] = beta;
}

// expected_log_likelihood for an arbitrary total (heavy entries), stats.rs:8-22.
CLX_HD double expected_ll_direct(uint64_t total, double alpha, double beta, double den,
                                 const double* __restrict__ lnfact) {
    double e = 0.0;
    for (uint64_t k = 0; k <= total; ++k) {
        const double c = ln_binomial(total, k, lnfact);
        const double ll = log_beta_binomial_pmf_den((double)k, (double)(total - k), alpha, beta, c, den);
        e = (k == 0) ? 2.0 * ll : logsumexp2(e, 2.0 * ll);
    }
    return e;
}

// order-preserving f64 -> u64 key (total order; -0.0 sorts just below +0.0, equal as doubles)
CLX_HD uint64_t f64_to_key(double x) {
    union { double d; uint64_t u; } v;
    v.d = x;
    return (v.u & 0x8000000000000000ull) ? ~v.u : (v.u | 0x8000000000000000ull);
}
CLX_HD double key_to_f64(uint64_t k) {
    union { double d; uint64_t u; } v;
    v.u = (k & 0x8000000000000000ull) ? (k & 0x7fffffffffffffffull) : ~k;
    return v.d;
}

}  // namespace clx

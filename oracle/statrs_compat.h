// oracle/statrs_compat.h — TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// CPU restatement of the statrs 0.16.0 routines that cellector calls.  statrs is a
// third-party crate that is NOT vendored under /root/reference (pinned at
// cellector/Cargo.lock:443-444, `statrs = "0.16"` cellector/Cargo.toml:13), so the
// published algorithm is restated here and anchored on the reference's call sites:
//   ln_gamma      <- stats.rs:49-51
//   ln_binomial   <- stats.rs:15,60 ; load_data.rs:163
//   Data::median / lower_quartile / upper_quartile <- main.rs:324-327, 442-443
//   Binomial::pmf <- main.rs:92-97
//
// PARITY UNPINNED: the reference ships no tests, golden vectors or runnable binary
// (no cargo/rustc here, cellector_linux is a missing blob), so these routines are
// pinned only by known-answer tests we wrote ourselves (mpmath / scipy / hand-worked
// vectors, tests/test_oracle_kat.py).
//
// Must be compiled with -ffp-contract=off: Rust never contracts a*b+c into an FMA.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <limits>
#include <vector>

namespace statrs {

// statrs::function::gamma — Lanczos approximation, g = 10.900511, 11 coefficients.
static const double GAMMA_R = 10.900511;
static const double GAMMA_DK[11] = {
    2.48574089138753565546e-5,  1.05142378581721974210,    -3.45687097222016235469,
    4.51227709466894823700,     -2.98285225323576655721,   1.05639711577126713077,
    -1.95428773191645869583e-1, 1.70970543404441224307e-2, -5.71926117404305781283e-4,
    4.63399473359905636708e-6,  -2.71994908488607703910e-9};
static const double LN_2_SQRT_E_OVER_PI = 0.6207822376352452223455184457816472122518527279025978;
static const double LN_PI = 1.1447298858494001741434273513530587116472948129153;
static const double K_PI = 3.14159265358979323846264338327950288;
static const double K_E = 2.71828182845904523536028747135266250;

inline double ln_gamma(double x) {
    if (x < 0.5) {
        double s = GAMMA_DK[0];
        for (int i = 1; i < 11; ++i) s = s + GAMMA_DK[i] / ((double)i - x);
        return LN_PI - std::log(std::sin(K_PI * x)) - std::log(s) - LN_2_SQRT_E_OVER_PI -
               (0.5 - x) * std::log((0.5 - x + GAMMA_R) / K_E);
    }
    double s = GAMMA_DK[0];
    for (int i = 1; i < 11; ++i) s = s + GAMMA_DK[i] / (x + (double)i - 1.0);
    return std::log(s) + LN_2_SQRT_E_OVER_PI + (x - 0.5) * std::log((x - 0.5 + GAMMA_R) / K_E);
}

// statrs::function::factorial — FCACHE[i] = FCACHE[i-1] * i in f64 for i <= 170.
struct FactorialCache {
    double f[171];
    FactorialCache() {
        f[0] = 1.0;
        for (int i = 1; i <= 170; ++i) f[i] = f[i - 1] * (double)i;
    }
};
inline const FactorialCache& fcache() {
    static FactorialCache c;
    return c;
}
inline double ln_factorial(uint64_t n) {
    if (n <= 170) return std::log(fcache().f[n]);
    return ln_gamma((double)n + 1.0);
}
inline double ln_binomial(uint64_t n, uint64_t k) {
    if (k > n) return -std::numeric_limits<double>::infinity();
    return ln_factorial(n) - ln_factorial(k) - ln_factorial(n - k);
}

// statrs::statistics::Data — order statistics by in-place selection.
// k-th smallest (0-based); NaN-free input assumed (argsort in main.rs:504 panics on NaN anyway).
inline double select_inplace(std::vector<double>& v, size_t k) {
    if (v.empty()) return std::numeric_limits<double>::quiet_NaN();
    if (k >= v.size()) k = v.size() - 1;
    std::nth_element(v.begin(), v.begin() + (ptrdiff_t)k, v.end());
    return v[k];
}
// R-8 quantile as statrs implements it:  h = (N + 1/3) tau + 1/3.
inline double quantile(std::vector<double>& v, double tau) {
    if (!(tau >= 0.0 && tau <= 1.0) || v.empty()) return std::numeric_limits<double>::quiet_NaN();
    const double n = (double)v.size();
    const double h = (n + 1.0 / 3.0) * tau + 1.0 / 3.0;
    const int64_t hf = (int64_t)h;
    if (hf <= 0 || tau == 0.0) return *std::min_element(v.begin(), v.end());
    if (hf >= (int64_t)v.size() || tau == 1.0) return *std::max_element(v.begin(), v.end());
    const double a = select_inplace(v, (size_t)hf - 1);
    const double b = select_inplace(v, std::min((size_t)hf, v.size() - 1));
    return a + (h - (double)hf) * (b - a);
}
inline double lower_quartile(std::vector<double>& v) { return quantile(v, 0.25); }
inline double upper_quartile(std::vector<double>& v) { return quantile(v, 0.75); }
inline double median(std::vector<double>& v) {
    if (v.empty()) return std::numeric_limits<double>::quiet_NaN();
    const size_t k = v.size() / 2;
    if (v.size() % 2 == 0) return (select_inplace(v, k - 1) + select_inplace(v, k)) / 2.0;
    return select_inplace(v, k);
}

// statrs::distribution::Binomial::pmf (VCF path only, main.rs:92-97).
inline double binomial_pmf(double p, uint64_t n, uint64_t k) {
    if (k > n) return 0.0;
    if (p == 0.0) return k == 0 ? 1.0 : 0.0;
    if (p == 1.0) return k == n ? 1.0 : 0.0;
    return std::exp(ln_binomial(n, k) + (double)k * std::log(p) + (double)(n - k) * std::log(1.0 - p));
}

}  // namespace statrs

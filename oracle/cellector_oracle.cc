// oracle/cellector_oracle.cc — TEST INFRASTRUCTURE ONLY.
//
// CPU restatement of the reference's hot path (wheaton5/cellector, cellector/src/{main,load_data,stats}.rs)
// plus the statrs 0.16.0 numerics it calls (statrs_compat.h).  It is the checker for the CUDA path and the
// timed CPU baseline; it is never linked into, imported by or called from the product (libcellector_cuda.so,
// the cellector CLI).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs may load it.
//
// PARITY UNPINNED: the reference has no tests, fixtures or golden vectors for this path and cannot be
// compiled or run here (no cargo/rustc; the prebuilt cellector_linux is a missing blob), so this
// restatement is pinned only by our own known-answer tests (tests/test_oracle_kat.py: mpmath/scipy values,
// hand-worked quantiles and a hand-worked toy matrix) and the committed goldens generated from it.
//
// Faithful by construction: same passes, same data structures (array of per-cell row vectors of 48-byte
// entries, 88-byte PMF records materialised per scoring call), same floating-point expression order, single
// thread — so that its run time is the reference's cost.  Build with -O2 -ffp-contract=off.
//
// Every function cites the reference lines it follows.
#include <array>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <chrono>
#include <fstream>
#include <numeric>
#include <string>
#include <vector>

#include "statrs_compat.h"
#include "../include/cellector_cuda.h"  // shares clx_iter_stats / clx_locus_arrays layouts

namespace {

// ---------------------------------------------------------------- stats.rs
// stats.rs:35-39
inline double logsumexp(double v1, double v2) {
    const double mx = std::fmax(v1, v2);
    const double sum = std::exp(v1 - mx) + std::exp(v2 - mx);
    return mx + std::log(sum);
}
// stats.rs:48-53
inline double log_beta_calc(double alpha, double beta) {
    const double lga = statrs::ln_gamma(alpha);
    const double lgb = statrs::ln_gamma(beta);
    const double lgab = statrs::ln_gamma(alpha + beta);
    return lga + lgb - lgab;
}
// stats.rs:41-46
inline double log_beta_binomial_pmf(double alt, double ref, double alpha, double beta, double ln_coefficient) {
    const double ln_numerator = log_beta_calc(alt + alpha, ref + beta);
    const double ln_denominator = log_beta_calc(alpha, beta);
    return ln_coefficient + ln_numerator - ln_denominator;
}
// stats.rs:55-65
std::vector<std::vector<double>> precompute_log_binomial_coefficients(size_t max_n) {
    std::vector<std::vector<double>> t;
    for (size_t n = 0; n <= max_n; ++n) {
        std::vector<double> row;
        for (size_t k = 0; k <= n; ++k) row.push_back(statrs::ln_binomial(n, k));
        t.push_back(row);
    }
    return t;
}
struct LogBetaBinomialExpectation {
    double expected_log_likelihood, expected_log_variance;
};
// stats.rs:8-33
LogBetaBinomialExpectation expected_log_beta_binomial_pmf(size_t total, double alpha, double beta,
                                                          const std::vector<std::vector<double>>& lbc) {
    std::vector<double> lls;
    for (size_t k = 0; k <= total; ++k) {
        double c;
        if (total < lbc.size()) c = lbc[total][k];
        else c = statrs::ln_binomial(total, k);
        lls.push_back(log_beta_binomial_pmf((double)k, (double)(total - k), alpha, beta, c));
    }
    double expectation = 2.0 * lls[0];
    for (size_t k = 1; k <= total; ++k) expectation = logsumexp(expectation, 2.0 * lls[k]);
    double variance = 0.0;
    for (double ll : lls) {
        const double d = ll - expectation;
        variance += std::exp(ll) * (d * d);
    }
    return {expectation, variance};
}

// ---------------------------------------------------------------- load_data.rs types
struct CellLocusData {  // load_data.rs:13-20 (48 bytes, like the Rust struct)
    size_t locus_index, locus_id;
    double log_binomial_coefficient, alt_count, ref_count;
    size_t total;
};
struct CellData {  // load_data.rs:22-27 (barcode / assignment strings live in the host CLI, not on this path)
    size_t cell_id;
    std::vector<CellLocusData> cell_loci_data;
};
struct AlphaBeta { double alpha, beta; };  // main.rs:593-596
struct PMFData {                            // main.rs:527-539 (88 bytes)
    size_t cell_id, locus_index, locus;
    bool excluded;
    double log_pmf;
    size_t alt_count, ref_count;
    double alpha, beta, expected_log_pmf, expected_log_variance;
};
struct CellLogLikelihoodData {  // main.rs:508-514
    std::vector<double> log_likelihoods, loci_used_per_cell, expected_log_likelihoods, expected_log_variances;
    std::vector<PMFData> all_pmfs;
};
struct AlleleCount { size_t ref_count, alt_count; };  // main.rs:500
struct LocusLogLikelihoodData {                       // main.rs:516-525
    std::vector<double> contrib_min, contrib_maj;
    std::vector<size_t> cells_min, cells_maj;
    std::vector<AlleleCount> alleles_min, alleles_maj;
    std::vector<double> expected_min, expected_maj;
};

}  // namespace

struct orc_ctx {
    std::string err;
    bool loaded = false;
    // load_cell_data outputs (load_data.rs:134-181)
    std::vector<bool> loci_used;
    std::vector<size_t> locus_ids;
    std::vector<CellData> cell_data;
    std::vector<std::array<double, 2>> locus_counts;  // [ref, alt]
    std::vector<std::vector<double>> lbc_table;
    size_t n_loci0 = 0;
    uint64_t nnz_used = 0;
    // retained COO for load_mtx_final (load_data.rs:109-132 re-reads the files; we keep the arrays)
    std::vector<uint32_t> coo_locus, coo_cell, coo_ref, coo_alt;
    // EM state (main.rs:36-46)
    std::vector<uint8_t> excluded;  // the HashSet<usize> as a membership vector
    size_t iteration = 0;
    // last-iteration outputs
    CellLogLikelihoodData cll;
    std::vector<double> normalized;
    LocusLogLikelihoodData locus;
    // timers (seconds) of the last iteration's phases
    double t_score = 0, t_locus = 0, t_total = 0;
    bool faithful_posteriors = false;  // also pay for the discarded expected-LL in orc_posteriors (timing only)
};

namespace {

// main.rs:598-611
std::vector<AlphaBeta> init_alpha_betas(const std::vector<std::array<double, 2>>& locus_counts,
                                        const std::vector<uint8_t>& excluded, const std::vector<CellData>& cells) {
    std::vector<AlphaBeta> ab;
    ab.reserve(locus_counts.size());
    for (const auto& c : locus_counts) ab.push_back({c[1] + 1.0, c[0] + 1.0});
    for (size_t cell = 0; cell < cells.size(); ++cell) {
        if (!excluded[cell]) continue;
        for (const auto& e : cells[cell].cell_loci_data) {
            ab[e.locus_index].alpha -= e.alt_count;
            ab[e.locus_index].beta -= e.ref_count;
        }
    }
    return ab;
}

// main.rs:541-591
CellLogLikelihoodData get_cell_log_likelihoods(const std::vector<bool>& loci_used, const std::vector<CellData>& cells,
                                               const std::vector<AlphaBeta>& ab, const std::vector<uint8_t>& excluded,
                                               const std::vector<std::vector<double>>& lbc, bool faithful) {
    CellLogLikelihoodData out;
    for (size_t cell_id = 0; cell_id < cells.size(); ++cell_id) {
        double ll = 0.0, ell = 0.0, evar = 0.0, n_used = 0.0;
        const bool excl = excluded[cell_id] != 0;
        for (const auto& locus : cells[cell_id].cell_loci_data) {
            if (!loci_used[locus.locus_index]) continue;
            const AlphaBeta& p = ab[locus.locus_index];
            const double log_pmf =
                log_beta_binomial_pmf(locus.alt_count, locus.ref_count, p.alpha, p.beta, locus.log_binomial_coefficient);
            ll += log_pmf;
            LogBetaBinomialExpectation e{0.0, 0.0};
            if (faithful) e = expected_log_beta_binomial_pmf(locus.total, p.alpha, p.beta, lbc);
            ell += e.expected_log_likelihood;
            evar += e.expected_log_variance;
            out.all_pmfs.push_back(PMFData{cell_id, locus.locus_index, locus.locus_id, excl, log_pmf,
                                           (size_t)locus.alt_count, (size_t)locus.ref_count, p.alpha, p.beta,
                                           e.expected_log_likelihood, e.expected_log_variance});
            n_used += 1.0;
        }
        out.log_likelihoods.push_back(ll);
        out.loci_used_per_cell.push_back(n_used);
        out.expected_log_likelihoods.push_back(ell);
        out.expected_log_variances.push_back(evar);
    }
    return out;
}

// main.rs:368-420
LocusLogLikelihoodData get_locus_log_likelihoods(const std::vector<PMFData>& all_pmfs, size_t n_cells, size_t n_loci,
                                                 const std::vector<uint8_t>& excluded) {
    LocusLogLikelihoodData d;
    d.contrib_min.assign(n_loci, 0.0);
    d.contrib_maj.assign(n_loci, 0.0);
    d.cells_min.assign(n_loci, 0);
    d.cells_maj.assign(n_loci, 0);
    d.expected_min.assign(n_loci, 0.0);
    d.expected_maj.assign(n_loci, 0.0);
    d.alleles_min.assign(n_loci, AlleleCount{0, 0});
    d.alleles_maj.assign(n_loci, AlleleCount{0, 0});
    (void)n_cells;
    for (const auto& p : all_pmfs) {
        const double log_pmf = p.log_pmf;
        const double expected_log_pmf = p.log_pmf;  // sic, main.rs:394
        if (excluded[p.cell_id]) {
            d.contrib_min[p.locus_index] += log_pmf;
            d.cells_min[p.locus_index] += 1;
            d.expected_min[p.locus_index] += expected_log_pmf;
            d.alleles_min[p.locus_index].ref_count += p.ref_count;
            d.alleles_min[p.locus_index].alt_count += p.alt_count;
        } else {
            d.contrib_maj[p.locus_index] += log_pmf;
            d.cells_maj[p.locus_index] += 1;
            d.expected_maj[p.locus_index] += expected_log_pmf;
            d.alleles_maj[p.locus_index].ref_count += p.ref_count;
            d.alleles_maj[p.locus_index].alt_count += p.alt_count;
        }
    }
    return d;
}

// The filter half of locus_filter_and_output_locus_data, main.rs:428-451.  Returns #loci newly switched off.
size_t locus_filter(std::vector<bool>& used_loci, const LocusLogLikelihoodData& d) {
    size_t n_filtered = 0;
    const double threshold = -80.0;  // main.rs:444
    for (size_t i = 0; i < used_loci.size(); ++i) {
        double per_cell_min = 0.0;
        if (d.cells_min[i] != 0) per_cell_min = d.contrib_min[i] / (double)d.cells_min[i];
        if (per_cell_min < threshold) {
            if (used_loci[i]) ++n_filtered;
            used_loci[i] = false;
        }
    }
    return n_filtered;
}

double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

}  // namespace

extern "C" {

// ---- scalar numerics exported for known-answer tests
double orc_ln_gamma(double x) { return statrs::ln_gamma(x); }
double orc_ln_factorial(uint64_t n) { return statrs::ln_factorial(n); }
double orc_ln_binomial(uint64_t n, uint64_t k) { return statrs::ln_binomial(n, k); }
double orc_logsumexp(double a, double b) { return logsumexp(a, b); }
double orc_log_beta_calc(double a, double b) { return log_beta_calc(a, b); }
double orc_log_beta_binomial_pmf(double alt, double ref, double alpha, double beta, double lbc) {
    return log_beta_binomial_pmf(alt, ref, alpha, beta, lbc);
}
void orc_expected_log_beta_binomial_pmf(uint64_t total, double alpha, double beta, double* e, double* var) {
    static const std::vector<std::vector<double>> lbc = precompute_log_binomial_coefficients(100);
    const auto r = expected_log_beta_binomial_pmf(total, alpha, beta, lbc);
    if (e) *e = r.expected_log_likelihood;
    if (var) *var = r.expected_log_variance;
}
double orc_quantile(const double* x, uint64_t n, double tau) {
    std::vector<double> v(x, x + n);
    return statrs::quantile(v, tau);
}
double orc_median(const double* x, uint64_t n) {
    std::vector<double> v(x, x + n);
    return statrs::median(v);
}
double orc_binomial_pmf(double p, uint64_t n, uint64_t k) { return statrs::binomial_pmf(p, n, k); }

// ---- context API, mirroring include/cellector_cuda.h one to one
orc_ctx* orc_create() { return new orc_ctx(); }
void orc_destroy(orc_ctx* c) { delete c; }
const char* orc_last_error(const orc_ctx* c) { return c ? c->err.c_str() : ""; }

// load_cell_data (load_data.rs:134-181) over COO arrays in file order; pass 1 = get_loci_used (:254-280).
int orc_load_coo(orc_ctx* c, uint64_t nnz0, uint32_t n_loci0, uint32_t n_cells, const uint32_t* locus,
                 const uint32_t* cell, const uint32_t* ref, const uint32_t* alt, uint32_t min_ref, uint32_t min_alt,
                 uint32_t* n_loci_used, uint64_t* nnz_used) {
    // PASS 1 — get_loci_used
    std::vector<std::array<size_t, 2>> counts(n_loci0, {0, 0});
    for (uint64_t i = 0; i < nnz0; ++i) {
        if (locus[i] >= n_loci0 || cell[i] >= n_cells) { c->err = "coordinate out of range"; return CLX_ERR_ARG; }
        if (ref[i] > 0) counts[locus[i]][0] += 1;
        if (alt[i] > 0) counts[locus[i]][1] += 1;
    }
    std::vector<bool> used0(n_loci0, false);
    std::vector<size_t> to_used(n_loci0, SIZE_MAX);
    size_t num_used = 0;
    for (size_t l = 0; l < n_loci0; ++l) {
        if (counts[l][0] >= min_ref && counts[l][1] >= min_alt) {
            used0[l] = true;
            to_used[l] = num_used++;
        }
    }
    // PASS 2
    c->n_loci0 = n_loci0;
    c->cell_data.clear();
    c->cell_data.resize(n_cells);
    for (size_t i = 0; i < n_cells; ++i) c->cell_data[i].cell_id = i;
    c->locus_counts.assign(num_used, {0.0, 0.0});
    c->locus_ids.clear();
    for (size_t l = 0; l < n_loci0; ++l) if (used0[l]) c->locus_ids.push_back(l);
    const size_t max_n = 100;
    c->lbc_table = precompute_log_binomial_coefficients(max_n);
    c->nnz_used = 0;
    for (uint64_t i = 0; i < nnz0; ++i) {
        if (!used0[locus[i]]) continue;
        const size_t r = ref[i], a = alt[i], total = r + a;
        const size_t ui = to_used[locus[i]];
        c->locus_counts[ui][0] += (double)r;
        c->locus_counts[ui][1] += (double)a;
        double coef;
        if (total <= max_n) coef = c->lbc_table[total][a];
        else coef = statrs::ln_binomial(total, a);
        c->cell_data[cell[i]].cell_loci_data.push_back(CellLocusData{ui, locus[i], coef, (double)a, (double)r, total});
        c->nnz_used++;
    }
    c->loci_used.assign(num_used, true);
    c->coo_locus.assign(locus, locus + nnz0);
    c->coo_cell.assign(cell, cell + nnz0);
    c->coo_ref.assign(ref, ref + nnz0);
    c->coo_alt.assign(alt, alt + nnz0);
    c->excluded.assign(n_cells, 0);
    c->iteration = 0;
    c->loaded = true;
    if (n_loci_used) *n_loci_used = (uint32_t)num_used;
    if (nnz_used) *nnz_used = c->nnz_used;
    return CLX_OK;
}

// Text front end: read_mtx_lines / consume_mtx_header (load_data.rs:190-223) for uncompressed files.
// Coordinates come from the alt file, the value is token 3 of each file, exactly three header lines.
int orc_read_mtx_pair(const char* alt_path, const char* ref_path, uint32_t* n_loci0, uint32_t* n_cells,
                      uint64_t* nnz, uint32_t* locus, uint32_t* cell, uint32_t* ref, uint32_t* alt, uint64_t cap) {
    std::ifstream fa(alt_path), fr(ref_path);
    if (!fa || !fr) return CLX_ERR_ARG;
    std::string la, lr;
    unsigned long long tl = 0, tc = 0;
    for (int x = 0; x < 3; ++x) {
        std::getline(fa, la);
        std::getline(fr, lr);
        if (x == 2) {
            if (sscanf(lr.c_str(), "%llu %llu", &tl, &tc) != 2) return CLX_ERR_ARG;
        }
    }
    *n_loci0 = (uint32_t)tl;
    *n_cells = (uint32_t)tc;
    uint64_t n = 0;
    while (std::getline(fa, la) && std::getline(fr, lr)) {
        unsigned long long l, ce, av, l2, c2, rv;
        if (sscanf(la.c_str(), "%llu %llu %llu", &l, &ce, &av) != 3) return CLX_ERR_ARG;
        if (sscanf(lr.c_str(), "%llu %llu %llu", &l2, &c2, &rv) != 3) return CLX_ERR_ARG;
        if (locus && n < cap) {
            locus[n] = (uint32_t)(l - 1);
            cell[n] = (uint32_t)(ce - 1);
            alt[n] = (uint32_t)av;
            ref[n] = (uint32_t)rv;
        }
        ++n;
    }
    *nnz = n;
    return CLX_OK;
}

void orc_set_faithful_posteriors(orc_ctx* c, int on) { c->faithful_posteriors = on != 0; }

int orc_get_locus_ids(orc_ctx* c, uint32_t* ids) {
    for (size_t i = 0; i < c->locus_ids.size(); ++i) ids[i] = (uint32_t)c->locus_ids[i];
    return CLX_OK;
}

int orc_reset(orc_ctx* c) {
    std::fill(c->excluded.begin(), c->excluded.end(), 0);
    c->loci_used.assign(c->loci_used.size(), true);
    c->iteration = 0;
    return CLX_OK;
}

// compute_new_excluded, main.rs:308-347.  `faithful` != 0 also evaluates the expected-log-likelihood
// diagnostic per entry like the reference does (stats.rs:8-33); 0 skips it (lean mode for big parity cases).
int orc_em_iteration_ex(orc_ctx* c, double iqr_multiple, clx_iter_stats* out, int faithful) {
    if (!c->loaded) { c->err = "not loaded"; return CLX_ERR_STATE; }
    const double t0 = now_s();
    const size_t n_cells = c->cell_data.size();
    const std::vector<AlphaBeta> ab = init_alpha_betas(c->locus_counts, c->excluded, c->cell_data);
    const double t1 = now_s();
    c->cll = get_cell_log_likelihoods(c->loci_used, c->cell_data, ab, c->excluded, c->lbc_table, faithful != 0);
    const double t2 = now_s();
    std::vector<double>& norm = c->normalized;
    norm.clear();
    for (size_t i = 0; i < n_cells; ++i) {
        if (c->cll.loci_used_per_cell[i] > 0.0) norm.push_back(c->cll.log_likelihoods[i] / c->cll.loci_used_per_cell[i]);
        else norm.push_back(0.0);
    }
    std::vector<double> tmp(norm);
    const double median = statrs::median(tmp);
    const double q1 = statrs::lower_quartile(tmp);
    const double q3 = statrs::upper_quartile(tmp);
    const double iqr = q3 - q1;
    const double threshold = q1 - iqr_multiple * iqr;
    std::vector<uint8_t> new_excluded(n_cells, 0);
    uint64_t n_new = 0, n_rescued = 0, n_excl = 0;
    for (size_t i = 0; i < n_cells; ++i) {
        if (norm[i] < threshold) new_excluded[i] = 1;
        if (new_excluded[i] && !c->excluded[i]) ++n_new;
        if (!new_excluded[i] && c->excluded[i]) ++n_rescued;
        n_excl += new_excluded[i];
    }
    const bool any_change = n_new > 0 || n_rescued > 0;
    const double t3 = now_s();
    c->locus = get_locus_log_likelihoods(c->cll.all_pmfs, n_cells, c->loci_used.size(), new_excluded);
    const size_t n_filtered = locus_filter(c->loci_used, c->locus);
    const double t4 = now_s();
    c->excluded.swap(new_excluded);
    if (out) {
        memset(out, 0, sizeof(*out));
        out->any_change = any_change ? 1 : 0;
        out->iteration = (uint32_t)c->iteration;
        out->n_new_excluded = n_new;
        out->n_rescued = n_rescued;
        out->n_excluded = n_excl;
        out->median = median;
        out->q1 = q1;
        out->q3 = q3;
        out->iqr = iqr;
        out->threshold = threshold;
        out->n_loci_filtered = n_filtered;
        size_t in_use = 0;
        for (bool b : c->loci_used) in_use += b;
        out->n_loci_in_use = in_use;
        out->device_ms = (float)((t4 - t0) * 1e3);
    }
    c->t_score = t2 - t1;
    c->t_locus = t4 - t3;
    c->t_total = t4 - t0;
    c->iteration++;
    return CLX_OK;
}
int orc_em_iteration(orc_ctx* c, double iqr_multiple, clx_iter_stats* out) {
    return orc_em_iteration_ex(c, iqr_multiple, out, 1);
}
// seconds spent in [score, locus pass, whole iteration] of the last iteration
void orc_last_timings(orc_ctx* c, double* t3) {
    t3[0] = c->t_score; t3[1] = c->t_locus; t3[2] = c->t_total;
}

int orc_get_cell_vectors(orc_ctx* c, double* ll, double* expected_ll, uint32_t* n_loci, double* norm, uint8_t* excluded) {
    const size_t n = c->cell_data.size();
    if (c->cll.log_likelihoods.size() != n) { c->err = "no iteration run yet"; return CLX_ERR_STATE; }
    for (size_t i = 0; i < n; ++i) {
        if (ll) ll[i] = c->cll.log_likelihoods[i];
        if (expected_ll) expected_ll[i] = c->cll.expected_log_likelihoods[i];
        if (n_loci) n_loci[i] = (uint32_t)c->cll.loci_used_per_cell[i];
        if (norm) norm[i] = c->normalized[i];
        if (excluded) excluded[i] = c->excluded[i];
    }
    return CLX_OK;
}

int orc_get_locus_vectors(orc_ctx* c, const clx_locus_arrays* o) {
    const size_t n = c->loci_used.size();
    if (c->locus.contrib_min.size() != n) { c->err = "no iteration run yet"; return CLX_ERR_STATE; }
    for (size_t i = 0; i < n; ++i) {
        if (o->ll_minority) o->ll_minority[i] = c->locus.contrib_min[i];
        if (o->ll_majority) o->ll_majority[i] = c->locus.contrib_maj[i];
        if (o->cells_minority) o->cells_minority[i] = (uint32_t)c->locus.cells_min[i];
        if (o->cells_majority) o->cells_majority[i] = (uint32_t)c->locus.cells_maj[i];
        if (o->ref_minority) o->ref_minority[i] = c->locus.alleles_min[i].ref_count;
        if (o->alt_minority) o->alt_minority[i] = c->locus.alleles_min[i].alt_count;
        if (o->ref_majority) o->ref_majority[i] = c->locus.alleles_maj[i].ref_count;
        if (o->alt_majority) o->alt_majority[i] = c->locus.alleles_maj[i].alt_count;
        if (o->loci_used) o->loci_used[i] = c->loci_used[i] ? 1 : 0;
    }
    return CLX_OK;
}

// calculate_posteriors (main.rs:228-280) + assignment rule of output_final_assignments (main.rs:141-170).
int orc_posteriors(orc_ctx* c, double posterior_threshold, uint32_t min_loci_used, double* posterior,
                   double* doublet_posterior, double* ll_major, double* ll_minor, uint8_t* label, uint8_t* anomaly,
                   uint8_t* qual) {
    if (!c->loaded) { c->err = "not loaded"; return CLX_ERR_STATE; }
    const size_t n_cells = c->cell_data.size();
    const size_t n_loci = c->loci_used.size();
    std::vector<uint8_t> included(n_cells, 0);
    size_t n_excluded = 0;
    for (size_t i = 0; i < n_cells; ++i) {
        included[i] = c->excluded[i] ? 0 : 1;
        n_excluded += c->excluded[i] ? 1 : 0;
    }
    std::vector<AlphaBeta> ab_major = init_alpha_betas(c->locus_counts, c->excluded, c->cell_data);
    double minority_fraction = ((double)n_excluded + 1.0) / ((double)n_cells + 1.0);
    const std::vector<AlphaBeta> ab_minor = init_alpha_betas(c->locus_counts, included, c->cell_data);
    std::vector<AlphaBeta> ab_doublet;
    for (size_t l = 0; l < n_loci; ++l) {
        const double alpha = (ab_major[l].alpha - 1.0) * minority_fraction + (ab_minor[l].alpha - 1.0) + 1.0;
        const double beta = (ab_major[l].beta - 1.0) * minority_fraction + (ab_minor[l].beta - 1.0) + 1.0;
        ab_doublet.push_back({alpha, beta});
    }
    minority_fraction = std::fmax(minority_fraction, 0.01);
    for (size_t l = 0; l < n_loci; ++l) {
        ab_major[l].alpha = (ab_major[l].alpha - 1.0) * minority_fraction + 1.0;
        ab_major[l].beta = (ab_major[l].beta - 1.0) * minority_fraction + 1.0;
    }
    // get_loci_used_for_posterior_calc (main.rs:282-306): both branches push true.
    const std::vector<bool> used_post(n_loci, true);
    // The reference evaluates the expected-log-likelihood here too and throws it away (main.rs:256-258);
    // it is not observable, so it is only evaluated when the context is in faithful-cost mode.
    const bool f = c->faithful_posteriors;
    const CellLogLikelihoodData lmin = get_cell_log_likelihoods(used_post, c->cell_data, ab_minor, c->excluded, c->lbc_table, f);
    const CellLogLikelihoodData lmaj = get_cell_log_likelihoods(used_post, c->cell_data, ab_major, included, c->lbc_table, f);
    const CellLogLikelihoodData ldbl = get_cell_log_likelihoods(used_post, c->cell_data, ab_doublet, included, c->lbc_table, f);
    const double log_prior_doublet = std::log((double)n_cells / 1000.0 / 100.0 * std::fmax(minority_fraction, 0.1));
    const double log_prior_minority = std::log(minority_fraction);
    const double log_prior_majority = std::log(1.0 - minority_fraction);
    for (size_t i = 0; i < n_cells; ++i) {
        const double log_num = log_prior_minority + lmin.log_likelihoods[i];
        double log_den = logsumexp(log_num, log_prior_majority + lmaj.log_likelihoods[i]);
        const double log_dbl_num = log_prior_doublet + ldbl.log_likelihoods[i];
        log_den = logsumexp(log_den, log_dbl_num);
        const double post = std::exp(log_num - log_den);
        const double dpost = std::exp(log_dbl_num - log_den);
        if (posterior) posterior[i] = post;
        if (doublet_posterior) doublet_posterior[i] = dpost;
        if (ll_major) ll_major[i] = lmaj.log_likelihoods[i];
        if (ll_minor) ll_minor[i] = lmin.log_likelihoods[i];
        // main.rs:143-153
        uint8_t lab = 3;
        if (post > posterior_threshold) lab = 0;
        else if (1.0 - post > posterior_threshold) lab = 1;
        if (dpost > 0.5) lab = 2;
        if (c->cell_data[i].cell_loci_data.size() < min_loci_used) lab = 3;
        if (label) label[i] = lab;
        if (anomaly) anomaly[i] = c->excluded[i] ? 0 : 1;  // main.rs:160-163
        if (qual) {
            // main.rs:167-169; Rust f64::max/min ignore a NaN operand, `as usize` saturates.
            const double p = std::fmax(post, 1.0 - post);
            double q = -10.0 * std::log10(1.0 - p);
            q = std::fmin(q, 255.0);
            qual[i] = (q != q || q < 0.0) ? 0 : (uint8_t)q;
        }
    }
    return CLX_OK;
}

// load_mtx_final (load_data.rs:109-132) over ALL loci, from the retained COO arrays.
int orc_final_locus_counts(orc_ctx* c, uint64_t* minority_ref_alt, uint64_t* majority_ref_alt) {
    for (size_t i = 0; i < 2 * c->n_loci0; ++i) { minority_ref_alt[i] = 0; majority_ref_alt[i] = 0; }
    for (size_t i = 0; i < c->coo_locus.size(); ++i) {
        uint64_t* dst = c->excluded[c->coo_cell[i]] ? minority_ref_alt : majority_ref_alt;
        dst[2 * (size_t)c->coo_locus[i] + 0] += c->coo_ref[i];
        dst[2 * (size_t)c->coo_locus[i] + 1] += c->coo_alt[i];
    }
    return CLX_OK;
}

}  // extern "C"

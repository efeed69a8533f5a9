// oracle/souporcell_oracle.cc — TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// CPU restatement of souporcell's k-cluster EM (the tool cellector_pipeline.py:233-240 runs right after cellector):
//   load_cell_data      souporcell/souporcell/src/main.rs:390-468   (locus filter with the UMI minima, one entry per
//                                                                   (locus, cell): HashMap insert = the LAST line wins,
//                                                                   zero-total entries skipped, rows in ascending locus)
//   EM                  main.rs:109-186   (9 temperature steps, loss-change limit 0.01 * cells, at most 1000 iterations)
//   binomial_loss       main.rs:199-213
//   log_sum_exp         main.rs:215-219
//   normalize_in_log_with_temp  main.rs:230-242
//   update_centers_average      main.rs:273-280
//   update_final / reset_sums_denoms   main.rs:244-262
//   souporcell_main's best-of-restarts and argmax assignment   main.rs:62-106
// The reference draws its initial centers from rand 0.7.2's StdRng (main.rs:354-364; Cargo.lock:319-321); that generator is
// a third-party crate absent from /root/reference, so the centers are an INPUT here (the caller's own generator).
// ln_binomial is statrs 0.11.0's here (souporcell/souporcell/Cargo.lock:514-515), not cellector's 0.16.0: the same published
// algorithm (ln of the cached f64 factorials up to 170, Lanczos ln_gamma beyond), restated in statrs_compat.h.
//
// Two arithmetic modes of the same statements:
//   faithful (float)   every variable f32 like the reference, sums in the reference's serial order
//                      (a row's loci ascending, cells ascending);
//   precise  (double)  the same statements with f64 variables (what the CUDA path computes); the f32 roundings that are
//                      part of the INPUT stay: ln_binomial(..) as f32 per entry, total_alleles as f32, the temperature
//                      in f32, the 0.01 / 0.99 clamps and the initial centers as f32 values.
//
// PARITY UNPINNED: the reference ships no tests, vectors or runnable binary for this path either; the restatement is
// pinned by an independent numpy statement (tests/test_soup_oracle.py).
// Must be compiled with -ffp-contract=off.
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <string>
#include <unordered_map>
#include <vector>

#include "statrs_compat.h"

namespace {

struct SoupCell {
    std::vector<uint32_t> loci, ref, alt;
    std::vector<float> lbc;   // ln_binomial(ref + alt, alt) as f32   main.rs:446-447
    float total_alleles = 0;  // main.rs:448
};

struct sorc_ctx {
    std::string err;
    uint32_t n_cells = 0, n_loci0 = 0, n_loci = 0;
    std::vector<uint32_t> locus_ids;
    std::vector<SoupCell> cells;
    uint64_t n_entries = 0;
};

template <typename F>
F log_sum_exp(const std::vector<F>& p) {  // main.rs:215-219
    F mx = -std::numeric_limits<F>::infinity();
    for (F x : p) mx = std::fmax(mx, x);
    F s = 0;
    for (F x : p) s += std::exp(x - mx);
    return mx + std::log(s);
}

template <typename F>
int em(const sorc_ctx& c, uint32_t k, const float* init, double* log_probs, double* centers_out, double* loss_out,
       uint32_t* iters_out, double* loss_trace, uint32_t trace_cap) {
    const size_t L = c.n_loci, N = c.n_cells;
    std::vector<std::vector<F>> centers(k, std::vector<F>(L)), sums(k, std::vector<F>(L)), denoms(k, std::vector<F>(L));
    std::vector<std::vector<F>> lnc(k, std::vector<F>(L)), ln1mc(k, std::vector<F>(L));
    for (uint32_t q = 0; q < k; ++q)
        for (size_t l = 0; l < L; ++l) centers[q][l] = (F)init[q * L + l];
    const F log_prior = std::log((F)(1.0f / (float)k));  // main.rs:122
    const F limit = (F)(0.01f * (float)N);               // main.rs:136
    const int temp_steps = 9;
    F last = -std::numeric_limits<F>::infinity(), total = -std::numeric_limits<F>::infinity();
    uint32_t iterations = 0;
    std::vector<std::vector<F>> final_lp(N, std::vector<F>(k, 0));
    std::vector<F> lp(k), nlp(k), prob(k);
    for (int ts = 0; ts < temp_steps; ++ts) {
        F change = 10000;
        while (change > limit && iterations < 1000) {
            for (uint32_t q = 0; q < k; ++q)
                for (size_t l = 0; l < L; ++l) {
                    sums[q][l] = 1;  // pseudocounts, main.rs:253-262
                    denoms[q][l] = 2;
                    lnc[q][l] = std::log(centers[q][l]);  // binomial_loss takes these logs per entry; same values
                    ln1mc[q][l] = std::log((F)1 - centers[q][l]);
                }
            F loss = 0;
            for (size_t i = 0; i < N; ++i) {
                const SoupCell& cell = c.cells[i];
                for (uint32_t q = 0; q < k; ++q) {  // binomial_loss main.rs:199-213
                    F s = log_prior;
                    for (size_t e = 0; e < cell.loci.size(); ++e) {
                        const uint32_t l = cell.loci[e];
                        s += (F)cell.lbc[e] + (F)(float)cell.alt[e] * lnc[q][l] + (F)(float)cell.ref[e] * ln1mc[q][l];
                    }
                    lp[q] = s;
                }
                loss += log_sum_exp(lp);
                float temp = std::fmax(cell.total_alleles / (20.0f * std::pow(2.0f, (float)ts)), 1.0f);  // main.rs:152
                if (ts == temp_steps - 1) temp = 1.0f;
                for (uint32_t q = 0; q < k; ++q) nlp[q] = lp[q] / (F)temp;  // normalize_in_log_with_temp
                const F z = log_sum_exp(nlp);
                for (uint32_t q = 0; q < k; ++q) prob[q] = std::exp(nlp[q] - z);
                for (size_t e = 0; e < cell.loci.size(); ++e) {  // update_centers_average main.rs:273-280
                    const uint32_t l = cell.loci[e];
                    for (uint32_t q = 0; q < k; ++q) {
                        sums[q][l] += prob[q] * (F)(float)cell.alt[e];
                        denoms[q][l] += prob[q] * (F)(float)(cell.alt[e] + cell.ref[e]);
                    }
                }
                final_lp[i] = lp;
            }
            total = loss;
            change = loss - last;
            last = loss;
            for (size_t l = 0; l < L; ++l)  // update_final main.rs:244-251
                for (uint32_t q = 0; q < k; ++q) {
                    const F u = sums[q][l] / denoms[q][l];
                    centers[q][l] = std::fmax(std::fmin(u, (F)0.99f), (F)0.01f);
                }
            if (loss_trace && iterations < trace_cap) loss_trace[iterations] = (double)loss;
            ++iterations;
        }
    }
    if (log_probs)
        for (size_t i = 0; i < N; ++i)
            for (uint32_t q = 0; q < k; ++q) log_probs[i * k + q] = (double)final_lp[i][q];
    if (centers_out)
        for (uint32_t q = 0; q < k; ++q)
            for (size_t l = 0; l < L; ++l) centers_out[q * L + l] = (double)centers[q][l];
    if (loss_out) *loss_out = (double)total;
    if (iters_out) *iters_out = iterations;
    return 0;
}

}  // namespace

extern "C" {

sorc_ctx* sorc_create() { return new sorc_ctx(); }
void sorc_destroy(sorc_ctx* c) { delete c; }
const char* sorc_last_error(const sorc_ctx* c) { return c ? c->err.c_str() : ""; }

// souporcell's load_cell_data on COO arrays in file order (0-based coordinates)
int sorc_load_coo(sorc_ctx* c, uint64_t nnz0, uint32_t n_loci0, uint32_t n_cells, const uint32_t* locus,
                  const uint32_t* cell, const uint32_t* ref, const uint32_t* alt, uint32_t min_ref, uint32_t min_alt,
                  uint32_t min_ref_umis, uint32_t min_alt_umis, uint32_t* n_loci_used, uint64_t* n_entries) {
    c->n_cells = n_cells;
    c->n_loci0 = n_loci0;
    std::vector<uint32_t> cc(2 * (size_t)n_loci0, 0), uc(2 * (size_t)n_loci0, 0);
    std::vector<uint8_t> seen(n_loci0, 0);
    for (uint64_t i = 0; i < nnz0; ++i) {
        if (locus[i] >= n_loci0 || cell[i] >= n_cells) {
            c->err = "coordinate out of range";
            return 2;
        }
        seen[locus[i]] = 1;  // all_loci main.rs:410
        if (ref[i] > 0) {    // main.rs:418-419: one count per LINE, duplicated coordinates included
            cc[2 * (size_t)locus[i]] += 1;
            uc[2 * (size_t)locus[i]] += ref[i];
        }
        if (alt[i] > 0) {
            cc[2 * (size_t)locus[i] + 1] += 1;
            uc[2 * (size_t)locus[i] + 1] += alt[i];
        }
    }
    std::vector<int64_t> index(n_loci0, -1);
    c->locus_ids.clear();
    for (uint32_t l = 0; l < n_loci0; ++l)  // all_loci.sort(); main.rs:436-440
        if (seen[l] && cc[2 * (size_t)l] >= min_ref && cc[2 * (size_t)l + 1] >= min_alt && uc[2 * (size_t)l] >= min_ref_umis &&
            uc[2 * (size_t)l + 1] >= min_alt_umis) {
            index[l] = (int64_t)c->locus_ids.size();
            c->locus_ids.push_back(l);
        }
    c->n_loci = (uint32_t)c->locus_ids.size();
    // cell_counts.insert(cell, [ref, alt]) main.rs:421: the last line of a coordinate replaces the earlier ones
    std::unordered_map<uint64_t, uint64_t> last_line;
    last_line.reserve(nnz0 * 2);
    for (uint64_t i = 0; i < nnz0; ++i)
        if (index[locus[i]] >= 0) last_line[(uint64_t)locus[i] << 32 | cell[i]] = i;
    c->cells.assign(n_cells, SoupCell());
    // rows in ascending locus order (the outer loop of main.rs:436 walks the sorted loci): bucket by locus, cells in any
    // order inside a locus (each cell gets at most one entry per locus)
    std::vector<std::vector<uint64_t>> by_locus(c->n_loci);
    for (const auto& kv : last_line) by_locus[(size_t)index[kv.first >> 32]].push_back(kv.second);
    c->n_entries = 0;
    for (uint32_t li = 0; li < c->n_loci; ++li)
        for (uint64_t i : by_locus[li]) {
            if (ref[i] + alt[i] == 0) continue;  // main.rs:441
            SoupCell& sc = c->cells[cell[i]];
            sc.loci.push_back(li);
            sc.ref.push_back(ref[i]);
            sc.alt.push_back(alt[i]);
            sc.lbc.push_back((float)statrs::ln_binomial((uint64_t)ref[i] + alt[i], (uint64_t)alt[i]));
            sc.total_alleles += (float)(ref[i] + alt[i]);
            ++c->n_entries;
        }
    if (n_loci_used) *n_loci_used = c->n_loci;
    if (n_entries) *n_entries = c->n_entries;
    return 0;
}

int sorc_get_locus_ids(sorc_ctx* c, uint32_t* ids) {
    std::memcpy(ids, c->locus_ids.data(), c->locus_ids.size() * sizeof(uint32_t));
    return 0;
}

// One EM run from the given centers ([k][n_loci_used], f32).  precise = 0: f32 like the reference; 1: f64 statements.
// log_probs [n_cells][k], centers_out [k][n_loci_used], loss_trace [trace_cap] (loss of every iteration) may be NULL.
int sorc_em(sorc_ctx* c, uint32_t k, const float* centers_init, int precise, double* log_probs, double* centers_out,
            double* total_log_loss, uint32_t* iterations, double* loss_trace, uint32_t trace_cap) {
    if (k < 1) {
        c->err = "k < 1";
        return 2;
    }
    return precise ? em<double>(*c, k, centers_init, log_probs, centers_out, total_log_loss, iterations, loss_trace, trace_cap)
                   : em<float>(*c, k, centers_init, log_probs, centers_out, total_log_loss, iterations, loss_trace, trace_cap);
}

}  // extern "C"

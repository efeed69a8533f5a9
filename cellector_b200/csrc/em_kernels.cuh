// em_kernels.cuh — kernels of one EM iteration (compute_new_excluded, cellector/src/main.rs:308-347) and of
// the posterior pass (calculate_posteriors, main.rs:228-280).  Included by em.cu only.
#pragma once
#include <cuda_runtime.h>

#include "ctx.h"
#include "numerics.cuh"

namespace clx {

// ------------------------------------------------------------------------------------------------
// Per-locus alpha/beta (init_alpha_betas, main.rs:598-611) + memo table (numerics.cuh layout).
//   MODE 0 (EM):        alpha = alt_total + 1 - alt_of_excluded_cells, beta likewise with ref.
//   MODE 1 (posterior): three tables per locus (main.rs:239-254): slot 0 minority, 1 majority, 2 doublet.
// tot[2l] = ref total, tot[2l+1] = alt total; raw_min_ref/alt = sums over ALL stored entries of excluded cells.
struct Lbc10 {
    double v[10];
};

template <int MODE>
__global__ void __launch_bounds__(128) k_tables(const unsigned long long* __restrict__ tot,
                                                const unsigned long long* __restrict__ raw_min_ref,
                                                const unsigned long long* __restrict__ raw_min_alt, uint32_t n_loci,
                                                Lbc10 lbc, double mf, double mf_clamped, int with_expected,
                                                double* __restrict__ tab) {
    const uint64_t idx = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    const uint32_t n_tab = MODE == 0 ? 1u : 3u;
    if (idx >= (uint64_t)n_loci * n_tab) return;
    const uint32_t t = (uint32_t)(idx / n_loci), l = (uint32_t)(idx % n_loci);
    const double maj_a = (double)tot[2 * (size_t)l + 1] + 1.0 - (double)raw_min_alt[l];
    const double maj_b = (double)tot[2 * (size_t)l] + 1.0 - (double)raw_min_ref[l];
    double alpha = maj_a, beta = maj_b;
    if (MODE == 1) {
        // init_alpha_betas(locus_counts, included): totals + 1 - majority counts = minority counts + 1 (exact integers)
        const double min_a = (double)raw_min_alt[l] + 1.0;
        const double min_b = (double)raw_min_ref[l] + 1.0;
        if (t == 0) {
            alpha = min_a;
            beta = min_b;
        } else if (t == 1) {  // main.rs:250-254
            alpha = (maj_a - 1.0) * mf_clamped + 1.0;
            beta = (maj_b - 1.0) * mf_clamped + 1.0;
        } else {  // main.rs:244-248, un-clamped fraction
            alpha = (maj_a - 1.0) * mf + (min_a - 1.0) + 1.0;
            beta = (maj_b - 1.0) * mf + (min_b - 1.0) + 1.0;
        }
    }
    double out[TAB_STRIDE];
    build_locus_table(alpha, beta, lbc.v, with_expected != 0, out);
    double4* dst = reinterpret_cast<double4*>(tab + ((size_t)t * n_loci + l) * TAB_STRIDE);
#pragma unroll
    for (int i = 0; i < 4; ++i) dst[i] = make_double4(out[4 * i], out[4 * i + 1], out[4 * i + 2], out[4 * i + 3]);
}

// ------------------------------------------------------------------------------------------------
// Heavy entries (total > K_LIGHT): direct evaluation with the same expressions (stats.rs:41-53, 8-22).
template <int NT>
__global__ void __launch_bounds__(128) k_heavy(const uint32_t* __restrict__ h_locus, const uint32_t* __restrict__ h_ref,
                                               const uint32_t* __restrict__ h_alt, uint32_t n_heavy,
                                               const double* __restrict__ tab, uint32_t n_loci,
                                               const uint8_t* __restrict__ used, const double* __restrict__ lnfact,
                                               int with_expected, double* __restrict__ hv_ll, double* __restrict__ hv_e) {
    const uint32_t h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= n_heavy) return;
    const uint32_t l = h_locus[h];
    const bool on = used == nullptr || used[l] != 0;
    const uint64_t a = h_alt[h], r = h_ref[h], n = a + r;
    const double lbc = on ? ln_binomial(n, a, lnfact) : 0.0;
#pragma unroll
    for (int t = 0; t < NT; ++t) {
        double v = 0.0;
        const double* rec = tab + ((size_t)t * n_loci + l) * TAB_STRIDE;
        if (on) v = log_beta_binomial_pmf_den((double)a, (double)r, rec[TAB_ALPHA], rec[TAB_BETA], lbc, rec[TAB_DEN]);
        hv_ll[(size_t)t * n_heavy + h] = v;
    }
    if (NT == 1) {
        double e = 0.0;
        if (on && with_expected) e = expected_ll_direct(n, tab[(size_t)l * TAB_STRIDE + TAB_ALPHA],
                                                        tab[(size_t)l * TAB_STRIDE + TAB_BETA],
                                                        tab[(size_t)l * TAB_STRIDE + TAB_DEN], lnfact);
        hv_e[h] = e;
    }
}

// ------------------------------------------------------------------------------------------------
// v0 scoring kernel (get_cell_log_likelihoods, main.rs:541-591): G lanes per CSR row, 128-bit loads of two
// packed entries, memo-table gathers served by L2, xor-shuffle tree, heavy values appended in row order.
template <int G, int NT, bool WITH_E>
__global__ void __launch_bounds__(256) k_score_rows(const RowEntry* __restrict__ csr, const uint64_t* __restrict__ row_ptr,
                                                    uint32_t n_cells, const double* __restrict__ tab, uint32_t n_loci,
                                                    const uint8_t* __restrict__ used, const uint32_t* __restrict__ h_row_ptr,
                                                    const double* __restrict__ hv_ll, const double* __restrict__ hv_e,
                                                    uint32_t n_heavy, double* __restrict__ ll, double* __restrict__ ell,
                                                    uint32_t* __restrict__ nloci, double* __restrict__ norm,
                                                    unsigned int* __restrict__ nan_flag) {
    const uint32_t gid = (blockIdx.x * blockDim.x + threadIdx.x) / G;
    const uint32_t g = threadIdx.x % G;
    const bool live = gid < n_cells;
    const uint32_t row = live ? gid : 0;
    const uint64_t begin = live ? row_ptr[row] : 0, end = live ? row_ptr[row + 1] : 0;
    double acc[NT];
#pragma unroll
    for (int t = 0; t < NT; ++t) acc[t] = 0.0;
    double acc_e = 0.0;
    uint32_t cnt = 0;
    for (uint64_t i = (begin & ~1ull) + 2ull * g; i < end; i += 2ull * G) {
        const uint4 q = __ldg(reinterpret_cast<const uint4*>(csr + i));
        const uint32_t loc[2] = {q.x, q.z};
        const uint32_t cnts[2] = {q.y, q.w};
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const uint64_t idx = i + k;
            if (idx < begin || idx >= end) continue;
            const uint32_t l = loc[k];
            if (used != nullptr && used[l] == 0) continue;
            ++cnt;
            const uint32_t r = cnts[k] & 0xFFFFu, a = cnts[k] >> 16, n = r + a;
            if (n == 0 || n > (uint32_t)K_LIGHT) continue;  // exact zero / heavy list
            const double* rec = tab + (size_t)l * TAB_STRIDE;
#pragma unroll
            for (int t = 0; t < NT; ++t) acc[t] += __ldg(rec + (size_t)t * n_loci * TAB_STRIDE + tab_idx_T((int)a, (int)n));
            if (WITH_E) acc_e += __ldg(rec + tab_idx_E((int)n));
        }
    }
#pragma unroll
    for (int off = G / 2; off > 0; off >>= 1) {
#pragma unroll
        for (int t = 0; t < NT; ++t) acc[t] += __shfl_xor_sync(0xffffffffu, acc[t], off);
        if (WITH_E) acc_e += __shfl_xor_sync(0xffffffffu, acc_e, off);
        cnt += __shfl_xor_sync(0xffffffffu, cnt, off);
    }
    if (live && g == 0) {
        const uint32_t hb = h_row_ptr[row], he = h_row_ptr[row + 1];
        for (uint32_t h = hb; h < he; ++h) {
#pragma unroll
            for (int t = 0; t < NT; ++t) acc[t] += hv_ll[(size_t)t * n_heavy + h];
            if (WITH_E) acc_e += hv_e[h];
        }
#pragma unroll
        for (int t = 0; t < NT; ++t) ll[(size_t)t * n_cells + row] = acc[t];
        if (NT == 1) {
            ell[row] = acc_e;
            nloci[row] = cnt;
            const double nv = cnt > 0 ? acc[0] / (double)cnt : 0.0;  // main.rs:315-322
            norm[row] = nv;
            if (nv != nv) atomicOr(nan_flag, 1u);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Exclusion (main.rs:330-335): new set = {i : norm_i < threshold}; one warp writes one 32-bit word.
__global__ void __launch_bounds__(256) k_exclude(const double* __restrict__ norm, uint32_t n_cells,
                                                 const uint32_t* __restrict__ old_bits, uint32_t* __restrict__ new_bits,
                                                 DeviceStats* __restrict__ st) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const double thr = st->threshold;
    const bool ex = i < n_cells && norm[i] < thr;
    const uint32_t word = __ballot_sync(0xffffffffu, ex);
    if ((threadIdx.x & 31) == 0 && i < n_cells) {
        const uint32_t w = i >> 5;
        const uint32_t old = old_bits[w];
        new_bits[w] = word;
        const uint32_t nn = __popc(word & ~old), nr = __popc(old & ~word), ne = __popc(word);
        if (nn) atomicAdd(&st->n_new, (unsigned long long)nn);
        if (nr) atomicAdd(&st->n_rescued, (unsigned long long)nr);
        if (ne) atomicAdd(&st->n_excluded, (unsigned long long)ne);
    }
}

// ------------------------------------------------------------------------------------------------
// Column pass (get_locus_log_likelihoods, main.rs:368-420) fused with the allele counts of the excluded cells
// that the next iteration's init_alpha_betas needs.  G lanes per CSC column, cells ascending, fixed xor tree
// (no floating-point atomics); heavy entries are appended by lane 0 in column order.
// loc_ll  : [0,L) minority sum log_pmf, [L,2L) majority.
// loc_int : 8 arrays of L: cells_min, cells_maj, ref_min, alt_min, ref_maj, alt_maj, raw_min_ref, raw_min_alt.
template <int G>
__global__ void __launch_bounds__(256) k_columns(const ColEntry* __restrict__ csc, const uint64_t* __restrict__ col_ptr,
                                                 uint32_t n_loci, const double* __restrict__ tab,
                                                 const uint8_t* __restrict__ used, const uint32_t* __restrict__ excl_bits,
                                                 const uint32_t* __restrict__ hc_col_ptr, const uint32_t* __restrict__ hc_perm,
                                                 const uint32_t* __restrict__ h_cell, const uint32_t* __restrict__ h_ref,
                                                 const uint32_t* __restrict__ h_alt, const double* __restrict__ hv_ll,
                                                 double* __restrict__ loc_ll, unsigned long long* __restrict__ loc_int) {
    const uint32_t gid = (blockIdx.x * blockDim.x + threadIdx.x) / G;
    const uint32_t g = threadIdx.x % G;
    const bool live = gid < n_loci;
    const uint32_t l = live ? gid : 0;
    const uint64_t begin = live ? col_ptr[l] : 0, end = live ? col_ptr[l + 1] : 0;
    const double* rec = tab + (size_t)l * TAB_STRIDE;
    double s_min = 0.0, s_maj = 0.0;
    uint32_t c_min = 0, c_maj = 0, r_min = 0, a_min = 0, r_maj = 0, a_maj = 0;
    for (uint64_t i = (begin & ~1ull) + 2ull * g; i < end; i += 2ull * G) {
        const uint4 q = __ldg(reinterpret_cast<const uint4*>(csc + i));
        const uint32_t cel[2] = {q.x, q.z};
        const uint32_t cnts[2] = {q.y, q.w};
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const uint64_t idx = i + k;
            if (idx < begin || idx >= end) continue;
            const uint32_t c = cel[k];
            const bool ex = (__ldg(excl_bits + (c >> 5)) >> (c & 31)) & 1u;
            const uint32_t r = cnts[k] & 0xFFFFu, a = cnts[k] >> 16, n = r + a;
            if (ex) ++c_min; else ++c_maj;
            if (n == 0 || n > (uint32_t)K_LIGHT) continue;
            const double v = __ldg(rec + tab_idx_T((int)a, (int)n));
            if (ex) { s_min += v; r_min += r; a_min += a; }
            else    { s_maj += v; r_maj += r; a_maj += a; }
        }
    }
#pragma unroll
    for (int off = G / 2; off > 0; off >>= 1) {
        s_min += __shfl_xor_sync(0xffffffffu, s_min, off);
        s_maj += __shfl_xor_sync(0xffffffffu, s_maj, off);
        c_min += __shfl_xor_sync(0xffffffffu, c_min, off);
        c_maj += __shfl_xor_sync(0xffffffffu, c_maj, off);
        r_min += __shfl_xor_sync(0xffffffffu, r_min, off);
        a_min += __shfl_xor_sync(0xffffffffu, a_min, off);
        r_maj += __shfl_xor_sync(0xffffffffu, r_maj, off);
        a_maj += __shfl_xor_sync(0xffffffffu, a_maj, off);
    }
    if (live && g == 0) {
        unsigned long long R_min = r_min, A_min = a_min, R_maj = r_maj, A_maj = a_maj;
        const uint32_t hb = hc_col_ptr[l], he = hc_col_ptr[l + 1];
        for (uint32_t j = hb; j < he; ++j) {
            const uint32_t h = hc_perm[j];
            const uint32_t c = h_cell[h];
            const bool ex = (excl_bits[c >> 5] >> (c & 31)) & 1u;
            const double v = hv_ll[h];
            if (ex) { s_min += v; R_min += h_ref[h]; A_min += h_alt[h]; }
            else    { s_maj += v; R_maj += h_ref[h]; A_maj += h_alt[h]; }
        }
        const bool on = used[l] != 0;  // entries were scored only if the locus was in use (main.rs:556)
        const size_t L = n_loci;
        loc_ll[l] = on ? s_min : 0.0;
        loc_ll[L + l] = on ? s_maj : 0.0;
        loc_int[0 * L + l] = on ? c_min : 0;
        loc_int[1 * L + l] = on ? c_maj : 0;
        loc_int[2 * L + l] = on ? R_min : 0;
        loc_int[3 * L + l] = on ? A_min : 0;
        loc_int[4 * L + l] = on ? R_maj : 0;
        loc_int[5 * L + l] = on ? A_maj : 0;
        loc_int[6 * L + l] = R_min;  // init_alpha_betas subtracts every stored entry of an excluded cell (main.rs:603-608)
        loc_int[7 * L + l] = A_min;
    }
}

// The -80 rule (main.rs:431-451): per-cell minority log likelihood of the locus below -80 switches it off.
__global__ void __launch_bounds__(256) k_locus_filter(const double* __restrict__ loc_ll,
                                                      const unsigned long long* __restrict__ loc_int, uint32_t n_loci,
                                                      uint8_t* __restrict__ used, DeviceStats* __restrict__ st) {
    const uint32_t l = blockIdx.x * blockDim.x + threadIdx.x;
    bool filtered = false, in_use = false;
    if (l < n_loci) {
        const unsigned long long cm = loc_int[l];
        const double per_cell = cm != 0 ? loc_ll[l] / (double)cm : 0.0;
        const bool was = used[l] != 0;
        if (per_cell < -80.0) {
            filtered = was;
            used[l] = 0;
        } else {
            in_use = was;
        }
    }
    const uint32_t f = __popc(__ballot_sync(0xffffffffu, filtered)), u = __popc(__ballot_sync(0xffffffffu, in_use));
    if ((threadIdx.x & 31) == 0) {
        if (f) atomicAdd(&st->n_loci_filtered, (unsigned long long)f);
        if (u) atomicAdd(&st->n_loci_in_use, (unsigned long long)u);
    }
}

}  // namespace clx

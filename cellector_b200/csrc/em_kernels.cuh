// em_kernels.cuh — kernels of one EM iteration (compute_new_excluded, cellector/src/main.rs:308-347) and of
// the posterior pass (calculate_posteriors, main.rs:228-280).  Included by em.cu only.
#pragma once
#include <cuda_runtime.h>

#include "ctx.h"
#include "numerics.cuh"
#include "p2p_dev.cuh"

namespace clx {

// ------------------------------------------------------------------------------------------------
// Per-locus alpha/beta (init_alpha_betas, main.rs:598-611) + memo table (numerics.cuh layout).
//   MODE 0 (EM):        alpha = alt_total + 1 - alt_of_excluded_cells, beta likewise with ref.
//   MODE 1 (posterior): three tables per locus (main.rs:239-254): slot 0 minority, 1 majority, 2 doublet.
// tot[2l] = ref total, tot[2l+1] = alt total; raw_min_ref/alt = sums over ALL stored entries of excluded cells.
struct Lbc10 {
    double v[10];
};

__device__ __forceinline__ double shfl_f64(double v, int src) { return __shfl_sync(0xffffffffu, v, src); }
__device__ __forceinline__ double quiet_nan() { return __longlong_as_double(0x7ff8000000000000ll); }

// The table build is fp64-issue bound (a Lanczos ln_gamma is ~320 instructions), so the ln_gamma evaluations of
// a CTA's TAB_LPB loci are laid out as one flat task list with every lane busy, results go to shared memory,
// and a second phase combines them in the reference's expression order (stats.rs:41-53), so every value is
// bit-identical to a one-thread evaluation.  Tasks of one locus:
//   0..3   ln_gamma(i + alpha)          4..7   ln_gamma(i + beta)
//   MODE 0 (alpha, beta integer-valued: (a + alpha) + (r + beta) is the same double for every a)
//     8..10  ln_gamma((0 + alpha) + (n + beta)), n = 1..3        11  ln_gamma(alpha + beta)
//   MODE 1 (alpha, beta scaled by the minority fraction: keep every (a | n) argument as the reference forms it)
//     8..16  ln_gamma((a + alpha) + (r + beta)) for the nine (a | n) pairs        17  ln_gamma(alpha + beta)
// A locus switched off by the -80 rule gets NaN in every T/E slot (slot 3 = 0.0 / NaN is the "used" marker
// zero-total and heavy entries read): the scorer then needs no separate loci_used gather.  Block 0 also
// clears the per-iteration DeviceStats.
constexpr int TAB_LPB = 64;  // loci per CTA
template <int MODE>
__global__ void __launch_bounds__(256) k_tables(const unsigned long long* __restrict__ tot,
                                                const unsigned long long* __restrict__ raw_min_ref,
                                                const unsigned long long* __restrict__ raw_min_alt, uint32_t n_loci,
                                                Lbc10 lbc, double mf, double mf_clamped, int with_expected,
                                                const uint8_t* __restrict__ used, double* __restrict__ tab,
                                                double* __restrict__ tab16, uint32_t n_loci_pad,
                                                DeviceStats* __restrict__ st_clear, const double* __restrict__ lng,
                                                uint32_t lng_n) {
    static_assert(K_LIGHT == 3, "task layout below is for K_LIGHT == 3");
    constexpr int NTASK = MODE == 0 ? 12 : 18;
    constexpr int TDEN = NTASK - 1;  // task index of ln_gamma(alpha + beta)
    __shared__ double s_ab[TAB_LPB][2];
    __shared__ unsigned char s_on[TAB_LPB];
    __shared__ double s_g[TAB_LPB][NTASK];
    __shared__ double s_rec[TAB_LPB][TAB_STRIDE];
    __shared__ double s_lbc[10];
    if (st_clear != nullptr && blockIdx.x == 0 && threadIdx.x < sizeof(DeviceStats) / 8)
        reinterpret_cast<unsigned long long*>(st_clear)[threadIdx.x] = 0ull;
    const uint32_t blocks_per_tab = (n_loci + TAB_LPB - 1) / TAB_LPB;
    const uint32_t t = blockIdx.x / blocks_per_tab;
    const uint32_t l0 = (blockIdx.x % blocks_per_tab) * TAB_LPB;
    const int tid = threadIdx.x;
    if (tid < 10) s_lbc[tid] = lbc.v[tid];
    if (tid < TAB_LPB) {
        const uint32_t l = l0 + tid;
        double alpha = 1.0, beta = 1.0;
        bool on = false;
        if (l < n_loci) {
            const double maj_a = (double)tot[2 * (size_t)l + 1] + 1.0 - (double)raw_min_alt[l];
            const double maj_b = (double)tot[2 * (size_t)l] + 1.0 - (double)raw_min_ref[l];
            alpha = maj_a;
            beta = maj_b;
            if (MODE == 1) {
                // init_alpha_betas(locus_counts, included): totals + 1 - majority counts = minority counts + 1
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
            on = used == nullptr || used[l] != 0;
        }
        s_ab[tid][0] = alpha;
        s_ab[tid][1] = beta;
        s_on[tid] = on ? 1 : 0;
    }
    __syncthreads();
    // phase 1: flat ln_gamma task list
    for (int task = tid; task < TAB_LPB * NTASK; task += blockDim.x) {
        const int ll = task / NTASK, j = task % NTASK;
        if (!s_on[ll]) continue;
        const double alpha = s_ab[ll][0], beta = s_ab[ll][1];
        double x;
        if (j < 4) x = (double)j + alpha;
        else if (j < 8) x = (double)(j - 4) + beta;
        else if (j == TDEN) x = alpha + beta;
        else if (MODE == 0) x = (0.0 + alpha) + ((double)(j - 7) + beta);
        else {
            const int p = j - 8;
            const int pn = p < 2 ? 1 : (p < 5 ? 2 : 3);
            const int pa = p < 2 ? p : (p < 5 ? p - 2 : p - 5);
            x = ((double)pa + alpha) + ((double)(pn - pa) + beta);
        }
        s_g[ll][j] = MODE == 0 ? ln_gamma_int(x, lng, lng_n) : ln_gamma(x);
    }
    __syncthreads();
    // phase 2a: T(a | n) = lbc + (lga[a] + lgb[r] - lgab) - den for the nine pairs; record tail
    for (int task = tid; task < TAB_LPB * 16; task += blockDim.x) {
        const int ll = task >> 4, slot = task & 15;
        const bool on = s_on[ll] != 0;
        const double* g = s_g[ll];
        double out = quiet_nan();
        if (slot == TAB_ALPHA) out = s_ab[ll][0];
        else if (slot == TAB_BETA) out = s_ab[ll][1];
        else if (on) {
            // log_beta_calc(alpha, beta): ln_gamma(0 + alpha), ln_gamma(0 + beta) share their argument bits with tasks 0 / 4
            const double den = g[0] + g[4] - g[TDEN];
            if (slot == TAB_DEN) out = den;
            else if (slot == 3) out = 0.0;
            else if (slot < 12) {
                const int pn = (slot >> 2) + 1, pa = slot & 3;
                if (pa <= pn) {
                    const int third = MODE == 0 ? 7 + pn : 8 + (pn == 1 ? 0 : (pn == 2 ? 2 : 5)) + pa;
                    const double num = g[pa] + g[4 + (pn - pa)] - g[third];
                    out = s_lbc[lbc_idx(pa, pn)] + num - den;
                }
            }
        }
        s_rec[ll][slot] = out;
    }
    __syncthreads();
    // phase 2b: expected_log_likelihood (stats.rs:17-22), one thread per (locus, n), left-to-right chain
    if (tid < TAB_LPB * K_LIGHT) {
        const int ll = tid / K_LIGHT, n = tid % K_LIGHT + 1;
        if (s_on[ll]) {
            double e = 0.0;
            if (with_expected) {
                e = 2.0 * s_rec[ll][tab_idx_T(0, n)];
                for (int k = 1; k <= n; ++k) e = logsumexp2(e, 2.0 * s_rec[ll][tab_idx_T(k, n)]);
            }
            s_rec[ll][tab_idx_E(n)] = e;
        }
    }
    __syncthreads();
    // coalesced store of the CTA's records (slots 2, 7, 12 were written by phase 2b after 2a left NaN there)
    const uint32_t n_here = n_loci - l0 < (uint32_t)TAB_LPB ? n_loci - l0 : (uint32_t)TAB_LPB;
    double* dst = tab + ((size_t)t * n_loci + l0) * TAB_STRIDE;
    for (uint32_t i = tid; i < n_here * TAB_STRIDE; i += blockDim.x) dst[i] = (&s_rec[0][0])[i];
    // 16-byte slots of the run-stream scoring kernel (numerics.cuh).  MODE 0: {T, E(n)}, zeros for a switched-off
    // locus.  MODE 1 (posterior pass): table 0 / 1 are the two halves of pass 0, table 2 is the first half of pass 1
    // (its second half stays 0.0 from the allocation).
    if (tab16 != nullptr) {
        const uint32_t pass = (MODE == 1 && t == 2) ? 1u : 0u, half = (MODE == 1 && t == 1) ? 1u : 0u;
        double* d16 = tab16 + (((size_t)pass * n_loci_pad + l0) * TAB_SLOTS) * 2;
        for (uint32_t i = tid; i < n_here * TAB_SLOTS; i += blockDim.x) {
            const uint32_t ll = i / TAB_SLOTS, p = i % TAB_SLOTS;
            const int n = p < 2 ? 1 : (p < 5 ? 2 : 3);
            const bool on = s_on[ll] != 0;
            const double v = on ? s_rec[ll][tab_idx_T((int)p - (n * (n + 1)) / 2 + 1, n)] : 0.0;
            if (MODE == 0) {
                d16[2 * i] = v;
                d16[2 * i + 1] = on ? s_rec[ll][tab_idx_E(n)] : 0.0;
            } else {
                d16[2 * i + half] = v;
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Heavy entries (total > K_LIGHT): direct evaluation with the same expressions (stats.rs:41-53, 8-22).
//
// EM pass (k_heavy_em): alpha, beta are integer-valued, so every ln_gamma argument is an integer and comes from the
// lng memo (numerics.cuh ln_gamma_int: bit-identical to the Lanczos series), and the third argument
// (a + alpha) + (r + beta) is the same double for every split k of n.  One thread per entry for n <= HEAVY_CHAIN_MAX:
//   ll_k = ln C(n,k) + (lnG(k + alpha) + lnG(n - k + beta) - lnG(n + alpha + beta)) - den,   own log_pmf = ll_alt,
// and the expected-log-likelihood chain runs left to right like stats.rs:17-22.  Deeper entries are taken by the
// whole warp, one after the other, with a max / sum logsumexp over k (same value to ~1e-16 relative; diagnostic
// column only).  Entries with n > HEAVY_GIANT are left to k_heavy_giant (one CTA each).
// alpha, beta and ln B(alpha, beta) of the EM pass straight from the counts (same expressions as k_tables<0>), so the
// heavy kernels do not wait for the table build: the two run as parallel branches of the iteration graph.
struct EmCounts {
    const unsigned long long* tot;          // [2L] (ref, alt) totals
    const unsigned long long* raw_min_ref;  // [L]
    const unsigned long long* raw_min_alt;  // [L]
    const uint8_t* used;                    // [L]
    const double* lng;
    uint32_t lng_n;
};
__device__ __forceinline__ void em_alpha_beta(const EmCounts& k, uint32_t l, double* alpha, double* beta, double* den) {
    const double a = (double)k.tot[2 * (size_t)l + 1] + 1.0 - (double)k.raw_min_alt[l];
    const double b = (double)k.tot[2 * (size_t)l] + 1.0 - (double)k.raw_min_ref[l];
    *alpha = a;
    *beta = b;
    if (k.used[l] != 0) *den = ln_gamma_int(0.0 + a, k.lng, k.lng_n) + ln_gamma_int(0.0 + b, k.lng, k.lng_n) - ln_gamma_int(a + b, k.lng, k.lng_n);
    else *den = quiet_nan();  // switched off by the -80 rule: contributes nothing
}

constexpr int HEAVY_HB = 256;           // entries per CTA (one thread each)
constexpr uint32_t HEAVY_GIANT = HEAVY_GIANT_TOTAL;
constexpr uint32_t HEAVY_CHAIN_MAX = HEAVY_CHAIN_TOTAL;

__device__ __forceinline__ double ln_factorial_int(uint64_t n, const double* __restrict__ lnfact, const double* __restrict__ lng,
                                                   uint32_t lng_n) {
    return n <= 170 ? lnfact[n] : ln_gamma_int((double)n + 1.0, lng, lng_n);  // statrs ln_factorial
}
__device__ __forceinline__ double ln_binomial_int(uint64_t n, uint64_t k, const double* __restrict__ lnfact,
                                                  const double* __restrict__ lng, uint32_t lng_n) {
    return ln_factorial_int(n, lnfact, lng, lng_n) - ln_factorial_int(k, lnfact, lng, lng_n) - ln_factorial_int(n - k, lnfact, lng, lng_n);
}
// ll_k of stats.rs:12-16 for an entry of total n at a locus with integer-valued (alpha, beta)
static __device__ __noinline__ double heavy_ll_k(uint64_t n, uint64_t k, double alpha, double beta, double lgab, double den,
                                             const double* __restrict__ lnfact, const double* __restrict__ lng, uint32_t lng_n) {
    const double num = ln_gamma_int((double)k + alpha, lng, lng_n) + ln_gamma_int((double)(n - k) + beta, lng, lng_n) - lgab;
    return ln_binomial_int(n, k, lnfact, lng, lng_n) + num - den;
}

__global__ void __launch_bounds__(HEAVY_HB) k_heavy_em(const uint32_t* __restrict__ deep_idx, uint32_t n_deep,
                                                       const uint32_t* __restrict__ h_locus, const uint32_t* __restrict__ h_ref,
                                                       const uint32_t* __restrict__ h_alt, EmCounts cnts,
                                                       const double* __restrict__ lnfact, int with_expected,
                                                       double* __restrict__ hv_ll, double* __restrict__ hv_e,
                                                       const uint32_t* __restrict__ hm_pos, double* __restrict__ hm_val) {
    const double* __restrict__ lng = cnts.lng;
    const uint32_t lng_n = cnts.lng_n;
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t lane = threadIdx.x & 31u;
    const bool live = i < n_deep;
    const uint32_t h = live ? deep_idx[i] : 0u;
    uint64_t a = 0, n = 0;
    double alpha = 1.0, beta = 1.0, den = 0.0, lgab = 0.0;
    bool on = false;
    if (live) {
        a = h_alt[h];
        n = a + (uint64_t)h_ref[h];
        em_alpha_beta(cnts, h_locus[h], &alpha, &beta, &den);
        on = den == den;  // NaN den marks a locus that is switched off: contributes nothing
        if (!on) {
            hv_ll[h] = 0.0;
            hv_e[h] = 0.0;
            hm_val[hm_pos[h]] = 0.0;
        } else {
            lgab = ln_gamma_int(((double)a + alpha) + ((double)(n - a) + beta), lng, lng_n);
        }
    }
    // deep entries of this warp, one at a time over all 32 lanes
    unsigned int deep = __ballot_sync(0xffffffffu, live && on);
    while (deep) {
        const int src = __ffs(deep) - 1;
        deep &= deep - 1;
        const uint64_t dn = __shfl_sync(0xffffffffu, (unsigned long long)n, src), da = __shfl_sync(0xffffffffu, (unsigned long long)a, src);
        const double dal = __shfl_sync(0xffffffffu, alpha, src), dbe = __shfl_sync(0xffffffffu, beta, src);
        const double dden = __shfl_sync(0xffffffffu, den, src), dlgab = __shfl_sync(0xffffffffu, lgab, src);
        const uint32_t dh = __shfl_sync(0xffffffffu, h, src);
        double ex = 0.0;
        if (with_expected) {
            double m = -INFINITY;
            for (uint64_t k = lane; k <= dn; k += 32) m = fmax(m, 2.0 * heavy_ll_k(dn, k, dal, dbe, dlgab, dden, lnfact, lng, lng_n));
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, off));
            double sum = 0.0;
            for (uint64_t k = lane; k <= dn; k += 32) sum += exp(2.0 * heavy_ll_k(dn, k, dal, dbe, dlgab, dden, lnfact, lng, lng_n) - m);
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, off);
            ex = m + log(sum);
        }
        if ((int)lane == src) {
            const double own = heavy_ll_k(dn, da, dal, dbe, dlgab, dden, lnfact, lng, lng_n);
            hv_ll[dh] = own;
            hv_e[dh] = ex;
            hm_val[hm_pos[dh]] = own;
        }
    }
}

// EM-pass table build (init_alpha_betas main.rs:598-611 + the memo record of numerics.cuh) as lookups: one thread
// per (locus, n), n = 1..3.  alpha and beta are integer-valued here, so every ln_gamma comes out of the lng memo and
// the record costs 2(n+1)+4 loads, n+1 additions chains and n logsumexp steps — no Lanczos series, no shared-memory
// phases.  Expressions are those of k_tables<0> / build_locus_table, operation for operation.  Thread 0.. of block
// 0 also clears the per-iteration DeviceStats.
__global__ void __launch_bounds__(256) k_tables_em(EmCounts cnts, uint32_t n_loci, Lbc10 lbc, int with_expected,
                                                   double* __restrict__ tab, double* __restrict__ tab16,
                                                   DeviceStats* __restrict__ st_clear, unsigned int* __restrict__ zero_buf,
                                                   uint32_t zero_words) {
    __shared__ double s_lbc[10];
    if (threadIdx.x < 10) s_lbc[threadIdx.x] = lbc.v[threadIdx.x];
    if (st_clear != nullptr && blockIdx.x == 0 && threadIdx.x < sizeof(DeviceStats) / 8)
        reinterpret_cast<unsigned long long*>(st_clear)[threadIdx.x] = 0ull;
    __syncthreads();
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    // the select's histogram scratch of this iteration (saves a memset node in the iteration graph)
    for (uint32_t i = t; i < zero_words; i += gridDim.x * blockDim.x) zero_buf[i] = 0u;
    const uint32_t l = t / 3u;
    const int n = (int)(t % 3u) + 1;
    if (l >= n_loci) return;
    double alpha, beta, den;
    em_alpha_beta(cnts, l, &alpha, &beta, &den);
    const bool on = den == den;
    double* rec = tab + (size_t)l * TAB_STRIDE;
    double* r16 = tab16 != nullptr ? tab16 + (size_t)l * TAB_SLOTS * 2 : nullptr;  // {T(a | n), E(n)} slots (numerics.cuh)
    if (n == 1) {
        rec[3] = on ? 0.0 : quiet_nan();
        rec[TAB_DEN] = den;
        rec[TAB_ALPHA] = alpha;
        rec[TAB_BETA] = beta;
    }
    // rolled loops on purpose (small code: the kernel runs for microseconds and instruction fetch showed up first)
    double pm0 = 0.0, pm1 = 0.0, pm2 = 0.0, pm3 = 0.0;
    double e = on ? 0.0 : quiet_nan();
    if (on) {
        const double lgab = ln_gamma_int((0.0 + alpha) + ((double)n + beta), cnts.lng, cnts.lng_n);
#pragma unroll 1
        for (int a = 0; a <= n; ++a) {
            const double num = ln_gamma_int((double)a + alpha, cnts.lng, cnts.lng_n) +
                               ln_gamma_int((double)(n - a) + beta, cnts.lng, cnts.lng_n) - lgab;
            const double v = s_lbc[lbc_idx(a, n)] + num - den;
            if (a == 0) pm0 = v;
            else if (a == 1) pm1 = v;
            else if (a == 2) pm2 = v;
            else pm3 = v;
        }
        if (with_expected) {  // stats.rs:17-22, left to right
            e = 2.0 * pm0;
#pragma unroll 1
            for (int k = 1; k <= n; ++k) e = logsumexp2(e, 2.0 * (k == 1 ? pm1 : (k == 2 ? pm2 : pm3)));
        }
    }
#pragma unroll 1
    for (int a = 0; a <= n; ++a) {
        const double v = a == 0 ? pm0 : (a == 1 ? pm1 : (a == 2 ? pm2 : pm3));
        rec[tab_idx_T(a, n)] = on ? v : quiet_nan();
        if (r16) reinterpret_cast<double2*>(r16)[tab_slot(a, n)] = on ? make_double2(v, e) : make_double2(0.0, 0.0);
    }
    rec[tab_idx_E(n)] = e;
}

// Chain entries by (locus, total) group: the total + 1 values ll_k and the expected-log-likelihood chain (stats.rs:17-22)
// depend only on the locus and the total, so they are evaluated once per group and handed to every member (own log_pmf
// = ll_alt).  Same expressions, same order as the per-entry chain of k_heavy_em: bit-identical values, ~4x fewer of them
// at 14 heavy entries per locus (C3: 287 -> measured in profiles/).
__global__ void __launch_bounds__(128) k_heavy_groups(const uint32_t* __restrict__ hg_start, const uint4* __restrict__ hm_ent,
                                                      const uint32_t* __restrict__ hg_count, uint32_t n_heavy,
                                                      const uint32_t* __restrict__ h_locus, EmCounts cnts,
                                                      const double* __restrict__ lnfact, int with_expected,
                                                      double* __restrict__ hv_ll, double* __restrict__ hv_e, double* __restrict__ hm_val) {
    const double* __restrict__ lng = cnts.lng;
    const uint32_t lng_n = cnts.lng_n;
    const uint32_t n_groups = hg_count[0];
    for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < n_groups; g += gridDim.x * blockDim.x) {
        const uint32_t m0 = hg_start[g], m1 = g + 1 < n_groups ? hg_start[g + 1] : n_heavy;
        const uint4 e0 = hm_ent[m0];
        const uint64_t n = (uint64_t)e0.y + (uint64_t)e0.z;
        if (n > (uint64_t)HEAVY_CHAIN_MAX) continue;  // deep and giant entries: k_heavy_em / k_heavy_giant
        double alpha, beta, den;
        em_alpha_beta(cnts, h_locus[e0.w], &alpha, &beta, &den);
        if (!(den == den)) {  // switched off by the -80 rule: contributes nothing
            for (uint32_t m = m0; m < m1; ++m) {
                const uint32_t h = hm_ent[m].w;
                hv_ll[h] = 0.0;
                hv_e[h] = 0.0;
                hm_val[m] = 0.0;
            }
            continue;
        }
        const double lgab = ln_gamma_int(((double)0 + alpha) + ((double)n + beta), lng, lng_n);
        double ll[HEAVY_CHAIN_MAX + 1];
        double ex = 0.0;
        for (uint64_t k = 0; k <= n; ++k) {
            const double v = heavy_ll_k(n, k, alpha, beta, lgab, den, lnfact, lng, lng_n);
            ll[k] = v;
            if (with_expected) ex = k == 0 ? 2.0 * v : logsumexp2(ex, 2.0 * v);
        }
        for (uint32_t m = m0; m < m1; ++m) {
            const uint4 e = hm_ent[m];
            const double own = ll[e.z];
            hv_ll[e.w] = own;
            hv_e[e.w] = ex;
            hm_val[m] = own;
        }
    }
}

// One CTA per giant entry (total > HEAVY_GIANT): streaming max / sum logsumexp over k = 0..n in fixed chunks.
__global__ void __launch_bounds__(256) k_heavy_giant(const uint32_t* __restrict__ giant_idx, uint32_t n_giant,
                                                     const uint32_t* __restrict__ h_locus, const uint32_t* __restrict__ h_ref,
                                                     const uint32_t* __restrict__ h_alt, EmCounts cnts,
                                                     const double* __restrict__ lnfact, int with_expected,
                                                     double* __restrict__ hv_ll, double* __restrict__ hv_e,
                                                     const uint32_t* __restrict__ hm_pos, double* __restrict__ hm_val) {
    __shared__ double s_m[8], s_s[8];
    const uint32_t h = giant_idx[blockIdx.x];
    double alpha, beta, den;
    em_alpha_beta(cnts, h_locus[h], &alpha, &beta, &den);
    if (!(den == den)) {  // switched off by the -80 rule: contributes nothing
        if (threadIdx.x == 0) {
            hv_ll[h] = 0.0;
            hv_e[h] = 0.0;
            hm_val[hm_pos[h]] = 0.0;
        }
        return;
    }
    const uint64_t a = h_alt[h], r = h_ref[h], n = a + r;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) {
        const double v = log_beta_binomial_pmf_den((double)a, (double)r, alpha, beta, ln_binomial(n, a, lnfact), den);
        hv_ll[h] = v;
        hm_val[hm_pos[h]] = v;
    }
    if (!with_expected) {
        if (tid == 0) hv_e[h] = 0.0;
        return;
    }
    double m = -INFINITY, sum = 0.0;  // running max and sum of exp(2 ll_k - m)
    for (uint64_t k = tid; k <= n; k += blockDim.x) {
        const double v = 2.0 * log_beta_binomial_pmf_den((double)k, (double)(n - k), alpha, beta, ln_binomial(n, k, lnfact), den);
        if (v > m) {
            sum = sum * exp(m - v) + 1.0;
            m = v;
        } else {
            sum += exp(v - m);
        }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        const double m2 = __shfl_xor_sync(0xffffffffu, m, off), s2 = __shfl_xor_sync(0xffffffffu, sum, off);
        const double mm = fmax(m, m2);
        sum = (m == -INFINITY ? 0.0 : sum * exp(m - mm)) + (m2 == -INFINITY ? 0.0 : s2 * exp(m2 - mm));
        m = mm;
    }
    if (lane == 0) { s_m[warp] = m; s_s[warp] = sum; }
    __syncthreads();
    if (tid == 0) {
        double M = s_m[0], S = s_s[0];
        for (int w = 1; w < 8; ++w) {
            const double mm = fmax(M, s_m[w]);
            S = (M == -INFINITY ? 0.0 : S * exp(M - mm)) + (s_m[w] == -INFINITY ? 0.0 : s_s[w] * exp(s_m[w] - mm));
            M = mm;
        }
        hv_e[h] = M + log(S);
    }
}

// Posterior pass (three tables, no expected-log-likelihood): one thread per (entry, table).
__global__ void __launch_bounds__(256) k_heavy_post(const uint32_t* __restrict__ h_locus, const uint32_t* __restrict__ h_ref,
                                                    const uint32_t* __restrict__ h_alt, uint32_t n_heavy,
                                                    const double* __restrict__ tab, uint32_t n_loci,
                                                    const double* __restrict__ lnfact, double* __restrict__ hv_ll) {
    const uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    if (i >= 3ull * n_heavy) return;
    const uint32_t t = (uint32_t)(i / n_heavy), h = (uint32_t)(i % n_heavy);
    const double* rec = tab + ((size_t)t * n_loci + h_locus[h]) * TAB_STRIDE;
    const uint64_t a = h_alt[h], r = h_ref[h];
    hv_ll[(size_t)t * n_heavy + h] =
        log_beta_binomial_pmf_den((double)a, (double)r, rec[TAB_ALPHA], rec[TAB_BETA], ln_binomial(a + r, a, lnfact), rec[TAB_DEN]);
}

// Heavy values summed per cell, one warp per cell (lanes stride over the cell's heavy entries in row order, fixed xor
// tree: deterministic).  The scoring kernels add these sums instead of walking the heavy list with one lane: a cell of a
// skewed matrix (C5: the deepest cell holds 390 000 entries, thousands of them heavy) made that walk the whole pass.
// v0 / v1 / v2: up to three value arrays (EM: log_pmf, expected; posteriors: minority, majority, doublet); out[t][cell].
__global__ void __launch_bounds__(256) k_heavy_row_sums(const uint32_t* __restrict__ h_row_ptr, uint32_t n_cells,
                                                        const double* __restrict__ v0, const double* __restrict__ v1,
                                                        const double* __restrict__ v2, double* __restrict__ out) {
    const uint32_t row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
    if (row >= n_cells) return;
    const uint32_t hb = h_row_ptr[row], he = h_row_ptr[row + 1];
    double s0 = 0.0, s1 = 0.0, s2 = 0.0;
    for (uint32_t h = hb + lane; h < he; h += 32) {
        s0 += v0[h];
        if (v1) s1 += v1[h];
        if (v2) s2 += v2[h];
    }
    if (he - hb > 1) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            s0 += __shfl_xor_sync(0xffffffffu, s0, off);
            s1 += __shfl_xor_sync(0xffffffffu, s1, off);
            s2 += __shfl_xor_sync(0xffffffffu, s2, off);
        }
    }
    if (lane == 0) {
        out[row] = s0;
        if (v1) out[(size_t)n_cells + row] = s1;
        if (v2) out[2 * (size_t)n_cells + row] = s2;
    }
}

// ------------------------------------------------------------------------------------------------
// Gather scoring kernel (get_cell_log_likelihoods, main.rs:541-591) for matrices whose memo table stays in L2:
// G lanes per CSR row, U independent 128-bit loads of two packed entries per lane in flight, one memo-table
// sector per entry (NaN = locus switched off), xor-shuffle tree, heavy values appended in row order.
constexpr int SCORE_UNROLL = 4;
constexpr uint32_t ROW_SEG = 4096;  // entries per row segment: a longer row is split over several lane groups
// segments of every row (at least one, so that an empty row still gets its outputs): ceil(len / ROW_SEG)
__global__ void k_row_seg_count(const uint64_t* __restrict__ row_ptr, uint32_t n_cells, uint32_t* __restrict__ n_seg) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i > n_cells) return;
    if (i == n_cells) {
        n_seg[i] = 0;
        return;
    }
    const uint64_t len = row_ptr[i + 1] - row_ptr[i];
    n_seg[i] = len ? (uint32_t)((len + ROW_SEG - 1) / ROW_SEG) : 1u;
}
__global__ void k_row_seg_fill(const uint32_t* __restrict__ seg_first, uint32_t n_cells, uint32_t* __restrict__ seg_row) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_cells) return;
    for (uint32_t s = seg_first[i]; s < seg_first[i + 1]; ++s) seg_row[s] = i;
}

// G lanes per row SEGMENT (seg_row / seg_first: a row of a skewed matrix can hold 10^5 entries; one lane group walking
// it alone was the whole pass at C5).  A row with one segment is finished here; the segments of a longer row leave
// partial sums in seg_part ([NT + 1][n_seg] doubles + counts) and k_score_rows_combine adds them in segment order.
template <int G, int NT, bool WITH_E>
__global__ void __launch_bounds__(256) k_score_rows(const RowEntry* __restrict__ csr, const uint64_t* __restrict__ row_ptr,
                                                    uint32_t n_cells, const double* __restrict__ tab, uint32_t n_loci,
                                                    const double* __restrict__ hrow, const uint32_t* __restrict__ seg_row,
                                                    const uint32_t* __restrict__ seg_first, uint32_t n_seg,
                                                    double* __restrict__ seg_part, uint32_t* __restrict__ seg_cnt,
                                                    double* __restrict__ ll, double* __restrict__ ell,
                                                    uint32_t* __restrict__ nloci, double* __restrict__ norm,
                                                    unsigned int* __restrict__ nan_flag) {
    constexpr int U = SCORE_UNROLL;
    const uint32_t gid = (blockIdx.x * blockDim.x + threadIdx.x) / G;
    const uint32_t g = threadIdx.x % G;
    const bool live = gid < n_seg;
    const uint32_t row = live ? seg_row[gid] : 0;
    const uint32_t sf = live ? seg_first[row] : 0, sn = live ? seg_first[row + 1] - sf : 1;
    const uint64_t rb = live ? row_ptr[row] : 0, re = live ? row_ptr[row + 1] : 0;
    const uint64_t begin = live ? rb + (uint64_t)(gid - sf) * ROW_SEG : 0;
    const uint64_t end = live ? (begin + ROW_SEG < re ? begin + ROW_SEG : re) : 0;
    double acc[NT];
#pragma unroll
    for (int t = 0; t < NT; ++t) acc[t] = 0.0;
    double acc_e = 0.0;
    uint32_t cnt = 0;
    const size_t tstride = (size_t)n_loci * TAB_STRIDE;
    for (uint64_t i0 = (begin & ~1ull) + 2ull * g; i0 < end; i0 += 2ull * G * U) {
        uint4 q[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint64_t i = i0 + 2ull * G * u;
            q[u] = i < end ? __ldg(reinterpret_cast<const uint4*>(csr + i)) : make_uint4(0u, 0u, 0u, 0u);
        }
        double v[U][2][NT], ev[U][2];
        bool ok[U][2];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint32_t loc[2] = {q[u].x, q[u].z};
            const uint32_t cnts[2] = {q[u].y, q[u].w};
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                const uint64_t idx = i0 + 2ull * G * u + k;
                ok[u][k] = idx >= begin && idx < end;
                const uint32_t r = cnts[k] & 0xFFFFu, a = cnts[k] >> 16, n = r + a;
                const bool light = (n - 1u) < (uint32_t)K_LIGHT;  // zero-total and heavy entries read the marker slot
                const double* rec = tab + (size_t)(ok[u][k] ? loc[k] : 0u) * TAB_STRIDE;
                const int st = light ? tab_idx_T((int)a, (int)n) : 3;
#pragma unroll
                for (int t = 0; t < NT; ++t) v[u][k][t] = ok[u][k] ? __ldg(rec + (size_t)t * tstride + st) : 0.0;
                if (WITH_E) ev[u][k] = ok[u][k] ? __ldg(rec + (light ? 5 * (int)n - 3 : 3)) : 0.0;
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                const bool used = ok[u][k] && (v[u][k][0] == v[u][k][0]);
                cnt += used ? 1u : 0u;
#pragma unroll
                for (int t = 0; t < NT; ++t) acc[t] += used ? v[u][k][t] : 0.0;
                if (WITH_E) acc_e += used ? ev[u][k] : 0.0;
            }
        }
    }
#pragma unroll
    for (int off = G / 2; off > 0; off >>= 1) {
#pragma unroll
        for (int t = 0; t < NT; ++t) acc[t] += __shfl_xor_sync(0xffffffffu, acc[t], off);
        if (WITH_E) acc_e += __shfl_xor_sync(0xffffffffu, acc_e, off);
        cnt += __shfl_xor_sync(0xffffffffu, cnt, off);
    }
    if (live && g == 0 && sn > 1) {  // one segment of a long row: partial sums for k_score_rows_combine
#pragma unroll
        for (int t = 0; t < NT; ++t) seg_part[(size_t)t * n_seg + gid] = acc[t];
        if (WITH_E) seg_part[(size_t)NT * n_seg + gid] = acc_e;
        seg_cnt[gid] = cnt;
    } else if (live && g == 0) {
        // the cell's heavy entries, already summed (k_heavy_row_sums): [t][cell], EM: [0] log_pmf, [1] expected
#pragma unroll
        for (int t = 0; t < NT; ++t) acc[t] += hrow[(size_t)t * n_cells + row];
        if (WITH_E) acc_e += hrow[(size_t)n_cells + row];
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

// rows of more than one segment: partial sums added in segment order, then the tail of k_score_rows
template <int NT, bool WITH_E>
__global__ void __launch_bounds__(256) k_score_rows_combine(const uint32_t* __restrict__ seg_first, uint32_t n_cells, uint32_t n_seg,
                                                            const double* __restrict__ seg_part, const uint32_t* __restrict__ seg_cnt,
                                                            const double* __restrict__ hrow, double* __restrict__ ll,
                                                            double* __restrict__ ell, uint32_t* __restrict__ nloci,
                                                            double* __restrict__ norm, unsigned int* __restrict__ nan_flag) {
    const uint32_t row = blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= n_cells) return;
    const uint32_t s0 = seg_first[row], s1 = seg_first[row + 1];
    if (s1 - s0 <= 1) return;
    double acc[NT], acc_e = 0.0;
    uint32_t cnt = 0;
#pragma unroll
    for (int t = 0; t < NT; ++t) acc[t] = 0.0;
    for (uint32_t s = s0; s < s1; ++s) {
#pragma unroll
        for (int t = 0; t < NT; ++t) acc[t] += seg_part[(size_t)t * n_seg + s];
        if (WITH_E) acc_e += seg_part[(size_t)NT * n_seg + s];
        cnt += seg_cnt[s];
    }
#pragma unroll
    for (int t = 0; t < NT; ++t) {
        acc[t] += hrow[(size_t)t * n_cells + row];
        ll[(size_t)t * n_cells + row] = acc[t];
    }
    if (NT == 1) {
        if (WITH_E) acc_e += hrow[(size_t)n_cells + row];
        ell[row] = acc_e;
        nloci[row] = cnt;
        const double nv = cnt > 0 ? acc[0] / (double)cnt : 0.0;  // main.rs:315-322
        norm[row] = nv;
        if (nv != nv) atomicOr(nan_flag, 1u);
    }
}

// ------------------------------------------------------------------------------------------------
// Exclusion (main.rs:330-335): new set = {i : norm_i < threshold}; one warp rewrites one 32-bit word in place.
__device__ __forceinline__ void exclude_word(const double* __restrict__ norm, uint32_t n_cells, uint32_t i, double thr,
                                             uint32_t* __restrict__ bits, DeviceStats* __restrict__ st) {
    const bool ex = i < n_cells && norm[i] < thr;
    const uint32_t word = __ballot_sync(0xffffffffu, ex);
    if ((threadIdx.x & 31) == 0 && i < n_cells) {
        const uint32_t w = i >> 5;
        const uint32_t old = bits[w];
        bits[w] = word;
        const uint32_t nn = __popc(word & ~old), nr = __popc(old & ~word), ne = __popc(word);
        if (nn) atomicAdd(&st->n_new, (unsigned long long)nn);
        if (nr) atomicAdd(&st->n_rescued, (unsigned long long)nr);
        if (ne) atomicAdd(&st->n_excluded, (unsigned long long)ne);
    }
}
__global__ void __launch_bounds__(256) k_exclude(const double* __restrict__ norm, uint32_t n_cells,
                                                 uint32_t* __restrict__ bits, DeviceStats* __restrict__ st) {
    exclude_word(norm, n_cells, blockIdx.x * blockDim.x + threadIdx.x, st->threshold, bits, st);
}

// ------------------------------------------------------------------------------------------------
// Column pass (get_locus_log_likelihoods, main.rs:368-420) fused with the allele counts of the excluded cells
// that the next iteration's init_alpha_betas needs and, on one GPU, with the -80 locus filter (main.rs:431-451).
// G lanes per CSC column, cells ascending, fixed xor tree (no floating-point atomics); the column's memo record
// is one L1-resident line; heavy entries are appended by lane 0 in column order.
// loc_ll  : [0,L) minority sum log_pmf, [L,2L) majority.
// loc_int : 8 arrays of L: cells_min, cells_maj, ref_min, alt_min, ref_maj, alt_maj, raw_min_ref, raw_min_alt.
template <int G, bool FUSE_FILTER>
__global__ void __launch_bounds__(256) k_columns(const ColEntry* __restrict__ csc, const uint64_t* __restrict__ col_ptr,
                                                 uint32_t n_loci, const double* __restrict__ tab,
                                                 uint8_t* __restrict__ used, const uint32_t* __restrict__ excl_bits,
                                                 const uint32_t* __restrict__ hc_col_ptr, const uint32_t* __restrict__ hc_perm,
                                                 const uint32_t* __restrict__ h_cell, const uint32_t* __restrict__ h_ref,
                                                 const uint32_t* __restrict__ h_alt, const double* __restrict__ hv_ll,
                                                 double* __restrict__ loc_ll, unsigned long long* __restrict__ loc_int,
                                                 uint32_t* __restrict__ n_used, DeviceStats* __restrict__ st) {
    constexpr int U = SCORE_UNROLL;
    const uint32_t gid = (blockIdx.x * blockDim.x + threadIdx.x) / G;
    const uint32_t g = threadIdx.x % G;
    const bool live = gid < n_loci;
    const uint32_t l = live ? gid : 0;
    const uint64_t begin = live ? col_ptr[l] : 0, end = live ? col_ptr[l + 1] : 0;
    const double* rec = tab + (size_t)l * TAB_STRIDE;
    double s_min = 0.0, s_maj = 0.0;
    uint32_t c_min = 0, c_maj = 0, r_min = 0, a_min = 0, r_maj = 0, a_maj = 0;
    for (uint64_t i0 = (begin & ~1ull) + 2ull * g; i0 < end; i0 += 2ull * G * U) {
        uint4 q[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint64_t i = i0 + 2ull * G * u;
            q[u] = i < end ? __ldg(reinterpret_cast<const uint4*>(csc + i)) : make_uint4(0u, 0u, 0u, 0u);
        }
        // excluded-bitmap words of the two cells; a slot outside [begin, end) may hold anything (the neighbouring
        // column, or bytes past the end of the array): never turn it into an address
        uint32_t w[U][2];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint64_t idx = i0 + 2ull * G * u;
            const bool ok0 = idx >= begin && idx < end, ok1 = idx + 1 >= begin && idx + 1 < end;
            w[u][0] = ok0 ? __ldg(excl_bits + (q[u].x >> 5)) : 0u;
            w[u][1] = ok1 ? __ldg(excl_bits + (q[u].z >> 5)) : 0u;
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint32_t cel[2] = {q[u].x, q[u].z};
            const uint32_t cnts[2] = {q[u].y, q[u].w};
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                const uint64_t idx = i0 + 2ull * G * u + k;
                if (idx < begin || idx >= end) continue;
                const bool ex = (w[u][k] >> (cel[k] & 31)) & 1u;
                const uint32_t r = cnts[k] & 0xFFFFu, a = cnts[k] >> 16, n = r + a;
                if (ex) ++c_min; else ++c_maj;
                if ((n - 1u) >= (uint32_t)K_LIGHT) continue;  // exact zero / heavy list
                const double v = __ldg(rec + tab_idx_T((int)a, (int)n));  // this column's record: one hot L1 line
                if (ex) { s_min += v; r_min += r; a_min += a; }
                else    { s_maj += v; r_maj += r; a_maj += a; }
            }
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
    bool filtered = false, in_use = false;
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
        if (FUSE_FILTER) {  // main.rs:431-451 on this locus
            const double per_cell = (on && c_min != 0) ? s_min / (double)c_min : 0.0;
            if (per_cell < -80.0) {
                filtered = on;
                used[l] = 0;
            } else {
                in_use = on;
            }
        }
    }
    if (FUSE_FILTER) {
        // a locus switched off now stops counting toward loci_used_per_cell (main.rs:575) of the cells it covers
        const bool grp_f = __shfl_sync(0xffffffffu, filtered ? 1 : 0, (int)((threadIdx.x & 31u) & ~(uint32_t)(G - 1))) != 0;
        if (grp_f)
            for (uint64_t i = begin + g; i < end; i += G) atomicSub(&n_used[csc[i].cell], 1u);
        const uint32_t f = __popc(__ballot_sync(0xffffffffu, filtered)), u = __popc(__ballot_sync(0xffffffffu, in_use));
        if ((threadIdx.x & 31) == 0) {
            if (f) atomicAdd(&st->n_loci_filtered, (unsigned long long)f);
            if (u) atomicAdd(&st->n_loci_in_use, (unsigned long long)u);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Column pass without the column walk.  A light entry's log_pmf depends only on (locus, alt, total), so the per-locus
// sums of get_locus_log_likelihoods (main.rs:368-420) are dot products of per-locus CLASS COUNTS with the locus's
// memo record: cls_all (all stored entries, fixed at load) and cls_min (entries of the excluded cells).  cls_min is
// kept up to date incrementally: only the rows of cells whose excluded bit changed since the last pass are walked
// (integer atomics, so the result does not depend on the order), which after the first iterations is a handful of rows
// instead of every nonzero of the matrix.
//
// One warp per row SEGMENT (the segments of the gather scoring kernel: at most ROW_SEG entries, so the deep cells of a
// skewed matrix are spread over many warps); a warp whose cell did not change leaves at once.  prev is brought up to date
// by k_locus_finish, which runs after every warp here has read it.
__global__ void __launch_bounds__(256) k_excl_delta(const RowEntry* __restrict__ csr, const uint64_t* __restrict__ row_ptr,
                                                    const uint32_t* __restrict__ seg_row, const uint32_t* __restrict__ seg_first,
                                                    uint32_t n_seg, const uint32_t* __restrict__ bits,
                                                    const uint32_t* __restrict__ prev, uint32_t* __restrict__ cls_min) {
    const uint32_t sg = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
    if (sg >= n_seg) return;
    const uint32_t cell = seg_row[sg];
    const uint32_t now = (bits[cell >> 5] >> (cell & 31u)) & 1u, was = (prev[cell >> 5] >> (cell & 31u)) & 1u;
    if (now == was) return;
    const uint32_t delta = now ? 1u : 0xFFFFFFFFu;  // joined the excluded set : left it
    const uint64_t rb = row_ptr[cell], re = row_ptr[cell + 1];
    const uint64_t b = rb + (uint64_t)(sg - seg_first[cell]) * ROW_SEG;
    const uint64_t e = b + ROW_SEG < re ? b + ROW_SEG : re;
    for (uint64_t i = b + lane; i < e; i += 32) {
        const RowEntry x = csr[i];
        const uint32_t n = (uint32_t)x.ref + (uint32_t)x.alt;
        const uint32_t cls = (n - 1u) < (uint32_t)K_LIGHT ? (uint32_t)tab_slot((int)x.alt, (int)n) : (uint32_t)CLS_OTHER;
        atomicAdd(&cls_min[(size_t)x.locus * CLS_STRIDE + cls], delta);
    }
}

// One thread per locus: the sums from the class counts, the heavy entries of the column appended in column order, the
// outputs of k_columns (same arrays, same meaning) and, on one GPU, the -80 locus filter (main.rs:431-451).
template <bool FUSE_FILTER>
__global__ void __launch_bounds__(256) k_locus_finish(const uint32_t* __restrict__ cls_all, const uint32_t* __restrict__ cls_min,
                                                      uint32_t n_loci, const double* __restrict__ tab, uint8_t* __restrict__ used,
                                                      const uint32_t* __restrict__ excl_bits,
                                                      const uint32_t* __restrict__ hm_col_ptr, const uint4* __restrict__ hm_ent,
                                                      const double* __restrict__ hm_val,
                                                      const ColEntry* __restrict__ csc, const uint64_t* __restrict__ col_ptr,
                                                      double* __restrict__ loc_ll, unsigned long long* __restrict__ loc_int,
                                                      uint32_t* __restrict__ n_used, DeviceStats* __restrict__ st,
                                                      uint32_t* __restrict__ excl_prev, uint32_t n_words) {
    const uint32_t l = blockIdx.x * blockDim.x + threadIdx.x;
    // cls_min now matches excl_bits (k_excl_delta has finished): remember the set it matches
    for (uint32_t i = l; i < n_words; i += gridDim.x * blockDim.x) excl_prev[i] = excl_bits[i];
    bool filtered = false, in_use = false;
    if (l < n_loci) {
        uint32_t ca[CLS_STRIDE], cm[CLS_STRIDE];
#pragma unroll
        for (int q = 0; q < CLS_STRIDE / 4; ++q) {
            const uint4 a = __ldg(reinterpret_cast<const uint4*>(cls_all + (size_t)l * CLS_STRIDE) + q);
            const uint4 m = __ldg(reinterpret_cast<const uint4*>(cls_min + (size_t)l * CLS_STRIDE) + q);
            ca[4 * q] = a.x; ca[4 * q + 1] = a.y; ca[4 * q + 2] = a.z; ca[4 * q + 3] = a.w;
            cm[4 * q] = m.x; cm[4 * q + 1] = m.y; cm[4 * q + 2] = m.z; cm[4 * q + 3] = m.w;
        }
        const double* rec = tab + (size_t)l * TAB_STRIDE;
        double s_min = 0.0, s_maj = 0.0;
        unsigned long long c_min = cm[CLS_OTHER], c_maj = ca[CLS_OTHER] - cm[CLS_OTHER];
        unsigned long long R_min = 0, A_min = 0, R_maj = 0, A_maj = 0;
#pragma unroll
        for (int n = 1; n <= K_LIGHT; ++n) {
#pragma unroll
            for (int a = 0; a <= n; ++a) {
                const int k = tab_slot(a, n);
                const uint32_t mi = cm[k], mj = ca[k] - cm[k];
                const double v = __ldg(rec + tab_idx_T(a, n));
                if (mi) s_min += (double)mi * v;  // a class nobody is in never touches its value
                if (mj) s_maj += (double)mj * v;
                c_min += mi;
                c_maj += mj;
                R_min += (unsigned long long)mi * (unsigned)(n - a);
                A_min += (unsigned long long)mi * (unsigned)a;
                R_maj += (unsigned long long)mj * (unsigned)(n - a);
                A_maj += (unsigned long long)mj * (unsigned)a;
            }
        }
        // the locus's heavy entries: one contiguous range in member order ({cell, ref, alt, .} + this pass's log_pmf)
        const uint32_t hb = hm_col_ptr[l], he = hm_col_ptr[l + 1];
        for (uint32_t m = hb; m < he; ++m) {
            const uint4 e = __ldg(hm_ent + m);
            const bool ex = (excl_bits[e.x >> 5] >> (e.x & 31)) & 1u;
            const double v = hm_val[m];
            if (ex) { s_min += v; R_min += e.y; A_min += e.z; }
            else    { s_maj += v; R_maj += e.y; A_maj += e.z; }
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
        if (FUSE_FILTER) {  // main.rs:431-451 on this locus
            const double per_cell = (on && c_min != 0) ? s_min / (double)c_min : 0.0;
            if (per_cell < -80.0) {
                filtered = on;
                used[l] = 0;
                if (on)  // rare: the cells of the locus stop counting it (loci_used_per_cell, main.rs:575)
                    for (uint64_t i = col_ptr[l]; i < col_ptr[l + 1]; ++i) atomicSub(&n_used[csc[i].cell], 1u);
            } else {
                in_use = on;
            }
        }
    }
    if (FUSE_FILTER) {
        const uint32_t f = __popc(__ballot_sync(0xffffffffu, filtered)), u = __popc(__ballot_sync(0xffffffffu, in_use));
        if ((threadIdx.x & 31) == 0) {
            if (f) atomicAdd(&st->n_loci_filtered, (unsigned long long)f);
            if (u) atomicAdd(&st->n_loci_in_use, (unsigned long long)u);
        }
    }
}

// The -80 rule (main.rs:431-451): per-cell minority log likelihood of the locus below -80 switches it off.
__global__ void __launch_bounds__(256) k_locus_filter(const double* __restrict__ loc_ll,
                                                      const unsigned long long* __restrict__ loc_int, uint32_t n_loci,
                                                      uint8_t* __restrict__ used, const ColEntry* __restrict__ csc,
                                                      const uint64_t* __restrict__ col_ptr, uint32_t* __restrict__ n_used,
                                                      DeviceStats* __restrict__ st) {
    const uint32_t l = blockIdx.x * blockDim.x + threadIdx.x;
    bool filtered = false, in_use = false;
    if (l < n_loci) {
        const unsigned long long cm = loc_int[l];
        const double per_cell = cm != 0 ? loc_ll[l] / (double)cm : 0.0;
        const bool was = used[l] != 0;
        if (per_cell < -80.0) {
            filtered = was;
            used[l] = 0;
            if (was)  // rare: this rank's cells of the locus stop counting it (loci_used_per_cell, main.rs:575)
                for (uint64_t i = col_ptr[l]; i < col_ptr[l + 1]; ++i) atomicSub(&n_used[csc[i].cell], 1u);
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

namespace clx {
// Sharded over peer memory: waits for barrier 2 (every slice owner has stored its totals into this rank's window),
// copies the totals and the summed change counters into the arrays the rest of the code reads, and applies the -80
// rule — k_locus_filter with the exchange fused in front.
__global__ void __launch_bounds__(256) k_locus_filter_p2p(unsigned char* const* __restrict__ win, int rank, int world,
                                                          size_t off_counters, size_t off_tot_ll, size_t off_tot_int,
                                                          double* __restrict__ loc_ll, unsigned long long* __restrict__ loc_int,
                                                          uint32_t n_loci, uint8_t* __restrict__ used, const ColEntry* __restrict__ csc,
                                                          const uint64_t* __restrict__ col_ptr, uint32_t* __restrict__ n_used,
                                                          DeviceStats* __restrict__ st) {
    p2p_sync_head(win, rank, world, 2, off_counters, &st->pad);
    const double* tot_ll = reinterpret_cast<const double*>(win[rank] + off_tot_ll);
    const unsigned long long* tot_int = reinterpret_cast<const unsigned long long*>(win[rank] + off_tot_int);
    if (blockIdx.x == 0 && threadIdx.x < 3)
        (&st->n_new)[threadIdx.x] = reinterpret_cast<const unsigned long long*>(win[rank] + off_counters)[4 + threadIdx.x];
    const uint32_t l = blockIdx.x * blockDim.x + threadIdx.x;
    const size_t L = n_loci;
    bool filtered = false, in_use = false;
    if (l < n_loci) {
        const double s_min = tot_ll[l];
        loc_ll[l] = s_min;
        loc_ll[L + l] = tot_ll[L + l];
        unsigned long long cm = 0;
#pragma unroll
        for (int a = 0; a < 8; ++a) {
            const unsigned long long v = tot_int[(size_t)a * L + l];
            loc_int[(size_t)a * L + l] = v;
            if (a == 0) cm = v;
        }
        const double per_cell = cm != 0 ? s_min / (double)cm : 0.0;
        const bool was = used[l] != 0;
        if (per_cell < -80.0) {
            filtered = was;
            used[l] = 0;
            if (was)
                for (uint64_t i = col_ptr[l]; i < col_ptr[l + 1]; ++i) atomicSub(&n_used[csc[i].cell], 1u);
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

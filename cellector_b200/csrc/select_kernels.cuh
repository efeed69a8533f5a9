// select_kernels.cuh — exact order statistics for the IQR threshold (main.rs:324-329) and the posterior
// elementwise pass (main.rs:259-278 + assignment rule main.rs:141-170).  Included by em.cu only.
//
// statrs 0.16.0 Data::{median, lower_quartile, upper_quartile} (R-8 quantile, h = (N + 1/3) tau + 1/3) need at
// most six order statistics.  They are selected exactly on order-preserving u64 keys:
//   * N <= SELECT_SMALL_MAX : one CTA sorts the keys in shared memory (bitonic network) — one launch;
//   * larger N             : MSB-first radix select, 8 passes of 8 bits, all six ranks in flight at once,
//                            one launch per pass (histogram + last-block digit pick).
#pragma once
#include <cuda_runtime.h>

#include "ctx.h"
#include "numerics.cuh"

namespace clx {

constexpr uint32_t SELECT_SMALL_MAX = 16384;
constexpr int SELECT_TARGETS = 6;

struct SelectPlan {            // ranks (0-based order statistics) and interpolation data, computed on the host
    unsigned long long rank[SELECT_TARGETS];  // med_lo, med_hi, q1_a, q1_b, q3_a, q3_b
    double frac_q1, frac_q3;                  // h - trunc(h)
    int q1_mode, q3_mode;                     // 0 interpolate, 1 = min, 2 = max   (statrs edge rules)
    int median_even;
    unsigned long long n;
};

// x = value at rank[], in plan order -> DeviceStats (same fp expressions as statrs / main.rs:328-329)
__device__ __forceinline__ void finish_stats(const double* x, const SelectPlan& p, double iqr_multiple,
                                             DeviceStats* st) {
    double median;
    if (p.n == 0) median = nan("");
    else if (p.median_even) median = (x[0] + x[1]) / 2.0;
    else median = x[1];
    double q1, q3;
    if (p.n == 0) { q1 = nan(""); q3 = nan(""); }
    else {
        q1 = p.q1_mode == 0 ? x[2] + p.frac_q1 * (x[3] - x[2]) : x[2];
        q3 = p.q3_mode == 0 ? x[4] + p.frac_q3 * (x[5] - x[4]) : x[4];
    }
    const double iqr = q3 - q1;
    st->median = median;
    st->q1 = q1;
    st->q3 = q3;
    st->iqr = iqr;
    st->threshold = q1 - iqr_multiple * iqr;
}

// ---- small N: bitonic sort in shared memory, single CTA of 1024 threads
__global__ void __launch_bounds__(1024) k_select_small(const double* __restrict__ vals, uint32_t n, uint32_t n_pow2,
                                                       SelectPlan plan, double iqr_multiple, DeviceStats* __restrict__ st) {
    extern __shared__ unsigned long long skeys[];
    for (uint32_t i = threadIdx.x; i < n_pow2; i += blockDim.x)
        skeys[i] = i < n ? f64_to_key(vals[i]) : 0xFFFFFFFFFFFFFFFFull;
    __syncthreads();
    for (uint32_t k = 2; k <= n_pow2; k <<= 1) {
        for (uint32_t j = k >> 1; j > 0; j >>= 1) {
            for (uint32_t t = threadIdx.x; t < (n_pow2 >> 1); t += blockDim.x) {
                // index of the lower element of the t-th compare-exchange pair at distance j
                const uint32_t lo = ((t & ~(j - 1)) << 1) | (t & (j - 1));
                const uint32_t hi = lo | j;
                const bool up = (lo & k) == 0;
                const unsigned long long a = skeys[lo], b = skeys[hi];
                if ((a > b) == up) {
                    skeys[lo] = b;
                    skeys[hi] = a;
                }
            }
            __syncthreads();
        }
    }
    if (threadIdx.x == 0) {
        double x[SELECT_TARGETS];
#pragma unroll
        for (int t = 0; t < SELECT_TARGETS; ++t) x[t] = n ? key_to_f64(skeys[plan.rank[t]]) : 0.0;
        finish_stats(x, plan, iqr_multiple, st);
    }
}

// ---- large N: radix select.  state[t] = {prefix, remaining rank}; hist[pass][t][256]; ticket[pass].
struct SelectState {
    unsigned long long prefix[SELECT_TARGETS];
    unsigned long long rank[SELECT_TARGETS];
};

__global__ void k_select_keys(const double* __restrict__ vals, uint64_t n, unsigned long long* __restrict__ keys) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
        keys[i] = f64_to_key(vals[i]);
}

__global__ void __launch_bounds__(512) k_select_pass(const unsigned long long* __restrict__ keys, uint64_t n, int pass,
                                                     SelectState* __restrict__ state, unsigned int* __restrict__ hist,
                                                     unsigned int* __restrict__ ticket, SelectPlan plan,
                                                     double iqr_multiple, DeviceStats* __restrict__ st) {
    __shared__ unsigned int sh[SELECT_TARGETS][256];
    __shared__ unsigned long long spre[SELECT_TARGETS];
    __shared__ int sgrp[SELECT_TARGETS];  // first target with the same prefix (histograms are shared)
    __shared__ bool last;
    const int shift = 56 - 8 * pass;
    for (int i = threadIdx.x; i < SELECT_TARGETS * 256; i += blockDim.x) (&sh[0][0])[i] = 0;
    if (threadIdx.x < SELECT_TARGETS) spre[threadIdx.x] = state->prefix[threadIdx.x];
    __syncthreads();
    if (threadIdx.x < SELECT_TARGETS) {
        int g = threadIdx.x;
        for (int t = threadIdx.x - 1; t >= 0; --t)
            if (spre[t] == spre[threadIdx.x]) g = t;
        sgrp[threadIdx.x] = g;
    }
    __syncthreads();
    // the high (8 * pass) bits of a candidate must equal the target's prefix
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        const unsigned long long k = keys[i];
        const unsigned long long hi = pass == 0 ? 0ull : (k >> (shift + 8));
        const unsigned int d = (unsigned int)(k >> shift) & 0xFFu;
#pragma unroll
        for (int t = 0; t < SELECT_TARGETS; ++t)
            if (sgrp[t] == t && hi == spre[t]) atomicAdd(&sh[t][d], 1u);
    }
    __syncthreads();
    unsigned int* gh = hist + (size_t)pass * SELECT_TARGETS * 256;
    for (int i = threadIdx.x; i < SELECT_TARGETS * 256; i += blockDim.x) {
        const unsigned int v = (&sh[0][0])[i];
        if (v) atomicAdd(&gh[i], v);
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) last = atomicAdd(&ticket[pass], 1u) == gridDim.x - 1;
    __syncthreads();
    if (!last) return;
    __threadfence();
    if (threadIdx.x < SELECT_TARGETS) {
        const int t = threadIdx.x;
        const volatile unsigned int* h = gh + (size_t)sgrp[t] * 256;
        unsigned long long r = state->rank[t];
        unsigned int d = 0;
        for (; d < 255; ++d) {
            const unsigned int c = h[d];
            if (r < c) break;
            r -= c;
        }
        state->prefix[t] = (spre[t] << 8) | d;
        state->rank[t] = r;
        spre[t] = (spre[t] << 8) | d;
    }
    __syncthreads();
    if (pass == 7 && threadIdx.x == 0) {
        double x[SELECT_TARGETS];
#pragma unroll
        for (int t = 0; t < SELECT_TARGETS; ++t) x[t] = key_to_f64(spre[t]);
        finish_stats(x, plan, iqr_multiple, st);
    }
}

// ------------------------------------------------------------------------------------------------
// Posterior pass, per cell (main.rs:266-278) + assignment rule (main.rs:143-169).
// ll3 = [minority | majority | doublet] log likelihoods, each n_cells long.
struct PosteriorParams {
    double log_prior_minority, log_prior_majority, log_prior_doublet;  // main.rs:259-265 (computed on the host)
    double posterior_threshold;
    uint32_t min_loci_used;
};
__global__ void __launch_bounds__(256) k_posterior(const double* __restrict__ ll3, uint32_t n_cells,
                                                   const uint64_t* __restrict__ row_ptr,
                                                   const uint32_t* __restrict__ excl_bits, PosteriorParams p,
                                                   double* __restrict__ post /*[posterior | doublet]*/,
                                                   uint8_t* __restrict__ out_u8 /*[label | anomaly | qual]*/) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_cells) return;
    const double ll_min = ll3[i], ll_maj = ll3[(size_t)n_cells + i], ll_dbl = ll3[2 * (size_t)n_cells + i];
    const double log_num = p.log_prior_minority + ll_min;
    double log_den = logsumexp2(log_num, p.log_prior_majority + ll_maj);
    const double log_dbl_num = p.log_prior_doublet + ll_dbl;
    log_den = logsumexp2(log_den, log_dbl_num);
    const double posterior = exp(log_num - log_den);
    const double dposterior = exp(log_dbl_num - log_den);
    post[i] = posterior;
    post[(size_t)n_cells + i] = dposterior;
    uint8_t lab = 3;
    if (posterior > p.posterior_threshold) lab = 0;
    else if (1.0 - posterior > p.posterior_threshold) lab = 1;
    if (dposterior > 0.5) lab = 2;
    if (row_ptr[i + 1] - row_ptr[i] < (uint64_t)p.min_loci_used) lab = 3;
    const bool ex = (excl_bits[i >> 5] >> (i & 31)) & 1u;
    const double pm = fmax(posterior, 1.0 - posterior);
    double q = -10.0 * log10(1.0 - pm);
    q = fmin(q, 255.0);
    out_u8[i] = lab;
    out_u8[(size_t)n_cells + i] = ex ? 0 : 1;
    out_u8[2 * (size_t)n_cells + i] = (q != q || q < 0.0) ? 0 : (uint8_t)q;
}

// Final per-locus allele counts over ALL loci (load_mtx_final, load_data.rs:109-132) from the retained COO.
__global__ void k_final_counts(const uint32_t* __restrict__ locus, const uint32_t* __restrict__ cell,
                               const uint32_t* __restrict__ ref, const uint32_t* __restrict__ alt, uint64_t nnz0,
                               const uint32_t* __restrict__ excl_bits, unsigned long long* __restrict__ mn,
                               unsigned long long* __restrict__ mj) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < nnz0; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t c = cell[i];
        unsigned long long* dst = ((excl_bits[c >> 5] >> (c & 31)) & 1u) ? mn : mj;
        if (ref[i]) atomicAdd(&dst[2 * (size_t)locus[i]], (unsigned long long)ref[i]);
        if (alt[i]) atomicAdd(&dst[2 * (size_t)locus[i] + 1], (unsigned long long)alt[i]);
    }
}

}  // namespace clx

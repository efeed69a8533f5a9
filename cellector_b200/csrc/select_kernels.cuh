// select_kernels.cuh — exact order statistics for the IQR threshold (main.rs:324-329) and the posterior
// elementwise pass (main.rs:259-278 + assignment rule main.rs:141-170).  Included by em.cu only.
//
// statrs 0.16.0 Data::{median, lower_quartile, upper_quartile} (R-8 quantile, h = (N + 1/3) tau + 1/3) need at
// most six order statistics.  They are selected exactly on order-preserving u64 keys:
//   * up to 16 keys per thread x 1024 threads x one CTA per SM (2.4 M values): k_select_coop, a cooperative grid
//     with the keys held in registers, which also applies the exclusion rule — one launch;
//   * larger N: MSB-first radix select, 8 passes of 8 bits, all six ranks in flight at once, one launch per pass
//     (histogram + last-block digit pick), then k_exclude.
#pragma once
#include <cooperative_groups.h>
#include <cuda_runtime.h>

#include "ctx.h"
#include "numerics.cuh"
#include "p2p_dev.cuh"

namespace clx {

constexpr uint32_t SELECT_SMALL_MAX = 16384;  // keys of one CTA: 1024 threads x at most 16 keys in registers
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

// ---- register-resident select: every thread keeps its keys in registers.  The radix select starts below the bits
// all keys share (normalized log likelihoods cluster: sign, exponent and leading mantissa bits are common), takes
// 8 bits per pass with all six ranks in flight (targets that still share a prefix share a histogram) and drops a
// key as soon as it matches no target prefix, so later passes touch a handful of keys.  The same launch then
// applies the exclusion rule (main.rs:330-335) to this rank's cells.
// The few keys left after the radix passes are sorted in shared memory (bitonic network over SELECT_LIST_PAD
// slots, pads = ~0) and target t is read at (start of its prefix group) + (its remaining rank): the keys of one
// prefix group are contiguous once sorted.  Called by every thread of the CTA; results land in spre[].
constexpr unsigned int SELECT_LIST_PAD = 256;
__device__ __forceinline__ void select_from_list(unsigned long long* __restrict__ s_list, unsigned int nl, int hi_bit,
                                                 unsigned long long* __restrict__ spre,
                                                 const unsigned long long* __restrict__ srank) {
    const unsigned int tid = threadIdx.x;
    // warp 0 alone runs the network (__syncwarp between steps) over the next power of two >= nl: the usual case is a
    // dozen survivors, where 36 CTA-wide barriers would cost more than the whole select
    unsigned int P = 32;
    while (P < nl) P <<= 1;
    if (tid < 32) {
        for (unsigned int i = tid; i < P; i += 32)
            if (i >= nl) s_list[i] = ~0ull;
        __syncwarp();
        for (unsigned int k = 2; k <= P; k <<= 1)
            for (unsigned int j = k >> 1; j > 0; j >>= 1) {
                for (unsigned int i = tid; i < P; i += 32) {
                    const unsigned int o = i ^ j;
                    if (o > i) {
                        const unsigned long long a = s_list[i], b = s_list[o];
                        const bool up = (i & k) == 0;
                        if ((a > b) == up) {
                            s_list[i] = b;
                            s_list[o] = a;
                        }
                    }
                }
                __syncwarp();
            }
    }
    __syncthreads();
    unsigned long long res = 0ull;
    if (tid < SELECT_TARGETS) {
        const unsigned long long lo_key = hi_bit >= 64 ? 0ull : (spre[tid] << hi_bit);
        unsigned int lo = 0, hi = nl;
        while (lo < hi) {
            const unsigned int mid = (lo + hi) >> 1;
            if (s_list[mid] < lo_key) lo = mid + 1;
            else hi = mid;
        }
        const unsigned long long at = (unsigned long long)lo + srank[tid];
        res = s_list[at < P ? at : P - 1];
    }
    __syncthreads();
    if (tid < SELECT_TARGETS) spre[tid] = res;
    __syncthreads();
}

constexpr int SELECT_SMALL_THREADS = 1024;
constexpr int SELECT_R = 16;
constexpr unsigned int SELECT_RANK_CAP = 192;  // finish by direct ranking once this few keys are left
static_assert(SELECT_SMALL_MAX == SELECT_SMALL_THREADS * SELECT_R, "keys per thread");
static_assert(SELECT_RANK_CAP <= SELECT_LIST_PAD, "survivor list fits the sort network");

// ---- up to SELECT_SMALL_MAX keys per CTA x one CTA per SM: a register-resident radix select spread over a
// cooperative grid (a one-CTA version measured slower even at 10 k values: 33 vs 28 us).  Every CTA keeps its keys in registers; per pass the CTA histograms are added
// into a global histogram (one region per pass, zeroed by the host), the grid synchronises, and every CTA picks
// the digits redundantly from the global histogram — one launch for the whole select, ~one grid barrier per
// pass.  Used when cells are sharded (the gathered values of all ranks are selected on every rank, so all ranks
// get the identical threshold) and on one GPU beyond SELECT_SMALL_MAX cells.  The exclusion rule
// (main.rs:330-335) for this rank's cells runs in the same launch.
constexpr int SELECT_COOP_HIST = 8 * SELECT_TARGETS * 256;  // u32 words
constexpr int SELECT_COOP_SCRATCH = SELECT_COOP_HIST + 8;   // + {or, nor} (2 x u64) + n_list + pad
// CL = true: the CTAs form ONE thread-block cluster (at most 16 CTAs, i.e. 16 k / 64 k keys at R = 1 / 4) and meet at
// the hardware cluster barrier (barrier.cluster arrive.release / wait.acquire) instead of the cooperative grid's
// barrier in global memory: the 10-k-cell select of C2 is four to five barriers and little else.
template <int R, int T, bool CL = false>
__global__ void __launch_bounds__(T) k_select_coop(
    const double* __restrict__ vals, uint64_t n, SelectPlan plan, const IterParams* __restrict__ prm,
    DeviceStats* __restrict__ st, unsigned int* __restrict__ scratch, unsigned long long* __restrict__ glist,
    const double* __restrict__ norm_local, uint32_t n_cells_local, uint32_t* __restrict__ bits,
    unsigned char* const* __restrict__ win, int rank, int world, size_t off_counters) {
    auto sync_all = [] {
        if constexpr (CL) cooperative_groups::this_cluster().sync();
        else cooperative_groups::this_grid().sync();
    };
    // sharded over peer memory: barrier 0 — every rank has stored its normalized values into this rank's window
    if (win != nullptr) p2p_sync_head(win, rank, world, 0, off_counters, &st->pad);
    __shared__ unsigned int sh[SELECT_TARGETS][256];
    __shared__ unsigned long long spre[SELECT_TARGETS], srank[SELECT_TARGETS];
    __shared__ int sgrp[SELECT_TARGETS];
    __shared__ unsigned long long s_or[32], s_nor[32];
    __shared__ unsigned int scnt[SELECT_TARGETS], s_total;
    __shared__ int s_nlead, s_leadt[SELECT_TARGETS];
    __shared__ unsigned long long s_leadpre[SELECT_TARGETS], s_result[SELECT_TARGETS];
    __shared__ unsigned long long s_list[SELECT_LIST_PAD];
    __shared__ double sthr;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned long long* g_or = reinterpret_cast<unsigned long long*>(scratch + SELECT_COOP_HIST);
    unsigned int* g_nlist = scratch + SELECT_COOP_HIST + 4;
    unsigned long long key[R];
    unsigned int valid = 0;
    unsigned long long k_or = 0ull, k_nor = 0ull;
#pragma unroll
    for (int j = 0; j < R; ++j) {
        const uint64_t i = ((uint64_t)blockIdx.x * R + j) * T + tid;
        key[j] = 0ull;
        if (i < n) {
            key[j] = f64_to_key(vals[i]);
            valid |= 1u << j;
            k_or |= key[j];
            k_nor |= ~key[j];
        }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        k_or |= __shfl_xor_sync(0xffffffffu, k_or, off);
        k_nor |= __shfl_xor_sync(0xffffffffu, k_nor, off);
    }
    if (lane == 0) { s_or[warp] = k_or; s_nor[warp] = k_nor; }
    if (tid < SELECT_TARGETS) srank[tid] = plan.rank[tid];
    __syncthreads();
    if (tid == 0) {
        for (int w = 1; w < T / 32; ++w) { k_or |= s_or[w]; k_nor |= s_nor[w]; }
        atomicOr(&g_or[0], k_or);
        atomicOr(&g_or[1], k_nor);
    }
    sync_all();
    k_or = __ldcg(&g_or[0]);
    k_nor = __ldcg(&g_or[1]);
    const unsigned long long diff = k_or & k_nor;  // bits that are 1 in some key and 0 in another
    int hi_bit = diff ? 64 - __clzll((long long)diff) : 0;
    if (tid < SELECT_TARGETS) {
        spre[tid] = hi_bit >= 64 ? 0ull : (k_or >> hi_bit);
        scnt[tid] = (unsigned int)(n > 0xFFFFFFFFull ? 0xFFFFFFFFull : n);
    }
    if (tid == 0) s_total = (unsigned int)(n > 0xFFFFFFFFull ? 0xFFFFFFFFull : n);
    __syncthreads();
    int pass = 0;
    while (hi_bit > 0 && s_total > SELECT_RANK_CAP) {
        const int lo_bit = hi_bit > 8 ? hi_bit - 8 : 0;
        const unsigned int dmask = (1u << (hi_bit - lo_bit)) - 1u;
        unsigned int* gh = scratch + (size_t)pass * SELECT_TARGETS * 256;
        for (int i = tid; i < SELECT_TARGETS * 256; i += T) (&sh[0][0])[i] = 0;
        if (tid < SELECT_TARGETS) {
            unsigned long long p[SELECT_TARGETS];
#pragma unroll
            for (int u = 0; u < SELECT_TARGETS; ++u) p[u] = spre[u];
            int g = tid;
#pragma unroll
            for (int u = SELECT_TARGETS - 1; u >= 0; --u)
                if (u < tid && p[u] == p[tid]) g = u;
            sgrp[tid] = g;
            int pos = 0, nl = 0;
#pragma unroll
            for (int u = 0; u < SELECT_TARGETS; ++u) {
                bool lead = true;
#pragma unroll
                for (int v = 0; v < SELECT_TARGETS; ++v) lead = lead && !(v < u && p[v] == p[u]);
                pos += (lead && u < tid) ? 1 : 0;
                nl += lead ? 1 : 0;
            }
            if (g == tid) { s_leadt[pos] = tid; s_leadpre[pos] = p[tid]; }
            if (tid == 0) s_nlead = nl;
        }
        __syncthreads();
        const int nlead = s_nlead;
        if (valid) {
            unsigned long long lpre[SELECT_TARGETS];
            int lrow[SELECT_TARGETS];
#pragma unroll
            for (int k = 0; k < SELECT_TARGETS; ++k) {
                lpre[k] = k < nlead ? s_leadpre[k] : ~0ull;
                lrow[k] = k < nlead ? s_leadt[k] : 0;
            }
#pragma unroll
            for (int j = 0; j < R; ++j) {
                if (!((valid >> j) & 1u)) continue;
                const unsigned long long pk = hi_bit >= 64 ? 0ull : (key[j] >> hi_bit);
                const unsigned int d = (unsigned int)(key[j] >> lo_bit) & dmask;
#pragma unroll
                for (int k = 0; k < SELECT_TARGETS; ++k)
                    if (k < nlead && pk == lpre[k]) atomicAdd(&sh[lrow[k]][d], 1u);
            }
        }
        __syncthreads();
        for (int i = tid; i < SELECT_TARGETS * 256; i += T) {
            const unsigned int v = (&sh[0][0])[i];
            if (v) atomicAdd(&gh[i], v);
        }
        sync_all();
        for (int i = tid; i < SELECT_TARGETS * 256; i += T) (&sh[0][0])[i] = __ldcg(&gh[i]);
        __syncthreads();
        if (warp < SELECT_TARGETS) {
            const int t = warp;
            const unsigned int* h = sh[sgrp[t]];
            unsigned int c[8], tot = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                c[j] = h[lane * 8 + j];
                tot += c[j];
            }
            unsigned int incl = tot;
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                const unsigned int o = __shfl_up_sync(0xffffffffu, incl, off);
                if (lane >= off) incl += o;
            }
            const unsigned long long r = srank[t];
            const unsigned long long excl = incl - tot;
            const bool here = r >= excl && r < (unsigned long long)incl;
            const unsigned int who = __ballot_sync(0xffffffffu, here);
            const int src = who ? __ffs(who) - 1 : 31;
            __syncwarp();
            if (lane == src) {
                unsigned long long rr = r - excl;
                int j = 0;
                for (; j < 7; ++j) {
                    if (rr < c[j]) break;
                    rr -= c[j];
                }
                spre[t] = (spre[t] << (hi_bit - lo_bit)) | (unsigned long long)(lane * 8 + j);
                srank[t] = rr;
                scnt[t] = c[j];
            }
        }
        __syncthreads();
        if (tid == 0) s_total = 0;
        __syncthreads();
        if (tid < SELECT_TARGETS) {
            bool first = true;
#pragma unroll
            for (int u = 0; u < SELECT_TARGETS; ++u) first = first && !(u < tid && spre[u] == spre[tid]);
            if (first) atomicAdd(&s_total, scnt[tid]);
        }
        if (valid) {
            unsigned long long pre[SELECT_TARGETS];
#pragma unroll
            for (int t = 0; t < SELECT_TARGETS; ++t) pre[t] = spre[t];
#pragma unroll
            for (int j = 0; j < R; ++j) {
                if (!((valid >> j) & 1u)) continue;
                const unsigned long long pk = key[j] >> lo_bit;
                bool keep = false;
#pragma unroll
                for (int t = 0; t < SELECT_TARGETS; ++t) keep = keep || pk == pre[t];
                if (!keep) valid &= ~(1u << j);
            }
        }
        hi_bit = lo_bit;
        ++pass;
        __syncthreads();
    }
    if (hi_bit > 0 && n > 0) {
        // few keys left over the whole grid: every CTA appends its survivors to one global list, then ranks them
        if (valid) {
#pragma unroll
            for (int j = 0; j < R; ++j)
                if ((valid >> j) & 1u) {
                    const unsigned int slot = atomicAdd(g_nlist, 1u);
                    if (slot < SELECT_RANK_CAP) glist[slot] = key[j];
                }
        }
        sync_all();
        const unsigned int nl = min(__ldcg(g_nlist), SELECT_RANK_CAP);
        for (unsigned int i = tid; i < nl; i += T) s_list[i] = __ldcg(&glist[i]);
        __syncthreads();
        select_from_list(s_list, nl, hi_bit, spre, srank);
    }
    if (tid == 0) {
        double x[SELECT_TARGETS];
#pragma unroll
        for (int t = 0; t < SELECT_TARGETS; ++t) x[t] = n ? key_to_f64(spre[t]) : 0.0;
        DeviceStats tmp;
        finish_stats(x, plan, prm->iqr_multiple, &tmp);
        sthr = tmp.threshold;
        if (blockIdx.x == 0) {
            st->median = tmp.median;
            st->q1 = tmp.q1;
            st->q3 = tmp.q3;
            st->iqr = tmp.iqr;
            st->threshold = tmp.threshold;
        }
    }
    __syncthreads();
    if (bits != nullptr) {  // new excluded set of this rank's cells, one bitmap word per warp step
        const double thr = sthr;
        const uint32_t n_words = (n_cells_local + 31) / 32;
        unsigned int nn = 0, nr = 0, ne = 0;
        for (uint32_t w = blockIdx.x * (T / 32) + warp; w < n_words; w += gridDim.x * (T / 32)) {
            const uint32_t i = w * 32 + lane;
            const bool ex = i < n_cells_local && norm_local[i] < thr;  // main.rs:331
            const unsigned int word = __ballot_sync(0xffffffffu, ex);
            if (lane == 0) {
                const unsigned int old = bits[w];
                bits[w] = word;
                nn += __popc(word & ~old);
                nr += __popc(old & ~word);
                ne += __popc(word);
            }
        }
        if (lane == 0) {
            if (nn) atomicAdd(&st->n_new, (unsigned long long)nn);
            if (nr) atomicAdd(&st->n_rescued, (unsigned long long)nr);
            if (ne) atomicAdd(&st->n_excluded, (unsigned long long)ne);
        }
    }
}

// ---- large N: radix select.  state[t] = {prefix, remaining rank}; hist[pass][t][256]; ticket[pass].
struct SelectState {
    unsigned long long prefix[SELECT_TARGETS];
    unsigned long long rank[SELECT_TARGETS];
};

__global__ void k_select_keys(const double* __restrict__ vals, uint64_t n, unsigned long long* __restrict__ keys,
                              SelectPlan plan, SelectState* __restrict__ state) {
    if (blockIdx.x == 0 && threadIdx.x < SELECT_TARGETS) {
        state->prefix[threadIdx.x] = 0ull;
        state->rank[threadIdx.x] = plan.rank[threadIdx.x];
    }
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
        keys[i] = f64_to_key(vals[i]);
}

__global__ void __launch_bounds__(512) k_select_pass(const unsigned long long* __restrict__ keys, uint64_t n, int pass,
                                                     SelectState* __restrict__ state, unsigned int* __restrict__ hist,
                                                     unsigned int* __restrict__ ticket, SelectPlan plan,
                                                     const IterParams* __restrict__ prm, DeviceStats* __restrict__ st) {
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
        finish_stats(x, plan, prm->iqr_multiple, st);
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

// soup.cu — souporcell's k-cluster EM on the matrix a context already holds, R restarts at a time.
// Reference: souporcell/souporcell/src/main.rs
//   souporcell_main :62-106 (best of `restarts` runs), EM :109-186 (9 temperature steps, loss-change limit
//   0.01 * cells, at most 1000 iterations), binomial_loss :199-213, log_sum_exp :215-219,
//   normalize_in_log_with_temp :230-242, update_centers_average :273-280, reset_sums_denoms / update_final :244-262.
// The reference runs in f32 with serial sums; here every sum is f64 with a fixed order (lane-strided partials, xor
// tree), so a run is reproducible bit for bit and lands within f32 rounding of the reference (tests/test_soup_gpu.py
// checks 1e-9 against the oracle's f64 statement of the same formulas and the assignments against its f32 one).
//
// Data: the packed CSR (rows = cells) and CSC (columns = used loci) of load.cu, as they are.  souporcell's loader
// differs from cellector's in two places, handled at set-up:
//   * zero-total entries are skipped (main.rs:441): they add exactly 0 to every sum here, so they may stay;
//   * a duplicated (locus, cell) coordinate keeps only its LAST line (HashMap insert, main.rs:421) where cellector keeps
//     both: k_soup_setup_loci finds duplicates (adjacent in a CSC column) and the call is refused;
//   counts above 65534 are clipped in the packed matrices and refused as well.
//
// Restarts are independent EM runs over the SAME matrix, so R of them ("slots") share every pass over it: the entry
// stream is read once per iteration for all R (the reference reads it once per restart and iteration, one restart per
// rayon thread).  R x K <= 16 f64 sums per lane.  The loop state of a slot (temperature step, iteration count, last
// loss, running or finished) lives in device memory and is advanced by the last kernel of an iteration with the
// reference's own conditions, so the host enqueues iterations in chunks without reading the loss back in between; a
// finished slot is skipped by every kernel and keeps its last log-probabilities and centres.
//
// One EM iteration of R slots = three launches on the context stream:
//   k_soup_cells<K, R>  one warp per cell: streams the row (8 B per entry, coalesced), gathers the locus's
//                       R x K x {ln c, ln(1-c)} record (16 R K bytes, L2), R K f64 sums; lane r finishes slot r of the
//                       cell: log-sum-exp, temperature, responsibilities.
//   k_soup_loci<K, R>   one warp per locus: streams the column (8 B per entry), gathers the cell's R x K
//                       responsibilities (8 R K bytes, L2), 2 R K f64 sums; lane (r, q) writes the new centre and logs.
//   k_soup_loss<R>      one CTA per slot, fixed tree over the per-cell log-sum-exps -> the slot's loss, then the
//                       reference's loop conditions (main.rs:138-147, 167-173) -> the slot's next state.
#include <math.h>

#include <vector>

#include "ctx.h"
#include "numerics.cuh"

namespace clx {
namespace {

constexpr int SOUP_TPB = 256;
constexpr int SOUP_MAX_K = 8;
constexpr int SOUP_TEMP_STEPS = 9;  // main.rs:137
constexpr size_t SOUP_BATCH_TABLE_MAX = 48u << 20;  // restarts share passes only while their table is well inside the 126 MB L2

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// per cell, once: sum of ln_binomial(total, alt) as f32 per entry (main.rs:446-447), total_alleles (main.rs:448);
// flags a clipped count
__global__ void k_soup_setup_cells(const RowEntry* __restrict__ csr, const uint64_t* __restrict__ row_ptr, uint32_t n_cells,
                                   const double* __restrict__ lnfact, double* __restrict__ lbc_sum,
                                   float* __restrict__ total_alleles, unsigned int* __restrict__ flags /* [2] */) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t cell = warp; cell < n_cells; cell += n_warps) {
        double s = 0.0;
        unsigned long long tot = 0;
        bool clipped = false;
        for (uint64_t e = row_ptr[cell] + lane, end = row_ptr[cell + 1]; e < end; e += 32) {
            const RowEntry en = csr[e];
            clipped |= en.ref == 0xFFFFu || en.alt == 0xFFFFu;
            const uint32_t n = (uint32_t)en.ref + en.alt;
            if (n) s += (double)(float)ln_binomial((uint64_t)n, (uint64_t)en.alt, lnfact);
            tot += n;
        }
        s = warp_sum(s);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
        if (__any_sync(0xffffffffu, clipped) && lane == 0) atomicOr(&flags[0], 1u);
        if (lane == 0) {
            lbc_sum[cell] = s;
            total_alleles[cell] = (float)tot;
        }
    }
}

// per locus, once: a (locus, cell) coordinate stored twice sits in adjacent CSC slots (cells ascend inside a column)
__global__ void k_soup_setup_loci(const ColEntry* __restrict__ csc, const uint64_t* __restrict__ col_ptr, uint32_t n_loci,
                                  unsigned int* __restrict__ flags) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
    bool dup = false;
    for (uint32_t l = warp; l < n_loci; l += n_warps)
        for (uint64_t j = col_ptr[l] + lane, end = col_ptr[l + 1]; j + 1 < end; j += 32) dup |= csc[j].cell == csc[j + 1].cell;
    if (dup) atomicOr(&flags[1], 1u);
}

// R restarts share a pass; R x K <= 16 accumulators per lane
constexpr int soup_slots(int K) { return K <= 2 ? 8 : 16 / K; }

// loop state of one restart slot (device memory; the host reads it back once per chunk of iterations)
struct SoupSlot {
    double last;    // last_log_loss            main.rs:138
    double total;   // total_log_loss           main.rs:167
    int ts;         // temperature step of the next iteration
    int active;     // 1 while the reference's loops would run another iteration
    uint32_t it;    // iterations               main.rs:125
    uint32_t pad;
};

// may an iteration start at temperature step ts?  `for temp_step in 0..9 { let mut log_loss_change = 10000.0;
// while log_loss_change > limit && iterations < 1000 {` (main.rs:139-144)
__host__ __device__ inline bool soup_step_admits(double limit, uint32_t it) { return 10000.0 > limit && it < 1000u; }

// centres -> the gather record: tab[i] = {ln c, ln(1 - c)}, i = (l * R + r) * K + q
__global__ void k_soup_tables(const double* __restrict__ cen, size_t n, double2* __restrict__ tab) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i < n) {
        const double c = cen[i];
        tab[i] = make_double2(log(c), log(1.0 - c));
    }
}

template <int K, int R>
__global__ void __launch_bounds__(SOUP_TPB)
k_soup_cells(const RowEntry* __restrict__ csr, const uint64_t* __restrict__ row_ptr, uint32_t n_cells,
             const double2* __restrict__ tab, const double* __restrict__ lbc_sum, const float* __restrict__ total_alleles,
             double log_prior, const SoupSlot* __restrict__ st, double* __restrict__ lp_out, double* __restrict__ prob,
             double* __restrict__ lse_out) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
    unsigned act = 0;
#pragma unroll
    for (int r = 0; r < R; ++r) act |= st[r].active ? 1u << r : 0u;
    if (!act) return;
    const int my_ts = lane < R ? st[lane].ts : 0;
    for (uint32_t cell = warp; cell < n_cells; cell += n_warps) {
        double acc[R][K];
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
            for (int q = 0; q < K; ++q) acc[r][q] = 0.0;
        // binomial_loss main.rs:204-209; two entries per lane in flight, added in the same order as one at a time
        uint64_t e = row_ptr[cell] + lane;
        const uint64_t end = row_ptr[cell + 1];
        for (; e + 32 < end; e += 64) {
            const RowEntry en0 = csr[e], en1 = csr[e + 32];
            const double a0 = (double)en0.alt, r0 = (double)en0.ref, a1 = (double)en1.alt, r1 = (double)en1.ref;
            const double2* __restrict__ t0 = tab + (size_t)en0.locus * (R * K);
            const double2* __restrict__ t1 = tab + (size_t)en1.locus * (R * K);
#pragma unroll
            for (int r = 0; r < R; ++r)
                if (act >> r & 1u) {
#pragma unroll
                    for (int q = 0; q < K; ++q) {
                        const double2 v0 = t0[r * K + q], v1 = t1[r * K + q];
                        acc[r][q] += a0 * v0.x + r0 * v0.y;
                        acc[r][q] += a1 * v1.x + r1 * v1.y;
                    }
                }
        }
        if (e < end) {
            const RowEntry en = csr[e];
            const double a = (double)en.alt, rf = (double)en.ref;
            const double2* __restrict__ t = tab + (size_t)en.locus * (R * K);
#pragma unroll
            for (int r = 0; r < R; ++r)
                if (act >> r & 1u) {
#pragma unroll
                    for (int q = 0; q < K; ++q) {
                        const double2 v = t[r * K + q];
                        acc[r][q] += a * v.x + rf * v.y;
                    }
                }
        }
        double mine[K];  // lane r keeps slot r
#pragma unroll
        for (int q = 0; q < K; ++q) mine[q] = 0.0;
#pragma unroll
        for (int r = 0; r < R; ++r)
            if (act >> r & 1u) {
#pragma unroll
                for (int q = 0; q < K; ++q) {
                    const double v = warp_sum(acc[r][q]);
                    if ((int)lane == r) mine[q] = v;
                }
            }
        if (lane < R && (act >> lane & 1u)) {
            const int r = (int)lane;
            const double base = log_prior + lbc_sum[cell];
            double lp[K], mx = -INFINITY;
#pragma unroll
            for (int q = 0; q < K; ++q) {
                lp[q] = base + mine[q];
                lp_out[((size_t)r * n_cells + cell) * K + q] = lp[q];
                mx = fmax(mx, lp[q]);
            }
            double s = 0.0;  // log_sum_exp main.rs:215-219
#pragma unroll
            for (int q = 0; q < K; ++q) s += exp(lp[q] - mx);
            lse_out[(size_t)r * n_cells + cell] = mx + log(s);
            // temperature main.rs:152-153 (f32 like the reference: it is a function of the input only)
            float temp = fmaxf(total_alleles[cell] / (20.0f * (float)(1 << my_ts)), 1.0f);
            if (my_ts == SOUP_TEMP_STEPS - 1) temp = 1.0f;
            double nlp[K], nmx = -INFINITY;  // normalize_in_log_with_temp main.rs:230-242
#pragma unroll
            for (int q = 0; q < K; ++q) {
                nlp[q] = lp[q] / (double)temp;
                nmx = fmax(nmx, nlp[q]);
            }
            double ns = 0.0;
#pragma unroll
            for (int q = 0; q < K; ++q) ns += exp(nlp[q] - nmx);
            const double z = nmx + log(ns);
#pragma unroll
            for (int q = 0; q < K; ++q) prob[((size_t)cell * R + r) * K + q] = exp(nlp[q] - z);
        }
    }
}

// update_centers_average + update_final (main.rs:273-280, 244-251) for one locus per warp.  finalize = false (cells
// sharded over ranks): only the partial sums are written, k_soup_finalize follows the all-reduce.
template <int K, int R>
__global__ void __launch_bounds__(SOUP_TPB)
k_soup_loci(const ColEntry* __restrict__ csc, const uint64_t* __restrict__ col_ptr, uint32_t n_loci,
            const double* __restrict__ prob, const SoupSlot* __restrict__ st, bool finalize, double2* __restrict__ sd,
            double* __restrict__ cen, double2* __restrict__ tab) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
    unsigned act = 0;
#pragma unroll
    for (int r = 0; r < R; ++r) act |= st[r].active ? 1u << r : 0u;
    if (!act) return;
    for (uint32_t l = warp; l < n_loci; l += n_warps) {
        double s[R][K], d[R][K];
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
            for (int q = 0; q < K; ++q) s[r][q] = d[r][q] = 0.0;
        for (uint64_t j = col_ptr[l] + lane, end = col_ptr[l + 1]; j < end; j += 32) {
            const ColEntry en = csc[j];
            const double a = (double)en.alt, n = (double)((uint32_t)en.alt + en.ref);
            const double* __restrict__ p = prob + (size_t)en.cell * (R * K);
#pragma unroll
            for (int r = 0; r < R; ++r)
                if (act >> r & 1u) {
#pragma unroll
                    for (int q = 0; q < K; ++q) {
                        const double pq = p[r * K + q];
                        s[r][q] += pq * a;
                        d[r][q] += pq * n;
                    }
                }
        }
        double my_s = 0.0, my_d = 0.0;  // lane r * K + q keeps that (slot, cluster) pair
#pragma unroll
        for (int r = 0; r < R; ++r)
            if (act >> r & 1u) {
#pragma unroll
                for (int q = 0; q < K; ++q) {
                    const double vs = warp_sum(s[r][q]), vd = warp_sum(d[r][q]);
                    if ((int)lane == r * K + q) {
                        my_s = vs;
                        my_d = vd;
                    }
                }
            }
        if (lane < R * K && (act >> (lane / K) & 1u)) {
            const size_t i = (size_t)l * (R * K) + lane;
            if (finalize) {
                const double u = fmax(fmin((1.0 + my_s) / (2.0 + my_d), (double)0.99f), (double)0.01f);
                cen[i] = u;
                tab[i] = make_double2(log(u), log(1.0 - u));
            } else {
                sd[i] = make_double2(my_s, my_d);
            }
        }
    }
}

// after the all-reduce of the partial sums (cells sharded over ranks); RK = R * K
__global__ void k_soup_finalize(const double2* __restrict__ sd, size_t n, int RK, int K, const SoupSlot* __restrict__ st,
                                double* __restrict__ cen, double2* __restrict__ tab) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i < n && st[(i % RK) / K].active) {
        const double u = fmax(fmin((1.0 + sd[i].x) / (2.0 + sd[i].y), (double)0.99f), (double)0.01f);
        cen[i] = u;
        tab[i] = make_double2(log(u), log(1.0 - u));
    }
}

// the reference's loop conditions after an iteration with this loss (main.rs:167-173, then :139-144)
__device__ inline void soup_advance(SoupSlot* sl, double loss, double limit) {
    const double change = loss - sl->last;
    sl->total = loss;
    sl->last = loss;
    sl->it += 1;
    if (!(change > limit && sl->it < 1000u)) {  // the while loop ends: next temperature step, change = 10000 again
        sl->ts += 1;
        if (sl->ts >= SOUP_TEMP_STEPS || !soup_step_admits(limit, sl->it)) sl->active = 0;
    }
}

// slot blockIdx.x: loss = sum over cells of their log-sum-exp (main.rs:149), fixed order; advance = false (cells
// sharded): the partial loss goes to loss_out[slot] and k_soup_control follows the all-reduce
__global__ void __launch_bounds__(1024) k_soup_loss(const double* __restrict__ lse, uint32_t n_cells, SoupSlot* __restrict__ st,
                                                    double limit, bool advance, double* __restrict__ loss_out) {
    __shared__ double sh[1024];
    const int r = blockIdx.x;
    if (!st[r].active) {
        if (threadIdx.x == 0 && !advance) loss_out[r] = 0.0;
        return;
    }
    double s = 0.0;
    for (uint32_t i = threadIdx.x; i < n_cells; i += 1024) s += lse[(size_t)r * n_cells + i];
    sh[threadIdx.x] = s;
    __syncthreads();
    for (int o = 512; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        if (advance) soup_advance(&st[r], sh[0], limit);
        else loss_out[r] = sh[0];
    }
}
__global__ void k_soup_control(SoupSlot* __restrict__ st, int R, const double* __restrict__ loss, double limit) {
    const int r = threadIdx.x;
    if (r < R && st[r].active) soup_advance(&st[r], loss[r], limit);
}

// [slots][k][L] f32 (the caller's layout per restart, main.rs:354-364) -> cen[(l * R + r) * K + q] f64; slots beyond
// n_used get 0.5 (never read: their state is inactive)
__global__ void k_soup_centers_in(const float* __restrict__ in, int K, int R, int n_used, uint32_t n_loci, double* __restrict__ cen) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i < (size_t)n_loci * R * K) {
        const int q = (int)(i % K), r = (int)((i / K) % R);
        const size_t l = i / ((size_t)R * K);
        cen[i] = r < n_used ? (double)in[((size_t)r * K + q) * n_loci + l] : 0.5;
    }
}
// centres of one slot -> [k][L]
__global__ void k_soup_centers_out(const double* __restrict__ cen, int K, int R, int r, uint32_t n_loci, double* __restrict__ out) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i < (size_t)n_loci * K) {
        const int q = (int)(i % K);
        const size_t l = i / K;
        out[(size_t)q * n_loci + l] = cen[(l * R + r) * K + q];
    }
}

inline unsigned warps_grid(const clx_ctx* c, uint64_t n_items) {  // one warp per item, at most 8 CTAs of 256 threads per SM
    const uint64_t want = (n_items + (SOUP_TPB / 32) - 1) / (SOUP_TPB / 32);
    const uint64_t cap = (uint64_t)c->sm_count * 8;
    return (unsigned)(want < 1 ? 1 : (want > cap ? cap : want));
}
inline unsigned flat_grid(size_t n) { return (unsigned)((n + SOUP_TPB - 1) / SOUP_TPB); }

template <int K, int R>
int soup_run(clx_ctx* c, uint32_t n_restarts, const float* centers_init, double* log_probs, double* centers_out,
             double* losses, uint32_t* iterations, uint32_t* best_out) {
    cudaStream_t s = c->stream;
    const uint32_t N = c->n_cells, L = c->n_loci;
    const size_t LK = (size_t)L * K, NK = (size_t)N * K, LRK = LK * R, NRK = NK * R;
    DevBuf<double> cen, prob, lp, lse, lbc_sum, d_loss, best_lp, best_cen;
    DevBuf<double2> tab, sd;
    DevBuf<float> tot, d_init;
    DevBuf<unsigned int> flags;
    DevBuf<SoupSlot> st;
    CLX_CUDA(c, cen.alloc(LRK + 1));
    CLX_CUDA(c, tab.alloc(LRK + 1));
    CLX_CUDA(c, prob.alloc(NRK + 1));
    CLX_CUDA(c, lp.alloc(NRK + 1));
    CLX_CUDA(c, lse.alloc((size_t)N * R + 1));
    CLX_CUDA(c, lbc_sum.alloc((size_t)N + 1));
    CLX_CUDA(c, tot.alloc((size_t)N + 1));
    CLX_CUDA(c, d_init.alloc(LRK + 1));
    CLX_CUDA(c, d_loss.alloc(R));
    CLX_CUDA(c, flags.alloc(2));
    CLX_CUDA(c, st.alloc(R));
    CLX_CUDA(c, best_lp.alloc(NK + 1));
    CLX_CUDA(c, best_cen.alloc(LK + 1));
    if (c->world > 1) CLX_CUDA(c, sd.alloc(LRK + 1));
    // pinned: slot states of a chunk read-back (R x 32 B) and the refusal flags
    if (!c->h_soup) CLX_CUDA(c, cudaMallocHost((void**)&c->h_soup, 512));
    SoupSlot* h_st = reinterpret_cast<SoupSlot*>(c->h_soup);
    unsigned int* h_flags = reinterpret_cast<unsigned int*>(c->h_soup + 48);
    static_assert(sizeof(SoupSlot) == 32 && R * sizeof(SoupSlot) <= 48 * sizeof(double), "slot read-back fits the pinned block");

    // ---- set-up: per-cell constants, refusals
    CLX_CUDA(c, cudaMemsetAsync(flags.p, 0, 8, s));
    if (NK) CLX_CUDA(c, cudaMemsetAsync(best_lp.p, 0, NK * 8, s));
    k_soup_setup_cells<<<warps_grid(c, N), SOUP_TPB, 0, s>>>(c->csr.p, c->row_ptr.p, N, c->lnfact.p, lbc_sum.p, tot.p, flags.p);
    k_soup_setup_loci<<<warps_grid(c, L), SOUP_TPB, 0, s>>>(c->csc.p, c->col_ptr.p, L, flags.p);
    c->launches += 2;
    CLX_TRY(comm_allreduce_u32(c, flags.p, 2));  // every rank refuses or none does
    CLX_CUDA(c, cudaMemcpyAsync(h_flags, flags.p, 8, cudaMemcpyDeviceToHost, s));
    CLX_CUDA(c, cudaStreamSynchronize(s));
    CLX_CUDA(c, cudaGetLastError());
    if (h_flags[0]) return fail(c, CLX_ERR_ARG, "clx_soup_em: an allele count above 65534 (clipped in the packed matrix) is not supported");
    if (h_flags[1])
        return fail(c, CLX_ERR_ARG,
                    "clx_soup_em: a (locus, cell) coordinate is stored more than once; souporcell keeps only its last line "
                    "(main.rs:421): deduplicate before clx_load_coo");

    const double log_prior = log((double)(1.0f / (float)K));             // main.rs:122
    const double limit = (double)(0.01f * (float)c->n_cells_global);     // main.rs:136
    double best_loss = -INFINITY;
    uint32_t best = 0;
    bool have_best = false;
    for (uint32_t b0 = 0; b0 < n_restarts; b0 += R) {
        const int used = (int)((n_restarts - b0) < (uint32_t)R ? (n_restarts - b0) : (uint32_t)R);
        // ---- slots of this batch
        for (int r = 0; r < R; ++r) {
            h_st[r].last = -INFINITY;
            h_st[r].total = -INFINITY;
            h_st[r].ts = 0;
            h_st[r].active = (r < used && soup_step_admits(limit, 0)) ? 1 : 0;
            h_st[r].it = 0;
            h_st[r].pad = 0;
        }
        CLX_CUDA(c, cudaMemcpyAsync(st.p, h_st, R * sizeof(SoupSlot), cudaMemcpyHostToDevice, s));
        if (LK) {
            CLX_CUDA(c, cudaMemcpyAsync(d_init.p, centers_init + (size_t)b0 * LK, (size_t)used * LK * sizeof(float), cudaMemcpyDefault, s));
            k_soup_centers_in<<<flat_grid(LRK), SOUP_TPB, 0, s>>>(d_init.p, K, R, used, L, cen.p);
            k_soup_tables<<<flat_grid(LRK), SOUP_TPB, 0, s>>>(cen.p, LRK, tab.p);
            c->launches += 2;
        }
        if (NRK) CLX_CUDA(c, cudaMemsetAsync(lp.p, 0, NRK * 8, s));  // a slot that never iterates (>= 1e6 cells: 10000 > limit fails) reports zeros
        CLX_CUDA(c, cudaStreamSynchronize(s));  // h_st is reused for the read-back below

        // ---- EM main.rs:139-174: iterations are enqueued in chunks; every kernel skips finished slots
        bool running = h_st[0].active != 0;
        uint32_t enqueued = 0;
        while (running) {
            const uint32_t chunk = enqueued == 0 ? (uint32_t)SOUP_TEMP_STEPS : 4u;  // no run ends before one iteration per step
            for (uint32_t i = 0; i < chunk; ++i) {
                k_soup_cells<K, R><<<warps_grid(c, N), SOUP_TPB, 0, s>>>(c->csr.p, c->row_ptr.p, N, tab.p, lbc_sum.p, tot.p, log_prior,
                                                                         st.p, lp.p, prob.p, lse.p);
                if (c->world > 1) {
                    k_soup_loci<K, R><<<warps_grid(c, L), SOUP_TPB, 0, s>>>(c->csc.p, c->col_ptr.p, L, prob.p, st.p, false, sd.p, cen.p, tab.p);
                    k_soup_loss<<<R, 1024, 0, s>>>(lse.p, N, st.p, limit, false, d_loss.p);
                    CLX_TRY(comm_allreduce_f64(c, reinterpret_cast<double*>(sd.p), 2 * LRK));
                    CLX_TRY(comm_allreduce_f64(c, d_loss.p, R));
                    if (LRK) k_soup_finalize<<<flat_grid(LRK), SOUP_TPB, 0, s>>>(sd.p, LRK, R * K, K, st.p, cen.p, tab.p);
                    k_soup_control<<<1, 32, 0, s>>>(st.p, R, d_loss.p, limit);
                    c->launches += 5;
                } else {
                    k_soup_loci<K, R><<<warps_grid(c, L), SOUP_TPB, 0, s>>>(c->csc.p, c->col_ptr.p, L, prob.p, st.p, true, nullptr, cen.p, tab.p);
                    k_soup_loss<<<R, 1024, 0, s>>>(lse.p, N, st.p, limit, true, nullptr);
                    c->launches += 3;
                }
            }
            enqueued += chunk;
            CLX_CUDA(c, cudaMemcpyAsync(h_st, st.p, R * sizeof(SoupSlot), cudaMemcpyDeviceToHost, s));
            CLX_CUDA(c, cudaStreamSynchronize(s));
            running = false;
            for (int r = 0; r < used; ++r) running |= h_st[r].active != 0;
            if (enqueued > 1000u + 8u) return fail(c, CLX_ERR_STATE, "clx_soup_em: loop state did not finish (internal)");
        }
        CLX_CUDA(c, cudaGetLastError());

        // ---- best of the restarts so far (strict >, first wins: main.rs:73, 84)
        int new_best = -1;
        for (int r = 0; r < used; ++r) {
            if (h_st[r].total != h_st[r].total) return fail(c, CLX_ERR_NAN, "clx_soup_em: the loss is NaN");
            if (losses) losses[b0 + r] = h_st[r].total;
            if (iterations) iterations[b0 + r] = h_st[r].it;
            if (!have_best || h_st[r].total > best_loss) {
                have_best = true;
                best_loss = h_st[r].total;
                best = b0 + (uint32_t)r;
                new_best = r;
            }
        }
        if (new_best >= 0) {
            if (NK) CLX_CUDA(c, cudaMemcpyAsync(best_lp.p, lp.p + (size_t)new_best * NK, NK * 8, cudaMemcpyDeviceToDevice, s));
            if (LK) {
                k_soup_centers_out<<<flat_grid(LK), SOUP_TPB, 0, s>>>(cen.p, K, R, new_best, L, best_cen.p);
                c->launches += 1;
            }
        }
    }

    // ---- results
    if (best_out) *best_out = best;
    if (log_probs && NK) CLX_CUDA(c, cudaMemcpyAsync(log_probs, best_lp.p, NK * 8, cudaMemcpyDeviceToHost, s));
    if (centers_out && LK) CLX_CUDA(c, cudaMemcpyAsync(centers_out, best_cen.p, LK * 8, cudaMemcpyDeviceToHost, s));
    CLX_CUDA(c, cudaStreamSynchronize(s));
    return CLX_OK;
}

template <int K = 1>
int soup_dispatch(clx_ctx* c, uint32_t k, uint32_t n, const float* init, double* lp, double* cen, double* losses,
                  uint32_t* its, uint32_t* best) {
    if constexpr (K > SOUP_MAX_K) {
        return fail(c, CLX_ERR_ARG, "clx_soup_em: k must be in [1, 8]");
    } else {
        // A single run gets one-slot kernels (16 K bytes of table per locus instead of 16 R K).  So do the restarts of a
        // matrix whose R-slot table would not stay in L2 (C3: 128 MB for R = 8, and 7.9 ms per restart and iteration
        // against 5.6 ms one at a time): sharing the passes only pays while the gathered records are L2 hits.
        if (k == (uint32_t)K) {
            const size_t table_bytes = (size_t)c->n_loci * soup_slots(K) * K * sizeof(double2);
            return (n == 1 || table_bytes > SOUP_BATCH_TABLE_MAX) ? soup_run<K, 1>(c, n, init, lp, cen, losses, its, best)
                                                                  : soup_run<K, soup_slots(K)>(c, n, init, lp, cen, losses, its, best);
        }
        return soup_dispatch<K + 1>(c, k, n, init, lp, cen, losses, its, best);
    }
}

}  // namespace
}  // namespace clx

extern "C" int clx_soup_restarts(clx_ctx* c, uint32_t k, uint32_t n_restarts, const float* centers_init, double* log_probs,
                                 double* centers_out, double* losses, uint32_t* iterations, uint32_t* best) {
    using namespace clx;
    if (!c) return CLX_ERR_ARG;
    cudaSetDevice(c->device);
    alloc_stream() = c->stream;
    if (!c->loaded) return fail(c, CLX_ERR_STATE, "clx_soup_em before clx_load_coo");
    if (k < 1 || k > (uint32_t)SOUP_MAX_K) return fail(c, CLX_ERR_ARG, "clx_soup_em: k must be in [1, 8]");
    if (n_restarts < 1) return fail(c, CLX_ERR_ARG, "clx_soup_em: no restarts");
    if (!centers_init && c->n_loci) return fail(c, CLX_ERR_ARG, "clx_soup_em: null initial centers");
    NvtxRange nvtx("clx:soup_em");
    return soup_dispatch<1>(c, k, n_restarts, centers_init, log_probs, centers_out, losses, iterations, best);
}

extern "C" int clx_soup_em(clx_ctx* c, uint32_t k, const float* centers_init, double* log_probs, double* centers_out,
                           double* total_log_loss, uint32_t* iterations) {
    return clx_soup_restarts(c, k, 1, centers_init, log_probs, centers_out, total_log_loss, iterations, nullptr);
}

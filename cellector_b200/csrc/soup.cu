// soup.cu — souporcell's k-cluster EM on the matrix a context already holds.
// Reference: souporcell/souporcell/src/main.rs
//   EM :109-186 (9 temperature steps, loss-change limit 0.01 * cells, at most 1000 iterations),
//   binomial_loss :199-213, log_sum_exp :215-219, normalize_in_log_with_temp :230-242,
//   update_centers_average :273-280, reset_sums_denoms / update_final :244-262.
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
// One EM iteration = three launches on the context stream:
//   k_soup_cells<K>  one warp per cell: streams the row (8 B per entry, coalesced), gathers the locus's K x {ln c, ln(1-c)}
//                    record (16 K bytes, L2 resident), K f64 sums; lane 0 finishes the cell: log-sum-exp, temperature,
//                    responsibilities.                                   bytes per entry: 8 (+ 16 K from L2)
//   k_soup_loss      one CTA, fixed tree over the per-cell log-sum-exps -> the iteration's loss (deterministic)
//   k_soup_loci<K>   one warp per locus: streams the column (8 B per entry), gathers the cell's K responsibilities
//                    (8 K bytes, L2), 2 K f64 sums; lane 0 writes the new centre and its two logs.
// The loss goes to the host (8 bytes, one stream synchronise) because the reference's loop condition needs it.
#include <math.h>

#include <vector>

#include "ctx.h"
#include "numerics.cuh"

namespace clx {
namespace {

constexpr int SOUP_TPB = 256;
constexpr int SOUP_MAX_K = 8;
constexpr int SOUP_TEMP_STEPS = 9;  // main.rs:137

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

// centres -> the gather record of a locus: tab[(l * K + q) * 2 + {0, 1}] = {ln c, ln(1 - c)}
__global__ void k_soup_tables(const double* __restrict__ cen, size_t n, double* __restrict__ tab) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i < n) {
        const double c = cen[i];
        tab[2 * i] = log(c);
        tab[2 * i + 1] = log(1.0 - c);
    }
}

template <int K>
__global__ void __launch_bounds__(SOUP_TPB)
k_soup_cells(const RowEntry* __restrict__ csr, const uint64_t* __restrict__ row_ptr, uint32_t n_cells,
             const double2* __restrict__ tab, const double* __restrict__ lbc_sum, const float* __restrict__ total_alleles,
             double log_prior, int temp_step, double* __restrict__ lp_out, double* __restrict__ prob,
             double* __restrict__ lse_out) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t cell = warp; cell < n_cells; cell += n_warps) {
        double acc[K];
#pragma unroll
        for (int q = 0; q < K; ++q) acc[q] = 0.0;
        for (uint64_t e = row_ptr[cell] + lane, end = row_ptr[cell + 1]; e < end; e += 32) {  // binomial_loss main.rs:204-209
            const RowEntry en = csr[e];
            const double a = (double)en.alt, r = (double)en.ref;
            const double2* __restrict__ t = tab + (size_t)en.locus * K;
#pragma unroll
            for (int q = 0; q < K; ++q) {
                const double2 v = t[q];
                acc[q] += a * v.x + r * v.y;
            }
        }
#pragma unroll
        for (int q = 0; q < K; ++q) acc[q] = warp_sum(acc[q]);
        if (lane == 0) {
            const double base = log_prior + lbc_sum[cell];
            double lp[K], mx = -INFINITY;
#pragma unroll
            for (int q = 0; q < K; ++q) {
                lp[q] = base + acc[q];
                lp_out[(size_t)cell * K + q] = lp[q];
                mx = fmax(mx, lp[q]);
            }
            double s = 0.0;  // log_sum_exp main.rs:215-219
#pragma unroll
            for (int q = 0; q < K; ++q) s += exp(lp[q] - mx);
            lse_out[cell] = mx + log(s);
            // temperature main.rs:152-153 (f32 like the reference: it is a function of the input only)
            float temp = fmaxf(total_alleles[cell] / (20.0f * (float)(1 << temp_step)), 1.0f);
            if (temp_step == SOUP_TEMP_STEPS - 1) temp = 1.0f;
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
            for (int q = 0; q < K; ++q) prob[(size_t)cell * K + q] = exp(nlp[q] - z);
        }
    }
}

// loss = sum over cells of their log-sum-exp (main.rs:149), one CTA, fixed order
__global__ void __launch_bounds__(1024) k_soup_loss(const double* __restrict__ lse, uint32_t n_cells, double* __restrict__ out) {
    __shared__ double sh[1024];
    double s = 0.0;
    for (uint32_t i = threadIdx.x; i < n_cells; i += 1024) s += lse[i];
    sh[threadIdx.x] = s;
    __syncthreads();
    for (int o = 512; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) *out = sh[0];
}

// update_centers_average + update_final (main.rs:273-280, 244-251) for one locus per warp.  finalize = false (cells
// sharded over ranks): only the partial sums are written, k_soup_finalize follows the all-reduce.
template <int K>
__global__ void __launch_bounds__(SOUP_TPB)
k_soup_loci(const ColEntry* __restrict__ csc, const uint64_t* __restrict__ col_ptr, uint32_t n_loci,
            const double* __restrict__ prob, bool finalize, double* __restrict__ sd, double* __restrict__ cen,
            double* __restrict__ tab) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t l = warp; l < n_loci; l += n_warps) {
        double s[K], d[K];
#pragma unroll
        for (int q = 0; q < K; ++q) s[q] = d[q] = 0.0;
        for (uint64_t j = col_ptr[l] + lane, end = col_ptr[l + 1]; j < end; j += 32) {
            const ColEntry en = csc[j];
            const double a = (double)en.alt, n = (double)((uint32_t)en.alt + en.ref);
            const double* __restrict__ p = prob + (size_t)en.cell * K;
#pragma unroll
            for (int q = 0; q < K; ++q) {
                const double pq = p[q];
                s[q] += pq * a;
                d[q] += pq * n;
            }
        }
#pragma unroll
        for (int q = 0; q < K; ++q) {
            s[q] = warp_sum(s[q]);
            d[q] = warp_sum(d[q]);
        }
        if (lane == 0) {
#pragma unroll
            for (int q = 0; q < K; ++q) {
                const size_t i = (size_t)l * K + q;
                if (finalize) {
                    const double u = fmax(fmin((1.0 + s[q]) / (2.0 + d[q]), (double)0.99f), (double)0.01f);
                    cen[i] = u;
                    tab[2 * i] = log(u);
                    tab[2 * i + 1] = log(1.0 - u);
                } else {
                    sd[2 * i] = s[q];
                    sd[2 * i + 1] = d[q];
                }
            }
        }
    }
}

__global__ void k_soup_finalize(const double* __restrict__ sd, size_t n, double* __restrict__ cen, double* __restrict__ tab) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i < n) {
        const double u = fmax(fmin((1.0 + sd[2 * i]) / (2.0 + sd[2 * i + 1]), (double)0.99f), (double)0.01f);
        cen[i] = u;
        tab[2 * i] = log(u);
        tab[2 * i + 1] = log(1.0 - u);
    }
}

// [k][L] f32 (the caller's layout, main.rs:354-364) -> [L][k] f64
__global__ void k_soup_centers_in(const float* __restrict__ in, uint32_t k, uint32_t n_loci, double* __restrict__ cen) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i < (size_t)k * n_loci) {
        const uint32_t l = (uint32_t)(i / k), q = (uint32_t)(i % k);
        cen[i] = (double)in[(size_t)q * n_loci + l];
    }
}
__global__ void k_soup_centers_out(const double* __restrict__ cen, uint32_t k, uint32_t n_loci, double* __restrict__ out) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i < (size_t)k * n_loci) {
        const uint32_t l = (uint32_t)(i / k), q = (uint32_t)(i % k);
        out[(size_t)q * n_loci + l] = cen[i];
    }
}

inline unsigned warps_grid(const clx_ctx* c, uint64_t n_items) {  // one warp per item, at most 8 CTAs of 256 threads per SM
    const uint64_t want = (n_items + (SOUP_TPB / 32) - 1) / (SOUP_TPB / 32);
    const uint64_t cap = (uint64_t)c->sm_count * 8;
    return (unsigned)(want < 1 ? 1 : (want > cap ? cap : want));
}

template <int K>
int soup_run(clx_ctx* c, const float* centers_init, double* log_probs, double* centers_out, double* total_log_loss,
             uint32_t* iterations) {
    cudaStream_t s = c->stream;
    const uint32_t N = c->n_cells, L = c->n_loci;
    const size_t LK = (size_t)L * K, NK = (size_t)N * K;
    DevBuf<double> cen, tab, sd, prob, lp, lse, lbc_sum, d_loss, out_t;
    DevBuf<float> tot, d_init;
    DevBuf<unsigned int> flags;
    CLX_CUDA(c, cen.alloc(LK + 1));
    CLX_CUDA(c, tab.alloc(2 * LK + 2));
    CLX_CUDA(c, prob.alloc(NK + 1));
    CLX_CUDA(c, lp.alloc(NK + 1));
    CLX_CUDA(c, lse.alloc((size_t)N + 1));
    CLX_CUDA(c, lbc_sum.alloc((size_t)N + 1));
    CLX_CUDA(c, tot.alloc((size_t)N + 1));
    CLX_CUDA(c, d_init.alloc(LK + 1));
    CLX_CUDA(c, d_loss.alloc(2));
    CLX_CUDA(c, flags.alloc(2));
    if (c->world > 1) CLX_CUDA(c, sd.alloc(2 * LK + 2));
    if (!c->h_soup) CLX_CUDA(c, cudaMallocHost((void**)&c->h_soup, 64));  // pinned: the per-iteration loss read-back
    CLX_CUDA(c, cudaMemsetAsync(flags.p, 0, 8, s));
    if (LK) CLX_CUDA(c, cudaMemcpyAsync(d_init.p, centers_init, LK * sizeof(float), cudaMemcpyDefault, s));

    // ---- set-up: per-cell constants, refusals
    k_soup_setup_cells<<<warps_grid(c, N), SOUP_TPB, 0, s>>>(c->csr.p, c->row_ptr.p, N, c->lnfact.p, lbc_sum.p, tot.p, flags.p);
    k_soup_setup_loci<<<warps_grid(c, L), SOUP_TPB, 0, s>>>(c->csc.p, c->col_ptr.p, L, flags.p);
    if (LK) {
        k_soup_centers_in<<<(unsigned)((LK + SOUP_TPB - 1) / SOUP_TPB), SOUP_TPB, 0, s>>>(d_init.p, K, L, cen.p);
        k_soup_tables<<<(unsigned)((LK + SOUP_TPB - 1) / SOUP_TPB), SOUP_TPB, 0, s>>>(cen.p, LK, tab.p);
        c->launches += 2;
    }
    c->launches += 2;
    CLX_TRY(comm_allreduce_u32(c, flags.p, 2));  // every rank refuses or none does
    unsigned int* h_flags = reinterpret_cast<unsigned int*>(c->h_soup + 4);
    CLX_CUDA(c, cudaMemcpyAsync(h_flags, flags.p, 8, cudaMemcpyDeviceToHost, s));
    CLX_CUDA(c, cudaStreamSynchronize(s));
    CLX_CUDA(c, cudaGetLastError());
    if (h_flags[0]) return fail(c, CLX_ERR_ARG, "clx_soup_em: an allele count above 65534 (clipped in the packed matrix) is not supported");
    if (h_flags[1])
        return fail(c, CLX_ERR_ARG,
                    "clx_soup_em: a (locus, cell) coordinate is stored more than once; souporcell keeps only its last line "
                    "(main.rs:421): deduplicate before clx_load_coo");

    // ---- EM main.rs:122-181
    const double log_prior = log((double)(1.0f / (float)K));
    const double limit = (double)(0.01f * (float)c->n_cells_global);
    double last = -INFINITY, total = -INFINITY;
    uint32_t it = 0;
    for (int ts = 0; ts < SOUP_TEMP_STEPS; ++ts) {
        double change = 10000.0;
        while (change > limit && it < 1000) {
            k_soup_cells<K><<<warps_grid(c, N), SOUP_TPB, 0, s>>>(c->csr.p, c->row_ptr.p, N, (const double2*)tab.p, lbc_sum.p, tot.p,
                                                                  log_prior, ts, lp.p, prob.p, lse.p);
            k_soup_loss<<<1, 1024, 0, s>>>(lse.p, N, d_loss.p);
            if (c->world > 1) {
                k_soup_loci<K><<<warps_grid(c, L), SOUP_TPB, 0, s>>>(c->csc.p, c->col_ptr.p, L, prob.p, false, sd.p, cen.p, tab.p);
                CLX_TRY(comm_allreduce_f64(c, sd.p, 2 * LK));
                CLX_TRY(comm_allreduce_f64(c, d_loss.p, 1));
                if (LK) k_soup_finalize<<<(unsigned)((LK + SOUP_TPB - 1) / SOUP_TPB), SOUP_TPB, 0, s>>>(sd.p, LK, cen.p, tab.p);
                c->launches += 1;
            } else {
                k_soup_loci<K><<<warps_grid(c, L), SOUP_TPB, 0, s>>>(c->csc.p, c->col_ptr.p, L, prob.p, true, nullptr, cen.p, tab.p);
            }
            c->launches += 3;
            CLX_CUDA(c, cudaMemcpyAsync(c->h_soup, d_loss.p, 8, cudaMemcpyDeviceToHost, s));
            CLX_CUDA(c, cudaStreamSynchronize(s));
            const double loss = c->h_soup[0];
            total = loss;
            change = loss - last;
            last = loss;
            ++it;
            if (loss != loss) return fail(c, CLX_ERR_NAN, "clx_soup_em: the loss is NaN");
        }
    }
    CLX_CUDA(c, cudaGetLastError());

    // ---- results
    if (total_log_loss) *total_log_loss = total;
    if (iterations) *iterations = it;
    if (log_probs && NK) CLX_CUDA(c, cudaMemcpyAsync(log_probs, lp.p, NK * 8, cudaMemcpyDeviceToHost, s));
    if (centers_out && LK) {
        CLX_CUDA(c, out_t.alloc(LK));
        k_soup_centers_out<<<(unsigned)((LK + SOUP_TPB - 1) / SOUP_TPB), SOUP_TPB, 0, s>>>(cen.p, K, L, out_t.p);
        c->launches += 1;
        CLX_CUDA(c, cudaMemcpyAsync(centers_out, out_t.p, LK * 8, cudaMemcpyDeviceToHost, s));
    }
    CLX_CUDA(c, cudaStreamSynchronize(s));
    return CLX_OK;
}

}  // namespace
}  // namespace clx

extern "C" int clx_soup_em(clx_ctx* c, uint32_t k, const float* centers_init, double* log_probs, double* centers_out,
                           double* total_log_loss, uint32_t* iterations) {
    using namespace clx;
    if (!c) return CLX_ERR_ARG;
    cudaSetDevice(c->device);
    alloc_stream() = c->stream;
    if (!c->loaded) return fail(c, CLX_ERR_STATE, "clx_soup_em before clx_load_coo");
    if (k < 1 || k > (uint32_t)SOUP_MAX_K) return fail(c, CLX_ERR_ARG, "clx_soup_em: k must be in [1, 8]");
    if (!centers_init && c->n_loci) return fail(c, CLX_ERR_ARG, "clx_soup_em: null initial centers");
    NvtxRange nvtx("clx:soup_em");
    switch (k) {
        case 1: return soup_run<1>(c, centers_init, log_probs, centers_out, total_log_loss, iterations);
        case 2: return soup_run<2>(c, centers_init, log_probs, centers_out, total_log_loss, iterations);
        case 3: return soup_run<3>(c, centers_init, log_probs, centers_out, total_log_loss, iterations);
        case 4: return soup_run<4>(c, centers_init, log_probs, centers_out, total_log_loss, iterations);
        case 5: return soup_run<5>(c, centers_init, log_probs, centers_out, total_log_loss, iterations);
        case 6: return soup_run<6>(c, centers_init, log_probs, centers_out, total_log_loss, iterations);
        case 7: return soup_run<7>(c, centers_init, log_probs, centers_out, total_log_loss, iterations);
        default: return soup_run<8>(c, centers_init, log_probs, centers_out, total_log_loss, iterations);
    }
}

// em.cu — host orchestration of one EM iteration (compute_new_excluded, cellector/src/main.rs:308-347), the
// posterior pass (calculate_posteriors, main.rs:228-280) and the final per-locus counts (load_data.rs:109-132).
// Everything runs on ctx->stream; one stream synchronisation per iteration (to read ~100 bytes of scalars).
#include <math.h>
#include <string.h>

#include <stdlib.h>

#include <algorithm>

#include <cub/cub.cuh>

#include "ctx.h"
#include "em_kernels.cuh"
#include "numerics.cuh"
#include "select_kernels.cuh"

namespace clx {
namespace {

constexpr int TPB = 256;

int env_int(const char* name, int dflt) {
    const char* v = getenv(name);
    return (v && *v) ? atoi(v) : dflt;
}

int pick_group(double avg_len) {
    // lanes per row/column: every lane should have about one unrolled batch (SCORE_UNROLL 128-bit loads of two
    // entries) of work, so that all of a row's loads are in flight at once
    // measured on B200 (column pass, C2 and a C3 shard): 4 lanes per ~100-entry column beat 2, 8, 16 and 32 — the
    // per-column tail (heavy entries, ten scalar stores, the -80 rule) runs once per lane GROUP, so narrow groups keep
    // more columns' tails in flight per warp; widen only when a lane would get more than ~4 unrolled batches
    int g = 4;
    while (g < 32 && avg_len > 8.0 * SCORE_UNROLL * g) g <<= 1;
    return g;
}

Lbc10 make_lbc(const clx_ctx* c) {
    Lbc10 l;
    for (int i = 0; i < 10; ++i) l.v[i] = c->lbc10[i];
    return l;
}

bool use_tiled(const clx_ctx* c) { return c->tiled_ready && !(c->flags & CLX_FLAG_KERNEL_V0); }


template <int NT, bool WITH_E>
void launch_score_rows(clx_ctx* c, int G) {
    const uint32_t n = c->n_cells;
    if (n == 0) return;
    double* out = NT == 1 ? c->ll.p : c->post_ll.p;
#define CLX_SCORE(GG)                                                                                                  \
    k_score_rows<GG, NT, WITH_E><<<(unsigned)(((uint64_t)c->n_seg * GG + TPB - 1) / TPB), TPB, 0, c->stream>>>(         \
        c->csr.p, c->row_ptr.p, n, c->tab.p, c->n_loci, c->hrow.p, c->seg_row.p, c->seg_first.p, c->n_seg, c->seg_part.p, \
        c->seg_cnt.p, out, c->ell.p, c->nloci.p, c->norm.p, &c->d_stats.p->nan_flag)
    switch (G) {
        case 4: CLX_SCORE(4); break;
        case 8: CLX_SCORE(8); break;
        case 16: CLX_SCORE(16); break;
        default: CLX_SCORE(32); break;
    }
#undef CLX_SCORE
    c->launches++;
    if (c->n_seg > n) {  // some row has more than one segment
        k_score_rows_combine<NT, WITH_E><<<(n + TPB - 1) / TPB, TPB, 0, c->stream>>>(c->seg_first.p, n, c->n_seg, c->seg_part.p, c->seg_cnt.p,
                                                                                  c->hrow.p, out, c->ell.p, c->nloci.p, c->norm.p,
                                                                                  &c->d_stats.p->nan_flag);
        c->launches++;
    }
}

template <bool FUSE>
void launch_columns(clx_ctx* c, int G) {
    const uint32_t n = c->n_loci;
    if (n == 0) return;
    // sharded over peer memory: the partial sums go into this rank's window, where the slice owners read them
    double* out_ll = c->p2p ? reinterpret_cast<double*>((unsigned char*)c->win + c->win_off_part_ll) : c->loc_ll.p;
    unsigned long long* out_int = c->p2p ? reinterpret_cast<unsigned long long*>((unsigned char*)c->win + c->win_off_part_int) : c->loc_int.p;
#define CLX_COLS(GG)                                                                                                   \
    k_columns<GG, FUSE><<<(unsigned)(((uint64_t)n * GG + TPB - 1) / TPB), TPB, 0, c->stream>>>(                          \
        c->csc.p, c->col_ptr.p, n, c->tab.p, c->loci_used.p, c->excl_bits.p, c->hc_col_ptr.p, c->hc_perm.p,            \
        c->h_cell.p, c->h_ref.p, c->h_alt.p, c->hv_ll.p, out_ll, out_int, c->n_used.p, c->d_stats.p)
    switch (G) {
        case 1: CLX_COLS(1); break;
        case 2: CLX_COLS(2); break;
        case 4: CLX_COLS(4); break;
        case 8: CLX_COLS(8); break;
        case 16: CLX_COLS(16); break;
        default: CLX_COLS(32); break;
    }
#undef CLX_COLS
    c->launches++;
}

// the column pass from per-locus class counts (k_excl_delta + k_locus_finish, em_kernels.cuh)
template <bool FUSE>
void launch_locus_classes(clx_ctx* c) {
    const uint32_t N = c->n_cells, L = c->n_loci;
    if (N) {
        k_excl_delta<<<(unsigned)(((uint64_t)c->n_seg * 32 + TPB - 1) / TPB), TPB, 0, c->stream>>>(c->csr.p, c->row_ptr.p, c->seg_row.p,
                                                                                             c->seg_first.p, c->n_seg, c->excl_bits.p,
                                                                                             c->excl_prev.p, c->cls_min.p);
        c->launches++;
    }
    if (L == 0) return;
    double* out_ll = c->p2p ? reinterpret_cast<double*>((unsigned char*)c->win + c->win_off_part_ll) : c->loc_ll.p;
    unsigned long long* out_int = c->p2p ? reinterpret_cast<unsigned long long*>((unsigned char*)c->win + c->win_off_part_int) : c->loc_int.p;
    k_locus_finish<FUSE><<<(L + TPB - 1) / TPB, TPB, 0, c->stream>>>(c->cls_all.p, c->cls_min.p, L, c->tab.p, c->loci_used.p, c->excl_bits.p,
                                                                    c->hm_col_ptr.p, c->hm_ent.p, c->hm_val.p, c->csc.p, c->col_ptr.p,
                                                                    out_ll, out_int, c->n_used.p,
                                                                    c->d_stats.p, c->excl_prev.p, (N + 31u) / 32u);
    c->launches++;
}

// ranks + interpolation weights of statrs' Data::{median, quantile} for n values
void quantile_plan(uint64_t n, double tau, unsigned long long* ra, unsigned long long* rb, double* frac, int* mode) {
    *ra = *rb = 0;
    *frac = 0.0;
    *mode = 1;
    if (n == 0) return;
    const double h = ((double)n + 1.0 / 3.0) * tau + 1.0 / 3.0;
    const long long hf = (long long)h;
    if (hf <= 0) { *mode = 1; return; }
    if (hf >= (long long)n) { *mode = 2; *ra = *rb = n - 1; return; }
    *mode = 0;
    *ra = (unsigned long long)hf - 1;
    *rb = (unsigned long long)hf < n - 1 ? (unsigned long long)hf : n - 1;
    *frac = h - (double)hf;
}

SelectPlan make_plan(uint64_t n) {
    SelectPlan p;
    memset(&p, 0, sizeof(p));
    p.n = n;
    if (n) {
        const uint64_t k = n / 2;
        p.median_even = (n % 2 == 0) ? 1 : 0;
        p.rank[0] = p.median_even ? k - 1 : k;
        p.rank[1] = k;
    }
    quantile_plan(n, 0.25, &p.rank[2], &p.rank[3], &p.frac_q1, &p.q1_mode);
    quantile_plan(n, 0.75, &p.rank[4], &p.rank[5], &p.frac_q3, &p.q3_mode);
    return p;
}

// the select as one thread-block cluster: one GPU, at most 16 CTAs of 1024 threads with 1 or 4 keys per thread
typedef void (*SelectClusterFn)(const double*, uint64_t, SelectPlan, const IterParams*, DeviceStats*, unsigned int*,
                                unsigned long long*, const double*, uint32_t, uint32_t*, unsigned char* const*, int, int, size_t);
SelectClusterFn select_cluster_fn(int R) {
    return R == 1 ? k_select_coop<1, SELECT_SMALL_THREADS, true> : k_select_coop<4, SELECT_SMALL_THREADS, true>;
}
bool select_cluster_shape(const clx_ctx* c, uint64_t n_vals, int* R_out, unsigned* grid_out) {
    if (c->world != 1 || n_vals > 16ull * 4 * SELECT_SMALL_THREADS) return false;
    const int R = n_vals <= 16ull * SELECT_SMALL_THREADS ? 1 : 4;
    const uint64_t per_cta = (uint64_t)SELECT_SMALL_THREADS * R;
    if (R_out) *R_out = R;
    if (grid_out) *grid_out = (unsigned)std::max<uint64_t>(1, (n_vals + per_cta - 1) / per_cta);
    return true;
}
// asked once per load, outside any stream capture: can this device co-schedule the cluster?
void select_cluster_probe(clx_ctx* c, uint64_t n_vals) {
    c->select_cluster_ok = 0;
    int R = 1;
    unsigned grid = 1;
    if (!select_cluster_shape(c, n_vals, &R, &grid)) return;
    const void* fn = (const void*)select_cluster_fn(R);
    if (cudaFuncSetAttribute(fn, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) {
        cudaGetLastError();
        return;
    }
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(SELECT_SMALL_THREADS);
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = grid;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int n_clusters = 0;
    if (cudaOccupancyMaxActiveClusters(&n_clusters, fn, &cfg) != cudaSuccess) {
        cudaGetLastError();
        return;
    }
    c->select_cluster_ok = n_clusters >= 1 ? 1 : 0;
}

// exact order statistics of `vals[0, n_vals)` (pads = +inf beyond the n real values) -> d_stats, then the new
// excluded set of this rank's cells (main.rs:330-335), in place in excl_bits
int run_select_exclude(clx_ctx* c, const double* vals, uint64_t n_vals, uint64_t n_real) {
    const SelectPlan plan = make_plan(n_real);
    cudaStream_t s = c->stream;
    const uint32_t N = c->n_cells;
    static const int mode = env_int("CLX_SELECT_MODE", 0);  // 0 cooperative kernel, 2 forces the multi-launch path
    // cooperative grid, keys in registers, one launch: as few keys per thread as one CTA per SM allows (1, 4 or 16);
    // 256-thread CTAs for small inputs measured slightly slower (29.8 vs 27.9 us at 10 k keys)
    const uint64_t cap1 = (uint64_t)SELECT_SMALL_THREADS * (uint64_t)c->sm_count;
    // up to 16 CTAs' worth of keys on one GPU: the same kernel as ONE thread-block cluster (hardware cluster barrier
    // instead of the cooperative grid barrier; em_alloc asked the device whether it can place it);
    // CLX_SELECT_MODE=1 keeps the cooperative launch for A/B runs
    if (mode == 0 && c->select_cluster_ok == 1 && select_cluster_shape(c, n_vals, nullptr, nullptr)) {
        int R = 1;
        unsigned grid = 1;
        select_cluster_shape(c, n_vals, &R, &grid);
        cudaLaunchConfig_t cfg;
        memset(&cfg, 0, sizeof(cfg));
        cfg.gridDim = dim3(grid);
        cfg.blockDim = dim3(SELECT_SMALL_THREADS);
        cfg.stream = s;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = grid;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        const double* norm_local = c->norm.p;
        CLX_CUDA(c, cudaLaunchKernelEx(&cfg, select_cluster_fn(R), vals, n_vals, plan, (const IterParams*)c->d_params.p, c->d_stats.p,
                                       c->sel_hist.p, (unsigned long long*)c->sel_keys.p, norm_local, N, c->excl_bits.p,
                                       (unsigned char* const*)nullptr, 0, 1, (size_t)0));
        c->launches++;
        return CLX_OK;
    }
    if (mode != 2 && n_vals <= 16 * cap1) {
        const int R = n_vals <= cap1 ? 1 : (n_vals <= 4 * cap1 ? 4 : 16);
        const int threads = SELECT_SMALL_THREADS;
        const uint64_t per_cta = (uint64_t)threads * R;
        // the histogram scratch was zeroed at the head of the iteration (k_tables_em)
        const unsigned grid = (unsigned)std::max<uint64_t>(1, (n_vals + per_cta - 1) / per_cta);
        const double* norm_local = c->norm.p;
        const IterParams* prm = c->d_params.p;
        DeviceStats* st = c->d_stats.p;
        unsigned int* scratch = c->sel_hist.p;
        unsigned long long* glist = (unsigned long long*)c->sel_keys.p;
        uint32_t* bits = c->excl_bits.p;
        uint32_t n_local = N;
        SelectPlan pl = plan;
        unsigned char* const* win = c->p2p ? (unsigned char* const*)c->d_peer_win.p : nullptr;
        int rank = c->rank, world = c->world;
        size_t off_counters = c->win_off_counters;
        void* args[] = {(void*)&vals, (void*)&n_vals, (void*)&pl, (void*)&prm, (void*)&st, (void*)&scratch,
                        (void*)&glist, (void*)&norm_local, (void*)&n_local, (void*)&bits, (void*)&win, (void*)&rank,
                        (void*)&world, (void*)&off_counters};
        const void* fn = R == 1 ? (const void*)k_select_coop<1, SELECT_SMALL_THREADS>
                                : (R == 4 ? (const void*)k_select_coop<4, SELECT_SMALL_THREADS>
                                          : (const void*)k_select_coop<16, SELECT_SMALL_THREADS>);
        CLX_CUDA(c, cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(threads), args, 0, s));
        c->launches++;
        return CLX_OK;
    }
    unsigned long long* keys = (unsigned long long*)c->sel_keys.p;
    SelectState* d_state = (SelectState*)c->sel_state.p;
    unsigned int* hist = c->sel_hist.p;
    unsigned int* ticket = hist + 8 * SELECT_TARGETS * 256;
    const unsigned grid = (unsigned)std::min<uint64_t>((n_vals + 512 * 8 - 1) / (512 * 8), (uint64_t)c->sm_count * 2);
    k_select_keys<<<grid, 512, 0, s>>>(vals, n_vals, keys, plan, d_state);
    for (int pass = 0; pass < 8; ++pass)
        k_select_pass<<<grid, 512, 0, s>>>(keys, n_vals, pass, d_state, hist, ticket, plan, c->d_params.p, c->d_stats.p);
    c->launches += 9;
    if (N) {
        k_exclude<<<(N + TPB - 1) / TPB, TPB, 0, s>>>(c->norm.p, N, c->excl_bits.p, c->d_stats.p);
        c->launches++;
    }
    return CLX_OK;
}

}  // namespace

int em_alloc(clx_ctx* c) {
    const size_t N = c->n_cells, L = c->n_loci;
    const size_t words = (N + 31) / 32 + 1;
    em_drop_graph(c);
    CLX_CUDA(c, c->excl_bits.alloc(words));
    CLX_CUDA(c, c->loci_used.alloc(L + 1));
    CLX_CUDA(c, c->cls_min.alloc((L + 1) * CLS_STRIDE));
    CLX_CUDA(c, c->excl_prev.alloc(words));
    CLX_CUDA(c, c->tab.alloc(3 * L * TAB_STRIDE + TAB_STRIDE));
    CLX_CUDA(c, c->ll.alloc(N + 1));
    CLX_CUDA(c, c->post_ll.alloc(3 * N + 1));
    CLX_CUDA(c, c->hrow.alloc(3 * N + 1));
    CLX_CUDA(c, cudaMemsetAsync(c->hrow.p, 0, c->hrow.bytes(), c->stream));
    CLX_CUDA(c, c->ell.alloc(N + 1));
    const size_t norm_n = std::max<size_t>(N, c->max_local_cells) + 1;
    CLX_CUDA(c, c->norm.alloc(norm_n));
    CLX_CUDA(c, c->nloci.alloc(N + 1));
    CLX_CUDA(c, c->loc_ll.alloc(2 * L + 1));
    CLX_CUDA(c, c->loc_int.alloc(8 * L + 1));
    CLX_CUDA(c, c->post.alloc(2 * N + 1));
    CLX_CUDA(c, c->post_u8.alloc(3 * N + 1));
    const size_t n_sel = c->world > 1 ? (size_t)c->world * c->max_local_cells : N;
    if (c->world > 1) CLX_CUDA(c, c->norm_all.alloc(n_sel + 1));
    CLX_CUDA(c, c->sel_keys.alloc(n_sel + 256));  // key array of the multi-launch path / survivor list
    CLX_CUDA(c, c->sel_hist.alloc(SELECT_COOP_SCRATCH + 8 * SELECT_TARGETS * 256 + 8));
    CLX_CUDA(c, c->sel_state.alloc(sizeof(SelectState) / 8 + 1));
    select_cluster_probe(c, n_sel);
    // pads of the gather buffer: +inf sorts after every real value
    {
        std::vector<double> inf(norm_n, INFINITY);
        CLX_CUDA(c, cudaMemcpyAsync(c->norm.p, inf.data(), norm_n * 8, cudaMemcpyHostToDevice, c->stream));
        CLX_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    CLX_TRY(p2p_setup(c, L, c->max_local_cells));
    {   // row segments of the gather scoring kernel
        DevBuf<uint32_t> cnt;
        DevBuf<uint8_t> tmp;
        size_t tb = 0;
        CLX_CUDA(c, cnt.alloc(N + 1));
        CLX_CUDA(c, c->seg_first.alloc(N + 1));
        k_row_seg_count<<<(unsigned)((N + 1 + TPB - 1) / TPB), TPB, 0, c->stream>>>(c->row_ptr.p, (uint32_t)N, cnt.p);
        CLX_CUDA(c, cub::DeviceScan::ExclusiveSum(nullptr, tb, cnt.p, c->seg_first.p, (int)(N + 1), c->stream));
        CLX_CUDA(c, tmp.alloc(tb));
        CLX_CUDA(c, cub::DeviceScan::ExclusiveSum(tmp.p, tb, cnt.p, c->seg_first.p, (int)(N + 1), c->stream));
        uint32_t n_seg = 0;
        CLX_CUDA(c, cudaMemcpyAsync(&n_seg, c->seg_first.p + N, 4, cudaMemcpyDeviceToHost, c->stream));
        CLX_CUDA(c, cudaStreamSynchronize(c->stream));
        c->n_seg = n_seg;
        CLX_CUDA(c, c->seg_row.alloc((size_t)n_seg + 1));
        CLX_CUDA(c, c->seg_cnt.alloc((size_t)n_seg + 1));
        CLX_CUDA(c, c->seg_part.alloc(4 * (size_t)n_seg + 1));
        if (N) k_row_seg_fill<<<(unsigned)((N + TPB - 1) / TPB), TPB, 0, c->stream>>>(c->seg_first.p, (uint32_t)N, c->seg_row.p);
        c->launches += 4;
    }
    c->score_G = env_int("CLX_SCORE_G", pick_group(N ? (double)c->nnz / N : 0.0));
    c->col_G = env_int("CLX_COL_G", pick_group(L ? (double)c->nnz / L : 0.0));
    // roofline accounting of the gather scoring kernel: packed entries + row pointers + outputs + table
    c->score_bytes = 8ull * c->nnz + 8ull * (N + 1) + 20ull * N + 128ull * L;
    // run-stream kernel: the stream (512-byte chunks of 2-byte words) + one read of the 144-byte slot records per
    // cell block + 16 B of partial sums per cell and work item
    if (use_tiled(c)) c->score_bytes = 512ull * c->tl_stream_chunks + 144ull * L * c->tl_Pc + 16ull * c->tl_n_items * c->tl_NB;
    return em_reset(c);
}

int em_reset(clx_ctx* c) {
    cudaStream_t s = c->stream;
    CLX_CUDA(c, cudaMemsetAsync(c->excl_bits.p, 0, c->excl_bits.bytes(), s));
    CLX_CUDA(c, cudaMemsetAsync(c->loci_used.p, 1, c->loci_used.bytes(), s));
    CLX_CUDA(c, cudaMemsetAsync(c->cls_min.p, 0, c->cls_min.bytes(), s));
    CLX_CUDA(c, cudaMemsetAsync(c->excl_prev.p, 0, c->excl_prev.bytes(), s));
    CLX_CUDA(c, cudaMemsetAsync(c->loc_int.p, 0, c->loc_int.bytes(), s));
    CLX_CUDA(c, cudaMemsetAsync(c->loc_ll.p, 0, c->loc_ll.bytes(), s));
    if (c->n_cells) CLX_CUDA(c, cudaMemcpyAsync(c->n_used.p, c->row_len.p, (size_t)c->n_cells * 4, cudaMemcpyDeviceToDevice, s));
    CLX_CUDA(c, cudaStreamSynchronize(s));
    c->iteration = 0;
    c->have_iteration = false;
    c->n_excluded_global = 0;
    return CLX_OK;
}

void em_drop_graph(clx_ctx* c) {
    if (c->iter_graph_exec) cudaGraphExecDestroy((cudaGraphExec_t)c->iter_graph_exec);
    c->iter_graph_exec = nullptr;
    for (auto& p : c->graph_events) {
        c->event_pool.push_back(p.a);
        c->event_pool.push_back(p.b);
    }
    c->graph_events.clear();
}

namespace {

// Every kernel of one compute_new_excluded pass, in stream order (also the body of the captured CUDA graph).
int enqueue_iteration(clx_ctx* c) {
    cudaStream_t s = c->stream;
    const uint32_t N = c->n_cells, L = c->n_loci;
    const int with_e = (c->flags & CLX_FLAG_NO_EXPECTED) ? 0 : 1;
    CLX_CUDA(c, cudaMemcpyAsync(c->d_params.p, c->h_params, sizeof(IterParams), cudaMemcpyHostToDevice, s));

    // 1. + 2a. alpha/beta + memo tables from the CURRENT excluded set (main.rs:309; also clears d_stats) and, as a
    //    parallel branch on the side stream, the direct evaluation of the heavy entries (it derives alpha/beta from
    //    the same counts).  With phase profiling on the two run back to back so that their event pairs mean something.
    const bool fork = c->n_heavy != 0 && !(c->flags & CLX_FLAG_PROFILE);
    cudaStream_t hs = fork ? c->stream2 : s;
    const EmCounts cnts{c->locus_tot.p, c->loc_int.p + 6 * (size_t)L, c->loc_int.p + 7 * (size_t)L, c->loci_used.p, c->lng.p, c->lng_n};
    if (fork) {
        CLX_CUDA(c, cudaEventRecord(c->ev_fork, s));
        CLX_CUDA(c, cudaStreamWaitEvent(hs, c->ev_fork, 0));
    }
    auto heavy = [&]() -> int {
        if (!c->n_heavy) return CLX_OK;
        if (c->n_deep) {  // HEAVY_CHAIN_TOTAL < total <= HEAVY_GIANT_TOTAL: a warp per entry
            k_heavy_em<<<(c->n_deep + HEAVY_HB - 1) / HEAVY_HB, TPB, 0, hs>>>(c->deep_idx.p, c->n_deep, c->h_locus.p, c->h_ref.p, c->h_alt.p,
                                                                              cnts, c->lnfact.p, with_e, c->hv_ll.p, c->hv_e.p, c->hm_pos.p, c->hm_val.p);
            c->launches++;
        }
        // entries up to HEAVY_CHAIN_TOTAL: one evaluation per (locus, total) group
        const unsigned gg = (unsigned)std::min<uint64_t>((c->n_heavy + 127) / 128, (uint64_t)c->sm_count * 16);
        k_heavy_groups<<<gg, 128, 0, hs>>>(c->hg_start.p, c->hm_ent.p, c->hg_count.p, (uint32_t)c->n_heavy, c->h_locus.p, cnts, c->lnfact.p,
                                           with_e, c->hv_ll.p, c->hv_e.p, c->hm_val.p);
        c->launches++;
        if (c->n_giant) {
            k_heavy_giant<<<c->n_giant, TPB, 0, hs>>>(c->giant_idx.p, c->n_giant, c->h_locus.p, c->h_ref.p, c->h_alt.p, cnts,
                                                      c->lnfact.p, with_e, c->hv_ll.p, c->hv_e.p, c->hm_pos.p, c->hm_val.p);
            c->launches++;
        }
        if (c->n_cells && !use_tiled(c)) {  // the gather kernel adds per-cell sums; the run-stream finish walks the lists itself
            k_heavy_row_sums<<<(unsigned)(((uint64_t)c->n_cells * 32 + TPB - 1) / TPB), TPB, 0, hs>>>(c->h_row_ptr.p, c->n_cells, c->hv_ll.p,
                                                                                                    c->hv_e.p, nullptr, c->hrow.p);
            c->launches++;
        }
        return CLX_OK;
    };
    if (fork) {
        CLX_TRY(heavy());
        CLX_CUDA(c, cudaEventRecord(c->ev_join, hs));
    }
    phase_begin(c, CLX_PHASE_TABLES);
    if (L) {
        k_tables_em<<<(unsigned)((3ull * L + TPB - 1) / TPB), TPB, 0, s>>>(cnts, L, make_lbc(c), with_e, c->tab.p,
                                                                           c->tiled_ready ? c->tab16.p : nullptr, c->d_stats.p,
                                                                           c->sel_hist.p, (uint32_t)c->sel_hist.n);
        c->launches++;
    } else {
        CLX_CUDA(c, cudaMemsetAsync(c->d_stats.p, 0, sizeof(DeviceStats), s));
        CLX_CUDA(c, cudaMemsetAsync(c->sel_hist.p, 0, c->sel_hist.bytes(), s));
    }
    phase_end(c, CLX_PHASE_TABLES);
    CLX_DEBUG_SYNC(c, "k_tables");
    phase_begin(c, CLX_PHASE_HEAVY);
    if (fork) CLX_CUDA(c, cudaStreamWaitEvent(s, c->ev_join, 0));
    else CLX_TRY(heavy());
    phase_end(c, CLX_PHASE_HEAVY);
    CLX_DEBUG_SYNC(c, "k_heavy");
    // 2b. the per-cell scoring pass (main.rs:312)
    phase_begin(c, CLX_PHASE_SCORE);
    if (use_tiled(c)) {
        CLX_TRY(score_tiled(c, SCORE_EM));
    } else {
        if (with_e) launch_score_rows<1, true>(c, c->score_G);
        else launch_score_rows<1, false>(c, c->score_G);
    }
    phase_end(c, CLX_PHASE_SCORE);
    CLX_DEBUG_SYNC(c, "score");

    // 3. threshold from exact order statistics of the normalized log likelihoods (main.rs:313-329)
    //    + new excluded set and change counts (main.rs:330-335)
    const double* sel_vals = c->norm.p;
    uint64_t n_vals = N;
    if (c->world > 1 && c->p2p) {
        phase_begin(c, CLX_PHASE_COMM);
        CLX_TRY(p2p_put_norm(c));   // all-gather = remote stores into every rank's window (also bumps the barrier epoch);
                                    // barrier 0 sits at the head of the cooperative select kernel
        if (env_int("CLX_SELECT_MODE", 0) == 2 ||
            (uint64_t)c->world * c->max_local_cells > 16ull * SELECT_SMALL_THREADS * (uint64_t)c->sm_count)
            CLX_TRY(p2p_barrier(c, 0));  // the multi-launch select has no barrier of its own
        phase_end(c, CLX_PHASE_COMM);
        sel_vals = reinterpret_cast<const double*>((unsigned char*)c->win + c->win_off_norm);
        n_vals = (uint64_t)c->world * c->max_local_cells;
    } else if (c->world > 1) {
        phase_begin(c, CLX_PHASE_COMM);
        CLX_TRY(comm_allgather_f64(c, c->norm.p, c->norm_all.p, c->max_local_cells));
        phase_end(c, CLX_PHASE_COMM);
        sel_vals = c->norm_all.p;
        n_vals = (uint64_t)c->world * c->max_local_cells;
    }
    phase_begin(c, CLX_PHASE_SELECT);
    CLX_TRY(run_select_exclude(c, sel_vals, n_vals, c->n_cells_global));
    phase_end(c, CLX_PHASE_SELECT);
    CLX_DEBUG_SYNC(c, "select");
    // 4. per-locus contributions split by the NEW set (main.rs:343) + minority allele counts for the next
    //    alpha/beta + the -80 locus filter (main.rs:444-451; after the allreduce when cells are sharded)
    phase_begin(c, CLX_PHASE_COLUMN);
    if (c->flags & CLX_FLAG_KERNEL_V0) {  // the column walk (k_columns): kept as the cross-check of the class-count pass
        if (c->world > 1) launch_columns<false>(c, c->col_G);
        else launch_columns<true>(c, c->col_G);
    } else {
        if (c->world > 1) launch_locus_classes<false>(c);
        else launch_locus_classes<true>(c);
    }
    phase_end(c, CLX_PHASE_COLUMN);
    CLX_DEBUG_SYNC(c, "k_columns");
    if (c->world > 1) {
        phase_begin(c, CLX_PHASE_COMM);
        if (c->p2p) {
            // barrier 1 + slice-owner reduction that stores the totals into every window (one launch); barrier 2, the
            // copy of the totals and the -80 rule are fused into k_locus_filter_p2p below
            CLX_TRY(p2p_reduce(c));
        } else {
            // one fused NCCL launch: per-locus sums and counts + the three change counters of the exclusion step
            CLX_TRY(comm_group_begin(c));
            CLX_TRY(comm_allreduce_u64(c, &c->d_stats.p->n_new, 3));
            CLX_TRY(comm_allreduce_f64(c, c->loc_ll.p, 2 * (size_t)L));
            CLX_TRY(comm_allreduce_u64(c, c->loc_int.p, 8 * (size_t)L));
            CLX_TRY(comm_group_end(c));
        }
        phase_end(c, CLX_PHASE_COMM);
        phase_begin(c, CLX_PHASE_COLUMN);
        if (c->p2p) {
            k_locus_filter_p2p<<<std::max(1u, (L + TPB - 1) / TPB), TPB, 0, s>>>(
                (unsigned char* const*)c->d_peer_win.p, c->rank, c->world, c->win_off_counters, c->win_off_tot_ll, c->win_off_tot_int,
                c->loc_ll.p, c->loc_int.p, L, c->loci_used.p, c->csc.p, c->col_ptr.p, c->n_used.p, c->d_stats.p);
            c->launches++;
        } else if (L) {
            k_locus_filter<<<(L + TPB - 1) / TPB, TPB, 0, s>>>(c->loc_ll.p, c->loc_int.p, L, c->loci_used.p, c->csc.p, c->col_ptr.p,
                                                               c->n_used.p, c->d_stats.p);
            c->launches++;
        }
        phase_end(c, CLX_PHASE_COLUMN);
    }
    CLX_CUDA(c, cudaMemcpyAsync(c->h_stats, c->d_stats.p, sizeof(DeviceStats), cudaMemcpyDeviceToHost, s));
    return CLX_OK;
}

}  // namespace

int em_iteration(clx_ctx* c, double iqr_multiple, clx_iter_stats* out) {
    if (!c->loaded) return fail(c, CLX_ERR_STATE, "clx_em_iteration before clx_load_coo");
    if (c->p2p_failed)
        return fail(c, CLX_ERR_NCCL, "a peer-memory barrier timed out earlier: the ranks are out of step, load again on every rank");
    NvtxRange nvtx("clx:em_iteration");
    cudaStream_t s = c->stream;
    c->h_params->iqr_multiple = iqr_multiple;
    // The iteration is replayed as a CUDA graph (a dozen launches, ~100 us of work at 10x scale: launch gaps matter),
    // also when cells are sharded and the exchanges go over peer memory (barrier kernels inside the graph).  Only
    // the NCCL fallback enqueues directly: collective calls sit between the kernels there.
    const bool use_graph = (c->world == 1 || c->p2p) && !(c->flags & CLX_FLAG_NO_GRAPH);
    if (use_graph && !c->iter_graph_exec) {
        cudaGraph_t graph = nullptr;
        const uint64_t launches0 = c->launches;
        CLX_CUDA(c, cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
        c->capturing = true;
        const int rc = enqueue_iteration(c);
        c->capturing = false;
        const cudaError_t e = cudaStreamEndCapture(s, &graph);
        c->graph_launches = c->launches - launches0;
        c->launches = launches0;
        if (rc != CLX_OK) {
            if (graph) cudaGraphDestroy(graph);
            return rc;
        }
        CLX_CUDA(c, e);
        cudaGraphExec_t exec = nullptr;
        const cudaError_t e2 = cudaGraphInstantiate(&exec, graph, 0);
        cudaGraphDestroy(graph);
        CLX_CUDA(c, e2);
        c->iter_graph_exec = exec;
    }
    CLX_CUDA(c, cudaEventRecord(c->ev_iter_a, s));
    if (use_graph) {
        CLX_CUDA(c, cudaGraphLaunch((cudaGraphExec_t)c->iter_graph_exec, s));
        c->launches += c->graph_launches;
        c->graph_events_fresh = true;
    } else {
        CLX_TRY(enqueue_iteration(c));
    }
    CLX_CUDA(c, cudaEventRecord(c->ev_iter_b, s));
    CLX_CUDA(c, cudaStreamSynchronize(s));
    CLX_CUDA(c, cudaGetLastError());
    phase_resolve(c);
    const DeviceStats& st = *c->h_stats;
    if (st.pad) {
        c->p2p_failed = true;
        return fail(c, CLX_ERR_NCCL, "a peer-memory barrier timed out (a rank is missing or failed)");
    }
    if (st.nan_flag) return fail(c, CLX_ERR_NAN, "NaN normalized log likelihood (the reference panics in argsort, main.rs:504)");
    c->n_excluded_global = st.n_excluded;
    if (out) {
        memset(out, 0, sizeof(*out));
        out->any_change = (st.n_new > 0 || st.n_rescued > 0) ? 1 : 0;
        out->iteration = c->iteration;
        out->n_new_excluded = st.n_new;
        out->n_rescued = st.n_rescued;
        out->n_excluded = st.n_excluded;
        out->median = st.median;
        out->q1 = st.q1;
        out->q3 = st.q3;
        out->iqr = st.iqr;
        out->threshold = st.threshold;
        out->n_loci_filtered = st.n_loci_filtered;
        out->n_loci_in_use = st.n_loci_in_use;
        float ms = 0.f;
        cudaEventElapsedTime(&ms, c->ev_iter_a, c->ev_iter_b);
        out->device_ms = ms;
    }
    c->iteration++;
    c->have_iteration = true;
    return CLX_OK;
}

int em_posteriors(clx_ctx* c, double posterior_threshold, uint32_t min_loci_used) {
    if (!c->loaded) return fail(c, CLX_ERR_STATE, "clx_posteriors before clx_load_coo");
    NvtxRange nvtx("clx:posteriors");
    cudaStream_t s = c->stream;
    const uint32_t N = c->n_cells, L = c->n_loci;
    // main.rs:240, 250, 259-265 — scalars on the host in the reference's expression order
    const double n_cells = (double)c->n_cells_global;
    double mf = ((double)c->n_excluded_global + 1.0) / (n_cells + 1.0);
    const double mf_unclamped = mf;
    mf = fmax(mf, 0.01);
    PosteriorParams pp;
    pp.log_prior_doublet = log(n_cells / 1000.0 / 100.0 * fmax(mf, 0.1));
    pp.log_prior_minority = log(mf);
    pp.log_prior_majority = log(1.0 - mf);
    pp.posterior_threshold = posterior_threshold;
    pp.min_loci_used = min_loci_used;

    phase_begin(c, CLX_PHASE_TABLES);
    if (L) {
        k_tables<1><<<3 * ((L + TAB_LPB - 1) / TAB_LPB), TPB, 0, s>>>(
            c->locus_tot.p, c->loc_int.p + 6 * (size_t)L, c->loc_int.p + 7 * (size_t)L, L, make_lbc(c), mf_unclamped, mf, 0,
            nullptr, c->tab.p, c->tiled_ready ? c->tab16.p : nullptr, c->tl_nT * TILE_L, nullptr, c->lng.p, c->lng_n);
        c->launches++;
    }
    phase_end(c, CLX_PHASE_TABLES);
    phase_begin(c, CLX_PHASE_HEAVY);
    if (c->n_heavy) {
        k_heavy_post<<<(unsigned)((3 * c->n_heavy + TPB - 1) / TPB), TPB, 0, s>>>(
            c->h_locus.p, c->h_ref.p, c->h_alt.p, (uint32_t)c->n_heavy, c->tab.p, L, c->lnfact.p, c->hv_ll.p);
        c->launches++;
        if (!use_tiled(c)) {
            k_heavy_row_sums<<<(unsigned)(((uint64_t)N * 32 + TPB - 1) / TPB), TPB, 0, s>>>(c->h_row_ptr.p, N, c->hv_ll.p, c->hv_ll.p + c->n_heavy,
                                                                                         c->hv_ll.p + 2 * c->n_heavy, c->hrow.p);
            c->launches++;
        }
    }
    phase_end(c, CLX_PHASE_HEAVY);
    // get_loci_used_for_posterior_calc pushes true on both branches (main.rs:301-303): every locus is scored
    phase_begin(c, CLX_PHASE_SCORE);
    if (use_tiled(c)) {
        // two passes over the run stream instead of three: {minority, majority} share the 16-byte slots of pass 0
        CLX_TRY(score_tiled(c, SCORE_POST_AB));
        CLX_TRY(score_tiled(c, SCORE_POST_C));
    } else {
        launch_score_rows<3, false>(c, c->score_G);
    }
    phase_end(c, CLX_PHASE_SCORE);
    if (N) {
        k_posterior<<<(N + TPB - 1) / TPB, TPB, 0, s>>>(c->post_ll.p, N, c->row_ptr.p, c->excl_bits.p, pp, c->post.p, c->post_u8.p);
        c->launches++;
    }
    CLX_CUDA(c, cudaStreamSynchronize(s));
    CLX_CUDA(c, cudaGetLastError());
    phase_resolve(c);
    // the EM memo table in slot 0 is stale now; the next clx_em_iteration rebuilds it anyway
    return CLX_OK;
}

int em_final_locus_counts(clx_ctx* c, unsigned long long* d_min, unsigned long long* d_maj) {
    if (!c->coo_locus.p && c->nnz0) return fail(c, CLX_ERR_STATE, "clx_final_locus_counts needs CLX_FLAG_KEEP_COO at clx_create");
    cudaStream_t s = c->stream;
    CLX_CUDA(c, cudaMemsetAsync(d_min, 0, 2 * (size_t)c->n_loci0 * 8, s));
    CLX_CUDA(c, cudaMemsetAsync(d_maj, 0, 2 * (size_t)c->n_loci0 * 8, s));
    if (c->nnz0) {
        const unsigned grid = (unsigned)std::min<uint64_t>((c->nnz0 + TPB - 1) / TPB, (uint64_t)c->sm_count * 16);
        k_final_counts<<<grid, TPB, 0, s>>>(c->coo_locus.p, c->coo_cell.p, c->coo_ref.p, c->coo_alt.p, c->nnz0, c->excl_bits.p,
                                            d_min, d_maj);
        c->launches++;
    }
    CLX_TRY(comm_allreduce_u64(c, d_min, 2 * (size_t)c->n_loci0));
    CLX_TRY(comm_allreduce_u64(c, d_maj, 2 * (size_t)c->n_loci0));
    CLX_CUDA(c, cudaStreamSynchronize(s));
    return CLX_OK;
}

}  // namespace clx


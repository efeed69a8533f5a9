// ctx.h — internal context of libcellector_cuda.so (one GPU, one stream, optionally one NCCL rank).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include <string>
#include <vector>

#include "../../include/cellector_cuda.h"

namespace clx {

// Stream-ordered allocations (cudaMallocAsync on the context stream): the device pool keeps freed blocks
// (release threshold = max, set in clx_create), so the dozens of scratch buffers of a load cost no driver
// round trip after the first job.  The ABI entry points set the stream of the calling thread.
cudaStream_t& alloc_stream();

template <typename T>
struct DevBuf {
    T* p = nullptr;
    size_t n = 0;
    cudaStream_t st = nullptr;
    cudaError_t alloc(size_t count) {
        release();
        n = count;
        if (count == 0) return cudaSuccess;
        st = alloc_stream();
        const cudaError_t e = cudaMallocAsync((void**)&p, count * sizeof(T), st);
        // CLX_POISON=1 fills every fresh allocation with 0xAB: a read of memory nobody wrote shows up at once
        static const bool poison = getenv("CLX_POISON") != nullptr;
        if (e == cudaSuccess && poison) cudaMemsetAsync(p, 0xAB, count * sizeof(T), st);
        if (e != cudaSuccess) {
            p = nullptr;
            n = 0;
        }
        return e;
    }
    void release() {
        if (p) cudaFreeAsync(p, st);
        p = nullptr;
        n = 0;
    }
    ~DevBuf() { release(); }
    DevBuf() = default;
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    size_t bytes() const { return n * sizeof(T); }
};

// packed nonzero of the CSR (by cell) matrix: used-locus index + u16 counts (0xFFFF/0xFFFF = "see heavy list")
struct __align__(8) RowEntry {
    uint32_t locus;
    uint16_t ref, alt;
};
// packed nonzero of the CSC (by locus) matrix
struct __align__(8) ColEntry {
    uint32_t cell;
    uint16_t ref, alt;
};

struct DeviceStats {  // written by kernels, copied to pinned host memory once per iteration
    double median, q1, q3, iqr, threshold;
    unsigned long long n_new, n_rescued, n_excluded, n_loci_filtered, n_loci_in_use;
    unsigned int nan_flag, pad;
};

// Scalars that may change between iterations live in device memory so that a captured CUDA graph stays valid.
struct IterParams {
    double iqr_multiple;
};

struct PendingEvent {
    int phase;
    cudaEvent_t a, b;
};

}  // namespace clx

struct clx_ctx {
    std::string err;
    int device = 0;
    unsigned flags = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t stream2 = nullptr;               // side branch of an iteration (heavy entries next to the table build)
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    int sm_count = 148;
    uint64_t launches = 0;
    uint64_t score_bytes = 0;  // algorithmic bytes of one launch of the scoring kernel

    // --- communicator (comm.cu)
    int rank = 0, world = 1;
    void* nccl_comm = nullptr;

    // --- peer-memory window (p2p.cu): the per-iteration exchanges as NVLink stores/loads instead of NCCL launches
    bool p2p = false;
    void* win = nullptr;
    size_t win_bytes = 0, win_off_counters = 0, win_off_norm = 0, win_off_part_ll = 0, win_off_part_int = 0, win_off_tot_ll = 0,
           win_off_tot_int = 0;
    std::vector<void*> peer_win;          // [world] every rank's window as mapped here (own window at [rank])
    clx::DevBuf<void*> d_peer_win;        // the same table in device memory
    bool p2p_failed = false;              // a peer-memory barrier timed out: the ranks are out of step until clx_load_coo maps again
    unsigned long long p2p_epoch = 0;     // one per iteration; barrier flags only grow

    // --- matrix (load.cu)
    bool loaded = false;
    uint32_t n_cells = 0;          // local cells
    uint64_t n_cells_global = 0;   // over all ranks
    uint64_t cell_offset = 0;      // global id of local cell 0
    uint32_t n_loci0 = 0, n_loci = 0;
    uint64_t nnz0 = 0, nnz = 0;
    uint64_t n_heavy = 0;
    clx::DevBuf<uint32_t> locus_ids;          // [n_loci] used -> original
    clx::DevBuf<clx::RowEntry> csr;           // [nnz] rows = cells, file order inside a row
    clx::DevBuf<uint64_t> row_ptr;            // [n_cells + 1]
    clx::DevBuf<clx::ColEntry> csc;           // [nnz] columns = used loci, cells ascending inside a column
    clx::DevBuf<uint64_t> col_ptr;            // [n_loci + 1]
    clx::DevBuf<uint32_t> cls_all;            // [n_loci][CLS_STRIDE] stored entries of every locus by (alt | total) class
    clx::DevBuf<unsigned long long> locus_tot;  // [2 * n_loci] (ref, alt) totals over all ranks
    // heavy entries (total > K_LIGHT or a count that does not fit u16), CSR order
    clx::DevBuf<uint32_t> h_cell, h_locus, h_ref, h_alt;  // [n_heavy]
    clx::DevBuf<uint32_t> h_row_ptr;                      // [n_cells + 1]
    clx::DevBuf<uint32_t> hc_perm;                        // [n_heavy] CSC order -> index into the CSR-order lists
    clx::DevBuf<uint32_t> hc_col_ptr;                     // [n_loci + 1]
    // heavy entries in (locus, total) order ("member" order): equal keys form a group whose chain (total <= HEAVY_CHAIN_TOTAL)
    // is evaluated once (em_kernels.cuh: k_heavy_groups); a locus's members are one contiguous range for the column pass
    clx::DevBuf<uint32_t> hg_member;                      // [n_heavy] member -> index into the CSR-order lists
    clx::DevBuf<uint32_t> hg_start;                       // [n_heavy + 1] first member of every group (only [0, *hg_count) valid)
    clx::DevBuf<uint32_t> hg_count;                       // [2] device: number of groups
    clx::DevBuf<uint4> hm_ent;                            // [n_heavy] member order: {cell, ref, alt, CSR-order index}
    clx::DevBuf<uint32_t> hm_pos;                         // [n_heavy] CSR-order index -> member
    clx::DevBuf<double> hm_val;                           // [n_heavy] member order: log_pmf of the current EM pass
    clx::DevBuf<uint32_t> hm_col_ptr;                     // [n_loci + 1] members of every locus
    clx::DevBuf<uint32_t> giant_idx;                      // [n_giant] heavy entries with total > HEAVY_GIANT_TOTAL
    uint32_t n_giant = 0;
    clx::DevBuf<uint32_t> deep_idx;                       // [n_deep] heavy entries with HEAVY_CHAIN_TOTAL < total <= HEAVY_GIANT_TOTAL
    uint32_t n_deep = 0;
    clx::DevBuf<double> hv_ll, hv_e;                      // [3 * n_heavy] values of the current scoring pass
    clx::DevBuf<double> hrow;                             // [3][n_cells] the same summed per cell (k_heavy_row_sums)
    // --- run-stream scoring layout (tiled.cu): light entries (total 1..3) as 2-byte words, ordered by runs (one cell x
    //     one locus tile) in length-sorted groups of 32; one work item = one cell block x a range of locus tiles
    bool tiled_ready = false;
    uint32_t tl_NB = 0, tl_W = 0, tl_Pc = 0, tl_nT = 0, tl_n_items = 0, tl_n_units = 0;
    uint64_t tl_n_light = 0, tl_stream_chunks = 0;    // light entries; stream length in 512-byte warp chunks
    uint64_t tl_stream_steps = 0;                     // lookup steps x 32 lanes of one pass (padding included)
    clx::DevBuf<uint4> tl_stream;                     // [tl_stream_chunks][32] one 16-byte piece (8 words) per lane
    clx::DevBuf<uint4> tl_item_q;                     // [n_items] items in block-major order: cell block, first tile, n tiles, item id
    clx::DevBuf<uint4> tl_item_rng;                   // [n_items][W] per item and warp: first chunk, end chunk, leading tiles to skip
    uint32_t tl_grid = 0;                             // CTAs of k_score_runs (one per SM)
    clx::DevBuf<uint32_t> tl_cb_item;                 // [Pc + 1] first item of every cell block
    clx::DevBuf<uint32_t> tl_cta_first;               // [tl_grid + 1] first item of every CTA (static, contiguous pieces of equal cost)
    clx::DevBuf<double> tl_part;                      // [n_items][NB][2] per-item partial sums of the two slot halves
    clx::DevBuf<double> tab16;                        // [2][nT * TILE_L][TAB_SLOTS][2] staged tables (pass 0 / pass 1)
    bool tl_attr_set = false;                         // dynamic shared-memory limit set on this context's device
    clx::DevBuf<uint32_t> n_used, row_len;            // [n_cells] loci_used_per_cell (main.rs:575) / stored row length
    // retained COO for clx_final_locus_counts
    clx::DevBuf<uint32_t> coo_locus, coo_cell, coo_ref, coo_alt;
    clx::DevBuf<double> lnfact;  // [171]
    // lng[i] = ln_gamma((double)i), i in [1, lng_n): every ln_gamma argument of the EM pass is an integer-valued
    // double (alpha, beta = allele counts + 1), so the per-iteration table build is lookups, not Lanczos series
    clx::DevBuf<double> lng;
    uint32_t lng_n = 0;
    double lbc10[10];

    // --- EM state (em.cu)
    uint32_t iteration = 0;
    uint64_t n_excluded_global = 0;  // size of the current excluded set over all ranks
    bool have_iteration = false;
    clx::DevBuf<uint32_t> excl_bits;                  // [(n_cells + 31) / 32] current excluded set, updated in place
    clx::DevBuf<uint8_t> loci_used;                   // [n_loci]
    clx::DevBuf<uint32_t> cls_min;                    // [n_loci][CLS_STRIDE] the same counts over the excluded cells only
    clx::DevBuf<uint32_t> excl_prev;                  // excluded set cls_min was last brought up to date with
    clx::DevBuf<unsigned long long> min_cnt;          // [2 * n_loci] minority (ref, alt) over all stored entries
    clx::DevBuf<double> tab;                          // [3][n_loci][16] memo tables (slot 0 = EM, 0..2 = posteriors)
    clx::DevBuf<double> ll, ell, norm;                // [n_cells] outputs of the last EM iteration
    clx::DevBuf<double> post_ll;                      // [3 * n_cells] posterior pass: minority | majority | doublet log likelihoods
    clx::DevBuf<uint32_t> nloci;                      // [n_cells]
    clx::DevBuf<double> norm_all;                     // [world * max_local] gathered normalized values (world > 1)
    int select_cluster_ok = 0;                        // 1: this load's select runs as one thread-block cluster (asked in em_alloc)
    clx::DevBuf<uint64_t> sel_keys;                   // select scratch
    clx::DevBuf<unsigned int> sel_hist;               // select scratch
    clx::DevBuf<unsigned long long> sel_state;        // select scratch
    // per-locus outputs of the column pass: packed for one allreduce each
    clx::DevBuf<double> loc_ll;                       // [2 * n_loci] (minority, majority)
    clx::DevBuf<unsigned long long> loc_int;          // [6 * n_loci] cells_min, cells_maj, ref_min, alt_min, ref_maj, alt_maj
    clx::DevBuf<double> post;                         // [5 * n_cells] posterior pass outputs
    clx::DevBuf<uint8_t> post_u8;                     // [3 * n_cells]
    clx::DevBuf<clx::DeviceStats> d_stats;
    clx::DevBuf<clx::IterParams> d_params;
    clx::IterParams* h_params = nullptr;              // pinned; copied to d_params at the head of every iteration
    // gather scoring kernel: rows in segments of ROW_SEG entries
    clx::DevBuf<uint32_t> seg_first, seg_row, seg_cnt;  // [n_cells + 1] / [n_seg] / [n_seg]
    clx::DevBuf<double> seg_part;                     // [4][n_seg] partial sums of multi-segment rows
    uint32_t n_seg = 0;
    int score_G = 32, col_G = 8;                      // lanes per CSR row / CSC column (picked at load)
    void* iter_graph_exec = nullptr;                  // cudaGraphExec_t of one EM iteration (world == 1)
    uint64_t graph_launches = 0;                      // kernels inside that graph
    bool capturing = false, graph_events_fresh = false;
    std::vector<clx::PendingEvent> graph_events;      // phase events recorded by graph nodes (CLX_FLAG_PROFILE)
    clx::DeviceStats* h_stats = nullptr;              // pinned
    unsigned char* h_stage = nullptr;                 // pinned staging of the result getters (one sync per call instead of one per array)
    size_t h_stage_bytes = 0;
    double* h_soup = nullptr;                         // pinned, 64 bytes: loss / refusal flags of clx_soup_em (soup.cu)
    clx::DevBuf<uint8_t> flush_buf;
    uint32_t max_local_cells = 0;

    // --- profiling
    std::vector<clx::PendingEvent> pending;
    std::vector<cudaEvent_t> event_pool;
    double phase_ms[CLX_PHASE_COUNT] = {0};
    uint64_t phase_launches[CLX_PHASE_COUNT] = {0};
    cudaEvent_t ev_iter_a = nullptr, ev_iter_b = nullptr;
};

namespace clx {

int fail(clx_ctx* c, int code, const std::string& msg);
#define CLX_CUDA(ctx, call)                                                                              \
    do {                                                                                                 \
        cudaError_t e__ = (call);                                                                        \
        if (e__ != cudaSuccess)                                                                          \
            return clx::fail(ctx, CLX_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e__));    \
    } while (0)
#define CLX_TRY(call)                 \
    do {                              \
        int rc__ = (call);            \
        if (rc__ != CLX_OK) return rc__; \
    } while (0)

// CLX_SYNC_DEBUG=1 (with CLX_FLAG_NO_GRAPH): synchronise after a launch and name the kernel that faulted
#define CLX_DEBUG_SYNC(ctx, what)                                                                        \
    do {                                                                                                 \
        static const bool dbg__ = getenv("CLX_SYNC_DEBUG") != nullptr;                                   \
        if (dbg__ && !(ctx)->capturing) {                                                                \
            cudaError_t e__ = cudaStreamSynchronize((ctx)->stream);                                      \
            if (e__ == cudaSuccess) e__ = cudaGetLastError();                                            \
            if (e__ != cudaSuccess)                                                                      \
                return clx::fail(ctx, CLX_ERR_CUDA, std::string("after ") + what + ": " + cudaGetErrorString(e__)); \
        }                                                                                                \
    } while (0)

// NVTX range for the lifetime of the object (load stages, one EM iteration, the posterior pass)
// CLX_LOAD_TIMING=1: wall-clock of the load stages (a stream sync at every mark: diagnostics only)
struct StageClock {
    cudaStream_t s;
    bool on;
    double t0;
    const char* what;
    StageClock(cudaStream_t st, const char* w);
    void mark(const char* stage);
};

struct NvtxRange {
    explicit NvtxRange(const char* name);
    ~NvtxRange();
    NvtxRange(const NvtxRange&) = delete;
    NvtxRange& operator=(const NvtxRange&) = delete;
};

// profiling scopes: always an NVTX range; CUDA events only under CLX_FLAG_PROFILE
void phase_begin(clx_ctx* c, int phase);
void phase_end(clx_ctx* c, int phase);
void phase_resolve(clx_ctx* c);  // after a stream sync

// comm.cu — collectives on ctx->stream; no-ops when world == 1
int comm_allreduce_u32(clx_ctx* c, unsigned int* buf, size_t n);
int comm_allreduce_u64(clx_ctx* c, unsigned long long* buf, size_t n);
int comm_allreduce_f64(clx_ctx* c, double* buf, size_t n);
int comm_group_begin(clx_ctx* c);
int comm_group_end(clx_ctx* c);
int comm_allgather_f64(clx_ctx* c, const double* send, double* recv, size_t n_per_rank);
int comm_allgather_u64(clx_ctx* c, const unsigned long long* send, unsigned long long* recv, size_t n_per_rank);
int comm_allgather_u64_host(clx_ctx* c, uint64_t v, std::vector<uint64_t>* out);
void comm_destroy(clx_ctx* c);
int comm_unique_id(void* out128);
int comm_init(clx_ctx* c, int rank, int world, const void* id128);

// p2p.cu
int p2p_setup(clx_ctx* c, size_t L, size_t max_local);  // collective; leaves c->p2p false if peer mapping is not possible
void p2p_teardown(clx_ctx* c);
int p2p_barrier(clx_ctx* c, int phase);
int p2p_put_norm(clx_ctx* c);
int p2p_reduce(clx_ctx* c);
// load.cu
// `alt` == nullptr: `ref` holds both counts of an entry, ref | alt << 16 (clx_load_coo_packed)
int load_coo(clx_ctx* c, uint64_t nnz0, uint32_t n_loci0, uint32_t n_cells, const uint32_t* locus,
             const uint32_t* cell, const uint32_t* ref, const uint32_t* alt, uint32_t min_ref, uint32_t min_alt);
// tiled.cu
int build_tiled(clx_ctx* c, const uint32_t* d_cell_of_entry);  // CSR-order cell id of every entry
enum ScoreMode { SCORE_EM = 0, SCORE_POST_AB = 1, SCORE_POST_C = 2 };
int score_tiled(clx_ctx* c, ScoreMode mode);
// em.cu
int em_alloc(clx_ctx* c);
int em_reset(clx_ctx* c);
void em_drop_graph(clx_ctx* c);
int em_iteration(clx_ctx* c, double iqr_multiple, clx_iter_stats* out);
int em_posteriors(clx_ctx* c, double posterior_threshold, uint32_t min_loci_used);
int em_final_locus_counts(clx_ctx* c, unsigned long long* d_min, unsigned long long* d_maj);

}  // namespace clx

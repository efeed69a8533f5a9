/* cellector_cuda.h — C ABI of libcellector_cuda.so (sm_100a CUDA implementation of cellector's
 * per-cell genotype-scoring EM loop).
 *
 * The reference (wheaton5/cellector) is one Rust binary with no FFI of its own; this header cuts it at
 * the function seams a Rust host would bind with an `extern "C"` block:
 *
 *   clx_load_coo            replaces load_data::load_cell_data        cellector/src/load_data.rs:134-181
 *                           (+ get_loci_used                           cellector/src/load_data.rs:254-280)
 *   clx_load_coo_packed     the same replacement, ref and alt of an entry in one word (the two counts
 *                           load_data.rs:195-196 parses per line pair)
 *   clx_get_locus_ids       replaces the locus_ids vector              cellector/src/load_data.rs:144-147
 *   clx_em_iteration        replaces compute_new_excluded              cellector/src/main.rs:308-347
 *                           (init_alpha_betas :598-611, get_cell_log_likelihoods :541-591,
 *                            get_locus_log_likelihoods :368-420, -80 locus filter :428-451)
 *   clx_get_cell_vectors    replaces CellLogLikelihoodData + normalized vector   main.rs:508-514, 313-323
 *   clx_get_locus_vectors   replaces LocusLogLikelihoodData            cellector/src/main.rs:516-525
 *   clx_posteriors          replaces calculate_posteriors              cellector/src/main.rs:228-280
 *                           + the assignment rule of output_final_assignments  main.rs:141-170
 *   clx_final_locus_counts  replaces load_data::load_mtx_final         cellector/src/load_data.rs:109-132
 *
 * Conventions: every call returns 0 on success, non-zero on failure (message via clx_last_error);
 * nothing throws across the boundary; the caller owns every host buffer; a context is driven by one
 * host thread and owns one GPU (one process per GPU; ranks are joined with clx_comm_init);
 * calls are synchronous on return unless stated.  There is NO CPU fallback: without a CUDA device
 * clx_create fails.
 */
#ifndef CELLECTOR_CUDA_H
#define CELLECTOR_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CLX_OK 0
#define CLX_ERR_CUDA 1      /* a CUDA runtime call failed */
#define CLX_ERR_ARG 2       /* invalid argument / input out of range */
#define CLX_ERR_STATE 3     /* call order violated (e.g. iteration before load) */
#define CLX_ERR_NCCL 4      /* NCCL missing or failed */
#define CLX_ERR_NAN 5       /* NaN reached a comparison (reference panics: main.rs:504) */

/* clx_create flags */
#define CLX_FLAG_PROFILE 1u        /* record CUDA events around every kernel phase (clx_get_phase_ms) */
#define CLX_FLAG_NO_EXPECTED 2u    /* skip the expected_log_likelihood diagnostic column (stats.rs:8-33) */
#define CLX_FLAG_KERNEL_V0 4u      /* plain CSR gather scoring kernel + CSC column walk instead of the run stream + class counts */
#define CLX_FLAG_KEEP_COO 8u       /* keep the unfiltered COO on the device (needed by clx_final_locus_counts) */
#define CLX_FLAG_NO_GRAPH 16u      /* enqueue the kernels of an iteration one by one instead of replaying a CUDA graph */

typedef struct clx_ctx clx_ctx;

/* Per-iteration scalars: the two println! lines of main.rs:338-339 plus bookkeeping. */
typedef struct clx_iter_stats {
    int32_t any_change;        /* main.rs:335 */
    uint32_t iteration;        /* 0-based index of the iteration just run (file names main.rs:350) */
    uint64_t n_new_excluded;   /* main.rs:333 */
    uint64_t n_rescued;        /* main.rs:334 */
    uint64_t n_excluded;       /* size of the new excluded set */
    double median;             /* main.rs:325 */
    double q1;                 /* main.rs:326 */
    double q3;                 /* main.rs:327 */
    double iqr;                /* main.rs:328 */
    double threshold;          /* main.rs:329 */
    uint64_t n_loci_filtered;  /* loci newly switched off by the -80 rule, main.rs:444-451 */
    uint64_t n_loci_in_use;    /* loci still in use after the filter */
    float device_ms;           /* GPU time of this call (CUDA events on the context stream) */
    float reserved;
} clx_iter_stats;

/* Per-locus vectors of one iteration (LocusLogLikelihoodData, main.rs:516-525); every pointer is a
 * caller-owned array of n_loci_used elements or NULL to skip it. */
typedef struct clx_locus_arrays {
    double* ll_minority;
    double* ll_majority;
    uint32_t* cells_minority;
    uint32_t* cells_majority;
    uint64_t* ref_minority;
    uint64_t* alt_minority;
    uint64_t* ref_majority;
    uint64_t* alt_majority;
    uint8_t* loci_used; /* flags AFTER this iteration's -80 filter */
} clx_locus_arrays;

/* phases timed under CLX_FLAG_PROFILE */
enum clx_phase {
    CLX_PHASE_TABLES = 0,  /* per-locus alpha/beta + memo tables */
    CLX_PHASE_SCORE = 1,   /* per-cell scoring kernel (dominant) */
    CLX_PHASE_HEAVY = 2,   /* direct-Lanczos kernel for deep entries */
    CLX_PHASE_SELECT = 3,  /* radix select + threshold + exclusion */
    CLX_PHASE_COLUMN = 4,  /* per-locus column reduction + -80 filter */
    CLX_PHASE_COMM = 5,    /* collectives */
    CLX_PHASE_COUNT = 6
};

int clx_create(clx_ctx** out, int device, unsigned flags);
void clx_destroy(clx_ctx* ctx);
const char* clx_last_error(const clx_ctx* ctx); /* ctx may be NULL: last error of clx_create */

/* Multi-GPU (one process per GPU).  Rank 0 calls clx_nccl_unique_id and ships the 128 bytes to the other
 * ranks by any means (the host's own launcher); then every rank calls clx_comm_init before clx_load_coo.
 * Cells are sharded across ranks; per-locus quantities are allreduced, the threshold is global. */
int clx_nccl_unique_id(void* out128);
int clx_comm_init(clx_ctx* ctx, int rank, int world, const void* unique_id128);

/* Load this rank's cells.  COO arrays are in FILE ORDER (the order of the lines of alt.mtx/ref.mtx),
 * 0-based, caller-owned: pageable or pinned host memory, or device memory (copied with cudaMemcpyDefault).
 * `cell` holds rank-local cell ids in
 * [0, n_cells).  Duplicated coordinates are kept as separate entries and zero-total entries are kept,
 * exactly like load_data.rs:151-174.  The locus filter (load_data.rs:268-273) counts, per locus, the
 * entries with ref>0 and with alt>0 (over all ranks) and keeps loci with both counts >= min. */
int clx_load_coo(clx_ctx* ctx, uint64_t nnz0, uint32_t n_loci0, uint32_t n_cells, const uint32_t* locus,
                 const uint32_t* cell, const uint32_t* ref, const uint32_t* alt, uint32_t min_ref,
                 uint32_t min_alt, uint32_t* n_loci_used, uint64_t* nnz_used);

/* Same as clx_load_coo for hosts whose parser keeps both counts of an entry in one word: ref_alt[i] = ref | alt << 16
 * (each count <= 65535; a larger count must use clx_load_coo).  12 instead of 16 bytes per stored entry cross PCIe. */
int clx_load_coo_packed(clx_ctx* ctx, uint64_t nnz0, uint32_t n_loci0, uint32_t n_cells, const uint32_t* locus,
                        const uint32_t* cell, const uint32_t* ref_alt, uint32_t min_ref, uint32_t min_alt,
                        uint32_t* n_loci_used, uint64_t* nnz_used);

/* used-locus index -> original locus id, n_loci_used entries. */
int clx_get_locus_ids(clx_ctx* ctx, uint32_t* locus_ids);

/* Forget the EM state (excluded set, loci_used flags, iteration counter) but keep the loaded matrix. */
int clx_reset(clx_ctx* ctx);

/* One pass of compute_new_excluded (main.rs:308-347). */
int clx_em_iteration(clx_ctx* ctx, double iqr_multiple, clx_iter_stats* out);

/* Per-cell vectors of the LAST iteration; any pointer may be NULL.  n_loci is loci_used_per_cell
 * (main.rs:575), norm is the normalized log likelihood (main.rs:313-323), excluded is the NEW set. */
int clx_get_cell_vectors(clx_ctx* ctx, double* ll, double* expected_ll, uint32_t* n_loci, double* norm,
                         uint8_t* excluded);
int clx_get_locus_vectors(clx_ctx* ctx, const clx_locus_arrays* out);

/* calculate_posteriors (main.rs:228-280) with the current excluded set, plus the per-cell assignment rule
 * (main.rs:141-170).  label: 0 = "0" (minority), 1 = "1" (majority), 2 = "doublet", 3 = "unassigned".
 * anomaly: 0 if excluded else 1.  Any pointer may be NULL. */
int clx_posteriors(clx_ctx* ctx, double posterior_threshold, uint32_t min_loci_used, double* posterior,
                   double* doublet_posterior, double* ll_major, double* ll_minor, uint8_t* label,
                   uint8_t* anomaly, uint8_t* qual);

/* Allele counts over ALL n_loci0 loci split by the current excluded set; each output is
 * [n_loci0][2] = (ref, alt).  Replaces the third text pass of load_mtx_final. */
int clx_final_locus_counts(clx_ctx* ctx, uint64_t* minority_ref_alt, uint64_t* majority_ref_alt);

/* souporcell's k-cluster EM (the tool cellector_pipeline.py:233-240 runs right after cellector; replaces
 * EM(), souporcell/souporcell/src/main.rs:109-186, with binomial_loss :199-213, normalize_in_log_with_temp :230-242,
 * update_centers_average :273-280, update_final :244-251) on the matrix loaded with clx_load_coo (same locus filter as
 * souporcell's load_cell_data :436-440 with its default min_*_umis = 0).  One run from the caller's initial centers
 * (the draws of init_cluster_centers_uniform, main.rs:354-364, stay on the host side with the host's generator):
 *   k                 clusters, 1..8
 *   centers_init      HOST array [k][n_loci_used], f32 in (0, 1)
 *   log_probs         HOST array [n_cells][k]: the log_binoms of the last iteration (main.rs:164); may be NULL
 *   centers_out       HOST array [k][n_loci_used]: centers after the last update; may be NULL
 *   total_log_loss    loss of the last iteration (main.rs:167), what souporcell_main compares between restarts (:73)
 *   iterations        EM iterations run (all temperature steps)
 * Sums are f64 in a fixed order (the reference: f32, serial).  Refused with CLX_ERR_ARG: a (locus, cell) coordinate
 * stored more than once (the reference keeps its last line, main.rs:421) and counts above 65534.  With cells sharded
 * over ranks (clx_comm_init) every rank calls it with the same centers; sums and loss are all-reduced. */
int clx_soup_em(clx_ctx* ctx, uint32_t k, const float* centers_init, double* log_probs, double* centers_out,
                double* total_log_loss, uint32_t* iterations);

/* Best of n_restarts such runs (replaces the restart loop of souporcell_main, main.rs:69-88, which gives one restart
 * at a time to each rayon thread): restarts are run R at a time over the same pass of the matrix (R = 8 for k <= 2,
 * else 16 / k), each with its own loop state, results identical to n_restarts calls of clx_soup_em.
 *   centers_init      HOST array [n_restarts][k][n_loci_used]
 *   log_probs, centers_out   of the best restart (strictly largest total_log_loss, the first one on ties: main.rs:73, 84)
 *   losses, iterations       HOST arrays [n_restarts]; may be NULL
 *   best              index of the best restart; may be NULL */
int clx_soup_restarts(clx_ctx* ctx, uint32_t k, uint32_t n_restarts, const float* centers_init, double* log_probs,
                      double* centers_out, double* losses, uint32_t* iterations, uint32_t* best);

/* Page-locked host memory for the COO arrays handed to clx_load_coo (what a host parser fills; replaces the
 * row-list Vec growth of load_data.rs:165-173).  Works without a context; returns NULL on failure. */
void* clx_alloc_pinned(size_t bytes);
void clx_free_pinned(void* p);

/* The locus filter alone (the rule of get_loci_used), allocating nothing.  This entry point has NO counterpart in the
 * reference: it was added for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it).  Its scratch (two u32 counters per locus) is a
 * file-scope static __device__ array inside the library, sized for CLX_STATIC_MAX_LOCI loci: 16 MiB + 256 B that every
 * GPU the library is used on holds until the process exits.  n_loci0 > CLX_STATIC_MAX_LOCI is rejected with
 * CLX_ERR_ARG.  It needs no context; errors are read with clx_last_error(NULL).
 *   d_locus, d_ref, d_alt   DEVICE arrays of nnz0 entries on `device` (same meaning as in clx_load_coo)
 *   d_used                  DEVICE array of n_loci0 bytes: 1 where the locus passes min_ref / min_alt, else 0
 *   d_n_used                DEVICE word: number of loci kept (may be NULL)
 *   stream                  cudaStream_t of `device` to enqueue on (NULL = the default stream); synchronised on return
 * The call makes `device` current, and calls no cudaMalloc, cudaFree or allocating library routine.  All calls on a GPU
 * share the one array (it is cleared at the start of each call): calls that overlap on the same GPU race and must be
 * serialised by the caller. */
#define CLX_STATIC_MAX_LOCI (1u << 21)
int clx_locus_filter_static(int device, uint64_t nnz0, uint32_t n_loci0, const uint32_t* d_locus,
                            const uint32_t* d_ref, const uint32_t* d_alt, uint32_t min_ref, uint32_t min_alt,
                            uint8_t* d_used, uint32_t* d_n_used, void* stream);

/* Profiling helpers (not part of the reference surface). */
int clx_get_phase_ms(clx_ctx* ctx, int phase, double* total_ms, uint64_t* launches);
/* Switch CLX_FLAG_PROFILE on or off after clx_create (the phase events are extra nodes of the iteration graph and
 * cost tens of microseconds per iteration: a throughput run keeps them off, a per-kernel timing run turns them on). */
int clx_set_profiling(clx_ctx* ctx, int on);
int clx_reset_phase_ms(clx_ctx* ctx);
void* clx_get_stream(clx_ctx* ctx); /* cudaStream_t all kernels are launched on */
/* Algorithmic bytes one launch of the scoring kernel must move (roofline accounting). */
int clx_score_bytes(clx_ctx* ctx, uint64_t* bytes);
/* Number of kernels launched by this context so far. */
int clx_launch_count(clx_ctx* ctx, uint64_t* launches);
/* Overwrite a buffer larger than L2 on the context stream (bench hygiene between timed steps). */
int clx_flush_l2(clx_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* CELLECTOR_CUDA_H */

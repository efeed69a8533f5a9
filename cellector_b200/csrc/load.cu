// load.cu — device-side replacement of load_data::load_cell_data (cellector/src/load_data.rs:134-181):
//   pass 1  get_loci_used (load_data.rs:254-280): per locus count the entries with ref>0 and with alt>0,
//           keep loci with both counts >= min, renumber densely in increasing locus order  -> GPU histogram +
//           exclusive scan + stream compaction of the nonzeros;
//   pass 2  per-cell row lists in file order (load_data.rs:151-174) -> packed CSR (stable sort by cell) and,
//           for the per-locus passes, packed CSC (stable sort by locus: cells ascending inside a column,
//           the order get_locus_log_likelihoods (main.rs:392-407) visits them in);
//           locus_counts[l] = [sum ref, sum alt] (load_data.rs:157-158) -> u64 atomics (integers: order-free).
// Zero-total entries and duplicated coordinates are kept as separate entries, like the reference.
// Entries whose total exceeds K_LIGHT (or whose counts do not fit u16) are additionally listed in the
// "heavy" side lists with full 32-bit counts; the packed matrices only carry a marker for them.
#include <cub/cub.cuh>

#include "ctx.h"
#include "numerics.cuh"

namespace clx {
namespace {

constexpr int TPB = 256;
inline unsigned grid_for(uint64_t n, int tpb = TPB, unsigned cap = 148u * 16u) {
    uint64_t g = (n + tpb - 1) / tpb;
    if (g < 1) g = 1;
    return (unsigned)(g > cap ? cap : g);
}

// every cell index in range (the cell array arrives on the side stream while the locus filter runs)
__global__ void k_cell_check(const uint32_t* __restrict__ cell, uint64_t nnz0, uint32_t n_cells, unsigned int* __restrict__ err) {
    bool bad = false;
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < nnz0; i += (uint64_t)gridDim.x * blockDim.x)
        bad |= cell[i] >= n_cells;
    if (bad) atomicOr(err, 1u);
}

__global__ void k_locus_hist(const uint32_t* __restrict__ locus, const uint32_t* __restrict__ ref,
                             const uint32_t* __restrict__ alt, uint64_t nnz0, uint32_t n_loci0,
                             unsigned int* __restrict__ cnt, unsigned int* __restrict__ err) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < nnz0; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t l = locus[i];
        if (l >= n_loci0) {
            atomicOr(err, 1u);
            continue;
        }
        if (ref[i] > 0) atomicAdd(&cnt[2 * (size_t)l], 1u);
        if (alt[i] > 0) atomicAdd(&cnt[2 * (size_t)l + 1], 1u);
    }
}

__global__ void k_locus_flags(const unsigned int* __restrict__ cnt, uint32_t n_loci0, uint32_t min_ref,
                              uint32_t min_alt, uint32_t* __restrict__ flag) {
    const uint32_t l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l < n_loci0) flag[l] = (cnt[2 * (size_t)l] >= min_ref && cnt[2 * (size_t)l + 1] >= min_alt) ? 1u : 0u;
}

__global__ void k_locus_ids(const uint32_t* __restrict__ flag, const uint32_t* __restrict__ used_idx,
                            uint32_t n_loci0, uint32_t* __restrict__ locus_ids) {
    const uint32_t l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l < n_loci0 && flag[l]) locus_ids[used_idx[l]] = l;
}

struct KeepEntry {
    const uint32_t* locus;
    const uint32_t* flag;
    __device__ bool operator()(uint32_t i) const { return flag[locus[i]] != 0; }
};
struct FlagAt {
    const uint8_t* flag;
    const uint32_t* idx;  // optional indirection
    __device__ bool operator()(uint32_t i) const { return flag[idx ? idx[i] : i] != 0; }
};
struct GiantAt {
    const uint32_t* ref;
    const uint32_t* alt;
    uint64_t limit;
    __device__ bool operator()(uint32_t h) const { return (uint64_t)ref[h] + (uint64_t)alt[h] > limit; }
};
struct DeepAt {  // beyond the grouped chains, up to the giants
    const uint32_t* ref;
    const uint32_t* alt;
    __device__ bool operator()(uint32_t h) const {
        const uint64_t n = (uint64_t)ref[h] + (uint64_t)alt[h];
        return n > (uint64_t)HEAVY_CHAIN_TOTAL && n <= (uint64_t)HEAVY_GIANT_TOTAL;
    }
};
// (locus, total) key of a heavy entry: the sort puts the heavy entries of a locus next to each other (the column pass
// reads them as one contiguous range) and equal totals of a locus into one group (their chain is evaluated once)
__global__ void k_heavy_group_keys(const uint32_t* __restrict__ h_locus, const uint32_t* __restrict__ h_ref,
                                   const uint32_t* __restrict__ h_alt, uint32_t n_heavy, unsigned long long* __restrict__ keys,
                                   uint32_t* __restrict__ idx) {
    const uint32_t h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= n_heavy) return;
    const uint64_t n = (uint64_t)h_ref[h] + (uint64_t)h_alt[h];
    keys[h] = ((unsigned long long)h_locus[h] << 32) | (unsigned long long)(n > 0xFFFFFFFFull ? 0xFFFFFFFFull : n);
    idx[h] = h;
}
struct GroupHead {
    const unsigned long long* keys;
    __device__ bool operator()(uint32_t i) const { return i == 0 || keys[i] != keys[i - 1]; }
};
// member-order copies: {cell, ref, alt, row-order index} per member, the inverse permutation, per-locus ranges
__global__ void k_heavy_members(const uint32_t* __restrict__ member, const unsigned long long* __restrict__ keys_sorted,
                                uint32_t n_heavy, const uint32_t* __restrict__ h_cell, const uint32_t* __restrict__ h_ref,
                                const uint32_t* __restrict__ h_alt, uint4* __restrict__ ent, uint32_t* __restrict__ pos) {
    const uint32_t m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= n_heavy) return;
    const uint32_t h = member[m];
    ent[m] = make_uint4(h_cell[h], h_ref[h], h_alt[h], h);
    pos[h] = m;
    (void)keys_sorted;
}
__global__ void k_heavy_col_ptr(const unsigned long long* __restrict__ keys_sorted, uint32_t n_heavy, uint32_t n_loci,
                                uint32_t* __restrict__ col_ptr) {
    const uint32_t l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l > n_loci) return;
    const unsigned long long want = (unsigned long long)l << 32;
    uint32_t lo = 0, hi = n_heavy;
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (keys_sorted[mid] < want) lo = mid + 1;
        else hi = mid;
    }
    col_ptr[l] = lo;
}
struct ToU64 {
    __device__ uint64_t operator()(uint32_t v) const { return (uint64_t)v; }
};

__global__ void k_cell_keys(const uint32_t* __restrict__ kept_idx, const uint32_t* __restrict__ cell, uint64_t nnz,
                            uint32_t n_cells, uint32_t* __restrict__ keys, uint32_t* __restrict__ row_cnt) {
    for (uint64_t k = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; k < nnz; k += (uint64_t)gridDim.x * blockDim.x) {
        uint32_t c = cell[kept_idx[k]];
        if (c >= n_cells) c = n_cells - 1;  // k_cell_check has flagged it; the load fails at its next host sync
        keys[k] = c;
        atomicAdd(&row_cnt[c], 1u);
    }
}

__global__ void k_pack_csr(const uint32_t* __restrict__ csr_src, const uint32_t* __restrict__ locus,
                           const uint32_t* __restrict__ ref, const uint32_t* __restrict__ alt,
                           const uint32_t* __restrict__ used_idx, uint64_t nnz, RowEntry* __restrict__ csr,
                           uint32_t* __restrict__ keys2, uint32_t* __restrict__ vals2, uint8_t* __restrict__ heavy,
                           uint32_t* __restrict__ col_cnt, unsigned long long* __restrict__ locus_tot,
                           uint32_t* __restrict__ cls_all) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < nnz; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t src = csr_src[i];
        const uint32_t l = used_idx[locus[src]];
        const uint32_t r = ref[src], a = alt[src];
        const bool hv = ((uint64_t)r + (uint64_t)a > (uint64_t)K_LIGHT);
        RowEntry e;
        e.locus = l;
        e.ref = (uint16_t)(r > 0xFFFFu ? 0xFFFFu : r);
        e.alt = (uint16_t)(a > 0xFFFFu ? 0xFFFFu : a);
        csr[i] = e;
        keys2[i] = l;
        vals2[i] = (uint32_t)i;
        heavy[i] = hv ? 1 : 0;
        atomicAdd(&col_cnt[l], 1u);
        const uint32_t n = e.ref + e.alt;  // clipped counts: anything above K_LIGHT stays above it
        atomicAdd(&cls_all[(size_t)l * CLS_STRIDE + ((n - 1u) < (uint32_t)K_LIGHT ? (uint32_t)tab_slot((int)e.alt, (int)n) : (uint32_t)CLS_OTHER)], 1u);
        if (r) atomicAdd(&locus_tot[2 * (size_t)l], (unsigned long long)r);
        if (a) atomicAdd(&locus_tot[2 * (size_t)l + 1], (unsigned long long)a);
    }
}

__global__ void k_pack_csc(const uint32_t* __restrict__ csc_src, const uint32_t* __restrict__ csr_src,
                           const uint32_t* __restrict__ cell, const RowEntry* __restrict__ csr, uint64_t nnz,
                           ColEntry* __restrict__ csc) {
    for (uint64_t j = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; j < nnz; j += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t p = csc_src[j];
        const RowEntry e = csr[p];
        ColEntry o;
        o.cell = cell[csr_src[p]];
        o.ref = e.ref;
        o.alt = e.alt;
        csc[j] = o;
    }
}

__global__ void k_heavy_gather(const uint32_t* __restrict__ hpos, uint64_t n_heavy, const uint32_t* __restrict__ csr_src,
                               const RowEntry* __restrict__ csr, const uint32_t* __restrict__ cell,
                               const uint32_t* __restrict__ ref, const uint32_t* __restrict__ alt,
                               uint32_t* __restrict__ h_cell, uint32_t* __restrict__ h_locus,
                               uint32_t* __restrict__ h_ref, uint32_t* __restrict__ h_alt) {
    const uint64_t h = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    if (h >= n_heavy) return;
    const uint32_t p = hpos[h];
    const uint32_t src = csr_src[p];
    h_cell[h] = cell[src];
    h_locus[h] = csr[p].locus;
    h_ref[h] = ref[src];
    h_alt[h] = alt[src];
}

__device__ __forceinline__ uint32_t lower_bound_u32(const uint32_t* __restrict__ a, uint32_t n, uint64_t v) {
    uint32_t lo = 0, hi = n;
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if ((uint64_t)a[mid] < v) lo = mid + 1;
        else hi = mid;
    }
    return lo;
}
// out[i] = number of (sorted) positions < ptr[i], for i in [0, n]
__global__ void k_segment_ptr(const uint32_t* __restrict__ pos, uint32_t n_pos, const uint64_t* __restrict__ ptr,
                              uint32_t n, uint32_t* __restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i <= n) out[i] = lower_bound_u32(pos, n_pos, ptr[i]);
}
__global__ void k_heavy_perm(const uint32_t* __restrict__ hcpos, const uint32_t* __restrict__ csc_src,
                             const uint32_t* __restrict__ hpos, uint32_t n_heavy, uint32_t* __restrict__ perm) {
    const uint32_t h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h < n_heavy) perm[h] = lower_bound_u32(hpos, n_heavy, (uint64_t)csc_src[hcpos[h]]);
}

__global__ void k_max_total(const unsigned long long* __restrict__ locus_tot, uint32_t n_loci, unsigned long long* __restrict__ out) {
    unsigned long long m = 0;
    for (uint32_t l = blockIdx.x * blockDim.x + threadIdx.x; l < n_loci; l += gridDim.x * blockDim.x) {
        const unsigned long long t = locus_tot[2 * (size_t)l] + locus_tot[2 * (size_t)l + 1];
        m = t > m ? t : m;
    }
    for (int off = 16; off > 0; off >>= 1) {
        const unsigned long long o = __shfl_xor_sync(0xffffffffu, m, off);
        m = o > m ? o : m;
    }
    if ((threadIdx.x & 31) == 0 && m) atomicMax(out, m);
}
__global__ void k_lng_build(double* __restrict__ lng, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) lng[i] = i == 0 ? INFINITY : ln_gamma((double)i);
}


__global__ void k_unpack_counts(const uint32_t* __restrict__ packed, uint64_t n, uint32_t* __restrict__ ref, uint32_t* __restrict__ alt) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t v = packed[i];
        ref[i] = v & 0xFFFFu;
        alt[i] = v >> 16;
    }
}

int bits_for(uint64_t n) {
    int b = 1;
    while (b < 32 && (1ull << b) < n) ++b;
    return b;
}

}  // namespace

int load_coo(clx_ctx* c, uint64_t nnz0, uint32_t n_loci0, uint32_t n_cells, const uint32_t* locus,
             const uint32_t* cell, const uint32_t* ref, const uint32_t* alt, uint32_t min_ref, uint32_t min_alt) {
    NvtxRange nvtx("clx:load_coo");
    cudaStream_t s = c->stream;
    StageClock clk(s, "load");
    c->loaded = false;
    // every cub call below takes a 32-bit signed item count
    if (nnz0 >= 0x7FFFFFFFull) return fail(c, CLX_ERR_ARG, "more than 2^31-2 nonzeros per GPU: shard the cells over more ranks");
    if (nnz0 > 0 && (!locus || !cell || !ref)) return fail(c, CLX_ERR_ARG, "null COO array");
    if (nnz0 > 0 && n_cells == 0) return fail(c, CLX_ERR_ARG, "COO coordinate out of range (locus >= n_loci0 or cell >= n_cells)");
    c->n_cells = n_cells;
    c->n_loci0 = n_loci0;
    c->nnz0 = nnz0;

    // ---- shard bookkeeping (cells are sharded across ranks; loci are global)
    {
        std::vector<uint64_t> all;
        CLX_TRY(comm_allgather_u64_host(c, (uint64_t)n_cells, &all));
        c->n_cells_global = 0;
        c->cell_offset = 0;
        uint64_t mx = 0;
        for (int r = 0; r < c->world; ++r) {
            if (r < c->rank) c->cell_offset += all[r];
            c->n_cells_global += all[r];
            mx = all[r] > mx ? all[r] : mx;
        }
        c->max_local_cells = (uint32_t)mx;
    }

    // ---- COO to device
    DevBuf<uint32_t>&d_locus = c->coo_locus, &d_cell = c->coo_cell, &d_ref = c->coo_ref, &d_alt = c->coo_alt;
    CLX_CUDA(c, d_locus.alloc(nnz0));
    CLX_CUDA(c, d_cell.alloc(nnz0));
    CLX_CUDA(c, d_ref.alloc(nnz0));
    CLX_CUDA(c, d_alt.alloc(nnz0));
    if (nnz0) {
        // cudaMemcpyDefault: the caller's arrays may be pageable or pinned host memory, or device memory (a host that
        // parses or generates on the GPU); the copy engine picks the path from the pointer
        CLX_CUDA(c, cudaMemcpyAsync(d_locus.p, locus, nnz0 * 4, cudaMemcpyDefault, s));
        if (alt) {
            CLX_CUDA(c, cudaMemcpyAsync(d_ref.p, ref, nnz0 * 4, cudaMemcpyDefault, s));
            CLX_CUDA(c, cudaMemcpyAsync(d_alt.p, alt, nnz0 * 4, cudaMemcpyDefault, s));
        } else {  // both counts in one word: 4 bytes less per entry over PCIe, split on the device
            DevBuf<uint32_t> d_packed;
            CLX_CUDA(c, d_packed.alloc(nnz0));
            CLX_CUDA(c, cudaMemcpyAsync(d_packed.p, ref, nnz0 * 4, cudaMemcpyDefault, s));
            k_unpack_counts<<<grid_for(nnz0), TPB, 0, s>>>(d_packed.p, nnz0, d_ref.p, d_alt.p);
            c->launches++;
        }
        // the cell indices travel on the side stream, after the arrays the locus filter needs have landed: the filter,
        // the scan and the compaction of the kept entries run while they are still on the wire
        CLX_CUDA(c, cudaEventRecord(c->ev_fork, s));
        CLX_CUDA(c, cudaStreamWaitEvent(c->stream2, c->ev_fork, 0));
        CLX_CUDA(c, cudaMemcpyAsync(d_cell.p, cell, nnz0 * 4, cudaMemcpyDefault, c->stream2));
        CLX_CUDA(c, cudaEventRecord(c->ev_join, c->stream2));
    }

    clk.mark("COO to device");
    // ---- pass 1: locus filter
    DevBuf<unsigned int> cnt, err;
    DevBuf<uint32_t> flag, used_idx;
    CLX_CUDA(c, cnt.alloc(2 * (size_t)n_loci0 + 2));
    CLX_CUDA(c, err.alloc(1));
    CLX_CUDA(c, flag.alloc((size_t)n_loci0 + 1));
    CLX_CUDA(c, used_idx.alloc((size_t)n_loci0 + 1));
    CLX_CUDA(c, cudaMemsetAsync(cnt.p, 0, cnt.bytes(), s));
    CLX_CUDA(c, cudaMemsetAsync(err.p, 0, 4, s));
    CLX_CUDA(c, cudaMemsetAsync(flag.p, 0, flag.bytes(), s));
    if (nnz0) {
        k_locus_hist<<<grid_for(nnz0), TPB, 0, s>>>(d_locus.p, d_ref.p, d_alt.p, nnz0, n_loci0, cnt.p, err.p);
        c->launches++;
    }
    CLX_TRY(comm_allreduce_u32(c, cnt.p, 2 * (size_t)n_loci0));
    if (n_loci0) {
        k_locus_flags<<<(n_loci0 + TPB - 1) / TPB, TPB, 0, s>>>(cnt.p, n_loci0, min_ref, min_alt, flag.p);
        c->launches++;
    }
    DevBuf<uint8_t> tmp;
    size_t tmp_bytes = 0;
    // exclusive scan over n_loci0 + 1 flags (last one is 0) so that used_idx[n_loci0] = n_loci
    CLX_CUDA(c, cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, flag.p, used_idx.p, (int)(n_loci0 + 1), s));
    CLX_CUDA(c, tmp.alloc(tmp_bytes));
    CLX_CUDA(c, cub::DeviceScan::ExclusiveSum(tmp.p, tmp_bytes, flag.p, used_idx.p, (int)(n_loci0 + 1), s));
    c->launches += 2;
    uint32_t n_loci = 0;
    unsigned int h_err = 0;
    CLX_CUDA(c, cudaMemcpyAsync(&n_loci, used_idx.p + n_loci0, 4, cudaMemcpyDeviceToHost, s));
    CLX_CUDA(c, cudaMemcpyAsync(&h_err, err.p, 4, cudaMemcpyDeviceToHost, s));
    CLX_CUDA(c, cudaStreamSynchronize(s));
    if (h_err) return fail(c, CLX_ERR_ARG, "COO coordinate out of range (locus >= n_loci0 or cell >= n_cells)");
    c->n_loci = n_loci;
    CLX_CUDA(c, c->locus_ids.alloc(n_loci));
    if (n_loci0) {
        k_locus_ids<<<(n_loci0 + TPB - 1) / TPB, TPB, 0, s>>>(flag.p, used_idx.p, n_loci0, c->locus_ids.p);
        c->launches++;
    }

    clk.mark("locus filter");
    // ---- compaction of the kept nonzeros (file order preserved)
    DevBuf<uint32_t> kept_idx, n_sel;
    CLX_CUDA(c, kept_idx.alloc(nnz0));
    CLX_CUDA(c, n_sel.alloc(1));
    uint32_t nnz = 0;
    if (nnz0) {
        cub::CountingInputIterator<uint32_t> counting(0);
        KeepEntry keep{d_locus.p, flag.p};
        CLX_CUDA(c, cub::DeviceSelect::If(nullptr, tmp_bytes, counting, kept_idx.p, n_sel.p, (int)nnz0, keep, s));
        CLX_CUDA(c, tmp.alloc(tmp_bytes));
        CLX_CUDA(c, cub::DeviceSelect::If(tmp.p, tmp_bytes, counting, kept_idx.p, n_sel.p, (int)nnz0, keep, s));
        c->launches += 2;
        CLX_CUDA(c, cudaMemcpyAsync(&nnz, n_sel.p, 4, cudaMemcpyDeviceToHost, s));
        CLX_CUDA(c, cudaStreamSynchronize(s));
    }
    c->nnz = nnz;

    clk.mark("compaction");
    // ---- CSR: stable sort by cell (the cell indices have arrived by now, or we wait for them here)
    if (nnz0) {
        CLX_CUDA(c, cudaStreamWaitEvent(s, c->ev_join, 0));
        k_cell_check<<<grid_for(nnz0), TPB, 0, s>>>(d_cell.p, nnz0, n_cells, err.p);
        c->launches++;
    }
    DevBuf<uint32_t> keys, keys_sorted, csr_src, row_cnt;
    CLX_CUDA(c, keys.alloc(nnz));
    CLX_CUDA(c, keys_sorted.alloc(nnz));
    CLX_CUDA(c, csr_src.alloc(nnz));
    CLX_CUDA(c, row_cnt.alloc((size_t)n_cells + 1));
    CLX_CUDA(c, cudaMemsetAsync(row_cnt.p, 0, row_cnt.bytes(), s));
    CLX_CUDA(c, c->row_ptr.alloc((size_t)n_cells + 1));
    if (nnz) {
        k_cell_keys<<<grid_for(nnz), TPB, 0, s>>>(kept_idx.p, d_cell.p, nnz, n_cells, keys.p, row_cnt.p);
        c->launches++;
        const int eb = bits_for(n_cells);
        CLX_CUDA(c, cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys.p, keys_sorted.p, kept_idx.p, csr_src.p, (int)nnz, 0, eb, s));
        CLX_CUDA(c, tmp.alloc(tmp_bytes));
        CLX_CUDA(c, cub::DeviceRadixSort::SortPairs(tmp.p, tmp_bytes, keys.p, keys_sorted.p, kept_idx.p, csr_src.p, (int)nnz, 0, eb, s));
        c->launches += 4;
    }
    {
        cub::TransformInputIterator<uint64_t, ToU64, const uint32_t*> it(row_cnt.p, ToU64());
        CLX_CUDA(c, cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, it, c->row_ptr.p, (int)(n_cells + 1), s));
        CLX_CUDA(c, tmp.alloc(tmp_bytes));
        CLX_CUDA(c, cub::DeviceScan::ExclusiveSum(tmp.p, tmp_bytes, it, c->row_ptr.p, (int)(n_cells + 1), s));
        c->launches += 2;
    }

    clk.mark("CSR sort");
    // ---- pack CSR, locus totals, CSC keys
    DevBuf<uint8_t> heavy;
    DevBuf<uint32_t> col_cnt, vals2, csc_src;
    CLX_CUDA(c, c->csr.alloc((size_t)nnz + 2));  // +2: the 128-bit loads of the gather kernels may touch one entry past the end
    CLX_CUDA(c, heavy.alloc(nnz));
    CLX_CUDA(c, col_cnt.alloc((size_t)n_loci + 1));
    CLX_CUDA(c, vals2.alloc(nnz));
    CLX_CUDA(c, csc_src.alloc(nnz));
    CLX_CUDA(c, c->locus_tot.alloc(2 * (size_t)n_loci + 2));
    CLX_CUDA(c, cudaMemsetAsync(col_cnt.p, 0, col_cnt.bytes(), s));
    CLX_CUDA(c, cudaMemsetAsync(c->locus_tot.p, 0, c->locus_tot.bytes(), s));
    CLX_CUDA(c, c->cls_all.alloc(((size_t)n_loci + 1) * CLS_STRIDE));
    CLX_CUDA(c, cudaMemsetAsync(c->cls_all.p, 0, c->cls_all.bytes(), s));
    CLX_CUDA(c, c->col_ptr.alloc((size_t)n_loci + 1));
    CLX_CUDA(c, c->csc.alloc((size_t)nnz + 2));
    if (nnz) {
        // keys / keys_sorted are reused as the locus keys of the second sort
        k_pack_csr<<<grid_for(nnz), TPB, 0, s>>>(csr_src.p, d_locus.p, d_ref.p, d_alt.p, used_idx.p, nnz, c->csr.p, keys.p,
                                                 vals2.p, heavy.p, col_cnt.p, c->locus_tot.p, c->cls_all.p);
        c->launches++;
        clk.mark("pack CSR");
        // tiled scoring layout: keys_sorted still holds the cell of every CSR entry here
        CLX_TRY(build_tiled(c, keys_sorted.p));
        clk.mark("run stream");
        const int eb = bits_for(n_loci);
        CLX_CUDA(c, cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys.p, keys_sorted.p, vals2.p, csc_src.p, (int)nnz, 0, eb, s));
        CLX_CUDA(c, tmp.alloc(tmp_bytes));
        CLX_CUDA(c, cub::DeviceRadixSort::SortPairs(tmp.p, tmp_bytes, keys.p, keys_sorted.p, vals2.p, csc_src.p, (int)nnz, 0, eb, s));
        c->launches += 4;
        k_pack_csc<<<grid_for(nnz), TPB, 0, s>>>(csc_src.p, csr_src.p, d_cell.p, c->csr.p, nnz, c->csc.p);
        c->launches++;
    }
    if (!nnz) CLX_TRY(build_tiled(c, nullptr));
    {
        cub::TransformInputIterator<uint64_t, ToU64, const uint32_t*> it(col_cnt.p, ToU64());
        CLX_CUDA(c, cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, it, c->col_ptr.p, (int)(n_loci + 1), s));
        CLX_CUDA(c, tmp.alloc(tmp_bytes));
        CLX_CUDA(c, cub::DeviceScan::ExclusiveSum(tmp.p, tmp_bytes, it, c->col_ptr.p, (int)(n_loci + 1), s));
        c->launches += 2;
    }
    CLX_TRY(comm_allreduce_u64(c, c->locus_tot.p, 2 * (size_t)n_loci));
    // ln_gamma memo over the integer arguments the EM pass can form: k + alpha, j + beta <= 2 * total + 2 and
    // (a + alpha) + (r + beta) <= 2 * total + 2 for the deepest locus (capped; beyond the cap ln_gamma is direct)
    {
        DevBuf<unsigned long long> d_max;
        CLX_CUDA(c, d_max.alloc(1));
        CLX_CUDA(c, cudaMemsetAsync(d_max.p, 0, 8, s));
        if (n_loci) {
            k_max_total<<<grid_for(n_loci), TPB, 0, s>>>(c->locus_tot.p, n_loci, d_max.p);
            c->launches++;
        }
        unsigned long long h_max = 0;
        CLX_CUDA(c, cudaMemcpyAsync(&h_max, d_max.p, 8, cudaMemcpyDeviceToHost, s));
        CLX_CUDA(c, cudaMemcpyAsync(&h_err, err.p, 4, cudaMemcpyDeviceToHost, s));
        CLX_CUDA(c, cudaStreamSynchronize(s));
        if (h_err) return fail(c, CLX_ERR_ARG, "COO coordinate out of range (locus >= n_loci0 or cell >= n_cells)");
        const unsigned long long want = 2 * h_max + 16, cap = 1ull << 22;  // at most 32 MB
        c->lng_n = (uint32_t)(want < cap ? want : cap);
        CLX_CUDA(c, c->lng.alloc(c->lng_n));
        k_lng_build<<<(c->lng_n + TPB - 1) / TPB, TPB, 0, s>>>(c->lng.p, c->lng_n);
        c->launches++;
    }

    clk.mark("CSC + ln_gamma memo");
    // ---- heavy side lists
    DevBuf<uint32_t> hpos, hcpos;
    CLX_CUDA(c, hpos.alloc(nnz));
    uint32_t n_heavy = 0;
    if (nnz) {
        cub::CountingInputIterator<uint32_t> counting(0);
        FlagAt f1{heavy.p, nullptr};
        CLX_CUDA(c, cub::DeviceSelect::If(nullptr, tmp_bytes, counting, hpos.p, n_sel.p, (int)nnz, f1, s));
        CLX_CUDA(c, tmp.alloc(tmp_bytes));
        CLX_CUDA(c, cub::DeviceSelect::If(tmp.p, tmp_bytes, counting, hpos.p, n_sel.p, (int)nnz, f1, s));
        c->launches += 2;
        CLX_CUDA(c, cudaMemcpyAsync(&n_heavy, n_sel.p, 4, cudaMemcpyDeviceToHost, s));
        CLX_CUDA(c, cudaStreamSynchronize(s));
    }
    c->n_heavy = n_heavy;
    CLX_CUDA(c, c->h_cell.alloc(n_heavy));
    CLX_CUDA(c, c->h_locus.alloc(n_heavy));
    CLX_CUDA(c, c->h_ref.alloc(n_heavy));
    CLX_CUDA(c, c->h_alt.alloc(n_heavy));
    CLX_CUDA(c, c->hc_perm.alloc(n_heavy));
    CLX_CUDA(c, c->h_row_ptr.alloc((size_t)n_cells + 1));
    CLX_CUDA(c, c->hc_col_ptr.alloc((size_t)n_loci + 1));
    CLX_CUDA(c, c->hv_ll.alloc(3 * (size_t)n_heavy));
    CLX_CUDA(c, c->hv_e.alloc(n_heavy));
    CLX_CUDA(c, hcpos.alloc(n_heavy));
    if (n_heavy) {
        k_heavy_gather<<<(n_heavy + TPB - 1) / TPB, TPB, 0, s>>>(hpos.p, n_heavy, csr_src.p, c->csr.p, d_cell.p, d_ref.p,
                                                               d_alt.p, c->h_cell.p, c->h_locus.p, c->h_ref.p, c->h_alt.p);
        cub::CountingInputIterator<uint32_t> counting(0);
        FlagAt f2{heavy.p, csc_src.p};
        CLX_CUDA(c, cub::DeviceSelect::If(nullptr, tmp_bytes, counting, hcpos.p, n_sel.p, (int)nnz, f2, s));
        CLX_CUDA(c, tmp.alloc(tmp_bytes));
        CLX_CUDA(c, cub::DeviceSelect::If(tmp.p, tmp_bytes, counting, hcpos.p, n_sel.p, (int)nnz, f2, s));
        k_heavy_perm<<<(n_heavy + TPB - 1) / TPB, TPB, 0, s>>>(hcpos.p, csc_src.p, hpos.p, n_heavy, c->hc_perm.p);
        c->launches += 4;
    }
    // heavy entries in (locus, total) order: groups of equal keys, member-order copies, per-locus ranges
    CLX_CUDA(c, c->hg_member.alloc(n_heavy));
    CLX_CUDA(c, c->hg_start.alloc((size_t)n_heavy + 1));
    CLX_CUDA(c, c->hg_count.alloc(2));
    CLX_CUDA(c, c->hm_ent.alloc((size_t)n_heavy + 1));
    CLX_CUDA(c, c->hm_pos.alloc((size_t)n_heavy + 1));
    CLX_CUDA(c, c->hm_val.alloc((size_t)n_heavy + 1));
    CLX_CUDA(c, c->hm_col_ptr.alloc((size_t)n_loci + 2));
    CLX_CUDA(c, cudaMemsetAsync(c->hg_count.p, 0, 8, s));
    CLX_CUDA(c, cudaMemsetAsync(c->hm_col_ptr.p, 0, c->hm_col_ptr.bytes(), s));
    if (n_heavy) {
        DevBuf<unsigned long long> gk, gk_sorted;
        DevBuf<uint32_t> gi;
        CLX_CUDA(c, gk.alloc(n_heavy));
        CLX_CUDA(c, gk_sorted.alloc(n_heavy));
        CLX_CUDA(c, gi.alloc(n_heavy));
        k_heavy_group_keys<<<(n_heavy + TPB - 1) / TPB, TPB, 0, s>>>(c->h_locus.p, c->h_ref.p, c->h_alt.p, n_heavy, gk.p, gi.p);
        CLX_CUDA(c, cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, gk.p, gk_sorted.p, gi.p, c->hg_member.p, (int)n_heavy, 0, 64, s));
        CLX_CUDA(c, tmp.alloc(tmp_bytes));
        CLX_CUDA(c, cub::DeviceRadixSort::SortPairs(tmp.p, tmp_bytes, gk.p, gk_sorted.p, gi.p, c->hg_member.p, (int)n_heavy, 0, 64, s));
        cub::CountingInputIterator<uint32_t> counting(0);
        GroupHead gh{gk_sorted.p};
        CLX_CUDA(c, cub::DeviceSelect::If(nullptr, tmp_bytes, counting, c->hg_start.p, c->hg_count.p, (int)n_heavy, gh, s));
        CLX_CUDA(c, tmp.alloc(tmp_bytes));
        CLX_CUDA(c, cub::DeviceSelect::If(tmp.p, tmp_bytes, counting, c->hg_start.p, c->hg_count.p, (int)n_heavy, gh, s));
        k_heavy_members<<<(n_heavy + TPB - 1) / TPB, TPB, 0, s>>>(c->hg_member.p, gk_sorted.p, n_heavy, c->h_cell.p, c->h_ref.p, c->h_alt.p,
                                                                c->hm_ent.p, c->hm_pos.p);
        k_heavy_col_ptr<<<(n_loci + 1 + TPB - 1) / TPB, TPB, 0, s>>>(gk_sorted.p, n_heavy, n_loci, c->hm_col_ptr.p);
        c->launches += 9;
    }
    // entries too deep for the batched heavy kernel get a CTA each (em_kernels.cuh: k_heavy_giant)
    c->n_giant = 0;
    c->n_deep = 0;
    CLX_CUDA(c, c->giant_idx.alloc(n_heavy));
    CLX_CUDA(c, c->deep_idx.alloc(n_heavy));
    if (n_heavy) {
        cub::CountingInputIterator<uint32_t> counting(0);
        GiantAt gp{c->h_ref.p, c->h_alt.p, (uint64_t)HEAVY_GIANT_TOTAL};
        CLX_CUDA(c, cub::DeviceSelect::If(nullptr, tmp_bytes, counting, c->giant_idx.p, n_sel.p, (int)n_heavy, gp, s));
        CLX_CUDA(c, tmp.alloc(tmp_bytes));
        CLX_CUDA(c, cub::DeviceSelect::If(tmp.p, tmp_bytes, counting, c->giant_idx.p, n_sel.p, (int)n_heavy, gp, s));
        c->launches += 2;
        uint32_t ng = 0, nd = 0;
        CLX_CUDA(c, cudaMemcpyAsync(&ng, n_sel.p, 4, cudaMemcpyDeviceToHost, s));
        // deep entries (HEAVY_CHAIN_TOTAL < total <= HEAVY_GIANT_TOTAL): a warp each in k_heavy_em
        DeepAt dp{c->h_ref.p, c->h_alt.p};
        CLX_CUDA(c, cub::DeviceSelect::If(nullptr, tmp_bytes, counting, c->deep_idx.p, n_sel.p, (int)n_heavy, dp, s));
        CLX_CUDA(c, tmp.alloc(tmp_bytes));
        CLX_CUDA(c, cub::DeviceSelect::If(tmp.p, tmp_bytes, counting, c->deep_idx.p, n_sel.p, (int)n_heavy, dp, s));
        c->launches += 2;
        CLX_CUDA(c, cudaMemcpyAsync(&nd, n_sel.p, 4, cudaMemcpyDeviceToHost, s));
        CLX_CUDA(c, cudaStreamSynchronize(s));
        c->n_giant = ng;
        c->n_deep = nd;
    }
    k_segment_ptr<<<(n_cells + 1 + TPB - 1) / TPB, TPB, 0, s>>>(hpos.p, n_heavy, c->row_ptr.p, n_cells, c->h_row_ptr.p);
    k_segment_ptr<<<(n_loci + 1 + TPB - 1) / TPB, TPB, 0, s>>>(hcpos.p, n_heavy, c->col_ptr.p, n_loci, c->hc_col_ptr.p);
    c->launches += 2;
    CLX_CUDA(c, cudaStreamSynchronize(s));
    CLX_CUDA(c, cudaGetLastError());

    if (!(c->flags & CLX_FLAG_KEEP_COO)) {
        d_locus.release();
        d_cell.release();
        d_ref.release();
        d_alt.release();
    }
    clk.mark("heavy lists");
    CLX_TRY(em_alloc(c));
    clk.mark("EM buffers");
    c->loaded = true;
    return CLX_OK;
}

}  // namespace clx

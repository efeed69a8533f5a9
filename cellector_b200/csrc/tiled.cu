// tiled.cu — the per-cell scoring path (get_cell_log_likelihoods, cellector/src/main.rs:541-591) as a run-stream
// kernel over shared-memory table tiles.
//
// Why this shape: a memo-table gather served by L1/L2 costs one L1 wavefront per nonzero (measured in round 1: the
// gather kernel k_score_rows stops at ~20 % of the HBM roofline), so the per-locus memo slots are staged in SHARED
// memory by TMA bulk copies (cp.async.bulk + mbarrier), tile by tile, and the nonzeros are re-laid out so that every
// CTA streams exactly the entries that hit the tile it holds.  The kernel is then bound by shared-memory wavefronts
// (one random 16-byte lookup per nonzero), and the layout below is built to spend them on nothing else:
//
//   * only "light" entries (total 1..3) are kept: zero-total entries contribute exactly 0.0 and heavy entries come
//     from the heavy lists; loci_used_per_cell (main.rs:575) is kept as a counter array instead (n_used);
//   * a table slot is 16 bytes {T(alt | total), E(total)} (numerics.cuh): ONE LDS.128 per nonzero fetches the log-pmf
//     and the expected log-likelihood term;
//   * a RUN = the entries of one cell inside one locus tile (TILE_L loci); an entry is one 16-bit word: the slot
//     index inside the staged tile (locus_local * 9 + slot(alt, total));
//   * a cell belongs to ONE warp of the CTA for the whole pass (cell_local % W): its shared-memory accumulator is
//     updated with plain read-modify-writes, no atomics, and there is no CTA barrier between tiles;
//   * the runs of one (cell block, warp, tile) SUB-UNIT are dealt into GROUPS of 32, one run per lane: lane 8j + v
//     always gets a run whose cell has accumulator class v = (cell_local / W) & 7, the runs of a class are sorted by
//     length and taken 4 at a time, so the runs of a group have almost equal lengths and the 8 lanes of a quarter warp
//     own accumulator slots in distinct 16-byte bank groups (the update at the end of a group is conflict-free);
//   * a group runs for its longest run, in PIECES of 4 steps; word 0 of a piece carries its number of valid steps,
//     word 1 the "group ends here" flag (both warp-uniform), so the decoder is straight-line code per piece.  A lane
//     with nothing to do in a step reads one of 8 all-zero slots behind the tile, the one in a bank group no other
//     lane of its quarter warp uses in that step; the ORDER in which a lane takes its run's entries is chosen at load
//     time (k_sub_fill) so that the lookups of a quarter warp spread over the bank groups;
//   * chunk c of a warp is 32 x 16 bytes (8 words per lane), read with one coalesced LDG.128 per lane, STREAM_PF chunks
//     ahead of the one being summed (registers, reloaded in place).  A CELL chunk carries the accumulator indices of the
//     lane's next 8 runs (and "this sub-unit starts the next tile"); ENTRY chunks are two pieces;
//   * sub-units are numbered warp-major ((cell block, warp), then tile) and start on chunk boundaries, so the words a
//     warp consumes over the consecutive tiles of a work item are ONE range of chunks;
//   * a work ITEM = one cell block x a range of (tile, warp) sub-units; the sequence of all sub-units is cut into one
//     contiguous piece of equal cost per SM (build_tiled), static, no queue.  The CTA keeps the block's accumulators in
//     shared memory across the range and writes one partial per cell at the end; k_score_finish adds the partials of
//     a cell block in item order (fixed order: deterministic), adds the heavy values and normalises;
//   * matrices with less than one entry per cell and tile do not get a stream at all: the gather kernel scores them.
#include <cub/cub.cuh>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "ctx.h"
#include "numerics.cuh"

namespace clx {
namespace {

constexpr int TPB = 256;
constexpr uint32_t NB_MAX = 4096;        // cells per block: 16 B of accumulators each = 64 KB of shared memory
constexpr uint32_t W_MAX = 32;           // warps per CTA
constexpr uint32_t CHUNK_WORDS = 8;      // 16-bit words per lane and chunk (one LDG.128)
// Stream format.  A chunk is 8 words per lane; every chunk is either a CELL chunk or an ENTRY chunk, and which one is
// warp-uniform (bit 15 of word 0):
//   cell chunk   word i = the accumulator index of the lane's cell in the i-th of the next (up to) 8 groups
//                (WC_IDLE: the lane has no run in that group); word 0 also carries WC_CELLS and, for the first chunk of
//                a sub-unit, WC_NEWTILE.  The entry chunks of those groups follow; the next cell chunk comes after
//                them, on a chunk boundary.
//   entry chunk  two PIECES of 4 words.  A group is a whole number of pieces, so the decoder is straight-line code per
//                piece: word = slot inside the staged tile; word 0 also carries (valid words of the piece - 1) in
//                WD_CNT, word 1 carries WD_LAST if the group ends with this piece (both warp-uniform).  A lane with no
//                entry in a step reads one of the 8 all-zero slots behind the tile, the one in a bank group no other
//                lane of its quarter warp uses in that step: padding costs no shared-memory wavefront and no branch.
constexpr uint32_t WC_CELLS = 0x8000u;    // word 0 of a cell chunk
constexpr uint32_t WC_NEWTILE = 0x4000u;  // word 0 of a cell chunk: this sub-unit belongs to the next tile of the item
constexpr uint32_t WC_IDLE = 0x1000u;     // cell word: this lane has no run in the group
constexpr uint32_t WC_QMASK = 0x01FFu;    // cell word: index of the lane's cell among the warp's cells
constexpr uint32_t WD_CNT_SHIFT = 13;     // word 0 of a piece: valid words - 1 (2 bits)
constexpr uint32_t WD_LAST = 0x4000u;     // word 1 of a piece: the group ends with this piece
constexpr uint32_t WD_SLOT = 0x1FFFu;     // entry word: slot inside the staged tile
constexpr uint32_t STREAM_PF = 3;          // 512-byte stream chunks in flight per warp (registers)
constexpr uint32_t PIECE_WORDS = 4;
constexpr uint32_t ZERO_SLOTS = 8;        // all-zero slots behind a staged tile, one per 16-byte bank group
constexpr uint32_t GROUPS_PER_CELL_CHUNK = 8;
constexpr uint32_t NULL_SLOT = (uint32_t)TILE_SLOTS;
static_assert(TILE_SLOTS < 0x1FFF, "slot index fits 13 bits");

inline unsigned grid_for(uint64_t n, int tpb = TPB, unsigned cap = 148u * 16u) {
    uint64_t g = (n + tpb - 1) / tpb;
    if (g < 1) g = 1;
    return (unsigned)(g > cap ? cap : g);
}

// ---------------------------------------------------------------------------------------------- builder
// key = ((cell block * nT + tile) << 12 | cell_local) << 3 | bank group of the slot   (sorts to unit, then cell, then
// bank group: the entries of a run come out grouped by the 16-byte bank group their lookup hits, which is what the
// conflict-aware step order of k_sub_fill needs); non-light entries get ~0 and sort to the end
__global__ void k_tile_keys(const RowEntry* __restrict__ csr, const uint32_t* __restrict__ cell_of, uint64_t nnz,
                            uint32_t NB, uint32_t nT, unsigned long long* __restrict__ keys, uint16_t* __restrict__ vals,
                            unsigned long long* __restrict__ n_light, uint32_t* __restrict__ row_len) {
    unsigned long long mine = 0;
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < nnz; i += (uint64_t)gridDim.x * blockDim.x) {
        const RowEntry e = csr[i];
        const uint32_t c = cell_of[i];
        atomicAdd(&row_len[c], 1u);
        const uint32_t n = (uint32_t)e.ref + (uint32_t)e.alt;
        unsigned long long k = ~0ull;
        uint16_t v = 0;
        if (n >= 1 && n <= (uint32_t)K_LIGHT) {
            const uint32_t cb = c / NB, cl = c % NB, t = e.locus / TILE_L, ll = e.locus % TILE_L;
            v = (uint16_t)(ll * (uint32_t)TAB_SLOTS + (uint32_t)tab_slot((int)e.alt, (int)n));
            k = (((((unsigned long long)cb * nT + t) << 12) | cl) << 3) | (unsigned long long)(v & 7u);
            ++mine;
        }
        keys[i] = k;
        vals[i] = v;
    }
    for (int off = 16; off > 0; off >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, off);
    if ((threadIdx.x & 31) == 0 && mine) atomicAdd(n_light, mine);
}

// stored entries per cell (the start value of loci_used_per_cell) when the run stream is not built
__global__ void k_row_len(const uint64_t* __restrict__ row_ptr, uint32_t n_cells, uint32_t* __restrict__ row_len) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_cells) row_len[i] = (uint32_t)(row_ptr[i + 1] - row_ptr[i]);
}

struct DropBank {  // run key of an entry key
    __device__ unsigned long long operator()(unsigned long long k) const { return k >> 3; }
};
struct IntToU64 {
    __device__ unsigned long long operator()(int v) const { return (unsigned long long)v; }
};

template <typename T>
__device__ __forceinline__ uint64_t lower_bound_dev(const T* __restrict__ a, uint64_t lo, uint64_t hi, T v) {
    while (lo < hi) {
        const uint64_t mid = (lo + hi) >> 1;
        if (a[mid] < v) lo = mid + 1;
        else hi = mid;
    }
    return lo;
}

// Sub-unit = the runs of one warp in one (cell block, tile) unit.  Sub-units are numbered warp-major,
// su = (cell block * W + warp) * nT + tile, so that the chunks a warp consumes over consecutive tiles of a work item
// are consecutive in the stream.
__device__ __forceinline__ unsigned long long sub_unit_of(unsigned long long unit, uint32_t w, uint32_t nT, uint32_t W) {
    const unsigned long long cb = unit / nT, t = unit % nT;
    return (cb * W + w) * nT + t;
}
// runs (unique keys + counts, from cub::DeviceRunLengthEncode) -> sort key ((sub-unit * 8 + class) << 16 | 0xFFFF -
// min(count, 0xFFFF)), class = q mod 8 = the 16-byte bank group of the run's accumulator slot: longest first inside a
// (sub-unit, class).  The order only decides how much padding a group carries, never the result.
__global__ void k_run_keys(const unsigned long long* __restrict__ ukeys, const int* __restrict__ counts, uint32_t n_runs,
                           uint32_t nT, uint32_t W, unsigned long long* __restrict__ key2, uint32_t* __restrict__ idx) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_runs) return;
    const uint32_t cnt = (uint32_t)counts[r];
    const unsigned long long unit = ukeys[r] >> 12;
    const uint32_t cl = (uint32_t)(ukeys[r] & 0xFFFull);
    const unsigned long long sc = sub_unit_of(unit, cl % W, nT, W) * 8ull + ((cl / W) & 7u);
    key2[r] = (sc << 16) | (unsigned long long)(0xFFFFu - (cnt > 0xFFFFu ? 0xFFFFu : cnt));
    idx[r] = r;
}
// first sorted run of every (sub-unit, class), i in [0, 8 * n_sub]
__global__ void k_sub_runs(const unsigned long long* __restrict__ key2_sorted, uint32_t n_runs, uint64_t n_sc,
                           uint32_t* __restrict__ sub_run) {
    const uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    if (i <= n_sc) sub_run[i] = (uint32_t)lower_bound_dev<unsigned long long>(key2_sorted, 0, n_runs, i << 16);
}

__device__ __forceinline__ uint32_t warp_max_u32(uint32_t v) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, off));
    return v;
}

// GROUPS of a sub-unit.  Lane l of a group sums a run of class v = l & 7 (accumulator slots of class v live in 16-byte
// bank group v): group g gives lane l = 8 j + v the (4 g + j)-th longest run of class v, so the 8 lanes of every quarter
// warp hold accumulator slots in 8 distinct bank groups and the group's accumulator update is conflict-free.  Classes
// are iid samples of the cells, so the runs of a group still have nearly equal lengths.
struct GroupWalk {
    uint32_t b, e;  // this lane's class: sorted runs [b, e)
    uint32_t G;     // groups of the sub-unit
    __device__ __forceinline__ GroupWalk(const uint32_t* __restrict__ sub_run, uint64_t su, uint32_t lane) {
        b = sub_run[su * 8 + (lane & 7u)];
        e = sub_run[su * 8 + (lane & 7u) + 1];
        G = warp_max_u32((e - b + 3u) / 4u);
    }
    // index into the sorted run list of this lane's run in group g, or 0xFFFFFFFF
    __device__ __forceinline__ uint32_t run_of(uint32_t g, uint32_t lane) const {
        const uint32_t r = b + 4u * g + (lane >> 3);
        return r < e ? r : 0xFFFFFFFFu;
    }
};

// one warp per sub-unit: its chunks.  steps = lookup steps (padding included) of the sub-unit.
__global__ void k_sub_count(const uint32_t* __restrict__ sub_run, const uint32_t* __restrict__ idx_sorted,
                            const int* __restrict__ counts, uint64_t n_sub, uint32_t* __restrict__ n_chunks,
                            unsigned long long* __restrict__ total_steps) {
    const uint64_t su = (blockIdx.x * (uint64_t)blockDim.x + threadIdx.x) >> 5;
    const uint32_t lane = threadIdx.x & 31;
    if (su > n_sub) return;
    if (su == n_sub) {
        if (lane == 0) n_chunks[su] = 0;
        return;
    }
    const GroupWalk gw(sub_run, su, lane);
    // per block of 8 groups: one cell chunk + the entry words of the groups rounded up to whole chunks
    uint64_t chunks = 0, steps = 0, block_words = 0;
    for (uint32_t g = 0; g < gw.G; ++g) {
        const uint32_t r = gw.run_of(g, lane);
        const uint32_t M = warp_max_u32(r != 0xFFFFFFFFu ? (uint32_t)counts[idx_sorted[r]] : 0u);
        block_words += (M + PIECE_WORDS - 1) / PIECE_WORDS * PIECE_WORDS;
        steps += M;
        if ((g + 1) % GROUPS_PER_CELL_CHUNK == 0 || g + 1 == gw.G) {
            chunks += 1 + (block_words + CHUNK_WORDS - 1) / CHUNK_WORDS;
            block_words = 0;
        }
    }
    if (gw.G == 0) chunks = 1;  // the cell chunk that announces the tile
    if (lane == 0) {
        n_chunks[su] = (uint32_t)chunks;
        if (steps) atomicAdd(total_steps, (unsigned long long)steps);
    }
}
// cost of every unit for the work-item partition: the chunks of its W sub-units
__global__ void k_unit_cost(const uint32_t* __restrict__ n_chunks, uint32_t n_units, uint32_t nT, uint32_t W,
                            unsigned long long* __restrict__ unit_cost) {
    const uint32_t u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= n_units) return;
    unsigned long long sum = 0;
    for (uint32_t w = 0; w < W; ++w) sum += n_chunks[sub_unit_of(u, w, nT, W)];
    unit_cost[u] = sum;
}

// one warp per sub-unit writes its chunks.  Lanes emit in lock step; a lane collects 8 words and stores one uint4.
struct WordSink {
    uint4* dst;  // this lane's piece of the current chunk
    unsigned long long lo, hi;
    uint32_t pos;
    __device__ __forceinline__ void put(uint32_t w16) {
        if (pos < 4) lo |= (unsigned long long)w16 << (16 * pos);
        else hi |= (unsigned long long)w16 << (16 * (pos - 4));
        if (++pos == CHUNK_WORDS) {
            *dst = make_uint4((uint32_t)lo, (uint32_t)(lo >> 32), (uint32_t)hi, (uint32_t)(hi >> 32));
            dst += 32;
            lo = hi = 0;
            pos = 0;
        }
    }
};

// STEP ORDER inside a group.  The entries of a run may be summed in any order, and a lookup costs one shared-memory
// wavefront per distinct address in the most loaded 16-byte bank group of its quarter warp: random order averages 2.4
// wavefronts per quarter-warp step.  The builder therefore picks, step by step and quarter by quarter, which of its
// remaining entries each lane takes: lanes that must take one now (entries left >= steps left) go first and take the
// entry whose bank group is least used in this step; the others take an entry only if its bank group is still free,
// else they sit the step out with a padding word (no lookup is issued for it).  ~1.6 wavefronts per step on
// 5-entry runs.  The run's entries arrive sorted by bank group (k_tile_keys), so "an entry of bank group b" is a
// cursor per bank group; cursors live in shared memory ([bank][lane]: conflict-free).
__global__ void __launch_bounds__(TPB) k_sub_fill(const uint32_t* __restrict__ sub_run, const uint32_t* __restrict__ chunk_off,
                                                  const uint32_t* __restrict__ idx_sorted, const unsigned long long* __restrict__ ukeys,
                                                  const int* __restrict__ counts, const unsigned long long* __restrict__ run_start,
                                                  const uint16_t* __restrict__ vals_sorted, uint64_t n_sub, uint32_t W,
                                                  uint4* __restrict__ stream) {
    __shared__ uint32_t s_left[TPB / 32][8][32];  // entries of bank group b this lane has not placed yet
    __shared__ uint32_t s_cur[TPB / 32][8][32];   // offset inside the run of the next one
    const uint64_t su = (blockIdx.x * (uint64_t)blockDim.x + threadIdx.x) >> 5;
    const uint32_t lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    if (su >= n_sub) return;
    const GroupWalk gw(sub_run, su, lane);
    WordSink sink;
    sink.dst = stream + (size_t)chunk_off[su] * 32 + lane;
    sink.lo = sink.hi = 0;
    sink.pos = 0;
    if (gw.G == 0) {
        sink.put(WC_CELLS | WC_NEWTILE | WC_IDLE);
        while (sink.pos != 0) sink.put(WC_IDLE);
    }
    const int qbase = (int)(lane & ~7u);
    for (uint32_t g = 0; g < gw.G; ++g) {
        if (g % GROUPS_PER_CELL_CHUNK == 0) {
            // close the previous block on a chunk boundary, then the cell chunk of groups g .. g + 7
            while (sink.pos != 0) sink.put(NULL_SLOT);  // a filler piece: one valid word that reads a zero slot, no WD_LAST
            for (uint32_t i = 0; i < GROUPS_PER_CELL_CHUNK; ++i) {
                const uint32_t ri = g + i < gw.G ? gw.run_of(g + i, lane) : 0xFFFFFFFFu;
                uint32_t cw = WC_IDLE;
                if (ri != 0xFFFFFFFFu) cw = ((uint32_t)(ukeys[idx_sorted[ri]] & 0xFFFull) / W) & WC_QMASK;  // slot = warp * Q + q
                sink.put(cw | (i == 0 ? (WC_CELLS | (g == 0 ? WC_NEWTILE : 0u)) : 0u));
            }
        }
        const uint32_t r = gw.run_of(g, lane);
        const bool has = r != 0xFFFFFFFFu;
        uint32_t cnt = 0;
        unsigned long long src = 0;
        if (has) {
            const uint32_t rid = idx_sorted[r];
            cnt = (uint32_t)counts[rid];
            src = run_start[rid];
        }
        const uint32_t M = warp_max_u32(cnt);
        // per bank group: how many entries, where they start (the run is sorted by bank group)
        uint32_t avail = 0;  // bit b: entries of bank group b left
#pragma unroll
        for (int b = 0; b < 8; ++b) s_left[wib][b][lane] = 0;
        for (uint32_t j = 0; j < cnt; ++j) s_left[wib][vals_sorted[src + j] & 7u][lane] += 1;
        {
            uint32_t off = 0;
#pragma unroll
            for (int b = 0; b < 8; ++b) {
                const uint32_t n = s_left[wib][b][lane];
                s_cur[wib][b][lane] = off;
                off += n;
                if (n) avail |= 1u << b;
            }
        }
        uint32_t rem = cnt;
        for (uint32_t st = 0; st < M; ++st) {
            const uint32_t left = M - st;
            const bool must = rem >= left;  // rem <= left always; rem > 0 here because left > 0
            uint32_t used = 0;              // 8 x 4-bit use counts of this quarter's bank groups in this step
            int pick = -1;
            // round 1: the lanes that must place an entry now, in lane order; round 2: the others, free bank groups only
#pragma unroll 1
            for (int round = 0; round < 2; ++round) {
#pragma unroll 1
                for (int i = 0; i < 8; ++i) {
                    int mine = -1;
                    if ((int)(lane & 7u) == i && rem > 0 && pick < 0 && (round == 0 ? must : !must)) {
                        uint32_t best_use = 16;
#pragma unroll
                        for (int b = 0; b < 8; ++b) {
                            const uint32_t u = (used >> (4 * b)) & 15u;
                            if (((avail >> b) & 1u) && u < best_use) {
                                best_use = u;
                                mine = b;
                            }
                        }
                        if (round == 1 && best_use != 0) mine = -1;
                        pick = mine;
                    }
                    const int pb = __shfl_sync(0xffffffffu, mine, qbase + i);
                    if (pb >= 0) used += 1u << (4 * pb);
                }
            }
            uint32_t w16;
            if (pick >= 0) {
                const uint32_t at = s_cur[wib][pick][lane];
                w16 = (uint32_t)vals_sorted[src + at];
                s_cur[wib][pick][lane] = at + 1;
                const uint32_t l2 = s_left[wib][pick][lane] - 1;
                s_left[wib][pick][lane] = l2;
                if (l2 == 0) avail &= ~(1u << pick);
                --rem;
            } else {  // sit the step out on the zero slot of a bank group nobody in this quarter uses now
                uint32_t f = 0;
                while (f < 7 && ((used >> (4 * f)) & 15u)) ++f;
                w16 = NULL_SLOT + f;
            }
            const uint32_t pos = st % PIECE_WORDS, piece0 = st - pos;
            if (pos == 0) w16 |= (min(PIECE_WORDS, M - piece0) - 1u) << WD_CNT_SHIFT;
            if (pos == 1 && piece0 + PIECE_WORDS >= M) w16 |= WD_LAST;
            sink.put(w16);
        }
        // complete the group's last piece (word 1 of a one-word piece still carries WD_LAST)
        for (uint32_t st = M; st % PIECE_WORDS != 0; ++st) sink.put(NULL_SLOT | (st % PIECE_WORDS == 1 ? WD_LAST : 0u));
    }
    while (sink.pos != 0) sink.put(NULL_SLOT);
}

// per item and warp: the warp's chunk range over its tiles of the item (one consecutive range, see sub_unit_of) and how
// many of the item's leading tiles it skips.  An item is a range of (tile, warp) sub-units of one cell block in tile-major
// order: it starts at warp w_a of tile t_a and ends before warp w_e of tile t_e, so warp w owns the tiles
// [t_a + (w < w_a), t_e + (w < w_e)).
__global__ void k_item_rng(const uint4* __restrict__ item_q, const uint4* __restrict__ item_x, uint32_t n_items,
                           const uint32_t* __restrict__ chunk_off, uint32_t nT, uint32_t W, uint4* __restrict__ item_rng) {
    const uint32_t q = blockIdx.x, w = threadIdx.x;
    if (q >= n_items || w >= W) return;
    const uint4 it = item_q[q], x = item_x[q];
    const uint32_t first = it.y + (w < x.x ? 1u : 0u), end = max(first, x.y + (w < x.z ? 1u : 0u));
    const size_t su0 = ((size_t)it.x * W + w) * nT;
    item_rng[(size_t)q * W + w] = make_uint4(chunk_off[su0 + first], chunk_off[su0 + end], first - it.y, 0u);
}

}  // namespace

// ---------------------------------------------------------------------------------------------- kernel
namespace {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// TMA bulk copy global -> shared (1-D, 16-byte granules), completion counted on an mbarrier
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile(
            "{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
            : "=r"(ok)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
    } while (!ok);
}
// streaming 128-bit load of the run stream: read once, keep it out of L1
__device__ __forceinline__ uint4 ld_stream(const uint4* p) {
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p));
    return v;
}

constexpr size_t TB_BYTES = (size_t)TILE_SLOTS * 16;                              // bytes one table bulk copy moves
constexpr size_t TB_BUF = (((size_t)(TILE_SLOTS + ZERO_SLOTS) * 16) + 127) / 128 * 128;  // + the all-zero slots
constexpr size_t TL_ACC_BYTES = (size_t)NB_MAX * 16;
constexpr size_t TL_SMEM = 2 * TB_BUF + TL_ACC_BYTES + 64;
static_assert(TL_SMEM <= 227 * 1024, "shared memory budget of the scoring kernel");

// Table buffers: tile g of a CTA (counted over all its items) lives in buffer g & 1; tab_full[b] is the mbarrier its
// bulk copy completes on (parity (g >> 1) & 1), tab_done[b] counts the warps that have finished with the buffer.
// Handing a buffer over needs no CTA barrier: the last warp to leave tile k starts the copy of tile k + 2 into it
// (cells never change warps, so accumulators need no barrier either, and warps drift by at most two tiles).
// Both are out of line: they run once per warp and tile, and inlined at each of the 24 word positions of the
// unrolled loop they made the kernel 50 KB of code.
__device__ __noinline__ void leave_tile(unsigned char* smem, uint32_t k, uint32_t g_tile, uint32_t t0, uint32_t ntile, uint32_t nw,
                                        const double* tab16) {
    uint64_t* tab_full = reinterpret_cast<uint64_t*>(smem + 2 * TB_BUF + TL_ACC_BYTES);
    unsigned int* tab_done = reinterpret_cast<unsigned int*>(tab_full + 2);
    const uint32_t g = g_tile + k;
    __syncwarp();
    if ((threadIdx.x & 31) == 0) {
        const unsigned int done = atomicAdd(&tab_done[g & 1], 1u);
        if (done == nw - 1) {
            tab_done[g & 1] = 0;
            if (k + 2 < ntile) {
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                mbar_expect_tx(&tab_full[g & 1], (uint32_t)TB_BYTES);
                bulk_g2s(smem + (size_t)(g & 1) * TB_BUF, tab16 + (size_t)(t0 + k + 2) * TILE_SLOTS * 2, (uint32_t)TB_BYTES, &tab_full[g & 1]);
            }
        }
    }
}
// leaves tile k (if any), waits for tile k + 1; returns the byte offset of its buffer in shared memory
__device__ __noinline__ uint32_t enter_tile(unsigned char* smem, int k, uint32_t g_tile, uint32_t t0, uint32_t ntile, uint32_t nw,
                                            const double* tab16) {
    if (k >= 0) leave_tile(smem, (uint32_t)k, g_tile, t0, ntile, nw, tab16);
    uint64_t* tab_full = reinterpret_cast<uint64_t*>(smem + 2 * TB_BUF + TL_ACC_BYTES);
    const uint32_t g = g_tile + (uint32_t)(k + 1);
    mbar_wait(&tab_full[g & 1], (g >> 1) & 1u);
    return (g & 1) * (uint32_t)TB_BUF;
}

// Work queue: CTA b takes queue slot b first, later slots come from an atomic cursor that starts at gridDim.x; the
// cursor for the NEXT item is fetched while the current one is being scored, so no item waits for the atomic.
__global__ void __launch_bounds__(W_MAX * 32, 1)
    k_score_runs(const uint4* __restrict__ stream, const uint4* __restrict__ item_q, const uint4* __restrict__ item_rng,
                 const uint32_t* __restrict__ cta_first, const double* __restrict__ tab16, uint32_t NB,
                 double* __restrict__ part) {
    extern __shared__ __align__(128) unsigned char smem[];
    double2* acc = reinterpret_cast<double2*>(smem + 2 * TB_BUF);
    uint64_t* tab_full = reinterpret_cast<uint64_t*>(smem + 2 * TB_BUF + TL_ACC_BYTES);
    unsigned int* tab_done = reinterpret_cast<unsigned int*>(tab_full + 2);  // [2] warps done with a buffer
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const uint32_t nw = blockDim.x >> 5;
    const uint32_t Q = NB / nw;  // cells per warp (NB is a multiple of 8 * nw)
    if (tid == 0) {
        mbar_init(&tab_full[0], 1);
        mbar_init(&tab_full[1], 1);
        tab_done[0] = 0;
        tab_done[1] = 0;
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid < 2 * ZERO_SLOTS)  // the all-zero slots behind each table buffer (padding words point at them)
        reinterpret_cast<double2*>(smem + (size_t)(tid / ZERO_SLOTS) * TB_BUF)[TILE_SLOTS + tid % ZERO_SLOTS] = make_double2(0.0, 0.0);
    for (uint32_t i = tid; i < NB; i += blockDim.x) acc[i] = make_double2(0.0, 0.0);
    __syncthreads();
    uint32_t g_tile = 0;  // tiles consumed by this CTA so far (buffer = g & 1, parity = (g >> 1) & 1)
    double2* const my_acc = acc + (size_t)w * Q;
    const uint32_t q_end = cta_first[blockIdx.x + 1];
    for (uint32_t qk = cta_first[blockIdx.x]; qk < q_end; ++qk) {
        const uint4 it = item_q[qk];
        const uint4 rng = item_rng[(size_t)qk * nw + w];
        const uint32_t t0 = it.y, ntile = it.z, item = it.w;
        const uint4* p = stream + (size_t)rng.x * 32 + lane;
        const uint32_t n = rng.y - rng.x;
        // three chunks in flight per lane: registers, no staging ring (loads past the end re-read the last chunk)
        const uint32_t nl = n ? n - 1 : 0;
        uint4 pf[STREAM_PF];
#pragma unroll
        for (uint32_t j = 0; j < STREAM_PF; ++j) pf[j] = ld_stream(p + (size_t)min(j, nl) * 32);
        if (tid == 0) {
            for (uint32_t k = 0; k < 2 && k < ntile; ++k) {
                const uint32_t g = g_tile + k;
                mbar_expect_tx(&tab_full[g & 1], (uint32_t)TB_BYTES);
                bulk_g2s(smem + (size_t)(g & 1) * TB_BUF, tab16 + (size_t)(t0 + k) * TILE_SLOTS * 2, (uint32_t)TB_BYTES, &tab_full[g & 1]);
            }
        }
        int k = -1;                    // item-local tile this warp is in
        const double2* tab = nullptr;  // its staged slots
        if (rng.z) {                   // the item's first tile belongs to the previous piece for this warp: pass through it
            enter_tile(smem, k, g_tile, t0, ntile, nw, tab16);
            ++k;
        }
        double sT = 0.0, sE = 0.0;     // the lane's run
        uint4 cq = make_uint4(0u, 0u, 0u, 0u);  // accumulator indices of the lane's cells in the coming groups, 16 bits each
        // One chunk = 8 words per lane.  A cell chunk refills cq (and may announce the next tile).  In an entry chunk
        // all 8 lookups go out back to back (predicated off for padding words: eight independent LDS.128 in flight
        // per warp keep the shared-memory pipe full), then the words are consumed in order; the last step of a group
        // (warp-uniform) adds the lane's run to its cell's accumulator and takes the next cell off cq.
        // One piece = 4 words per lane, 1..4 of them valid (warp-uniform count): the lookups go out back to back, then
        // the sums; if the group ends here (warp-uniform flag) the lane's run is added to its cell's accumulator and
        // the next cell comes off cq.
        auto piece = [&](const uint32_t lo, const uint32_t hi) {
            const uint32_t cnt = (lo >> WD_CNT_SHIFT) & 3u;
            const double2 t0 = tab[lo & WD_SLOT];
            double2 t1 = make_double2(0.0, 0.0), t2 = t1, t3 = t1;
            if (cnt >= 1) t1 = tab[(lo >> 16) & WD_SLOT];
            if (cnt >= 2) t2 = tab[hi & WD_SLOT];
            if (cnt >= 3) t3 = tab[(hi >> 16) & WD_SLOT];
            sT += (t0.x + t1.x) + (t2.x + t3.x);
            sE += (t0.y + t1.y) + (t2.y + t3.y);
            if (lo & (WD_LAST << 16)) {
                const uint32_t cell = cq.x;
                cq.x = __funnelshift_r(cq.x, cq.y, 16);
                cq.y = __funnelshift_r(cq.y, cq.z, 16);
                cq.z = __funnelshift_r(cq.z, cq.w, 16);
                cq.w >>= 16;
                if (!(cell & WC_IDLE)) {
                    double2* cell_acc = my_acc + (cell & WC_QMASK);
                    double2 a = *cell_acc;
                    a.x += sT;
                    a.y += sE;
                    *cell_acc = a;
                }
                sT = 0.0;
                sE = 0.0;
            }
        };
        auto chunk = [&](const uint4 cw) {
            if (cw.x & WC_CELLS) {
                if (cw.x & WC_NEWTILE) {
                    tab = reinterpret_cast<const double2*>(smem + enter_tile(smem, k, g_tile, t0, ntile, nw, tab16));
                    ++k;
                }
                cq = cw;
                return;
            }
            piece(cw.x, cw.y);
            piece(cw.z, cw.w);
        };
        // three chunks per trip, each register set reloaded right after its chunk is consumed: moving an in-flight
        // load between registers would wait for it (measured: long-scoreboard stalls dominated the rotated form)
        uint32_t c = 0;
#pragma unroll 1
        for (; c + STREAM_PF <= n; c += STREAM_PF) {
#pragma unroll
            for (uint32_t j = 0; j < STREAM_PF; ++j) {
                const uint4 cur = pf[j];
                pf[j] = ld_stream(p + (size_t)min(c + STREAM_PF + j, nl) * 32);
                chunk(cur);
            }
        }
#pragma unroll
        for (uint32_t j = 0; j + 1 < STREAM_PF; ++j)
            if (c + j < n) chunk(pf[j]);
        // every warp walks every tile of the item (waits for its table, leaves it); the stream carries one NEWTILE
        // header per tile, so the loop body only runs for an item without chunks
        while (k + 1 < (int)ntile) {
            enter_tile(smem, k, g_tile, t0, ntile, nw, tab16);
            ++k;
        }
        leave_tile(smem, (uint32_t)k, g_tile, t0, ntile, nw, tab16);
        g_tile += ntile;
        __syncthreads();  // every warp has added its last run
        // partials out (cell_local = q * nw + warp lives in accumulator slot warp * Q + q), accumulators cleared
        double2* dst = reinterpret_cast<double2*>(part) + (size_t)item * NB;
        for (uint32_t i = tid; i < NB; i += blockDim.x) {
            const uint32_t slot = (i % nw) * Q + i / nw;
            dst[i] = acc[slot];
            acc[slot] = make_double2(0.0, 0.0);
        }
        __syncthreads();
    }
}

// Adds the partials of a cell's block in item order (FIN_LANES lanes per cell, each a fixed subsequence, combined
// by a fixed shuffle tree: deterministic), appends the heavy values in row order, and (EM pass) normalises by
// loci_used_per_cell (main.rs:313-323).
//   SCORE_EM       out0 = log_likelihood, out1 = expected_log_likelihood, + n_loci, normalised value
//   SCORE_POST_AB  out0 = minority, out1 = majority log likelihood (heavy values hv0 / hv1)
//   SCORE_POST_C   out0 = doublet log likelihood
constexpr int FIN_LANES = 8;
template <int MODE>
__global__ void __launch_bounds__(256) k_score_finish(const double* __restrict__ part, const uint32_t* __restrict__ cb_item,
                                                      uint32_t NB, uint32_t n_cells, const uint32_t* __restrict__ h_row_ptr,
                                                      const double* __restrict__ hv0, const double* __restrict__ hv1,
                                                      const uint32_t* __restrict__ n_used, double* __restrict__ out0,
                                                      double* __restrict__ out1, uint32_t* __restrict__ nloci,
                                                      double* __restrict__ norm, unsigned int* __restrict__ nan_flag) {
    // lanes 0..3 of a warp take cells c0..c0+3 for sub 0, lanes 4..7 the same cells for sub 1, ...: partial loads of one
    // sub are 4 consecutive double2 (64 B)
    const uint32_t warp_cell0 = ((blockIdx.x * blockDim.x + threadIdx.x) >> 5) * (32 / FIN_LANES);
    const uint32_t lane = threadIdx.x & 31, sub = lane / (32 / FIN_LANES);
    const uint32_t i = warp_cell0 + lane % (32 / FIN_LANES);
    const bool live = i < n_cells;
    const uint32_t ii = live ? i : 0;
    const uint32_t cb = ii / NB, cl = ii % NB;
    const uint32_t it0 = cb_item[cb], it1 = cb_item[cb + 1];
    const double2* p2 = reinterpret_cast<const double2*>(part);
    double s0 = 0.0, s1 = 0.0;
    for (uint32_t it = it0 + sub; it < it1; it += FIN_LANES) {
        const double2 p = __ldg(p2 + (size_t)it * NB + cl);
        s0 += p.x;
        s1 += p.y;
    }
    if (live) {  // the cell's heavy values, row order, strided over the cell's lanes (a skewed matrix has cells with
                 // thousands of them: one lane alone walking the list was the whole pass)
        const uint32_t hb = h_row_ptr[i], he = h_row_ptr[i + 1];
        for (uint32_t h = hb + sub; h < he; h += FIN_LANES) {
            s0 += hv0[h];
            if (MODE != SCORE_POST_C) s1 += hv1[h];
        }
    }
#pragma unroll
    for (int off = 32 / FIN_LANES; off < 32; off <<= 1) {
        s0 += __shfl_xor_sync(0xffffffffu, s0, off);
        if (MODE != SCORE_POST_C) s1 += __shfl_xor_sync(0xffffffffu, s1, off);
    }
    if (!live || sub != 0) return;
    out0[i] = s0;
    if (MODE != SCORE_POST_C) out1[i] = s1;
    if (MODE == SCORE_EM) {
        const uint32_t cnt = n_used[i];
        nloci[i] = cnt;
        const double nv = cnt > 0 ? s0 / (double)cnt : 0.0;
        norm[i] = nv;
        if (nv != nv) atomicOr(nan_flag, 1u);
    }
}

}  // namespace

// One scoring pass over the run stream.  SCORE_EM reads table pass 0 ({T, E} slots); the posterior pass reads
// {T_minority, T_majority} (pass 0 of tab16 as rebuilt by k_tables<1>) and {T_doublet, 0} (pass 1).
int score_tiled(clx_ctx* c, ScoreMode mode) {
    cudaStream_t s = c->stream;
    const uint32_t N = c->n_cells;
    if (N == 0) return CLX_OK;
    if (!c->tl_attr_set) {  // per context: the attribute is per device
        CLX_CUDA(c, cudaFuncSetAttribute(k_score_runs, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TL_SMEM));
        c->tl_attr_set = true;
    }
    const double* tab = c->tab16.p + (mode == SCORE_POST_C ? (size_t)c->tl_nT * TILE_SLOTS * 2 : 0);
    if (c->tl_n_items) {
        k_score_runs<<<c->tl_grid, c->tl_W * 32, TL_SMEM, s>>>(c->tl_stream.p, c->tl_item_q.p, c->tl_item_rng.p, c->tl_cta_first.p, tab,
                                                              c->tl_NB, c->tl_part.p);
        c->launches++;
        CLX_DEBUG_SYNC(c, "k_score_runs");
    }
    const size_t nh = c->n_heavy;
    const unsigned fgrid = (unsigned)(((uint64_t)N * FIN_LANES + 255) / 256);
    if (mode == SCORE_EM)
        k_score_finish<SCORE_EM><<<fgrid, 256, 0, s>>>(c->tl_part.p, c->tl_cb_item.p, c->tl_NB, N, c->h_row_ptr.p, c->hv_ll.p, c->hv_e.p,
                                                      c->n_used.p, c->ll.p, c->ell.p, c->nloci.p, c->norm.p, &c->d_stats.p->nan_flag);
    else if (mode == SCORE_POST_AB)
        k_score_finish<SCORE_POST_AB><<<fgrid, 256, 0, s>>>(c->tl_part.p, c->tl_cb_item.p, c->tl_NB, N, c->h_row_ptr.p, c->hv_ll.p,
                                                           c->hv_ll.p + nh, nullptr, c->post_ll.p, c->post_ll.p + N, nullptr, nullptr,
                                                           nullptr);
    else
        k_score_finish<SCORE_POST_C><<<fgrid, 256, 0, s>>>(c->tl_part.p, c->tl_cb_item.p, c->tl_NB, N, c->h_row_ptr.p, c->hv_ll.p + 2 * nh,
                                                          nullptr, nullptr, c->post_ll.p + 2 * (size_t)N, nullptr, nullptr, nullptr, nullptr);
    c->launches++;
    return CLX_OK;
}

int build_tiled(clx_ctx* c, const uint32_t* d_cell_of_entry) {
    NvtxRange nvtx("clx:build_run_stream");
    cudaStream_t s = c->stream;
    StageClock clk(s, "  run stream");
    c->tiled_ready = false;
    const uint32_t N = c->n_cells, L = c->n_loci;
    const uint64_t nnz = c->nnz;
    CLX_CUDA(c, c->row_len.alloc((size_t)N + 1));
    CLX_CUDA(c, c->n_used.alloc((size_t)N + 1));
    CLX_CUDA(c, cudaMemsetAsync(c->row_len.p, 0, c->row_len.bytes(), s));
    // cell blocks: few and large (every block re-reads the table tiles, and a sub-unit should pool >= 256 runs so that
    // the runs of a group have equal lengths) but enough (cell block, tile) units to fill the SMs
    c->tl_nT = L ? (L + TILE_L - 1) / TILE_L : 1;
    {
        uint32_t W;
        const char* env_w = getenv("CLX_TILED_W");
        uint32_t pc = N ? (N + NB_MAX - 1) / NB_MAX : 1;
        // about 1.5 (cell block, tile) units per SM, so that the one-item-per-SM split has something to balance; more cell
        // blocks than that only re-read the table tiles (C2: 4 blocks of 3072 cells 25.9 us, 5 of 2048 31.9 us)
        const uint32_t want = (3u * (uint32_t)c->sm_count + 2u * c->tl_nT - 1) / (2u * c->tl_nT);
        const uint32_t pc_cap = N ? std::max<uint32_t>(1u, N / 512u) : 1u;
        pc = std::max(pc, std::min(want, pc_cap));
        const char* env = getenv("CLX_TILED_PC");
        if (env && *env) pc = std::max<uint32_t>(pc, (uint32_t)atoi(env));
        uint32_t nb = N ? (N + pc - 1) / pc : 1;
        const char* env_nb = getenv("CLX_TILED_NB");  // experiments: force the cell-block size
        if (env_nb && *env_nb) nb = (uint32_t)atoi(env_nb);
        // warps: as many as keep >= 64 cells (8 per accumulator class) in every warp's pool; 8 warps per scheduler hide
        // the shared-memory latency that 4 do not (C3s: W=32 77 us, W=16 83-89 us)
        W = nb > 1536 ? 32u : (nb > 768 ? 16u : 8u);
        if (env_w && *env_w) W = std::min<uint32_t>(W_MAX, std::max<uint32_t>(1u, (uint32_t)atoi(env_w)));
        // Q = NB / W is a multiple of 32: the 8 accumulator classes of a warp's pool hold whole groups of 4 runs
        const uint32_t gran = 32u * W;
        nb = std::min<uint32_t>(NB_MAX, std::max<uint32_t>(gran, (nb + gran - 1) / gran * gran));
        c->tl_W = W;
        c->tl_NB = nb;
        c->tl_Pc = N ? (N + nb - 1) / nb : 1;
    }
    const uint32_t W = c->tl_W;
    c->tl_n_units = c->tl_Pc * c->tl_nT;
    {
        // Very sparse matrices: every (cell block, tile) unit stages a 72 KB table tile, an entry gathered straight from
        // the per-locus records costs ~72 B (entry + two sectors).  Below one entry per cell and tile the runs degenerate
        // to single entries, the 32-lane groups are mostly padding (C5: 10 lookup steps per entry) and the gather kernel
        // (k_score_rows) moves fewer bytes; the run stream is then not built at all.  Small matrices (tables that stay in
        // L2 anyway) always take the stream.  C5 (density 5e-4): 448 entries per unit, 8.0 ms per pass as a stream.
        const uint64_t units = (uint64_t)c->tl_Pc * c->tl_nT;
        bool gather = units * (uint64_t)TB_BYTES > (16ull << 20) && nnz < 4096ull * units;
        const char* path = getenv("CLX_SCORE_PATH");  // "stream" / "gather": force one (tests, experiments)
        if (path && !strcmp(path, "gather")) gather = true;
        if (path && !strcmp(path, "stream")) gather = false;
        if (gather) {
            if (N) k_row_len<<<(N + TPB - 1) / TPB, TPB, 0, s>>>(c->row_ptr.p, N, c->row_len.p);
            c->launches++;
            c->tl_n_items = c->tl_n_units = c->tl_grid = 0;
            c->tl_n_light = c->tl_stream_chunks = c->tl_stream_steps = 0;
            c->tl_stream.release();
            c->tl_part.release();
            c->tab16.release();
            if (getenv("CLX_TILED_STATS"))
                fprintf(stderr, "[clx] gather scoring: N %u L %u, %.0f entries per (cell block, tile) unit\n", N, L,
                        units ? (double)nnz / (double)units : 0.0);
            CLX_CUDA(c, cudaGetLastError());
            return CLX_OK;
        }
    }
    if ((uint64_t)c->tl_Pc * c->tl_nT >= (1ull << 31)) return fail(c, CLX_ERR_ARG, "too many (cell block, tile) units");
    const uint32_t n_units = c->tl_n_units;

    DevBuf<unsigned long long> keys, keys_sorted, d_nlight, ukeys, key2, key2_sorted, run_start, unit_cost, d_steps;
    DevBuf<uint16_t> vals, vals_sorted;
    DevBuf<int> counts, d_nruns;
    DevBuf<uint32_t> ridx, ridx_sorted;
    DevBuf<uint8_t> tmp;
    size_t tb = 0;
    CLX_CUDA(c, keys.alloc(nnz + 1));
    CLX_CUDA(c, keys_sorted.alloc(nnz + 1));
    CLX_CUDA(c, vals.alloc(nnz + 1));
    CLX_CUDA(c, vals_sorted.alloc(nnz + 1));
    CLX_CUDA(c, d_nlight.alloc(1));
    CLX_CUDA(c, d_nruns.alloc(1));
    CLX_CUDA(c, d_steps.alloc(1));
    CLX_CUDA(c, cudaMemsetAsync(d_nlight.p, 0, 8, s));
    CLX_CUDA(c, cudaMemsetAsync(d_nruns.p, 0, 4, s));
    CLX_CUDA(c, cudaMemsetAsync(d_steps.p, 0, 8, s));
    unsigned long long n_light = 0;
    if (nnz) {
        k_tile_keys<<<grid_for(nnz), TPB, 0, s>>>(c->csr.p, d_cell_of_entry, nnz, c->tl_NB, c->tl_nT, keys.p, vals.p, d_nlight.p,
                                                  c->row_len.p);
        int bits = 16;
        while (bits < 63 && (1ull << bits) <= (((unsigned long long)n_units) << 15)) ++bits;
        // non-light keys are ~0: sorting on the low `bits + 1` bits still puts them last
        CLX_CUDA(c, cub::DeviceRadixSort::SortPairs(nullptr, tb, keys.p, keys_sorted.p, vals.p, vals_sorted.p, (int)nnz, 0, bits + 1, s));
        CLX_CUDA(c, tmp.alloc(tb));
        CLX_CUDA(c, cub::DeviceRadixSort::SortPairs(tmp.p, tb, keys.p, keys_sorted.p, vals.p, vals_sorted.p, (int)nnz, 0, bits + 1, s));
        c->launches += 5;
        CLX_CUDA(c, cudaMemcpyAsync(&n_light, d_nlight.p, 8, cudaMemcpyDeviceToHost, s));
        CLX_CUDA(c, cudaStreamSynchronize(s));
    }
    c->tl_n_light = n_light;
    clk.mark("entry keys + sort");
    keys.release();
    vals.release();
    // ---- runs
    int n_runs = 0;
    CLX_CUDA(c, ukeys.alloc(n_light + 1));
    CLX_CUDA(c, counts.alloc(n_light + 1));
    if (n_light) {
        cub::TransformInputIterator<unsigned long long, DropBank, const unsigned long long*> run_key(keys_sorted.p, DropBank());
        CLX_CUDA(c, cub::DeviceRunLengthEncode::Encode(nullptr, tb, run_key, ukeys.p, counts.p, d_nruns.p, (int)n_light, s));
        CLX_CUDA(c, tmp.alloc(tb));
        CLX_CUDA(c, cub::DeviceRunLengthEncode::Encode(tmp.p, tb, run_key, ukeys.p, counts.p, d_nruns.p, (int)n_light, s));
        CLX_CUDA(c, cudaMemcpyAsync(&n_runs, d_nruns.p, 4, cudaMemcpyDeviceToHost, s));
        CLX_CUDA(c, cudaStreamSynchronize(s));
        c->launches += 2;
    }
    keys_sorted.release();
    const uint32_t R = (uint32_t)n_runs;
    clk.mark("runs (RLE)");
    const uint64_t n_sub = (uint64_t)n_units * W;
    if (n_sub >= (1ull << 28) - 2) return fail(c, CLX_ERR_ARG, "too many (cell block, tile, warp) sub-units");
    DevBuf<uint32_t> sub_run, n_chunks, chunk_off;
    CLX_CUDA(c, run_start.alloc((size_t)R + 1));
    CLX_CUDA(c, key2.alloc((size_t)R + 1));
    CLX_CUDA(c, key2_sorted.alloc((size_t)R + 1));
    CLX_CUDA(c, ridx.alloc((size_t)R + 1));
    CLX_CUDA(c, ridx_sorted.alloc((size_t)R + 1));
    CLX_CUDA(c, sub_run.alloc(8 * n_sub + 2));
    CLX_CUDA(c, n_chunks.alloc(n_sub + 2));
    CLX_CUDA(c, chunk_off.alloc(n_sub + 2));
    CLX_CUDA(c, unit_cost.alloc((size_t)n_units + 2));
    if (R) {
        cub::TransformInputIterator<unsigned long long, IntToU64, const int*> cnt64(counts.p, IntToU64());
        CLX_CUDA(c, cub::DeviceScan::ExclusiveSum(nullptr, tb, cnt64, run_start.p, (int)R, s));
        CLX_CUDA(c, tmp.alloc(tb));
        CLX_CUDA(c, cub::DeviceScan::ExclusiveSum(tmp.p, tb, cnt64, run_start.p, (int)R, s));
        k_run_keys<<<(R + TPB - 1) / TPB, TPB, 0, s>>>(ukeys.p, counts.p, R, c->tl_nT, W, key2.p, ridx.p);
        int bits2 = 20;
        while (bits2 < 63 && (1ull << bits2) <= (n_sub << 19)) ++bits2;
        CLX_CUDA(c, cub::DeviceRadixSort::SortPairs(nullptr, tb, key2.p, key2_sorted.p, ridx.p, ridx_sorted.p, (int)R, 0, bits2, s));
        CLX_CUDA(c, tmp.alloc(tb));
        CLX_CUDA(c, cub::DeviceRadixSort::SortPairs(tmp.p, tb, key2.p, key2_sorted.p, ridx.p, ridx_sorted.p, (int)R, 0, bits2, s));
        c->launches += 7;
    }
    k_sub_runs<<<(unsigned)((8 * n_sub + 1 + TPB - 1) / TPB), TPB, 0, s>>>(key2_sorted.p, R, 8 * n_sub, sub_run.p);
    k_sub_count<<<(unsigned)(((n_sub + 1) * 32 + TPB - 1) / TPB), TPB, 0, s>>>(sub_run.p, ridx_sorted.p, counts.p, n_sub, n_chunks.p,
                                                                               d_steps.p);
    k_unit_cost<<<(n_units + TPB - 1) / TPB, TPB, 0, s>>>(n_chunks.p, n_units, c->tl_nT, W, unit_cost.p);
    CLX_CUDA(c, cub::DeviceScan::ExclusiveSum(nullptr, tb, n_chunks.p, chunk_off.p, (int)(n_sub + 1), s));
    CLX_CUDA(c, tmp.alloc(tb));
    CLX_CUDA(c, cub::DeviceScan::ExclusiveSum(tmp.p, tb, n_chunks.p, chunk_off.p, (int)(n_sub + 1), s));
    c->launches += 5;
    uint32_t total_chunks = 0;
    unsigned long long h_steps = 0;
    std::vector<unsigned long long> h_unit((size_t)n_units + 1);
    CLX_CUDA(c, cudaMemcpyAsync(&total_chunks, chunk_off.p + n_sub, 4, cudaMemcpyDeviceToHost, s));
    CLX_CUDA(c, cudaMemcpyAsync(&h_steps, d_steps.p, 8, cudaMemcpyDeviceToHost, s));
    CLX_CUDA(c, cudaMemcpyAsync(h_unit.data(), unit_cost.p, (size_t)n_units * 8, cudaMemcpyDeviceToHost, s));
    CLX_CUDA(c, cudaStreamSynchronize(s));
    {
        unsigned long long sum = 0;
        for (uint32_t u = 0; u < n_units; ++u) sum += h_unit[u];
        if (sum >= (1ull << 32)) return fail(c, CLX_ERR_ARG, "run stream longer than 2^32 chunks (2 TB): shard the cells over more ranks");
    }
    c->tl_stream_chunks = total_chunks;
    clk.mark("run sort + chunk counts");
    c->tl_stream_steps = h_steps * 32ull;
    if (getenv("CLX_TILED_STATS"))
        fprintf(stderr, "[clx] run stream: N %u L %u NB %u W %u Pc %u nT %u | light entries %llu runs %u | chunks %u (%.2f B/entry) lookup steps %llu (%.3f per entry)\n",
                N, L, c->tl_NB, W, c->tl_Pc, c->tl_nT, (unsigned long long)n_light, R, total_chunks,
                n_light ? 512.0 * total_chunks / (double)n_light : 0.0, (unsigned long long)c->tl_stream_steps,
                n_light ? (double)c->tl_stream_steps / (double)n_light : 0.0);
    CLX_CUDA(c, c->tl_stream.alloc((size_t)total_chunks * 32 + 32));
    if (total_chunks) {
        k_sub_fill<<<(unsigned)((n_sub * 32 + TPB - 1) / TPB), TPB, 0, s>>>(sub_run.p, chunk_off.p, ridx_sorted.p, ukeys.p, counts.p,
                                                                          run_start.p, vals_sorted.p, n_sub, W, c->tl_stream.p);
        c->launches++;
    }

    clk.mark("stream fill");
    // ---- work items: the (cell block, tile, warp) sub-units in that order are one sequence of costs (a unit's chunks
    //      spread evenly over its warps: cells are dealt to warps round-robin); it is cut into one contiguous piece of equal
    //      cost per SM, and a piece that crosses cell-block boundaries becomes one item per block (same CTA, one after the
    //      other).  Static: no queue, every launch does the same work on the same CTA.  A piece boundary inside a tile
    //      gives the tile's first warps to one CTA and the rest to the next (both stage the tile).
    {
        std::vector<uint4> queue, extra;  // {cell block, first tile, n tiles, item id} / {w_a, t_e, w_e, 0}
        std::vector<uint32_t> cb_item(c->tl_Pc + 1, 0), cta_first;
        unsigned long long cost_total = 0;
        for (uint32_t u = 0; u < n_units; ++u) cost_total += h_unit[u];
        const uint32_t G = (uint32_t)std::max(1, c->sm_count);
        const uint64_t S = (uint64_t)n_units * W, per_block = (uint64_t)c->tl_nT * W;
        std::vector<uint64_t> cut(G + 1, S);
        cut[0] = 0;
        {
            unsigned long long run = 0;
            uint32_t u = 0;
            for (uint32_t g = 1; g < G; ++g) {
                const long double target = (long double)cost_total * g / G;
                while (u < n_units && (long double)(run + h_unit[u]) <= target) run += h_unit[u++];
                if (u >= n_units) break;  // the remaining cuts stay at S
                const long double f = h_unit[u] ? (target - (long double)run) / (long double)h_unit[u] : 0.0L;
                cut[g] = (uint64_t)u * W + std::min<uint64_t>(W, (uint64_t)(f * W + 0.5L));
            }
            for (uint32_t g = 1; g <= G; ++g) cut[g] = std::max(cut[g], cut[g - 1]);
        }
        cta_first.assign(G + 1, 0);
        for (uint32_t g = 0; g < G; ++g) {
            cta_first[g] = (uint32_t)queue.size();
            uint64_t sa = cut[g];
            const uint64_t sb = cut[g + 1];
            while (sa < sb) {
                const uint32_t cb = (uint32_t)(sa / per_block);
                const uint64_t base = (uint64_t)cb * per_block, se = std::min(sb, base + per_block);
                const uint32_t t_a = (uint32_t)((sa - base) / W), w_a = (uint32_t)((sa - base) % W);
                const uint32_t t_e = (uint32_t)((se - base) / W), w_e = (uint32_t)((se - base) % W);
                const uint32_t t_hi = t_e + (w_e ? 1u : 0u);
                queue.push_back(make_uint4(cb, t_a, t_hi - t_a, (uint32_t)queue.size()));
                extra.push_back(make_uint4(w_a, t_e, w_e, 0u));
                sa = se;
            }
        }
        cta_first[G] = (uint32_t)queue.size();
        c->tl_n_items = (uint32_t)queue.size();
        {   // first item of every cell block (items are in block-major order)
            uint32_t q = 0;
            for (uint32_t cb = 0; cb <= c->tl_Pc; ++cb) {
                while (q < c->tl_n_items && queue[q].x < cb) ++q;
                cb_item[cb] = q;
            }
        }
        c->tl_grid = c->tl_n_items ? G : 0;
        DevBuf<uint4> item_x;
        CLX_CUDA(c, c->tl_item_q.alloc((size_t)c->tl_n_items + 1));
        CLX_CUDA(c, item_x.alloc((size_t)c->tl_n_items + 1));
        CLX_CUDA(c, c->tl_item_rng.alloc((size_t)c->tl_n_items * W + 1));
        CLX_CUDA(c, c->tl_cb_item.alloc(cb_item.size()));
        CLX_CUDA(c, c->tl_cta_first.alloc(cta_first.size()));
        CLX_CUDA(c, cudaMemcpyAsync(c->tl_cta_first.p, cta_first.data(), cta_first.size() * 4, cudaMemcpyHostToDevice, s));
        if (!queue.empty()) {
            CLX_CUDA(c, cudaMemcpyAsync(c->tl_item_q.p, queue.data(), queue.size() * sizeof(uint4), cudaMemcpyHostToDevice, s));
            CLX_CUDA(c, cudaMemcpyAsync(item_x.p, extra.data(), extra.size() * sizeof(uint4), cudaMemcpyHostToDevice, s));
            k_item_rng<<<c->tl_n_items, W, 0, s>>>(c->tl_item_q.p, item_x.p, c->tl_n_items, chunk_off.p, c->tl_nT, W, c->tl_item_rng.p);
            c->launches++;
        }
        CLX_CUDA(c, cudaMemcpyAsync(c->tl_cb_item.p, cb_item.data(), cb_item.size() * 4, cudaMemcpyHostToDevice, s));
        CLX_CUDA(c, cudaStreamSynchronize(s));  // the host vectors above go out of scope
        if (getenv("CLX_TILED_STATS")) fprintf(stderr, "[clx] run stream: %u items on %u CTAs\n", c->tl_n_items, G);
    }
    clk.mark("work items");
    CLX_CUDA(c, c->tl_part.alloc((size_t)c->tl_n_items * c->tl_NB * 2 + 2));
    CLX_CUDA(c, c->tab16.alloc(2ull * c->tl_nT * TILE_SLOTS * 2 + 2));
    CLX_CUDA(c, cudaMemsetAsync(c->tab16.p, 0, c->tab16.bytes(), s));
    CLX_CUDA(c, cudaGetLastError());
    c->tiled_ready = true;
    return CLX_OK;
}

}  // namespace clx

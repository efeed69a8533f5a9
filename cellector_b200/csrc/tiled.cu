// tiled.cu — the tiled per-cell scoring path (get_cell_log_likelihoods, cellector/src/main.rs:541-591).
//
// Why: a memo-table gather served by L1/L2 costs one L1 wavefront per nonzero (measured: the gather kernel is
// bound at ~1 distinct line per SM-cycle, ~25 % of the HBM roofline).  Here the per-locus memo records are staged
// in SHARED memory by TMA bulk copies (cp.async.bulk + mbarrier), tile by tile, and the nonzeros are re-laid out
// so that every CTA streams exactly the entries that hit the tile it holds:
//
//   * only "light" entries (total 1..3) are kept: zero-total entries contribute exactly 0.0 and heavy entries
//     come from the heavy lists; loci_used_per_cell (main.rs:575) is kept as a counter array instead (n_used);
//   * a RUN = the entries of one cell inside one locus tile (TILE_L loci), in file order; an entry is 2 bytes:
//     locus_local << 6 | (total - 1) << 4 | slot, slot = index of T(alt | total) in the 13-double compact record
//     of numerics.cuh (E(total) sits at 9 + total - 1);
//   * a cell belongs to ONE warp of the CTA for the whole pass (cell_local % TL_WARPS), so the read-modify-writes of
//     its shared-memory accumulator never race: no atomics, and no CTA barrier between tiles;
//   * the runs of one (cell block, tile, warp) SUB-UNIT are ordered by length and taken 32 at a time: a GROUP.  A
//     group is stored TRANSPOSED — entry j of lane l at j * 32 + l, shorter runs padded with a null entry that
//     points at an all-zero record — so one lane sums one run with conflict-free 64-byte warp loads, no per-lane
//     trip counts, no shuffles; runs of a group have similar lengths, so padding is small (~2.6 B per nonzero);
//   * groups are cut into BATCHES of at most BATCH_ITERS iterations ([16 B header][32 one-byte cell indices if first][entries])
//     and batches are packed into PACKETS of at most STAGE_BYTES — one TMA bulk copy each into the warp's ring;
//   * sub-units are numbered warp-major ((cell block, warp), then tile): the packets one warp consumes over the
//     consecutive tiles of a work item are ONE consecutive range of the descriptor list {start, size}, which the
//     warp reads 32 at a time with one coalesced load and hands round by shuffle (no dependent global load sits
//     between two packets); a packet's header names its tile, and that is how the warp learns a tile has ended;
//   * a work ITEM = one cell block x a range of consecutive tiles; the CTA keeps the block's accumulators in
//     shared memory across the range and writes one partial per cell at the end; k_score_finish adds the
//     partials of a cell block in item order (fixed order: deterministic), appends the heavy values and normalises.
#include <cub/cub.cuh>
#include <stdio.h>

#include <algorithm>
#include <vector>

#include "ctx.h"
#include "numerics.cuh"

namespace clx {
namespace {

constexpr int TPB = 256;
constexpr uint32_t NB_MAX = 3072;        // cells per block: 16 B of accumulators each = 48 KB of shared memory
constexpr uint32_t NB_MIN = 256;
constexpr int RING_STAGES = 2;           // packets per warp: one being summed, one in flight
constexpr uint32_t STAGE_BYTES = 1024;
constexpr uint32_t BATCH_ITERS = 15;     // 16 + 16 + 32 + 64 * 15 = 1024 <= STAGE_BYTES
constexpr uint32_t NULL_ENTRY = (uint32_t)TILE_L << 6;  // record TILE_L of a staged tile is all zeros
static_assert(16 + 16 + 32 + 64 * BATCH_ITERS <= STAGE_BYTES, "a packet holds at least one full batch");
static_assert(TL_WARPS == 32, "one CTA of 1024 threads");

inline unsigned grid_for(uint64_t n, int tpb = TPB, unsigned cap = 148u * 16u) {
    uint64_t g = (n + tpb - 1) / tpb;
    if (g < 1) g = 1;
    return (unsigned)(g > cap ? cap : g);
}

// ---------------------------------------------------------------------------------------------- builder
// key = (cell block * nT + tile) << 12 | cell_local   (sorts to unit, then cell; stable: file order inside a run);
// non-light entries get ~0 and sort to the end
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
            k = (((unsigned long long)cb * nT + t) << 12) | cl;
            v = (uint16_t)((ll << 6) | ((n - 1u) << 4) | (uint32_t)tab13_T((int)e.alt, (int)n));
            ++mine;
        }
        keys[i] = k;
        vals[i] = v;
    }
    for (int off = 16; off > 0; off >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, off);
    if ((threadIdx.x & 31) == 0 && mine) atomicAdd(n_light, mine);
}

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
// su = (cell block * TL_WARPS + warp) * nT + tile, so that the packets a warp consumes over consecutive tiles of a
// work item are consecutive in the packet list and in the stream.
__device__ __forceinline__ unsigned long long sub_unit_of(unsigned long long unit, uint32_t w, uint32_t nT) {
    const unsigned long long cb = unit / nT, t = unit % nT;
    return (cb * TL_WARPS + w) * nT + t;
}
// runs (unique keys + counts, from cub::DeviceRunLengthEncode) -> sort key (sub-unit << 16 | 0xFFFF - count): longest
// first inside a sub-unit
__global__ void k_run_keys(const unsigned long long* __restrict__ ukeys, const int* __restrict__ counts, uint32_t n_runs,
                           uint32_t nT, unsigned long long* __restrict__ key2, uint32_t* __restrict__ idx,
                           unsigned int* __restrict__ too_long) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_runs) return;
    const uint32_t cnt = (uint32_t)counts[r];
    if (cnt > 0xFFFFu) atomicOr(too_long, 1u);
    const unsigned long long unit = ukeys[r] >> 12, cl = ukeys[r] & 0xFFFull;
    key2[r] = (sub_unit_of(unit, (uint32_t)(cl % TL_WARPS), nT) << 16) | (unsigned long long)(0xFFFFu - (cnt > 0xFFFFu ? 0xFFFFu : cnt));
    idx[r] = r;
}
// first sorted run of every sub-unit, su in [0, n_sub]
__global__ void k_sub_runs(const unsigned long long* __restrict__ key2_sorted, uint32_t n_runs, uint64_t n_sub,
                           uint32_t* __restrict__ sub_run) {
    const uint64_t su = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    if (su <= n_sub) sub_run[su] = (uint32_t)lower_bound_dev<unsigned long long>(key2_sorted, 0, n_runs, su << 16);
}
// cost of every unit for the work-item partition: the stream bytes (in quads) of its 32 sub-units — padded groups
// included, so a ragged cell block whose groups are mostly padding is not mistaken for cheap work
__global__ void k_unit_cost(const uint32_t* __restrict__ n_quads, uint32_t n_units, uint32_t nT,
                            unsigned long long* __restrict__ unit_cost) {
    const uint32_t u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= n_units) return;
    unsigned long long sum = 0;
    for (uint32_t w = 0; w < (uint32_t)TL_WARPS; ++w) sum += n_quads[sub_unit_of(u, w, nT)];
    unit_cost[u] = sum;
}

// The packets of one sub-unit, generated in order; F is called once per batch with
//   (packet index within the sub-unit, byte offset of the batch inside the sub-unit's stream, group, batch, m, first, last)
// and P once per packet with (packet index, start offset, size, n_batches).  Offsets are multiples of 16.
template <typename F, typename P>
__device__ __forceinline__ void walk_sub_unit(const uint32_t* __restrict__ idx_sorted, const int* __restrict__ counts,
                                              uint32_t r0, uint32_t r1, F&& on_batch, P&& on_packet) {
    uint32_t pk = 0, pk_start = 0, pk_bytes = 0, pk_batches = 0;  // current packet (bytes includes its 16-byte header)
    for (uint32_t g0 = r0; g0 < r1; g0 += 32) {
        const uint32_t M = (uint32_t)counts[idx_sorted[g0]];  // longest run of the group
        for (uint32_t j0 = 0; j0 < M; j0 += BATCH_ITERS) {
            const uint32_t m = min(BATCH_ITERS, M - j0);
            const bool first = j0 == 0, last = j0 + m >= M;
            const uint32_t bytes = 16u + (first ? 32u : 0u) + 64u * m;
            if (pk_batches && pk_bytes + bytes > STAGE_BYTES) {
                on_packet(pk, pk_start, pk_bytes, pk_batches);
                pk_start += pk_bytes;
                ++pk;
                pk_bytes = 0;
                pk_batches = 0;
            }
            if (!pk_batches) pk_bytes = 16u;
            on_batch(pk, pk_start + pk_bytes, g0, j0, m, first, last);
            pk_bytes += bytes;
            ++pk_batches;
        }
    }
    if (pk_batches) on_packet(pk, pk_start, pk_bytes, pk_batches);
}

__global__ void k_sub_count(const uint32_t* __restrict__ sub_run, const uint32_t* __restrict__ idx_sorted,
                            const int* __restrict__ counts, uint64_t n_sub, uint32_t* __restrict__ n_pk,
                            uint32_t* __restrict__ n_quads) {
    const uint64_t su = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    if (su > n_sub) return;
    uint32_t packets = 0, bytes = 0;
    if (su < n_sub)
        walk_sub_unit(idx_sorted, counts, sub_run[su], sub_run[su + 1],
                      [](uint32_t, uint32_t, uint32_t, uint32_t, uint32_t, bool, bool) {},
                      [&](uint32_t, uint32_t, uint32_t sz, uint32_t) { ++packets; bytes += sz; });
    n_pk[su] = packets;
    n_quads[su] = bytes >> 4;
}

// one warp per sub-unit writes its packets: packet header {n_batches, tile}, batch header {m | first << 8 | last << 9},
// the 32 cell ids of a first batch, and the transposed entries; and the packet's descriptor {start quad, size in quads}
__global__ void k_sub_fill(const uint32_t* __restrict__ sub_run, const uint32_t* __restrict__ pk_base,
                           const uint32_t* __restrict__ q_base, const uint32_t* __restrict__ idx_sorted,
                           const unsigned long long* __restrict__ ukeys, const int* __restrict__ counts,
                           const unsigned long long* __restrict__ run_start, const uint16_t* __restrict__ vals_sorted,
                           uint64_t n_sub, uint32_t nT, uint2* __restrict__ pk_d, uint32_t* __restrict__ stream) {
    const uint64_t su = (blockIdx.x * (uint64_t)blockDim.x + threadIdx.x) >> 5;
    const uint32_t lane = threadIdx.x & 31;
    if (su >= n_sub) return;
    const uint32_t r0 = sub_run[su], r1 = sub_run[su + 1];
    unsigned char* base = reinterpret_cast<unsigned char*>(stream) + (size_t)q_base[su] * 16;
    const uint32_t pkb = pk_base[su], qb = q_base[su];
    walk_sub_unit(
        idx_sorted, counts, r0, r1,
        [&](uint32_t, uint32_t off, uint32_t g0, uint32_t j0, uint32_t m, bool first, bool last) {
            uint32_t* hdr = reinterpret_cast<uint32_t*>(base + off);
            if (lane < 4) hdr[lane] = lane == 0 ? (m | (first ? 0x100u : 0u) | (last ? 0x200u : 0u)) : 0u;
            const uint32_t r = g0 + lane;
            uint32_t cnt = 0;
            unsigned long long src = 0;
            uint32_t cell = 0xFFFFu;
            if (r < r1) {
                const uint32_t rid = idx_sorted[r];
                cnt = (uint32_t)counts[rid];
                src = run_start[rid];
                cell = (uint32_t)(ukeys[rid] & 0xFFFull);
            }
            uint16_t* p16 = reinterpret_cast<uint16_t*>(base + off + 16);
            if (first) {
                // the run's cell as q = cell_local / 32 (cell_local % 32 is the owning warp): one byte, 0xFF = no run
                reinterpret_cast<uint8_t*>(p16)[lane] = cell == 0xFFFFu ? (uint8_t)0xFF : (uint8_t)(cell >> 5);
                p16 += 16;
            }
            for (uint32_t j = 0; j < m; ++j) {
                const uint32_t jj = j0 + j;
                p16[j * 32 + lane] = jj < cnt ? vals_sorted[src + jj] : (uint16_t)NULL_ENTRY;
            }
        },
        [&](uint32_t pk, uint32_t start, uint32_t sz, uint32_t nb) {
            if (lane == 0) {
                pk_d[pkb + pk] = make_uint2(qb + (start >> 4), sz >> 4);
                uint32_t* ph = reinterpret_cast<uint32_t*>(base + start);
                ph[0] = nb;
                ph[1] = (uint32_t)(su % nT);  // the locus tile this packet indexes
                ph[2] = 0;
                ph[3] = 0;
            }
        });
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

constexpr size_t TL_TAB_BYTES = (size_t)TILE_L * TAB13 * 8;                             // bytes one table bulk copy moves
constexpr size_t TL_TAB_BUF = (((size_t)(TILE_L + 1) * TAB13 * 8) + 127) / 128 * 128;   // + the all-zero record
constexpr size_t TL_ACC_BYTES = (size_t)NB_MAX * 16;
constexpr size_t TL_RING_BYTES = (size_t)TL_WARPS * RING_STAGES * STAGE_BYTES;
constexpr size_t TL_SMEM = 2 * TL_TAB_BUF + TL_ACC_BYTES + TL_RING_BYTES + (2 + TL_WARPS * RING_STAGES) * 8 + 16;
static_assert(TL_SMEM <= 227 * 1024, "shared memory budget of the tiled kernel");

// Per-lane state of the run a lane is summing (a group may span several batches and packets).
struct RunState {
    double sT, sE;
    uint32_t cell;
};

// One packet = n batches.  In a batch, lane l reads entry j at j * 32 + l (one conflict-free 64-byte warp load),
// looks T and E up in the staged table tile and adds them to its run's sums, in file order; the last batch of a
// group adds the run to the cell's accumulator (this warp owns the cell: plain read-modify-write).
// Accumulator slot of a cell: the cells of one warp are cell_local = 32 q + w, and a group touches 32 of them at
// once, so slot = w * Q + q puts them in neighbouring 16-byte slots (slot = cell_local would put all 32 in the
// same banks: an 8-way conflict on every read-modify-write).
__device__ __forceinline__ uint32_t acc_slot(uint32_t cell_local, uint32_t Q) { return (cell_local & 31u) * Q + (cell_local >> 5); }

template <bool WITH_E>
__device__ __forceinline__ void tiled_packet(const unsigned char* __restrict__ pk, const double* __restrict__ tab,
                                             double2* __restrict__ acc, int lane, int w, RunState& st, uint32_t Q) {
    const uint32_t nb = *reinterpret_cast<const uint32_t*>(pk);
    const unsigned char* p = pk + 16;
    for (uint32_t b = 0; b < nb; ++b) {
        const uint32_t hdr = *reinterpret_cast<const uint32_t*>(p);
        const uint32_t m = hdr & 0xFFu;
        const uint16_t* e = reinterpret_cast<const uint16_t*>(p + 16);
        if (hdr & 0x100u) {  // first batch of a group: 32 one-byte cell indices
            const uint32_t q = reinterpret_cast<const uint8_t*>(e)[lane];
            st.cell = q == 0xFFu ? 0xFFFFu : (uint32_t)w * Q + q;  // accumulator slot (acc_slot of cell 32 q + w)
            st.sT = 0.0;
            st.sE = 0.0;
            e += 16;
        }
        double sT = st.sT, sE = st.sE;
#pragma unroll 4
        for (uint32_t j = 0; j < m; ++j) {
            const uint32_t v = e[j * 32 + lane];
            const double* rec = tab + (v >> 6) * TAB13;
            sT += rec[v & 15u];
            if (WITH_E) sE += rec[9 + ((v >> 4) & 3u)];
        }
        st.sT = sT;
        st.sE = sE;
#ifdef CLX_TILED_DEBUG
        if (nb > 64 || m == 0 || m > BATCH_ITERS || ((hdr & 0x200u) && st.cell != 0xFFFFu && st.cell >= 32u * Q))
            printf("bad batch: blk %d w %d lane %d nb %u b %u hdr %x cell %u\n", blockIdx.x, threadIdx.x >> 5, lane, nb, b, hdr, st.cell);
#endif
        if ((hdr & 0x200u) && st.cell != 0xFFFFu) {  // last batch of the group
            double2 a = acc[st.cell];
            a.x += sT;
            a.y += sE;
            acc[st.cell] = a;
        }
        p += 16u + ((hdr & 0x100u) ? 32u : 0u) + 64u * m;
    }
}

// per queue slot and warp: the warp's packet range over the item's tiles (one consecutive range, see sub_unit_of)
__global__ void k_item_pk(const uint4* __restrict__ item_q, uint32_t n_items, const uint32_t* __restrict__ pk_base, uint32_t nT,
                          uint2* __restrict__ item_pk) {
    const uint32_t q = blockIdx.x, w = threadIdx.x;
    if (q >= n_items || w >= (uint32_t)TL_WARPS) return;
    const uint4 it = item_q[q];
    const size_t su0 = ((size_t)it.x * TL_WARPS + w) * nT + it.y;
    const uint32_t P0 = pk_base[su0];
    item_pk[(size_t)q * TL_WARPS + w] = make_uint2(P0, pk_base[su0 + it.z] - P0);
}

// Work queue: CTA b takes queue slot b first, later slots come from an atomic cursor that starts at gridDim.x; the
// cursor for the NEXT item is fetched while the current one is being scored, so no item waits for the atomic.  The
// start of an item costs two dependent global loads (slot record + packet range, then the descriptor window) before
// the first bulk copy goes out.
template <bool WITH_E>
__global__ void __launch_bounds__(TL_WARPS * 32, 1)
    k_score_tiled(const uint32_t* __restrict__ stream, const uint2* __restrict__ pk_d, const uint4* __restrict__ item_q,
                  const uint2* __restrict__ item_pk, uint32_t n_items, uint32_t* __restrict__ counter,
                  const double* __restrict__ tab13, uint32_t NB, double* __restrict__ part) {
    extern __shared__ __align__(128) unsigned char smem[];
    double2* acc = reinterpret_cast<double2*>(smem + 2 * TL_TAB_BUF);
    unsigned char* ring = smem + 2 * TL_TAB_BUF + TL_ACC_BYTES;
    uint64_t* tab_full = reinterpret_cast<uint64_t*>(smem + 2 * TL_TAB_BUF + TL_ACC_BYTES + TL_RING_BYTES);
    uint64_t* ring_full = tab_full + 2;
    unsigned int* tab_done = reinterpret_cast<unsigned int*>(ring_full + TL_WARPS * RING_STAGES);  // [2] warps done with a buffer
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const uint32_t Q = (NB + 31u) / 32u;
    if (tid == 0) {
        mbar_init(&tab_full[0], 1);
        mbar_init(&tab_full[1], 1);
        for (int i = 0; i < TL_WARPS * RING_STAGES; ++i) mbar_init(&ring_full[i], 1);
        tab_done[0] = 0;
        tab_done[1] = 0;
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid < 2 * TAB13) {  // the all-zero record behind each table buffer (null entries point at it)
        double* z = reinterpret_cast<double*>(smem + (size_t)(tid / TAB13) * TL_TAB_BUF) + (size_t)TILE_L * TAB13;
        z[tid % TAB13] = 0.0;
    }
    for (uint32_t i = tid; i < 32u * Q; i += blockDim.x) acc[i] = make_double2(0.0, 0.0);
    __syncthreads();
    uint32_t g_tile = 0;  // tiles consumed by this CTA so far (buffer = g & 1, parity = (g >> 1) & 1)
    uint32_t cons_c = 0;  // packets consumed by this warp so far (stage = c % STAGES, parity = (c / STAGES) & 1)
    unsigned char* my_ring = ring + (size_t)w * RING_STAGES * STAGE_BYTES;
    uint64_t* my_bar = ring_full + w * RING_STAGES;
    RunState st;
    st.sT = 0.0;
    st.sE = 0.0;
    st.cell = 0xFFFFu;
    __shared__ uint32_t s_next;
    uint32_t qk = blockIdx.x;
    while (qk < n_items) {
        uint32_t nxt = 0;
        if (tid == 0) nxt = atomicAdd(counter, 1u);  // used at the end of this item
        const uint4 it = item_q[qk];
        const uint2 rng = item_pk[(size_t)qk * TL_WARPS + w];
        const uint32_t t0 = it.y, ntile = it.z, item = it.w;
        const uint32_t P0 = rng.x, np = rng.y;
        // descriptor window: lane l holds packet win0 + l (and win0 + 32 + l prefetched), one coalesced load per 32
        // packets; nothing on the packet path waits for a dependent global load
        uint32_t win0 = 0;
        uint2 dwin = lane < np ? pk_d[P0 + lane] : make_uint2(0u, 0u);
        uint2 dnxt = 32u + lane < np ? pk_d[P0 + 32u + lane] : make_uint2(0u, 0u);
        if (tid == 0) {
            for (uint32_t k = 0; k < 2 && k < ntile; ++k) {
                const uint32_t g = g_tile + k;
                mbar_expect_tx(&tab_full[g & 1], (uint32_t)TL_TAB_BYTES);
                bulk_g2s(smem + (size_t)(g & 1) * TL_TAB_BUF, tab13 + (size_t)(t0 + k) * TILE_L * TAB13, (uint32_t)TL_TAB_BYTES,
                         &tab_full[g & 1]);
            }
        }
        // packet i of this warp goes to ring stage (cons_c at item start + i) % STAGES; issue runs STAGES ahead
        const uint32_t c_base = cons_c;
        auto issue = [&](uint32_t i) {
            if (i >= np) return;
            if (i >= win0 + 32u) {
                win0 += 32u;
                dwin = dnxt;
                dnxt = win0 + 32u + lane < np ? pk_d[P0 + win0 + 32u + lane] : make_uint2(0u, 0u);
            }
            const uint32_t q0 = __shfl_sync(0xffffffffu, dwin.x, (int)(i - win0));
            const uint32_t bytes = __shfl_sync(0xffffffffu, dwin.y, (int)(i - win0)) * 16u;
            const uint32_t s = (c_base + i) % RING_STAGES;
#ifdef CLX_TILED_DEBUG
            if (lane == 0 && (bytes == 0 || bytes > STAGE_BYTES))
                printf("bad packet: blk %d w %d item %u i %u q0 %u bytes %u\n", blockIdx.x, w, item, i, q0, bytes);
#endif
            if (lane == 0) {
                mbar_expect_tx(&my_bar[s], bytes);
                bulk_g2s(my_ring + (size_t)s * STAGE_BYTES, stream + (size_t)q0 * 4, bytes, &my_bar[s]);
            }
        };
        for (uint32_t i = 0; i < (uint32_t)RING_STAGES; ++i) issue(i);
        // hand a table buffer over: the last warp to leave tile k starts the copy of tile k + 2 into it.  No CTA
        // barrier: cells never change warps, so accumulators need none, and warps drift by at most two tiles.
        auto leave_tile = [&](uint32_t k) {
            const uint32_t g = g_tile + k;
            __syncwarp();
            if (lane == 0) {
                const unsigned int done = atomicAdd(&tab_done[g & 1], 1u);
                if (done == TL_WARPS - 1) {
                    tab_done[g & 1] = 0;
                    if (k + 2 < ntile) {
                        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                        mbar_expect_tx(&tab_full[g & 1], (uint32_t)TL_TAB_BYTES);
                        bulk_g2s(smem + (size_t)(g & 1) * TL_TAB_BUF, tab13 + (size_t)(t0 + k + 2) * TILE_L * TAB13,
                                 (uint32_t)TL_TAB_BYTES, &tab_full[g & 1]);
                    }
                }
            }
        };
        // every warp walks every tile of the item in order (waits for its table, leaves it), with or without packets
        uint32_t k = 0;
        mbar_wait(&tab_full[g_tile & 1], (g_tile >> 1) & 1u);
        for (uint32_t i = 0; i < np; ++i) {
            const uint32_t s = cons_c % RING_STAGES;
            mbar_wait(&my_bar[s], (cons_c / RING_STAGES) & 1u);
            const unsigned char* pk = my_ring + (size_t)s * STAGE_BYTES;
            const uint32_t kt = reinterpret_cast<const uint32_t*>(pk)[1] - t0;  // tile of this packet, item-local
#ifdef CLX_TILED_DEBUG
            if (lane == 0 && (kt >= ntile || kt < k))
                printf("bad tile: blk %d w %d item %u i %u kt %u k %u ntile %u\n", blockIdx.x, w, item, i, kt, k, ntile);
#endif
            while (k < kt) {
                leave_tile(k);
                ++k;
                mbar_wait(&tab_full[(g_tile + k) & 1], ((g_tile + k) >> 1) & 1u);
            }
            const double* tab = reinterpret_cast<const double*>(smem + (size_t)((g_tile + k) & 1) * TL_TAB_BUF);
            tiled_packet<WITH_E>(pk, tab, acc, lane, w, st, Q);
            __syncwarp();
            ++cons_c;
            issue(i + RING_STAGES);  // refill the stage just drained
        }
        for (;;) {
            leave_tile(k);
            if (++k >= ntile) break;
            mbar_wait(&tab_full[(g_tile + k) & 1], ((g_tile + k) >> 1) & 1u);
        }
        g_tile += ntile;
        if (tid == 0) s_next = nxt;
        __syncthreads();  // every warp has added its last run; the next queue slot is posted
        // partials out, accumulators cleared for the next item in the same sweep
        double2* dst = reinterpret_cast<double2*>(part) + (size_t)item * NB;
        for (uint32_t i = tid; i < 32u * Q; i += blockDim.x) {
            const uint32_t slot = acc_slot(i, Q);
            if (i < NB) dst[i] = acc[slot];
            acc[slot] = make_double2(0.0, 0.0);
        }
        qk = s_next;
        __syncthreads();
    }
}

// Adds the partials of a cell's block in item order (FIN_LANES lanes per cell, each a fixed subsequence, combined
// by a fixed shuffle tree: deterministic), appends the heavy values in row order, and (EM pass) normalises by
// loci_used_per_cell (main.rs:313-323).
constexpr int FIN_LANES = 8;
template <bool EM>
__global__ void __launch_bounds__(256) k_score_finish(const double* __restrict__ part, const uint32_t* __restrict__ cb_item,
                                                      uint32_t NB, uint32_t n_cells, const uint32_t* __restrict__ h_row_ptr,
                                                      const double* __restrict__ hv_ll, const double* __restrict__ hv_e,
                                                      const uint32_t* __restrict__ n_used, double* __restrict__ ll,
                                                      double* __restrict__ ell, uint32_t* __restrict__ nloci,
                                                      double* __restrict__ norm, unsigned int* __restrict__ nan_flag,
                                                      uint32_t* __restrict__ queue_counter, uint32_t first_round) {
    if (blockIdx.x == 0 && threadIdx.x == 0) *queue_counter = first_round;  // re-arm the work queue of k_score_tiled
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
    double sT = 0.0, sE = 0.0;
    for (uint32_t it = it0 + sub; it < it1; it += FIN_LANES) {
        const double2 p = __ldg(p2 + (size_t)it * NB + cl);
        sT += p.x;
        sE += p.y;
    }
#pragma unroll
    for (int off = 32 / FIN_LANES; off < 32; off <<= 1) {
        sT += __shfl_xor_sync(0xffffffffu, sT, off);
        if (EM) sE += __shfl_xor_sync(0xffffffffu, sE, off);
    }
    if (!live || sub != 0) return;
    const uint32_t hb = h_row_ptr[i], he = h_row_ptr[i + 1];
    for (uint32_t h = hb; h < he; ++h) {
        sT += hv_ll[h];
        if (EM) sE += hv_e[h];
    }
    ll[i] = sT;
    if (EM) {
        ell[i] = sE;
        const uint32_t cnt = n_used[i];
        nloci[i] = cnt;
        const double nv = cnt > 0 ? sT / (double)cnt : 0.0;
        norm[i] = nv;
        if (nv != nv) atomicOr(nan_flag, 1u);
    }
}

}  // namespace

// Scores every cell against memo table `table` (0 in the EM pass; 0..2 in the posterior pass).
int score_tiled(clx_ctx* c, int table, bool with_e, bool em_outputs) {
    cudaStream_t s = c->stream;
    const uint32_t N = c->n_cells;
    if (N == 0) return CLX_OK;
    static bool attr_done = false;
    if (!attr_done) {
        CLX_CUDA(c, cudaFuncSetAttribute(k_score_tiled<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TL_SMEM));
        CLX_CUDA(c, cudaFuncSetAttribute(k_score_tiled<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TL_SMEM));
        attr_done = true;
    }
    const double* tab = c->tab13.p + (size_t)table * c->tl_nT * TILE_L * TAB13;
    if (c->tl_n_items) {
        const unsigned grid = c->tl_grid;
        if (with_e)
            k_score_tiled<true><<<grid, TL_WARPS * 32, TL_SMEM, s>>>(c->tl_stream.p, c->tl_pk_d.p, c->tl_item_q.p, c->tl_item_pk.p,
                                                                    c->tl_n_items, c->tl_counter.p, tab, c->tl_NB, c->tl_part.p);
        else
            k_score_tiled<false><<<grid, TL_WARPS * 32, TL_SMEM, s>>>(c->tl_stream.p, c->tl_pk_d.p, c->tl_item_q.p, c->tl_item_pk.p,
                                                                     c->tl_n_items, c->tl_counter.p, tab, c->tl_NB, c->tl_part.p);
        c->launches++;
        CLX_DEBUG_SYNC(c, "k_score_tiled");
    }
    const double* hv = c->hv_ll.p + (size_t)table * c->n_heavy;
    const unsigned fgrid = (unsigned)(((uint64_t)N * FIN_LANES + 255) / 256);
    if (em_outputs)
        k_score_finish<true><<<fgrid, 256, 0, s>>>(c->tl_part.p, c->tl_cb_item.p, c->tl_NB, N, c->h_row_ptr.p, hv, c->hv_e.p,
                                                  c->n_used.p, c->ll.p, c->ell.p, c->nloci.p, c->norm.p, &c->d_stats.p->nan_flag,
                                                  c->tl_counter.p, c->tl_grid);
    else
        k_score_finish<false><<<fgrid, 256, 0, s>>>(c->tl_part.p, c->tl_cb_item.p, c->tl_NB, N, c->h_row_ptr.p, hv, nullptr, nullptr,
                                                   c->ll.p + (size_t)table * N, nullptr, nullptr, nullptr, nullptr, c->tl_counter.p,
                                                   c->tl_grid);
    c->launches++;
    return CLX_OK;
}

int build_tiled(clx_ctx* c, const uint32_t* d_cell_of_entry) {
    cudaStream_t s = c->stream;
    c->tiled_ready = false;
    const uint32_t N = c->n_cells, L = c->n_loci;
    const uint64_t nnz = c->nnz;
    CLX_CUDA(c, c->row_len.alloc((size_t)N + 1));
    CLX_CUDA(c, c->n_used.alloc((size_t)N + 1));
    CLX_CUDA(c, cudaMemsetAsync(c->row_len.p, 0, c->row_len.bytes(), s));
    // cell blocks: few (every block re-reads the table tiles) but enough (cell block, tile) units to fill the SMs
    c->tl_nT = L ? (L + TILE_L - 1) / TILE_L : 1;
    {
        uint32_t pc = N ? (N + NB_MAX - 1) / NB_MAX : 1;
        // one work item per SM, each at least ~4 consecutive tiles long (table copies pipeline inside an item, and the
        // cut points of an item are whole tiles: short items balance badly)
        const uint32_t want = (4u * (uint32_t)c->sm_count + c->tl_nT - 1) / c->tl_nT;
        const uint32_t pc_cap = N ? std::max<uint32_t>(1u, N / NB_MIN) : 1u;
        pc = std::max(pc, std::min(want, pc_cap));
        const char* env = getenv("CLX_TILED_PC");
        if (env && *env) pc = std::max<uint32_t>(pc, (uint32_t)atoi(env));
        c->tl_Pc = pc;
        uint32_t nb = N ? (N + pc - 1) / pc : 1;
        // whole groups: 32 warps x 32 lanes = 1024 cells give every (tile, warp) exactly 32 runs; a ragged tail of a few
        // runs per warp costs a whole padded group per tile
        const uint32_t nb_even = (nb + 31) / 32 * 32;
        if (nb >= 1024) {
            const uint32_t nb_whole = std::min<uint32_t>(NB_MAX, nb / 1024 * 1024);
            const uint32_t tail = N % nb_whole;  // cells of the ragged last block
            nb = (tail == 0 || 2 * tail >= nb_whole) ? nb_whole : nb_even;
        } else {
            nb = nb_even;
        }
        const char* env_nb = getenv("CLX_TILED_NB");  // experiments: force the cell-block size (tools/tiled_model.py)
        if (env_nb && *env_nb) nb = std::min<uint32_t>(NB_MAX, std::max<uint32_t>(32u, ((uint32_t)atoi(env_nb) + 31u) / 32u * 32u));
        c->tl_NB = nb;
        c->tl_Pc = N ? (N + c->tl_NB - 1) / c->tl_NB : 1;
    }
    c->tl_n_units = c->tl_Pc * c->tl_nT;
    if ((uint64_t)c->tl_Pc * c->tl_nT >= (1ull << 31)) return fail(c, CLX_ERR_ARG, "too many (cell block, tile) units");
    const uint32_t n_units = c->tl_n_units;

    DevBuf<unsigned long long> keys, keys_sorted, d_nlight, ukeys, key2, key2_sorted, run_start, unit_entries;
    DevBuf<uint16_t> vals, vals_sorted;
    DevBuf<int> counts, d_nruns;
    DevBuf<uint32_t> ridx, ridx_sorted, n_pk;
    DevBuf<unsigned int> too_long;
    DevBuf<uint8_t> tmp;
    size_t tb = 0;
    CLX_CUDA(c, keys.alloc(nnz + 1));
    CLX_CUDA(c, keys_sorted.alloc(nnz + 1));
    CLX_CUDA(c, vals.alloc(nnz + 1));
    CLX_CUDA(c, vals_sorted.alloc(nnz + 1));
    CLX_CUDA(c, d_nlight.alloc(1));
    CLX_CUDA(c, d_nruns.alloc(1));
    CLX_CUDA(c, too_long.alloc(1));
    CLX_CUDA(c, cudaMemsetAsync(d_nlight.p, 0, 8, s));
    CLX_CUDA(c, cudaMemsetAsync(d_nruns.p, 0, 4, s));
    CLX_CUDA(c, cudaMemsetAsync(too_long.p, 0, 4, s));
    unsigned long long n_light = 0;
    if (nnz) {
        k_tile_keys<<<grid_for(nnz), TPB, 0, s>>>(c->csr.p, d_cell_of_entry, nnz, c->tl_NB, c->tl_nT, keys.p, vals.p, d_nlight.p,
                                                  c->row_len.p);
        int bits = 13;
        while (bits < 63 && (1ull << bits) <= (((unsigned long long)n_units) << 12)) ++bits;
        // non-light keys are ~0: sorting on the low `bits + 1` bits still puts them last
        CLX_CUDA(c, cub::DeviceRadixSort::SortPairs(nullptr, tb, keys.p, keys_sorted.p, vals.p, vals_sorted.p, (int)nnz, 0, bits + 1, s));
        CLX_CUDA(c, tmp.alloc(tb));
        CLX_CUDA(c, cub::DeviceRadixSort::SortPairs(tmp.p, tb, keys.p, keys_sorted.p, vals.p, vals_sorted.p, (int)nnz, 0, bits + 1, s));
        c->launches += 5;
        CLX_CUDA(c, cudaMemcpyAsync(&n_light, d_nlight.p, 8, cudaMemcpyDeviceToHost, s));
        CLX_CUDA(c, cudaStreamSynchronize(s));
    }
    c->tl_n_light = n_light;
    keys.release();
    vals.release();
    // ---- runs
    int n_runs = 0;
    CLX_CUDA(c, ukeys.alloc(n_light + 1));
    CLX_CUDA(c, counts.alloc(n_light + 1));
    if (n_light) {
        CLX_CUDA(c, cub::DeviceRunLengthEncode::Encode(nullptr, tb, keys_sorted.p, ukeys.p, counts.p, d_nruns.p, (int)n_light, s));
        CLX_CUDA(c, tmp.alloc(tb));
        CLX_CUDA(c, cub::DeviceRunLengthEncode::Encode(tmp.p, tb, keys_sorted.p, ukeys.p, counts.p, d_nruns.p, (int)n_light, s));
        CLX_CUDA(c, cudaMemcpyAsync(&n_runs, d_nruns.p, 4, cudaMemcpyDeviceToHost, s));
        CLX_CUDA(c, cudaStreamSynchronize(s));
        c->launches += 2;
    }
    keys_sorted.release();
    const uint32_t R = (uint32_t)n_runs;
    const uint64_t n_sub = (uint64_t)n_units * TL_WARPS;
    if (n_sub >= (1ull << 32) - 2) return fail(c, CLX_ERR_ARG, "too many (cell block, tile, warp) sub-units");
    DevBuf<uint32_t> sub_run, n_quads, q_base;
    CLX_CUDA(c, run_start.alloc((size_t)R + 1));
    CLX_CUDA(c, key2.alloc((size_t)R + 1));
    CLX_CUDA(c, key2_sorted.alloc((size_t)R + 1));
    CLX_CUDA(c, ridx.alloc((size_t)R + 1));
    CLX_CUDA(c, ridx_sorted.alloc((size_t)R + 1));
    CLX_CUDA(c, sub_run.alloc(n_sub + 2));
    CLX_CUDA(c, n_pk.alloc(n_sub + 2));
    CLX_CUDA(c, n_quads.alloc(n_sub + 2));
    CLX_CUDA(c, q_base.alloc(n_sub + 2));
    CLX_CUDA(c, unit_entries.alloc((size_t)n_units + 2));
    CLX_CUDA(c, c->tl_pk_base.alloc(n_sub + 2));
    if (R) {
        cub::TransformInputIterator<unsigned long long, IntToU64, const int*> cnt64(counts.p, IntToU64());
        CLX_CUDA(c, cub::DeviceScan::ExclusiveSum(nullptr, tb, cnt64, run_start.p, (int)R, s));
        CLX_CUDA(c, tmp.alloc(tb));
        CLX_CUDA(c, cub::DeviceScan::ExclusiveSum(tmp.p, tb, cnt64, run_start.p, (int)R, s));
        k_run_keys<<<(R + TPB - 1) / TPB, TPB, 0, s>>>(ukeys.p, counts.p, R, c->tl_nT, key2.p, ridx.p, too_long.p);
        int bits2 = 17;
        while (bits2 < 63 && (1ull << bits2) <= (n_sub << 16)) ++bits2;
        CLX_CUDA(c, cub::DeviceRadixSort::SortPairs(nullptr, tb, key2.p, key2_sorted.p, ridx.p, ridx_sorted.p, (int)R, 0, bits2, s));
        CLX_CUDA(c, tmp.alloc(tb));
        CLX_CUDA(c, cub::DeviceRadixSort::SortPairs(tmp.p, tb, key2.p, key2_sorted.p, ridx.p, ridx_sorted.p, (int)R, 0, bits2, s));
        c->launches += 7;
    }
    k_sub_runs<<<(unsigned)((n_sub + 1 + TPB - 1) / TPB), TPB, 0, s>>>(key2_sorted.p, R, n_sub, sub_run.p);
    k_sub_count<<<(unsigned)((n_sub + 1 + TPB - 1) / TPB), TPB, 0, s>>>(sub_run.p, ridx_sorted.p, counts.p, n_sub, n_pk.p, n_quads.p);
    k_unit_cost<<<(n_units + TPB - 1) / TPB, TPB, 0, s>>>(n_quads.p, n_units, c->tl_nT, unit_entries.p);
    CLX_CUDA(c, cub::DeviceScan::ExclusiveSum(nullptr, tb, n_pk.p, c->tl_pk_base.p, (int)(n_sub + 1), s));
    CLX_CUDA(c, tmp.alloc(tb));
    CLX_CUDA(c, cub::DeviceScan::ExclusiveSum(tmp.p, tb, n_pk.p, c->tl_pk_base.p, (int)(n_sub + 1), s));
    CLX_CUDA(c, cub::DeviceScan::ExclusiveSum(nullptr, tb, n_quads.p, q_base.p, (int)(n_sub + 1), s));
    CLX_CUDA(c, tmp.alloc(tb));
    CLX_CUDA(c, cub::DeviceScan::ExclusiveSum(tmp.p, tb, n_quads.p, q_base.p, (int)(n_sub + 1), s));
    c->launches += 7;
    uint32_t n_packets = 0, total_q = 0;
    unsigned int h_too_long = 0;
    std::vector<unsigned long long> h_unit((size_t)n_units + 1);
    CLX_CUDA(c, cudaMemcpyAsync(&n_packets, c->tl_pk_base.p + n_sub, 4, cudaMemcpyDeviceToHost, s));
    CLX_CUDA(c, cudaMemcpyAsync(&total_q, q_base.p + n_sub, 4, cudaMemcpyDeviceToHost, s));
    CLX_CUDA(c, cudaMemcpyAsync(&h_too_long, too_long.p, 4, cudaMemcpyDeviceToHost, s));
    CLX_CUDA(c, cudaMemcpyAsync(h_unit.data(), unit_entries.p, (size_t)n_units * 8, cudaMemcpyDeviceToHost, s));
    CLX_CUDA(c, cudaStreamSynchronize(s));
    if (h_too_long) {
        // more than 65535 entries of one cell inside one locus tile (massively duplicated coordinates): the batch
        // format does not hold such a run; the gather kernel scores this matrix instead
        CLX_CUDA(c, c->tab13.alloc(TAB13));
        return CLX_OK;
    }
    c->tl_stream_len = (uint64_t)total_q * 4;  // in 4-byte words
    CLX_CUDA(c, c->tl_stream.alloc(c->tl_stream_len + 4));
    CLX_CUDA(c, c->tl_pk_d.alloc((size_t)n_packets + 2));
    if (n_packets) {
        k_sub_fill<<<(unsigned)((n_sub * 32 + TPB - 1) / TPB), TPB, 0, s>>>(sub_run.p, c->tl_pk_base.p, q_base.p, ridx_sorted.p, ukeys.p,
                                                                          counts.p, run_start.p, vals_sorted.p, n_sub, c->tl_nT,
                                                                          c->tl_pk_d.p, c->tl_stream.p);
        c->launches++;
    }
    CLX_CUDA(c, cudaStreamSynchronize(s));

    const char* env_it = getenv("CLX_TILED_ITEMS_PER_SM");
    const int env_items = env_it && *env_it ? std::max(1, atoi(env_it)) : 2;
    // ---- work items: every cell block gets a share of (items per SM x SMs) items proportional to its entries; its tiles are cut
    //      into that many consecutive ranges of about equal entry counts
    {
        std::vector<uint32_t> items, cb_item;
        std::vector<unsigned long long> item_entries;
        unsigned long long cost_total = 0;
        for (uint32_t u = 0; u < n_units; ++u) cost_total += h_unit[u];
        const uint64_t total_items = std::max<uint64_t>(1, (uint64_t)env_items * (uint64_t)c->sm_count);
        for (uint32_t cb = 0; cb < c->tl_Pc; ++cb) {
            cb_item.push_back((uint32_t)(items.size() / 4));
            const unsigned long long* ue = h_unit.data() + (size_t)cb * c->tl_nT;
            unsigned long long cb_total = 0;
            for (uint32_t t = 0; t < c->tl_nT; ++t) cb_total += ue[t];
            if (cb_total == 0) continue;
            uint64_t want = cost_total ? (cb_total * total_items + cost_total / 2) / cost_total : 1;
            want = std::max<uint64_t>(1, std::min<uint64_t>(want, c->tl_nT));
            unsigned long long acc = 0;
            uint32_t t0 = 0;
            for (uint64_t j = 1; j <= want && t0 < c->tl_nT; ++j) {
                const unsigned long long goal = cb_total * j / want;
                uint32_t t1 = t0;
                while (t1 < c->tl_nT && (acc < goal || t1 == t0)) acc += ue[t1++];
                if (j == want) {
                    while (t1 < c->tl_nT) acc += ue[t1++];
                }
                unsigned long long in_range = 0;
                for (uint32_t t = t0; t < t1; ++t) in_range += ue[t];
                if (in_range) {
                    items.push_back(cb);
                    items.push_back(t0);
                    items.push_back(t1);
                    items.push_back(0);
                    item_entries.push_back(in_range);
                }
                t0 = t1;
            }
        }
        cb_item.push_back((uint32_t)(items.size() / 4));
        c->tl_n_items = (uint32_t)(items.size() / 4);
        c->tl_grid = std::min<uint32_t>(c->tl_n_items, (uint32_t)c->sm_count);
        CLX_CUDA(c, c->tl_item_q.alloc((size_t)c->tl_n_items + 1));
        CLX_CUDA(c, c->tl_item_pk.alloc((size_t)c->tl_n_items * TL_WARPS + 1));
        CLX_CUDA(c, c->tl_cb_item.alloc(cb_item.size()));
        std::vector<uint32_t> order(c->tl_n_items);
        for (uint32_t i = 0; i < c->tl_n_items; ++i) order[i] = i;
        std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return item_entries[a] > item_entries[b]; });
        std::vector<uint4> queue(c->tl_n_items);  // largest first
        for (uint32_t q = 0; q < c->tl_n_items; ++q) {
            const uint32_t i = order[q];
            queue[q] = make_uint4(items[4 * i], items[4 * i + 1], items[4 * i + 2] - items[4 * i + 1], i);
        }
        CLX_CUDA(c, c->tl_counter.alloc(1));
        CLX_CUDA(c, cudaMemcpyAsync(c->tl_counter.p, &c->tl_grid, 4, cudaMemcpyHostToDevice, s));
        if (!queue.empty()) {
            CLX_CUDA(c, cudaMemcpyAsync(c->tl_item_q.p, queue.data(), queue.size() * sizeof(uint4), cudaMemcpyHostToDevice, s));
            k_item_pk<<<c->tl_n_items, TL_WARPS, 0, s>>>(c->tl_item_q.p, c->tl_n_items, c->tl_pk_base.p, c->tl_nT, c->tl_item_pk.p);
            c->launches++;
        }
        CLX_CUDA(c, cudaMemcpyAsync(c->tl_cb_item.p, cb_item.data(), cb_item.size() * 4, cudaMemcpyHostToDevice, s));
        CLX_CUDA(c, cudaStreamSynchronize(s));
    }
    CLX_CUDA(c, c->tl_part.alloc((size_t)c->tl_n_items * c->tl_NB * 2 + 2));
    CLX_CUDA(c, c->tab13.alloc(3ull * c->tl_nT * TILE_L * TAB13 + TAB13));
    CLX_CUDA(c, cudaMemsetAsync(c->tab13.p, 0, c->tab13.bytes(), s));
    CLX_CUDA(c, cudaStreamSynchronize(s));
    CLX_CUDA(c, cudaGetLastError());
    c->tiled_ready = true;
    return CLX_OK;
}

}  // namespace clx

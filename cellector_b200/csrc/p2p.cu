// p2p.cu — the per-iteration exchanges of the cell-sharded path over NVLink peer memory instead of NCCL launches.
//
// Every rank owns a WINDOW (plain cudaMalloc, exported with cudaIpcGetMemHandle, mapped by every peer at load):
//   flags    [3 phases][world] u64   epoch numbers written by the peers (release.sys stores), polled locally
//   counters [4] u64                 this rank's new / rescued / excluded counts of the iteration
//   norm_all [world * max_local] f64 every rank's normalized log likelihoods (each rank stores its slice into
//                                    every window: the all-gather is remote stores, no collective)
//   part_ll  [2L] f64, part_int [8L] u64   this rank's column-pass partial sums (peers read them)
//   tot_ll   [2L] f64, tot_int  [8L] u64   the sums over ranks (the owner of a locus slice stores them everywhere)
// Per EM iteration: put_norm -> barrier 0 -> select -> columns -> barrier 1 -> reduce (rank r sums the loci of
// slice r over all ranks in rank order — identical on every rank, integers exact — and stores the totals into
// every window) -> barrier 2 -> filter.  A barrier is one tiny kernel: thread t signals peer t and spins on the
// flag peer t writes here; a spin that lasts longer than ~2 s sets an error flag and gives up (no hang).
#include <string.h>

#include <algorithm>
#include <vector>

#include "ctx.h"
#include "p2p_dev.cuh"

namespace clx {
namespace {

constexpr size_t WIN_FLAGS_BYTES = 4096;
constexpr int P2P_PHASES = 3;
// the counters block sits 256 bytes before the norm area (p2p_setup)
__host__ __device__ constexpr size_t off_norm_to_counters(size_t off_norm) { return off_norm - 256; }

// window base pointers of all ranks (this rank's own included), in device memory
// The epoch lives in the window (slot 8 of the counters block) and is bumped by k_p2p_put_norm at the head of every
// iteration, so the launch arguments never change and the whole sharded iteration replays as one CUDA graph.
__global__ void k_p2p_barrier(unsigned char* const* __restrict__ win, int rank, int world, int phase, size_t off_counters,
                              unsigned int* __restrict__ err) {
    const int t = threadIdx.x;
    if (t >= world) return;
    const unsigned long long epoch = reinterpret_cast<const unsigned long long*>(win[rank] + off_counters)[8];
    const long long limit = (long long)reinterpret_cast<const unsigned long long*>(win[rank] + off_counters)[9];
    __threadfence_system();  // everything this rank stored before the barrier is visible before the flag is
    unsigned long long* theirs = reinterpret_cast<unsigned long long*>(win[t]) + (size_t)phase * world + rank;
    st_release_sys(theirs, epoch);
    const unsigned long long* mine = reinterpret_cast<const unsigned long long*>(win[rank]) + (size_t)phase * world + t;
    const long long t0 = clock64();
    while (ld_acquire_sys(mine) < epoch) {
        if (clock64() - t0 > limit) {  // a peer died or never got here
            atomicOr(err, 1u);
            break;
        }
    }
}

// all-gather by remote stores: this rank's normalized values (pads = +inf) into slot `rank` of every window
__global__ void k_p2p_put_norm(unsigned char* const* __restrict__ win, int rank, int world, size_t off_norm,
                               const double* __restrict__ norm, uint32_t n_local, uint32_t max_local) {
    const double inf = __longlong_as_double(0x7ff0000000000000ll);
    if (blockIdx.x == 0 && threadIdx.x == 0) reinterpret_cast<unsigned long long*>(win[rank] + off_norm_to_counters(off_norm))[8] += 1ull;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < max_local; i += gridDim.x * blockDim.x) {
        const double v = i < n_local ? norm[i] : inf;
        for (int r = 0; r < world; ++r) {
            double* dst = reinterpret_cast<double*>(win[r] + off_norm) + (size_t)rank * max_local + i;
            *dst = v;
        }
    }
}

constexpr int P2P_BATCH = 8;  // peer loads of one element in flight together

// reduce-scatter + all-gather in one kernel: rank r owns loci [L r / world, L (r+1) / world) of each of the ten arrays
__global__ void k_p2p_reduce(unsigned char* const* __restrict__ win, int rank, int world, size_t off_part_ll, size_t off_part_int,
                             size_t off_tot_ll, size_t off_tot_int, size_t off_counters, uint32_t L,
                             const unsigned long long* __restrict__ my_counters, unsigned int* __restrict__ err) {
    // this rank's change counters go next to its partial sums, then barrier 1 (every rank's column pass is done)
    if (blockIdx.x == 0 && threadIdx.x < 3) reinterpret_cast<unsigned long long*>(win[rank] + off_counters)[threadIdx.x] = my_counters[threadIdx.x];
    __syncthreads();
    p2p_sync_head(win, rank, world, 1, off_counters, err);
    const uint32_t l0 = (uint32_t)((uint64_t)L * rank / world), l1 = (uint32_t)((uint64_t)L * (rank + 1) / world);
    const uint32_t n = l1 - l0;
    // ten arrays of L: a = 0, 1 are the f64 arrays, 2..9 the u64 arrays
    for (uint64_t k = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; k < 10ull * n; k += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t a = (uint32_t)(k / n), l = l0 + (uint32_t)(k % n);
        // a peer load takes ~1 us: all of an element's loads are issued before the first one is used (up to P2P_BATCH
        // ranks at a time), then added in rank order — the order every rank uses, so the doubles agree everywhere
        if (a < 2) {
            double sum = 0.0;
            for (int r0 = 0; r0 < world; r0 += P2P_BATCH) {
                double v[P2P_BATCH];
#pragma unroll
                for (int j = 0; j < P2P_BATCH; ++j)
                    v[j] = r0 + j < world ? reinterpret_cast<const double*>(win[r0 + j] + off_part_ll)[(size_t)a * L + l] : 0.0;
#pragma unroll
                for (int j = 0; j < P2P_BATCH; ++j)
                    if (r0 + j < world) sum += v[j];
            }
            for (int r = 0; r < world; ++r) reinterpret_cast<double*>(win[r] + off_tot_ll)[(size_t)a * L + l] = sum;
        } else {
            unsigned long long sum = 0;
            for (int r0 = 0; r0 < world; r0 += P2P_BATCH) {
                unsigned long long v[P2P_BATCH];
#pragma unroll
                for (int j = 0; j < P2P_BATCH; ++j)
                    v[j] = r0 + j < world ? reinterpret_cast<const unsigned long long*>(win[r0 + j] + off_part_int)[(size_t)(a - 2) * L + l] : 0ull;
#pragma unroll
                for (int j = 0; j < P2P_BATCH; ++j) sum += v[j];
            }
            for (int r = 0; r < world; ++r) reinterpret_cast<unsigned long long*>(win[r] + off_tot_int)[(size_t)(a - 2) * L + l] = sum;
        }
    }
    // the three change counters: rank 0 sums them and stores the totals into slot [4..7) of every window
    if (rank == 0 && blockIdx.x == 0 && threadIdx.x < 3) {
        unsigned long long sum = 0;
        for (int r = 0; r < world; ++r) sum += reinterpret_cast<const unsigned long long*>(win[r] + off_counters)[threadIdx.x];
        for (int r = 0; r < world; ++r) reinterpret_cast<unsigned long long*>(win[r] + off_counters)[4 + threadIdx.x] = sum;
    }
}

}  // namespace

void p2p_teardown(clx_ctx* c) {
    for (size_t r = 0; r < c->peer_win.size(); ++r)
        if ((int)r != c->rank && c->peer_win[r]) cudaIpcCloseMemHandle(c->peer_win[r]);
    c->peer_win.clear();
    if (c->win) cudaFree(c->win);
    c->win = nullptr;
    c->p2p = false;
}

// Collective over the NCCL communicator: (re)allocates the window for L loci and max_local cells per rank and maps
// every peer's window.  Any failure (no peer access, IPC refused) leaves c->p2p false: the NCCL path is used.
int p2p_setup(clx_ctx* c, size_t L, size_t max_local) {
    if (c->world <= 1) return CLX_OK;
    const char* env = getenv("CLX_P2P");
    const bool want = !(env && env[0] == '0');
    auto up = [](size_t v) { return (v + 255) / 256 * 256; };
    c->win_off_counters = WIN_FLAGS_BYTES;
    c->win_off_norm = c->win_off_counters + 256;
    c->win_off_part_ll = c->win_off_norm + up((size_t)c->world * max_local * 8 + 8);
    c->win_off_part_int = c->win_off_part_ll + up(2 * L * 8 + 8);
    c->win_off_tot_ll = c->win_off_part_int + up(8 * L * 8 + 8);
    c->win_off_tot_int = c->win_off_tot_ll + up(2 * L * 8 + 8);
    const size_t need = c->win_off_tot_int + up(8 * L * 8 + 8);
    // barrier time limit in SM cycles, slot 9 of the counters block (slot 8 is the epoch)
    auto put_limit = [&]() -> int {
        const char* tenv = getenv("CLX_P2P_TIMEOUT_MS");
        const double ms = tenv && *tenv ? atof(tenv) : 2000.0;
        int khz = 1900000;
        cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, c->device);
        const unsigned long long cycles = (unsigned long long)(ms * (double)khz);
        CLX_CUDA(c, cudaMemcpyAsync((unsigned char*)c->win + c->win_off_counters + 9 * 8, &cycles, 8, cudaMemcpyHostToDevice, c->stream));
        CLX_CUDA(c, cudaStreamSynchronize(c->stream));
        return CLX_OK;
    };
    // a window that is already mapped everywhere and large enough is kept (mapping costs milliseconds): the layout
    // depends only on global sizes, flags and epoch simply continue.  Unanimous or not at all.
    DevBuf<unsigned long long> sync;
    CLX_CUDA(c, sync.alloc(2));
    {
        const unsigned long long keep = (c->p2p && !c->p2p_failed && want && c->win && need <= c->win_bytes) ? 1ull : 0ull;
        CLX_CUDA(c, cudaMemcpyAsync(sync.p, &keep, 8, cudaMemcpyHostToDevice, c->stream));
        CLX_TRY(comm_allreduce_u64(c, sync.p, 1));
        unsigned long long n_keep = 0;
        CLX_CUDA(c, cudaMemcpyAsync(&n_keep, sync.p, 8, cudaMemcpyDeviceToHost, c->stream));
        CLX_CUDA(c, cudaStreamSynchronize(c->stream));
        if (n_keep == (unsigned long long)c->world) {
            c->p2p_failed = false;
            return put_limit();
        }
    }
    // peers may still be reading the old window: everyone closes its mappings, then a collective, then the free
    for (size_t r = 0; r < c->peer_win.size(); ++r)
        if ((int)r != c->rank && c->peer_win[r]) cudaIpcCloseMemHandle(c->peer_win[r]);
    c->peer_win.clear();
    c->p2p = false;
    CLX_CUDA(c, cudaMemsetAsync(sync.p, 0, 16, c->stream));
    CLX_TRY(comm_allreduce_u64(c, sync.p, 1));
    CLX_CUDA(c, cudaStreamSynchronize(c->stream));
    if (c->win) cudaFree(c->win);
    c->win = nullptr;
    c->win_bytes = need + need / 4;  // slack: the next load of a similar matrix reuses the mapping
    unsigned long long ok = 0;
    cudaIpcMemHandle_t mine;
    memset(&mine, 0, sizeof(mine));
    if (want && cudaMalloc(&c->win, c->win_bytes) == cudaSuccess && cudaMemset(c->win, 0, c->win_bytes) == cudaSuccess &&
        cudaIpcGetMemHandle(&mine, c->win) == cudaSuccess)
        ok = 1;
    cudaGetLastError();
    // exchange the handles (64 bytes each) + the per-rank verdict through the communicator we already have
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    DevBuf<unsigned long long> d_h;
    const size_t per = 9;  // 8 words of handle + 1 verdict
    CLX_CUDA(c, d_h.alloc(per * ((size_t)c->world + 1)));
    unsigned long long hbuf[9];
    memcpy(hbuf, &mine, 64);
    hbuf[8] = ok;
    CLX_CUDA(c, cudaMemcpyAsync(d_h.p + per * c->world, hbuf, sizeof(hbuf), cudaMemcpyHostToDevice, c->stream));
    CLX_TRY(comm_allgather_u64(c, d_h.p + per * c->world, d_h.p, per));
    std::vector<unsigned long long> all(per * c->world);
    CLX_CUDA(c, cudaMemcpyAsync(all.data(), d_h.p, all.size() * 8, cudaMemcpyDeviceToHost, c->stream));
    CLX_CUDA(c, cudaStreamSynchronize(c->stream));
    bool all_ok = true;
    for (int r = 0; r < c->world; ++r) all_ok = all_ok && all[per * r + 8] == 1;
    unsigned long long mapped = all_ok ? 1 : 0;
    c->peer_win.assign((size_t)c->world, nullptr);
    if (all_ok) {
        for (int r = 0; r < c->world && mapped; ++r) {
            if (r == c->rank) {
                c->peer_win[r] = c->win;
                continue;
            }
            cudaIpcMemHandle_t h;
            memcpy(&h, &all[per * r], 64);
            void* p = nullptr;
            if (cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
                cudaGetLastError();
                mapped = 0;
            } else {
                c->peer_win[r] = p;
            }
        }
    }
    // everyone must agree, or nobody uses the window
    CLX_CUDA(c, cudaMemcpyAsync(sync.p, &mapped, 8, cudaMemcpyHostToDevice, c->stream));
    CLX_TRY(comm_allreduce_u64(c, sync.p, 1));
    unsigned long long n_mapped = 0;
    CLX_CUDA(c, cudaMemcpyAsync(&n_mapped, sync.p, 8, cudaMemcpyDeviceToHost, c->stream));
    CLX_CUDA(c, cudaStreamSynchronize(c->stream));
    if (n_mapped != (unsigned long long)c->world) {
        for (int r = 0; r < c->world; ++r)
            if (r != c->rank && c->peer_win[r]) cudaIpcCloseMemHandle(c->peer_win[r]);
        c->peer_win.clear();
        if (c->win) cudaFree(c->win);
        c->win = nullptr;
        return CLX_OK;  // NCCL path
    }
    CLX_CUDA(c, c->d_peer_win.alloc((size_t)c->world));
    CLX_CUDA(c, cudaMemcpyAsync(c->d_peer_win.p, c->peer_win.data(), (size_t)c->world * sizeof(void*), cudaMemcpyHostToDevice, c->stream));
    CLX_CUDA(c, cudaStreamSynchronize(c->stream));
    c->p2p_epoch = 0;
    c->p2p = true;
    c->p2p_failed = false;
    return put_limit();
}

int p2p_barrier(clx_ctx* c, int phase) {
    // one epoch per iteration, one flag row per phase: flags only ever grow, so a slow rank never sees a stale "go"
    k_p2p_barrier<<<1, 32, 0, c->stream>>>((unsigned char* const*)c->d_peer_win.p, c->rank, c->world, phase, c->win_off_counters, &c->d_stats.p->pad);
    c->launches++;
    return CLX_OK;
}
int p2p_put_norm(clx_ctx* c) {
    const uint32_t m = c->max_local_cells;
    if (!m) return CLX_OK;
    k_p2p_put_norm<<<std::min<unsigned>((m + 255) / 256, (unsigned)c->sm_count), 256, 0, c->stream>>>(
        (unsigned char* const*)c->d_peer_win.p, c->rank, c->world, c->win_off_norm, c->norm.p, c->n_cells, m);
    c->launches++;
    return CLX_OK;
}
int p2p_reduce(clx_ctx* c) {
    const uint32_t L = c->n_loci;
    const uint64_t work = 10ull * (L / c->world + 1);
    // latency-bound on NVLink round trips: as many threads in flight as the SMs hold (8 CTAs of 256 each)
    k_p2p_reduce<<<(unsigned)std::min<uint64_t>((work + 255) / 256, (uint64_t)c->sm_count * 8), 256, 0, c->stream>>>(
        (unsigned char* const*)c->d_peer_win.p, c->rank, c->world, c->win_off_part_ll, c->win_off_part_int, c->win_off_tot_ll,
        c->win_off_tot_int, c->win_off_counters, L, &c->d_stats.p->n_new, &c->d_stats.p->pad);
    c->launches++;
    return CLX_OK;
}

}  // namespace clx

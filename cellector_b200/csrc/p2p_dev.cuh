// p2p_dev.cuh — device side of the peer-memory barrier (see p2p.cu for the window layout).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>

namespace clx {

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

// Cross-rank barrier at the head of a kernel: block 0 signals every peer (flag row `phase`, epoch from this rank's
// window), every block waits until all peers have signalled this rank, then the CTA proceeds.  Flags only grow.  A
// wait longer than the limit in counters[9] (SM cycles; CLX_P2P_TIMEOUT_MS, default 2000 ms) sets *err and gives up (no hang).  blockDim.x must be >= world.
__device__ __forceinline__ void p2p_sync_head(unsigned char* const* __restrict__ win, int rank, int world, int phase,
                                              size_t off_counters, unsigned int* __restrict__ err) {
    const int t = threadIdx.x;
    if (t < world) {
        const unsigned long long epoch = reinterpret_cast<const unsigned long long*>(win[rank] + off_counters)[8];
        const long long limit = (long long)reinterpret_cast<const unsigned long long*>(win[rank] + off_counters)[9];
        if (blockIdx.x == 0) {
            __threadfence_system();  // what this rank stored before the barrier is visible before the flag is
            st_release_sys(reinterpret_cast<unsigned long long*>(win[t]) + (size_t)phase * world + rank, epoch);
        }
        const unsigned long long* mine = reinterpret_cast<const unsigned long long*>(win[rank]) + (size_t)phase * world + t;
        const long long t0 = clock64();
        while (ld_acquire_sys(mine) < epoch) {
            if (clock64() - t0 > limit) {
                atomicOr(err, 1u);
                break;
            }
        }
    }
    __syncthreads();
}

}  // namespace clx

// noalloc.cu — clx_locus_filter_static: the locus filter (get_loci_used, cellector/src/load_data.rs:254-280) as an
// entry point that allocates nothing.  It has no counterpart in the reference; it exists for callers that cannot call
// cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// The filter needs scratch: two u32 counters per locus (entries with ref > 0, entries with alt > 0), the same counters
// load.cu's k_locus_hist fills in a cudaMalloc'ed buffer.  Here they live in ONE file-scope __device__ array of a
// fixed maximum size; every GPU that loads the library holds these 16 MiB + 256 B until the process exits.
//   * In host code `g_static_scratch` is a host-side stand-in for the symbol, NOT a device address: the address the
//     kernels get comes from cudaGetSymbolAddress, asked once per device with that device current (scratch_for()).
//   * Every call shares the array and finds the previous call's counters in it: the words this call uses are cleared
//     on the caller's stream before the histogram kernel.  Two calls that overlap on one GPU (different streams or
//     host threads) race on the counters; the caller serialises them.
//   * No cudaMalloc / cudaFree / cub temp storage / torch allocation anywhere on this path: cudaSetDevice,
//     cudaGetSymbolAddress, cudaMemsetAsync, three kernel launches, one 8-byte cudaMemcpyAsync to the host stack.
#include <mutex>
#include <string>

#include "ctx.h"

#define CLX_STATIC_MAX_LOCI_ (1u << 21)  // must equal CLX_STATIC_MAX_LOCI of include/cellector_cuda.h

constexpr uint32_t kMaxLoci = CLX_STATIC_MAX_LOCI_;
constexpr size_t kTailWords = 64;  // [0] coordinate-error flag, [1] number of loci kept; padded to 256 B

// The scratch: [2 * l] = entries of locus l with ref > 0, [2 * l + 1] = with alt > 0; then the tail words.
static __device__ __align__(256) unsigned int g_static_scratch[2 * (size_t)kMaxLoci + kTailWords];

namespace clx {
namespace {

constexpr int TPB = 256;
constexpr int kMaxDevices = 64;

std::mutex g_mu;
unsigned int* g_scratch_addr[kMaxDevices] = {nullptr};  // device address of g_static_scratch on each device

// Device address of the static array on `device` (which must be current).  cudaGetSymbolAddress loads the module on
// that device the first time; the answer differs per device, hence the table.
cudaError_t scratch_for(int device, unsigned int** out) {
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_scratch_addr[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess) return e;
        g_scratch_addr[device] = static_cast<unsigned int*>(p);
    }
    *out = g_scratch_addr[device];
    return cudaSuccess;
}

// same counting rule as load.cu:k_locus_hist (load_data.rs:268-273); cnt is the static array's device address
__global__ void k_static_locus_hist(const uint32_t* __restrict__ locus, const uint32_t* __restrict__ ref,
                                    const uint32_t* __restrict__ alt, uint64_t nnz0, uint32_t n_loci0,
                                    unsigned int* __restrict__ cnt, unsigned int* __restrict__ tail) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < nnz0; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t l = locus[i];
        if (l >= n_loci0) {
            atomicOr(&tail[0], 1u);
            continue;
        }
        if (ref[i] > 0) atomicAdd(&cnt[2 * (size_t)l], 1u);
        if (alt[i] > 0) atomicAdd(&cnt[2 * (size_t)l + 1], 1u);
    }
}

// flag[l] = both counts reach their minimum; the kept loci are counted per warp into tail[1]
__global__ void k_static_locus_flags(const unsigned int* __restrict__ cnt, uint32_t n_loci0, uint32_t min_ref,
                                     uint32_t min_alt, uint8_t* __restrict__ used, unsigned int* __restrict__ tail) {
    const uint32_t l = blockIdx.x * blockDim.x + threadIdx.x;
    bool keep = false;
    if (l < n_loci0) {
        keep = cnt[2 * (size_t)l] >= min_ref && cnt[2 * (size_t)l + 1] >= min_alt;
        used[l] = keep ? 1 : 0;
    }
    const unsigned m = __ballot_sync(0xffffffffu, keep);
    if ((threadIdx.x & 31) == 0 && m) atomicAdd(&tail[1], (unsigned)__popc(m));
}

__global__ void k_static_publish(const unsigned int* __restrict__ tail, uint32_t* __restrict__ n_used) {
    if (n_used) *n_used = tail[1];
}

}  // namespace
}  // namespace clx

extern "C" int clx_locus_filter_static(int device, uint64_t nnz0, uint32_t n_loci0, const uint32_t* d_locus,
                                       const uint32_t* d_ref, const uint32_t* d_alt, uint32_t min_ref, uint32_t min_alt,
                                       uint8_t* d_used, uint32_t* d_n_used, void* stream) {
    using namespace clx;
    if (device < 0 || device >= kMaxDevices) return fail(nullptr, CLX_ERR_ARG, "clx_locus_filter_static: device index out of range");
    if (n_loci0 > kMaxLoci)
        return fail(nullptr, CLX_ERR_ARG,
                    "clx_locus_filter_static: n_loci0 = " + std::to_string(n_loci0) + " exceeds CLX_STATIC_MAX_LOCI = " +
                        std::to_string(kMaxLoci) + " (the static scratch array is fixed-size; use clx_load_coo)");
    if (nnz0 > 0 && (!d_locus || !d_ref || !d_alt)) return fail(nullptr, CLX_ERR_ARG, "clx_locus_filter_static: null COO array");
    if (n_loci0 > 0 && !d_used) return fail(nullptr, CLX_ERR_ARG, "clx_locus_filter_static: null output array");
    CLX_CUDA(nullptr, cudaSetDevice(device));
    unsigned int* cnt = nullptr;
    CLX_CUDA(nullptr, scratch_for(device, &cnt));
    unsigned int* tail = cnt + 2 * (size_t)kMaxLoci;
    cudaStream_t s = static_cast<cudaStream_t>(stream);

    // the array still holds the previous call's counters: clear the words this call reads
    if (n_loci0) CLX_CUDA(nullptr, cudaMemsetAsync(cnt, 0, 2 * (size_t)n_loci0 * sizeof(unsigned int), s));
    CLX_CUDA(nullptr, cudaMemsetAsync(tail, 0, kTailWords * sizeof(unsigned int), s));
    if (nnz0) {
        uint64_t g = (nnz0 + TPB - 1) / TPB;
        const uint64_t cap = 148u * 16u;  // 16 resident CTAs of 256 threads on each of the 148 SMs
        k_static_locus_hist<<<(unsigned)(g > cap ? cap : g), TPB, 0, s>>>(d_locus, d_ref, d_alt, nnz0, n_loci0, cnt, tail);
    }
    if (n_loci0) k_static_locus_flags<<<(n_loci0 + TPB - 1) / TPB, TPB, 0, s>>>(cnt, n_loci0, min_ref, min_alt, d_used, tail);
    k_static_publish<<<1, 1, 0, s>>>(tail, d_n_used);
    CLX_CUDA(nullptr, cudaGetLastError());
    unsigned int h_tail[2] = {0, 0};  // pageable stack memory: the copy is staged by the driver, nothing is allocated here
    CLX_CUDA(nullptr, cudaMemcpyAsync(h_tail, tail, sizeof(h_tail), cudaMemcpyDeviceToHost, s));
    CLX_CUDA(nullptr, cudaStreamSynchronize(s));
    if (h_tail[0]) return fail(nullptr, CLX_ERR_ARG, "clx_locus_filter_static: COO coordinate out of range (locus >= n_loci0)");
    return CLX_OK;
}

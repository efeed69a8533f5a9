// comm.cu — NCCL plumbing for the cell-sharded multi-GPU path (one process per GPU).
// NCCL is resolved at run time with dlopen so that the single-GPU path has no NCCL dependency; in a process that
// already loaded torch's bundled libnccl.so.2 the same copy is reused (same SONAME).
// Collectives per EM iteration: all-gather of the normalized log likelihoods (global, exact
// threshold), all-reduce of the per-locus contribution / allele-count arrays (integers: bit-exact, order-free).
#include <dlfcn.h>

#include <cstring>

#include "ctx.h"

namespace clx {
namespace {

// minimal NCCL ABI (nccl.h 2.x): stable since 2.0
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
enum { ncclSuccess = 0 };
enum { ncclInt8 = 0, ncclChar = 0, ncclUint8 = 1, ncclInt32 = 2, ncclUint32 = 3, ncclInt64 = 4, ncclUint64 = 5,
       ncclFloat16 = 6, ncclFloat32 = 7, ncclFloat64 = 8 };
enum { ncclSum = 0 };

struct Api {
    void* h = nullptr;
    int (*GetUniqueId)(ncclUniqueId*) = nullptr;
    int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    std::string err;
    bool ok = false;
};

Api& api() {
    static Api a;
    if (a.h || !a.err.empty()) return a;
    const char* names[] = {"libnccl.so.2", "libnccl.so", nullptr};
    for (int i = 0; names[i] && !a.h; ++i) a.h = dlopen(names[i], RTLD_NOW | RTLD_GLOBAL);
    if (!a.h) {
        a.err = std::string("cannot dlopen libnccl.so.2: ") + dlerror();
        return a;
    }
#define CLX_SYM(field, name)                                                   \
    *(void**)(&a.field) = dlsym(a.h, name);                                    \
    if (!a.field) { a.err = std::string("missing NCCL symbol ") + name; return a; }
    CLX_SYM(GetUniqueId, "ncclGetUniqueId")
    CLX_SYM(CommInitRank, "ncclCommInitRank")
    CLX_SYM(CommDestroy, "ncclCommDestroy")
    CLX_SYM(AllReduce, "ncclAllReduce")
    CLX_SYM(AllGather, "ncclAllGather")
    CLX_SYM(GetErrorString, "ncclGetErrorString")
    CLX_SYM(GroupStart, "ncclGroupStart")
    CLX_SYM(GroupEnd, "ncclGroupEnd")
#undef CLX_SYM
    a.ok = true;
    return a;
}

int nccl_fail(clx_ctx* c, const char* what, int rc) {
    Api& a = api();
    return fail(c, CLX_ERR_NCCL, std::string(what) + ": " + (a.GetErrorString ? a.GetErrorString(rc) : "?"));
}

int allreduce(clx_ctx* c, void* buf, size_t n, int dtype) {
    if (c->world <= 1 || n == 0) return CLX_OK;
    Api& a = api();
    const int rc = a.AllReduce(buf, buf, n, dtype, ncclSum, (ncclComm_t)c->nccl_comm, c->stream);
    if (rc != ncclSuccess) return nccl_fail(c, "ncclAllReduce", rc);
    c->launches++;
    return CLX_OK;
}

}  // namespace

// Collectives issued between group_begin and group_end are fused by NCCL into one launch.
int comm_group_begin(clx_ctx* c) {
    if (c->world <= 1) return CLX_OK;
    const int rc = api().GroupStart();
    return rc == ncclSuccess ? CLX_OK : nccl_fail(c, "ncclGroupStart", rc);
}
int comm_group_end(clx_ctx* c) {
    if (c->world <= 1) return CLX_OK;
    const int rc = api().GroupEnd();
    return rc == ncclSuccess ? CLX_OK : nccl_fail(c, "ncclGroupEnd", rc);
}

int comm_unique_id(void* out128) {
    Api& a = api();
    if (!a.ok) return CLX_ERR_NCCL;
    ncclUniqueId id;
    memset(&id, 0, sizeof(id));
    if (a.GetUniqueId(&id) != ncclSuccess) return CLX_ERR_NCCL;
    memcpy(out128, &id, 128);
    return CLX_OK;
}

int comm_init(clx_ctx* c, int rank, int world, const void* id128) {
    if (world < 1 || rank < 0 || rank >= world) return fail(c, CLX_ERR_ARG, "bad rank/world");
    if (c->loaded) return fail(c, CLX_ERR_STATE, "clx_comm_init must precede clx_load_coo");
    c->rank = rank;
    c->world = world;
    if (world == 1) return CLX_OK;
    Api& a = api();
    if (!a.ok) return fail(c, CLX_ERR_NCCL, a.err);
    ncclUniqueId id;
    memcpy(&id, id128, 128);
    ncclComm_t comm = nullptr;
    CLX_CUDA(c, cudaSetDevice(c->device));
    const int rc = a.CommInitRank(&comm, world, id, rank);
    if (rc != ncclSuccess) return nccl_fail(c, "ncclCommInitRank", rc);
    c->nccl_comm = comm;
    return CLX_OK;
}

void comm_destroy(clx_ctx* c) {
    if (c->nccl_comm) {
        api().CommDestroy((ncclComm_t)c->nccl_comm);
        c->nccl_comm = nullptr;
    }
}

int comm_allreduce_u32(clx_ctx* c, unsigned int* buf, size_t n) { return allreduce(c, buf, n, ncclUint32); }
int comm_allreduce_u64(clx_ctx* c, unsigned long long* buf, size_t n) { return allreduce(c, buf, n, ncclUint64); }
int comm_allreduce_f64(clx_ctx* c, double* buf, size_t n) { return allreduce(c, buf, n, ncclFloat64); }

int comm_allgather_f64(clx_ctx* c, const double* send, double* recv, size_t n_per_rank) {
    if (c->world <= 1 || n_per_rank == 0) return CLX_OK;
    Api& a = api();
    const int rc = a.AllGather(send, recv, n_per_rank, ncclFloat64, (ncclComm_t)c->nccl_comm, c->stream);
    if (rc != ncclSuccess) return nccl_fail(c, "ncclAllGather", rc);
    c->launches++;
    return CLX_OK;
}

int comm_allgather_u64(clx_ctx* c, const unsigned long long* send, unsigned long long* recv, size_t n_per_rank) {
    if (c->world <= 1 || n_per_rank == 0) return CLX_OK;
    Api& a = api();
    const int rc = a.AllGather(send, recv, n_per_rank, ncclUint64, (ncclComm_t)c->nccl_comm, c->stream);
    if (rc != ncclSuccess) return nccl_fail(c, "ncclAllGather", rc);
    c->launches++;
    return CLX_OK;
}

int comm_allgather_u64_host(clx_ctx* c, uint64_t v, std::vector<uint64_t>* out) {
    out->assign((size_t)c->world, 0);
    if (c->world <= 1) {
        (*out)[0] = v;
        return CLX_OK;
    }
    Api& a = api();
    DevBuf<unsigned long long> d;
    CLX_CUDA(c, d.alloc((size_t)c->world + 1));
    CLX_CUDA(c, cudaMemcpyAsync(d.p + c->world, &v, 8, cudaMemcpyHostToDevice, c->stream));
    const int rc = a.AllGather(d.p + c->world, d.p, 1, ncclUint64, (ncclComm_t)c->nccl_comm, c->stream);
    if (rc != ncclSuccess) { d.release(); return nccl_fail(c, "ncclAllGather", rc); }
    CLX_CUDA(c, cudaMemcpyAsync(out->data(), d.p, 8 * (size_t)c->world, cudaMemcpyDeviceToHost, c->stream));
    CLX_CUDA(c, cudaStreamSynchronize(c->stream));
    d.release();
    return CLX_OK;
}

}  // namespace clx

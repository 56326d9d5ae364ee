// Multi-GPU step of the filter behind the C ABI: an opaque per-rank context holding an NCCL
// communicator, the all-reduce of the convergence sums and the final gather of the outputs.
//
// The reference fans chunks out with dask (kafka_test_Py36.py:241-255) and every chunk decides on
// its own when to stop (kafka/linear_kf.py:293).  Here the image is sharded by pixel, one rank per
// GPU, and the only coupling is the scalar of linear_kf.py:293: every rank adds its
// [sum of squares, element count, bad-pixel count] into one all-reduce per iteration and all
// ranks take the same stop decision as one process holding the whole image.
//
// NCCL is not linked: libnccl.so.2 is opened when the first context is made -- the copy the
// process has already loaded (torch ships one) if there is one, the system's otherwise -- so the
// library itself has no NCCL dependency and single-GPU users never touch it.
#include <dlfcn.h>
#include <nccl.h>
#include <string.h>

#include <mutex>

#include "common.cuh"

namespace kafka {
namespace {

struct NcclApi {
    ncclResult_t (*GetUniqueId)(ncclUniqueId*);
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int);
    ncclResult_t (*CommDestroy)(ncclComm_t);
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t);
    ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
    ncclResult_t (*GroupStart)();
    ncclResult_t (*GroupEnd)();
    const char* (*GetErrorString)(ncclResult_t);
    ncclResult_t (*GetVersion)(int*);
    bool ok;
};

NcclApi g_nccl;
std::mutex g_nccl_mutex;

int load_nccl() {
    std::lock_guard<std::mutex> lock(g_nccl_mutex);
    if (g_nccl.ok) return 0;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);       // the copy already in the process, if any
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) {
        set_error("kafka_ctx: cannot open libnccl.so.2 (%s)", dlerror());
        return KAFKA_E_UNSUPPORTED;
    }
#define KAFKA_NCCL_SYM(field, name)                                                   \
    g_nccl.field = reinterpret_cast<decltype(g_nccl.field)>(dlsym(h, name));          \
    if (!g_nccl.field) {                                                              \
        set_error("kafka_ctx: libnccl.so.2 lacks %s", name);                          \
        return KAFKA_E_UNSUPPORTED;                                                   \
    }
    KAFKA_NCCL_SYM(GetUniqueId, "ncclGetUniqueId")
    KAFKA_NCCL_SYM(CommInitRank, "ncclCommInitRank")
    KAFKA_NCCL_SYM(CommDestroy, "ncclCommDestroy")
    KAFKA_NCCL_SYM(AllReduce, "ncclAllReduce")
    KAFKA_NCCL_SYM(Broadcast, "ncclBroadcast")
    KAFKA_NCCL_SYM(GroupStart, "ncclGroupStart")
    KAFKA_NCCL_SYM(GroupEnd, "ncclGroupEnd")
    KAFKA_NCCL_SYM(GetErrorString, "ncclGetErrorString")
    KAFKA_NCCL_SYM(GetVersion, "ncclGetVersion")
#undef KAFKA_NCCL_SYM
    g_nccl.ok = true;
    return 0;
}

#define KAFKA_NCCL(call)                                                                       \
    do {                                                                                       \
        ncclResult_t r_ = (call);                                                              \
        if (r_ != ncclSuccess) {                                                               \
            set_error("%s failed: %s (%s:%d)", #call, g_nccl.GetErrorString(r_), __FILE__, __LINE__); \
            return 1000 + (int)r_;                                                             \
        }                                                                                      \
    } while (0)

}  // namespace
}  // namespace kafka

using namespace kafka;

struct kafka_ctx {
    ncclComm_t comm;
    int rank;
    int n_ranks;
    int device;
};

extern "C" {

int kafka_comm_unique_id(void* id128) {
    KAFKA_CHECK_ARG(id128, "kafka_comm_unique_id: null pointer");
    int rc = load_nccl();
    if (rc) return rc;
    static_assert(sizeof(ncclUniqueId) == KAFKA_COMM_ID_BYTES, "unique id size");
    ncclUniqueId id;
    KAFKA_NCCL(g_nccl.GetUniqueId(&id));
    memcpy(id128, &id, sizeof(id));
    return 0;
}

int kafka_ctx_create(int rank, int n_ranks, const void* id128, kafka_ctx** ctx) {
    KAFKA_CHECK_ARG(ctx && id128, "kafka_ctx_create: null pointer");
    KAFKA_CHECK_ARG(n_ranks >= 1 && rank >= 0 && rank < n_ranks, "kafka_ctx_create: rank %d of %d", rank, n_ranks);
    *ctx = nullptr;
    int rc = load_nccl();
    if (rc) return rc;
    int dev = 0;
    KAFKA_CUDA(cudaGetDevice(&dev));
    ncclUniqueId id;
    memcpy(&id, id128, sizeof(id));
    ncclComm_t comm = nullptr;
    KAFKA_NCCL(g_nccl.CommInitRank(&comm, n_ranks, id, rank));
    kafka_ctx* c = new kafka_ctx;
    c->comm = comm;
    c->rank = rank;
    c->n_ranks = n_ranks;
    c->device = dev;
    *ctx = c;
    return 0;
}

int kafka_ctx_destroy(kafka_ctx* ctx) {
    if (!ctx) return 0;
    ncclResult_t r = g_nccl.ok ? g_nccl.CommDestroy(ctx->comm) : ncclSuccess;
    delete ctx;
    if (r != ncclSuccess) {
        set_error("ncclCommDestroy failed: %s", g_nccl.GetErrorString(r));
        return 1000 + (int)r;
    }
    return 0;
}

int kafka_ctx_info(const kafka_ctx* ctx, int* rank, int* n_ranks, int* nccl_version) {
    KAFKA_CHECK_ARG(ctx, "kafka_ctx_info: null context");
    if (rank) *rank = ctx->rank;
    if (n_ranks) *n_ranks = ctx->n_ranks;
    if (nccl_version) KAFKA_NCCL(g_nccl.GetVersion(nccl_version));
    return 0;
}

int kafka_allreduce_norm(kafka_ctx* ctx, double* sums, int count, void* stream) {
    KAFKA_CHECK_ARG(ctx && sums, "kafka_allreduce_norm: null pointer");
    KAFKA_CHECK_ARG(count >= 1 && count <= 64, "kafka_allreduce_norm: count=%d outside 1..64", count);
    KAFKA_NCCL(g_nccl.AllReduce(sums, sums, (size_t)count, ncclDouble, ncclSum, ctx->comm, (cudaStream_t)stream));
    return 0;
}

int kafka_gather_outputs(kafka_ctx* ctx, const double* send, const int64_t* counts, double* recv, void* stream) {
    KAFKA_CHECK_ARG(ctx && counts && recv, "kafka_gather_outputs: null pointer");
    int64_t off = 0;
    for (int r = 0; r < ctx->n_ranks; ++r) KAFKA_CHECK_ARG(counts[r] >= 0, "kafka_gather_outputs: negative count");
    KAFKA_CHECK_ARG(send || counts[ctx->rank] == 0, "kafka_gather_outputs: send is null");
    // one broadcast per rank inside a group: every rank ends up with the concatenation in rank
    // order, whatever the shard sizes are
    KAFKA_NCCL(g_nccl.GroupStart());
    for (int r = 0; r < ctx->n_ranks; ++r) {
        if (counts[r] > 0) {
            ncclResult_t res = g_nccl.Broadcast(r == ctx->rank ? (const void*)send : (const void*)(recv + off), recv + off,
                                                (size_t)counts[r], ncclDouble, r, ctx->comm, (cudaStream_t)stream);
            if (res != ncclSuccess) {
                g_nccl.GroupEnd();
                set_error("ncclBroadcast failed: %s", g_nccl.GetErrorString(res));
                return 1000 + (int)res;
            }
        }
        off += counts[r];
    }
    KAFKA_NCCL(g_nccl.GroupEnd());
    return 0;
}

}  // extern "C"

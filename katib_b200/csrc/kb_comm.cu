// kb_comm.cu -- multi-GPU plumbing: one process per GPU, NCCL over NVLink / NVSwitch.
// The reference has nothing to replace here (its service is one sequential process, service.go:77-88).
// NCCL is resolved at run time with dlopen so that single-GPU deployments need no libnccl: if the
// process already loaded one (e.g. torch's bundled copy) that copy is used.
#include <dlfcn.h>

#include "kb_internal.cuh"

typedef struct ncclComm *ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;  // 0 = ncclSuccess
typedef int ncclDataType_t;
#define KB_NCCL_INT8 0
#define KB_NCCL_FLOAT64 8
#define KB_NCCL_SUM 0

struct KbNccl {
    void *handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    std::string why;
};

static KbNccl *nccl_api() {
    static KbNccl api;
    static std::once_flag once;
    std::call_once(once, [] {
        // KB_NCCL_LIBRARY names the library to load instead (a deployment with its own NCCL build; the tests use it
        // to reach the "not found" path)
        const char *over = getenv("KB_NCCL_LIBRARY");
        const char *names[] = {over && over[0] ? over : "libnccl.so.2", over && over[0] ? nullptr : "libnccl.so", nullptr};
        for (int i = 0; names[i] && !api.handle; i++) api.handle = dlopen(names[i], RTLD_NOW | RTLD_GLOBAL);
        if (!api.handle) {
            const char *e = dlerror();  // one call: dlerror() clears the state it returns
            api.why = e ? e : "libnccl.so.2 not found";
            return;
        }
        api.GetUniqueId = (decltype(api.GetUniqueId))dlsym(api.handle, "ncclGetUniqueId");
        api.CommInitRank = (decltype(api.CommInitRank))dlsym(api.handle, "ncclCommInitRank");
        api.CommDestroy = (decltype(api.CommDestroy))dlsym(api.handle, "ncclCommDestroy");
        api.AllGather = (decltype(api.AllGather))dlsym(api.handle, "ncclAllGather");
        api.AllReduce = (decltype(api.AllReduce))dlsym(api.handle, "ncclAllReduce");
        api.GetErrorString = (decltype(api.GetErrorString))dlsym(api.handle, "ncclGetErrorString");
        if (!api.GetUniqueId || !api.CommInitRank || !api.CommDestroy || !api.AllGather) {
            api.why = "libnccl is missing a required symbol";
            api.handle = nullptr;
        }
    });
    return &api;
}

struct KbComm {
    ncclComm_t comm = nullptr;
    int world = 1, rank = 0;
};

extern "C" int kb_comm_unique_id(void *out128) {
    if (!out128) return KB_ERR_INVALID;
    KbNccl *a = nccl_api();
    if (!a->handle) return kb_fail(nullptr, KB_ERR_COMM, "kb_comm_unique_id: NCCL unavailable: %s", a->why.c_str());
    ncclUniqueId id;
    ncclResult_t r = a->GetUniqueId(&id);
    if (r) return kb_fail(nullptr, KB_ERR_COMM, "ncclGetUniqueId: %s", a->GetErrorString ? a->GetErrorString(r) : "?");
    memcpy(out128, id.internal, 128);
    return KB_OK;
}

extern "C" int kb_comm_init(kb_ctx *ctx, int world, int rank, const void *unique_id128) {
    if (!ctx) return KB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    KbDeviceGuard g(ctx->device);
    if (world < 1 || rank < 0 || rank >= world || !unique_id128)
        return kb_fail(ctx, KB_ERR_INVALID, "kb_comm_init: bad world/rank/id");
    kb_comm_free(ctx);
    KbComm *c = new KbComm();
    c->world = world; c->rank = rank;
    if (world > 1) {
        KbNccl *a = nccl_api();
        if (!a->handle) { delete c; return kb_fail(ctx, KB_ERR_COMM, "kb_comm_init: NCCL unavailable: %s", a->why.c_str()); }
        ncclUniqueId id;
        memcpy(id.internal, unique_id128, 128);
        ncclResult_t r = a->CommInitRank(&c->comm, world, id, rank);
        if (r) { delete c; return kb_fail(ctx, KB_ERR_COMM, "ncclCommInitRank: %s", a->GetErrorString ? a->GetErrorString(r) : "?"); }
    }
    ctx->comm = c;
    return KB_OK;
}

extern "C" int kb_comm_world(const kb_ctx *ctx, int *world, int *rank) {
    if (!ctx) return KB_ERR_INVALID;
    if (world) *world = ctx->comm ? ctx->comm->world : 1;
    if (rank) *rank = ctx->comm ? ctx->comm->rank : 0;
    return KB_OK;
}

int kb_comm_free(kb_ctx *ctx) {
    if (ctx->comm) {
        if (ctx->comm->comm) nccl_api()->CommDestroy(ctx->comm->comm);
        delete ctx->comm;
        ctx->comm = nullptr;
    }
    return KB_OK;
}

int kb_comm_allgather(kb_ctx *ctx, const void *send, void *recv, size_t bytes, cudaStream_t s) {
    if (!ctx->comm || !ctx->comm->comm) return kb_fail(ctx, KB_ERR_COMM, "kb_comm_allgather: no communicator");
    ncclResult_t r = nccl_api()->AllGather(send, recv, bytes, KB_NCCL_INT8, ctx->comm->comm, s);
    if (r) return kb_fail(ctx, KB_ERR_COMM, "ncclAllGather: %s", nccl_api()->GetErrorString ? nccl_api()->GetErrorString(r) : "?");
    ctx->launches++;
    return KB_OK;
}

// in-place sum of `count` doubles over the ranks (the CMA-ES rank-mu partials, kb_cmaes.cu)
int kb_comm_allreduce_sum_f64(kb_ctx *ctx, double *buf, size_t count, cudaStream_t s) {
    if (!ctx->comm || !ctx->comm->comm) return kb_fail(ctx, KB_ERR_COMM, "kb_comm_allreduce_sum_f64: no communicator");
    if (!nccl_api()->AllReduce) return kb_fail(ctx, KB_ERR_COMM, "kb_comm_allreduce_sum_f64: libnccl has no ncclAllReduce");
    ncclResult_t r = nccl_api()->AllReduce(buf, buf, count, KB_NCCL_FLOAT64, KB_NCCL_SUM, ctx->comm->comm, s);
    if (r) return kb_fail(ctx, KB_ERR_COMM, "ncclAllReduce: %s", nccl_api()->GetErrorString ? nccl_api()->GetErrorString(r) : "?");
    ctx->launches++;
    return KB_OK;
}

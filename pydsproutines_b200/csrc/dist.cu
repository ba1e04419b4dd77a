// dist.cu -- the one exchange step of the shift-sharded multi-GPU search for hosts that are not
// Python: an allgather of the per-rank xc_peak tuples over NCCL (NVLink / NVSwitch inside a box), followed by the same
// deterministic arg max on every rank (largest QF^2, lowest shift on ties).  NCCL is loaded at run time
// (dlopen "libnccl.so.2"), so the library itself has no link dependency on it and a process that already carries
// NCCL (torch) shares the one copy.  One communicator per process (one process per GPU is the model).
#include <dlfcn.h>

#include "internal.cuh"

namespace xc {

// the few NCCL declarations used here (nccl.h, ABI-stable since 2.x)
typedef struct ncclComm* ncclComm_t;
typedef struct {
    char internal[128];
} ncclUniqueId;
typedef int ncclResult_t;       // 0 == ncclSuccess
constexpr int NCCL_INT8 = 0;    // ncclInt8 / ncclChar

struct NcclApi {
    void* handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
};
static NcclApi g_nccl;
static ncclComm_t g_comm = nullptr;
static int g_rank = -1, g_world = 0;
static xc_peak* g_dpeaks = nullptr;     // device staging: [1 + world] peaks

static int32_t nccl_load() {
    if (g_nccl.handle) return XC_OK;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) return fail(XC_ERR_UNSUPPORTED, "NCCL is not available: %s", dlerror());
    NcclApi a;
    a.handle = h;
    a.GetUniqueId = (decltype(a.GetUniqueId))dlsym(h, "ncclGetUniqueId");
    a.CommInitRank = (decltype(a.CommInitRank))dlsym(h, "ncclCommInitRank");
    a.AllGather = (decltype(a.AllGather))dlsym(h, "ncclAllGather");
    a.CommDestroy = (decltype(a.CommDestroy))dlsym(h, "ncclCommDestroy");
    a.GetErrorString = (decltype(a.GetErrorString))dlsym(h, "ncclGetErrorString");
    if (!a.GetUniqueId || !a.CommInitRank || !a.AllGather || !a.CommDestroy)
        return fail(XC_ERR_UNSUPPORTED, "libnccl.so.2 lacks the expected entry points");
    g_nccl = a;
    return XC_OK;
}
static int32_t nccl_fail(const char* what, ncclResult_t r) {
    return fail(XC_ERR_CUDA, "%s failed: %s", what, g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "NCCL error");
}

// largest QF^2, lowest shift on ties; entries with a negative shift (empty shards) never win
int peak_argmax(const xc_peak* all, int world) {
    int best = -1;
    for (int r = 0; r < world; ++r) {
        if (all[r].shift < 0) continue;
        if (best < 0 || all[r].qf2 > all[best].qf2 || (all[r].qf2 == all[best].qf2 && all[r].shift < all[best].shift)) best = r;
    }
    return best;
}

}  // namespace xc

using namespace xc;

extern "C" {

int32_t xc_dist_unique_id(void* out128) {
    XC_API_GUARD;
    if (!out128) return fail(XC_ERR_ARG, "out is NULL");
    XC_TRY(nccl_load());
    ncclUniqueId id;
    const ncclResult_t r = g_nccl.GetUniqueId(&id);
    if (r != 0) return nccl_fail("ncclGetUniqueId", r);
    memcpy(out128, &id, sizeof(id));
    return XC_OK;
}

int32_t xc_dist_init(int32_t rank, int32_t world, const void* unique_id128) {
    XC_API_GUARD;
    if (world < 1 || rank < 0 || rank >= world) return fail(XC_ERR_ARG, "need 0 <= rank < world (rank %d, world %d)", rank, world);
    if (!unique_id128) return fail(XC_ERR_ARG, "unique id is NULL");
    if (g_comm) return fail(XC_ERR_ARG, "xc_dist_init: already initialised (call xc_dist_finalize first)");
    XC_TRY(nccl_load());
    ncclUniqueId id;
    memcpy(&id, unique_id128, sizeof(id));
    const ncclResult_t r = g_nccl.CommInitRank(&g_comm, world, id, rank);   // binds to the current device
    if (r != 0) {
        g_comm = nullptr;
        return nccl_fail("ncclCommInitRank", r);
    }
    g_rank = rank;
    g_world = world;
    XC_CUDA_TRY(cudaMalloc(&g_dpeaks, sizeof(xc_peak) * (size_t)(1 + world)));
    return XC_OK;
}

int32_t xc_dist_allgather_peak(const xc_peak* mine, xc_peak* all, int32_t* best_rank, void* stream) {
    XC_API_GUARD;
    if (!mine || !all) return fail(XC_ERR_ARG, "mine/all is NULL");
    if (!g_comm) return fail(XC_ERR_ARG, "xc_dist_allgather_peak: call xc_dist_init first");
    cudaStream_t st = (cudaStream_t)stream;
    xc_peak p = *mine;
    p.rank = g_rank;
    XC_CUDA_TRY(cudaMemcpyAsync(g_dpeaks, &p, sizeof(p), cudaMemcpyHostToDevice, st));
    const ncclResult_t r = g_nccl.AllGather(g_dpeaks, g_dpeaks + 1, sizeof(xc_peak), NCCL_INT8, g_comm, st);
    if (r != 0) return nccl_fail("ncclAllGather", r);
    XC_CUDA_TRY(cudaMemcpyAsync(all, g_dpeaks + 1, sizeof(xc_peak) * (size_t)g_world, cudaMemcpyDeviceToHost, st));
    XC_CUDA_TRY(cudaStreamSynchronize(st));
    if (best_rank) *best_rank = peak_argmax(all, g_world);
    return XC_OK;
}

int32_t xc_dist_finalize(void) {
    XC_API_GUARD;
    if (g_comm) {
        g_nccl.CommDestroy(g_comm);
        g_comm = nullptr;
    }
    if (g_dpeaks) {
        cudaFree(g_dpeaks);
        g_dpeaks = nullptr;
    }
    g_rank = -1;
    g_world = 0;
    return XC_OK;
}

int32_t xc_peak_argmax(const xc_peak* all, int32_t world) {
    if (!all || world < 1) return -1;
    return peak_argmax(all, world);
}

}  // extern "C"

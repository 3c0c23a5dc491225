// NCCL through dlopen: librwkv.so keeps loading on machines without NCCL, and inside a PyTorch
// process the already-loaded libnccl.so.2 is reused.
#include "tp.h"
#include <dlfcn.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

namespace rb {

namespace {
typedef int ncclResult;                       // ncclSuccess == 0
struct NcclId { char internal[128]; };
constexpr int NCCL_FLOAT32 = 7, NCCL_SUM = 0; // ncclDataType_t / ncclRedOp_t (nccl.h)

struct Api {
    void * h = nullptr;
    ncclResult (*GetUniqueId)(NcclId *) = nullptr;
    ncclResult (*CommInitRank)(void **, int, NcclId, int) = nullptr;
    ncclResult (*CommDestroy)(void *) = nullptr;
    ncclResult (*AllReduce)(const void *, void *, size_t, int, int, void *, cudaStream_t) = nullptr;
    ncclResult (*AllGather)(const void *, void *, size_t, int, void *, cudaStream_t) = nullptr;
    const char * (*GetErrorString)(ncclResult) = nullptr;
    bool ok = false;
};

Api & api() {
    static Api a = [] {
        Api x;
        for (const char * name : {"libnccl.so.2", "libnccl.so"}) { x.h = dlopen(name, RTLD_NOW | RTLD_LOCAL); if (x.h) break; }
        if (!x.h) return x;
        x.GetUniqueId    = (decltype(x.GetUniqueId))    dlsym(x.h, "ncclGetUniqueId");
        x.CommInitRank   = (decltype(x.CommInitRank))   dlsym(x.h, "ncclCommInitRank");
        x.CommDestroy    = (decltype(x.CommDestroy))    dlsym(x.h, "ncclCommDestroy");
        x.AllReduce      = (decltype(x.AllReduce))      dlsym(x.h, "ncclAllReduce");
        x.AllGather      = (decltype(x.AllGather))      dlsym(x.h, "ncclAllGather");
        x.GetErrorString = (decltype(x.GetErrorString)) dlsym(x.h, "ncclGetErrorString");
        x.ok = x.GetUniqueId && x.CommInitRank && x.CommDestroy && x.AllReduce && x.AllGather;
        return x;
    }();
    return a;
}
} // namespace

bool tp_available() { return api().ok; }

bool tp_unique_id(void * out128) {
    if (!api().ok) return false;
    return api().GetUniqueId(reinterpret_cast<NcclId *>(out128)) == 0;
}

bool tp_comm_init(TpInfo & tp, const void * id128) {
    if (!api().ok || tp.world < 2) return false;
    NcclId id = *reinterpret_cast<const NcclId *>(id128);
    const ncclResult r = api().CommInitRank(&tp.comm, tp.world, id, tp.rank);
    if (r != 0) { fprintf(stderr, "ncclCommInitRank failed: %s\n", api().GetErrorString ? api().GetErrorString(r) : "?"); tp.comm = nullptr; return false; }
    return true;
}

void tp_comm_destroy(TpInfo & tp) {
    tp_p2p_destroy(tp);
    if (tp.comm && api().ok) api().CommDestroy(tp.comm);
    tp.comm = nullptr;
}

static bool use_p2p(const TpInfo & tp, size_t n) {
    static const bool off = [] { const char * e = getenv("RWKV_B200_TP_NCCL_ONLY"); return e && e[0] == '1'; }();
    return !off && tp.p2p.connected && n <= TP_NMAX;
}

cudaError_t tp_allreduce(const TpInfo & tp, const float * send, float * recv, size_t n, cudaStream_t st, bool low_latency) {
    if (low_latency && use_p2p(tp, n) && tp_ll_allreduce_usable(tp, n)) return tp_ll_allreduce(tp, send, recv, n, nullptr, st);
    if (use_p2p(tp, n)) return tp_p2p_collective(tp, send, recv, n, false, st);
    if (!tp.comm) return cudaErrorInvalidValue;
    return api().AllReduce(send, recv, n, NCCL_FLOAT32, NCCL_SUM, tp.comm, st) == 0 ? cudaSuccess : cudaErrorUnknown;
}

cudaError_t tp_allgather(const TpInfo & tp, const float * send, float * recv, size_t n, cudaStream_t st) {
    if (use_p2p(tp, n)) return tp_p2p_collective(tp, send, recv, n, true, st);
    if (!tp.comm) return cudaErrorInvalidValue;
    return api().AllGather(send, recv, n, NCCL_FLOAT32, tp.comm, st) == 0 ? cudaSuccess : cudaErrorUnknown;
}

} // namespace rb

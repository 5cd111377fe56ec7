// Queue hand-off of the sharded FIFO loop (replaces the mp.Queue / CUDA-IPC exchange of
// scripts/evaluation/funcs_mp.py:197-236): one NCCL all-gather of the updated half-windows per iteration, issued from inside
// the library on the caller's stream.  NCCL is resolved with dlopen at moca_comm_init (libnccl.so.2: in a PyTorch process this
// is the copy torch already loaded, so both use one NCCL), not linked: a single-GPU user never needs it, and the library keeps
// no link-time dependency besides the CUDA runtime.
#include <dlfcn.h>
#include "common.cuh"

namespace moca {

struct NcclUniqueId { char internal[128]; };           // ncclUniqueId (nccl.h: NCCL_UNIQUE_ID_BYTES = 128)
typedef void* NcclComm;
constexpr int kNcclFloat32 = 7;                        // ncclDataType_t::ncclFloat32

struct NcclApi {
    void* handle = nullptr;
    int (*GetUniqueId)(NcclUniqueId*) = nullptr;
    int (*CommInitRank)(NcclComm*, int, NcclUniqueId, int) = nullptr;
    int (*CommDestroy)(NcclComm) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, NcclComm, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};

static int nccl_api(NcclApi** out) {
    static NcclApi api;
    if (!api.handle) {
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char* nm : names) {
            api.handle = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
            if (api.handle) break;
        }
        MOCA_REQUIRE(api.handle, MOCA_E_NCCL, "moca_comm: libnccl.so.2 not found (%s)", dlerror());
        api.GetUniqueId = (int (*)(NcclUniqueId*))dlsym(api.handle, "ncclGetUniqueId");
        api.CommInitRank = (int (*)(NcclComm*, int, NcclUniqueId, int))dlsym(api.handle, "ncclCommInitRank");
        api.CommDestroy = (int (*)(NcclComm))dlsym(api.handle, "ncclCommDestroy");
        api.AllGather = (int (*)(const void*, void*, size_t, int, NcclComm, cudaStream_t))dlsym(api.handle, "ncclAllGather");
        api.GetErrorString = (const char* (*)(int))dlsym(api.handle, "ncclGetErrorString");
        if (!(api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.AllGather && api.GetErrorString)) {
            api.handle = nullptr;
            MOCA_REQUIRE(false, MOCA_E_NCCL, "moca_comm: libnccl lacks an expected symbol");
        }
    }
    *out = &api;
    return MOCA_OK;
}

#define MOCA_CHECK_NCCL(api, expr)                                                                \
    do {                                                                                          \
        int _r = (expr);                                                                          \
        if (_r != 0) {                                                                            \
            set_error("%s failed: %s", #expr, (api)->GetErrorString(_r));                         \
            return MOCA_E_NCCL;                                                                   \
        }                                                                                         \
    } while (0)

}  // namespace moca

using namespace moca;

// rank 0 calls this and hands the 128 bytes to every rank (any channel: the torch.distributed store, MPI, a file)
extern "C" int32_t moca_comm_unique_id(void* id_out) {
    MOCA_REQUIRE(id_out, MOCA_E_ARG, "moca_comm_unique_id: NULL");
    NcclApi* api = nullptr;
    int rc = nccl_api(&api);
    if (rc) return rc;
    MOCA_CHECK_NCCL(api, api->GetUniqueId((NcclUniqueId*)id_out));
    return MOCA_OK;
}

extern "C" int32_t moca_comm_init(moca_ctx* ctx, const void* id, int32_t nranks, int32_t rank) {
    MOCA_REQUIRE(ctx && id && nranks >= 1 && rank >= 0 && rank < nranks, MOCA_E_ARG, "moca_comm_init: bad argument");
    MOCA_REQUIRE(!ctx->c.nccl_comm, MOCA_E_STATE, "moca_comm_init: this context already has a communicator");
    NcclApi* api = nullptr;
    int rc = nccl_api(&api);
    if (rc) return rc;
    int prev = 0;
    MOCA_CHECK_CUDA(cudaGetDevice(&prev));
    MOCA_CHECK_CUDA(cudaSetDevice(ctx->c.device));
    NcclUniqueId uid;
    memcpy(&uid, id, sizeof(uid));
    NcclComm comm = nullptr;
    const int r = api->CommInitRank(&comm, nranks, uid, rank);
    cudaSetDevice(prev);
    if (r != 0) {
        set_error("ncclCommInitRank failed: %s", api->GetErrorString(r));
        return MOCA_E_NCCL;
    }
    ctx->c.nccl_comm = comm; ctx->c.nranks = nranks; ctx->c.rank = rank;
    return MOCA_OK;
}

// every rank contributes `count` floats (its updated half-windows, packed by moca_window_scatter); gathered holds nranks * count
extern "C" int32_t moca_queue_handoff(moca_ctx* ctx, const float* packed, float* gathered, int64_t count, void* stream) {
    MOCA_REQUIRE(ctx && packed && gathered && count > 0, MOCA_E_ARG, "moca_queue_handoff: bad argument");
    MOCA_REQUIRE(ctx->c.nccl_comm, MOCA_E_STATE, "moca_queue_handoff: call moca_comm_init first");
    NcclApi* api = nullptr;
    int rc = nccl_api(&api);
    if (rc) return rc;
    MOCA_CHECK_NCCL(api, api->AllGather(packed, gathered, (size_t)count, kNcclFloat32, (NcclComm)ctx->c.nccl_comm, (cudaStream_t)stream));
    return MOCA_OK;
}

extern "C" int32_t moca_comm_destroy(moca_ctx* ctx) {
    if (!ctx || !ctx->c.nccl_comm) return MOCA_OK;
    NcclApi* api = nullptr;
    int rc = nccl_api(&api);
    if (rc) return rc;
    MOCA_CHECK_NCCL(api, api->CommDestroy((NcclComm)ctx->c.nccl_comm));
    ctx->c.nccl_comm = nullptr;
    return MOCA_OK;
}

extern "C" int32_t moca_comm_size(moca_ctx* ctx) { return ctx && ctx->c.nccl_comm ? ctx->c.nranks : 0; }

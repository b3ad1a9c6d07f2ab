// NCCL baseline collective, bound at run time with dlopen/dlsym so that the
// library uses the SAME libnccl.so.2 the host process (torch) already loaded -
// the box also has a second, older system copy, and mixing the two is a hazard.
#include <dlfcn.h>

#include <cstdlib>
#include <cstring>
#include <mutex>

#include "internal.h"

namespace b200ddp {

namespace {

// Minimal mirror of nccl.h (stable ABI since 2.x).
typedef struct ncclComm *ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;
constexpr int kNcclSum = 0, kNcclFloat32 = 7, kNcclBfloat16 = 9;

struct Api {
    void *handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GetVersion)(int *) = nullptr;
};

Api g_api;
std::once_flag g_once;
char g_load_err[256] = "";

void load() {
    // 1. whatever the process already has mapped (torch's pip nvidia-nccl copy)
    void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
    // 2. an explicit path, 3. the loader's search path
    if (!h) {
        const char *p = getenv("B200DDP_NCCL_LIB");
        if (p && *p) h = dlopen(p, RTLD_NOW | RTLD_GLOBAL);
    }
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) {
        const char *e = dlerror();
        strncpy(g_load_err, e ? e : "dlopen(libnccl.so.2) failed", sizeof(g_load_err) - 1);
        return;
    }
    g_api.handle = h;
#define B2_SYM(field, name)                                              \
    do {                                                                 \
        *(void **)(&g_api.field) = dlsym(h, name);                       \
        if (!g_api.field) {                                              \
            strncpy(g_load_err, "missing symbol " name, sizeof(g_load_err) - 1); \
            g_api.handle = nullptr;                                      \
            return;                                                      \
        }                                                                \
    } while (0)
    B2_SYM(GetUniqueId, "ncclGetUniqueId");
    B2_SYM(CommInitRank, "ncclCommInitRank");
    B2_SYM(CommDestroy, "ncclCommDestroy");
    B2_SYM(AllReduce, "ncclAllReduce");
    B2_SYM(GetErrorString, "ncclGetErrorString");
    B2_SYM(GetVersion, "ncclGetVersion");
#undef B2_SYM
}

int ensure() {
    std::call_once(g_once, load);
    B2_REQUIRE(g_api.handle != nullptr, B200DDP_ENCCL, "cannot load NCCL: %s", g_load_err);
    return B200DDP_OK;
}

#define B2_CHECK_NCCL(expr)                                                            \
    do {                                                                               \
        ncclResult_t _r = (expr);                                                      \
        if (_r != 0) {                                                                 \
            set_error("%s failed: %s", #expr, g_api.GetErrorString(_r));               \
            return B200DDP_ENCCL;                                                      \
        }                                                                              \
    } while (0)

}  // namespace

int nccl_unique_id(void *out128) {
    if (int rc = ensure()) return rc;
    static_assert(sizeof(ncclUniqueId) == B200DDP_NCCL_UID_BYTES, "uid size");
    ncclUniqueId id;
    B2_CHECK_NCCL(g_api.GetUniqueId(&id));
    memcpy(out128, &id, sizeof(id));
    return B200DDP_OK;
}

int nccl_init(b200ddp_ctx *ctx, const void *uid) {
    if (int rc = ensure()) return rc;
    B2_REQUIRE(ctx->nccl_comm == nullptr, B200DDP_ESTATE, "NCCL communicator already initialised");
    ncclUniqueId id;
    memcpy(&id, uid, sizeof(id));
    ncclComm_t comm = nullptr;
    B2_CHECK_NCCL(g_api.CommInitRank(&comm, ctx->world, id, ctx->rank));
    ctx->nccl_comm = comm;
    return B200DDP_OK;
}

int nccl_allreduce(b200ddp_ctx *ctx, void *buf, int64_t nelem, int dtype, cudaStream_t s) {
    if (int rc = ensure()) return rc;
    B2_CHECK_NCCL(g_api.AllReduce(buf, buf, (size_t)nelem, dtype == B200DDP_BF16 ? kNcclBfloat16 : kNcclFloat32,
                                  kNcclSum, (ncclComm_t)ctx->nccl_comm, s));
    return B200DDP_OK;
}

void nccl_destroy(b200ddp_ctx *ctx) {
    if (ctx->nccl_comm && g_api.handle) g_api.CommDestroy((ncclComm_t)ctx->nccl_comm);
    ctx->nccl_comm = nullptr;
}

}  // namespace b200ddp

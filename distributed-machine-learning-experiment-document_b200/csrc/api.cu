// extern "C" surface of libb200ddp.so: contexts, plans, argument checking and
// dispatch.  Declarations and the reference lines each entry point replaces
// are in include/b200ddp.h.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <new>

#include "internal.h"

namespace b200ddp {

static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

static size_t dtype_size(int dtype) { return dtype == B200DDP_BF16 ? 2 : 4; }

// The caller's pointer arrays are read on the host at launch time and copied into the
// kernel parameters; only their shape is checked here.
int check_ptrs(const b200ddp_plan *plan, PtrKind kind, const void *const *host_ptrs) {
    if (plan->T == 0) return B200DDP_OK;
    B2_REQUIRE(host_ptrs != nullptr, B200DDP_EINVAL, "pointer table %d is NULL", (int)kind);
    for (int t = 0; t < plan->T; ++t)
        B2_REQUIRE(host_ptrs[t] != nullptr || plan->meta[t].numel == 0, B200DDP_EINVAL,
                   "tensor %d of pointer table %d is NULL", t, (int)kind);
    return B200DDP_OK;
}

int sgd_scalars(const b200ddp_sgd_args *a, int div_mode, SgdScalars *o) {
    memset(o, 0, sizeof(*o));
    B2_REQUIRE(a != nullptr, B200DDP_EINVAL, "sgd args is NULL");
    B2_REQUIRE(std::isfinite(a->lr) && std::isfinite(a->momentum) && std::isfinite(a->dampening) &&
                   std::isfinite(a->weight_decay) && std::isfinite(a->grad_divisor),
               B200DDP_EINVAL, "non-finite SGD hyper-parameter");
    B2_REQUIRE(a->grad_divisor != 0.0, B200DDP_EINVAL, "grad_divisor must not be 0");
    B2_REQUIRE(!a->nesterov || (a->momentum > 0 && a->dampening == 0), B200DDP_EINVAL,
               "Nesterov momentum requires a momentum and zero dampening");
    o->neg_lr = (float)(-a->lr);
    o->momentum = (float)a->momentum;
    o->one_minus_damp = (float)(1.0 - a->dampening);
    o->weight_decay = (float)a->weight_decay;
    o->dv = make_divisor((float)a->grad_divisor, div_mode);
    o->rule = kRuleSgd;
    o->nesterov = a->nesterov != 0;
    o->maximize = a->maximize != 0;
    o->first_step = a->first_step != 0;
    o->has_momentum = a->momentum != 0.0;
    o->has_wd = a->weight_decay != 0.0;
    return B200DDP_OK;
}

int opt_scalars(const b200ddp_opt_args *a, int div_mode, SgdScalars *o) {
    memset(o, 0, sizeof(*o));
    B2_REQUIRE(a != nullptr, B200DDP_EINVAL, "optimizer args is NULL");
    B2_REQUIRE(a->rule == B200DDP_RULE_ADAM || a->rule == B200DDP_RULE_GD, B200DDP_EINVAL, "bad update rule %d", a->rule);
    B2_REQUIRE(std::isfinite(a->lr) && std::isfinite(a->grad_divisor) && a->grad_divisor != 0.0, B200DDP_EINVAL,
               "bad lr or grad_divisor");
    o->dv = make_divisor((float)a->grad_divisor, div_mode);
    o->lr = (float)a->lr;
    o->neg_lr = (float)(-a->lr);
    if (a->rule == B200DDP_RULE_ADAM) {
        B2_REQUIRE(std::isfinite(a->beta1) && std::isfinite(a->beta2) && std::isfinite(a->eps), B200DDP_EINVAL,
                   "non-finite Adam hyper-parameter");
        o->rule = kRuleAdam;
        // Python evaluates (1 - beta) in double before the product with an fp32 tensor casts it
        o->beta1 = (float)a->beta1;
        o->one_minus_beta1 = (float)(1.0 - a->beta1);
        o->beta2 = (float)a->beta2;
        o->one_minus_beta2 = (float)(1.0 - a->beta2);
        o->eps = (float)a->eps;
        o->has_momentum = 1;
    } else {
        o->rule = kRuleGd;
    }
    return B200DDP_OK;
}

constexpr int kPdlDefault = 1;

// value checks only; the caller holds whatever lock it needs
static int set_option(b200ddp_ctx *ctx, int option, long long v) {
    switch (option) {
        case B200DDP_OPT_DIV_MODE:
            B2_REQUIRE(v == B200DDP_DIV_IEEE || v == B200DDP_DIV_RECIPROCAL, B200DDP_EINVAL, "bad division mode %lld", v);
            ctx->div_mode = (int)v;
            return B200DDP_OK;
        case B200DDP_OPT_PDL: ctx->pdl = v != 0; return B200DDP_OK;
        case B200DDP_OPT_NVLS_MIN_BYTES: ctx->nvls_min_bytes = v < -2 ? -2 : v; return B200DDP_OK;
        case B200DDP_OPT_ONESHOT_MAX_BYTES: ctx->oneshot_max_bytes = v < -2 ? -2 : v; return B200DDP_OK;
        case B200DDP_OPT_PEER_TIMEOUT_MS: ctx->peer_timeout_ms = v < 0 ? 0 : v; return B200DDP_OK;
        case B200DDP_OPT_P2P_BLOCKS:
            B2_REQUIRE(v >= 0 && v <= kMaxBlocksP2P, B200DDP_EINVAL, "block cap %lld not in [0, %d]", v, kMaxBlocksP2P);
            ctx->p2p_blocks = (int)v;
            return B200DDP_OK;
        case B200DDP_OPT_NVLS_DEPTH:
            B2_REQUIRE(v == 0 || v == 1 || v == 2 || v == 4 || v == 8, B200DDP_EINVAL, "NVLS depth must be 1, 2, 4 or 8");
            ctx->nvls_depth = (int)v;
            return B200DDP_OK;
        case B200DDP_OPT_COPY_UNROLL:
            B2_REQUIRE(v == 2 || v == 4 || v == 8, B200DDP_EINVAL, "copy unroll must be 2, 4 or 8");
            ctx->copy_unroll = (int)v;
            return B200DDP_OK;
        case B200DDP_OPT_NVLS_BLOCKS:
            B2_REQUIRE(v >= 0 && v <= kMaxBlocksP2P, B200DDP_EINVAL, "block cap %lld not in [0, %d]", v, kMaxBlocksP2P);
            ctx->nvls_blocks = (int)v;
            return B200DDP_OK;
        case B200DDP_OPT_HBM_BLOCKS_PER_SM:
            B2_REQUIRE(v >= 0 && v <= 32, B200DDP_EINVAL, "blocks per SM %lld not in [0, 32]", v);
            ctx->hbm_blocks_per_sm = (int)v;
            return B200DDP_OK;
        case B200DDP_OPT_TRACE: {
            if (v != 0 && ctx->trace == nullptr) {
                B2_CHECK_CUDA(cudaSetDevice(ctx->device));
                B2_CHECK_CUDA(cudaMalloc((void **)&ctx->trace, sizeof(unsigned long long) * 4 * kMaxBlocksP2P));
                B2_CHECK_CUDA(cudaMemset(ctx->trace, 0, sizeof(unsigned long long) * 4 * kMaxBlocksP2P));
            } else if (v == 0 && ctx->trace != nullptr) {
                B2_CHECK_CUDA(cudaSetDevice(ctx->device));
                cudaDeviceSynchronize();
                cudaFree(ctx->trace);
                ctx->trace = nullptr;
            }
            return B200DDP_OK;
        }
        default:
            set_error("unknown option %d", option);
            return B200DDP_EINVAL;
    }
}

static int check_plan(b200ddp_ctx *ctx, b200ddp_plan *plan) {
    B2_REQUIRE(ctx != nullptr && plan != nullptr, B200DDP_EINVAL, "ctx or plan is NULL");
    B2_REQUIRE(plan->ctx == ctx, B200DDP_EINVAL, "plan belongs to another context");
    return B200DDP_OK;
}

static int in_staging(const b200ddp_ctx *ctx, size_t off, int64_t nelem, int dtype) {
    B2_REQUIRE(dtype == B200DDP_F32 || dtype == B200DDP_BF16, B200DDP_EINVAL, "bad dtype %d", dtype);
    B2_REQUIRE(nelem >= 0, B200DDP_EINVAL, "negative element count");
    B2_REQUIRE(off % 256 == 0, B200DDP_EINVAL, "bucket offset %zu is not 256-byte aligned", off);
    B2_REQUIRE((nelem * (int64_t)dtype_size(dtype)) % 16 == 0, B200DDP_EINVAL,
               "bucket of %lld elements is not a whole number of 16-byte vectors", (long long)nelem);
    B2_REQUIRE(off + (size_t)nelem * dtype_size(dtype) <= ctx->staging_bytes, B200DDP_ETOOBIG,
               "bucket [%zu, +%lld x %zu B) exceeds the %zu-byte staging area", off, (long long)nelem,
               dtype_size(dtype), ctx->staging_bytes);
    return B200DDP_OK;
}

}  // namespace b200ddp

using namespace b200ddp;

extern "C" {

const char *b200ddp_last_error(void) { return g_err; }
int b200ddp_version(void) { return 100; }

int b200ddp_create(int rank, int world, int device, size_t staging_bytes, b200ddp_ctx **out) {
    B2_REQUIRE(out != nullptr, B200DDP_EINVAL, "out is NULL");
    *out = nullptr;
    B2_REQUIRE(world >= 1 && world <= B200DDP_MAX_WORLD, B200DDP_EINVAL, "world size %d not in [1, %d]", world,
               B200DDP_MAX_WORLD);
    B2_REQUIRE(rank >= 0 && rank < world, B200DDP_EINVAL, "rank %d not in [0, %d)", rank, world);
    int ndev = 0;
    B2_CHECK_CUDA(cudaGetDeviceCount(&ndev));
    B2_REQUIRE(device >= 0 && device < ndev, B200DDP_EINVAL, "device %d not in [0, %d)", device, ndev);
    B2_CHECK_CUDA(cudaSetDevice(device));
    b200ddp_ctx *ctx = new (std::nothrow) b200ddp_ctx();
    B2_REQUIRE(ctx != nullptr, B200DDP_ESTATE, "out of host memory");
    ctx->rank = rank;
    ctx->world = world;
    ctx->device = device;
    ctx->staging_bytes = (staging_bytes + 255) / 256 * 256;
    int sms = 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) == cudaSuccess && sms > 0)
        ctx->sm_count = sms;
    // environment defaults of the per-context options (b200ddp_set_option overrides)
    auto env_ll = [](const char *name, long long dflt) {
        const char *v = getenv(name);
        return (v && *v) ? atoll(v) : dflt;
    };
    if (const char *m = getenv("B200DDP_DIV_MODE"))
        ctx->div_mode = (m[0] == 'i' || m[0] == 'I' || m[0] == '0') ? B200DDP_DIV_IEEE : B200DDP_DIV_RECIPROCAL;
    set_option(ctx, B200DDP_OPT_COPY_UNROLL, env_ll("B200DDP_COPY_UNROLL", ctx->copy_unroll));
    set_option(ctx, B200DDP_OPT_PDL, env_ll("B200DDP_PDL", kPdlDefault));
    set_option(ctx, B200DDP_OPT_NVLS_MIN_BYTES, env_ll("B200DDP_NVLS_MIN_BYTES", -2));
    set_option(ctx, B200DDP_OPT_ONESHOT_MAX_BYTES, env_ll("B200DDP_ONESHOT_MAX_BYTES", -2));
    set_option(ctx, B200DDP_OPT_PEER_TIMEOUT_MS, env_ll("B200DDP_PEER_TIMEOUT_MS", 0));
    set_option(ctx, B200DDP_OPT_P2P_BLOCKS, env_ll("B200DDP_P2P_BLOCKS", 0));
    set_option(ctx, B200DDP_OPT_NVLS_DEPTH, env_ll("B200DDP_NVLS_DEPTH", 0));
    set_option(ctx, B200DDP_OPT_NVLS_BLOCKS, env_ll("B200DDP_NVLS_BLOCKS", 0));
    const size_t flag_bytes = sizeof(uint32_t) * kMaxBlocksP2P * B200DDP_MAX_WORLD;
    cudaError_t e = cudaMalloc((void **)&ctx->staging, 2 * ctx->staging_bytes + 256);
    if (e == cudaSuccess) e = cudaMalloc((void **)&ctx->flags, flag_bytes);
    if (e == cudaSuccess) e = cudaMalloc((void **)&ctx->block_gen, sizeof(uint32_t) * kMaxBlocksP2P);
    if (e == cudaSuccess) e = cudaMemset(ctx->staging, 0, 2 * ctx->staging_bytes + 256);
    if (e == cudaSuccess) e = cudaMemset(ctx->flags, 0, flag_bytes);
    if (e == cudaSuccess) e = cudaMemset(ctx->block_gen, 0, sizeof(uint32_t) * kMaxBlocksP2P);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
        set_error("allocating %zu bytes of staging on device %d failed: %s", 2 * ctx->staging_bytes, device,
                  cudaGetErrorString(e));
        cudaFree(ctx->staging);
        cudaFree(ctx->flags);
        cudaFree(ctx->block_gen);
        delete ctx;
        return B200DDP_ECUDA;
    }
    for (int r = 0; r < B200DDP_MAX_WORLD; ++r) {
        ctx->peers.staging[r] = nullptr;
        ctx->peers.flags[r] = nullptr;
    }
    ctx->peers.staging[rank] = ctx->staging;
    ctx->peers.flags[rank] = ctx->flags;
    ctx->peers_ready = (world == 1);
    *out = ctx;
    return B200DDP_OK;
}

static void close_peer_mappings(b200ddp_ctx *ctx) {
    if (ctx->peers_ipc) {
        for (int r = 0; r < ctx->world; ++r) {
            if (r == ctx->rank) continue;
            if (ctx->peers.staging[r] && !ctx->staging_external) cudaIpcCloseMemHandle(ctx->peers.staging[r]);
            if (ctx->peers.flags[r]) cudaIpcCloseMemHandle(ctx->peers.flags[r]);
            ctx->peers.staging[r] = nullptr;
            ctx->peers.flags[r] = nullptr;
        }
        ctx->peers_ipc = false;
        ctx->peers_ready = false;
    }
}

int b200ddp_close_peers(b200ddp_ctx *ctx) {
    B2_REQUIRE(ctx != nullptr, B200DDP_EINVAL, "ctx is NULL");
    std::lock_guard<std::mutex> lk(ctx->mu);
    B2_CHECK_CUDA(cudaSetDevice(ctx->device));
    B2_CHECK_CUDA(cudaDeviceSynchronize());
    close_peer_mappings(ctx);
    cudaGetLastError();
    return B200DDP_OK;
}

int b200ddp_free_retired(b200ddp_ctx *ctx) {
    B2_REQUIRE(ctx != nullptr, B200DDP_EINVAL, "ctx is NULL");
    std::lock_guard<std::mutex> lk(ctx->mu);
    if (ctx->retired_staging) {
        B2_CHECK_CUDA(cudaSetDevice(ctx->device));
        cudaFree(ctx->retired_staging);
        ctx->retired_staging = nullptr;
        cudaGetLastError();
    }
    return B200DDP_OK;
}

int b200ddp_destroy(b200ddp_ctx *ctx) {
    if (!ctx) return B200DDP_OK;
    cudaSetDevice(ctx->device);
    cudaDeviceSynchronize();
    nccl_destroy(ctx);
    close_peer_mappings(ctx);
    cudaFree(ctx->retired_staging);
    cudaFree(ctx->sampler_scratch);
    cudaFree(ctx->trace);
    if (!ctx->staging_external) cudaFree(ctx->staging);
    cudaFree(ctx->flags);
    cudaFree(ctx->block_gen);
    cudaGetLastError();
    delete ctx;
    return B200DDP_OK;
}

int b200ddp_set_option(b200ddp_ctx *ctx, int option, long long value) {
    B2_REQUIRE(ctx != nullptr, B200DDP_EINVAL, "ctx is NULL");
    std::lock_guard<std::mutex> lk(ctx->mu);
    return set_option(ctx, option, value);
}

int b200ddp_get_option(const b200ddp_ctx *ctx, int option, long long *value) {
    B2_REQUIRE(ctx != nullptr && value != nullptr, B200DDP_EINVAL, "NULL argument");
    switch (option) {
        case B200DDP_OPT_DIV_MODE: *value = ctx->div_mode; break;
        case B200DDP_OPT_PDL: *value = ctx->pdl ? 1 : 0; break;
        case B200DDP_OPT_NVLS_MIN_BYTES: *value = ctx->nvls_min_bytes; break;
        case B200DDP_OPT_ONESHOT_MAX_BYTES: *value = ctx->oneshot_max_bytes; break;
        case B200DDP_OPT_PEER_TIMEOUT_MS: *value = ctx->peer_timeout_ms; break;
        case B200DDP_OPT_P2P_BLOCKS: *value = ctx->p2p_blocks; break;
        case B200DDP_OPT_NVLS_DEPTH: *value = ctx->nvls_depth; break;
        case B200DDP_OPT_COPY_UNROLL: *value = ctx->copy_unroll; break;
        case B200DDP_OPT_TRACE: *value = ctx->trace != nullptr; break;
        case B200DDP_OPT_HBM_BLOCKS_PER_SM: *value = ctx->hbm_blocks_per_sm; break;
        case B200DDP_OPT_NVLS_BLOCKS: *value = ctx->nvls_blocks; break;
        default: set_error("unknown option %d", option); return B200DDP_EINVAL;
    }
    return B200DDP_OK;
}

int b200ddp_trace_read(b200ddp_ctx *ctx, uint64_t *out, int nblocks) {
    B2_REQUIRE(ctx != nullptr && out != nullptr, B200DDP_EINVAL, "NULL argument");
    B2_REQUIRE(ctx->trace != nullptr, B200DDP_ESTATE, "tracing is off (B200DDP_OPT_TRACE)");
    B2_REQUIRE(nblocks >= 0 && nblocks <= kMaxBlocksP2P, B200DDP_EINVAL, "bad block count");
    B2_CHECK_CUDA(cudaSetDevice(ctx->device));
    B2_CHECK_CUDA(cudaDeviceSynchronize());
    B2_CHECK_CUDA(cudaMemcpy(out, ctx->trace, sizeof(uint64_t) * 4 * (size_t)nblocks, cudaMemcpyDeviceToHost));
    return B200DDP_OK;
}

int b200ddp_staging(b200ddp_ctx *ctx, void **ptr, size_t *bytes) {
    B2_REQUIRE(ctx != nullptr, B200DDP_EINVAL, "ctx is NULL");
    if (ptr) *ptr = ctx->staging;
    if (bytes) *bytes = ctx->staging_bytes;
    return B200DDP_OK;
}

int b200ddp_flags(b200ddp_ctx *ctx, void **ptr) {
    B2_REQUIRE(ctx != nullptr && ptr != nullptr, B200DDP_EINVAL, "NULL argument");
    *ptr = ctx->flags;
    return B200DDP_OK;
}

int b200ddp_peer_staging(b200ddp_ctx *ctx, int peer, void **ptr) {
    B2_REQUIRE(ctx != nullptr && ptr != nullptr, B200DDP_EINVAL, "NULL argument");
    B2_REQUIRE(peer >= 0 && peer < ctx->world, B200DDP_EINVAL, "peer %d not in [0, %d)", peer, ctx->world);
    B2_REQUIRE(ctx->peers_ready && ctx->peers.staging[peer] != nullptr, B200DDP_ESTATE, "peer memory not opened");
    *ptr = ctx->peers.staging[peer];
    return B200DDP_OK;
}

int b200ddp_nccl_unique_id(void *out_uid) {
    B2_REQUIRE(out_uid != nullptr, B200DDP_EINVAL, "out is NULL");
    return nccl_unique_id(out_uid);
}

int b200ddp_nccl_init(b200ddp_ctx *ctx, const void *uid) {
    B2_REQUIRE(ctx != nullptr && uid != nullptr, B200DDP_EINVAL, "NULL argument");
    std::lock_guard<std::mutex> lk(ctx->mu);
    B2_CHECK_CUDA(cudaSetDevice(ctx->device));
    return nccl_init(ctx, uid);
}

int b200ddp_ipc_export(b200ddp_ctx *ctx, void *out) {
    B2_REQUIRE(ctx != nullptr && out != nullptr, B200DDP_EINVAL, "NULL argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    cudaIpcMemHandle_t h[2];
    B2_CHECK_CUDA(cudaSetDevice(ctx->device));
    B2_CHECK_CUDA(cudaIpcGetMemHandle(&h[0], ctx->staging));
    B2_CHECK_CUDA(cudaIpcGetMemHandle(&h[1], ctx->flags));
    memcpy(out, h, sizeof(h));
    return B200DDP_OK;
}

int b200ddp_ipc_open_peers(b200ddp_ctx *ctx, const void *handles) {
    B2_REQUIRE(ctx != nullptr && handles != nullptr, B200DDP_EINVAL, "NULL argument");
    std::lock_guard<std::mutex> lk(ctx->mu);
    B2_REQUIRE(!ctx->peers_ready || ctx->world == 1, B200DDP_ESTATE, "peers already set");
    B2_CHECK_CUDA(cudaSetDevice(ctx->device));
    const char *hp = (const char *)handles;
    for (int r = 0; r < ctx->world; ++r) {
        if (r == ctx->rank) continue;
        cudaIpcMemHandle_t h[2];
        memcpy(h, hp + (size_t)r * B200DDP_IPC_BYTES, sizeof(h));
        void *ps = nullptr, *pf = nullptr;
        B2_CHECK_CUDA(cudaIpcOpenMemHandle(&ps, h[0], cudaIpcMemLazyEnablePeerAccess));
        B2_CHECK_CUDA(cudaIpcOpenMemHandle(&pf, h[1], cudaIpcMemLazyEnablePeerAccess));
        ctx->peers.staging[r] = (char *)ps;
        ctx->peers.flags[r] = (uint32_t *)pf;
    }
    ctx->peers_ipc = true;
    ctx->peers_ready = true;
    return B200DDP_OK;
}

int b200ddp_set_peers_direct(b200ddp_ctx *ctx, void *const *staging_ptrs, void *const *flag_ptrs) {
    B2_REQUIRE(ctx != nullptr && staging_ptrs != nullptr && flag_ptrs != nullptr, B200DDP_EINVAL, "NULL argument");
    std::lock_guard<std::mutex> lk(ctx->mu);
    B2_CHECK_CUDA(cudaSetDevice(ctx->device));
    for (int r = 0; r < ctx->world; ++r) {
        B2_REQUIRE(staging_ptrs[r] != nullptr && flag_ptrs[r] != nullptr, B200DDP_EINVAL, "peer %d pointer is NULL", r);
        if (r == ctx->rank) {
            B2_REQUIRE(staging_ptrs[r] == ctx->staging && flag_ptrs[r] == ctx->flags, B200DDP_EINVAL,
                       "entry %d must be this context's own staging and flags", r);
            continue;
        }
        cudaPointerAttributes at;
        B2_CHECK_CUDA(cudaPointerGetAttributes(&at, staging_ptrs[r]));
        if (at.device != ctx->device) {
            cudaError_t e = cudaDeviceEnablePeerAccess(at.device, 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) {
                cudaGetLastError();
            } else if (e != cudaSuccess) {
                set_error("cudaDeviceEnablePeerAccess(%d -> %d) failed: %s", ctx->device, at.device,
                          cudaGetErrorString(e));
                return B200DDP_ECUDA;
            }
        }
        ctx->peers.staging[r] = (char *)staging_ptrs[r];
        ctx->peers.flags[r] = (uint32_t *)flag_ptrs[r];
    }
    ctx->peers_ready = true;
    return B200DDP_OK;
}

int b200ddp_adopt_symmetric(b200ddp_ctx *ctx, void *const *peer_ptrs, void *mc_ptr, size_t bytes) {
    B2_REQUIRE(ctx != nullptr && peer_ptrs != nullptr, B200DDP_EINVAL, "NULL argument");
    std::lock_guard<std::mutex> lk(ctx->mu);
    B2_REQUIRE(ctx->world > 1, B200DDP_ESTATE, "symmetric memory is for more than one rank");
    B2_REQUIRE(!ctx->staging_external, B200DDP_ESTATE, "symmetric memory already adopted");
    B2_REQUIRE(bytes >= 2 * ctx->staging_bytes, B200DDP_EINVAL, "symmetric buffer of %zu bytes is smaller than 2 x %zu",
               bytes, ctx->staging_bytes);
    for (int r = 0; r < ctx->world; ++r)
        B2_REQUIRE(peer_ptrs[r] != nullptr && ((uintptr_t)peer_ptrs[r] & 255) == 0, B200DDP_EINVAL,
                   "peer buffer %d is NULL or not 256-byte aligned", r);
    B2_REQUIRE(((uintptr_t)mc_ptr & 255) == 0, B200DDP_EINVAL, "multicast address is not 256-byte aligned");
    B2_CHECK_CUDA(cudaSetDevice(ctx->device));
    B2_CHECK_CUDA(cudaDeviceSynchronize());
    if (ctx->peers_ipc)
        for (int r = 0; r < ctx->world; ++r)
            if (r != ctx->rank && ctx->peers.staging[r]) cudaIpcCloseMemHandle(ctx->peers.staging[r]);
    // The old staging area was exported over CUDA IPC and a peer may still have it mapped: it is
    // only retired here and freed by b200ddp_free_retired, which the caller invokes after a
    // cross-rank barrier (every rank has closed its mappings by then).
    ctx->retired_staging = ctx->staging;
    ctx->staging_external = true;
    ctx->staging = (char *)peer_ptrs[ctx->rank];
    for (int r = 0; r < ctx->world; ++r) ctx->peers.staging[r] = (char *)peer_ptrs[r];
    ctx->mc_ptr = (char *)mc_ptr;
    return B200DDP_OK;
}

// ---------------------------------------------------------------------- plan
int b200ddp_plan_create(b200ddp_ctx *ctx, int T, const int64_t *numels, int bucket_dtype, b200ddp_plan **out) {
    B2_REQUIRE(ctx != nullptr && out != nullptr, B200DDP_EINVAL, "NULL argument");
    *out = nullptr;
    B2_REQUIRE(T >= 0 && (T == 0 || numels != nullptr), B200DDP_EINVAL, "bad tensor list");
    B2_REQUIRE(bucket_dtype == B200DDP_F32 || bucket_dtype == B200DDP_BF16, B200DDP_EINVAL, "bad dtype %d",
               bucket_dtype);
    B2_CHECK_CUDA(cudaSetDevice(ctx->device));
    b200ddp_plan *p = new (std::nothrow) b200ddp_plan();
    B2_REQUIRE(p != nullptr, B200DDP_ESTATE, "out of host memory");
    p->ctx = ctx;
    p->device = ctx->device;
    p->T = T;
    p->dtype = bucket_dtype;
    p->copy_unroll = ctx->copy_unroll;
    const int64_t copy_chunk = (int64_t)kThreads * kVec * p->copy_unroll;
    int64_t off = 0;
    Segment sg{0, 0, 0, 0, 0, 0};
    for (int t = 0; t < T; ++t) {
        if (numels[t] < 0) {
            set_error("tensor %d has negative numel", t);
            delete p;
            return B200DDP_EINVAL;
        }
        if (t - sg.t0 == kTabLarge) {            // table full: close the segment, open the next
            sg.t1 = t;
            sg.c1 = (int)p->chunks.size();
            sg.k1 = (int)p->copy_chunks.size();
            p->segments.push_back(sg);
            sg = Segment{t, t, sg.c1, sg.c1, sg.k1, sg.k1};
        }
        p->meta.push_back({numels[t], off});
        const int64_t nch = (numels[t] + kChunk - 1) / kChunk;
        const int64_t ncp = (numels[t] + copy_chunk - 1) / copy_chunk;
        if (nch > INT32_MAX || p->chunks.size() + (size_t)nch > (size_t)INT32_MAX) {
            set_error("tensor list too large for one plan");
            delete p;
            return B200DDP_ETOOBIG;
        }
        for (int64_t c = 0; c < nch; ++c) p->chunks.push_back({t - sg.t0, (int)c});
        for (int64_t c = 0; c < ncp; ++c) p->copy_chunks.push_back({t - sg.t0, (int)c});
        off += (numels[t] + kBucketAlign - 1) / kBucketAlign * kBucketAlign;
    }
    if (T > 0) {
        sg.t1 = T;
        sg.c1 = (int)p->chunks.size();
        sg.k1 = (int)p->copy_chunks.size();
        p->segments.push_back(sg);
    }
    p->bucket_elems = off;
    cudaError_t e = cudaSuccess;
    const size_t nU = p->chunks.size(), nK = p->copy_chunks.size();
    e = cudaMalloc((void **)&p->d_chunks, sizeof(Chunk) * (nU + nK + 1));
    p->d_copy_chunks = p->d_chunks ? p->d_chunks + nU : nullptr;
    if (e == cudaSuccess && nU)
        e = cudaMemcpy(p->d_chunks, p->chunks.data(), sizeof(Chunk) * nU, cudaMemcpyHostToDevice);
    if (e == cudaSuccess && nK)
        e = cudaMemcpy(p->d_copy_chunks, p->copy_chunks.data(), sizeof(Chunk) * nK, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        set_error("plan allocation failed: %s", cudaGetErrorString(e));
        b200ddp_plan_destroy(p);
        return B200DDP_ECUDA;
    }
    *out = p;
    return B200DDP_OK;
}

int b200ddp_plan_destroy(b200ddp_plan *p) {
    if (!p) return B200DDP_OK;
    int prev = -1;
    cudaGetDevice(&prev);
    cudaSetDevice(p->device);
    cudaFree(p->d_chunks);
    if (prev >= 0) cudaSetDevice(prev);
    cudaGetLastError();   // teardown never leaves a stale error behind for the next launch check
    delete p;
    return B200DDP_OK;
}

int64_t b200ddp_plan_bucket_elems(const b200ddp_plan *plan) { return plan ? plan->bucket_elems : -1; }

int b200ddp_plan_offsets(const b200ddp_plan *plan, int64_t *out) {
    B2_REQUIRE(plan != nullptr && (out != nullptr || plan->T == 0), B200DDP_EINVAL, "NULL argument");
    for (int t = 0; t < plan->T; ++t) out[t] = plan->meta[t].bucket_off;
    return B200DDP_OK;
}

// ------------------------------------------------------------ K1 / unpack / K3
int b200ddp_pack_scale(b200ddp_ctx *ctx, b200ddp_plan *plan, const void *const *grads, void *bucket, float scale,
                       b200ddp_stream stream) {
    if (int rc = check_plan(ctx, plan)) return rc;
    B2_REQUIRE(bucket != nullptr || plan->bucket_elems == 0, B200DDP_EINVAL, "bucket is NULL");
    B2_REQUIRE(((uintptr_t)bucket & 15) == 0, B200DDP_EINVAL, "bucket must be 16-byte aligned");
    std::lock_guard<std::mutex> lk(ctx->mu);
    cudaStream_t s = (cudaStream_t)stream;
    if (int rc = check_ptrs(plan, kGradIn, grads)) return rc;
    return launch_pack(plan, grads, bucket, scale, ctx->lc(), s);
}

int b200ddp_unpack(b200ddp_ctx *ctx, b200ddp_plan *plan, const void *bucket, void *const *grads_out, float divisor,
                   b200ddp_stream stream) {
    if (int rc = check_plan(ctx, plan)) return rc;
    B2_REQUIRE(bucket != nullptr || plan->bucket_elems == 0, B200DDP_EINVAL, "bucket is NULL");
    B2_REQUIRE(((uintptr_t)bucket & 15) == 0, B200DDP_EINVAL, "bucket must be 16-byte aligned");
    B2_REQUIRE(divisor != 0.0f && std::isfinite(divisor), B200DDP_EINVAL, "bad divisor");
    std::lock_guard<std::mutex> lk(ctx->mu);
    cudaStream_t s = (cudaStream_t)stream;
    if (int rc = check_ptrs(plan, kGradOut, (const void *const *)grads_out)) return rc;
    return launch_unpack(plan, bucket, (const void *const *)grads_out, divisor, ctx->lc(), s);
}

static int update_tables(const b200ddp_plan *plan, const SgdScalars &hp, const void *const *grads,
                         void *const *params, void *const *moms, void *const *grads_out, PtrArrays *pa,
                         void *const *vels = nullptr) {
    if (hp.rule == kRuleAdam) {
        if (int rc = check_ptrs(plan, kVel, (const void *const *)vels)) return rc;
        pa->k[kVel] = (const void *const *)vels;
    }
    if (int rc = check_ptrs(plan, kParam, (const void *const *)params)) return rc;
    pa->k[kParam] = (const void *const *)params;
    if (grads) {
        if (int rc = check_ptrs(plan, kGradIn, grads)) return rc;
        pa->k[kGradIn] = grads;
    }
    if (hp.has_momentum) {
        if (int rc = check_ptrs(plan, kMom, (const void *const *)moms)) return rc;
        pa->k[kMom] = (const void *const *)moms;
    }
    if (grads_out) {
        if (int rc = check_ptrs(plan, kGradOut, (const void *const *)grads_out)) return rc;
        pa->k[kGradOut] = (const void *const *)grads_out;
    }
    return B200DDP_OK;
}

int b200ddp_unpack_sgd(b200ddp_ctx *ctx, b200ddp_plan *plan, const void *bucket, void *const *params,
                       void *const *moms, void *const *grads_out, const b200ddp_sgd_args *args,
                       b200ddp_stream stream) {
    if (int rc = check_plan(ctx, plan)) return rc;
    B2_REQUIRE(bucket != nullptr || plan->bucket_elems == 0, B200DDP_EINVAL, "bucket is NULL");
    B2_REQUIRE(((uintptr_t)bucket & 15) == 0, B200DDP_EINVAL, "bucket must be 16-byte aligned");
    SgdScalars hp;
    if (int rc = sgd_scalars(args, ctx->div_mode, &hp)) return rc;
    if (plan->bucket_elems == 0) return B200DDP_OK;
    std::lock_guard<std::mutex> lk(ctx->mu);
    cudaStream_t s = (cudaStream_t)stream;
    PtrArrays pa;
    if (int rc = update_tables(plan, hp, nullptr, params, moms, grads_out, &pa)) return rc;
    return launch_update(plan, bucket, pa, hp, ctx->lc(), s);
}

int b200ddp_sgd_step(b200ddp_ctx *ctx, b200ddp_plan *plan, const void *const *grads, void *const *params,
                     void *const *moms, const b200ddp_sgd_args *args, b200ddp_stream stream) {
    if (int rc = check_plan(ctx, plan)) return rc;
    SgdScalars hp;
    if (int rc = sgd_scalars(args, ctx->div_mode, &hp)) return rc;
    if (plan->bucket_elems == 0) return B200DDP_OK;
    std::lock_guard<std::mutex> lk(ctx->mu);
    cudaStream_t s = (cudaStream_t)stream;
    B2_REQUIRE(grads != nullptr, B200DDP_EINVAL, "pointer table %d is NULL", (int)kGradIn);
    PtrArrays pa;
    if (int rc = update_tables(plan, hp, grads, params, moms, nullptr, &pa)) return rc;
    // direct mode: bucket == nullptr makes the kernel read the gradient tensors themselves
    return launch_update(plan, nullptr, pa, hp, ctx->lc(), s);
}

int b200ddp_unpack_opt(b200ddp_ctx *ctx, b200ddp_plan *plan, const void *bucket, void *const *params,
                       void *const *state1, void *const *state2, void *const *grads_out,
                       const b200ddp_opt_args *args, b200ddp_stream stream) {
    if (int rc = check_plan(ctx, plan)) return rc;
    B2_REQUIRE(bucket != nullptr || plan->bucket_elems == 0, B200DDP_EINVAL, "bucket is NULL");
    B2_REQUIRE(((uintptr_t)bucket & 15) == 0, B200DDP_EINVAL, "bucket must be 16-byte aligned");
    SgdScalars hp;
    if (int rc = opt_scalars(args, ctx->div_mode, &hp)) return rc;
    if (plan->bucket_elems == 0) return B200DDP_OK;
    std::lock_guard<std::mutex> lk(ctx->mu);
    PtrArrays pa;
    if (int rc = update_tables(plan, hp, nullptr, params, state1, grads_out, &pa, state2)) return rc;
    return launch_update(plan, bucket, pa, hp, ctx->lc(), (cudaStream_t)stream);
}

int b200ddp_opt_step(b200ddp_ctx *ctx, b200ddp_plan *plan, const void *const *grads, void *const *params,
                     void *const *state1, void *const *state2, const b200ddp_opt_args *args,
                     b200ddp_stream stream) {
    if (int rc = check_plan(ctx, plan)) return rc;
    SgdScalars hp;
    if (int rc = opt_scalars(args, ctx->div_mode, &hp)) return rc;
    if (plan->bucket_elems == 0) return B200DDP_OK;
    std::lock_guard<std::mutex> lk(ctx->mu);
    B2_REQUIRE(grads != nullptr, B200DDP_EINVAL, "pointer table %d is NULL", (int)kGradIn);
    PtrArrays pa;
    if (int rc = update_tables(plan, hp, grads, params, state1, nullptr, &pa, state2)) return rc;
    return launch_update(plan, nullptr, pa, hp, ctx->lc(), (cudaStream_t)stream);
}

// ------------------------------------------------------------------------ K2
int b200ddp_allreduce_pick(const b200ddp_ctx *ctx, int64_t nelem, int dtype) {
    if (!ctx) return B200DDP_EINVAL;
    if (ctx->world == 1) return B200DDP_ALGO_ONESHOT;
    const size_t bytes = (size_t)nelem * dtype_size(dtype);
    if (!ctx->peers_ready) return ctx->nccl_comm ? (int)B200DDP_ALGO_NCCL : (int)B200DDP_ESTATE;
    // Crossovers measured on 2, 4 and 8 B200s in round 2 (profiles/r02_sweep_n{2,4,8}_fp32_variants.jsonl,
    // r02_sweep_n{4,8}_fp32_bf16.jsonl).  One-shot moves (N-1) x S per GPU and only pays off while the
    // bucket is so small that saving the separate update launch (fused one-shot + SGD) matters: at 1 MB
    // two-shot already wins on 8 GPUs (20.7 vs 29.7 us) and on 4 (19.7 vs 22.0 us).  Two-shot moves
    // 2 (N-1)/N x S and beats NCCL at every size up to 256 MB on 2 and 4 GPUs (117 vs 139 us at the
    // ResNet-18 bucket on 4 GPUs) and up to 64 MB on 8.
    const long long oneshot_max = ctx->oneshot_max_bytes != -2
                                      ? ctx->oneshot_max_bytes
                                      : (long long)(ctx->world <= 2 ? 1024 : ctx->world <= 4 ? 512 : 256) << 10;
    if ((long long)bytes <= oneshot_max) return B200DDP_ALGO_ONESHOT;
    // In-switch reduction, when a multicast mapping was adopted.  With the round-2 kernel (one multimem
    // load in flight per thread, 74 blocks) on 8 GPUs: behind two-shot at 4 MB (31.7 vs 29.0 us), ahead from
    // 16 MB (59.7 vs 62.1 us), 121.9 vs 146.1 us (642 vs 536 GB/s busbw) at the ResNet-18 bucket, 816 vs 635
    // GB/s at 256 MB, 841 GB/s at 1 GiB (NCCL 726).  On 4 GPUs it draws level at 44 MB (119.4 vs 117.2 us) and
    // leads from 256 MB (606 vs 623 us); on 2 GPUs it loses everywhere (134 vs 83 us): the switch has nothing
    // to save there.  B200DDP_OPT_NVLS_MIN_BYTES overrides the threshold (-1: never).
    if (ctx->mc_ptr) {
        const long long dflt = ctx->world >= 8 ? (12ll << 20) : ctx->world >= 4 ? (128ll << 20) : -1;
        const long long nvls_min = ctx->nvls_min_bytes != -2 ? ctx->nvls_min_bytes : dflt;
        if (nvls_min >= 0 && (long long)bytes >= nvls_min) return B200DDP_ALGO_NVLS;
    }
    // NCCL only for fp32 (ncclAllReduce on bf16 rounds at every step, the kernels here once) and only
    // where it was measured ahead: 8 GPUs without a multicast mapping, above 128 MB (716 vs 740 us at 256 MB).
    if (ctx->world > 4 && ctx->nccl_comm && dtype == B200DDP_F32 && bytes > ((size_t)128 << 20)) return B200DDP_ALGO_NCCL;
    return B200DDP_ALGO_TWOSHOT;
}

int b200ddp_allreduce(b200ddp_ctx *ctx, size_t offset_bytes, int64_t nelem, int dtype, int algo,
                      b200ddp_stream stream, size_t *result_offset_bytes) {
    B2_REQUIRE(ctx != nullptr, B200DDP_EINVAL, "ctx is NULL");
    if (int rc = in_staging(ctx, offset_bytes, nelem, dtype)) return rc;
    if (result_offset_bytes) *result_offset_bytes = offset_bytes;
    if (ctx->world == 1 || nelem == 0) return B200DDP_OK;
    std::lock_guard<std::mutex> lk(ctx->mu);
    cudaStream_t s = (cudaStream_t)stream;
    if (algo == B200DDP_ALGO_AUTO) {
        algo = b200ddp_allreduce_pick(ctx, nelem, dtype);
        B2_REQUIRE(algo > 0, B200DDP_ESTATE, "no collective is set up: call b200ddp_ipc_open_peers or b200ddp_nccl_init");
    }
    if (algo == B200DDP_ALGO_NCCL) {
        B2_REQUIRE(ctx->nccl_comm != nullptr, B200DDP_ESTATE, "NCCL communicator not initialised");
        return nccl_allreduce(ctx, ctx->staging + offset_bytes, nelem, dtype, s);
    }
    B2_REQUIRE(ctx->peers_ready, B200DDP_ESTATE, "peer memory not opened");
    if (algo == B200DDP_ALGO_NVLS) {
        B2_REQUIRE(ctx->mc_ptr != nullptr, B200DDP_ESTATE, "no multicast mapping: NVLS needs b200ddp_adopt_symmetric");
        return launch_nvls_allreduce(ctx, offset_bytes, nelem, dtype, s);
    }
    B2_REQUIRE(algo == B200DDP_ALGO_ONESHOT || algo == B200DDP_ALGO_TWOSHOT, B200DDP_EINVAL, "bad algo %d", algo);
    if (algo == B200DDP_ALGO_ONESHOT && result_offset_bytes) *result_offset_bytes = ctx->staging_bytes + offset_bytes;
    return launch_p2p_allreduce(ctx, offset_bytes, nelem, dtype, algo, s);
}

int b200ddp_barrier(b200ddp_ctx *ctx, b200ddp_stream stream) {
    B2_REQUIRE(ctx != nullptr, B200DDP_EINVAL, "ctx is NULL");
    if (ctx->world == 1) return B200DDP_OK;
    B2_REQUIRE(ctx->peers_ready, B200DDP_ESTATE, "peer memory not opened");
    std::lock_guard<std::mutex> lk(ctx->mu);
    return launch_barrier(ctx, (cudaStream_t)stream);
}

int b200ddp_allreduce_sgd(b200ddp_ctx *ctx, b200ddp_plan *plan, size_t offset_bytes, void *const *params,
                          void *const *moms, void *const *grads_out, const b200ddp_sgd_args *args,
                          b200ddp_stream stream) {
    if (int rc = check_plan(ctx, plan)) return rc;
    if (int rc = in_staging(ctx, offset_bytes, plan->bucket_elems, plan->dtype)) return rc;
    SgdScalars hp;
    if (int rc = sgd_scalars(args, ctx->div_mode, &hp)) return rc;
    if (plan->bucket_elems == 0) return B200DDP_OK;
    B2_REQUIRE(ctx->peers_ready, B200DDP_ESTATE, "peer memory not opened");
    std::lock_guard<std::mutex> lk(ctx->mu);
    cudaStream_t s = (cudaStream_t)stream;
    PtrArrays pa;
    if (int rc = update_tables(plan, hp, nullptr, params, moms, grads_out, &pa)) return rc;
    if (ctx->world == 1) return launch_update(plan, ctx->staging + offset_bytes, pa, hp, ctx->lc(), s);
    return launch_oneshot_update(ctx, plan, offset_bytes, pa, hp, s);
}

// ----------------------------------------------------------------- emulation
static int check_emu(b200ddp_ctx *const *ctxs, int n) {
    B2_REQUIRE(ctxs != nullptr && n >= 2 && n <= B200DDP_MAX_WORLD, B200DDP_EINVAL, "bad emulated world %d", n);
    for (int r = 0; r < n; ++r) {
        B2_REQUIRE(ctxs[r] != nullptr, B200DDP_EINVAL, "ctx %d is NULL", r);
        B2_REQUIRE(ctxs[r]->world == n && ctxs[r]->rank == r, B200DDP_EINVAL, "ctx %d has rank %d / world %d", r,
                   ctxs[r]->rank, ctxs[r]->world);
        B2_REQUIRE(ctxs[r]->device == ctxs[0]->device, B200DDP_EINVAL, "emulated ranks must share one device");
        B2_REQUIRE(ctxs[r]->peers_ready, B200DDP_ESTATE, "ctx %d has no peers", r);
        B2_REQUIRE(ctxs[r]->staging_bytes == ctxs[0]->staging_bytes, B200DDP_ESTATE,
                   "emulated contexts differ in staging size");
    }
    return B200DDP_OK;
}

int b200ddp_emu_allreduce(b200ddp_ctx *const *ctxs, int n, size_t offset_bytes, int64_t nelem, int dtype, int algo,
                          b200ddp_stream stream, size_t *result_offset_bytes) {
    if (int rc = check_emu(ctxs, n)) return rc;
    if (int rc = in_staging(ctxs[0], offset_bytes, nelem, dtype)) return rc;
    if (algo == B200DDP_ALGO_AUTO) algo = b200ddp_allreduce_pick(ctxs[0], nelem, dtype);
    B2_REQUIRE(algo == B200DDP_ALGO_ONESHOT || algo == B200DDP_ALGO_TWOSHOT, B200DDP_EINVAL,
               "emulation covers the P2P algorithms only (got %d)", algo);
    if (result_offset_bytes)
        *result_offset_bytes = algo == B200DDP_ALGO_ONESHOT ? ctxs[0]->staging_bytes + offset_bytes : offset_bytes;
    if (nelem == 0) return B200DDP_OK;
    B2_CHECK_CUDA(cudaSetDevice(ctxs[0]->device));
    return launch_p2p_allreduce_emu(ctxs, n, offset_bytes, nelem, dtype, algo, (cudaStream_t)stream);
}

int b200ddp_emu_allreduce_sgd(b200ddp_ctx *const *ctxs, b200ddp_plan *const *plans, int n, size_t offset_bytes,
                              void *const *const *params, void *const *const *moms,
                              void *const *const *grads_out, const b200ddp_sgd_args *args, b200ddp_stream stream) {
    if (int rc = check_emu(ctxs, n)) return rc;
    B2_REQUIRE(plans != nullptr && params != nullptr, B200DDP_EINVAL, "NULL argument");
    SgdScalars hp;
    if (int rc = sgd_scalars(args, ctxs[0]->div_mode, &hp)) return rc;
    B2_REQUIRE(!hp.has_momentum || moms != nullptr, B200DDP_EINVAL, "momentum buffers missing");
    cudaStream_t s = (cudaStream_t)stream;
    B2_CHECK_CUDA(cudaSetDevice(ctxs[0]->device));
    PtrArrays pas[B200DDP_MAX_WORLD];
    for (int r = 0; r < n; ++r) {
        if (int rc = check_plan(ctxs[r], plans[r])) return rc;
        B2_REQUIRE(plans[r]->bucket_elems == plans[0]->bucket_elems && plans[r]->T == plans[0]->T &&
                       plans[r]->dtype == plans[0]->dtype,
                   B200DDP_EINVAL, "plans differ across emulated ranks");
        if (int rc = update_tables(plans[r], hp, nullptr, params[r], moms ? moms[r] : nullptr,
                                   grads_out ? grads_out[r] : nullptr, &pas[r]))
            return rc;
    }
    if (int rc = in_staging(ctxs[0], offset_bytes, plans[0]->bucket_elems, plans[0]->dtype)) return rc;
    if (plans[0]->bucket_elems == 0) return B200DDP_OK;
    return launch_oneshot_update_emu(ctxs, plans, n, offset_bytes, pas, hp, s);
}

// ------------------------------------------------------------------------ K4
int b200ddp_sampler_sizes(int64_t n, int world, int drop_last, int64_t *num_samples, int64_t *total_size) {
    B2_REQUIRE(n >= 0 && world >= 1, B200DDP_EINVAL, "bad sampler geometry n=%lld world=%d", (long long)n, world);
    int64_t num;
    if (drop_last && n % world != 0) {
        const int64_t a = n - world;   // ceil((n - world) / world), torch/utils/data/distributed.py:98-100
        num = a >= 0 ? (a + world - 1) / world : -((-a) / world);
    } else {
        num = (n + world - 1) / world;   // :102
    }
    if (num_samples) *num_samples = num;
    if (total_size) *total_size = num * world;
    return B200DDP_OK;
}

int b200ddp_sampler_shard(b200ddp_ctx *ctx, int64_t n, int world, int rank, uint64_t seed, uint64_t epoch,
                          int shuffle, int drop_last, int64_t *out_dev, b200ddp_stream stream) {
    B2_REQUIRE(ctx != nullptr, B200DDP_EINVAL, "ctx is NULL");
    B2_REQUIRE(world >= 1, B200DDP_EINVAL, "num_replicas must be positive");
    B2_REQUIRE(rank >= 0 && rank < world, B200DDP_EINVAL, "Invalid rank %d, rank should be in the interval [0, %d]",
               rank, world - 1);
    B2_REQUIRE(n >= 0 && n < (int64_t)(0xffffffffu / 20), B200DDP_ETOOBIG,
               "dataset length %lld is outside the 32-bit randperm range", (long long)n);
    int64_t num = 0;
    if (int rc = b200ddp_sampler_sizes(n, world, drop_last, &num, nullptr)) return rc;
    if (num <= 0 || n == 0) return B200DDP_OK;
    B2_REQUIRE(out_dev != nullptr, B200DDP_EINVAL, "out_dev is NULL");
    std::lock_guard<std::mutex> lk(ctx->mu);
    B2_CHECK_CUDA(cudaSetDevice(ctx->device));
    return launch_sampler(ctx, n, world, rank, seed + epoch, shuffle, num, out_dev, (cudaStream_t)stream);
}

}  // extern "C"

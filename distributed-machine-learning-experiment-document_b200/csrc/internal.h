// Internal declarations shared by the translation units of libb200ddp.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include <mutex>
#include <vector>

#include "../../include/b200ddp.h"

namespace b200ddp {

// ---- error plumbing ---------------------------------------------------------
void set_error(const char *fmt, ...);
#define B2_CHECK_CUDA(expr)                                                              \
    do {                                                                                 \
        cudaError_t _e = (expr);                                                         \
        if (_e != cudaSuccess) {                                                         \
            ::b200ddp::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                                 __FILE__, __LINE__);                                    \
            return B200DDP_ECUDA;                                                        \
        }                                                                                \
    } while (0)
#define B2_REQUIRE(cond, code, ...)            \
    do {                                       \
        if (!(cond)) {                         \
            ::b200ddp::set_error(__VA_ARGS__); \
            return (code);                     \
        }                                      \
    } while (0)

// ---- geometry of the multi-tensor kernels ------------------------------------
constexpr int kThreads = 256;          // threads per block, pack / unpack / update
constexpr int kVec = 4;                // fp32 elements per 128-bit access
constexpr int kUnroll = 2;             // 128-bit accesses per thread per stream per chunk
constexpr int kChunk = kThreads * kVec * kUnroll;  // 2048 elements of one tensor
constexpr int kBucketAlign = 32;       // tensor offsets in the bucket, in elements
constexpr int kMaxBlocksP2P = 1024;    // flag slots: [block][src rank]
constexpr int kSmCount = 148;

struct TensorMeta {       // immutable per plan
    int64_t numel;
    int64_t bucket_off;   // elements
};
struct Chunk {            // immutable per plan: one kChunk-sized piece of one tensor
    int tensor;
    int index;            // chunk number inside the tensor
};

enum PtrKind { kGradIn = 0, kParam = 1, kMom = 2, kGradOut = 3, kNumPtrKinds = 4 };

// Device view of a plan, passed to kernels by value.
struct PlanView {
    const TensorMeta *meta;
    const Chunk *chunks;
    int nchunks;
    int T;
    void *const *ptrs[kNumPtrKinds];   // each: T device pointers (device array)
};

struct SgdScalars {   // host-side casts follow torch: scalars are cast to fp32 once
    float neg_lr, momentum, one_minus_damp, weight_decay, divisor;
    int use_div, nesterov, maximize, first_step, has_momentum, has_wd;
};

struct PeerTable {    // kernel parameter of the P2P kernels
    char *staging[B200DDP_MAX_WORLD];
    uint32_t *flags[B200DDP_MAX_WORLD];
};

}  // namespace b200ddp

struct b200ddp_ctx {
    int rank = 0, world = 1, device = 0;
    size_t staging_bytes = 0;          // one half
    char *staging = nullptr;           // 2 x staging_bytes
    uint32_t *flags = nullptr;         // kMaxBlocksP2P x MAX_WORLD
    bool peers_ready = false;
    bool peers_ipc = false;
    b200ddp::PeerTable peers{};
    uint32_t epoch = 0;                // barrier generation, identical on every rank
    void *nccl_comm = nullptr;
    int sm_count = b200ddp::kSmCount;
    // sampler scratch
    void *sampler_scratch = nullptr;
    size_t sampler_scratch_bytes = 0;
    std::mutex mu;
};

struct b200ddp_plan {
    b200ddp_ctx *ctx = nullptr;        // identity check only; never dereferenced after creation
    int device = 0;
    int T = 0;
    int dtype = B200DDP_F32;
    int64_t bucket_elems = 0;
    std::vector<b200ddp::TensorMeta> meta;
    std::vector<b200ddp::Chunk> chunks;
    // device tables
    b200ddp::TensorMeta *d_meta = nullptr;
    b200ddp::Chunk *d_chunks = nullptr;
    void **d_ptrs = nullptr;                    // kNumPtrKinds x T
    // host mirrors + upload ring
    std::vector<void *> h_ptrs;                 // last uploaded values, kNumPtrKinds x T
    bool uploaded[b200ddp::kNumPtrKinds] = {false, false, false, false};
    static constexpr int kRing = 8;
    void **pinned[kRing] = {};
    cudaEvent_t pinned_ev[kRing] = {};
    int ring_pos = 0;
};

namespace b200ddp {

// api.cu
int plan_upload(b200ddp_plan *plan, PtrKind kind, const void *const *host_ptrs, cudaStream_t s);
PlanView plan_view(const b200ddp_plan *plan);
int sgd_scalars(const b200ddp_sgd_args *a, SgdScalars *out);

// pack_update.cu
int launch_pack(const PlanView &pv, int dtype, void *bucket, float scale, int sm_count, cudaStream_t s);
int launch_unpack(const PlanView &pv, int dtype, const void *bucket, float divisor, int sm_count, cudaStream_t s);
int launch_update(const PlanView &pv, int dtype, const void *bucket /*null: direct*/, bool write_grad,
                  const SgdScalars &hp, int sm_count, cudaStream_t s);

// p2p.cu
int p2p_grid_for(int64_t nelem, int dtype, int algo, int world, int sm_count);
int launch_p2p_allreduce(b200ddp_ctx *ctx, size_t off, int64_t nelem, int dtype, int algo, cudaStream_t s);
int launch_p2p_allreduce_emu(b200ddp_ctx *const *ctxs, int n, size_t off, int64_t nelem, int dtype,
                             int algo, cudaStream_t s);
int launch_oneshot_update(b200ddp_ctx *ctx, const PlanView &pv, int dtype, size_t off, bool write_grad,
                          const SgdScalars &hp, cudaStream_t s);
int launch_oneshot_update_emu(b200ddp_ctx *const *ctxs, const PlanView *pvs, int n, int dtype, size_t off,
                              bool write_grad, const SgdScalars &hp, cudaStream_t s);

// nccl_dyn.cpp
int nccl_unique_id(void *out128);
int nccl_init(b200ddp_ctx *ctx, const void *uid);
int nccl_allreduce(b200ddp_ctx *ctx, void *buf, int64_t nelem, int dtype, cudaStream_t s);
void nccl_destroy(b200ddp_ctx *ctx);

// sampler.cu
int launch_sampler(b200ddp_ctx *ctx, int64_t n, int world, int rank, uint64_t seed, int shuffle,
                   int64_t num_samples, int64_t *out_dev, cudaStream_t s);

}  // namespace b200ddp

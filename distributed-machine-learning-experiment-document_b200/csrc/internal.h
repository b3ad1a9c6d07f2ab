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

// One launch path for every kernel: cudaLaunchKernelEx, with the programmatic-dependent-launch
// attribute when the context asks for it (see pdl_prologue in device_utils.cuh).
template <typename... KArgs, typename... Args>
inline cudaError_t launch_ex(void (*kernel)(KArgs...), int grid, int block, cudaStream_t s, bool pdl, Args &&...args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3((unsigned)block);
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// ---- geometry of the multi-tensor kernels ------------------------------------
constexpr int kThreads = 256;          // threads per block, pack / unpack / update
constexpr int kVec = 4;                // fp32 elements per 128-bit access
constexpr int kUnroll = 2;             // 128-bit accesses per thread per stream per chunk
constexpr int kChunk = kThreads * kVec * kUnroll;  // 2048 elements of one tensor (update kernels)
// pack / unpack move two streams, not five, and have the registers to keep more loads in flight per
// thread: their chunks are kThreads * kVec * U elements, U in {2, 4, 8} chosen per context
// (B200DDP_COPY_UNROLL).  Measured on ResNet-18 (profiles/r02_kernels_resnet18_variants.json): 2, 4 and 8 are
// within 2 % of each other - the pack is not short of loads in flight; what it was short of is the launch ramp,
// which programmatic dependent launch removes (0.74 -> 0.82 of the measured HBM peak)
constexpr int kCopyUnrollDefault = 2;
constexpr int kBucketAlign = 32;       // tensor offsets in the bucket, in elements
constexpr int kMaxBlocksP2P = 1024;    // flag slots: [block][src rank]
constexpr int kSmCount = 148;
constexpr int kTabSmall = 32;          // tensors per launch, small table (1.5 KB of kernel parameters)
constexpr int kTabLarge = 256;         // tensors per launch, large table (12 KB); longer lists go in segments

struct TensorMeta {       // immutable per plan
    int64_t numel;
    int64_t bucket_off;   // elements
};
struct Chunk {            // immutable per plan: one kChunk-sized piece of one tensor
    int tensor;           // index inside the chunk's segment
    int index;            // chunk number inside the tensor
};
struct Segment {          // at most kTabLarge consecutive tensors = one kernel launch
    int t0, t1;           // tensors [t0, t1)
    int c0, c1;           // chunks [c0, c1) of the plan's update chunk table
    int k0, k1;           // chunks [k0, k1) of the plan's copy (pack / unpack) chunk table
};

// kMom: SGD momentum buffer / Adam first moment; kVel: Adam second moment
enum PtrKind { kGradIn = 0, kParam = 1, kMom = 2, kGradOut = 3, kVel = 4, kNumPtrKinds = 5 };

// Per-launch tensor table, passed BY VALUE as a __grid_constant__ kernel
// parameter: sizes, bucket offsets and raw pointers sit in the constant bank.
// No device-side pointer table has to be kept in step with torch's allocator
// (p.grad moves every step), nothing is uploaded, and a launch captured in a
// CUDA graph carries its pointers with it.
template <int CAP, int NK> struct TensorTab {
    long long numel[CAP];
    long long boff[CAP];
    void *ptr[NK][CAP];
};

// `param.grad.data /= world_size` (codes/task2/dist_utils.py:42), per element.
//   kDivRcp  : x * RN(1/d) - what torch computes ON CUDA, where the reference's scripts run (nccl backend,
//              codes/task2/dist_utils.py:12): ATen's div_true_kernel_cuda multiplies by the reciprocal when the
//              divisor is a CPU scalar.  Default (B200DDP_DIV_RECIPROCAL).
//   kDivFast / kDivIeee : the correctly rounded quotient - what torch computes on the CPU (the lab's gloo
//              run): a 3-instruction sequence for small integer divisors, __fdiv_rn otherwise
//              (B200DDP_DIV_IEEE).
// For power-of-two divisors (worlds 2, 4, 8) all of them give the same bits.
enum DivKind { kDivNone = 0, kDivRcp = 1, kDivFast = 2, kDivIeee = 3 };
struct Divisor { float d, rcp; int kind; };

// Divisors for which the short division sequence was checked bit for bit against IEEE division
// (every integer 2..64, tests/test_host_logic.py); world sizes are in this range.
inline bool fast_divisor(float d) { return d >= 2.0f && d <= 64.0f && d == (float)(int)d; }
inline Divisor make_divisor(float d, int div_mode) {
    Divisor v{d, 1.0f / d, kDivNone};
    if (d != 1.0f) v.kind = div_mode == B200DDP_DIV_RECIPROCAL ? kDivRcp : (fast_divisor(d) ? kDivFast : kDivIeee);
    return v;
}

enum UpdateRule { kRuleSgd = 0, kRuleAdam = 1, kRuleGd = 2 };

struct SgdScalars {   // host-side casts follow torch: scalars are cast to fp32 once
    float neg_lr, momentum, one_minus_damp, weight_decay;
    Divisor dv;
    int nesterov, maximize, first_step, has_momentum, has_wd;
    // the lab's own optimizers (codes/task1/pytorch/MyOptimizer.py:18-43)
    int rule;
    float lr, beta1, one_minus_beta1, beta2, one_minus_beta2, eps;
};

struct LaunchCfg {    // per-context launch settings, read at every call
    int sm_count;
    bool pdl;         // launch with the programmatic-stream-serialization attribute
    int div_mode;     // B200DDP_DIV_*
    int blocks_per_sm;   // grid cap of pack / unpack / update in blocks per SM; 0: as many as fit (occupancy)
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
    bool staging_external = false;     // adopted symmetric memory: owned by the caller
    char *retired_staging = nullptr;   // own IPC-exported staging replaced by symmetric memory, not yet freed
    char *mc_ptr = nullptr;            // multicast address over every rank's staging (NVLS), or null
    uint32_t *flags = nullptr;         // peer-visible: kMaxBlocksP2P x MAX_WORLD
    uint32_t *block_gen = nullptr;     // local: barrier generation of each block index, advanced by the kernels
    bool peers_ready = false;
    bool peers_ipc = false;
    b200ddp::PeerTable peers{};
    void *nccl_comm = nullptr;
    int sm_count = b200ddp::kSmCount;
    int copy_unroll = b200ddp::kCopyUnrollDefault;   // 128-bit accesses per thread per chunk in pack / unpack
    bool pdl = false;
    int div_mode = B200DDP_DIV_RECIPROCAL;
    long long nvls_min_bytes = -2;     // -2: built-in threshold, -1: never, >= 0: from this size
    long long oneshot_max_bytes = -2;  // -2: built-in threshold
    long long peer_timeout_ms = 0;     // 0: wait for ever, as an NCCL collective does
    int p2p_blocks = 0;                // 0: built-in grid cap
    int nvls_depth = 0;                // 0: built-in number of multimem loads in flight per thread
    int nvls_blocks = 0;               // 0: built-in grid cap of the in-switch kernel
    unsigned long long *trace = nullptr;   // B200DDP_OPT_TRACE: 4 time stamps per block of the last P2P launch
    int hbm_blocks_per_sm = 0;         // 0: pack / unpack / update fill the GPU; n: at most n blocks per SM
    b200ddp::LaunchCfg lc() const { return b200ddp::LaunchCfg{sm_count, pdl, div_mode, hbm_blocks_per_sm}; }
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
    int copy_unroll = b200ddp::kCopyUnrollDefault;
    std::vector<b200ddp::Chunk> chunks;        // update kernels: kChunk elements each
    std::vector<b200ddp::Chunk> copy_chunks;   // pack / unpack: kThreads * kVec * copy_unroll elements each
    std::vector<b200ddp::Segment> segments;
    b200ddp::Chunk *d_chunks = nullptr;        // the only device-side table: static, written once;
    b200ddp::Chunk *d_copy_chunks = nullptr;   //   the copy table follows the update table in the same allocation
};

namespace b200ddp {

// Host view of one call: the plan plus the caller's pointer arrays (nullable per kind).
struct PtrArrays {
    const void *const *k[kNumPtrKinds] = {nullptr, nullptr, nullptr, nullptr, nullptr};
};

// api.cu
int sgd_scalars(const b200ddp_sgd_args *a, int div_mode, SgdScalars *out);
int opt_scalars(const b200ddp_opt_args *a, int div_mode, SgdScalars *out);
int check_ptrs(const b200ddp_plan *plan, PtrKind kind, const void *const *host_ptrs);

// pack_update.cu
int launch_pack(const b200ddp_plan *plan, const void *const *grads, void *bucket, float scale, const LaunchCfg &lc,
                cudaStream_t s);
int launch_unpack(const b200ddp_plan *plan, const void *bucket, const void *const *grads_out, float divisor,
                  const LaunchCfg &lc, cudaStream_t s);
int launch_update(const b200ddp_plan *plan, const void *bucket /*null: direct*/, const PtrArrays &pa,
                  const SgdScalars &hp, const LaunchCfg &lc, cudaStream_t s);

// p2p.cu
int p2p_grid_for(int64_t nelem, int dtype, int algo, int world, int block_cap);
int launch_p2p_allreduce(b200ddp_ctx *ctx, size_t off, int64_t nelem, int dtype, int algo, cudaStream_t s);
int launch_barrier(b200ddp_ctx *ctx, cudaStream_t s);
int launch_nvls_allreduce(b200ddp_ctx *ctx, size_t off, int64_t nelem, int dtype, cudaStream_t s);
int launch_p2p_allreduce_emu(b200ddp_ctx *const *ctxs, int n, size_t off, int64_t nelem, int dtype,
                             int algo, cudaStream_t s);
int launch_oneshot_update(b200ddp_ctx *ctx, const b200ddp_plan *plan, size_t off, const PtrArrays &pa,
                          const SgdScalars &hp, cudaStream_t s);
int launch_oneshot_update_emu(b200ddp_ctx *const *ctxs, b200ddp_plan *const *plans, int n, size_t off,
                              const PtrArrays *pas, const SgdScalars &hp, cudaStream_t s);

// nccl_dyn.cpp
int nccl_unique_id(void *out128);
int nccl_init(b200ddp_ctx *ctx, const void *uid);
int nccl_allreduce(b200ddp_ctx *ctx, void *buf, int64_t nelem, int dtype, cudaStream_t s);
void nccl_destroy(b200ddp_ctx *ctx);

// sampler.cu
int launch_sampler(b200ddp_ctx *ctx, int64_t n, int world, int rank, uint64_t seed, int shuffle,
                   int64_t num_samples, int64_t *out_dev, cudaStream_t s);

}  // namespace b200ddp

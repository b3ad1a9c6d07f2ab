/* b200ddp - C ABI of the B200-native data-parallel gradient-sync path.
 *
 * One process per GPU.  Plain pointers and sizes only; no torch types.  Every
 * function returns 0 on success or a negative B200DDP_E* code and never
 * throws; b200ddp_last_error() gives the message of the calling thread's last
 * failure.  All device work is enqueued on the stream the caller passes; no
 * entry point synchronises the host unless its comment says so.
 *
 * The reference (/root/reference, a PyTorch lab handbook) has no native code
 * and no FFI: its interface for this path is three Python call sites.  Each
 * group of entry points below names the reference lines it stands in for;
 * INTEGRATION.md shows the ctypes stubs a maintainer adds to dist_utils.py.
 */
#ifndef B200DDP_H_
#define B200DDP_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200ddp_ctx b200ddp_ctx;    /* per-process, per-GPU state        */
typedef struct b200ddp_plan b200ddp_plan;  /* bucket layout of one tensor list  */
typedef void *b200ddp_stream;              /* a cudaStream_t                    */

enum { B200DDP_F32 = 0, B200DDP_BF16 = 1 };                 /* bucket / wire dtype */
enum { B200DDP_ALGO_AUTO = 0, B200DDP_ALGO_NCCL = 1,
       B200DDP_ALGO_ONESHOT = 2, B200DDP_ALGO_TWOSHOT = 3,
       B200DDP_ALGO_NVLS = 4 /* in-switch reduction, needs b200ddp_adopt_symmetric with a multicast address */ };
enum { B200DDP_OK = 0, B200DDP_EINVAL = -1, B200DDP_ECUDA = -2, B200DDP_ENCCL = -3,
       B200DDP_ESTATE = -4, B200DDP_ETOOBIG = -5 };

/* how `param.grad.data /= world_size` (codes/task2/dist_utils.py:42) is rounded when world_size is not a
 * power of two (for 2, 4, 8 both give the same bits):
 *   RECIPROCAL  x * (float)(1/N): torch on CUDA - the reference's scripts run with backend "nccl" on .cuda()
 *               tensors (codes/task2/dist_utils.py:12, codes/task2/model.py:128), where ATen divides by a
 *               host scalar by multiplying with its reciprocal.  Default.
 *   IEEE        the correctly rounded quotient: torch on the CPU (the gloo run). */
enum { B200DDP_DIV_IEEE = 0, B200DDP_DIV_RECIPROCAL = 1 };
/* b200ddp_set_option keys */
enum { B200DDP_OPT_DIV_MODE = 1,          /* B200DDP_DIV_*                                                   */
       B200DDP_OPT_PDL = 2,               /* 1: launch kernels with programmatic dependent launch            */
       B200DDP_OPT_NVLS_MIN_BYTES = 3,    /* in-switch all-reduce from this size; -1 never; -2 built-in rule */
       B200DDP_OPT_ONESHOT_MAX_BYTES = 4, /* one-shot all-reduce up to this size; -2 built-in rule           */
       B200DDP_OPT_PEER_TIMEOUT_MS = 5,   /* cross-GPU barrier gives up (traps) after this long; 0: never     */
       B200DDP_OPT_P2P_BLOCKS = 6,        /* grid cap of the P2P / NVLS kernels; 0: built-in (2 per SM)       */
       B200DDP_OPT_NVLS_DEPTH = 7,        /* multimem loads in flight per thread (1, 2, 4, 8); 0: built-in    */
       B200DDP_OPT_COPY_UNROLL = 8,       /* 128-bit accesses per thread in pack / unpack (2, 4, 8); new plans */
       B200DDP_OPT_TRACE = 9,             /* 1: the cross-GPU kernels stamp their phases (b200ddp_trace_read)   */
       B200DDP_OPT_NVLS_BLOCKS = 11,      /* grid cap of the in-switch kernel; 0: built-in                        */
       B200DDP_OPT_HBM_BLOCKS_PER_SM = 10 }; /* grid cap of pack / unpack / update, blocks per SM; 0: fill the GPU.
                                              A small value leaves SM slots to kernels of other streams.          */

#define B200DDP_MAX_WORLD 8
#define B200DDP_IPC_BYTES 128   /* staging handle (64 B) + flag handle (64 B) */
#define B200DDP_NCCL_UID_BYTES 128

const char *b200ddp_last_error(void);
int b200ddp_version(void);

/* ---- context ---------------------------------------------------------------
 * Replaces: dist_init / get_local_rank / get_world_size,
 *   codes/task2/dist_utils.py:6-30 (rendezvous stays torch.distributed; this
 *   only adds what the kernels need: peer-visible staging, flags, NCCL comm).
 * b200ddp_create makes `device` current, allocates 2 x staging_bytes of
 * peer-visible staging (the second half receives one-shot results) and the
 * barrier flags; synchronises the host.
 */
int b200ddp_create(int rank, int world, int device, size_t staging_bytes, b200ddp_ctx **out);
int b200ddp_destroy(b200ddp_ctx *ctx);
/* Per-context settings (B200DDP_OPT_*); each has an environment default read at b200ddp_create
 * (B200DDP_DIV_MODE=ieee|reciprocal, B200DDP_PDL, B200DDP_NVLS_MIN_BYTES, B200DDP_ONESHOT_MAX_BYTES,
 * B200DDP_PEER_TIMEOUT_MS, B200DDP_P2P_BLOCKS, B200DDP_NVLS_DEPTH, B200DDP_NVLS_BLOCKS, B200DDP_COPY_UNROLL).  Settings that
 * change the shape of a collective launch (grid, algorithm thresholds) must be equal on all ranks. */
int b200ddp_set_option(b200ddp_ctx *ctx, int option, long long value);
int b200ddp_get_option(const b200ddp_ctx *ctx, int option, long long *value);
/* With B200DDP_OPT_TRACE on: 4 x nblocks nanosecond stamps (%globaltimer) of the LAST cross-GPU launch,
 * per block: entry, first barrier passed, data phase done, last barrier passed.  Synchronises the device.
 * (ncu cannot replay a kernel that waits for other GPUs; this is the library's own phase profile.) */
int b200ddp_trace_read(b200ddp_ctx *ctx, uint64_t *out /* 4 x nblocks */, int nblocks);
int b200ddp_staging(b200ddp_ctx *ctx, void **ptr, size_t *bytes);
int b200ddp_flags(b200ddp_ctx *ctx, void **ptr);
/* this process's mapping of rank `peer`'s staging area (after the peers are opened); used by
 * the flat initial-weight broadcast that replaces init_parameters, codes/task2/dist_utils.py:33-37:
 * rank 0 packs, the others unpack straight out of rank 0's memory over NVLink. */
int b200ddp_peer_staging(b200ddp_ctx *ctx, int peer, void **ptr);

/* NCCL baseline collective: rank 0 makes the id, every rank joins (collective).
 * Join it only when B200DDP_ALGO_NCCL is really wanted: measured on 8 x B200, a second NVLS-capable
 * communicator in a process that already holds one (torch.distributed's) puts ~0.3 us of idle time
 * between consecutive kernels of every later CUDA-graph replay (profiles/r02_second_nccl_comm_n8.md).
 * The Python side therefore joins on the first all-reduce that asks for the algorithm (B200DDP_NCCL=lazy). */
int b200ddp_nccl_unique_id(void *out_uid /* B200DDP_NCCL_UID_BYTES */);
int b200ddp_nccl_init(b200ddp_ctx *ctx, const void *uid);

/* Peer memory for the P2P kernels.  Across processes: export, all-gather the
 * handles with torch.distributed, open.  Inside one process (tests, emulation):
 * pass the raw pointers of every rank's ctx. */
int b200ddp_ipc_export(b200ddp_ctx *ctx, void *out /* B200DDP_IPC_BYTES */);
int b200ddp_ipc_open_peers(b200ddp_ctx *ctx, const void *handles /* world x B200DDP_IPC_BYTES */);
int b200ddp_set_peers_direct(b200ddp_ctx *ctx, void *const *staging_ptrs, void *const *flag_ptrs);

/* Replace the staging area by caller-provided SYMMETRIC memory: peer_ptrs[r] is this process's
 * mapping of rank r's buffer (peer_ptrs[rank] the local one), all of `bytes` >= 2 x staging_bytes
 * and zeroed; mc_ptr (nullable) is a multicast address over the same buffers (NVLink SHARP:
 * a multimem.ld_reduce on it returns the sum over all ranks, a multimem.st on it stores to all).
 * The Python side gets these from torch.distributed._symmetric_memory; ownership stays with the
 * caller, which must keep the memory alive until b200ddp_destroy.  Collective in effect: every
 * rank must adopt, or none. */
int b200ddp_adopt_symmetric(b200ddp_ctx *ctx, void *const *peer_ptrs, void *mc_ptr, size_t bytes);
/* Memory exported over CUDA IPC must not be freed while a peer still has it mapped, so giving it up
 * takes two steps with a cross-rank barrier (the caller's: torch.distributed.barrier) between them:
 *   adopt:    b200ddp_adopt_symmetric (closes this rank's mappings of the peers' old staging and retires
 *             its own) | barrier | b200ddp_free_retired (frees the retired area)
 *   teardown: b200ddp_close_peers (synchronises the device, closes every imported mapping) | barrier |
 *             b200ddp_destroy.
 * b200ddp_destroy alone still closes and frees everything; the barrier is what makes the order safe. */
int b200ddp_free_retired(b200ddp_ctx *ctx);
int b200ddp_close_peers(b200ddp_ctx *ctx);

/* ---- plan: bucket layout of a list of T fp32 tensors ------------------------
 * Tensor t occupies bucket elements [offset[t], offset[t] + numel[t]); offsets
 * are aligned to 32 elements, the padding stays zero.  A plan is immutable and
 * owns one device-side table (the chunk list).  Pointer arrays passed to the calls below are HOST
 * arrays read at call time and copied into the kernel parameters: nothing is retained, a plan can
 * be used from any stream, and a launch captured in a CUDA graph carries its pointers with it. */
int b200ddp_plan_create(b200ddp_ctx *ctx, int T, const int64_t *numels, int bucket_dtype,
                        b200ddp_plan **out);
int b200ddp_plan_destroy(b200ddp_plan *plan);
int64_t b200ddp_plan_bucket_elems(const b200ddp_plan *plan);
int b200ddp_plan_offsets(const b200ddp_plan *plan, int64_t *out /* T */);

/* ---- K1 pack + scale (+ bf16 cast) -----------------------------------------
 * Replaces the read side of `dist.all_reduce(param.grad.data)` per parameter,
 *   codes/task2/dist_utils.py:40-41 / codes/task3/dist_utils.py:44-45, and,
 *   when scale = 1/world, the division at :42 / :46.
 * bucket[offset[t] + i] = cast(grads[t][i] * scale).  `grads` is a HOST array
 * of T device pointers to contiguous fp32 tensors. */
int b200ddp_pack_scale(b200ddp_ctx *ctx, b200ddp_plan *plan, const void *const *grads,
                       void *bucket, float scale, b200ddp_stream stream);

/* ---- K2 all-reduce (SUM) of a bucket that lives in this ctx's staging -------
 * Replaces `dist.all_reduce(..., op=SUM)`, codes/task2/dist_utils.py:41.
 * offset_bytes / nelem address the first staging half.  NCCL, two-shot and NVLS
 * reduce in place; one-shot writes its result into the second half.
 * *result_offset_bytes receives where the sums are, relative to the staging
 * base.  Every rank must make the same calls in the same order.  world == 1
 * is a no-op.  bf16 buckets are summed in fp32 and rounded once by the one-shot, two-shot and NVLS
 * kernels; B200DDP_ALGO_NCCL is ncclAllReduce on ncclBfloat16, which rounds at every reduction step,
 * so AUTO never picks NCCL for a bf16 bucket. */
int b200ddp_allreduce(b200ddp_ctx *ctx, size_t offset_bytes, int64_t nelem, int dtype, int algo,
                      b200ddp_stream stream, size_t *result_offset_bytes);
/* Device-side barrier over all ranks, enqueued on `stream` (one tiny kernel, no host synchronisation):
 * what each rank enqueued before it is visible to what every rank enqueues after it.  Collective.
 * Stands in for the implicit ordering of the per-parameter `dist.broadcast` calls of init_parameters,
 * codes/task2/dist_utils.py:33-37, in the flat broadcast (rank 0 packs | barrier | peers read). */
int b200ddp_barrier(b200ddp_ctx *ctx, b200ddp_stream stream);
/* which algorithm AUTO resolves to for this size (host-only, no device work) */
int b200ddp_allreduce_pick(const b200ddp_ctx *ctx, int64_t nelem, int dtype);

/* ---- unpack (bucket -> p.grad) ----------------------------------------------
 * Replaces the write side of the in-place all_reduce and `param.grad.data /=
 * world_size`, codes/task2/dist_utils.py:41-42: grads_out[t][i] =
 * float(bucket[offset[t] + i]) / divisor (rounded as B200DDP_OPT_DIV_MODE says; divisor 1 skips it). */
int b200ddp_unpack(b200ddp_ctx *ctx, b200ddp_plan *plan, const void *bucket,
                   void *const *grads_out, float divisor, b200ddp_stream stream);

/* ---- K3 fused unpack + momentum-SGD -----------------------------------------
 * Replaces `optimizer.step()` of torch.optim.SGD, codes/task2/model.py:131,62
 *   (arithmetic: torch/optim/sgd.py:322-380), fused with the unpack above.
 * Rounding sequence is pinned to torch's: add_(x, alpha=a) is one fma. */
typedef struct {
    double lr, momentum, dampening, weight_decay;
    double grad_divisor; /* g <- g / grad_divisor before anything else; 1 = skip */
    int nesterov, maximize;
    int first_step;      /* momentum buffers are uninitialised: buf <- g */
} b200ddp_sgd_args;

/* gradients come from `bucket`; grads_out (nullable) also receives them, so
 * that p.grad holds the average as it does after the reference's sync. */
int b200ddp_unpack_sgd(b200ddp_ctx *ctx, b200ddp_plan *plan, const void *bucket,
                       void *const *params, void *const *moms, void *const *grads_out,
                       const b200ddp_sgd_args *args, b200ddp_stream stream);
/* gradients come from T separate tensors (world == 1, or after b200ddp_unpack) */
int b200ddp_sgd_step(b200ddp_ctx *ctx, b200ddp_plan *plan, const void *const *grads,
                     void *const *params, void *const *moms, const b200ddp_sgd_args *args,
                     b200ddp_stream stream);

/* ---- K3 with the lab's own optimizers ----------------------------------------
 * Replaces GdOptimizer.step / AdamOptimizer.step, codes/task1/pytorch/MyOptimizer.py:18-43, fused with the
 * unpack like b200ddp_unpack_sgd.  GD: p = p - lr * g.  ADAM (no bias correction, as the lab writes it):
 * m = b1 m + (1 - b1) g;  v = b2 v + (1 - b2) g^2;  p = p - (lr / (sqrt(v) + eps)) m, every tensor
 * operation rounded on its own as torch evaluates that expression.  m and v start as zeros (:32-33) and are
 * the caller's tensors (state1, state2; unused and nullable for GD). */
enum { B200DDP_RULE_ADAM = 1, B200DDP_RULE_GD = 2 };
typedef struct {
    int rule;                      /* B200DDP_RULE_*                                  */
    double lr, beta1, beta2, eps;  /* betas and eps unused for GD                     */
    double grad_divisor;           /* g <- g / grad_divisor first; 1 = skip           */
} b200ddp_opt_args;
int b200ddp_unpack_opt(b200ddp_ctx *ctx, b200ddp_plan *plan, const void *bucket, void *const *params,
                       void *const *state1, void *const *state2, void *const *grads_out,
                       const b200ddp_opt_args *args, b200ddp_stream stream);
int b200ddp_opt_step(b200ddp_ctx *ctx, b200ddp_plan *plan, const void *const *grads, void *const *params,
                     void *const *state1, void *const *state2, const b200ddp_opt_args *args,
                     b200ddp_stream stream);

/* ---- K2+K3 fused: one-shot peer reduction feeding the update directly -------
 * Replaces codes/task2/dist_utils.py:39-42 + codes/task2/model.py:62 in one
 * launch: every rank reads the bucket at offset_bytes from all peers' staging,
 * sums in rank order, divides, updates.  For buckets below the one-shot
 * crossover. */
int b200ddp_allreduce_sgd(b200ddp_ctx *ctx, b200ddp_plan *plan, size_t offset_bytes,
                          void *const *params, void *const *moms, void *const *grads_out,
                          const b200ddp_sgd_args *args, b200ddp_stream stream);

/* ---- single-GPU emulation of the N-rank P2P kernels (tests) -----------------
 * Kernels of different ranks that spin on each other's flags must not be
 * launched separately on one GPU; these run all `n` ranks' blocks as ONE
 * cooperative launch.  ctxs[r] are contexts created on the same device with
 * rank r / world n and linked with b200ddp_set_peers_direct. */
int b200ddp_emu_allreduce(b200ddp_ctx *const *ctxs, int n, size_t offset_bytes, int64_t nelem,
                          int dtype, int algo, b200ddp_stream stream, size_t *result_offset_bytes);
int b200ddp_emu_allreduce_sgd(b200ddp_ctx *const *ctxs, b200ddp_plan *const *plans, int n,
                              size_t offset_bytes, void *const *const *params,
                              void *const *const *moms, void *const *const *grads_out,
                              const b200ddp_sgd_args *args, b200ddp_stream stream);

/* ---- K4 DistributedSampler index shard, built on device ----------------------
 * Replaces `DistributedSampler(dataset, num_replicas, rank)` iteration,
 *   codes/task2/model.py:124 (arithmetic: torch/utils/data/distributed.py:93-141
 *   over randperm_cpu / mt19937), API shape of codes/task3/sampler.py:5-25.
 * out_dev receives num_samples int64 indices, bit-exact with torch.
 * n must be below 2^32/20 (torch switches to 64-bit draws there).
 * Uses scratch owned by ctx (grown with cudaMalloc on first use of a larger n,
 * which synchronises).  The scratch is shared by all calls on this ctx: keep them on one stream (or
 * order the streams) - calls that overlap on the GPU race on it.
 * The call is asynchronous, so a failure inside the kernel cannot come back as a return code: if the
 * shuffle does not finish within its round budget (4096 rounds; the expected depth is ~2.5 log2 n, 37
 * at n = 60,000 - never observed) every element of out_dev is set to -1.  The Python sampler checks
 * for that when it copies the shard to the host. */
int b200ddp_sampler_sizes(int64_t n, int world, int drop_last, int64_t *num_samples,
                          int64_t *total_size);
int b200ddp_sampler_shard(b200ddp_ctx *ctx, int64_t n, int world, int rank, uint64_t seed,
                          uint64_t epoch, int shuffle, int drop_last, int64_t *out_dev,
                          b200ddp_stream stream);

/* ---- allocation-free variant of the sampler shard ----------------------------
 * No counterpart in the reference (its sampler is host Python); added for
 * callers that must not allocate device memory: a host program whose own pool
 * already owns the GPU's memory, or a loop that cannot take cudaMalloc's
 * device-wide stall (say, refilling next epoch's indices while other streams
 * train).  Needs no context.  Same output as b200ddp_sampler_shard, written to
 * the caller's out_dev, on the CURRENT device and the given stream.
 * Scratch is a fixed array inside the library (about 1 MiB per GPU, held until
 * the process exits), hence the limit: n <= B200DDP_SAMPLER_STATIC_MAX_N, beyond
 * which the call fails with B200DDP_ETOOBIG.  It calls no cudaMalloc, cudaFree
 * or allocating library routine.  The first call on each device resolves the
 * array's device address; make that call outside any stream capture.  All calls
 * on a device share the array: keep them on one stream (or order the streams),
 * because calls that overlap on the GPU race on it. */
#define B200DDP_SAMPLER_STATIC_MAX_N 65536
int b200ddp_sampler_shard_static(int64_t n, int world, int rank, uint64_t seed, uint64_t epoch,
                                 int shuffle, int drop_last, int64_t *out_dev, b200ddp_stream stream);

#ifdef __cplusplus
}
#endif
#endif /* B200DDP_H_ */

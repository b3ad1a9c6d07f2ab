// K4: DistributedSampler's per-rank index shard, built on device, bit-exact with
// torch (torch/utils/data/distributed.py:107-141 over randperm_cpu / mt19937).
//
// The permutation is a forward Fisher-Yates: for i in 0..n-2, swap(r[i], r[i + z_i])
// with z_i = draw_i % (n - i), draw_i the i-th tempered MT19937 output.  Taken
// literally that is a chain of n-1 dependent swaps.  Here:
//
//   1. block 0 produces the draws 227 at a time (word k+624 of the MT sequence
//      depends on words k, k+1 and k+397, so 227 consecutive words are
//      independent) and turns them into swap targets t_i = i + z_i, while the
//      other blocks initialise r = identity;
//   2. the swaps run as deterministic reservations (Shun, Gu, Blelloch, Fineman,
//      Gibbons, SODA 2015): in each round every pending iterate i writes its
//      priority to positions i and t_i with atomicMax, and commits its swap only
//      if it holds both - i.e. only if no earlier pending iterate touches either
//      position.  That yields exactly the sequential result; the dependence
//      depth, hence the number of rounds, is O(log n) w.h.p.  Reservation words
//      carry the round number in their high half, so later rounds override
//      earlier ones and the array is never cleared;
//   3. out[k] = r[pad(rank + k * world)], the pad repeating the list from its head.
//
// One cooperative launch; rounds are separated by a grid barrier on a global
// counter.  All cross-block data is read with ld.cg (L2), never from L1.
#include <cooperative_groups.h>

#include <cstdlib>

#include "internal.h"

namespace b200ddp {

namespace {

constexpr int kSamplerThreads = 1024;
constexpr int kMaxRounds = 4096;
constexpr uint32_t kDone = 0xffffffffu;
constexpr int kMtN = 624, kMtM = 397, kMtPar = kMtN - kMtM;   // 227 independent words
constexpr int kRing = 1024;                                    // >= 624 + 227, power of two

struct SamplerScratch {
    uint32_t *tgt;               // n: swap target of iterate i, kDone once committed
    int32_t *perm;               // n
    unsigned long long *resv;    // n: (round << 32) | (0xffffffff - iterate)
    unsigned int *pending;       // kMaxRounds: iterates still pending after round r
    unsigned int *barrier;       // grid barrier arrival counter (zeroed before launch)
};

__device__ __forceinline__ void grid_barrier(unsigned int *counter, unsigned int &phase) {
    __syncthreads();
    if (threadIdx.x == 0) {
        phase += 1;
        const unsigned int target = phase * gridDim.x;
        __threadfence();
        atomicAdd(counter, 1u);
        // all blocks are co-resident (cooperative launch), so this terminates; the poll budget only
        // turns a would-be hang (a bug) into a failed launch
        for (unsigned long long spins = 0; *reinterpret_cast<volatile unsigned int *>(counter) < target; ++spins)
            if (spins > (1ull << 27)) __trap();   // about a minute of L2 polls
        __threadfence();
    }
    __syncthreads();
}

__device__ __forceinline__ uint32_t mt_temper(uint32_t y) {
    y ^= y >> 11;
    y ^= (y << 7) & 0x9d2c5680u;
    y ^= (y << 15) & 0xefc60000u;
    y ^= y >> 18;
    return y;
}

// Block 0 only.  x[] is the MT19937 word sequence: x[0..623] the seeded state,
// x[k+624] = x[k+397] ^ twist(x[k], x[k+1]); draw k = temper(x[k+624]).
__device__ void mt_targets(uint32_t *ring, uint32_t seed, uint32_t n, uint32_t *tgt) {
    const int tid = threadIdx.x;
    if (tid == 0) {
        uint32_t s = seed;
        ring[0] = s;
        for (int i = 1; i < kMtN; ++i) {
            s = 1812433253u * (s ^ (s >> 30)) + (uint32_t)i;
            ring[i] = s;
        }
    }
    __syncthreads();
    const uint32_t ndraw = n - 1;
    for (uint32_t base = 0; base < ndraw; base += kMtPar) {
        const uint32_t k = base + tid;
        if (tid < kMtPar && k < ndraw) {
            const uint32_t a = ring[k & (kRing - 1)], b = ring[(k + 1) & (kRing - 1)], c = ring[(k + kMtM) & (kRing - 1)];
            const uint32_t y = (a & 0x80000000u) | (b & 0x7fffffffu);
            const uint32_t x = c ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
            ring[(k + kMtN) & (kRing - 1)] = x;
            tgt[k] = mt_temper(x);   // raw draw; `% (n - k)` is off this serial path (done in parallel later)
        }
        __syncthreads();
    }
}

__device__ __forceinline__ int64_t padded_source(int64_t t, int64_t n) { return t < n ? t : (t - n) % n; }

__global__ void __launch_bounds__(kSamplerThreads) sampler_shuffle_kernel(SamplerScratch sc, uint32_t n, uint32_t seed,
                                                                         int world, int rank, int64_t num_samples,
                                                                         int64_t *__restrict__ out) {
    __shared__ uint32_t ring[kRing];
    __shared__ unsigned int block_pending;
    const unsigned int nthreads = gridDim.x * kSamplerThreads;
    const unsigned int gtid = blockIdx.x * kSamplerThreads + threadIdx.x;
    unsigned int phase = 0;

    // ---- phase 0: identity permutation, cleared reservations, swap targets
    for (uint32_t i = gtid; i < n; i += nthreads) {
        sc.perm[i] = (int32_t)i;
        sc.resv[i] = 0ull;
    }
    if (blockIdx.x == gridDim.x - 1)
        for (int i = threadIdx.x; i < kMaxRounds; i += kSamplerThreads) sc.pending[i] = 0u;
    if (blockIdx.x == 0) mt_targets(ring, seed, n, sc.tgt);
    grid_barrier(sc.barrier, phase);

    // ---- rounds of reserve / commit over the n-1 iterates
    const uint32_t niter = n - 1;
    // draw -> swap target, in parallel; iterate i is only ever touched by this same thread
    for (uint32_t i = gtid; i < niter; i += nthreads) sc.tgt[i] = i + __ldcg(sc.tgt + i) % (n - i);
    bool overflow = false;
    for (uint32_t round = 1;; ++round) {
        if (round >= (uint32_t)kMaxRounds) { overflow = true; break; }
        for (uint32_t i = gtid; i < niter; i += nthreads) {
            const uint32_t t = __ldcg(sc.tgt + i);
            if (t == kDone) continue;
            const unsigned long long key = ((unsigned long long)round << 32) | (unsigned long long)(0xffffffffu - i);
            atomicMax(sc.resv + i, key);
            if (t != i) atomicMax(sc.resv + t, key);
        }
        grid_barrier(sc.barrier, phase);
        if (round > 1 && __ldcg(sc.pending + (round - 1)) == 0u) break;   // previous round finished the job
        if (threadIdx.x == 0) block_pending = 0u;
        __syncthreads();
        unsigned int mine = 0;
        for (uint32_t i = gtid; i < niter; i += nthreads) {
            const uint32_t t = __ldcg(sc.tgt + i);
            if (t == kDone) continue;
            const unsigned long long key = ((unsigned long long)round << 32) | (unsigned long long)(0xffffffffu - i);
            if (__ldcg(sc.resv + i) == key && __ldcg(sc.resv + t) == key) {
                if (t != i) {
                    const int32_t a = __ldcg(sc.perm + i), b = __ldcg(sc.perm + t);
                    __stcg(sc.perm + i, b);
                    __stcg(sc.perm + t, a);
                }
                __stcg(sc.tgt + i, kDone);
            } else {
                ++mine;
            }
        }
        if (mine) atomicAdd(&block_pending, mine);
        __syncthreads();
        if (threadIdx.x == 0) atomicAdd(sc.pending + round, block_pending);
        // no barrier here: a block that runs ahead into the next reserve phase can only make a
        // slower block's commit check fail spuriously, and that iterate simply retries.
    }

    // ---- this rank's strided, padded slice
    for (int64_t k = gtid; k < num_samples; k += nthreads) {
        const int64_t src = padded_source((int64_t)rank + (int64_t)world * k, (int64_t)n);
        out[k] = overflow ? (int64_t)-1 : (int64_t)__ldcg(sc.perm + src);
    }
}

// ---- n <= 65,535: the same rounds inside one thread-block cluster
//
// The rounds above are latency chains: ~37 of them at n = 60,000 (the reference's MNIST case), each
// paying a software grid barrier (~2 us) and three dependent L2 round trips - 230 us in all, no better
// than the host.  For short lists one cluster of 8 CTAs is enough parallelism, and then the round
// barrier is the hardware cluster barrier and every thread keeps its <= 8 iterates (swap target,
// pending bit) in registers.  Permutation and reservation words stay in global memory (L2).
// Measured dead end: keeping them in distributed shared memory instead is SLOWER (309 us): DSMEM
// sustains ~20 B/cycle/SM, and ~120,000 random 4-byte remote atomics per round saturate it.
constexpr int kClusterCtas = 8;
constexpr int kClusterMaxN = 65535;
constexpr int kClusterIters = (kClusterMaxN + kClusterCtas * kSamplerThreads - 1) / (kClusterCtas * kSamplerThreads);   // 8

__device__ __forceinline__ void bar_sync_256() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

// Warps 0-7 of block 0 only (the other warps go straight to the cluster barrier).
__device__ void mt_targets_256(uint32_t *ring, uint32_t seed, uint32_t n, uint32_t *tgt) {
    const int tid = threadIdx.x;
    if (tid == 0) {
        uint32_t s = seed;
        ring[0] = s;
        for (int i = 1; i < kMtN; ++i) {
            s = 1812433253u * (s ^ (s >> 30)) + (uint32_t)i;
            ring[i] = s;
        }
    }
    bar_sync_256();
    const uint32_t ndraw = n - 1;
    for (uint32_t base = 0; base < ndraw; base += kMtPar) {
        const uint32_t k = base + tid;
        if (tid < kMtPar && k < ndraw) {
            const uint32_t a = ring[k & (kRing - 1)], b = ring[(k + 1) & (kRing - 1)], c = ring[(k + kMtM) & (kRing - 1)];
            const uint32_t y = (a & 0x80000000u) | (b & 0x7fffffffu);
            const uint32_t x = c ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
            ring[(k + kMtN) & (kRing - 1)] = x;
            tgt[k] = mt_temper(x);   // raw draw; `% (n - k)` is off this serial path (done in parallel later)
        }
        bar_sync_256();
    }
}

__global__ void __cluster_dims__(kClusterCtas, 1, 1) __launch_bounds__(kSamplerThreads)
sampler_cluster_kernel(uint32_t *draws, uint32_t *resv, int32_t *perm, uint32_t n, uint32_t seed, int world, int rank,
                       int64_t num_samples, int64_t *__restrict__ out) {
    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    __shared__ uint32_t ring[kRing];
    __shared__ unsigned int pend[4];                                   // block 0's copy is the cluster's
    const unsigned int cta = cluster.block_rank();
    const unsigned int tid = threadIdx.x;
    const unsigned int nthreads = kClusterCtas * kSamplerThreads;
    const unsigned int gtid = cta * kSamplerThreads + tid;

    for (uint32_t pos = gtid; pos < n; pos += nthreads) {
        __stcg(perm + pos, (int32_t)pos);
        __stcg(resv + pos, 0u);
    }
    if (tid < 4) pend[tid] = 0u;
    if (cta == 0 && tid < 256) mt_targets_256(ring, seed, n, draws);
    cluster.sync();

    const uint32_t niter = n - 1;
    uint32_t tg[kClusterIters];
    unsigned int mask = 0;
#pragma unroll
    for (int k = 0; k < kClusterIters; ++k) {
        const uint32_t i = gtid + k * nthreads;
        tg[k] = 0;
        if (i < niter) {
            tg[k] = i + __ldcg(draws + i) % (n - i);
            mask |= 1u << k;
        }
    }
    unsigned int *pend0 = cluster.map_shared_rank(pend, 0);

    bool overflow = false;
    for (uint32_t round = 1;; ++round) {
        // the round number shares the reservation word with the priority (16 bits each); the same bound
        // on every thread, so all CTAs leave together.  Never reached in practice (depth is O(log n)).
        if (round >= (uint32_t)kMaxRounds) { overflow = true; break; }
#pragma unroll
        for (int k = 0; k < kClusterIters; ++k)
            if (mask & (1u << k)) {
                const uint32_t i = gtid + k * nthreads;
                const uint32_t key = (round << 16) | (0xffffu - i);
                atomicMax(resv + i, key);
                if (tg[k] != i) atomicMax(resv + tg[k], key);
            }
        cluster.sync();
        if (round > 1 && *reinterpret_cast<volatile unsigned int *>(pend0 + ((round - 1) & 3)) == 0u) break;
        if (cta == 0 && tid == 0) pend[(round + 1) & 3] = 0u;   // the slot the NEXT round will count into
        // all of this thread's checks in flight together, then all of its swaps: two L2 round trips
        // per round instead of two per iterate.  Measured dead end (round 2): also reading the two
        // permutation entries speculatively with the reservation words (one round trip less for
        // winners) is SLOWER - 218 vs 195 us at n = 60,000: every pending iterate then makes four
        // random L2 reads per round instead of two, and most iterates lose most rounds.
        uint32_t ri[kClusterIters], rt[kClusterIters];
#pragma unroll
        for (int k = 0; k < kClusterIters; ++k)
            if (mask & (1u << k)) {
                ri[k] = __ldcg(resv + gtid + k * nthreads);
                rt[k] = __ldcg(resv + tg[k]);
            }
        unsigned int win = 0;
#pragma unroll
        for (int k = 0; k < kClusterIters; ++k)
            if (mask & (1u << k)) {
                const uint32_t key = (round << 16) | (0xffffu - (gtid + k * nthreads));
                if (ri[k] == key && rt[k] == key) win |= 1u << k;
            }
        int32_t pa[kClusterIters], pb[kClusterIters];
#pragma unroll
        for (int k = 0; k < kClusterIters; ++k)
            if ((win & (1u << k)) && tg[k] != gtid + k * nthreads) {
                pa[k] = __ldcg(perm + gtid + k * nthreads);
                pb[k] = __ldcg(perm + tg[k]);
            }
#pragma unroll
        for (int k = 0; k < kClusterIters; ++k)
            if ((win & (1u << k)) && tg[k] != gtid + k * nthreads) {
                __stcg(perm + gtid + k * nthreads, pb[k]);
                __stcg(perm + tg[k], pa[k]);
            }
        mask &= ~win;
        const unsigned int mine = mask;
        const int waiting = __syncthreads_count(mine != 0);
        if (tid == 0 && waiting) atomicAdd(pend0 + (round & 3), (unsigned int)waiting);
        // as in the grid version, no barrier here: running ahead only causes spurious retries
    }

    for (int64_t k = gtid; k < num_samples; k += nthreads) {
        const uint32_t src = (uint32_t)padded_source((int64_t)rank + (int64_t)world * k, (int64_t)n);
        out[k] = overflow ? (int64_t)-1 : (int64_t)__ldcg(perm + src);
    }
    cluster.sync();   // block 0's `pend` may still be read by the others until here
}

__global__ void sampler_sequential_kernel(int64_t n, int world, int rank, int64_t num_samples, int64_t *__restrict__ out) {
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < num_samples; k += (int64_t)gridDim.x * blockDim.x)
        out[k] = padded_source((int64_t)rank + (int64_t)world * k, n);
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

}  // namespace

size_t sampler_scratch_bytes(int64_t n) {
    return align_up((size_t)n * 4, 256) * 2 + align_up((size_t)n * 8, 256) + align_up(kMaxRounds * 4, 256) + 256;
}

SamplerScratch carve_scratch(void *base, int64_t n) {
    char *p = (char *)base;
    SamplerScratch sc;
    sc.resv = (unsigned long long *)p;  p += align_up((size_t)n * 8, 256);
    sc.tgt = (uint32_t *)p;             p += align_up((size_t)n * 4, 256);
    sc.perm = (int32_t *)p;             p += align_up((size_t)n * 4, 256);
    sc.pending = (unsigned int *)p;     p += align_up(kMaxRounds * 4, 256);
    sc.barrier = (unsigned int *)p;
    return sc;
}

int launch_sampler_on(void *scratch, int sm_count, int64_t n, int world, int rank, uint64_t seed, int64_t num_samples,
                      int64_t *out_dev, cudaStream_t s) {
    SamplerScratch sc = carve_scratch(scratch, n);
    uint32_t n32 = (uint32_t)n, seed32 = (uint32_t)(seed & 0xffffffffu);   // mt19937 uses the low 32 bits
    if (n <= kClusterMaxN && !getenv("B200DDP_SAMPLER_NO_CLUSTER")) {
        // one cluster of 8 CTAs; sc.tgt carries the draws, sc.resv is used as 32-bit words
        cudaGetLastError();
        sampler_cluster_kernel<<<kClusterCtas, kSamplerThreads, 0, s>>>(sc.tgt, reinterpret_cast<uint32_t *>(sc.resv), sc.perm,
                                                                        n32, seed32, world, rank, num_samples, out_dev);
        B2_CHECK_CUDA(cudaGetLastError());
        return B200DDP_OK;
    }
    int per_sm = 0;
    B2_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, sampler_shuffle_kernel, kSamplerThreads, 0));
    B2_REQUIRE(per_sm >= 1, B200DDP_ESTATE, "sampler kernel does not fit on an SM");
    int64_t want = (n + 2 * kSamplerThreads - 1) / (2 * kSamplerThreads);   // >= 2 iterates per thread
    int grid = (int)(want < 1 ? 1 : want);
    if (grid > sm_count * per_sm) grid = sm_count * per_sm;
    if (grid > sm_count) grid = sm_count;   // one block per SM keeps the barrier cheap
    B2_CHECK_CUDA(cudaMemsetAsync(sc.barrier, 0, sizeof(unsigned int), s));
    void *params[] = {&sc, &n32, &seed32, &world, &rank, &num_samples, &out_dev};
    B2_CHECK_CUDA(cudaLaunchCooperativeKernel((const void *)sampler_shuffle_kernel, dim3(grid), dim3(kSamplerThreads),
                                              params, 0, s));
    return B200DDP_OK;
}

int launch_sampler(b200ddp_ctx *ctx, int64_t n, int world, int rank, uint64_t seed, int shuffle, int64_t num_samples,
                   int64_t *out_dev, cudaStream_t s) {
    if (!shuffle) {
        int grid = (int)((num_samples + 255) / 256);
        if (grid > 4 * ctx->sm_count) grid = 4 * ctx->sm_count;
        cudaGetLastError();
        sampler_sequential_kernel<<<grid, 256, 0, s>>>(n, world, rank, num_samples, out_dev);
        B2_CHECK_CUDA(cudaGetLastError());
        return B200DDP_OK;
    }
    const size_t need = sampler_scratch_bytes(n);
    if (need > ctx->sampler_scratch_bytes) {
        // grow: the old buffer may still be in use by an earlier call on another stream
        B2_CHECK_CUDA(cudaDeviceSynchronize());
        if (ctx->sampler_scratch) B2_CHECK_CUDA(cudaFree(ctx->sampler_scratch));
        ctx->sampler_scratch = nullptr;
        ctx->sampler_scratch_bytes = 0;
        B2_CHECK_CUDA(cudaMalloc(&ctx->sampler_scratch, need));
        ctx->sampler_scratch_bytes = need;
    }
    return launch_sampler_on(ctx->sampler_scratch, ctx->sm_count, n, world, rank, seed, num_samples, out_dev, s);
}

}  // namespace b200ddp

// =====================================================================================
// b200ddp_sampler_shard_static: the same shard, for callers that must not allocate.
//
// Who that is: a host program that has already handed the GPU's memory to its own
// pool (any further cudaMalloc would fail, or take memory the pool counts on), and
// a loop that cannot afford cudaMalloc's device-wide stall - e.g. refilling the
// sampler for the next epoch while training kernels of other streams are running.
// The regular entry point needs a context (which allocates staging) and grows its
// scratch with cudaMalloc; this one needs neither.
//
// Its scratch (swap targets, permutation, reservation words, round counters, the
// grid-barrier counter) lives in the fixed array below, which every GPU that loads
// the library carries until the process exits.  Limits and rules for the caller:
//   * n <= B200DDP_SAMPLER_STATIC_MAX_N (65,536: MNIST's 60,000 and CIFAR's 50,000
//     fit); a larger n is refused with B200DDP_ETOOBIG, never truncated.
//   * The first call on each device looks up the array's device address
//     (cudaGetSymbolAddress) and the SM count.  Make that first call outside any
//     stream capture.
//   * All calls on one device share the one array.  Nothing a previous call left in
//     it is relied on: the kernel rewrites targets and permutation and zeroes the
//     reservation words and round counters itself, and the barrier counter is
//     cleared on the stream just before the launch.  Two calls on one device whose
//     kernels overlap in time (issued on different streams with no ordering between
//     them) race on the array and corrupt each other: issue them on one stream, or
//     order the streams yourself.  Another process has its own copy of the array.
//   * No cudaMalloc / cudaFree / allocating library call is made, here or below.
// =====================================================================================
namespace {

constexpr int64_t kStaticMaxN = B200DDP_SAMPLER_STATIC_MAX_N;
constexpr size_t kStaticScratchBytes =
    ((size_t)kStaticMaxN * 8 + 255) / 256 * 256 + 2 * (((size_t)kStaticMaxN * 4 + 255) / 256 * 256) +
    ((size_t)b200ddp::kMaxRounds * 4 + 255) / 256 * 256 + 256;

// Device storage.  In host code this name is only a handle for the runtime's symbol
// table - it is NOT a device address and must never be passed to a kernel as one.
static __device__ __align__(256) unsigned char g_static_sampler_scratch[kStaticScratchBytes];

constexpr int kMaxDevices = 64;
struct StaticSlot {
    void *scratch = nullptr;   // device address of g_static_sampler_scratch on this device
    int sm_count = 0;
};
StaticSlot g_static_slots[kMaxDevices];
std::mutex g_static_mu;

// Once per device, with that device current.
int static_slot_for_current_device(StaticSlot *out) {
    int dev = -1;
    B2_CHECK_CUDA(cudaGetDevice(&dev));
    B2_REQUIRE(dev >= 0 && dev < kMaxDevices, B200DDP_EINVAL, "device ordinal %d outside [0, %d)", dev, kMaxDevices);
    std::lock_guard<std::mutex> lk(g_static_mu);
    StaticSlot &slot = g_static_slots[dev];
    if (slot.scratch == nullptr) {
        void *ptr = nullptr;
        B2_CHECK_CUDA(cudaGetSymbolAddress(&ptr, g_static_sampler_scratch));
        B2_REQUIRE(ptr != nullptr, B200DDP_ECUDA, "cudaGetSymbolAddress returned a null address");
        int sms = 0;
        B2_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        B2_REQUIRE(sms > 0, B200DDP_ECUDA, "device %d reports %d SMs", dev, sms);
        slot.sm_count = sms;
        slot.scratch = ptr;
    }
    *out = slot;
    return B200DDP_OK;
}

}  // namespace

extern "C" int b200ddp_sampler_shard_static(int64_t n, int world, int rank, uint64_t seed, uint64_t epoch,
                                            int shuffle, int drop_last, int64_t *out_dev, b200ddp_stream stream) {
    using namespace b200ddp;
    B2_REQUIRE(world >= 1, B200DDP_EINVAL, "num_replicas must be positive");
    B2_REQUIRE(rank >= 0 && rank < world, B200DDP_EINVAL, "Invalid rank %d, rank should be in the interval [0, %d]",
               rank, world - 1);
    B2_REQUIRE(n >= 0, B200DDP_EINVAL, "negative dataset length");
    B2_REQUIRE(n <= kStaticMaxN, B200DDP_ETOOBIG,
               "dataset length %lld exceeds the %lld indices the allocation-free sampler has room for; "
               "use b200ddp_sampler_shard",
               (long long)n, (long long)kStaticMaxN);
    int64_t num = 0;
    if (int rc = b200ddp_sampler_sizes(n, world, drop_last, &num, nullptr)) return rc;
    if (num <= 0 || n == 0) return B200DDP_OK;
    B2_REQUIRE(out_dev != nullptr, B200DDP_EINVAL, "out_dev is NULL");
    B2_REQUIRE(sampler_scratch_bytes(n) <= kStaticScratchBytes, B200DDP_ETOOBIG, "static scratch too small for n=%lld",
               (long long)n);
    StaticSlot slot;
    if (int rc = static_slot_for_current_device(&slot)) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    if (!shuffle) {
        int grid = (int)((num + 255) / 256);
        if (grid > 4 * slot.sm_count) grid = 4 * slot.sm_count;
        cudaGetLastError();
        sampler_sequential_kernel<<<grid, 256, 0, s>>>(n, world, rank, num, out_dev);
        B2_CHECK_CUDA(cudaGetLastError());
        return B200DDP_OK;
    }
    return launch_sampler_on(slot.scratch, slot.sm_count, n, world, rank, seed + epoch, num, out_dev, s);
}

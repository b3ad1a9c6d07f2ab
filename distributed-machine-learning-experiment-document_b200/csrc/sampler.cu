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
        while (*reinterpret_cast<volatile unsigned int *>(counter) < target) {
        }
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
            tgt[k] = k + mt_temper(x) % (n - k);
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
    int per_sm = 0;
    B2_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, sampler_shuffle_kernel, kSamplerThreads, 0));
    B2_REQUIRE(per_sm >= 1, B200DDP_ESTATE, "sampler kernel does not fit on an SM");
    int64_t want = (n + 2 * kSamplerThreads - 1) / (2 * kSamplerThreads);   // >= 2 iterates per thread
    int grid = (int)(want < 1 ? 1 : want);
    if (grid > sm_count * per_sm) grid = sm_count * per_sm;
    if (grid > sm_count) grid = sm_count;   // one block per SM keeps the barrier cheap
    B2_CHECK_CUDA(cudaMemsetAsync(sc.barrier, 0, sizeof(unsigned int), s));
    uint32_t n32 = (uint32_t)n, seed32 = (uint32_t)(seed & 0xffffffffu);   // mt19937 uses the low 32 bits
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

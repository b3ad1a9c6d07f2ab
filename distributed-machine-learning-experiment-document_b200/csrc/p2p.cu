// K2: all-reduce of a staging bucket by direct peer loads/stores over NVLink,
// and the fused one-shot-reduce + momentum-SGD kernel.
//
// Every rank launches the same grid.  Block b of rank r synchronises with block
// b of every peer through flags in peer memory: flag[b][src] on rank dst is
// written only by block b of rank src, with a generation number that increases
// by two per launch (never reset, so no host sync between launches).
//
//   one-shot : barrier | every rank reads the whole bucket of all N ranks, sums
//              in rank order (so all ranks get identical bits), writes the result
//              out of place | barrier (peers are done reading my bucket)
//   two-shot : barrier | rank r reduces slice r of all N buckets and stores the
//              sums into slice r of all N buckets | barrier (all slices landed)
//
// EMU = true runs all N ranks as blockIdx.y of ONE cooperative launch on one
// GPU: kernels that spin on each other's flags must not be separate launches
// on the same device.
#include <cuda_bf16.h>

#include "device_utils.cuh"

namespace b200ddp {

namespace {

constexpr int kP2PThreads = 512;
constexpr int kNvlsDepthDefault = 1;   // multimem loads in flight per thread  } profiles/r02_sweep_n{2,4,8}_fp32_variants.jsonl:
constexpr int kNvlsBlocksDefault = 74;   // grid cap of the in-switch kernel      } fewer and shallower is faster

__device__ __forceinline__ void st_release_sys(uint32_t *p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// Peer data: plain weak loads, ordered behind the barrier's acquire; no L1 allocation.
__device__ __forceinline__ uint4 ld_peer_u4(const void *p) {
    uint4 r;
    asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p)
                 : "memory");
    return r;
}
__device__ __forceinline__ void st_peer_u4(void *p, const uint4 &v) {
    asm volatile("st.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// All threads of the block call this.  Orders every earlier write of the block
// (release, cumulative over the bar.sync) before the flag, and every later read
// after the peers' flags (acquire, then bar.sync).
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// By default a block waits for its peers for as long as it takes, as an NCCL collective does: a
// legitimately slow rank (evaluation on rank 0, a data-loader stall, the lab's straggler experiment)
// must not crash the others.  B200DDP_OPT_PEER_TIMEOUT_MS > 0 turns a wait longer than that into a
// trap, which fails the launch with an error the host sees (hang detection for debugging).
__device__ __forceinline__ void peer_barrier(const PeerTable &pt, int rank, int world, int bid, uint32_t gen,
                                             unsigned long long timeout_ns) {
    __syncthreads();
    if ((int)threadIdx.x < world) {
        const int peer = threadIdx.x;
        st_release_sys(pt.flags[peer] + bid * B200DDP_MAX_WORLD + rank, gen);
        const uint32_t *mine = pt.flags[rank] + bid * B200DDP_MAX_WORLD + peer;
        unsigned long long t0 = 0;
        for (unsigned int spins = 0; (int32_t)(ld_acquire_sys(mine) - gen) < 0; ++spins) {
            if (timeout_ns != 0 && (spins & 0xfffu) == 0xfffu) {          // look at the clock every 4096 polls
                const unsigned long long now = global_ns();
                if (t0 == 0) t0 = now;
                else if (now - t0 > timeout_ns) __trap();
            }
        }
    }
    __syncthreads();
}

// Optional per-block time stamps (B200DDP_OPT_TRACE): entry, first barrier passed, data moved,
// last barrier passed.  ncu cannot replay a kernel that waits for other GPUs, so this is how the
// phases of the cross-GPU kernels are measured (tools/p2p_trace.py).
__device__ __forceinline__ void trace_stamp(unsigned long long *trace, int bid, int slot) {
    if (trace != nullptr && threadIdx.x == 0) trace[bid * 4 + slot] = global_ns();
}

// ---- 16-byte vector arithmetic in fp32 ---------------------------------------
template <typename BT> struct Vec16;
template <> struct Vec16<float> {
    static constexpr int kAcc = 4;
    static __device__ __forceinline__ void zero(float *a) { a[0] = a[1] = a[2] = a[3] = 0.f; }
    static __device__ __forceinline__ void set(float *a, const uint4 &v) {
        a[0] = __uint_as_float(v.x); a[1] = __uint_as_float(v.y); a[2] = __uint_as_float(v.z); a[3] = __uint_as_float(v.w);
    }
    static __device__ __forceinline__ void add(float *a, const uint4 &v) {
        a[0] = __fadd_rn(a[0], __uint_as_float(v.x)); a[1] = __fadd_rn(a[1], __uint_as_float(v.y));
        a[2] = __fadd_rn(a[2], __uint_as_float(v.z)); a[3] = __fadd_rn(a[3], __uint_as_float(v.w));
    }
    static __device__ __forceinline__ uint4 pack(const float *a) {
        return make_uint4(__float_as_uint(a[0]), __float_as_uint(a[1]), __float_as_uint(a[2]), __float_as_uint(a[3]));
    }
};
template <> struct Vec16<__nv_bfloat16> {
    static constexpr int kAcc = 8;
    static __device__ __forceinline__ void set(float *a, const uint4 &v) {
        const unsigned w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) { a[2 * i] = __uint_as_float(w[i] << 16); a[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u); }
    }
    static __device__ __forceinline__ void add(float *a, const uint4 &v) {
        const unsigned w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            a[2 * i] = __fadd_rn(a[2 * i], __uint_as_float(w[i] << 16));
            a[2 * i + 1] = __fadd_rn(a[2 * i + 1], __uint_as_float(w[i] & 0xffff0000u));
        }
    }
    static __device__ __forceinline__ uint4 pack(const float *a) {   // one rounding, after the fp32 sum
        unsigned w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            __nv_bfloat162 h = __floats2bfloat162_rn(a[2 * i], a[2 * i + 1]);
            w[i] = *reinterpret_cast<unsigned *>(&h);
        }
        return make_uint4(w[0], w[1], w[2], w[3]);
    }
};

struct P2PArgs {
    PeerTable pt;
    int rank, world;
    size_t off_bytes;       // bucket inside the first staging half
    size_t result_off;      // one-shot: where the sums go (own staging)
    int64_t nvec;           // 16-byte vectors in the bucket
    // Barrier generation of block b of rank r lives in gen[r][b] (device memory, local to rank r)
    // and is advanced by the kernel itself: a launch replayed from a CUDA graph must not carry
    // a generation number frozen at capture time.  EMU fills all ranks, a real launch only its own.
    uint32_t *gen[B200DDP_MAX_WORLD];
    unsigned long long timeout_ns;   // 0: wait for ever
    unsigned long long *trace;       // nullable: 4 stamps per block
};

// Sum of vector i over all ranks, in rank order, in fp32.
// N > 0: unrolled for that world size; N == 0: runtime loop over args.world.
// V vectors per thread are in flight together: V x N independent 16-byte peer loads, all issued
// before the first add, so that one NVLink round trip (~2 us) covers V x N x 16 bytes per thread.
template <typename BT, int N, int V>
__device__ __forceinline__ void reduce_vectors(const P2PArgs &a, int world, const int64_t (&idx)[V], int64_t hi,
                                               uint4 (&out)[V]) {
    if (N > 0) {
        uint4 v[V][N > 0 ? N : 1];
#pragma unroll
        for (int u = 0; u < V; ++u)
            if (idx[u] < hi) {
#pragma unroll
                for (int r = 0; r < N; ++r)
                    v[u][r] = ld_peer_u4(reinterpret_cast<const uint4 *>(a.pt.staging[r] + a.off_bytes) + idx[u]);
            }
#pragma unroll
        for (int u = 0; u < V; ++u)
            if (idx[u] < hi) {
                float acc[Vec16<BT>::kAcc];
                Vec16<BT>::set(acc, v[u][0]);
#pragma unroll
                for (int r = 1; r < N; ++r) Vec16<BT>::add(acc, v[u][r]);
                out[u] = Vec16<BT>::pack(acc);
            }
    } else {
#pragma unroll
        for (int u = 0; u < V; ++u)
            if (idx[u] < hi) {
                float acc[Vec16<BT>::kAcc];
                Vec16<BT>::set(acc, ld_peer_u4(reinterpret_cast<const uint4 *>(a.pt.staging[0] + a.off_bytes) + idx[u]));
                for (int r = 1; r < world; ++r)
                    Vec16<BT>::add(acc, ld_peer_u4(reinterpret_cast<const uint4 *>(a.pt.staging[r] + a.off_bytes) + idx[u]));
                out[u] = Vec16<BT>::pack(acc);
            }
    }
}

// vectors in flight per thread, by world size: V x N loads of 16 bytes
template <int N> struct P2PUnroll { static constexpr int kV = N == 2 ? 4 : N == 4 ? 4 : N == 8 ? 2 : 1; };

template <typename BT, int N, bool EMU>
__global__ void __launch_bounds__(kP2PThreads) oneshot_kernel(P2PArgs a) {
    constexpr int V = P2PUnroll<N>::kV;
    const int rank = EMU ? (int)blockIdx.y : a.rank;
    const int world = N > 0 ? N : a.world;
    const int bid = blockIdx.x;
    unsigned long long *trace = EMU ? nullptr : a.trace;
    pdl_prologue();
    trace_stamp(trace, bid, 0);
    const uint32_t gen = a.gen[rank][bid];
    peer_barrier(a.pt, rank, world, bid, gen + 1, a.timeout_ns);
    trace_stamp(trace, bid, 1);
    uint4 *dst = reinterpret_cast<uint4 *>(a.pt.staging[rank] + a.result_off);
    const int64_t stride = (int64_t)gridDim.x * kP2PThreads;
    for (int64_t i0 = (int64_t)bid * kP2PThreads + threadIdx.x; i0 < a.nvec; i0 += stride * V) {
        int64_t idx[V];
        uint4 out[V];
#pragma unroll
        for (int u = 0; u < V; ++u) idx[u] = i0 + u * stride;
        reduce_vectors<BT, N, V>(a, world, idx, a.nvec, out);
#pragma unroll
        for (int u = 0; u < V; ++u)
            if (idx[u] < a.nvec) dst[idx[u]] = out[u];
    }
    trace_stamp(trace, bid, 2);
    peer_barrier(a.pt, rank, world, bid, gen + 2, a.timeout_ns);
    trace_stamp(trace, bid, 3);
    if (threadIdx.x == 0) a.gen[rank][bid] = gen + 2;
}

template <typename BT, int N, bool EMU>
__global__ void __launch_bounds__(kP2PThreads) twoshot_kernel(P2PArgs a) {
    constexpr int V = P2PUnroll<N>::kV;
    const int rank = EMU ? (int)blockIdx.y : a.rank;
    const int world = N > 0 ? N : a.world;
    const int bid = blockIdx.x;
    unsigned long long *trace = EMU ? nullptr : a.trace;
    pdl_prologue();
    trace_stamp(trace, bid, 0);
    const uint32_t gen = a.gen[rank][bid];
    peer_barrier(a.pt, rank, world, bid, gen + 1, a.timeout_ns);
    trace_stamp(trace, bid, 1);
    const int64_t per = (a.nvec + world - 1) / world;
    const int64_t lo = per * rank;
    const int64_t hi = lo + per < a.nvec ? lo + per : a.nvec;
    const int64_t stride = (int64_t)gridDim.x * kP2PThreads;
    for (int64_t i0 = lo + (int64_t)bid * kP2PThreads + threadIdx.x; i0 < hi; i0 += stride * V) {
        int64_t idx[V];
        uint4 out[V];
#pragma unroll
        for (int u = 0; u < V; ++u) idx[u] = i0 + u * stride;
        reduce_vectors<BT, N, V>(a, world, idx, hi, out);
#pragma unroll
        for (int u = 0; u < V; ++u)
            if (idx[u] < hi) {
                if (N > 0) {
#pragma unroll
                    for (int r = 0; r < N; ++r)
                        st_peer_u4(reinterpret_cast<uint4 *>(a.pt.staging[r] + a.off_bytes) + idx[u], out[u]);
                } else {
                    for (int r = 0; r < world; ++r)
                        st_peer_u4(reinterpret_cast<uint4 *>(a.pt.staging[r] + a.off_bytes) + idx[u], out[u]);
                }
            }
    }
    trace_stamp(trace, bid, 2);
    peer_barrier(a.pt, rank, world, bid, gen + 2, a.timeout_ns);
    trace_stamp(trace, bid, 3);
    if (threadIdx.x == 0) a.gen[rank][bid] = gen + 2;
}

// ---- in-switch reduction (NVLink SHARP) ---------------------------------------------
// mc is a multicast address over every rank's staging: one multimem.ld_reduce makes the switch
// fetch the 16 bytes from all N GPUs and return their sum (fp32 accumulation), one multimem.st
// stores to all N.  Rank r handles slice r, like the two-shot kernel, but every GPU receives
// S/N + S instead of 2 (N-1)/N x S and sends as little.  The order of the additions inside the
// switch is not specified: results are identical on all ranks (each element is reduced once and
// broadcast) and exact for N = 2, within the fp32 bound otherwise.
template <typename BT> struct Multimem;
template <> struct Multimem<float> {
    static __device__ __forceinline__ uint4 ld_reduce(const char *p) {
        uint4 r;
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
        return r;
    }
    static __device__ __forceinline__ void st(char *p, const uint4 &v) {
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                     :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
    }
};
template <> struct Multimem<__nv_bfloat16> {
    static __device__ __forceinline__ uint4 ld_reduce(const char *p) {
        uint4 r;
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                     : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
        return r;
    }
    static __device__ __forceinline__ void st(char *p, const uint4 &v) {
        asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1,%2,%3,%4};"
                     :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
    }
};

// Every thread keeps a ring of D loads in flight and, slot by slot, stores a result and at once
// issues the load that reuses the slot, so reduce traffic (outbound: every GPU feeds S to the
// switch) and broadcast traffic (inbound: every GPU receives S) share the links instead of taking
// turns.  What the round-2 sweeps on 2, 4 and 8 GPUs say about D and the grid
// (profiles/r02_sweep_n*_fp32_variants.jsonl, with per-phase time stamps): LESS is faster.  Round 1
// ran D = 4 on 296 blocks: 146.9 us for the 44.7 MB ResNet-18 bucket on 8 GPUs; D = 1 on 74 blocks:
// 121.6 us (37 blocks: 119.4), data phase 105 us = 430 GB/s per GPU through the switch, which is
// also where NCCL's own NVLS path levels off at 1 GiB.  More requests in flight only queue in the
// switch, and their barriers (one flag word per block per peer) get dearer.
template <typename BT, int D>
__global__ void __launch_bounds__(kP2PThreads) nvls_kernel(P2PArgs a, char *mc) {
    const int rank = a.rank, world = a.world, bid = blockIdx.x;
    pdl_prologue();
    trace_stamp(a.trace, bid, 0);
    const uint32_t gen = a.gen[rank][bid];
    peer_barrier(a.pt, rank, world, bid, gen + 1, a.timeout_ns);
    trace_stamp(a.trace, bid, 1);
    const int64_t per = (a.nvec + world - 1) / world;
    const int64_t lo = per * rank;
    const int64_t hi = lo + per < a.nvec ? lo + per : a.nvec;
    const int64_t stride = (int64_t)gridDim.x * kP2PThreads;
    char *base = mc + a.off_bytes;
    const int64_t first = lo + (int64_t)bid * kP2PThreads + threadIdx.x;
    uint4 v[D];
#pragma unroll
    for (int u = 0; u < D; ++u) {
        const int64_t i = first + u * stride;
        if (i < hi) v[u] = Multimem<BT>::ld_reduce(base + i * 16);
    }
    for (int64_t i0 = first; i0 < hi; i0 += stride * D) {
#pragma unroll
        for (int u = 0; u < D; ++u) {
            const int64_t i = i0 + u * stride;
            if (i < hi) {
                Multimem<BT>::st(base + i * 16, v[u]);
                const int64_t nxt = i + stride * D;
                if (nxt < hi) v[u] = Multimem<BT>::ld_reduce(base + nxt * 16);
            }
        }
    }
    __threadfence_system();   // multimem stores are weak: order them before the barrier's release
    trace_stamp(a.trace, bid, 2);
    peer_barrier(a.pt, rank, world, bid, gen + 2, a.timeout_ns);
    trace_stamp(a.trace, bid, 3);
    if (threadIdx.x == 0) a.gen[rank][bid] = gen + 2;
}

// ---- a barrier on its own ------------------------------------------------------------------
// One block: "everything this rank enqueued on the stream before this launch is visible to every
// rank's work enqueued after it".  Used by the flat initial-weight broadcast (rank 0 packs, barrier,
// the others read rank 0's staging, barrier) in place of host synchronisation + a process-group barrier.
__global__ void __launch_bounds__(32) barrier_kernel(P2PArgs a) {
    pdl_prologue();
    const uint32_t gen = a.gen[a.rank][0];
    peer_barrier(a.pt, a.rank, a.world, 0, gen + 1, a.timeout_ns);
    if (threadIdx.x == 0) a.gen[a.rank][0] = gen + 1;
}

// ---- fused one-shot reduce + update --------------------------------------------
// Tensor tables come by value in the kernel parameters (see TensorTab); under EMU every
// emulated rank has its own table and its own chunk list.
template <bool EMU, int CAP> struct TabSet {
    TensorTab<CAP, kNumPtrKinds> t[EMU ? B200DDP_MAX_WORLD : 1];
    const Chunk *chunks[EMU ? B200DDP_MAX_WORLD : 1];
};

template <typename BT, bool WRITE_GRAD, bool EMU, int CAP>
__global__ void __launch_bounds__(kThreads) oneshot_update_kernel(P2PArgs a, const __grid_constant__ TabSet<EMU, CAP> ts,
                                                                  int nchunks, SgdScalars hp) {
    const int rank = EMU ? (int)blockIdx.y : a.rank;
    const int world = a.world;
    const int bid = blockIdx.x;
    const TensorTab<CAP, kNumPtrKinds> &tab = ts.t[EMU ? rank : 0];
    const Chunk *chunks = ts.chunks[EMU ? rank : 0];
    const int tid = threadIdx.x;
    const bool load_m = hp.has_momentum && !hp.first_step;
    unsigned long long *trace = EMU ? nullptr : a.trace;
    pdl_prologue();
    trace_stamp(trace, bid, 0);
    const uint32_t gen = a.gen[rank][bid];
    peer_barrier(a.pt, rank, world, bid, gen + 1, a.timeout_ns);
    trace_stamp(trace, bid, 1);
    for (int c = bid; c < nchunks; c += gridDim.x) {
        const Chunk ck = chunks[c];
        const int64_t e0 = (int64_t)ck.index * kChunk;
        const int64_t left = tab.numel[ck.tensor] - e0;
        const int count = left < kChunk ? (int)left : kChunk;
        const int64_t boff = tab.boff[ck.tensor] + e0;
        float *p = (float *)tab.ptr[kParam][ck.tensor] + e0;
        float *m = hp.has_momentum ? (float *)tab.ptr[kMom][ck.tensor] + e0 : nullptr;
        float *go = WRITE_GRAD ? (float *)tab.ptr[kGradOut][ck.tensor] + e0 : nullptr;
        const bool aligned = (((uintptr_t)p | (uintptr_t)m | (uintptr_t)go) & 15) == 0;   // bucket side is aligned by layout
        const int nvec = aligned ? count / kVec : 0;
        for (int i = tid; i < nvec; i += kThreads) {
            const int o = i * kVec;
            // all peers' loads are issued before the first add (NVLink round trips overlap)
            float4 gr[B200DDP_MAX_WORLD];
#pragma unroll
            for (int r = 0; r < B200DDP_MAX_WORLD; ++r)
                if (r < world) gr[r] = Bucket<BT>::load4c(reinterpret_cast<const BT *>(a.pt.staging[r] + a.off_bytes) + boff + o);
            float4 gs = gr[0];
#pragma unroll
            for (int r = 1; r < B200DDP_MAX_WORLD; ++r)
                if (r < world) {
                    gs.x = __fadd_rn(gs.x, gr[r].x); gs.y = __fadd_rn(gs.y, gr[r].y);
                    gs.z = __fadd_rn(gs.z, gr[r].z); gs.w = __fadd_rn(gs.w, gr[r].w);
                }
            float4 p4 = ld_rmw_f4(p + o), m4, ga;
            if (load_m) m4 = ld_rmw_f4(m + o);
            sgd_elem4(gs, p4, m4, ga, hp);
            st_f4(p + o, p4);
            if (hp.has_momentum) st_f4(m + o, m4);
            if (WRITE_GRAD) st_f4(go + o, ga);
        }
        for (int i = nvec * kVec + tid; i < count; i += kThreads) {
            float g1[B200DDP_MAX_WORLD];
#pragma unroll
            for (int r = 0; r < B200DDP_MAX_WORLD; ++r)
                if (r < world) g1[r] = Bucket<BT>::load1c(reinterpret_cast<const BT *>(a.pt.staging[r] + a.off_bytes) + boff + i);
            float gs = g1[0];
#pragma unroll
            for (int r = 1; r < B200DDP_MAX_WORLD; ++r)
                if (r < world) gs = __fadd_rn(gs, g1[r]);
            float ps1 = p[i], ms = load_m ? m[i] : 0.f, ga;
            sgd_elem(gs, ps1, ms, ga, hp);
            p[i] = ps1;
            if (hp.has_momentum) m[i] = ms;
            if (WRITE_GRAD) go[i] = ga;
        }
    }
    trace_stamp(trace, bid, 2);
    peer_barrier(a.pt, rank, world, bid, gen + 2, a.timeout_ns);
    trace_stamp(trace, bid, 3);
    if (threadIdx.x == 0) a.gen[rank][bid] = gen + 2;
}

#define B2_LAUNCH_RC(expr)                                                                              \
    do {                                                                                                \
        cudaError_t _e = (expr);                                                                        \
        if (_e != cudaSuccess) {                                                                        \
            set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e), __FILE__, __LINE__);  \
            cudaGetLastError();                                                                         \
            return B200DDP_ECUDA;                                                                       \
        }                                                                                               \
    } while (0)

typedef void (*P2PKernel)(P2PArgs);
template <typename BT, int N, bool EMU> P2PKernel pick_kernel(int algo) {
    return algo == B200DDP_ALGO_ONESHOT ? oneshot_kernel<BT, N, EMU> : twoshot_kernel<BT, N, EMU>;
}
template <typename BT, bool EMU> P2PKernel pick_world(int algo, int world) {
    switch (world) {
        case 2: return pick_kernel<BT, 2, EMU>(algo);
        case 4: return pick_kernel<BT, 4, EMU>(algo);
        case 8: return pick_kernel<BT, 8, EMU>(algo);
        default: return pick_kernel<BT, 0, EMU>(algo);
    }
}
template <bool EMU> P2PKernel pick(int dtype, int algo, int world) {
    return dtype == B200DDP_F32 ? pick_world<float, EMU>(algo, world) : pick_world<__nv_bfloat16, EMU>(algo, world);
}
P2PArgs make_args(const b200ddp_ctx *ctx, size_t off, int64_t nelem, int dtype) {
    P2PArgs a;
    a.pt = ctx->peers;
    a.rank = ctx->rank;
    a.world = ctx->world;
    a.off_bytes = off;
    a.result_off = ctx->staging_bytes + off;
    a.nvec = nelem * (dtype == B200DDP_BF16 ? 2 : 4) / 16;
    for (int r = 0; r < B200DDP_MAX_WORLD; ++r) a.gen[r] = nullptr;
    a.gen[ctx->rank] = ctx->block_gen;
    a.timeout_ns = ctx->peer_timeout_ms > 0 ? (unsigned long long)ctx->peer_timeout_ms * 1000000ull : 0ull;
    a.trace = ctx->trace;
    return a;
}

template <int CAP>
void fill_update_tab(TensorTab<CAP, kNumPtrKinds> &tab, const b200ddp_plan *plan, const Segment &sg, const PtrArrays &pa) {
    const int n = sg.t1 - sg.t0;
    for (int i = 0; i < n; ++i) {
        tab.numel[i] = plan->meta[sg.t0 + i].numel;
        tab.boff[i] = plan->meta[sg.t0 + i].bucket_off;
    }
    for (int k = 0; k < kNumPtrKinds; ++k)
        for (int i = 0; i < n; ++i) tab.ptr[k][i] = pa.k[k] ? const_cast<void *>(pa.k[k][sg.t0 + i]) : nullptr;
}

int block_cap(const b200ddp_ctx *ctx) {
    int cap = ctx->p2p_blocks > 0 ? ctx->p2p_blocks : 2 * ctx->sm_count;
    return cap < kMaxBlocksP2P ? cap : kMaxBlocksP2P;
}
// Blocks that spin on their peers should all be resident: never launch more than the GPU can hold
// at once (same GPUs on every rank, so every rank computes the same clamp).
template <typename K> int resident_clamp(const b200ddp_ctx *ctx, K kernel, int threads, int grid) {
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, 0) != cudaSuccess || per_sm < 1) per_sm = 1;
    const int fit = per_sm * ctx->sm_count;
    return grid < fit ? grid : fit;
}
int update_grid(const b200ddp_ctx *ctx, int nchunks) {
    const int cap = block_cap(ctx);
    return nchunks < cap ? nchunks : cap;
}

template <typename BT, bool WG, int CAP>
int run_fused(b200ddp_ctx *ctx, const b200ddp_plan *plan, const Segment &sg, size_t off, const PtrArrays &pa,
              const SgdScalars &hp, cudaStream_t s) {
    P2PArgs a = make_args(ctx, off, 0, plan->dtype);
    TabSet<false, CAP> ts;
    fill_update_tab(ts.t[0], plan, sg, pa);
    ts.chunks[0] = plan->d_chunks + sg.c0;
    const int n = sg.c1 - sg.c0;
    const int grid = resident_clamp(ctx, oneshot_update_kernel<BT, WG, false, CAP>, kThreads, update_grid(ctx, n));
    B2_LAUNCH_RC(launch_ex(oneshot_update_kernel<BT, WG, false, CAP>, grid, kThreads, s, ctx->pdl, a, ts, n, hp));
    return B200DDP_OK;
}

int nvls_depth(const b200ddp_ctx *ctx) {
    const int d = ctx->nvls_depth;
    return (d == 1 || d == 2 || d == 4 || d == 8) ? d : kNvlsDepthDefault;
}

}  // namespace

// Same on every rank: a pure function of the bucket, the world size and settings that the caller
// keeps equal on all ranks.
int p2p_grid_for(int64_t nelem, int dtype, int algo, int world, int cap) {
    int64_t nvec = nelem * (dtype == B200DDP_BF16 ? 2 : 4) / 16;
    if (algo != B200DDP_ALGO_ONESHOT) nvec = (nvec + world - 1) / world;
    // small buckets: two vectors per thread, so that few blocks (few flags) take part in the barrier
    int64_t g = (nvec + 2 * kP2PThreads - 1) / (2 * kP2PThreads);
    if (cap > kMaxBlocksP2P) cap = kMaxBlocksP2P;
    if (g < 1) g = 1;
    if (g > cap) g = cap;
    return (int)g;
}

int launch_p2p_allreduce(b200ddp_ctx *ctx, size_t off, int64_t nelem, int dtype, int algo, cudaStream_t s) {
    P2PArgs a = make_args(ctx, off, nelem, dtype);
    P2PKernel k = pick<false>(dtype, algo, ctx->world);
    const int grid = resident_clamp(ctx, k, kP2PThreads, p2p_grid_for(nelem, dtype, algo, ctx->world, block_cap(ctx)));
    B2_LAUNCH_RC(launch_ex(k, grid, kP2PThreads, s, ctx->pdl, a));
    return B200DDP_OK;
}

int launch_barrier(b200ddp_ctx *ctx, cudaStream_t s) {
    P2PArgs a = make_args(ctx, 0, 0, B200DDP_F32);
    B2_LAUNCH_RC(launch_ex(barrier_kernel, 1, 32, s, ctx->pdl, a));
    return B200DDP_OK;
}

int launch_nvls_allreduce(b200ddp_ctx *ctx, size_t off, int64_t nelem, int dtype, cudaStream_t s) {
    P2PArgs a = make_args(ctx, off, nelem, dtype);
    // The in-switch kernel wants FEWER blocks than the peer-load kernels: measured on 2, 4 and 8 GPUs
    // (profiles/r02_sweep_n*_fp32_variants.jsonl) a grid of 74-148 blocks beats 296 at every size from 4 MB up.
    int cap = ctx->nvls_blocks > 0 ? ctx->nvls_blocks : (ctx->p2p_blocks > 0 ? ctx->p2p_blocks : kNvlsBlocksDefault);
    const int want = p2p_grid_for(nelem, dtype, B200DDP_ALGO_NVLS, ctx->world, cap);
    cudaError_t e;
#define B2_NVLS_T(BT, D) \
    launch_ex(nvls_kernel<BT, D>, resident_clamp(ctx, nvls_kernel<BT, D>, kP2PThreads, want), kP2PThreads, s, ctx->pdl, a, ctx->mc_ptr)
#define B2_NVLS(D) (dtype == B200DDP_F32 ? B2_NVLS_T(float, D) : B2_NVLS_T(__nv_bfloat16, D))
    switch (nvls_depth(ctx)) {
        case 1: e = B2_NVLS(1); break;
        case 2: e = B2_NVLS(2); break;
        case 8: e = B2_NVLS(8); break;
        default: e = B2_NVLS(4); break;
    }
#undef B2_NVLS
#undef B2_NVLS_T
    B2_LAUNCH_RC(e);
    return B200DDP_OK;
}

int launch_p2p_allreduce_emu(b200ddp_ctx *const *ctxs, int n, size_t off, int64_t nelem, int dtype, int algo,
                             cudaStream_t s) {
    P2PArgs a = make_args(ctxs[0], off, nelem, dtype);
    for (int r = 0; r < n; ++r) a.gen[r] = ctxs[r]->block_gen;
    const void *k = (const void *)pick<true>(dtype, algo, n);
    int per_sm = 0;
    B2_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, kP2PThreads, 0));
    int grid = p2p_grid_for(nelem, dtype, algo, n, block_cap(ctxs[0]));
    const int fit = per_sm * ctxs[0]->sm_count / n;   // all ranks' blocks must be co-resident
    B2_REQUIRE(fit >= 1, B200DDP_ESTATE, "device cannot co-schedule %d emulated ranks", n);
    if (grid > fit) grid = fit;
    void *params[] = {&a};
    B2_CHECK_CUDA(cudaLaunchCooperativeKernel(k, dim3(grid, n), dim3(kP2PThreads), params, 0, s));
    return B200DDP_OK;
}

int launch_oneshot_update(b200ddp_ctx *ctx, const b200ddp_plan *plan, size_t off, const PtrArrays &pa,
                          const SgdScalars &hp, cudaStream_t s) {
    const bool wg = pa.k[kGradOut] != nullptr;
    for (const Segment &sg : plan->segments) {
        if (sg.c1 == sg.c0) continue;
        const bool small = sg.t1 - sg.t0 <= kTabSmall;
        int rc;
#define B2_FUSED(BT, WG) (small ? run_fused<BT, WG, kTabSmall>(ctx, plan, sg, off, pa, hp, s) \
                                : run_fused<BT, WG, kTabLarge>(ctx, plan, sg, off, pa, hp, s))
        if (plan->dtype == B200DDP_F32) rc = wg ? B2_FUSED(float, true) : B2_FUSED(float, false);
        else rc = wg ? B2_FUSED(__nv_bfloat16, true) : B2_FUSED(__nv_bfloat16, false);
#undef B2_FUSED
        if (rc) return rc;
    }
    return B200DDP_OK;
}

int launch_oneshot_update_emu(b200ddp_ctx *const *ctxs, b200ddp_plan *const *plans, int n, size_t off,
                              const PtrArrays *pas, const SgdScalars &hp, cudaStream_t s) {
    B2_REQUIRE(plans[0]->segments.size() <= 1 && plans[0]->T <= kTabSmall, B200DDP_ETOOBIG,
               "the emulated fused kernel takes at most %d tensors", kTabSmall);
    if (plans[0]->segments.empty() || plans[0]->chunks.empty()) return B200DDP_OK;
    const Segment &sg = plans[0]->segments[0];
    P2PArgs a = make_args(ctxs[0], off, 0, plans[0]->dtype);
    static TabSet<true, kTabSmall> ts;   // 14 KB: keep it off the stack
    for (int r = 0; r < B200DDP_MAX_WORLD; ++r) {
        const int q = r < n ? r : 0;
        a.gen[r] = ctxs[q]->block_gen;
        fill_update_tab(ts.t[r], plans[q], plans[q]->segments[0], pas[q]);
        ts.chunks[r] = plans[q]->d_chunks;
    }
    const bool wg = pas[0].k[kGradOut] != nullptr;
    const void *k;
    if (plans[0]->dtype == B200DDP_F32)
        k = wg ? (const void *)oneshot_update_kernel<float, true, true, kTabSmall>
               : (const void *)oneshot_update_kernel<float, false, true, kTabSmall>;
    else
        k = wg ? (const void *)oneshot_update_kernel<__nv_bfloat16, true, true, kTabSmall>
               : (const void *)oneshot_update_kernel<__nv_bfloat16, false, true, kTabSmall>;
    int per_sm = 0;
    B2_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, kThreads, 0));
    int nchunks = sg.c1 - sg.c0;
    int grid = update_grid(ctxs[0], nchunks);
    const int fit = per_sm * ctxs[0]->sm_count / n;
    B2_REQUIRE(fit >= 1, B200DDP_ESTATE, "device cannot co-schedule %d emulated ranks", n);
    if (grid > fit) grid = fit;
    SgdScalars h = hp;
    void *params[] = {&a, &ts, &nchunks, &h};
    B2_CHECK_CUDA(cudaLaunchCooperativeKernel(k, dim3(grid, n), dim3(kThreads), params, 0, s));
    return B200DDP_OK;
}

}  // namespace b200ddp

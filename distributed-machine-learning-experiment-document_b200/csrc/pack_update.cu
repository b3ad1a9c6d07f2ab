// K1 pack+scale, unpack, K3 unpack + momentum-SGD: multi-tensor, HBM-bound kernels.
//
// Work unit: one chunk = kChunk (2048) consecutive elements of one tensor; the
// chunk table is built once per plan.  A block takes NB consecutive chunks per
// iteration (grid stride over batches): first the metadata of the whole batch
// (chunk entry -> tensor meta + pointers, independent across the batch), then
// every 128-bit load of every full, aligned chunk of the batch, and only then
// the arithmetic and the stores.  One memory round trip therefore covers
// NB x kUnroll vectors per stream and thread; at the sizes of the lab's models
// (tens of microseconds per launch) the kernels are bound by that latency chain,
// not by bandwidth (profiles/r01_ncu_update_pack_summary.md).
// Ragged chunks (tensor tails, tiny tensors, misaligned bases) take a
// vector-then-scalar path afterwards.
#include <cstdlib>

#include "device_utils.cuh"

namespace b200ddp {

namespace {

struct ChunkCtx {
    int tensor;
    int count;       // valid elements in this chunk (0: past the end of the table)
    int64_t e0;      // first element inside the tensor
    int64_t boff;    // first element inside the bucket
};

__device__ __forceinline__ ChunkCtx chunk_ctx(const PlanView &pv, int c) {
    ChunkCtx r;
    if (c >= pv.nchunks) {
        r.tensor = 0; r.count = 0; r.e0 = 0; r.boff = 0;
        return r;
    }
    const Chunk ck = pv.chunks[c];
    const TensorMeta tm = pv.meta[ck.tensor];
    r.tensor = ck.tensor;
    r.e0 = (int64_t)ck.index * kChunk;
    const int64_t left = tm.numel - r.e0;
    r.count = left < kChunk ? (int)left : kChunk;
    r.boff = tm.bucket_off + r.e0;
    return r;
}

__device__ __forceinline__ float4 scale4(float4 v, float s) {
    v.x = __fmul_rn(v.x, s); v.y = __fmul_rn(v.y, s); v.z = __fmul_rn(v.z, s); v.w = __fmul_rn(v.w, s);
    return v;
}
__device__ __forceinline__ float4 div4(float4 v, float d) {
    v.x = __fdiv_rn(v.x, d); v.y = __fdiv_rn(v.y, d); v.z = __fdiv_rn(v.z, d); v.w = __fdiv_rn(v.w, d);
    return v;
}
__device__ __forceinline__ int vec_offset(int u, int tid) { return u * (kThreads * kVec) + tid * kVec; }

// ------------------------------------------------------------------------ K1
template <typename BT, int NB>
__global__ void __launch_bounds__(kThreads) pack_kernel(PlanView pv, BT *__restrict__ bucket, float scale) {
    const int tid = threadIdx.x;
    for (int c0 = blockIdx.x * NB; c0 < pv.nchunks; c0 += gridDim.x * NB) {
        ChunkCtx cc[NB];
        const float *g[NB];
        BT *b[NB];
        bool fast[NB];
#pragma unroll
        for (int j = 0; j < NB; ++j) {
            cc[j] = chunk_ctx(pv, c0 + j);
            g[j] = cc[j].count ? (const float *)pv.ptrs[kGradIn][cc[j].tensor] + cc[j].e0 : nullptr;
            b[j] = bucket + cc[j].boff;
            fast[j] = cc[j].count == kChunk && ((((uintptr_t)g[j] & 15) | ((uintptr_t)b[j] & Bucket<BT>::kAlignMask)) == 0);
        }
        float4 v[NB][kUnroll];
#pragma unroll
        for (int j = 0; j < NB; ++j)
            if (fast[j]) {
#pragma unroll
                for (int u = 0; u < kUnroll; ++u) v[j][u] = ld_stream_f4(g[j] + vec_offset(u, tid));
            }
#pragma unroll
        for (int j = 0; j < NB; ++j)
            if (fast[j]) {
#pragma unroll
                for (int u = 0; u < kUnroll; ++u) Bucket<BT>::store4(b[j] + vec_offset(u, tid), scale4(v[j][u], scale));
            }
#pragma unroll
        for (int j = 0; j < NB; ++j)
            if (!fast[j] && cc[j].count > 0) {
                const bool aligned = (((uintptr_t)g[j] & 15) | ((uintptr_t)b[j] & Bucket<BT>::kAlignMask)) == 0;
                const int nvec = aligned ? cc[j].count / kVec : 0;
                for (int i = tid; i < nvec; i += kThreads)
                    Bucket<BT>::store4(b[j] + i * kVec, scale4(ld_stream_f4(g[j] + i * kVec), scale));
                for (int i = nvec * kVec + tid; i < cc[j].count; i += kThreads)
                    Bucket<BT>::store1(b[j] + i, __fmul_rn(__ldg(g[j] + i), scale));
            }
    }
}

// -------------------------------------------------------------------- unpack
template <typename BT, int NB>
__global__ void __launch_bounds__(kThreads) unpack_kernel(PlanView pv, const BT *__restrict__ bucket, float divisor,
                                                          int use_div) {
    const int tid = threadIdx.x;
    for (int c0 = blockIdx.x * NB; c0 < pv.nchunks; c0 += gridDim.x * NB) {
        ChunkCtx cc[NB];
        float *g[NB];
        const BT *b[NB];
        bool fast[NB];
#pragma unroll
        for (int j = 0; j < NB; ++j) {
            cc[j] = chunk_ctx(pv, c0 + j);
            g[j] = cc[j].count ? (float *)pv.ptrs[kGradOut][cc[j].tensor] + cc[j].e0 : nullptr;
            b[j] = bucket + cc[j].boff;
            fast[j] = cc[j].count == kChunk && ((((uintptr_t)g[j] & 15) | ((uintptr_t)b[j] & Bucket<BT>::kAlignMask)) == 0);
        }
        float4 v[NB][kUnroll];
#pragma unroll
        for (int j = 0; j < NB; ++j)
            if (fast[j]) {
#pragma unroll
                for (int u = 0; u < kUnroll; ++u) v[j][u] = Bucket<BT>::load4(b[j] + vec_offset(u, tid));
            }
#pragma unroll
        for (int j = 0; j < NB; ++j)
            if (fast[j]) {
#pragma unroll
                for (int u = 0; u < kUnroll; ++u)
                    st_f4(g[j] + vec_offset(u, tid), use_div ? div4(v[j][u], divisor) : v[j][u]);
            }
#pragma unroll
        for (int j = 0; j < NB; ++j)
            if (!fast[j] && cc[j].count > 0) {
                const bool aligned = (((uintptr_t)g[j] & 15) | ((uintptr_t)b[j] & Bucket<BT>::kAlignMask)) == 0;
                const int nvec = aligned ? cc[j].count / kVec : 0;
                for (int i = tid; i < nvec; i += kThreads) {
                    const float4 x = Bucket<BT>::load4(b[j] + i * kVec);
                    st_f4(g[j] + i * kVec, use_div ? div4(x, divisor) : x);
                }
                for (int i = nvec * kVec + tid; i < cc[j].count; i += kThreads) {
                    const float x = Bucket<BT>::load1(b[j] + i);
                    g[j][i] = use_div ? __fdiv_rn(x, divisor) : x;
                }
            }
    }
}

// ------------------------------------------------------------------------ K3
// BT: element type the gradients are read in (bucket dtype; float in direct mode).
// FROM_BUCKET: gradients at bucket + offset[t]; otherwise from ptrs[kGradIn][t].
// WRITE_GRAD: also store the averaged gradient to ptrs[kGradOut][t].
template <typename BT, bool FROM_BUCKET, bool WRITE_GRAD, int NB>
__global__ void __launch_bounds__(kThreads) update_kernel(PlanView pv, const BT *__restrict__ bucket, SgdScalars hp) {
    const int tid = threadIdx.x;
    const bool load_m = hp.has_momentum && !hp.first_step;
    for (int c0 = blockIdx.x * NB; c0 < pv.nchunks; c0 += gridDim.x * NB) {
        ChunkCtx cc[NB];
        float *p[NB], *m[NB], *go[NB];
        const BT *g[NB];
        bool fast[NB];
#pragma unroll
        for (int j = 0; j < NB; ++j) {
            cc[j] = chunk_ctx(pv, c0 + j);
            const bool live = cc[j].count > 0;
            p[j] = live ? (float *)pv.ptrs[kParam][cc[j].tensor] + cc[j].e0 : nullptr;
            m[j] = live && hp.has_momentum ? (float *)pv.ptrs[kMom][cc[j].tensor] + cc[j].e0 : nullptr;
            go[j] = live && WRITE_GRAD ? (float *)pv.ptrs[kGradOut][cc[j].tensor] + cc[j].e0 : nullptr;
            g[j] = FROM_BUCKET ? bucket + cc[j].boff
                               : (live ? (const BT *)pv.ptrs[kGradIn][cc[j].tensor] + cc[j].e0 : nullptr);
            fast[j] = cc[j].count == kChunk && (((((uintptr_t)p[j] | (uintptr_t)m[j] | (uintptr_t)go[j]) & 15) |
                                                 ((uintptr_t)g[j] & Bucket<BT>::kAlignMask)) == 0);
        }
        float4 gv[NB][kUnroll], pv4[NB][kUnroll], mv[NB][kUnroll];
#pragma unroll
        for (int j = 0; j < NB; ++j)
            if (fast[j]) {
#pragma unroll
                for (int u = 0; u < kUnroll; ++u) {
                    const int o = vec_offset(u, tid);
                    gv[j][u] = Bucket<BT>::load4(g[j] + o);
                    pv4[j][u] = ld_rmw_f4(p[j] + o);
                    if (load_m) mv[j][u] = ld_rmw_f4(m[j] + o);
                }
            }
#pragma unroll
        for (int j = 0; j < NB; ++j)
            if (fast[j]) {
#pragma unroll
                for (int u = 0; u < kUnroll; ++u) {
                    const int o = vec_offset(u, tid);
                    float4 ga;
                    sgd_elem4(gv[j][u], pv4[j][u], mv[j][u], ga, hp);
                    st_f4(p[j] + o, pv4[j][u]);
                    if (hp.has_momentum) st_f4(m[j] + o, mv[j][u]);
                    if (WRITE_GRAD) st_f4(go[j] + o, ga);
                }
            }
#pragma unroll
        for (int j = 0; j < NB; ++j)
            if (!fast[j] && cc[j].count > 0) {
                const bool aligned = ((((uintptr_t)p[j] | (uintptr_t)m[j] | (uintptr_t)go[j]) & 15) |
                                      ((uintptr_t)g[j] & Bucket<BT>::kAlignMask)) == 0;
                const int nvec = aligned ? cc[j].count / kVec : 0;
                for (int i = tid; i < nvec; i += kThreads) {
                    const int o = i * kVec;
                    float4 g4 = Bucket<BT>::load4(g[j] + o), p4 = ld_rmw_f4(p[j] + o), m4, ga;
                    if (load_m) m4 = ld_rmw_f4(m[j] + o);
                    sgd_elem4(g4, p4, m4, ga, hp);
                    st_f4(p[j] + o, p4);
                    if (hp.has_momentum) st_f4(m[j] + o, m4);
                    if (WRITE_GRAD) st_f4(go[j] + o, ga);
                }
                for (int i = nvec * kVec + tid; i < cc[j].count; i += kThreads) {
                    float gs = Bucket<BT>::load1(g[j] + i), ps = p[j][i], ms = load_m ? m[j][i] : 0.f, ga;
                    sgd_elem(gs, ps, ms, ga, hp);
                    p[j][i] = ps;
                    if (hp.has_momentum) m[j][i] = ms;
                    if (WRITE_GRAD) go[j][i] = ga;
                }
            }
    }
}

// Occupancy is a property of the kernel, not of the launch: query once.
struct Geometry { int blocks_per_sm = 0; };
template <typename K> int grid_for(K kernel, Geometry &geo, int nchunks, int nb, int sm_count) {
    if (geo.blocks_per_sm == 0) {
        int n = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, kThreads, 0) != cudaSuccess || n < 1) n = 1;
        geo.blocks_per_sm = n;
    }
    const int batches = (nchunks + nb - 1) / nb;
    const int g = sm_count * geo.blocks_per_sm;
    return batches < g ? batches : g;
}

// Chunks per batch; overridable for experiments (B200DDP_PACK_NB / B200DDP_UPDATE_NB = 1, 2 or 4).
int env_nb(const char *name, int dflt) {
    const char *v = getenv(name);
    if (!v) return dflt;
    const int n = atoi(v);
    return (n == 1 || n == 2 || n == 4) ? n : dflt;
}
// Measured on B200 (ResNet-18 / ResNet-50 lists): NB = 1 wins everywhere - e.g. K3 39 us vs 42 us
// (NB 2) vs 54 us (NB 4): the larger batch costs registers, hence resident warps, and the warps
// hide the latency chain better than the batch does.
int pack_nb() { static const int n = env_nb("B200DDP_PACK_NB", 1); return n; }
int update_nb() { static const int n = env_nb("B200DDP_UPDATE_NB", 1); return n; }

#define B2_LAUNCH_CHECK()                                                                   \
    do {                                                                                    \
        cudaError_t _e = cudaGetLastError();                                                \
        if (_e != cudaSuccess) {                                                            \
            set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e), __FILE__, \
                      __LINE__);                                                            \
            return B200DDP_ECUDA;                                                           \
        }                                                                                   \
    } while (0)

template <typename BT, int NB> void run_pack(const PlanView &pv, void *bucket, float scale, int sm, cudaStream_t s) {
    static Geometry geo;
    pack_kernel<BT, NB><<<grid_for(pack_kernel<BT, NB>, geo, pv.nchunks, NB, sm), kThreads, 0, s>>>(pv, (BT *)bucket, scale);
}
template <typename BT, int NB>
void run_unpack(const PlanView &pv, const void *bucket, float divisor, int use_div, int sm, cudaStream_t s) {
    static Geometry geo;
    unpack_kernel<BT, NB><<<grid_for(unpack_kernel<BT, NB>, geo, pv.nchunks, NB, sm), kThreads, 0, s>>>(
        pv, (const BT *)bucket, divisor, use_div);
}
template <typename BT, bool FB, bool WG, int NB>
void run_update(const PlanView &pv, const void *bucket, const SgdScalars &hp, int sm, cudaStream_t s) {
    static Geometry geo;
    update_kernel<BT, FB, WG, NB><<<grid_for(update_kernel<BT, FB, WG, NB>, geo, pv.nchunks, NB, sm), kThreads, 0, s>>>(
        pv, (const BT *)bucket, hp);
}
template <typename BT, bool FB, bool WG>
void run_update_nb(int nb, const PlanView &pv, const void *bucket, const SgdScalars &hp, int sm, cudaStream_t s) {
    if (nb == 1) run_update<BT, FB, WG, 1>(pv, bucket, hp, sm, s);
    else if (nb == 2) run_update<BT, FB, WG, 2>(pv, bucket, hp, sm, s);
    else run_update<BT, FB, WG, 4>(pv, bucket, hp, sm, s);
}

}  // namespace

int launch_pack(const PlanView &pv, int dtype, void *bucket, float scale, int sm_count, cudaStream_t s) {
    if (pv.nchunks == 0) return B200DDP_OK;
    cudaGetLastError();   // a stale error of an unrelated earlier call must not be blamed on this launch
    const int nb = pack_nb();
    if (dtype == B200DDP_F32) {
        if (nb == 1) run_pack<float, 1>(pv, bucket, scale, sm_count, s);
        else if (nb == 2) run_pack<float, 2>(pv, bucket, scale, sm_count, s);
        else run_pack<float, 4>(pv, bucket, scale, sm_count, s);
    } else {
        if (nb == 1) run_pack<__nv_bfloat16, 1>(pv, bucket, scale, sm_count, s);
        else if (nb == 2) run_pack<__nv_bfloat16, 2>(pv, bucket, scale, sm_count, s);
        else run_pack<__nv_bfloat16, 4>(pv, bucket, scale, sm_count, s);
    }
    B2_LAUNCH_CHECK();
    return B200DDP_OK;
}

int launch_unpack(const PlanView &pv, int dtype, const void *bucket, float divisor, int sm_count, cudaStream_t s) {
    if (pv.nchunks == 0) return B200DDP_OK;
    cudaGetLastError();
    const int use_div = divisor != 1.0f;
    const int nb = pack_nb();
    if (dtype == B200DDP_F32) {
        if (nb == 1) run_unpack<float, 1>(pv, bucket, divisor, use_div, sm_count, s);
        else if (nb == 2) run_unpack<float, 2>(pv, bucket, divisor, use_div, sm_count, s);
        else run_unpack<float, 4>(pv, bucket, divisor, use_div, sm_count, s);
    } else {
        if (nb == 1) run_unpack<__nv_bfloat16, 1>(pv, bucket, divisor, use_div, sm_count, s);
        else if (nb == 2) run_unpack<__nv_bfloat16, 2>(pv, bucket, divisor, use_div, sm_count, s);
        else run_unpack<__nv_bfloat16, 4>(pv, bucket, divisor, use_div, sm_count, s);
    }
    B2_LAUNCH_CHECK();
    return B200DDP_OK;
}

int launch_update(const PlanView &pv, int dtype, const void *bucket, bool write_grad, const SgdScalars &hp,
                  int sm_count, cudaStream_t s) {
    if (pv.nchunks == 0) return B200DDP_OK;
    cudaGetLastError();
    const int nb = update_nb();
    if (bucket == nullptr) {
        run_update_nb<float, false, false>(nb, pv, bucket, hp, sm_count, s);
    } else if (dtype == B200DDP_F32) {
        if (write_grad) run_update_nb<float, true, true>(nb, pv, bucket, hp, sm_count, s);
        else run_update_nb<float, true, false>(nb, pv, bucket, hp, sm_count, s);
    } else {
        if (write_grad) run_update_nb<__nv_bfloat16, true, true>(nb, pv, bucket, hp, sm_count, s);
        else run_update_nb<__nv_bfloat16, true, false>(nb, pv, bucket, hp, sm_count, s);
    }
    B2_LAUNCH_CHECK();
    return B200DDP_OK;
}

}  // namespace b200ddp

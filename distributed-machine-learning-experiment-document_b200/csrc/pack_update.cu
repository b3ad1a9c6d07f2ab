// K1 pack+scale, unpack, K3 unpack + momentum-SGD: multi-tensor, HBM-bound kernels.
//
// Work unit: one chunk = kChunk (2048) consecutive elements of one tensor.  The
// chunk table is built once per plan; blocks walk it with a grid stride, so at
// any instant the resident blocks stream one contiguous window of every array.
// Per chunk a thread moves kUnroll 128-bit vectors of every stream, all loads
// issued before the first use.  Ragged chunks (tensor tails, tiny tensors,
// misaligned bases) take a vector-then-scalar path.
#include "device_utils.cuh"

namespace b200ddp {

namespace {

struct ChunkCtx {
    int tensor;
    int count;       // valid elements in this chunk
    int64_t e0;      // first element inside the tensor
    int64_t boff;    // first element inside the bucket
};

__device__ __forceinline__ ChunkCtx chunk_ctx(const PlanView &pv, int c) {
    Chunk ck = pv.chunks[c];
    TensorMeta tm = pv.meta[ck.tensor];
    ChunkCtx r;
    r.tensor = ck.tensor;
    r.e0 = (int64_t)ck.index * kChunk;
    int64_t left = tm.numel - r.e0;
    r.count = left < kChunk ? (int)left : kChunk;
    r.boff = tm.bucket_off + r.e0;
    return r;
}

// ------------------------------------------------------------------------ K1
template <typename BT>
__global__ void __launch_bounds__(kThreads) pack_kernel(PlanView pv, BT *__restrict__ bucket, float scale) {
    const int tid = threadIdx.x;
    for (int c = blockIdx.x; c < pv.nchunks; c += gridDim.x) {
        ChunkCtx cc = chunk_ctx(pv, c);
        const float *g = (const float *)pv.ptrs[kGradIn][cc.tensor] + cc.e0;
        BT *b = bucket + cc.boff;
        const bool aligned = (((uintptr_t)g & 15) | ((uintptr_t)b & Bucket<BT>::kAlignMask)) == 0;
        if (cc.count == kChunk && aligned) {
            float4 v[kUnroll];
#pragma unroll
            for (int u = 0; u < kUnroll; ++u) v[u] = ld_stream_f4(g + u * (kThreads * kVec) + tid * kVec);
#pragma unroll
            for (int u = 0; u < kUnroll; ++u) {
                float4 o;
                o.x = __fmul_rn(v[u].x, scale);
                o.y = __fmul_rn(v[u].y, scale);
                o.z = __fmul_rn(v[u].z, scale);
                o.w = __fmul_rn(v[u].w, scale);
                Bucket<BT>::store4(b + u * (kThreads * kVec) + tid * kVec, o);
            }
        } else {
            const int nvec = aligned ? cc.count / kVec : 0;
            for (int i = tid; i < nvec; i += kThreads) {
                float4 v = ld_stream_f4(g + i * kVec);
                v.x = __fmul_rn(v.x, scale);
                v.y = __fmul_rn(v.y, scale);
                v.z = __fmul_rn(v.z, scale);
                v.w = __fmul_rn(v.w, scale);
                Bucket<BT>::store4(b + i * kVec, v);
            }
            for (int i = nvec * kVec + tid; i < cc.count; i += kThreads)
                Bucket<BT>::store1(b + i, __fmul_rn(__ldg(g + i), scale));
        }
    }
}

// -------------------------------------------------------------------- unpack
template <typename BT>
__global__ void __launch_bounds__(kThreads) unpack_kernel(PlanView pv, const BT *__restrict__ bucket, float divisor,
                                                          int use_div) {
    const int tid = threadIdx.x;
    for (int c = blockIdx.x; c < pv.nchunks; c += gridDim.x) {
        ChunkCtx cc = chunk_ctx(pv, c);
        float *g = (float *)pv.ptrs[kGradOut][cc.tensor] + cc.e0;
        const BT *b = bucket + cc.boff;
        const bool aligned = (((uintptr_t)g & 15) | ((uintptr_t)b & Bucket<BT>::kAlignMask)) == 0;
        if (cc.count == kChunk && aligned) {
            float4 v[kUnroll];
#pragma unroll
            for (int u = 0; u < kUnroll; ++u) v[u] = Bucket<BT>::load4(b + u * (kThreads * kVec) + tid * kVec);
#pragma unroll
            for (int u = 0; u < kUnroll; ++u) {
                if (use_div) {
                    v[u].x = __fdiv_rn(v[u].x, divisor);
                    v[u].y = __fdiv_rn(v[u].y, divisor);
                    v[u].z = __fdiv_rn(v[u].z, divisor);
                    v[u].w = __fdiv_rn(v[u].w, divisor);
                }
                st_f4(g + u * (kThreads * kVec) + tid * kVec, v[u]);
            }
        } else {
            const int nvec = aligned ? cc.count / kVec : 0;
            for (int i = tid; i < nvec; i += kThreads) {
                float4 v = Bucket<BT>::load4(b + i * kVec);
                if (use_div) {
                    v.x = __fdiv_rn(v.x, divisor);
                    v.y = __fdiv_rn(v.y, divisor);
                    v.z = __fdiv_rn(v.z, divisor);
                    v.w = __fdiv_rn(v.w, divisor);
                }
                st_f4(g + i * kVec, v);
            }
            for (int i = nvec * kVec + tid; i < cc.count; i += kThreads) {
                float v = Bucket<BT>::load1(b + i);
                g[i] = use_div ? __fdiv_rn(v, divisor) : v;
            }
        }
    }
}

// ------------------------------------------------------------------------ K3
// BT: element type the gradients are read in (bucket dtype; float in direct mode).
// FROM_BUCKET: gradients at bucket + offset[t]; otherwise from ptrs[kGradIn][t].
// WRITE_GRAD: also store the averaged gradient to ptrs[kGradOut][t].
template <typename BT, bool FROM_BUCKET, bool WRITE_GRAD>
__global__ void __launch_bounds__(kThreads) update_kernel(PlanView pv, const BT *__restrict__ bucket, SgdScalars hp) {
    const int tid = threadIdx.x;
    const bool load_m = hp.has_momentum && !hp.first_step;
    for (int c = blockIdx.x; c < pv.nchunks; c += gridDim.x) {
        ChunkCtx cc = chunk_ctx(pv, c);
        float *p = (float *)pv.ptrs[kParam][cc.tensor] + cc.e0;
        float *m = hp.has_momentum ? (float *)pv.ptrs[kMom][cc.tensor] + cc.e0 : nullptr;
        float *go = WRITE_GRAD ? (float *)pv.ptrs[kGradOut][cc.tensor] + cc.e0 : nullptr;
        const BT *g = FROM_BUCKET ? bucket + cc.boff : (const BT *)pv.ptrs[kGradIn][cc.tensor] + cc.e0;
        const bool aligned = ((((uintptr_t)p | (uintptr_t)m | (uintptr_t)go) & 15) |
                              ((uintptr_t)g & Bucket<BT>::kAlignMask)) == 0;
        if (cc.count == kChunk && aligned) {
            float4 gv[kUnroll], pv4[kUnroll], mv[kUnroll];
#pragma unroll
            for (int u = 0; u < kUnroll; ++u) {
                const int o = u * (kThreads * kVec) + tid * kVec;
                gv[u] = Bucket<BT>::load4(g + o);
                pv4[u] = ld_rmw_f4(p + o);
                if (load_m) mv[u] = ld_rmw_f4(m + o);
            }
#pragma unroll
            for (int u = 0; u < kUnroll; ++u) {
                const int o = u * (kThreads * kVec) + tid * kVec;
                float4 ga;
                sgd_elem4(gv[u], pv4[u], mv[u], ga, hp);
                st_f4(p + o, pv4[u]);
                if (hp.has_momentum) st_f4(m + o, mv[u]);
                if (WRITE_GRAD) st_f4(go + o, ga);
            }
        } else {
            const int nvec = aligned ? cc.count / kVec : 0;
            for (int i = tid; i < nvec; i += kThreads) {
                const int o = i * kVec;
                float4 gv = Bucket<BT>::load4(g + o), p4 = ld_rmw_f4(p + o), m4, ga;
                if (load_m) m4 = ld_rmw_f4(m + o);
                sgd_elem4(gv, p4, m4, ga, hp);
                st_f4(p + o, p4);
                if (hp.has_momentum) st_f4(m + o, m4);
                if (WRITE_GRAD) st_f4(go + o, ga);
            }
            for (int i = nvec * kVec + tid; i < cc.count; i += kThreads) {
                float gs = Bucket<BT>::load1(g + i), ps = p[i], ms = load_m ? m[i] : 0.f, ga;
                sgd_elem(gs, ps, ms, ga, hp);
                p[i] = ps;
                if (hp.has_momentum) m[i] = ms;
                if (WRITE_GRAD) go[i] = ga;
            }
        }
    }
}

template <typename K> int blocks_per_sm(K kernel) {
    int n = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, kThreads, 0) != cudaSuccess || n < 1) n = 1;
    return n;
}
template <typename K> int grid_for(K kernel, int nchunks, int sm_count) {
    int g = sm_count * blocks_per_sm(kernel);
    return nchunks < g ? nchunks : g;
}

#define B2_LAUNCH_CHECK()                                                                   \
    do {                                                                                    \
        cudaError_t _e = cudaGetLastError();                                                \
        if (_e != cudaSuccess) {                                                            \
            set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e), __FILE__, \
                      __LINE__);                                                            \
            return B200DDP_ECUDA;                                                           \
        }                                                                                   \
    } while (0)

}  // namespace

int launch_pack(const PlanView &pv, int dtype, void *bucket, float scale, int sm_count, cudaStream_t s) {
    if (pv.nchunks == 0) return B200DDP_OK;
    cudaGetLastError();   // a stale error of an unrelated earlier call must not be blamed on this launch
    if (dtype == B200DDP_F32) {
        pack_kernel<float><<<grid_for(pack_kernel<float>, pv.nchunks, sm_count), kThreads, 0, s>>>(pv, (float *)bucket,
                                                                                                  scale);
    } else {
        pack_kernel<__nv_bfloat16><<<grid_for(pack_kernel<__nv_bfloat16>, pv.nchunks, sm_count), kThreads, 0, s>>>(
            pv, (__nv_bfloat16 *)bucket, scale);
    }
    B2_LAUNCH_CHECK();
    return B200DDP_OK;
}

int launch_unpack(const PlanView &pv, int dtype, const void *bucket, float divisor, int sm_count, cudaStream_t s) {
    if (pv.nchunks == 0) return B200DDP_OK;
    cudaGetLastError();
    const int use_div = divisor != 1.0f;
    if (dtype == B200DDP_F32) {
        unpack_kernel<float><<<grid_for(unpack_kernel<float>, pv.nchunks, sm_count), kThreads, 0, s>>>(
            pv, (const float *)bucket, divisor, use_div);
    } else {
        unpack_kernel<__nv_bfloat16><<<grid_for(unpack_kernel<__nv_bfloat16>, pv.nchunks, sm_count), kThreads, 0, s>>>(
            pv, (const __nv_bfloat16 *)bucket, divisor, use_div);
    }
    B2_LAUNCH_CHECK();
    return B200DDP_OK;
}

int launch_update(const PlanView &pv, int dtype, const void *bucket, bool write_grad, const SgdScalars &hp,
                  int sm_count, cudaStream_t s) {
    if (pv.nchunks == 0) return B200DDP_OK;
    cudaGetLastError();
#define B2_UPD(BT, FB, WG)                                                                              \
    update_kernel<BT, FB, WG><<<grid_for(update_kernel<BT, FB, WG>, pv.nchunks, sm_count), kThreads, 0, s>>>( \
        pv, (const BT *)bucket, hp)
    if (bucket == nullptr) {
        B2_UPD(float, false, false);
    } else if (dtype == B200DDP_F32) {
        if (write_grad) B2_UPD(float, true, true); else B2_UPD(float, true, false);
    } else {
        if (write_grad) B2_UPD(__nv_bfloat16, true, true); else B2_UPD(__nv_bfloat16, true, false);
    }
#undef B2_UPD
    B2_LAUNCH_CHECK();
    return B200DDP_OK;
}

}  // namespace b200ddp

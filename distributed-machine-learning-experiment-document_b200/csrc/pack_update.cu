// K1 pack+scale, unpack, K3 unpack + momentum-SGD: multi-tensor, HBM-bound kernels.
//
// Work unit: one chunk = kChunk (2048) consecutive elements of one tensor.  The
// chunk table (device memory, written once per plan) is walked with a grid
// stride, so at any instant the resident blocks stream one contiguous window of
// every array.  Sizes, bucket offsets and tensor pointers come from a
// __grid_constant__ table in the kernel parameters (constant bank): the only
// global-memory access before the first data load is the 8-byte chunk entry.
// Per chunk a thread moves kUnroll 128-bit vectors of every stream, all loads
// issued before the first use.  Ragged chunks (tensor tails, tiny tensors,
// misaligned bases) take a vector-then-scalar path.
//
// Measured dead end (B200, ResNet-18/50 lists): handling 2 or 4 chunks per
// iteration to batch more loads per latency round trip is SLOWER (K3: 39 us ->
// 42 us -> 54 us): the registers it costs remove resident warps, and warps hide
// the latency better than the batch does.
#include "device_utils.cuh"

namespace b200ddp {

namespace {

struct ChunkCtx {
    int tensor;
    int count;       // valid elements in this chunk
    int64_t e0;      // first element inside the tensor
    int64_t boff;    // first element inside the bucket
};

template <int CHUNK, int CAP, int NK>
__device__ __forceinline__ ChunkCtx chunk_ctx(const TensorTab<CAP, NK> &tab, const Chunk *chunks, int c) {
    const Chunk ck = chunks[c];
    ChunkCtx r;
    r.tensor = ck.tensor;
    r.e0 = (int64_t)ck.index * CHUNK;
    const int64_t left = tab.numel[ck.tensor] - r.e0;
    r.count = left < CHUNK ? (int)left : CHUNK;
    r.boff = tab.boff[ck.tensor] + r.e0;
    return r;
}

__device__ __forceinline__ float4 scale4(float4 v, float s) {
    v.x = __fmul_rn(v.x, s); v.y = __fmul_rn(v.y, s); v.z = __fmul_rn(v.z, s); v.w = __fmul_rn(v.w, s);
    return v;
}
__device__ __forceinline__ float4 div4(float4 v, const Divisor &dv) {
    v.x = apply_div(v.x, dv); v.y = apply_div(v.y, dv); v.z = apply_div(v.z, dv); v.w = apply_div(v.w, dv);
    return v;
}
__device__ __forceinline__ int vec_offset(int u, int tid) { return u * (kThreads * kVec) + tid * kVec; }

// ------------------------------------------------------------------------ K1
template <typename BT, int CAP, int U>
__global__ void __launch_bounds__(kThreads) pack_kernel(const __grid_constant__ TensorTab<CAP, 1> tab,
                                                        const Chunk *__restrict__ chunks, int nchunks,
                                                        BT *__restrict__ bucket, float scale) {
    constexpr int CHUNK = kThreads * kVec * U;
    const int tid = threadIdx.x;
    pdl_prologue();
    for (int c = blockIdx.x; c < nchunks; c += gridDim.x) {
        const ChunkCtx cc = chunk_ctx<CHUNK>(tab, chunks, c);
        const float *g = (const float *)tab.ptr[0][cc.tensor] + cc.e0;
        BT *b = bucket + cc.boff;
        const bool aligned = (((uintptr_t)g & 15) | ((uintptr_t)b & Bucket<BT>::kAlignMask)) == 0;
        if (cc.count == CHUNK && aligned) {
            float4 v[U];
#pragma unroll
            for (int u = 0; u < U; ++u) v[u] = ld_stream_f4(g + vec_offset(u, tid));
#pragma unroll
            for (int u = 0; u < U; ++u) Bucket<BT>::store4(b + vec_offset(u, tid), scale4(v[u], scale));
        } else {
            const int nvec = aligned ? cc.count / kVec : 0;
            for (int i = tid; i < nvec; i += kThreads)
                Bucket<BT>::store4(b + i * kVec, scale4(ld_stream_f4(g + i * kVec), scale));
            for (int i = nvec * kVec + tid; i < cc.count; i += kThreads)
                Bucket<BT>::store1(b + i, __fmul_rn(__ldg(g + i), scale));
        }
    }
}

// -------------------------------------------------------------------- unpack
template <typename BT, int CAP, int U>
__global__ void __launch_bounds__(kThreads) unpack_kernel(const __grid_constant__ TensorTab<CAP, 1> tab,
                                                          const Chunk *__restrict__ chunks, int nchunks,
                                                          const BT *__restrict__ bucket, Divisor dv) {
    constexpr int CHUNK = kThreads * kVec * U;
    const int tid = threadIdx.x;
    pdl_prologue();
    for (int c = blockIdx.x; c < nchunks; c += gridDim.x) {
        const ChunkCtx cc = chunk_ctx<CHUNK>(tab, chunks, c);
        float *g = (float *)tab.ptr[0][cc.tensor] + cc.e0;
        const BT *b = bucket + cc.boff;
        const bool aligned = (((uintptr_t)g & 15) | ((uintptr_t)b & Bucket<BT>::kAlignMask)) == 0;
        if (cc.count == CHUNK && aligned) {
            float4 v[U];
#pragma unroll
            for (int u = 0; u < U; ++u) v[u] = Bucket<BT>::load4(b + vec_offset(u, tid));
#pragma unroll
            for (int u = 0; u < U; ++u) st_f4(g + vec_offset(u, tid), div4(v[u], dv));
        } else {
            const int nvec = aligned ? cc.count / kVec : 0;
            for (int i = tid; i < nvec; i += kThreads) {
                const float4 x = Bucket<BT>::load4(b + i * kVec);
                st_f4(g + i * kVec, div4(x, dv));
            }
            for (int i = nvec * kVec + tid; i < cc.count; i += kThreads) {
                const float x = Bucket<BT>::load1(b + i);
                g[i] = apply_div(x, dv);
            }
        }
    }
}

// ------------------------------------------------------------------------ K3
// BT: element type the gradients are read in (bucket dtype; float in direct mode).
// FROM_BUCKET: gradients at bucket + offset[t]; otherwise from tab.ptr[kGradIn][t].
// WRITE_GRAD: also store the averaged gradient to tab.ptr[kGradOut][t].
// RULE: kRuleSgd (torch.optim.SGD), kRuleAdam / kRuleGd (the lab's own optimizers).
template <typename BT, bool FROM_BUCKET, bool WRITE_GRAD, int RULE, int CAP>
__global__ void __launch_bounds__(kThreads) update_kernel(const __grid_constant__ TensorTab<CAP, kNumPtrKinds> tab,
                                                          const Chunk *__restrict__ chunks, int nchunks,
                                                          const BT *__restrict__ bucket, SgdScalars hp) {
    constexpr bool ADAM = RULE == kRuleAdam;
    const int tid = threadIdx.x;
    const bool has_m = ADAM || (RULE == kRuleSgd && hp.has_momentum);
    const bool load_m = ADAM || (RULE == kRuleSgd && hp.has_momentum && !hp.first_step);
    pdl_prologue();
    for (int c = blockIdx.x; c < nchunks; c += gridDim.x) {
        const ChunkCtx cc = chunk_ctx<kChunk>(tab, chunks, c);
        float *p = (float *)tab.ptr[kParam][cc.tensor] + cc.e0;
        float *m = has_m ? (float *)tab.ptr[kMom][cc.tensor] + cc.e0 : nullptr;
        float *vv = ADAM ? (float *)tab.ptr[kVel][cc.tensor] + cc.e0 : nullptr;
        float *go = WRITE_GRAD ? (float *)tab.ptr[kGradOut][cc.tensor] + cc.e0 : nullptr;
        const BT *g = FROM_BUCKET ? bucket + cc.boff : (const BT *)tab.ptr[kGradIn][cc.tensor] + cc.e0;
        const bool aligned = ((((uintptr_t)p | (uintptr_t)m | (uintptr_t)vv | (uintptr_t)go) & 15) |
                              ((uintptr_t)g & Bucket<BT>::kAlignMask)) == 0;
        if (cc.count == kChunk && aligned) {
            float4 gv[kUnroll], pv4[kUnroll], mv[kUnroll], sv[kUnroll];
#pragma unroll
            for (int u = 0; u < kUnroll; ++u) {
                const int o = vec_offset(u, tid);
                gv[u] = Bucket<BT>::load4(g + o);
                pv4[u] = ld_rmw_f4(p + o);
                if (load_m) mv[u] = ld_rmw_f4(m + o);
                if (ADAM) sv[u] = ld_rmw_f4(vv + o);
            }
#pragma unroll
            for (int u = 0; u < kUnroll; ++u) {
                const int o = vec_offset(u, tid);
                float4 ga;
                opt_elem4<RULE>(gv[u], pv4[u], mv[u], sv[u], ga, hp);
                st_f4(p + o, pv4[u]);
                if (has_m) st_f4(m + o, mv[u]);
                if (ADAM) st_f4(vv + o, sv[u]);
                if (WRITE_GRAD) st_f4(go + o, ga);
            }
        } else {
            const int nvec = aligned ? cc.count / kVec : 0;
            for (int i = tid; i < nvec; i += kThreads) {
                const int o = i * kVec;
                float4 g4 = Bucket<BT>::load4(g + o), p4 = ld_rmw_f4(p + o), m4, v4, ga;
                if (load_m) m4 = ld_rmw_f4(m + o);
                if (ADAM) v4 = ld_rmw_f4(vv + o);
                opt_elem4<RULE>(g4, p4, m4, v4, ga, hp);
                st_f4(p + o, p4);
                if (has_m) st_f4(m + o, m4);
                if (ADAM) st_f4(vv + o, v4);
                if (WRITE_GRAD) st_f4(go + o, ga);
            }
            for (int i = nvec * kVec + tid; i < cc.count; i += kThreads) {
                float gs = Bucket<BT>::load1(g + i), ps = p[i], ms = load_m ? m[i] : 0.f, vs = ADAM ? vv[i] : 0.f, ga;
                opt_elem<RULE>(gs, ps, ms, vs, ga, hp);
                p[i] = ps;
                if (has_m) m[i] = ms;
                if (ADAM) vv[i] = vs;
                if (WRITE_GRAD) go[i] = ga;
            }
        }
    }
}

// Occupancy is a property of the kernel, not of the launch: query once per instantiation.
template <typename K> int grid_for(K kernel, int &blocks_per_sm, int nchunks, const LaunchCfg &lc) {
    if (blocks_per_sm == 0) {
        int n = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, kThreads, 0) != cudaSuccess || n < 1) n = 1;
        blocks_per_sm = n;
    }
    const int per = lc.blocks_per_sm > 0 && lc.blocks_per_sm < blocks_per_sm ? lc.blocks_per_sm : blocks_per_sm;
    const int g = lc.sm_count * per;
    return nchunks < g ? nchunks : g;
}

template <int CAP, int NK>
void fill_tab(TensorTab<CAP, NK> &tab, const b200ddp_plan *plan, const Segment &sg, const void *const *const *kinds) {
    const int n = sg.t1 - sg.t0;
    for (int i = 0; i < n; ++i) {
        tab.numel[i] = plan->meta[sg.t0 + i].numel;
        tab.boff[i] = plan->meta[sg.t0 + i].bucket_off;
    }
    for (int k = 0; k < NK; ++k)
        for (int i = 0; i < n; ++i) tab.ptr[k][i] = kinds[k] ? const_cast<void *>(kinds[k][sg.t0 + i]) : nullptr;
}

template <typename BT, int CAP, int U>
cudaError_t run_pack(const b200ddp_plan *plan, const Segment &sg, const void *const *grads, void *bucket, float scale,
                     const LaunchCfg &lc, cudaStream_t s) {
    static int occ = 0;
    TensorTab<CAP, 1> tab;
    const void *const *kinds[1] = {grads};
    fill_tab(tab, plan, sg, kinds);
    const int n = sg.k1 - sg.k0;
    return launch_ex(pack_kernel<BT, CAP, U>, grid_for(pack_kernel<BT, CAP, U>, occ, n, lc), kThreads, s, lc.pdl,
                     tab, (const Chunk *)(plan->d_copy_chunks + sg.k0), n, (BT *)bucket, scale);
}
template <typename BT, int CAP, int U>
cudaError_t run_unpack(const b200ddp_plan *plan, const Segment &sg, const void *bucket, const void *const *grads_out,
                       float divisor, const LaunchCfg &lc, cudaStream_t s) {
    static int occ = 0;
    TensorTab<CAP, 1> tab;
    const void *const *kinds[1] = {grads_out};
    fill_tab(tab, plan, sg, kinds);
    const int n = sg.k1 - sg.k0;
    const Divisor dv = make_divisor(divisor, lc.div_mode);
    return launch_ex(unpack_kernel<BT, CAP, U>, grid_for(unpack_kernel<BT, CAP, U>, occ, n, lc), kThreads, s,
                     lc.pdl, tab, (const Chunk *)(plan->d_copy_chunks + sg.k0), n, (const BT *)bucket, dv);
}
template <typename BT, int CAP>
cudaError_t run_pack_u(const b200ddp_plan *plan, const Segment &sg, const void *const *grads, void *bucket, float scale,
                       const LaunchCfg &lc, cudaStream_t s) {
    switch (plan->copy_unroll) {
        case 2: return run_pack<BT, CAP, 2>(plan, sg, grads, bucket, scale, lc, s);
        case 8: return run_pack<BT, CAP, 8>(plan, sg, grads, bucket, scale, lc, s);
        default: return run_pack<BT, CAP, 4>(plan, sg, grads, bucket, scale, lc, s);
    }
}
template <typename BT, int CAP>
cudaError_t run_unpack_u(const b200ddp_plan *plan, const Segment &sg, const void *bucket, const void *const *grads_out,
                         float divisor, const LaunchCfg &lc, cudaStream_t s) {
    switch (plan->copy_unroll) {
        case 2: return run_unpack<BT, CAP, 2>(plan, sg, bucket, grads_out, divisor, lc, s);
        case 8: return run_unpack<BT, CAP, 8>(plan, sg, bucket, grads_out, divisor, lc, s);
        default: return run_unpack<BT, CAP, 4>(plan, sg, bucket, grads_out, divisor, lc, s);
    }
}
template <typename BT, bool FB, bool WG, int OPT, int CAP>
cudaError_t run_update(const b200ddp_plan *plan, const Segment &sg, const void *bucket, const PtrArrays &pa,
                       const SgdScalars &hp, const LaunchCfg &lc, cudaStream_t s) {
    static int occ = 0;
    TensorTab<CAP, kNumPtrKinds> tab;
    fill_tab(tab, plan, sg, pa.k);
    const int n = sg.c1 - sg.c0;
    return launch_ex(update_kernel<BT, FB, WG, OPT, CAP>, grid_for(update_kernel<BT, FB, WG, OPT, CAP>, occ, n, lc),
                     kThreads, s, lc.pdl, tab, (const Chunk *)(plan->d_chunks + sg.c0), n, (const BT *)bucket, hp);
}
template <typename BT, bool FB, bool WG, int OPT>
cudaError_t run_update_cap(const b200ddp_plan *plan, const Segment &sg, const void *bucket, const PtrArrays &pa,
                           const SgdScalars &hp, const LaunchCfg &lc, cudaStream_t s) {
    if (sg.t1 - sg.t0 <= kTabSmall) return run_update<BT, FB, WG, OPT, kTabSmall>(plan, sg, bucket, pa, hp, lc, s);
    return run_update<BT, FB, WG, OPT, kTabLarge>(plan, sg, bucket, pa, hp, lc, s);
}
template <typename BT, bool FB, bool WG>
cudaError_t run_update_opt(const b200ddp_plan *plan, const Segment &sg, const void *bucket, const PtrArrays &pa,
                           const SgdScalars &hp, const LaunchCfg &lc, cudaStream_t s) {
    if (hp.rule == kRuleAdam) return run_update_cap<BT, FB, WG, kRuleAdam>(plan, sg, bucket, pa, hp, lc, s);
    if (hp.rule == kRuleGd) return run_update_cap<BT, FB, WG, kRuleGd>(plan, sg, bucket, pa, hp, lc, s);
    return run_update_cap<BT, FB, WG, kRuleSgd>(plan, sg, bucket, pa, hp, lc, s);
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

}  // namespace

int launch_pack(const b200ddp_plan *plan, const void *const *grads, void *bucket, float scale, const LaunchCfg &lc,
                cudaStream_t s) {
    for (const Segment &sg : plan->segments) {
        if (sg.k1 == sg.k0) continue;
        const bool small = sg.t1 - sg.t0 <= kTabSmall;
        cudaError_t e;
        if (plan->dtype == B200DDP_F32)
            e = small ? run_pack_u<float, kTabSmall>(plan, sg, grads, bucket, scale, lc, s)
                      : run_pack_u<float, kTabLarge>(plan, sg, grads, bucket, scale, lc, s);
        else
            e = small ? run_pack_u<__nv_bfloat16, kTabSmall>(plan, sg, grads, bucket, scale, lc, s)
                      : run_pack_u<__nv_bfloat16, kTabLarge>(plan, sg, grads, bucket, scale, lc, s);
        B2_LAUNCH_RC(e);
    }
    return B200DDP_OK;
}

int launch_unpack(const b200ddp_plan *plan, const void *bucket, const void *const *grads_out, float divisor,
                  const LaunchCfg &lc, cudaStream_t s) {
    for (const Segment &sg : plan->segments) {
        if (sg.k1 == sg.k0) continue;
        const bool small = sg.t1 - sg.t0 <= kTabSmall;
        cudaError_t e;
        if (plan->dtype == B200DDP_F32)
            e = small ? run_unpack_u<float, kTabSmall>(plan, sg, bucket, grads_out, divisor, lc, s)
                      : run_unpack_u<float, kTabLarge>(plan, sg, bucket, grads_out, divisor, lc, s);
        else
            e = small ? run_unpack_u<__nv_bfloat16, kTabSmall>(plan, sg, bucket, grads_out, divisor, lc, s)
                      : run_unpack_u<__nv_bfloat16, kTabLarge>(plan, sg, bucket, grads_out, divisor, lc, s);
        B2_LAUNCH_RC(e);
    }
    return B200DDP_OK;
}

int launch_update(const b200ddp_plan *plan, const void *bucket, const PtrArrays &pa, const SgdScalars &hp,
                  const LaunchCfg &lc, cudaStream_t s) {
    const bool write_grad = pa.k[kGradOut] != nullptr;
    for (const Segment &sg : plan->segments) {
        if (sg.c1 == sg.c0) continue;
        cudaError_t e;
        if (bucket == nullptr) {
            e = run_update_opt<float, false, false>(plan, sg, bucket, pa, hp, lc, s);
        } else if (plan->dtype == B200DDP_F32) {
            e = write_grad ? run_update_opt<float, true, true>(plan, sg, bucket, pa, hp, lc, s)
                           : run_update_opt<float, true, false>(plan, sg, bucket, pa, hp, lc, s);
        } else {
            e = write_grad ? run_update_opt<__nv_bfloat16, true, true>(plan, sg, bucket, pa, hp, lc, s)
                           : run_update_opt<__nv_bfloat16, true, false>(plan, sg, bucket, pa, hp, lc, s);
        }
        B2_LAUNCH_RC(e);
    }
    return B200DDP_OK;
}

}  // namespace b200ddp

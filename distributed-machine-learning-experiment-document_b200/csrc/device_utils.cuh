// Device helpers: 128-bit streaming loads/stores, bf16 conversion, the per-element
// momentum-SGD arithmetic shared by the update kernels.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include "internal.h"

namespace b200ddp {

// Read-once data (gradients, the bucket): non-coherent path, do not allocate in L1.
__device__ __forceinline__ float4 ld_stream_f4(const float *p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(p));
    return r;
}
// Read-modify-write data (params, momentum): coherent path, still no L1 allocation.
__device__ __forceinline__ float4 ld_rmw_f4(const float *p) {
    float4 r;
    asm volatile("ld.global.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(p)
                 : "memory");
    return r;
}
__device__ __forceinline__ uint2 ld_stream_u2(const void *p) {
    uint2 r;
    asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
    return r;
}
__device__ __forceinline__ uint2 ld_coh_u2(const void *p) {
    uint2 r;
    asm volatile("ld.global.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p) : "memory");
    return r;
}
__device__ __forceinline__ void st_f4(float *p, const float4 &v) {
    asm volatile("st.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}
__device__ __forceinline__ void st_u2(void *p, const uint2 &v) {
    asm volatile("st.global.v2.u32 [%0], {%1,%2};" ::"l"(p), "r"(v.x), "r"(v.y) : "memory");
}

// ---- programmatic dependent launch -------------------------------------------------------
// Every kernel of the library starts with these two.  Launched with the programmatic-stream-
// serialization attribute (ctx->pdl), a kernel may be scheduled while its predecessor on the stream
// drains: its blocks take the SM slots the predecessor's finished blocks leave and park in
// griddepcontrol.wait until the predecessor has completed and flushed, which removes the launch gap
// and the ramp between pack -> reduce -> update.  Without the attribute both are no-ops.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_prologue() { pdl_launch_dependents(); pdl_wait(); }

// ---- bucket element access, 4 elements at a time -------------------------------
__device__ __forceinline__ float4 bf16x4_to_f4(uint2 raw) {
    float4 r;
    r.x = __uint_as_float(raw.x << 16);
    r.y = __uint_as_float(raw.x & 0xffff0000u);
    r.z = __uint_as_float(raw.y << 16);
    r.w = __uint_as_float(raw.y & 0xffff0000u);
    return r;
}
__device__ __forceinline__ uint2 f4_to_bf16x4(const float4 &v) {
    __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y);
    __nv_bfloat162 hi = __floats2bfloat162_rn(v.z, v.w);
    uint2 r;
    r.x = *reinterpret_cast<unsigned *>(&lo);
    r.y = *reinterpret_cast<unsigned *>(&hi);
    return r;
}
template <typename BT> struct Bucket;
template <> struct Bucket<float> {
    static constexpr uintptr_t kAlignMask = 15;
    static __device__ __forceinline__ float4 load4(const float *p) { return ld_stream_f4(p); }
    static __device__ __forceinline__ void store4(float *p, const float4 &v) { st_f4(p, v); }
    static __device__ __forceinline__ float load1(const float *p) { return __ldg(p); }
    // coherent variants, for peer memory that other GPUs wrote before the barrier
    static __device__ __forceinline__ float4 load4c(const float *p) { return ld_rmw_f4(p); }
    static __device__ __forceinline__ float load1c(const float *p) { return *reinterpret_cast<const volatile float *>(p); }
    static __device__ __forceinline__ void store1(float *p, float v) { *p = v; }
};
template <> struct Bucket<__nv_bfloat16> {
    static constexpr uintptr_t kAlignMask = 7;
    static __device__ __forceinline__ float4 load4(const __nv_bfloat16 *p) { return bf16x4_to_f4(ld_stream_u2(p)); }
    static __device__ __forceinline__ void store4(__nv_bfloat16 *p, const float4 &v) { st_u2(p, f4_to_bf16x4(v)); }
    static __device__ __forceinline__ float load1(const __nv_bfloat16 *p) {
        return __uint_as_float((unsigned)(*reinterpret_cast<const unsigned short *>(p)) << 16);
    }
    static __device__ __forceinline__ float4 load4c(const __nv_bfloat16 *p) { return bf16x4_to_f4(ld_coh_u2(p)); }
    static __device__ __forceinline__ float load1c(const __nv_bfloat16 *p) {
        return __uint_as_float((unsigned)(*reinterpret_cast<const volatile unsigned short *>(p)) << 16);
    }
    static __device__ __forceinline__ void store1(__nv_bfloat16 *p, float v) { *p = __float2bfloat16_rn(v); }
};

// ---- x / d ------------------------------------------------------------------------------
// See DivKind in internal.h for which rule stands for which run of the reference.  The IEEE rule:
// __fdiv_rn costs ~50 instructions per element and makes the kernel issue-bound (ncu, round 1:
// unpack 27 us vs pack 14 us for the same bytes).  For a divisor known on the host, Markstein's
// sequence gives the same correctly rounded quotient in three: q = RN(x * rcp), r = fma(-d, q, x)
// (exact), q' = fma(r, rcp, q), valid while quotient and remainder stay in the normal range.
__device__ __forceinline__ float apply_div(float x, const Divisor &dv) {
    if (dv.kind == kDivNone) return x;
    if (dv.kind == kDivRcp) return __fmul_rn(x, dv.rcp);
    if (dv.kind == kDivFast) {
        const float ax = fabsf(x);
        if (ax > 1e-30f && ax < 1e30f) {
            const float q = __fmul_rn(x, dv.rcp);
            return __fmaf_rn(__fmaf_rn(-dv.d, q, x), dv.rcp, q);
        }
        if (ax == 0.0f) return x;          // +-0 / d = +-0
    }
    return __fdiv_rn(x, dv.d);
}

// ---- the update rule on one element --------------------------------------------------------
// g_avg returns what p.grad holds after the reference's sync (codes/task2/dist_utils.py:41-42):
// the averaged gradient, before the optimizer's own transforms (those are out of place).
//
// kRuleSgd: rounding sequence of torch/optim/sgd.py:322-380 on fp32 tensors, where every
//   add_(x, alpha=a) / add(x, alpha=a) is a single fused multiply-add and mul_ is a rounded multiply.
// kRuleGd:  codes/task1/pytorch/MyOptimizer.py:23  p = p - lr * g      (two roundings: product, difference)
// kRuleAdam: codes/task1/pytorch/MyOptimizer.py:38-43 (no bias correction), every tensor op rounded on its own:
//   m = b1 * m + (1 - b1) * g;  v = b2 * v + (1 - b2) * g**2;  p = p - (lr / (sqrt(v) + eps)) * m
//   where `lr / tensor` is torch's Tensor.__rtruediv__: reciprocal(tensor) * lr.
template <int RULE>
__device__ __forceinline__ void opt_elem(float g, float &p, float &m, float &v, float &g_avg, const SgdScalars &hp) {
    g = apply_div(g, hp.dv);
    g_avg = g;
    if (RULE == kRuleGd) {
        p = __fsub_rn(p, __fmul_rn(hp.lr, g));
    } else if (RULE == kRuleAdam) {
        m = __fadd_rn(__fmul_rn(hp.beta1, m), __fmul_rn(hp.one_minus_beta1, g));
        v = __fadd_rn(__fmul_rn(hp.beta2, v), __fmul_rn(hp.one_minus_beta2, __fmul_rn(g, g)));
        const float den = __fadd_rn(__fsqrt_rn(v), hp.eps);
        p = __fsub_rn(p, __fmul_rn(__fmul_rn(__frcp_rn(den), hp.lr), m));
    } else {
        if (hp.maximize) g = -g;
        if (hp.has_wd) g = __fmaf_rn(hp.weight_decay, p, g);
        if (hp.has_momentum) {
            m = hp.first_step ? g : __fmaf_rn(hp.one_minus_damp, g, __fmul_rn(m, hp.momentum));
            g = hp.nesterov ? __fmaf_rn(hp.momentum, m, g) : m;
        }
        p = __fmaf_rn(hp.neg_lr, g, p);
    }
}
template <int RULE>
__device__ __forceinline__ void opt_elem4(const float4 &g, float4 &p, float4 &m, float4 &v, float4 &ga,
                                          const SgdScalars &hp) {
    opt_elem<RULE>(g.x, p.x, m.x, v.x, ga.x, hp);
    opt_elem<RULE>(g.y, p.y, m.y, v.y, ga.y, hp);
    opt_elem<RULE>(g.z, p.z, m.z, v.z, ga.z, hp);
    opt_elem<RULE>(g.w, p.w, m.w, v.w, ga.w, hp);
}
// the momentum-SGD rule, for the fused one-shot kernel
__device__ __forceinline__ void sgd_elem(float g, float &p, float &m, float &g_avg, const SgdScalars &hp) {
    float v = 0.f;
    opt_elem<kRuleSgd>(g, p, m, v, g_avg, hp);
}
__device__ __forceinline__ void sgd_elem4(const float4 &g, float4 &p, float4 &m, float4 &ga, const SgdScalars &hp) {
    sgd_elem(g.x, p.x, m.x, ga.x, hp);
    sgd_elem(g.y, p.y, m.y, ga.y, hp);
    sgd_elem(g.z, p.z, m.z, ga.z, hp);
    sgd_elem(g.w, p.w, m.w, ga.w, hp);
}

}  // namespace b200ddp

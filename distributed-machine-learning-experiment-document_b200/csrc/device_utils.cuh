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

// ---- x / d, correctly rounded ----------------------------------------------------------
// `param.grad.data /= world_size` (codes/task2/dist_utils.py:42) is an IEEE division, and for
// world sizes that are not powers of two it cannot be folded into an exact pre-scale.
// __fdiv_rn costs ~50 instructions per element and makes the kernel issue-bound (ncu, round 1:
// unpack 27 us vs pack 14 us for the same bytes).  For a divisor known on the host, Markstein's
// sequence gives the same correctly rounded quotient in three: q = RN(x * rcp), r = fma(-d, q, x)
// (exact), q' = fma(r, rcp, q), valid while quotient and remainder stay in the normal range.
__device__ __forceinline__ float div_rn(float x, float d, float rcp, int fast) {
    if (fast) {
        const float ax = fabsf(x);
        if (ax > 1e-30f && ax < 1e30f) {
            const float q = __fmul_rn(x, rcp);
            return __fmaf_rn(__fmaf_rn(-d, q, x), rcp, q);
        }
        if (ax == 0.0f) return x;          // +-0 / d = +-0
    }
    return __fdiv_rn(x, d);
}

// ---- momentum SGD on one element ------------------------------------------------
// Rounding sequence of torch/optim/sgd.py:322-380 on fp32 tensors, where every
// add_(x, alpha=a) / add(x, alpha=a) is a single fused multiply-add and mul_ is
// a rounded multiply.  g_avg returns what p.grad holds after the reference's
// sync (codes/task2/dist_utils.py:41-42): the averaged gradient, before the
// optimizer's own weight-decay / maximize transforms (those are out of place).
__device__ __forceinline__ void sgd_elem(float g, float &p, float &m, float &g_avg, const SgdScalars &hp) {
    if (hp.use_div) g = div_rn(g, hp.divisor, hp.div_rcp, hp.div_fast);
    g_avg = g;
    if (hp.maximize) g = -g;
    if (hp.has_wd) g = __fmaf_rn(hp.weight_decay, p, g);
    if (hp.has_momentum) {
        m = hp.first_step ? g : __fmaf_rn(hp.one_minus_damp, g, __fmul_rn(m, hp.momentum));
        g = hp.nesterov ? __fmaf_rn(hp.momentum, m, g) : m;
    }
    p = __fmaf_rn(hp.neg_lr, g, p);
}
__device__ __forceinline__ void sgd_elem4(const float4 &g, float4 &p, float4 &m, float4 &ga, const SgdScalars &hp) {
    sgd_elem(g.x, p.x, m.x, ga.x, hp);
    sgd_elem(g.y, p.y, m.y, ga.y, hp);
    sgd_elem(g.z, p.z, m.z, ga.z, hp);
    sgd_elem(g.w, p.w, m.w, ga.w, hp);
}

}  // namespace b200ddp

// Element updates of the fused optimizer steps, shared by the step kernels (rk_step.cu) and
// the sharded step kernel (rk_comm.cu): per-group constant tables, the AdamW / Adam / SGD
// update spelled in torch-CPU rounding order, zero-safe IEEE div/sqrt, explicit
// ld.global / st.global helpers.
#pragma once
#include "rk_common.cuh"

// ---- per-group constants, passed by value in kernel-parameter space ------------------
struct AdamDev {
  float decay;     // (float)(1 - lr*wd)         decoupled
  float wd;        // (float)wd                  coupled (L2)
  float w1;        // (float)(1 - beta1)
  float w1m1;      // w1 - 1.0f                  (lerp, weight >= 0.5 branch)
  float beta2;
  float w2;        // (float)(1 - beta2)
  float s2;        // (float)sqrt(bias_correction2)
  float eps;
  float neg_step;  // (float)(-lr / bias_correction1)
  int flags;
};
#define ADAM_DECOUPLED 1
#define ADAM_HAS_WD 2
#define GROUP_SKIP 4
#define ADAM_LERP_SMALL 8
struct AdamTable { AdamDev g[RK_MAX_GROUPS]; };

struct SgdDev {
  float wd, mu, one_minus_damp, neg_lr;
  int flags;
};
#define SGD_HAS_WD 1
#define SGD_HAS_MOM 2
#define SGD_NESTEROV 8
#define SGD_FIRST 16
struct SgdTable { SgdDev g[RK_MAX_GROUPS]; };

struct Prolog {       // gradient pre-scaling shared by all optimizers
  float scale;        // grad_scale (1/world_size)
  float coef;         // clip coefficient
  int has_scale, has_clip;
};

__device__ __forceinline__ float prolog_g(float g, const Prolog& pr) {
  if (pr.has_scale) g = __fmul_rn(g, pr.scale);
  if (pr.has_clip) g = __fmul_rn(g, pr.coef);  // clip_grad.py:182  g.mul_(clip_coef_clamped)
  return g;
}

// ---- IEEE division / square root without the slow-path subroutine on zeros ------------
// __fdiv_rn / __fsqrt_rn branch to a ~100-instruction subroutine when an operand is zero
// (or denormal), and the whole warp waits.  Real gradients are full of exact zeros (dead
// ReLUs, unused embedding rows): in a ResNet-18 step 74 % of warps took that path.  These
// wrappers give the same bits -- 0/b = signed zero, sqrt(+-0) = +-0 -- while feeding the
// hardware sequence an operand that keeps it on its fast path.
__device__ __forceinline__ float div_rn(float a, float b) {
  const bool zero_num = (a == 0.0f) && (b != 0.0f) && (b == b);
  const float q = __fdiv_rn(zero_num ? 1.0f : a, b);
  const float z = __int_as_float((__float_as_int(a) ^ __float_as_int(b)) & 0x80000000);
  return zero_num ? z : q;
}
__device__ __forceinline__ float sqrt_rn(float x) {
  const bool zero = (x == 0.0f);
  const float r = __fsqrt_rn(zero ? 1.0f : x);
  return zero ? x : r;
}

// ---- element updates -----------------------------------------------------------------
struct AdamOp {
  static constexpr bool kHasV = true;
  AdamDev h;
  __device__ __forceinline__ void operator()(float& p, float g, float& m, float& v) const {
    if (h.flags & ADAM_HAS_WD) {
      if (h.flags & ADAM_DECOUPLED) p = __fmul_rn(p, h.decay);  // adam.py:419
      else g = __fmaf_rn(h.wd, p, g);                           // adam.py:430
    }
    const float d = __fsub_rn(g, m);                            // adam.py:457 lerp_
    m = (h.flags & ADAM_LERP_SMALL) ? __fmaf_rn(h.w1, d, m) : __fmaf_rn(h.w1m1, d, g);
    v = __fmaf_rn(__fmul_rn(h.w2, g), g, __fmul_rn(v, h.beta2));              // adam.py:476
    const float denom = __fadd_rn(div_rn(sqrt_rn(v), h.s2), h.eps);           // adam.py:545
    p = __fadd_rn(p, div_rn(__fmul_rn(h.neg_step, m), denom));                // adam.py:547
  }
};

struct SgdOp {
  static constexpr bool kHasV = false;
  SgdDev h;
  __device__ __forceinline__ void operator()(float& p, float g, float& m, float&) const {
    if (h.flags & SGD_HAS_WD) g = __fmaf_rn(h.wd, p, g);                        // sgd.py:356
    if (h.flags & SGD_HAS_MOM) {
      if (h.flags & SGD_FIRST) m = g;                                           // sgd.py:362
      else m = __fmaf_rn(h.one_minus_damp, g, __fmul_rn(m, h.mu));              // sgd.py:365
      g = (h.flags & SGD_NESTEROV) ? __fmaf_rn(h.mu, m, g) : m;                 // sgd.py:368-370
    }
    p = __fmaf_rn(h.neg_lr, g, p);                                              // sgd.py:380
  }
};

// ---- global memory access: explicit ld.global / st.global (pointers come out of the
// descriptor table, so the compiler would otherwise emit generic LD/ST) ------------------
template <int HINT> __device__ __forceinline__ float4 ldg4(const float* p) {
  float4 r;
  if (HINT == 1)
    asm volatile("ld.global.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
  else if (HINT == 2)
    asm volatile("ld.global.cs.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
  else
    asm volatile("ld.global.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
  return r;
}
template <int HINT> __device__ __forceinline__ void stg4(float* p, const float4& v) {
  if (HINT == 2)
    asm volatile("st.global.cs.v4.f32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
  else
    asm volatile("st.global.v4.f32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
// 256-bit accesses (sm_100, PTX 8.8): half the LSU instructions per byte of the 128-bit form.
struct float8 { float4 lo, hi; };
__device__ __forceinline__ float8 ldg8(const float* p) {
  float8 r;
  asm volatile("ld.global.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(r.lo.x), "=f"(r.lo.y), "=f"(r.lo.z), "=f"(r.lo.w),
                 "=f"(r.hi.x), "=f"(r.hi.y), "=f"(r.hi.z), "=f"(r.hi.w) : "l"(p));
  return r;
}
__device__ __forceinline__ void stg8(float* p, const float8& v) {
  asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
               :: "l"(p), "f"(v.lo.x), "f"(v.lo.y), "f"(v.lo.z), "f"(v.lo.w),
                  "f"(v.hi.x), "f"(v.hi.y), "f"(v.hi.z), "f"(v.hi.w) : "memory");
}
__device__ __forceinline__ float ldg1(const float* p) {
  float r;
  asm volatile("ld.global.f32 %0, [%1];" : "=f"(r) : "l"(p));
  return r;
}
__device__ __forceinline__ void stg1(float* p, float v) {
  asm volatile("st.global.f32 [%0], %1;" :: "l"(p), "f"(v) : "memory");
}

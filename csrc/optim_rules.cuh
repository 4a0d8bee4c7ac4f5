// Optimizer update rules as device functors.  Shared by the stand-alone fused update kernel
// (optim.cu) and by the epilogues of the fused all-reduce kernels (comm_p2p.cu), so the five
// rules x {one-shot, two-shot, NVLS} x {fp32, bf16} matrix is a template product, not copies.
//
// Reference: the optimizer table at /root/reference/src/trainer.py:123-138 (torch.optim SGD / Adam /
// Adagrad / Adamax / AdamW with torch defaults).  The arithmetic below matches torch's single-tensor
// implementations statement by statement; ml_trainer_b200/optim/rules.py is the tested oracle.
#pragma once
#include "common.cuh"

namespace mlt {

// hyper vector layout (device memory, refreshed by the host every step; keeps graphs replayable)
struct Hyper {
  float lr, wd, b1, b2, eps, bc1, bc2_sqrt, grad_scale;
};
static_assert(sizeof(Hyper) == 32, "hyper vector is 8 floats");

enum RuleId : int { RULE_SGD = 0, RULE_ADAM = 1, RULE_ADAMW = 2, RULE_ADAGRAD = 3, RULE_ADAMAX = 4, RULE_MEAN_ONLY = 5 };

// Each rule: update(p, g, s1, s2, h) in registers.  g arrives already scaled.
struct SgdRule {
  static constexpr int kStates = 1;
  __device__ __forceinline__ static void apply(float& p, float g, float& s1, float& s2, const Hyper& h) {
    g = fmaf(h.wd, p, g);
    if (h.b1 != 0.0f) {           // momentum buffer; zero-initialised so step 1 gives buf = g
      s1 = fmaf(h.b1, s1, g);
      g = s1;
    }
    p = fmaf(-h.lr, g, p);
  }
};

__device__ __forceinline__ void adam_core(float& p, float g, float& m, float& v, const Hyper& h) {
  m = fmaf(1.0f - h.b1, g - m, m);                 // lerp(m, g, 1-b1)
  v = fmaf(h.b2, v, (1.0f - h.b2) * g * g);
  const float denom = sqrtf(v) / h.bc2_sqrt + h.eps;
  p = fmaf(-(h.lr / h.bc1), m / denom, p);
}

struct AdamRule {
  static constexpr int kStates = 2;
  __device__ __forceinline__ static void apply(float& p, float g, float& s1, float& s2, const Hyper& h) {
    g = fmaf(h.wd, p, g);                          // L2 penalty folded into the gradient
    adam_core(p, g, s1, s2, h);
  }
};

struct AdamWRule {
  static constexpr int kStates = 2;
  __device__ __forceinline__ static void apply(float& p, float g, float& s1, float& s2, const Hyper& h) {
    p *= (1.0f - h.lr * h.wd);                     // decoupled decay
    adam_core(p, g, s1, s2, h);
  }
};

struct AdagradRule {
  static constexpr int kStates = 1;
  __device__ __forceinline__ static void apply(float& p, float g, float& s1, float& s2, const Hyper& h) {
    g = fmaf(h.wd, p, g);
    s1 = fmaf(g, g, s1);
    p = fmaf(-h.lr, g / (sqrtf(s1) + h.eps), p);
  }
};

struct AdamaxRule {
  static constexpr int kStates = 2;
  __device__ __forceinline__ static void apply(float& p, float g, float& s1, float& s2, const Hyper& h) {
    g = fmaf(h.wd, p, g);
    s1 = fmaf(1.0f - h.b1, g - s1, s1);
    s2 = fmaxf(h.b2 * s2, fabsf(g) + h.eps);
    p = fmaf(-(h.lr / h.bc1), s1 / s2, p);
  }
};

// "Rule" that only hands back the (scaled) reduced gradient: turns the fused kernels into a plain mean all-reduce
// (the semantics of the reference's _average_gradients, src/trainer.py:152-158).
struct MeanOnlyRule {
  static constexpr int kStates = 1;
  __device__ __forceinline__ static void apply(float& p, float g, float& s1, float& s2, const Hyper& h) { p = g; }
};

// Dispatch a runtime rule id onto a functor-templated callable:  MLT_DISPATCH_RULE(id, R, { foo<R>(...); })
#define MLT_DISPATCH_RULE(rule_id, R, ...)                         \
  switch (rule_id) {                                               \
    case ::mlt::RULE_SGD:     { using R = ::mlt::SgdRule;     __VA_ARGS__; break; } \
    case ::mlt::RULE_ADAM:    { using R = ::mlt::AdamRule;    __VA_ARGS__; break; } \
    case ::mlt::RULE_ADAMW:   { using R = ::mlt::AdamWRule;   __VA_ARGS__; break; } \
    case ::mlt::RULE_ADAGRAD: { using R = ::mlt::AdagradRule; __VA_ARGS__; break; } \
    case ::mlt::RULE_ADAMAX:  { using R = ::mlt::AdamaxRule;  __VA_ARGS__; break; } \
    case ::mlt::RULE_MEAN_ONLY: { using R = ::mlt::MeanOnlyRule; __VA_ARGS__; break; } \
    default: ::mlt::set_error("unknown optimizer rule id %d", (int)(rule_id)); return 1; \
  }

// Apply a rule to 4 consecutive elements held in registers and write everything back.
// G is the gradient storage type (float or bf16); M the optional compute-dtype shadow (bf16) or void.
template <typename R>
__device__ __forceinline__ void update4(float4& p, const float4& g, float4& a, float4& b, const Hyper& h) {
  R::apply(p.x, g.x, a.x, b.x, h);
  R::apply(p.y, g.y, a.y, b.y, h);
  R::apply(p.z, g.z, a.z, b.z, h);
  R::apply(p.w, g.w, a.w, b.w, h);
}

__device__ __forceinline__ float4 load_grad4(const float* g, long long i) {
  return *reinterpret_cast<const float4*>(g + i);
}
__device__ __forceinline__ float4 load_grad4(const __nv_bfloat16* g, long long i) {
  const uint2 raw = *reinterpret_cast<const uint2*>(g + i);
  const __nv_bfloat162 lo = *reinterpret_cast<const __nv_bfloat162*>(&raw.x);
  const __nv_bfloat162 hi = *reinterpret_cast<const __nv_bfloat162*>(&raw.y);
  const float2 a = __bfloat1622float2(lo), b = __bfloat1622float2(hi);
  return make_float4(a.x, a.y, b.x, b.y);
}
__device__ __forceinline__ uint2 pack_bf16x4(const float4& v) {
  __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
  uint2 r;
  r.x = *reinterpret_cast<uint32_t*>(&lo);
  r.y = *reinterpret_cast<uint32_t*>(&hi);
  return r;
}

}  // namespace mlt

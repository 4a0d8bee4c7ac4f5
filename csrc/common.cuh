// Shared host/device helpers for libmlt_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cstdint>
#include <cstdio>
#include <cstdarg>

namespace mlt {

// ---- error reporting --------------------------------------------------------------------
// Every extern "C" entry point returns 0 on success, otherwise a non-zero code and a message
// retrievable through mlt_last_error() (thread-local).
void set_error(const char* fmt, ...);
const char* last_error();

#define MLT_CHECK_CUDA(expr)                                                            \
  do {                                                                                  \
    cudaError_t _e = (expr);                                                            \
    if (_e != cudaSuccess) {                                                            \
      ::mlt::set_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
      return (int)_e;                                                                   \
    }                                                                                   \
  } while (0)

#define MLT_REQUIRE(cond, ...)                       \
  do {                                               \
    if (!(cond)) {                                   \
      ::mlt::set_error(__VA_ARGS__);                 \
      return 1;                                      \
    }                                                \
  } while (0)

#define MLT_LAUNCH_CHECK() MLT_CHECK_CUDA(cudaGetLastError())

constexpr int kNumSMs = 148;   // B200: 2 dies x 74

__host__ __device__ inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
__host__ __device__ inline long long ceil_div64(long long a, long long b) { return (a + b - 1) / b; }

// ---- device helpers -----------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

template <typename T> __device__ __forceinline__ float to_float(T v);
template <> __device__ __forceinline__ float to_float<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_float<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_float(float v);
template <> __device__ __forceinline__ float from_float<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_float<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

// activation ids shared by every epilogue in the library
enum Act : int { ACT_NONE = 0, ACT_RELU = 1, ACT_GELU = 2, ACT_TANH = 3 };

__device__ __forceinline__ float gelu_fwd(float x) {   // exact (erf) GELU, what torch.nn.GELU() computes
  return 0.5f * x * (1.0f + erff(x * 0.70710678118654752f));
}
__device__ __forceinline__ float gelu_bwd(float x) {
  const float cdf = 0.5f * (1.0f + erff(x * 0.70710678118654752f));
  const float pdf = 0.39894228040143268f * __expf(-0.5f * x * x);
  return cdf + x * pdf;
}
// GELU and its derivative for the GEMM epilogues, one MUFU op per element: Phi(z) = 0.5 (1 + erf(z / sqrt 2)) with
// erf(x) ~ tanh(x (c1 + c2 x^2 + c3 x^4)), coefficients fitted (minimax, scipy) to |erf error| < 3.7e-5; tanh.approx.f32
// adds ~2^-11 relative.  Total |GELU error| < 3e-4 max(1, |z|), below the bf16 rounding of the stored result, and the
// derivative is the exact derivative of the function that is evaluated (what training needs).
__device__ __forceinline__ void gelu_both(float z, float& g, float& dg) {
  constexpr float c1 = 1.128143378f, c2 = 0.104081189f, c3 = -0.001786476f;
  const float x = z * 0.70710678118654752f;
  const float x2 = x * x;
  const float u = x * fmaf(fmaf(c3, x2, c2), x2, c1);
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(u));
  const float cdf = fmaf(0.5f, t, 0.5f);
  g = z * cdf;
  const float du = fmaf(fmaf(5.0f * c3, x2, 3.0f * c2), x2, c1);          // d u / d x
  dg = fmaf(z * 0.35355339059327376f * fmaf(-t, t, 1.0f), du, cdf);       // cdf + z * 0.5 sech^2(u) u'(x) / sqrt 2
}
template <int ACT> __device__ __forceinline__ float act_fwd(float x) {
  if (ACT == ACT_RELU) return fmaxf(x, 0.0f);
  if (ACT == ACT_GELU) return gelu_fwd(x);
  if (ACT == ACT_TANH) return tanhf(x);
  return x;
}
__device__ __forceinline__ float act_fwd_rt(int act, float x) {
  switch (act) {
    case ACT_RELU: return fmaxf(x, 0.0f);
    case ACT_GELU: return gelu_fwd(x);
    case ACT_TANH: return tanhf(x);
    default: return x;
  }
}

}  // namespace mlt

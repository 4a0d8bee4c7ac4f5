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
// GELU and its derivative from one exponential (GEMM epilogues): erfc via Abramowitz-Stegun 7.1.26 (|error| < 1.5e-7, far
// below the bf16 the results are rounded to); exp(-z^2/2) is shared between the erfc and the density term of the derivative.
__device__ __forceinline__ void gelu_both(float z, float& g, float& dg) {
  const float az = fabsf(z) * 0.70710678118654752f;
  const float e = __expf(-az * az);                       // exp(-z^2 / 2)
  const float t = __fdividef(1.0f, fmaf(0.3275911f, az, 1.0f));
  float poly = fmaf(1.061405429f, t, -1.453152027f);
  poly = fmaf(poly, t, 1.421413741f);
  poly = fmaf(poly, t, -0.284496736f);
  poly = fmaf(poly, t, 0.254829592f);
  const float half_erfc = 0.5f * poly * t * e;            // 0.5 * erfc(|z| / sqrt 2)
  const float cdf = z >= 0.f ? 1.0f - half_erfc : half_erfc;
  g = z * cdf;
  dg = fmaf(z * 0.39894228040143268f, e, cdf);
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

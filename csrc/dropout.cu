// Dropout with a counter-based generator (Philox4x32-10): the mask is a pure function of (seed, step, call site, element),
// so backward regenerates it instead of storing it, and a captured CUDA graph replays with fresh masks because `step` is
// read from device memory.  One Philox call covers 8 elements (16 random bits each).  y = x * keep / (1 - p).
// Not in the reference (it has no dropout layer); needed by the driver-named BERT workload.
#include "common.cuh"
#include "../include/mlt_b200.h"

namespace mlt {

__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) {
  constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
    const uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += W0;
    key.y += W1;
  }
  return ctr;
}

template <typename T>
__global__ void __launch_bounds__(256)
dropout_kernel(const T* __restrict__ x, T* __restrict__ y, long long n, uint32_t thresh16, float scale, unsigned long long seed,
               const long long* __restrict__ step_ptr, uint32_t call_id) {
  const long long step = step_ptr ? *step_ptr : 0;
  const long long groups = (n + 7) / 8;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < groups; g += stride) {
    const uint4 r = philox4x32_10(make_uint4((uint32_t)g, (uint32_t)(g >> 32), call_id, (uint32_t)step),
                                  make_uint2((uint32_t)seed, (uint32_t)(seed >> 32) ^ (uint32_t)(step >> 32)));
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
    const long long e0 = g * 8;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const long long e = e0 + j;
      if (e < n) {
        const uint32_t bits = (w[j >> 1] >> ((j & 1) * 16)) & 0xFFFFu;
        y[e] = from_float<T>(bits >= thresh16 ? to_float(x[e]) * scale : 0.f);
      }
    }
  }
}

}  // namespace mlt

using namespace mlt;

// y = dropout(x) with drop probability p (resolution 2^-16).  Calling it again with the same (seed, *step, call_id) on the
// upstream gradient is the backward pass.  x == y (in place) is allowed.
extern "C" int mlt_dropout(const void* x, void* y, long long n, float p, unsigned long long seed, const void* step_ptr, unsigned int call_id,
                           int is_bf16, void* stream) {
  MLT_REQUIRE(p >= 0.f && p < 1.f, "mlt_dropout: p must be in [0, 1)");
  if (n <= 0) return 0;
  const uint32_t thresh = (uint32_t)(p * 65536.0f + 0.5f);
  const float scale = 1.0f / (1.0f - (float)thresh / 65536.0f);
  long long blocks = ceil_div64((n + 7) / 8, 256);
  if (blocks > (long long)kNumSMs * 16) blocks = (long long)kNumSMs * 16;
  cudaStream_t st = (cudaStream_t)stream;
  if (is_bf16)
    dropout_kernel<__nv_bfloat16><<<(int)blocks, 256, 0, st>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, n, thresh, scale, seed,
                                                              (const long long*)step_ptr, call_id);
  else
    dropout_kernel<float><<<(int)blocks, 256, 0, st>>>((const float*)x, (float*)y, n, thresh, scale, seed, (const long long*)step_ptr, call_id);
  MLT_LAUNCH_CHECK();
  return 0;
}

// Fused optimizer update over a flat arena slice (SURVEY §2.3 N14, §8 K1-K3 epilogue when world == 1).
// One launch replaces torch's foreach mul_/add_ chains: grad scale + bf16->fp32 cast + rule + fp32 master
// write + compute-dtype shadow write.  Memory-bound: 16-28 B/element; the roofline is bytes / HBM bandwidth.
#include "optim_rules.cuh"
#include "../include/mlt_b200.h"

namespace mlt {

template <typename R, typename G>
__global__ void __launch_bounds__(256)
fused_update_kernel(float* __restrict__ master, const G* __restrict__ grad, __nv_bfloat16* __restrict__ shadow,
                    float* __restrict__ s1, float* __restrict__ s2, const Hyper* __restrict__ hyper, long long n4) {
  const Hyper h = *hyper;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x; v < n4; v += stride) {
    const long long i = v * 4;
    float4 p = *reinterpret_cast<const float4*>(master + i);
    float4 g = load_grad4(grad, i);
    g.x *= h.grad_scale; g.y *= h.grad_scale; g.z *= h.grad_scale; g.w *= h.grad_scale;
    float4 a = *reinterpret_cast<const float4*>(s1 + i);
    float4 b = make_float4(0.f, 0.f, 0.f, 0.f);
    if (R::kStates > 1) b = *reinterpret_cast<const float4*>(s2 + i);
    update4<R>(p, g, a, b, h);
    *reinterpret_cast<float4*>(master + i) = p;
    *reinterpret_cast<float4*>(s1 + i) = a;
    if (R::kStates > 1) *reinterpret_cast<float4*>(s2 + i) = b;
    if (shadow != nullptr) *reinterpret_cast<uint2*>(shadow + i) = pack_bf16x4(p);
  }
}

template <typename R, typename G>
static int launch_update(float* master, const void* grad, void* shadow, float* s1, float* s2, const void* hyper,
                         long long n, cudaStream_t st) {
  const long long n4 = n / 4;
  const int threads = 256;
  long long blocks = ceil_div64(n4, threads);
  const long long cap = (long long)kNumSMs * 8;       // 8 resident CTAs/SM of 256 threads; grid-stride beyond
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  fused_update_kernel<R, G><<<(int)blocks, threads, 0, st>>>(
      master, reinterpret_cast<const G*>(grad), reinterpret_cast<__nv_bfloat16*>(shadow), s1, s2,
      reinterpret_cast<const Hyper*>(hyper), n4);
  MLT_LAUNCH_CHECK();
  return 0;
}

}  // namespace mlt

extern "C" int mlt_fused_update(int rule, void* master_f32, const void* grad, int grad_is_bf16, void* shadow_bf16,
                                void* s1, void* s2, const void* hyper_dev, long long n, void* stream) {
  using namespace mlt;
  MLT_REQUIRE(n % 4 == 0, "mlt_fused_update: n (%lld) must be a multiple of 4", n);
  MLT_REQUIRE(((uintptr_t)master_f32 % 16) == 0 && ((uintptr_t)grad % 8) == 0, "mlt_fused_update: unaligned arena slice");
  cudaStream_t st = (cudaStream_t)stream;
  if (n == 0) return 0;
  MLT_DISPATCH_RULE(rule, R, {
    if (grad_is_bf16) return launch_update<R, __nv_bfloat16>((float*)master_f32, grad, shadow_bf16, (float*)s1, (float*)s2, hyper_dev, n, st);
    return launch_update<R, float>((float*)master_f32, grad, shadow_bf16, (float*)s1, (float*)s2, hyper_dev, n, st);
  });
  return 0;
}

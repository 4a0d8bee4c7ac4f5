// Fused softmax cross-entropy forward+backward and arg-max accuracy.
//
// One pass produces: the mean loss (scalar), dlogits = (softmax - onehot) / B, and, when asked, the batch
// accuracy; loss / accuracy / step-count are also added into the trainer's 3-float device accumulator, so
// nothing is read back per step (the reference does loss.item() + two .cpu() per step, trainer.py:186,166).
// Softmax is monotone, so accuracy is the arg-max of the raw logits (ties -> lowest index, as torch.argmax).
//
// Cross-CTA reduction uses a 16-byte workspace {float loss_sum, float correct, uint ticket, pad}: it must be
// zero before the first launch; the last CTA to finish consumes it and zeroes it again for the next launch.
#include "common.cuh"
#include "../include/mlt_b200.h"

namespace mlt {

struct XentWorkspace {
  float loss_sum;
  float correct;
  unsigned int ticket;
  unsigned int pad;
};

template <typename T>
__global__ void __launch_bounds__(256)
xent_fwd_bwd_kernel(const T* __restrict__ logits, const long long* __restrict__ targets, T* __restrict__ dlogits,
                    float* __restrict__ loss_out, float* __restrict__ acc3, XentWorkspace* __restrict__ ws,
                    int B, int C, float inv_B, int want_acc) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  float loss_part = 0.f, correct_part = 0.f;
  for (int row = blockIdx.x * nwarp + warp; row < B; row += gridDim.x * nwarp) {
    const T* x = logits + (size_t)row * C;
    const int tgt = (int)targets[row];
    // pass 1: max and first arg-max
    float m = -INFINITY;
    int am = 0x7fffffff;
    for (int c = lane; c < C; c += 32) {
      const float v = to_float(x[c]);
      if (v > m) { m = v; am = c; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float om = __shfl_xor_sync(0xffffffffu, m, o);
      const int oa = __shfl_xor_sync(0xffffffffu, am, o);
      if (om > m || (om == m && oa < am)) { m = om; am = oa; }
    }
    // pass 2: sum of exp
    float s = 0.f;
    for (int c = lane; c < C; c += 32) s += expf(to_float(x[c]) - m);
    s = warp_sum(s);
    const float log_z = logf(s) + m;
    const float inv_s = 1.0f / s;
    // pass 3: gradient
    if (dlogits != nullptr) {
      T* dx = dlogits + (size_t)row * C;
      for (int c = lane; c < C; c += 32) {
        float p = expf(to_float(x[c]) - m) * inv_s;
        if (c == tgt) p -= 1.0f;
        dx[c] = from_float<T>(p * inv_B);
      }
    }
    if (lane == 0) {
      const float xt = (tgt >= 0 && tgt < C) ? to_float(x[tgt]) : log_z;   // out-of-range target contributes 0
      loss_part += log_z - xt;
      correct_part += (am == tgt) ? 1.0f : 0.0f;
    }
  }
  __shared__ float sh_loss[8], sh_corr[8];
  __shared__ bool is_last;
  if (lane == 0) { sh_loss[warp] = loss_part; sh_corr[warp] = correct_part; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float l = 0.f, c = 0.f;
    for (int w = 0; w < nwarp; ++w) { l += sh_loss[w]; c += sh_corr[w]; }
    atomicAdd(&ws->loss_sum, l);
    atomicAdd(&ws->correct, c);
    __threadfence();
    const unsigned int t = atomicAdd(&ws->ticket, 1u);
    is_last = (t == gridDim.x - 1);
    if (is_last) {
      __threadfence();
      const float lsum = atomicExch(&ws->loss_sum, 0.0f);   // consume and re-zero for the next launch
      const float csum = atomicExch(&ws->correct, 0.0f);
      ws->ticket = 0u;
      const float loss = lsum * inv_B;
      if (loss_out) *loss_out = loss;
      if (acc3) {
        acc3[0] += loss;
        if (want_acc) acc3[1] += csum * inv_B;
        acc3[2] += 1.0f;
      }
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(1024)
argmax_accuracy_kernel(const T* __restrict__ logits, const long long* __restrict__ targets, float* __restrict__ out,
                       int B, int C) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  float correct = 0.f;
  for (int row = warp; row < B; row += nwarp) {
    const T* x = logits + (size_t)row * C;
    float m = -INFINITY;
    int am = 0x7fffffff;
    for (int c = lane; c < C; c += 32) {
      const float v = to_float(x[c]);
      if (v > m) { m = v; am = c; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float om = __shfl_xor_sync(0xffffffffu, m, o);
      const int oa = __shfl_xor_sync(0xffffffffu, am, o);
      if (om > m || (om == m && oa < am)) { m = om; am = oa; }
    }
    if (lane == 0 && am == (int)targets[row]) correct += 1.0f;
  }
  __shared__ float sh[32];
  if (lane == 0) sh[warp] = correct;
  __syncthreads();
  if (warp == 0) {
    float v = lane < nwarp ? sh[lane] : 0.f;
    v = warp_sum(v);
    if (lane == 0) *out = v / (float)B;
  }
}

}  // namespace mlt

extern "C" int mlt_softmax_xent(const void* logits, int is_bf16, const void* targets_i64, void* dlogits, void* loss_out,
                                void* acc3, void* workspace16, int B, int C, int want_acc, void* stream) {
  using namespace mlt;
  MLT_REQUIRE(B > 0 && C > 0, "mlt_softmax_xent: empty problem (B=%d, C=%d)", B, C);
  MLT_REQUIRE(workspace16 != nullptr, "mlt_softmax_xent: workspace is required");
  const int warps = 8;
  int grid = ceil_div(B, warps);
  if (grid > kNumSMs * 4) grid = kNumSMs * 4;
  cudaStream_t st = (cudaStream_t)stream;
  const float inv_B = 1.0f / (float)B;
  if (is_bf16)
    xent_fwd_bwd_kernel<__nv_bfloat16><<<grid, warps * 32, 0, st>>>(
        (const __nv_bfloat16*)logits, (const long long*)targets_i64, (__nv_bfloat16*)dlogits, (float*)loss_out,
        (float*)acc3, (XentWorkspace*)workspace16, B, C, inv_B, want_acc);
  else
    xent_fwd_bwd_kernel<float><<<grid, warps * 32, 0, st>>>(
        (const float*)logits, (const long long*)targets_i64, (float*)dlogits, (float*)loss_out, (float*)acc3,
        (XentWorkspace*)workspace16, B, C, inv_B, want_acc);
  MLT_LAUNCH_CHECK();
  return 0;
}

extern "C" int mlt_argmax_accuracy(const void* logits, int is_bf16, const void* targets_i64, void* out_scalar, int B, int C,
                                   void* stream) {
  using namespace mlt;
  MLT_REQUIRE(B > 0 && C > 0, "mlt_argmax_accuracy: empty problem");
  cudaStream_t st = (cudaStream_t)stream;
  if (is_bf16)
    argmax_accuracy_kernel<__nv_bfloat16><<<1, 1024, 0, st>>>((const __nv_bfloat16*)logits, (const long long*)targets_i64,
                                                             (float*)out_scalar, B, C);
  else
    argmax_accuracy_kernel<float><<<1, 1024, 0, st>>>((const float*)logits, (const long long*)targets_i64,
                                                      (float*)out_scalar, B, C);
  MLT_LAUNCH_CHECK();
  return 0;
}

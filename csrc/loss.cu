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

// =====================================================================================================================
// The reference's other criteria (src/trainer.py:140-150) and its second metric (src/trainer.py:160-163), fused the same way:
//   l1 / l2 / custom   mean |p - t| / mean (p - t)^2 over ALL elements (torch L1Loss / MSELoss, and custom_loss_function
//                      = mean((out - target)^2), src/utils/functions.py:15-17), forward + d(loss)/d(pred) in one pass
//   neg-loss           NLLLoss on log-probabilities: -mean_i logp[i, t_i], gradient -1/B at the target column, + accuracy
//   mcrmse             mean over columns of sqrt(column-wise mean squared error), accumulated next to the loss
// Loss, metric and the step count go into the trainer's 3-float device accumulator: nothing is read back per step.
// Workspace (floats): [0] loss sum, [1] ticket (uint), [2..3] pad, [4 .. 4 + C) column sums of squared error; zero before
// the first launch, the last CTA consumes it and leaves it zero.
// =====================================================================================================================
namespace mlt {

constexpr int kRegMaxCols = 1024;

template <typename P, typename T>
__global__ void __launch_bounds__(256)
regression_loss_kernel(int kind, const P* __restrict__ pred, const T* __restrict__ target, P* __restrict__ dpred,
                       float* __restrict__ loss_out, float* __restrict__ acc3, float* __restrict__ ws, long long n, int C,
                       int want_mcrmse, float inv_n) {
  __shared__ float sh_col[kRegMaxCols];
  __shared__ float sh_part[8];
  __shared__ bool is_last;
  if (want_mcrmse)
    for (int c = threadIdx.x; c < C; c += blockDim.x) sh_col[c] = 0.f;
  __syncthreads();
  float part = 0.f;
  const float gscale = kind == 0 ? inv_n : 2.0f * inv_n;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float d = to_float(pred[i]) - to_float(target[i]);
    const float sq = d * d;
    part += kind == 0 ? fabsf(d) : sq;
    if (dpred != nullptr) dpred[i] = from_float<P>(kind == 0 ? (d > 0.f ? gscale : (d < 0.f ? -gscale : 0.f)) : gscale * d);
    if (want_mcrmse) atomicAdd(&sh_col[(int)(i % C)], sq);
  }
  part = warp_sum(part);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) sh_part[warp] = part;
  __syncthreads();
  if (want_mcrmse)
    for (int c = threadIdx.x; c < C; c += blockDim.x)
      if (sh_col[c] != 0.f) atomicAdd(&ws[4 + c], sh_col[c]);
  if (threadIdx.x == 0) {
    float l = 0.f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) l += sh_part[w];
    atomicAdd(&ws[0], l);
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int t = atomicAdd(reinterpret_cast<unsigned int*>(&ws[1]), 1u);
    is_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  // the last CTA: fold the column sums into the metric, publish, re-zero the workspace
  float m = 0.f;
  if (want_mcrmse) {
    const float inv_rows = (float)C / (float)n;                // rows = n / C
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      m += sqrtf(atomicExch(&ws[4 + c], 0.f) * inv_rows);
    }
    m = warp_sum(m);
    if (lane == 0) sh_part[warp] = m;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float msum = 0.f;
    if (want_mcrmse)
      for (int w = 0; w < (int)(blockDim.x >> 5); ++w) msum += sh_part[w];
    const float loss = atomicExch(&ws[0], 0.f) * inv_n;
    *reinterpret_cast<unsigned int*>(&ws[1]) = 0u;
    if (loss_out) *loss_out = loss;
    if (acc3) {
      acc3[0] += loss;
      if (want_mcrmse) acc3[1] += msum / (float)C;
      acc3[2] += 1.0f;
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
nll_fwd_bwd_kernel(const T* __restrict__ logp, const long long* __restrict__ targets, T* __restrict__ dlogp,
                   float* __restrict__ loss_out, float* __restrict__ acc3, XentWorkspace* __restrict__ ws, int B, int C,
                   float inv_B, int want_acc) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  float loss_part = 0.f, correct_part = 0.f;
  for (int row = blockIdx.x * nwarp + warp; row < B; row += gridDim.x * nwarp) {
    const T* x = logp + (size_t)row * C;
    const int tgt = (int)targets[row];
    float m = -INFINITY;
    int am = 0x7fffffff;
    if (want_acc) {
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
    }
    if (dlogp != nullptr) {
      T* dx = dlogp + (size_t)row * C;
      for (int c = lane; c < C; c += 32) dx[c] = from_float<T>(c == tgt ? -inv_B : 0.f);
    }
    if (lane == 0) {
      if (tgt >= 0 && tgt < C) loss_part -= to_float(x[tgt]);
      correct_part += (am == tgt) ? 1.0f : 0.0f;
    }
  }
  __shared__ float sh_loss[8], sh_corr[8];
  if (lane == 0) { sh_loss[warp] = loss_part; sh_corr[warp] = correct_part; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float l = 0.f, c = 0.f;
    for (int w = 0; w < nwarp; ++w) { l += sh_loss[w]; c += sh_corr[w]; }
    atomicAdd(&ws->loss_sum, l);
    atomicAdd(&ws->correct, c);
    __threadfence();
    const unsigned int t = atomicAdd(&ws->ticket, 1u);
    if (t == gridDim.x - 1) {
      __threadfence();
      const float lsum = atomicExch(&ws->loss_sum, 0.0f);
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

}  // namespace mlt

extern "C" long long mlt_regression_loss_ws_bytes(int C) { return (long long)(4 + (C > 0 ? C : 0)) * 4; }

// kind 0: L1 (mean |p - t|), 1: L2 / MSE (mean (p - t)^2; also the reference's custom_loss_function).  pred / dpred share a dtype
// (fp32 or bf16), target is fp32 or bf16.  n elements, viewed as rows of C for the optional mcrmse metric (C = 0: no metric).
extern "C" int mlt_regression_loss(int kind, const void* pred, int pred_bf16, const void* target, int target_bf16, void* dpred,
                                   void* loss_out, void* acc3, void* workspace, long long ws_bytes, long long n, int C,
                                   int want_mcrmse, void* stream) {
  using namespace mlt;
  MLT_REQUIRE(kind == 0 || kind == 1, "mlt_regression_loss: kind %d (0 = l1, 1 = l2)", kind);
  MLT_REQUIRE(n > 0, "mlt_regression_loss: empty problem");
  MLT_REQUIRE(!want_mcrmse || (C > 0 && C <= kRegMaxCols && n % C == 0), "mlt_regression_loss: mcrmse needs 1 <= C <= %d dividing n (C=%d)",
              kRegMaxCols, C);
  MLT_REQUIRE(workspace != nullptr && ws_bytes >= mlt_regression_loss_ws_bytes(want_mcrmse ? C : 0),
              "mlt_regression_loss: workspace of %lld bytes needed", mlt_regression_loss_ws_bytes(want_mcrmse ? C : 0));
  long long blocks = ceil_div64(n, 256 * 4);
  if (blocks > kNumSMs * 4) blocks = kNumSMs * 4;
  if (blocks < 1) blocks = 1;
  cudaStream_t st = (cudaStream_t)stream;
  const float inv_n = 1.0f / (float)n;
  const int cc = want_mcrmse ? C : 1;
#define MLT_REG_LAUNCH(P, T)                                                                                             \
  regression_loss_kernel<P, T><<<(int)blocks, 256, 0, st>>>(kind, (const P*)pred, (const T*)target, (P*)dpred, (float*)loss_out, \
                                                           (float*)acc3, (float*)workspace, n, cc, want_mcrmse, inv_n)
  if (pred_bf16 && target_bf16) MLT_REG_LAUNCH(__nv_bfloat16, __nv_bfloat16);
  else if (pred_bf16) MLT_REG_LAUNCH(__nv_bfloat16, float);
  else if (target_bf16) MLT_REG_LAUNCH(float, __nv_bfloat16);
  else MLT_REG_LAUNCH(float, float);
#undef MLT_REG_LAUNCH
  MLT_LAUNCH_CHECK();
  return 0;
}

// NLLLoss (mean) on log-probabilities + its gradient + arg-max accuracy; same workspace contract as mlt_softmax_xent.
extern "C" int mlt_nll_loss(const void* logp, int is_bf16, const void* targets_i64, void* dlogp, void* loss_out, void* acc3,
                            void* workspace16, int B, int C, int want_acc, void* stream) {
  using namespace mlt;
  MLT_REQUIRE(B > 0 && C > 0, "mlt_nll_loss: empty problem (B=%d, C=%d)", B, C);
  MLT_REQUIRE(workspace16 != nullptr, "mlt_nll_loss: workspace is required");
  const int warps = 8;
  int grid = ceil_div(B, warps);
  if (grid > kNumSMs * 4) grid = kNumSMs * 4;
  cudaStream_t st = (cudaStream_t)stream;
  const float inv_B = 1.0f / (float)B;
  if (is_bf16)
    nll_fwd_bwd_kernel<__nv_bfloat16><<<grid, warps * 32, 0, st>>>((const __nv_bfloat16*)logp, (const long long*)targets_i64,
                                                                  (__nv_bfloat16*)dlogp, (float*)loss_out, (float*)acc3,
                                                                  (XentWorkspace*)workspace16, B, C, inv_B, want_acc);
  else
    nll_fwd_bwd_kernel<float><<<grid, warps * 32, 0, st>>>((const float*)logp, (const long long*)targets_i64, (float*)dlogp,
                                                          (float*)loss_out, (float*)acc3, (XentWorkspace*)workspace16, B, C, inv_B,
                                                          want_acc);
  MLT_LAUNCH_CHECK();
  return 0;
}

// Training-mode BatchNorm for NHWC bf16 activations viewed as [M = B*H*W, C], with ReLU and the
// residual add fused into the normalisation pass, and the running-statistics update fused into the reduction.
//   forward :  stats (per-channel sum / sum of squares, two-level reduction)  ->  apply (+residual, +ReLU)
//   backward:  reduce (dgamma, dbeta with the ReLU mask applied on the fly)   ->  apply (dx, and the residual gradient)
// Channel reductions use the column-tile scheme of elementwise.cu: a lane owns 8 channels (16 bytes), warps stride
// over rows, CTAs of a channel tile combine through scratch partials + a ticket; the last CTA finalises.
// Bandwidth-bound; roofline = bytes / HBM bandwidth.
#include "common.cuh"
#include "colreduce.cuh"
#include "../include/mlt_b200.h"

namespace mlt {

__device__ __forceinline__ void ld8(const __nv_bfloat16* p, float (&v)[8]) {
  const uint4 raw = *reinterpret_cast<const uint4*>(p);
  const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[j]));
    v[2 * j] = f.x; v[2 * j + 1] = f.y;
  }
}
__device__ __forceinline__ void st8(__nv_bfloat16* p, const float (&v)[8]) {
  uint4 o;
  o.x = pack_bf16(v[0], v[1]); o.y = pack_bf16(v[2], v[3]); o.z = pack_bf16(v[4], v[5]); o.w = pack_bf16(v[6], v[7]);
  *reinterpret_cast<uint4*>(p) = o;
}
template <typename P> __device__ __forceinline__ float ldp(const P* p, int i) { return to_float(p[i]); }

__global__ void __launch_bounds__(256)
bn_stats_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ mean, float* __restrict__ rstd, float* __restrict__ run_mean,
                float* __restrict__ run_var, float momentum, float eps, float* __restrict__ partial, unsigned int* __restrict__ tickets,
                int M, int C) {
  __shared__ float tot[2][kBnTile];
  auto f = [&](int r, int c0, float (&a)[8], float (&b)[8]) {
    float v[8];
    ld8(x + (size_t)r * C + c0, v);
#pragma unroll
    for (int j = 0; j < 8; ++j) { a[j] += v[j]; b[j] = fmaf(v[j], v[j], b[j]); }
  };
  if (!column_reduce2(f, partial, tickets, M, C, tot)) return;
  const int c = blockIdx.x * kBnTile + threadIdx.x;
  if (c < C) {
    const float m = tot[0][threadIdx.x] / (float)M;
    const float var = fmaxf(tot[1][threadIdx.x] / (float)M - m * m, 0.f);
    mean[c] = m;
    rstd[c] = rsqrtf(var + eps);
    if (run_mean != nullptr) {
      run_mean[c] = (1.f - momentum) * run_mean[c] + momentum * m;
      run_var[c] = (1.f - momentum) * run_var[c] + momentum * var * ((float)M / (float)max(M - 1, 1));
    }
  }
}

// y = [relu]( (x - mean) * rstd * gamma + beta [+ residual] )
template <typename P>
__global__ void __launch_bounds__(256)
bn_apply_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ mean, const float* __restrict__ rstd,
                const P* __restrict__ gamma, const P* __restrict__ beta, const __nv_bfloat16* __restrict__ residual,
                __nv_bfloat16* __restrict__ y, long long M, int C, int relu) {
  const int C8 = C / 8;
  const long long total = M * C8;
  const long long step = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += step) {
    const int c0 = (int)(i % C8) * 8;
    float v[8], r[8];
    ld8(x + i * 8, v);
    if (residual != nullptr) ld8(residual + i * 8, r);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float o = (v[j] - mean[c0 + j]) * rstd[c0 + j] * ldp(gamma, c0 + j) + ldp(beta, c0 + j);
      if (residual != nullptr) o += r[j];
      v[j] = relu ? fmaxf(o, 0.f) : o;
    }
    st8(y + i * 8, v);
  }
}

// dgamma[c] = sum dz * xhat, dbeta[c] = sum dz, with dz = dy * (y > 0) when the forward had a ReLU
template <typename G>
__global__ void __launch_bounds__(256)
bn_bwd_reduce_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ x,
                     const float* __restrict__ mean, const float* __restrict__ rstd, G* __restrict__ dgamma, G* __restrict__ dbeta,
                     float* __restrict__ sums, float* __restrict__ partial, unsigned int* __restrict__ tickets, int M, int C, int relu) {
  __shared__ float tot[2][kBnTile];
  auto f = [&](int r, int c0, float (&a)[8], float (&b)[8]) {
    float g[8], v[8], o[8];
    ld8(dy + (size_t)r * C + c0, g);
    ld8(x + (size_t)r * C + c0, v);
    if (relu) ld8(y + (size_t)r * C + c0, o);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float dz = (relu && !(o[j] > 0.f)) ? 0.f : g[j];
      a[j] = fmaf(dz, (v[j] - mean[c0 + j]) * rstd[c0 + j], a[j]);
      b[j] += dz;
    }
  };
  if (!column_reduce2(f, partial, tickets, M, C, tot)) return;
  const int c = blockIdx.x * kBnTile + threadIdx.x;
  if (c < C) {
    dgamma[c] = from_float<G>(tot[0][threadIdx.x]);
    dbeta[c] = from_float<G>(tot[1][threadIdx.x]);
    sums[c] = tot[0][threadIdx.x];            // fp32 copies for the apply pass
    sums[C + c] = tot[1][threadIdx.x];
  }
}

// dx = gamma * rstd * (dz - dbeta/M - xhat * dgamma/M);  dres = dz (gradient of the fused residual input)
template <typename P>
__global__ void __launch_bounds__(256)
bn_bwd_apply_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ x,
                    const float* __restrict__ mean, const float* __restrict__ rstd, const P* __restrict__ gamma,
                    const float* __restrict__ sums, __nv_bfloat16* __restrict__ dx, __nv_bfloat16* __restrict__ dres, long long M,
                    int C, int relu) {
  const int C8 = C / 8;
  const long long total = M * C8;
  const long long step = (long long)gridDim.x * blockDim.x;
  const float inv_M = 1.0f / (float)M;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += step) {
    const int c0 = (int)(i % C8) * 8;
    float g[8], v[8], o[8], d[8];
    ld8(dy + i * 8, g);
    ld8(x + i * 8, v);
    if (relu) ld8(y + i * 8, o);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float dz = (relu && !(o[j] > 0.f)) ? 0.f : g[j];
      const float xh = (v[j] - mean[c0 + j]) * rstd[c0 + j];
      d[j] = ldp(gamma, c0 + j) * rstd[c0 + j] * (dz - sums[C + c0 + j] * inv_M - xh * sums[c0 + j] * inv_M);
      g[j] = dz;
    }
    st8(dx + i * 8, d);
    if (dres != nullptr) st8(dres + i * 8, g);
  }
}

static int bn_rows_split(int M, int C) {
  const int tiles = ceil_div(C, kBnTile);
  int R = (4 * kNumSMs) / tiles;
  if (R > ceil_div(M, 64)) R = ceil_div(M, 64);
  return R < 1 ? 1 : R;
}
static int bn_grid(long long items) {
  long long b = ceil_div64(items, 256);
  const long long cap = (long long)kNumSMs * 16;
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

}  // namespace mlt

using namespace mlt;

// scratch: [tickets 256-byte aligned][partials R*2*C floats][sums 2*C floats]
extern "C" long long mlt_bn_scratch_bytes(int M, int C) {
  return ((long long)ceil_div(C, kBnTile) * 4 + 255) / 256 * 256 + ((long long)bn_rows_split(M, C) * 2 * C + 2 * C) * 4;
}

static void bn_carve(void* scratch, int M, int C, unsigned int** tickets, float** partial, float** sums) {
  *tickets = (unsigned int*)scratch;
  *partial = (float*)((char*)scratch + ((long long)ceil_div(C, kBnTile) * 4 + 255) / 256 * 256);
  *sums = *partial + (size_t)bn_rows_split(M, C) * 2 * C;
}

extern "C" int mlt_bn_stats(const void* x, void* mean, void* rstd, void* run_mean, void* run_var, float momentum, float eps,
                            void* scratch, long long scratch_bytes, int M, int C, void* stream) {
  MLT_REQUIRE(C % 8 == 0 && M > 0, "mlt_bn_stats: channels (%d) must be a multiple of 8", C);
  MLT_REQUIRE(scratch && scratch_bytes >= mlt_bn_scratch_bytes(M, C), "mlt_bn_stats: scratch too small");
  unsigned int* tk; float *pa, *su;
  bn_carve(scratch, M, C, &tk, &pa, &su);
  dim3 grid(ceil_div(C, kBnTile), bn_rows_split(M, C));
  bn_stats_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, (float*)mean, (float*)rstd, (float*)run_mean,
                                                         (float*)run_var, momentum, eps, pa, tk, M, C);
  MLT_LAUNCH_CHECK();
  return 0;
}

extern "C" int mlt_bn_apply(const void* x, const void* mean, const void* rstd, const void* gamma, const void* beta, int params_bf16,
                            const void* residual, void* y, long long M, int C, int relu, void* stream) {
  MLT_REQUIRE(C % 8 == 0, "mlt_bn_apply: channels must be a multiple of 8");
  cudaStream_t st = (cudaStream_t)stream;
  const int g = bn_grid(M * (C / 8));
  if (params_bf16)
    bn_apply_kernel<__nv_bfloat16><<<g, 256, 0, st>>>((const __nv_bfloat16*)x, (const float*)mean, (const float*)rstd,
        (const __nv_bfloat16*)gamma, (const __nv_bfloat16*)beta, (const __nv_bfloat16*)residual, (__nv_bfloat16*)y, M, C, relu);
  else
    bn_apply_kernel<float><<<g, 256, 0, st>>>((const __nv_bfloat16*)x, (const float*)mean, (const float*)rstd, (const float*)gamma,
                                              (const float*)beta, (const __nv_bfloat16*)residual, (__nv_bfloat16*)y, M, C, relu);
  MLT_LAUNCH_CHECK();
  return 0;
}

extern "C" int mlt_bn_bwd(const void* dy, const void* y, const void* x, const void* mean, const void* rstd, const void* gamma,
                          int params_bf16, void* dx, void* dres, void* dgamma, void* dbeta, int grads_bf16, void* scratch,
                          long long scratch_bytes, int M, int C, int relu, void* stream) {
  MLT_REQUIRE(C % 8 == 0 && M > 0, "mlt_bn_bwd: channels must be a multiple of 8");
  MLT_REQUIRE(scratch && scratch_bytes >= mlt_bn_scratch_bytes(M, C), "mlt_bn_bwd: scratch too small");
  unsigned int* tk; float *pa, *su;
  bn_carve(scratch, M, C, &tk, &pa, &su);
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid(ceil_div(C, kBnTile), bn_rows_split(M, C));
  if (grads_bf16)
    bn_bwd_reduce_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)y, (const __nv_bfloat16*)x,
        (const float*)mean, (const float*)rstd, (__nv_bfloat16*)dgamma, (__nv_bfloat16*)dbeta, su, pa, tk, M, C, relu);
  else
    bn_bwd_reduce_kernel<float><<<grid, 256, 0, st>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)y, (const __nv_bfloat16*)x,
        (const float*)mean, (const float*)rstd, (float*)dgamma, (float*)dbeta, su, pa, tk, M, C, relu);
  MLT_LAUNCH_CHECK();
  const int g = bn_grid((long long)M * (C / 8));
  if (params_bf16)
    bn_bwd_apply_kernel<__nv_bfloat16><<<g, 256, 0, st>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)y, (const __nv_bfloat16*)x,
        (const float*)mean, (const float*)rstd, (const __nv_bfloat16*)gamma, su, (__nv_bfloat16*)dx, (__nv_bfloat16*)dres, M, C, relu);
  else
    bn_bwd_apply_kernel<float><<<g, 256, 0, st>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)y, (const __nv_bfloat16*)x,
        (const float*)mean, (const float*)rstd, (const float*)gamma, su, (__nv_bfloat16*)dx, (__nv_bfloat16*)dres, M, C, relu);
  MLT_LAUNCH_CHECK();
  return 0;
}

// RMSNorm forward / backward with the residual add fused in (the north star names LayerNorm/RMSNorm).
// The reference has no normalisation layer of its own (src/model.py:7-24 is a LeNet); RMSNorm is the norm of the
// "bring your own nn.Module" transformer a user plugs in (reference main.py:30), next to LayerNorm (norm.cu).
//
//   y = h * rstd * gamma,   h = x [+ res],   rstd = rsqrt(mean(h^2) + eps)
//   dx = rstd * (g - xh * mean(g * xh)),   g = dy * gamma,   xh = h * rstd;   dgamma = sum_rows dy * xh
//
// bf16 activations, fp32 statistics.  A row is handled by TPR threads (32: one warp, widths <= 1024; 256: one CTA, widths <=
// 8192), each thread owning the same <= 4 16-byte vectors of every row it visits, so the backward pass keeps its dgamma
// partials in registers across rows (grid sized to the machine), folds them per CTA in shared memory and per grid with fp32
// atomics; the last CTA (ticket) writes dgamma and re-zeroes the accumulator.  One read of dy and h, one write of dx: the
// roofline is bytes / HBM bandwidth.
#include "common.cuh"
#include "../include/mlt_b200.h"

namespace mlt {
namespace rms {

constexpr int kThreads = 256;
constexpr int kMaxNV = 4;

__device__ __forceinline__ void unpack8(const uint4& raw, float (&v)[8]) {
  const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[j]));
    v[2 * j] = f.x; v[2 * j + 1] = f.y;
  }
}
__device__ __forceinline__ uint4 pack8(const float (&v)[8]) {
  uint4 o;
  o.x = pack_bf16(v[0], v[1]); o.y = pack_bf16(v[2], v[3]); o.z = pack_bf16(v[4], v[5]); o.w = pack_bf16(v[6], v[7]);
  return o;
}
template <typename P> __device__ __forceinline__ void load_param8(const P* p, int i, float (&v)[8]);
template <> __device__ __forceinline__ void load_param8<float>(const float* p, int i, float (&v)[8]) {
  const float4 a = *reinterpret_cast<const float4*>(p + i), b = *reinterpret_cast<const float4*>(p + i + 4);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
template <> __device__ __forceinline__ void load_param8<__nv_bfloat16>(const __nv_bfloat16* p, int i, float (&v)[8]) {
  unpack8(*reinterpret_cast<const uint4*>(p + i), v);
}

// Sum over the TPR threads that share a row.  TPR == 256: `slot` is this CTA's 8-float exchange buffer; two barriers.
template <int TPR>
__device__ __forceinline__ float row_sum(float v, float* slot) {
  v = warp_sum(v);
  if (TPR == 32) return v;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();                       // the previous use of the buffer has been read by everyone
  if (lane == 0) slot[warp] = v;
  __syncthreads();
  float t = 0.f;
#pragma unroll
  for (int w = 0; w < kThreads / 32; ++w) t += slot[w];
  return t;
}

template <typename P, int TPR, int NV>
__global__ void __launch_bounds__(kThreads)
rmsnorm_fwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ res, const P* __restrict__ gamma,
                   __nv_bfloat16* __restrict__ y, __nv_bfloat16* __restrict__ hsum, float* __restrict__ rstd_out, int T, int H, float eps) {
  __shared__ float slot[8];
  constexpr int RPB = kThreads / TPR;                      // rows per CTA
  const int t = threadIdx.x % TPR;
  const int row = blockIdx.x * RPB + threadIdx.x / TPR;
  const bool live = row < T;                               // TPR == 256: RPB == 1 and the grid is exact, so the CTA is uniform
  if (TPR == 32 && !live) return;
  const int nvec = H / 8;
  float v[NV][8];
  float sq = 0.f;
#pragma unroll
  for (int c = 0; c < NV; ++c) {
    const int iv = t + TPR * c;
    if (iv < nvec) {
      unpack8(*reinterpret_cast<const uint4*>(x + (size_t)row * H + iv * 8), v[c]);
      if (res != nullptr) {
        float r[8];
        unpack8(*reinterpret_cast<const uint4*>(res + (size_t)row * H + iv * 8), r);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[c][j] += r[j];
        if (hsum != nullptr) {                             // round once so forward and backward see the same pre-norm values
          const uint4 pk = pack8(v[c]);
          *reinterpret_cast<uint4*>(hsum + (size_t)row * H + iv * 8) = pk;
          unpack8(pk, v[c]);
        }
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) sq = fmaf(v[c][j], v[c][j], sq);
    }
  }
  const float rstd = rsqrtf(row_sum<TPR>(sq, slot) / (float)H + eps);
#pragma unroll
  for (int c = 0; c < NV; ++c) {
    const int iv = t + TPR * c;
    if (iv < nvec) {
      float g[8], o[8];
      load_param8<P>(gamma, iv * 8, g);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = v[c][j] * rstd * g[j];
      *reinterpret_cast<uint4*>(y + (size_t)row * H + iv * 8) = pack8(o);
    }
  }
  if (t == 0) rstd_out[row] = rstd;
}

template <typename P, typename GO, int TPR, int NV>
__global__ void __launch_bounds__(kThreads, 2)
rmsnorm_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ h, const float* __restrict__ rstd,
                   const P* __restrict__ gamma, __nv_bfloat16* __restrict__ dx, GO* __restrict__ dgamma, float* __restrict__ acc,
                   unsigned int* __restrict__ ticket, int T, int H) {
  extern __shared__ float fold[];                          // [H]
  __shared__ float slot[8];
  __shared__ bool is_last;
  constexpr int RPB = kThreads / TPR;
  const int t = threadIdx.x % TPR;
  const int nvec = H / 8;
  for (int i = threadIdx.x; i < H; i += blockDim.x) fold[i] = 0.f;
  __syncthreads();
  float ag[NV][8];
#pragma unroll
  for (int c = 0; c < NV; ++c)
#pragma unroll
    for (int j = 0; j < 8; ++j) ag[c][j] = 0.f;
  // TPR == 256: every thread of the CTA walks the same rows (uniform trip count, the barriers inside row_sum are safe)
  for (int row = blockIdx.x * RPB + threadIdx.x / TPR; row < T; row += gridDim.x * RPB) {
    const float rs = rstd[row];
    float gd[NV][8], xh[NV][8];
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < NV; ++c) {
      const int iv = t + TPR * c;
      if (iv < nvec) {
        float gm[8];
        unpack8(*reinterpret_cast<const uint4*>(dy + (size_t)row * H + iv * 8), gd[c]);
        unpack8(*reinterpret_cast<const uint4*>(h + (size_t)row * H + iv * 8), xh[c]);
        load_param8<P>(gamma, iv * 8, gm);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          xh[c][j] *= rs;
          ag[c][j] = fmaf(gd[c][j], xh[c][j], ag[c][j]);
          gd[c][j] *= gm[j];
          s = fmaf(gd[c][j], xh[c][j], s);
        }
      }
    }
    s = row_sum<TPR>(s, slot) / (float)H;
#pragma unroll
    for (int c = 0; c < NV; ++c) {
      const int iv = t + TPR * c;
      if (iv < nvec) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = rs * (gd[c][j] - xh[c][j] * s);
        *reinterpret_cast<uint4*>(dx + (size_t)row * H + iv * 8) = pack8(o);
      }
    }
  }
#pragma unroll
  for (int c = 0; c < NV; ++c) {
    const int iv = t + TPR * c;
    if (iv < nvec) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (TPR == 32) atomicAdd(&fold[iv * 8 + j], ag[c][j]);      // 8 warps share the columns
        else fold[iv * 8 + j] = ag[c][j];                           // one owner per column
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < H; i += blockDim.x) atomicAdd(acc + i, fold[i]);
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int tk = atomicAdd(ticket, 1u);
    is_last = tk == gridDim.x - 1;
    if (is_last) *ticket = 0u;
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  for (int i = threadIdx.x; i < H; i += blockDim.x) {
    dgamma[i] = from_float<GO>(__ldcg(acc + i));
    acc[i] = 0.f;
  }
}

constexpr int kMaxWidth = 8 * kThreads * kMaxNV;          // 8192

}  // namespace rms
}  // namespace mlt

using namespace mlt;

// y = RMSNorm(x [+ res]) * gamma; x / res / y / hsum bf16 [T, H], gamma fp32 or bf16 [H], rstd fp32 [T].
// hsum (optional, with res): the rounded pre-norm sum h = x + res that backward (and the next residual branch) reads.
extern "C" int mlt_rmsnorm_fwd(const void* x, const void* res, const void* gamma, int params_bf16, void* y, void* hsum, void* rstd,
                               int T, int H, float eps, void* stream) {
  MLT_REQUIRE(H % 8 == 0 && H >= 8 && H <= rms::kMaxWidth, "mlt_rmsnorm_fwd: width %d unsupported (multiple of 8, <= %d)", H, rms::kMaxWidth);
  MLT_REQUIRE(T > 0, "mlt_rmsnorm_fwd: empty input");
  cudaStream_t st = (cudaStream_t)stream;
  const bool wide = H > 1024;
  const int nv = ceil_div(H / 8, wide ? 256 : 32);
  const int grid = wide ? T : ceil_div(T, 8);
#define RMS_FWD(P, TPR, NV)                                                                                                    \
  rms::rmsnorm_fwd_kernel<P, TPR, NV><<<grid, rms::kThreads, 0, st>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)res,          \
      (const P*)gamma, (__nv_bfloat16*)y, (__nv_bfloat16*)hsum, (float*)rstd, T, H, eps)
#define RMS_FWD_NV(P, TPR)                                                                                                      \
  do { if (nv == 1) RMS_FWD(P, TPR, 1); else if (nv == 2) RMS_FWD(P, TPR, 2); else if (nv == 3) RMS_FWD(P, TPR, 3); else RMS_FWD(P, TPR, 4); } while (0)
  if (params_bf16) { if (wide) RMS_FWD_NV(__nv_bfloat16, 256); else RMS_FWD_NV(__nv_bfloat16, 32); }
  else { if (wide) RMS_FWD_NV(float, 256); else RMS_FWD_NV(float, 32); }
#undef RMS_FWD_NV
#undef RMS_FWD
  MLT_LAUNCH_CHECK();
  return 0;
}

// scratch (zeroed once, left zero by every call): [ticket 256 B][H fp32 accumulators, sized for the widest row]
extern "C" long long mlt_rmsnorm_bwd_scratch_bytes(void) { return 256 + (long long)rms::kMaxWidth * 4; }

extern "C" int mlt_rmsnorm_bwd(const void* dy, const void* h, const void* rstd, const void* gamma, int params_bf16, void* dx,
                               void* dgamma, int grads_bf16, void* scratch, long long scratch_bytes, int T, int H, void* stream) {
  MLT_REQUIRE(H % 8 == 0 && H >= 8 && H <= rms::kMaxWidth, "mlt_rmsnorm_bwd: width %d unsupported", H);
  MLT_REQUIRE(T > 0, "mlt_rmsnorm_bwd: empty input");
  MLT_REQUIRE(scratch && scratch_bytes >= mlt_rmsnorm_bwd_scratch_bytes(), "mlt_rmsnorm_bwd: scratch too small");
  cudaStream_t st = (cudaStream_t)stream;
  unsigned int* ticket = (unsigned int*)scratch;
  float* acc = (float*)((char*)scratch + 256);
  const bool wide = H > 1024;
  const int nv = ceil_div(H / 8, wide ? 256 : 32);
  int grid = 2 * kNumSMs;
  const int row_groups = wide ? T : ceil_div(T, 8);
  if (grid > row_groups) grid = row_groups;
  const size_t smem = (size_t)H * sizeof(float);
#define RMS_BWD(P, GO, TPR, NV)                                                                                                 \
  rms::rmsnorm_bwd_kernel<P, GO, TPR, NV><<<grid, rms::kThreads, smem, st>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)h,    \
      (const float*)rstd, (const P*)gamma, (__nv_bfloat16*)dx, (GO*)dgamma, acc, ticket, T, H)
#define RMS_BWD_NV(P, GO, TPR)                                                                                                  \
  do { if (nv == 1) RMS_BWD(P, GO, TPR, 1); else if (nv == 2) RMS_BWD(P, GO, TPR, 2); else if (nv == 3) RMS_BWD(P, GO, TPR, 3);  \
       else RMS_BWD(P, GO, TPR, 4); } while (0)
#define RMS_BWD_T(P, GO) do { if (wide) RMS_BWD_NV(P, GO, 256); else RMS_BWD_NV(P, GO, 32); } while (0)
  if (params_bf16) { if (grads_bf16) RMS_BWD_T(__nv_bfloat16, __nv_bfloat16); else RMS_BWD_T(__nv_bfloat16, float); }
  else { if (grads_bf16) RMS_BWD_T(float, __nv_bfloat16); else RMS_BWD_T(float, float); }
#undef RMS_BWD_T
#undef RMS_BWD_NV
#undef RMS_BWD
  MLT_LAUNCH_CHECK();
  return 0;
}

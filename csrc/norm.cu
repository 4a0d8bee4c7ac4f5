// LayerNorm forward / backward with the residual add fused in, and row softmax forward / backward
// for attention scores.  One warp per row, 16-byte vector accesses, statistics in fp32.  Bandwidth-bound: the
// roofline is bytes moved / HBM bandwidth.  Replaces ATen's layer_norm / add / softmax kernels of the BERT path.
#include "common.cuh"
#include "colreduce.cuh"
#include "../include/mlt_b200.h"

namespace mlt {

constexpr int kMaxVecPerLane = 4;      // 4 vectors x 8 bf16 x 32 lanes -> rows up to 1024 wide

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
// parameters (gamma / beta) may be fp32 or bf16
template <typename P> __device__ __forceinline__ void load_param8(const P* p, int i, float (&v)[8]);
template <> __device__ __forceinline__ void load_param8<float>(const float* p, int i, float (&v)[8]) {
  const float4 a = *reinterpret_cast<const float4*>(p + i), b = *reinterpret_cast<const float4*>(p + i + 4);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
template <> __device__ __forceinline__ void load_param8<__nv_bfloat16>(const __nv_bfloat16* p, int i, float (&v)[8]) {
  unpack8(*reinterpret_cast<const uint4*>(p + i), v);
}

// y = LN(x [+ res]) * gamma + beta.  Writes y, optionally the pre-norm sum h = x + res, and mean / rstd per row.
template <typename P>
__global__ void __launch_bounds__(256, 3)
layernorm_fwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ res, const P* __restrict__ gamma,
                     const P* __restrict__ beta, __nv_bfloat16* __restrict__ y, __nv_bfloat16* __restrict__ hsum,
                     float* __restrict__ mean_out, float* __restrict__ rstd_out, int T, int H, float eps) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= T) return;
  const int nvec = H / 8;
  float v[kMaxVecPerLane][8];
  float sum = 0.f;
  // every load of the row first (a warp owns one row: its only latency hiding is the number of loads it has in flight; with the
  // loads sunk next to their consumers a lane waited for memory four times per row), then the arithmetic
  uint4 rx[kMaxVecPerLane], rr[kMaxVecPerLane];
#pragma unroll
  for (int c = 0; c < kMaxVecPerLane; ++c) {
    const int iv = lane + 32 * c;
    rx[c] = make_uint4(0u, 0u, 0u, 0u);
    rr[c] = make_uint4(0u, 0u, 0u, 0u);
    if (iv < nvec) {
      rx[c] = *reinterpret_cast<const uint4*>(x + (size_t)row * H + iv * 8);
      if (res != nullptr) rr[c] = *reinterpret_cast<const uint4*>(res + (size_t)row * H + iv * 8);
    }
  }
#pragma unroll
  for (int c = 0; c < kMaxVecPerLane; ++c) {
    const int iv = lane + 32 * c;
    if (iv < nvec) {
      unpack8(rx[c], v[c]);
      if (res != nullptr) {
        float r[8];
        unpack8(rr[c], r);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[c][j] += r[j];
        if (hsum != nullptr) {
          // round once to bf16 so forward and backward see the same pre-norm values
          const uint4 pk = pack8(v[c]);
          *reinterpret_cast<uint4*>(hsum + (size_t)row * H + iv * 8) = pk;
          unpack8(pk, v[c]);
        }
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) sum += v[c][j];
    }
  }
  const float mean = warp_sum(sum) / (float)H;
  float sq = 0.f;
#pragma unroll
  for (int c = 0; c < kMaxVecPerLane; ++c)
    if (lane + 32 * c < nvec)
#pragma unroll
      for (int j = 0; j < 8; ++j) { const float d = v[c][j] - mean; sq += d * d; }
  const float rstd = rsqrtf(warp_sum(sq) / (float)H + eps);
#pragma unroll
  for (int c = 0; c < kMaxVecPerLane; ++c) {
    const int iv = lane + 32 * c;
    if (iv < nvec) {
      float g[8], b[8], o[8];
      load_param8<P>(gamma, iv * 8, g);
      load_param8<P>(beta, iv * 8, b);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = (v[c][j] - mean) * rstd * g[j] + b[j];
      *reinterpret_cast<uint4*>(y + (size_t)row * H + iv * 8) = pack8(o);
    }
  }
  if (lane == 0) { mean_out[row] = mean; rstd_out[row] = rstd; }
}

// dx = rstd * (g*dy - mean(g*dy) - xhat * mean(g*dy*xhat)).  One warp per row, nothing carried between rows, so the
// register footprint stays small and many warps are resident (this pass is pure bandwidth).
template <typename P>
__global__ void __launch_bounds__(256)
layernorm_bwd_dx_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ h, const float* __restrict__ mean,
                        const float* __restrict__ rstd, const P* __restrict__ gamma, __nv_bfloat16* __restrict__ dx, int T, int H) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= T) return;
  const int nvec = H / 8;
  const float mu = mean[row], rs = rstd[row];
  float gd[kMaxVecPerLane][8], xh[kMaxVecPerLane][8];
  float s1 = 0.f, s2 = 0.f;
#pragma unroll
  for (int c = 0; c < kMaxVecPerLane; ++c) {
    const int iv = lane + 32 * c;
    if (iv < nvec) {
      float g[8];
      unpack8(*reinterpret_cast<const uint4*>(dy + (size_t)row * H + iv * 8), gd[c]);
      unpack8(*reinterpret_cast<const uint4*>(h + (size_t)row * H + iv * 8), xh[c]);
      load_param8<P>(gamma, iv * 8, g);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        xh[c][j] = (xh[c][j] - mu) * rs;
        gd[c][j] *= g[j];
        s1 += gd[c][j];
        s2 = fmaf(gd[c][j], xh[c][j], s2);
      }
    }
  }
  s1 = warp_sum(s1) / (float)H;
  s2 = warp_sum(s2) / (float)H;
#pragma unroll
  for (int c = 0; c < kMaxVecPerLane; ++c) {
    const int iv = lane + 32 * c;
    if (iv < nvec) {
      float o[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = rs * (gd[c][j] - s1 - xh[c][j] * s2);
      *reinterpret_cast<uint4*>(dx + (size_t)row * H + iv * 8) = pack8(o);
    }
  }
}

// One-pass LayerNorm backward: dx as above AND the parameter gradients.  Warps loop over rows (grid sized to the machine,
// two CTAs per SM); a lane always owns the same columns, so dgamma / dbeta partials stay in registers across rows, are folded
// per CTA with shared-memory atomics and per grid with fp32 atomics on a scratch accumulator (zero at rest); the last CTA
// (ticket) writes dgamma / dbeta and re-zeroes the accumulator.  dy and h are read once instead of twice.
template <typename P, typename GO, int NV>
__global__ void __launch_bounds__(256, 2)
layernorm_bwd_fused_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ h, const float* __restrict__ mean,
                           const float* __restrict__ rstd, const P* __restrict__ gamma, __nv_bfloat16* __restrict__ dx,
                           GO* __restrict__ dgamma, GO* __restrict__ dbeta, float* __restrict__ acc, unsigned int* __restrict__ ticket,
                           int T, int H) {
  extern __shared__ float fold[];                       // [2][H]
  __shared__ bool is_last;
  const int lane = threadIdx.x & 31;
  const int nvec = H / 8;
  for (int i = threadIdx.x; i < 2 * H; i += blockDim.x) fold[i] = 0.f;
  __syncthreads();
  float ag[NV][8], ab[NV][8];
#pragma unroll
  for (int c = 0; c < NV; ++c) {
#pragma unroll
    for (int j = 0; j < 8; ++j) { ag[c][j] = 0.f; ab[c][j] = 0.f; }
  }
  const int warps = gridDim.x * (blockDim.x >> 5);
  for (int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); row < T; row += warps) {
    const float mu = mean[row], rs = rstd[row];
    float gd[NV][8], xh[NV][8];
    float s1 = 0.f, s2 = 0.f;
    uint4 rg[NV], rh[NV];                               // the row's loads go out together, then the arithmetic
#pragma unroll
    for (int c = 0; c < NV; ++c) {
      const int iv = lane + 32 * c;
      rg[c] = make_uint4(0u, 0u, 0u, 0u);
      rh[c] = make_uint4(0u, 0u, 0u, 0u);
      if (iv < nvec) {
        rg[c] = *reinterpret_cast<const uint4*>(dy + (size_t)row * H + iv * 8);
        rh[c] = *reinterpret_cast<const uint4*>(h + (size_t)row * H + iv * 8);
      }
    }
#pragma unroll
    for (int c = 0; c < NV; ++c) {
      const int iv = lane + 32 * c;
      if (iv < nvec) {
        float gm[8];
        unpack8(rg[c], gd[c]);
        unpack8(rh[c], xh[c]);
        load_param8<P>(gamma, iv * 8, gm);               // L1-resident; keeping it in registers costs occupancy
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          xh[c][j] = (xh[c][j] - mu) * rs;
          ag[c][j] = fmaf(gd[c][j], xh[c][j], ag[c][j]);
          ab[c][j] += gd[c][j];
          gd[c][j] *= gm[j];
          s1 += gd[c][j];
          s2 = fmaf(gd[c][j], xh[c][j], s2);
        }
      }
    }
    s1 = warp_sum(s1) / (float)H;
    s2 = warp_sum(s2) / (float)H;
#pragma unroll
    for (int c = 0; c < NV; ++c) {
      const int iv = lane + 32 * c;
      if (iv < nvec) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = rs * (gd[c][j] - s1 - xh[c][j] * s2);
        *reinterpret_cast<uint4*>(dx + (size_t)row * H + iv * 8) = pack8(o);
      }
    }
  }
#pragma unroll
  for (int c = 0; c < NV; ++c) {
    const int iv = lane + 32 * c;
    if (iv < nvec) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        atomicAdd(&fold[iv * 8 + j], ag[c][j]);
        atomicAdd(&fold[H + iv * 8 + j], ab[c][j]);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 2 * H; i += blockDim.x) atomicAdd(acc + i, fold[i]);
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
    dbeta[i] = from_float<GO>(__ldcg(acc + H + i));
    acc[i] = 0.f;
    acc[H + i] = 0.f;
  }
}

// dgamma[c] = sum_rows dy * xhat;  dbeta[c] = sum_rows dy   (column reduction, see colreduce.cuh)
template <typename GO>
__global__ void __launch_bounds__(256)
layernorm_bwd_params_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ h, const float* __restrict__ mean,
                            const float* __restrict__ rstd, GO* __restrict__ dgamma, GO* __restrict__ dbeta, float* __restrict__ partial,
                            unsigned int* __restrict__ tickets, int T, int H) {
  __shared__ float tot[2][kBnTile];
  auto f = [&](int r, int c0, float (&a)[8], float (&b)[8]) {
    float d[8], x[8];
    unpack8(*reinterpret_cast<const uint4*>(dy + (size_t)r * H + c0), d);
    unpack8(*reinterpret_cast<const uint4*>(h + (size_t)r * H + c0), x);
    const float mu = mean[r], rs = rstd[r];
#pragma unroll
    for (int j = 0; j < 8; ++j) { a[j] = fmaf(d[j], (x[j] - mu) * rs, a[j]); b[j] += d[j]; }
  };
  if (!column_reduce2(f, partial, tickets, T, H, tot)) return;
  const int c = blockIdx.x * kBnTile + threadIdx.x;
  if (c < H) {
    dgamma[c] = from_float<GO>(tot[0][threadIdx.x]);
    dbeta[c] = from_float<GO>(tot[1][threadIdx.x]);
  }
}

// In-place row softmax of bf16 scores: p = softmax(scale * s).  One warp per row, L <= 1024, L % 8 == 0.
__global__ void __launch_bounds__(256)
softmax_rows_fwd_kernel(__nv_bfloat16* __restrict__ s, long long rows, int L, float scale) {
  const int lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int nvec = L / 8;
  float v[kMaxVecPerLane][8];
  float m = -INFINITY;
#pragma unroll
  for (int c = 0; c < kMaxVecPerLane; ++c) {
    const int iv = lane + 32 * c;
    if (iv < nvec) {
      unpack8(*reinterpret_cast<const uint4*>(s + row * L + iv * 8), v[c]);
#pragma unroll
      for (int j = 0; j < 8; ++j) { v[c][j] *= scale; m = fmaxf(m, v[c][j]); }
    }
  }
  m = warp_max(m);
  float sum = 0.f;
#pragma unroll
  for (int c = 0; c < kMaxVecPerLane; ++c)
    if (lane + 32 * c < nvec)
#pragma unroll
      for (int j = 0; j < 8; ++j) { v[c][j] = __expf(v[c][j] - m); sum += v[c][j]; }
  const float inv = 1.0f / warp_sum(sum);
#pragma unroll
  for (int c = 0; c < kMaxVecPerLane; ++c) {
    const int iv = lane + 32 * c;
    if (iv < nvec) {
#pragma unroll
      for (int j = 0; j < 8; ++j) v[c][j] *= inv;
      *reinterpret_cast<uint4*>(s + row * L + iv * 8) = pack8(v[c]);
    }
  }
}

// In place on dp: ds = scale * p * (dp - sum_j dp_j p_j)
__global__ void __launch_bounds__(256)
softmax_rows_bwd_kernel(__nv_bfloat16* __restrict__ dp, const __nv_bfloat16* __restrict__ p, long long rows, int L, float scale) {
  const int lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int nvec = L / 8;
  float d[kMaxVecPerLane][8], q[kMaxVecPerLane][8];
  float dot = 0.f;
#pragma unroll
  for (int c = 0; c < kMaxVecPerLane; ++c) {
    const int iv = lane + 32 * c;
    if (iv < nvec) {
      unpack8(*reinterpret_cast<const uint4*>(dp + row * L + iv * 8), d[c]);
      unpack8(*reinterpret_cast<const uint4*>(p + row * L + iv * 8), q[c]);
#pragma unroll
      for (int j = 0; j < 8; ++j) dot += d[c][j] * q[c][j];
    }
  }
  dot = warp_sum(dot);
#pragma unroll
  for (int c = 0; c < kMaxVecPerLane; ++c) {
    const int iv = lane + 32 * c;
    if (iv < nvec) {
      float o[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = scale * q[c][j] * (d[c][j] - dot);
      *reinterpret_cast<uint4*>(dp + row * L + iv * 8) = pack8(o);
    }
  }
}

static int ln_rows_split(int T, int H) {
  const int tiles = ceil_div(H, kBnTile);
  int R = (4 * kNumSMs) / tiles;
  if (R > ceil_div(T, 64)) R = ceil_div(T, 64);
  return R < 1 ? 1 : R;
}

}  // namespace mlt

using namespace mlt;

extern "C" int mlt_layernorm_fwd(const void* x, const void* res, const void* gamma, const void* beta, int params_bf16, void* y,
                                 void* hsum, void* mean, void* rstd, int T, int H, float eps, void* stream) {
  MLT_REQUIRE(H % 8 == 0 && H <= 8 * 32 * kMaxVecPerLane, "mlt_layernorm_fwd: width %d unsupported (multiple of 8, <= 1024)", H);
  MLT_REQUIRE(T > 0, "mlt_layernorm_fwd: empty input");
  const int grid = ceil_div(T, 8);
  cudaStream_t st = (cudaStream_t)stream;
  if (params_bf16)
    layernorm_fwd_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)res,
        (const __nv_bfloat16*)gamma, (const __nv_bfloat16*)beta, (__nv_bfloat16*)y, (__nv_bfloat16*)hsum, (float*)mean, (float*)rstd, T, H, eps);
  else
    layernorm_fwd_kernel<float><<<grid, 256, 0, st>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)res, (const float*)gamma,
        (const float*)beta, (__nv_bfloat16*)y, (__nv_bfloat16*)hsum, (float*)mean, (float*)rstd, T, H, eps);
  MLT_LAUNCH_CHECK();
  return 0;
}

// scratch (zeroed once, left zero by every call): [ticket 256 B][2*H fp32 accumulators]
extern "C" long long mlt_layernorm_bwd_scratch_bytes(int T, int H) {
  (void)T;
  return 256 + 2LL * 1024 * 4 + 0 * H;
}

extern "C" int mlt_layernorm_bwd(const void* dy, const void* h, const void* mean, const void* rstd, const void* gamma,
                                 int params_bf16, void* dx, void* dgamma, void* dbeta, int grads_bf16, void* scratch,
                                 long long scratch_bytes, int T, int H, void* stream) {
  MLT_REQUIRE(H % 8 == 0 && H <= 8 * 32 * kMaxVecPerLane, "mlt_layernorm_bwd: width %d unsupported", H);
  MLT_REQUIRE(scratch && scratch_bytes >= mlt_layernorm_bwd_scratch_bytes(T, H), "mlt_layernorm_bwd: scratch too small");
  cudaStream_t st = (cudaStream_t)stream;
  unsigned int* ticket = (unsigned int*)scratch;
  float* acc = (float*)((char*)scratch + 256);
  int grid = 2 * kNumSMs;
  if (grid > ceil_div(T, 8)) grid = ceil_div(T, 8);
  const size_t smem = (size_t)2 * H * sizeof(float);
  const int nv = ceil_div(H / 8, 32);                       // 16-byte vectors per lane: sizes the register arrays
#define LN_BWD_NV(P, GO, NV)                                                                                                   \
  layernorm_bwd_fused_kernel<P, GO, NV><<<grid, 256, smem, st>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)h,               \
      (const float*)mean, (const float*)rstd, (const P*)gamma, (__nv_bfloat16*)dx, (GO*)dgamma, (GO*)dbeta, acc, ticket, T, H)
#define LN_BWD(P, GO)                                                                                                          \
  do {                                                                                                                         \
    if (nv == 1) LN_BWD_NV(P, GO, 1); else if (nv == 2) LN_BWD_NV(P, GO, 2); else if (nv == 3) LN_BWD_NV(P, GO, 3);            \
    else LN_BWD_NV(P, GO, 4);                                                                                                  \
  } while (0)
  if (params_bf16) { if (grads_bf16) LN_BWD(__nv_bfloat16, __nv_bfloat16); else LN_BWD(__nv_bfloat16, float); }
  else { if (grads_bf16) LN_BWD(float, __nv_bfloat16); else LN_BWD(float, float); }
#undef LN_BWD
#undef LN_BWD_NV
  MLT_LAUNCH_CHECK();
  return 0;
}

extern "C" int mlt_softmax_rows(void* scores_bf16, long long rows, int L, float scale, void* stream) {
  MLT_REQUIRE(L % 8 == 0 && L <= 8 * 32 * kMaxVecPerLane, "mlt_softmax_rows: row length %d unsupported (multiple of 8, <= 1024)", L);
  if (rows <= 0) return 0;
  softmax_rows_fwd_kernel<<<(unsigned)ceil_div64(rows, 8), 256, 0, (cudaStream_t)stream>>>((__nv_bfloat16*)scores_bf16, rows, L, scale);
  MLT_LAUNCH_CHECK();
  return 0;
}

extern "C" int mlt_softmax_rows_bwd(void* dp_bf16, const void* p_bf16, long long rows, int L, float scale, void* stream) {
  MLT_REQUIRE(L % 8 == 0 && L <= 8 * 32 * kMaxVecPerLane, "mlt_softmax_rows_bwd: row length %d unsupported", L);
  if (rows <= 0) return 0;
  softmax_rows_bwd_kernel<<<(unsigned)ceil_div64(rows, 8), 256, 0, (cudaStream_t)stream>>>((__nv_bfloat16*)dp_bf16,
                                                                                        (const __nv_bfloat16*)p_bf16, rows, L, scale);
  MLT_LAUNCH_CHECK();
  return 0;
}

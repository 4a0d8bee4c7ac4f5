// Convolution support for the ResNet path (SURVEY §8 K7 conv, K10): NHWC bf16 im2col / col2im around the tcgen05 GEMM,
// 3x3/2 max-pool and global average pool, NCHW fp32 -> NHWC bf16 input conversion.
// A convolution is  out[B*OH*OW, Cout] = col[B*OH*OW, KH*KW*C] * W[Cout, KH*KW*C]^T  with the GEMM on tensor cores;
// 1x1 stride-1 convolutions need no col matrix at all (the activation tensor IS the A operand).
// All kernels move 16 bytes per thread (8 channels); the roofline is bytes / HBM bandwidth.
#include "common.cuh"
#include "../include/mlt_b200.h"

namespace mlt {

// col[(b,oh,ow)][(kh,kw,c)] = x[b][oh*s-p+kh][ow*s-p+kw][c]   (zero outside the image)
__global__ void __launch_bounds__(256)
im2col_nhwc_kernel(const uint4* __restrict__ x, uint4* __restrict__ col, int B, int H, int W, int C8, int KH, int KW, int stride,
                   int pad, int OH, int OW) {
  const long long total = (long long)B * OH * OW * KH * KW * C8;
  const long long step = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += step) {
    const int c = (int)(i % C8);
    long long r = i / C8;
    const int kw = (int)(r % KW); r /= KW;
    const int kh = (int)(r % KH); r /= KH;
    const int ow = (int)(r % OW); r /= OW;
    const int oh = (int)(r % OH);
    const int b = (int)(r / OH);
    const int h = oh * stride - pad + kh, w = ow * stride - pad + kw;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (h >= 0 && h < H && w >= 0 && w < W) v = x[(((size_t)b * H + h) * W + w) * C8 + c];
    col[i] = v;
  }
}

// The 8-channel (padded 3-channel) stem with a compile-time filter: consecutive threads copy consecutive 16-byte chunks of the
// patch matrix (coalesced stores, and loads that walk consecutive pixels of a filter row), four chunks per thread and iteration
// with the four loads issued first; the tap index comes from divisions by constants and everything is 32-bit.  (The generic
// kernel above spends five 64-bit divisions and one dependent load per 16 bytes: 1.1 ms for ResNet-50's 7x7/2 stem at batch 256,
// whose 2.5 GB patch matrix needs 0.4 ms to write.  A thread-per-filter-row variant was slower still, 1.27 ms: its stores are
// 112 bytes apart across a warp.)
template <int KH, int KW>
__global__ void __launch_bounds__(256, 4)
im2col_c8_fixed_kernel(const uint4* __restrict__ x, uint4* __restrict__ col, int B, int H, int W, int stride, int pad, int OH, int OW) {
  const unsigned int total = (unsigned int)B * OH * OW * (KH * KW);
  const unsigned int step = gridDim.x * blockDim.x;
  auto locate = [&](unsigned int i, bool& in) -> size_t {
    const unsigned int m = i / (unsigned int)(KH * KW);
    const int t = (int)(i - m * (unsigned int)(KH * KW));
    const int kh = t / KW, kw = t - kh * KW;
    const int ow = (int)(m % (unsigned int)OW);
    const unsigned int r = m / (unsigned int)OW;
    const int oh = (int)(r % (unsigned int)OH);
    const int b = (int)(r / (unsigned int)OH);
    const int h = oh * stride - pad + kh, w = ow * stride - pad + kw;
    in = h >= 0 && h < H && w >= 0 && w < W;
    return in ? ((size_t)b * H + h) * W + w : 0;
  };
  unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
  for (; i < total && total - i > 3 * step; i += 4 * step) {
    uint4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      bool in;
      const size_t src = locate(i + u * step, in);
      v[u] = in ? x[src] : make_uint4(0u, 0u, 0u, 0u);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) col[i + u * step] = v[u];
  }
  for (; i < total; i += step) {
    bool in;
    const size_t src = locate(i, in);
    col[i] = in ? x[src] : make_uint4(0u, 0u, 0u, 0u);
  }
}

__device__ __forceinline__ void add8(float (&acc)[8], const uint4& raw) {
  const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[j]));
    acc[2 * j] += f.x; acc[2 * j + 1] += f.y;
  }
}
__device__ __forceinline__ uint4 pack8f(const float (&v)[8]) {
  uint4 o;
  o.x = pack_bf16(v[0], v[1]); o.y = pack_bf16(v[2], v[3]); o.z = pack_bf16(v[4], v[5]); o.w = pack_bf16(v[6], v[7]);
  return o;
}

// dx[b][h][w][c] = sum over the (kh,kw) whose window covers (h,w) of dcol[(b,oh,ow)][(kh,kw,c)]   (gather: deterministic)
__global__ void __launch_bounds__(256)
col2im_nhwc_kernel(const uint4* __restrict__ dcol, uint4* __restrict__ dx, int B, int H, int W, int C8, int KH, int KW, int stride,
                   int pad, int OH, int OW) {
  const long long total = (long long)B * H * W * C8;
  const long long step = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += step) {
    const int c = (int)(i % C8);
    long long r = i / C8;
    const int w = (int)(r % W); r /= W;
    const int h = (int)(r % H);
    const int b = (int)(r / H);
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int kh = 0; kh < KH; ++kh) {
      const int th = h + pad - kh;
      if (th < 0 || th % stride != 0) continue;
      const int oh = th / stride;
      if (oh >= OH) continue;
      for (int kw = 0; kw < KW; ++kw) {
        const int tw = w + pad - kw;
        if (tw < 0 || tw % stride != 0) continue;
        const int ow = tw / stride;
        if (ow >= OW) continue;
        add8(acc, dcol[((((size_t)b * OH + oh) * OW + ow) * KH * KW + kh * KW + kw) * C8 + c]);
      }
    }
    dx[i] = pack8f(acc);
  }
}

// The same with a compile-time window (square K, stride S, pad P): at most ceil(K / S)^2 taps reach an input pixel; their loads are
// issued together, then summed in the generic kernel's order (kh ascending, kw ascending).
template <int K, int S, int P>
__global__ void __launch_bounds__(256)
col2im_nhwc_fixed_kernel(const uint4* __restrict__ dcol, uint4* __restrict__ dx, int B, int H, int W, int C8, int OH, int OW) {
  constexpr int T = (K + S - 1) / S;
  const long long total = (long long)B * H * W * C8;
  const long long step = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += step) {
    const int c = (int)(i % C8);
    long long r = i / C8;
    const int w = (int)(r % W); r /= W;
    const int h = (int)(r % H);
    const int b = (int)(r / H);
    const int kh0 = (h + P) % S, kw0 = (w + P) % S;
    uint4 raw[T * T];
    bool ok[T * T];
#pragma unroll
    for (int a = 0; a < T; ++a) {
#pragma unroll
      for (int e = 0; e < T; ++e) {
        const int kh = kh0 + a * S, kw = kw0 + e * S;
        const int oh = (h + P - kh) / S, ow = (w + P - kw) / S;
        const bool in = kh < K && kw < K && h + P - kh >= 0 && w + P - kw >= 0 && oh < OH && ow < OW;
        ok[a * T + e] = in;
        raw[a * T + e] = in ? dcol[((((size_t)b * OH + oh) * OW + ow) * K * K + kh * K + kw) * C8 + c] : make_uint4(0u, 0u, 0u, 0u);
      }
    }
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int q = 0; q < T * T; ++q)
      if (ok[q]) add8(acc, raw[q]);
    dx[i] = pack8f(acc);
  }
}

// NCHW fp32 -> NHWC bf16 with the channel dim zero-padded to Cp (a multiple of 8; the 3-channel stem is padded to 8).
// A thread produces 8 channels of one pixel: 8 plane reads that are coalesced across the warp (consecutive pixels) and one
// 16-byte store.
__global__ void __launch_bounds__(256)
nchw_to_nhwc_bf16_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ y, int B, int C, int HW, int Cp8) {
  const long long total = (long long)B * HW * Cp8;
  const long long step = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += step) {
    // pixel index fastest within a channel group so the warp's plane reads are contiguous
    const long long hw = i % HW;
    long long r = i / HW;
    const int g = (int)(r % Cp8);
    const long long b = r / Cp8;
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int c = g * 8 + j;
      v[j] = c < C ? __ldg(x + ((size_t)b * C + c) * HW + hw) : 0.f;
    }
    uint4 o;
    o.x = pack_bf16(v[0], v[1]); o.y = pack_bf16(v[2], v[3]); o.z = pack_bf16(v[4], v[5]); o.w = pack_bf16(v[6], v[7]);
    reinterpret_cast<uint4*>(y)[((size_t)b * HW + hw) * Cp8 + g] = o;
  }
}

// ---- pooling (NHWC bf16) ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
maxpool_nhwc_fwd_kernel(const uint4* __restrict__ x, uint4* __restrict__ y, unsigned char* __restrict__ idx, int B, int H, int W,
                        int C8, int K, int stride, int pad, int OH, int OW) {
  const long long total = (long long)B * OH * OW * C8;
  const long long step = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += step) {
    const int c = (int)(i % C8);
    long long r = i / C8;
    const int ow = (int)(r % OW); r /= OW;
    const int oh = (int)(r % OH);
    const int b = (int)(r / OH);
    float m[8];
    unsigned char am[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { m[j] = -INFINITY; am[j] = 0; }
    for (int kh = 0; kh < K; ++kh) {
      const int h = oh * stride - pad + kh;
      if (h < 0 || h >= H) continue;
      for (int kw = 0; kw < K; ++kw) {
        const int w = ow * stride - pad + kw;
        if (w < 0 || w >= W) continue;
        float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        add8(v, x[(((size_t)b * H + h) * W + w) * C8 + c]);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (v[j] > m[j]) { m[j] = v[j]; am[j] = (unsigned char)(kh * K + kw); }
      }
    }
    y[i] = pack8f(m);
    *reinterpret_cast<uint2*>(idx + i * 8) = *reinterpret_cast<const uint2*>(am);
  }
}

// grid = (ceil(W * C8 / 256), H, B): no per-element index decomposition; only the (at most ceil(K/stride)^2) windows that
// cover an input pixel are visited, their loads are independent.
__global__ void __launch_bounds__(256)
maxpool_nhwc_bwd_kernel(const uint4* __restrict__ dy, const unsigned char* __restrict__ idx, uint4* __restrict__ dx, int B, int H,
                        int W, int C8, int K, int stride, int pad, int OH, int OW) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= W * C8) return;
  const int w = t / C8, c = t - w * C8;
  const int h = blockIdx.y, b = blockIdx.z;
  // windows oh with oh*stride - pad <= h <= oh*stride - pad + K - 1
  const int oh_lo = max(0, (h + pad - K + stride) / stride), oh_hi = min(OH - 1, (h + pad) / stride);
  const int ow_lo = max(0, (w + pad - K + stride) / stride), ow_hi = min(OW - 1, (w + pad) / stride);
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  for (int oh = oh_lo; oh <= oh_hi; ++oh) {
    const int kh = h + pad - oh * stride;
    for (int ow = ow_lo; ow <= ow_hi; ++ow) {
      const int kw = w + pad - ow * stride;
      const size_t o = (((size_t)b * OH + oh) * OW + ow) * C8 + c;
      unsigned char am[8];
      *reinterpret_cast<uint2*>(am) = *reinterpret_cast<const uint2*>(idx + o * 8);
      float g[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      add8(g, dy[o]);
      const unsigned char me = (unsigned char)(kh * K + kw);
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (am[j] == me) acc[j] += g[j];
    }
  }
  dx[(((size_t)b * H + h) * W + w) * C8 + c] = pack8f(acc);
}

// Compile-time window (K, stride S, pad P): every load of a window is issued before the first comparison (the generic kernels
// above walk the window with run-time bounds, one dependent load at a time: 0.22 + 0.47 ms for ResNet-50's 3x3/2 pool at batch
// 256 where the bytes need 0.09 + 0.09).  Same visiting order as the generic kernels, so the recorded arg-max is identical.
template <int K, int S, int P>
__global__ void __launch_bounds__(256)
maxpool_nhwc_fwd_fixed_kernel(const uint4* __restrict__ x, uint4* __restrict__ y, unsigned char* __restrict__ idx, int B, int H, int W,
                              int C8, int OH, int OW) {
  const long long total = (long long)B * OH * OW * C8;
  const long long step = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += step) {
    const int c = (int)(i % C8);
    long long r = i / C8;
    const int ow = (int)(r % OW); r /= OW;
    const int oh = (int)(r % OH);
    const int b = (int)(r / OH);
    uint4 raw[K * K];
    bool ok[K * K];
#pragma unroll
    for (int kh = 0; kh < K; ++kh) {
#pragma unroll
      for (int kw = 0; kw < K; ++kw) {
        const int h = oh * S - P + kh, w = ow * S - P + kw;
        const bool in = h >= 0 && h < H && w >= 0 && w < W;
        ok[kh * K + kw] = in;
        raw[kh * K + kw] = in ? x[(((size_t)b * H + h) * W + w) * C8 + c] : make_uint4(0u, 0u, 0u, 0u);
      }
    }
    float m[8];
    unsigned char am[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { m[j] = -INFINITY; am[j] = 0; }
#pragma unroll
    for (int t = 0; t < K * K; ++t) {
      if (!ok[t]) continue;
      float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      add8(v, raw[t]);
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (v[j] > m[j]) { m[j] = v[j]; am[j] = (unsigned char)t; }
    }
    y[i] = pack8f(m);
    *reinterpret_cast<uint2*>(idx + i * 8) = *reinterpret_cast<const uint2*>(am);
  }
}

template <int K, int S, int P>
__global__ void __launch_bounds__(256)
maxpool_nhwc_bwd_fixed_kernel(const uint4* __restrict__ dy, const unsigned char* __restrict__ idx, uint4* __restrict__ dx, int B, int H,
                              int W, int C8, int OH, int OW) {
  constexpr int T = (K + S - 1) / S;                 // windows that can cover one input pixel, per dimension
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= W * C8) return;
  const int w = t / C8, c = t - w * C8;
  const int h = blockIdx.y, b = blockIdx.z;
  const int oh_hi = (h + P) / S, ow_hi = (w + P) / S;
  uint4 g[T * T];
  uint2 am[T * T];
  int me[T * T];                                     // the window-relative index this pixel has in each candidate, or -1
#pragma unroll
  for (int a = 0; a < T; ++a) {
#pragma unroll
    for (int e = 0; e < T; ++e) {
      const int oh = oh_hi - (T - 1 - a), ow = ow_hi - (T - 1 - e);      // ascending, like the generic kernel
      const int kh = h + P - oh * S, kw = w + P - ow * S;
      const bool in = oh >= 0 && oh < OH && ow >= 0 && ow < OW && kh >= 0 && kh < K && kw >= 0 && kw < K;
      const size_t o = in ? (((size_t)b * OH + oh) * OW + ow) * C8 + c : 0;
      me[a * T + e] = in ? kh * K + kw : -1;
      g[a * T + e] = in ? dy[o] : make_uint4(0u, 0u, 0u, 0u);
      am[a * T + e] = in ? *reinterpret_cast<const uint2*>(idx + o * 8) : make_uint2(0u, 0u);
    }
  }
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int q = 0; q < T * T; ++q) {
    if (me[q] < 0) continue;
    float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    add8(v, g[q]);
    const unsigned int lo = am[q].x, hi = am[q].y;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const unsigned int a = ((j < 4 ? lo : hi) >> (8 * (j & 3))) & 0xFFu;
      if ((int)a == me[q]) acc[j] += v[j];
    }
  }
  dx[(((size_t)b * H + h) * W + w) * C8 + c] = pack8f(acc);
}

// y[b][c] = mean over HW of x[b][hw][c]
__global__ void __launch_bounds__(256)
avgpool_nhwc_fwd_kernel(const uint4* __restrict__ x, uint4* __restrict__ y, int B, int HW, int C8) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * C8) return;
  const int b = i / C8, c = i - b * C8;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  for (int p = 0; p < HW; ++p) add8(acc, x[((size_t)b * HW + p) * C8 + c]);
  const float inv = 1.0f / (float)HW;
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] *= inv;
  y[i] = pack8f(acc);
}

__global__ void __launch_bounds__(256)
avgpool_nhwc_bwd_kernel(const uint4* __restrict__ dy, uint4* __restrict__ dx, int B, int HW, int C8) {
  const long long total = (long long)B * HW * C8;
  const long long step = (long long)gridDim.x * blockDim.x;
  const float inv = 1.0f / (float)HW;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += step) {
    const int c = (int)(i % C8);
    const int b = (int)(i / ((long long)HW * C8));
    float g[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    add8(g, dy[(size_t)b * C8 + c]);
#pragma unroll
    for (int j = 0; j < 8; ++j) g[j] *= inv;
    dx[i] = pack8f(g);
  }
}

static int grid_for(long long items) {
  long long b = ceil_div64(items, 256);
  const long long cap = (long long)kNumSMs * 16;
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

}  // namespace mlt

using namespace mlt;

extern "C" int mlt_im2col_nhwc(const void* x, void* col, int B, int H, int W, int C, int KH, int KW, int stride, int pad, int OH,
                               int OW, void* stream) {
  MLT_REQUIRE(C % 8 == 0, "mlt_im2col_nhwc: channels (%d) must be a multiple of 8", C);
  const long long total = (long long)B * OH * OW * KH * KW * (C / 8);
  if (C == 8 && KH == 7 && KW == 7 && total < (1LL << 32) - 4LL * kNumSMs * 16 * 256) {     // the (channel-padded) ResNet stem
    im2col_c8_fixed_kernel<7, 7><<<grid_for(total / 4), 256, 0, (cudaStream_t)stream>>>((const uint4*)x, (uint4*)col, B, H, W, stride, pad,
                                                                                       OH, OW);
    MLT_LAUNCH_CHECK();
    return 0;
  }
  im2col_nhwc_kernel<<<grid_for(total), 256, 0, (cudaStream_t)stream>>>((const uint4*)x, (uint4*)col, B, H, W, C / 8, KH, KW, stride,
                                                                       pad, OH, OW);
  MLT_LAUNCH_CHECK();
  return 0;
}

extern "C" int mlt_col2im_nhwc(const void* dcol, void* dx, int B, int H, int W, int C, int KH, int KW, int stride, int pad, int OH,
                               int OW, void* stream) {
  MLT_REQUIRE(C % 8 == 0, "mlt_col2im_nhwc: channels (%d) must be a multiple of 8", C);
  const long long total = (long long)B * H * W * (C / 8);
  if (KH == KW && stride == 2 && ((KH == 3 && pad == 1) || (KH == 1 && pad == 0))) {       // ResNet's two stride-2 convolutions
    if (KH == 3)
      col2im_nhwc_fixed_kernel<3, 2, 1><<<grid_for(total), 256, 0, (cudaStream_t)stream>>>((const uint4*)dcol, (uint4*)dx, B, H, W, C / 8, OH, OW);
    else
      col2im_nhwc_fixed_kernel<1, 2, 0><<<grid_for(total), 256, 0, (cudaStream_t)stream>>>((const uint4*)dcol, (uint4*)dx, B, H, W, C / 8, OH, OW);
    MLT_LAUNCH_CHECK();
    return 0;
  }
  col2im_nhwc_kernel<<<grid_for(total), 256, 0, (cudaStream_t)stream>>>((const uint4*)dcol, (uint4*)dx, B, H, W, C / 8, KH, KW, stride,
                                                                       pad, OH, OW);
  MLT_LAUNCH_CHECK();
  return 0;
}

extern "C" int mlt_nchw_to_nhwc_bf16(const void* x_f32, void* y_bf16, int B, int C, int H, int W, int Cp, void* stream) {
  MLT_REQUIRE(Cp >= C && Cp % 8 == 0, "mlt_nchw_to_nhwc_bf16: the padded channel count must be a multiple of 8 and >= C");
  nchw_to_nhwc_bf16_kernel<<<grid_for((long long)B * H * W * (Cp / 8)), 256, 0, (cudaStream_t)stream>>>(
      (const float*)x_f32, (__nv_bfloat16*)y_bf16, B, C, H * W, Cp / 8);
  MLT_LAUNCH_CHECK();
  return 0;
}

extern "C" int mlt_maxpool_nhwc_fwd(const void* x, void* y, void* idx_u8, int B, int H, int W, int C, int K, int stride, int pad,
                                    int OH, int OW, void* stream) {
  MLT_REQUIRE(C % 8 == 0 && K * K <= 255, "mlt_maxpool_nhwc_fwd: unsupported shape");
  if (K == 3 && stride == 2 && pad == 1) {            // the ResNet stem pool
    maxpool_nhwc_fwd_fixed_kernel<3, 2, 1><<<grid_for((long long)B * OH * OW * (C / 8)), 256, 0, (cudaStream_t)stream>>>(
        (const uint4*)x, (uint4*)y, (unsigned char*)idx_u8, B, H, W, C / 8, OH, OW);
    MLT_LAUNCH_CHECK();
    return 0;
  }
  maxpool_nhwc_fwd_kernel<<<grid_for((long long)B * OH * OW * (C / 8)), 256, 0, (cudaStream_t)stream>>>(
      (const uint4*)x, (uint4*)y, (unsigned char*)idx_u8, B, H, W, C / 8, K, stride, pad, OH, OW);
  MLT_LAUNCH_CHECK();
  return 0;
}

extern "C" int mlt_maxpool_nhwc_bwd(const void* dy, const void* idx_u8, void* dx, int B, int H, int W, int C, int K, int stride,
                                    int pad, int OH, int OW, void* stream) {
  MLT_REQUIRE(C % 8 == 0, "mlt_maxpool_nhwc_bwd: unsupported shape");
  MLT_REQUIRE(H <= 65535 && B <= 65535, "mlt_maxpool_nhwc_bwd: H and B must fit a grid dimension");
  if (K == 3 && stride == 2 && pad == 1) {
    maxpool_nhwc_bwd_fixed_kernel<3, 2, 1><<<dim3(ceil_div(W * (C / 8), 256), H, B), 256, 0, (cudaStream_t)stream>>>(
        (const uint4*)dy, (const unsigned char*)idx_u8, (uint4*)dx, B, H, W, C / 8, OH, OW);
    MLT_LAUNCH_CHECK();
    return 0;
  }
  maxpool_nhwc_bwd_kernel<<<dim3(ceil_div(W * (C / 8), 256), H, B), 256, 0, (cudaStream_t)stream>>>(
      (const uint4*)dy, (const unsigned char*)idx_u8, (uint4*)dx, B, H, W, C / 8, K, stride, pad, OH, OW);
  MLT_LAUNCH_CHECK();
  return 0;
}

extern "C" int mlt_avgpool_nhwc_fwd(const void* x, void* y, int B, int HW, int C, void* stream) {
  MLT_REQUIRE(C % 8 == 0, "mlt_avgpool_nhwc_fwd: channels must be a multiple of 8");
  avgpool_nhwc_fwd_kernel<<<ceil_div(B * (C / 8), 256), 256, 0, (cudaStream_t)stream>>>((const uint4*)x, (uint4*)y, B, HW, C / 8);
  MLT_LAUNCH_CHECK();
  return 0;
}

extern "C" int mlt_avgpool_nhwc_bwd(const void* dy, void* dx, int B, int HW, int C, void* stream) {
  MLT_REQUIRE(C % 8 == 0, "mlt_avgpool_nhwc_bwd: channels must be a multiple of 8");
  avgpool_nhwc_bwd_kernel<<<grid_for((long long)B * HW * (C / 8)), 256, 0, (cudaStream_t)stream>>>((const uint4*)dy, (uint4*)dx, B, HW,
                                                                                                C / 8);
  MLT_LAUNCH_CHECK();
  return 0;
}

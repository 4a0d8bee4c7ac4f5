// Training-mode BatchNorm for NHWC bf16 activations viewed as [M = B*H*W, C], with ReLU and the
// residual add fused into the normalisation pass, and the running-statistics update fused into the reduction.
//   forward :  stats (per-channel sum / sum of squares, two-level reduction)  ->  apply (+residual, +ReLU)
//   backward:  reduce (dgamma, dbeta with the ReLU mask applied on the fly)   ->  apply (dx, and the residual gradient)
// Bandwidth-bound; roofline = bytes / HBM bandwidth.  Everything is addressed as a flat stream of 16-byte groups (8
// channels): consecutive threads read consecutive groups whatever C is, and because the grid stride is a multiple of
// C/8 a thread sees the SAME 8 channels on every iteration - its per-channel constants live in registers and the inner
// loops are 16-byte loads, a handful of FMAs and 16-byte stores, several iterations in flight.
// Channel reductions: a CTA covers 256/(C/8) rows per iteration, folds them through shared memory, CTAs of a channel
// tile combine through fp32 atomics on a scratch accumulator + a ticket; the last CTA finalises.
#include "common.cuh"
#include "colreduce.cuh"
#include "../include/mlt_b200.h"

namespace mlt {

template <typename P> __device__ __forceinline__ float ldp(const P* p, int i) { return to_float(p[i]); }

// Which of the 8 elements at element offset `off` passed the forward ReLU: from the 1-bit-per-element mask the forward
// wrote (one byte instead of 16), or recomputed from y when no mask was kept.
template <bool RELU>
__device__ __forceinline__ unsigned int relu_bits(const __nv_bfloat16* y, const unsigned char* mask, size_t off) {
  if (!RELU) return 0xFFu;
  if (mask != nullptr) return mask[off >> 3];
  float o[8];
  ld8(y + off, o);
  unsigned int b = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) b |= (o[j] > 0.f ? 1u : 0u) << j;
  return b;
}

__global__ void __launch_bounds__(kBnThreads, 3)
bn_stats_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ mean, float* __restrict__ rstd, float* __restrict__ run_mean,
                float* __restrict__ run_var, float momentum, float eps, float* __restrict__ partial, unsigned int* __restrict__ tickets,
                int M, int C) {
  __shared__ float tot[2 * kBnGroups * 8];
  struct {
    const __nv_bfloat16* x;
    __device__ __forceinline__ uint4 load(size_t off) const { return ld8raw(x + off); }
    __device__ __forceinline__ void pin(const uint4& raw) const { asm volatile("" ::"r"(raw.x) : "memory"); }
    __device__ __forceinline__ void use(const uint4& raw, float (&a)[8], float (&b)[8]) const {
      float v[8];
      cvt8(raw, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) { a[j] += v[j]; b[j] = fmaf(v[j], v[j], b[j]); }
    }
  } f{x};
  if (!col_reduce_atomic<true, 8, true>(f, partial, tickets, M, C, tot)) return;
  const int g_lo = blockIdx.x * kBnGroups;
  const int nch = min(kBnGroups, (C >> 3) - g_lo) * 8;
  for (int ch = threadIdx.x; ch < nch; ch += kBnThreads) {
    const int c = g_lo * 8 + ch;
    const float m = tot[ch] / (float)M;
    const float var = fmaxf(tot[nch + ch] / (float)M - m * m, 0.f);
    mean[c] = m;
    rstd[c] = rsqrtf(var + eps);
    if (run_mean != nullptr) {
      run_mean[c] = (1.f - momentum) * run_mean[c] + momentum * m;
      run_var[c] = (1.f - momentum) * run_var[c] + momentum * var * ((float)M / (float)max(M - 1, 1));
    }
  }
}

// y = [relu]( (x - mean) * rstd * gamma + beta [+ residual] )
template <typename P, bool RES, bool RELU>
__global__ void __launch_bounds__(kBnThreads, 3)
bn_apply_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ mean, const float* __restrict__ rstd,
                const P* __restrict__ gamma, const P* __restrict__ beta, const __nv_bfloat16* __restrict__ residual,
                __nv_bfloat16* __restrict__ y, unsigned char* __restrict__ mask, long long total, int C8) {
  const long long step = (long long)gridDim.x * kBnThreads;         // a multiple of C8 (host guarantees it)
  long long i = (long long)blockIdx.x * kBnThreads + threadIdx.x;
  const int c0 = (int)(i % C8) * 8;
  float sc[8], sf[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    sc[j] = rstd[c0 + j] * ldp(gamma, c0 + j);
    sf[j] = fmaf(-mean[c0 + j], sc[j], ldp(beta, c0 + j));
  }
  auto one = [&](long long e) {
    float v[8], r[8];
    ld8(x + e * 8, v);
    if (RES) ld8(residual + e * 8, r);
    unsigned int bits = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float o = fmaf(v[j], sc[j], sf[j]);
      if (RES) o += r[j];
      if (RELU) bits |= (o > 0.f ? 1u : 0u) << j;
      v[j] = RELU ? fmaxf(o, 0.f) : o;
    }
    st8(y + e * 8, v);
    if (RELU && mask != nullptr) mask[e] = (unsigned char)bits;      // 1 bit per element: all backward needs of y
  };
  // four positions per iteration, all loads first (x, and the residual when there is one), then the arithmetic and the stores
  auto finish = [&](long long e, const uint4& xr, const uint4& rr) {
    float v[8], r[8];
    cvt8(xr, v);
    if (RES) cvt8(rr, r);
    unsigned int bits = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float o = fmaf(v[j], sc[j], sf[j]);
      if (RES) o += r[j];
      if (RELU) bits |= (o > 0.f ? 1u : 0u) << j;
      v[j] = RELU ? fmaxf(o, 0.f) : o;
    }
    st8(y + e * 8, v);
    if (RELU && mask != nullptr) mask[e] = (unsigned char)bits;
  };
  for (; i + 3 * step < total; i += 4 * step) {
    uint4 xr[4], rr[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      xr[u] = ld8raw(x + (i + u * step) * 8);
      if (RES) rr[u] = ld8raw(residual + (i + u * step) * 8); else rr[u] = make_uint4(0u, 0u, 0u, 0u);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      asm volatile("" ::"r"(xr[u].x) : "memory");
      if (RES) asm volatile("" ::"r"(rr[u].x) : "memory");
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) finish(i + u * step, xr[u], rr[u]);
  }
  for (; i < total; i += step) one(i);
}

// dgamma[c] = sum dz * xhat, dbeta[c] = sum dz, with dz = dy * (y > 0) when the forward had a ReLU
template <typename G, bool RELU>
__global__ void __launch_bounds__(kBnThreads)
bn_bwd_reduce_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ y, const unsigned char* __restrict__ mask,
                     const __nv_bfloat16* __restrict__ x, const float* __restrict__ mean, const float* __restrict__ rstd,
                     G* __restrict__ dgamma, G* __restrict__ dbeta, float* __restrict__ sums, float* __restrict__ partial,
                     unsigned int* __restrict__ tickets, int M, int C) {
  __shared__ float tot[2 * kBnGroups * 8];
  const int g_lo = blockIdx.x * kBnGroups;
  const int cw = min(kBnGroups, (C >> 3) - g_lo);
  const int c0 = (g_lo + (int)(threadIdx.x % cw)) * 8;
  float mu[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) mu[j] = mean[c0 + j];
  auto f = [&](size_t off, float (&a)[8], float (&b)[8]) {
    float g[8], v[8];
    ld8(dy + off, g);
    ld8(x + off, v);
    const unsigned int bits = relu_bits<RELU>(y, mask, off);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float dz = (RELU && !((bits >> j) & 1u)) ? 0.f : g[j];
      a[j] = fmaf(dz, v[j] - mu[j], a[j]);          // sum dz (x - mean); the 1/std factor is applied once at the end
      b[j] += dz;
    }
  };
  if (!col_reduce_atomic<true>(f, partial, tickets, M, C, tot)) return;
  const int nch = cw * 8;
  for (int ch = threadIdx.x; ch < nch; ch += kBnThreads) {
    const int c = g_lo * 8 + ch;
    const float dg = tot[ch] * rstd[c], db = tot[nch + ch];
    dgamma[c] = from_float<G>(dg);
    dbeta[c] = from_float<G>(db);
    sums[c] = dg;                                   // fp32 copies for the apply pass
    sums[C + c] = db;
  }
}

// dx = gamma * rstd * (dz - dbeta/M - xhat * dgamma/M);  dres = dz (gradient of the fused residual input)
template <typename P, bool RES, bool RELU>
__global__ void __launch_bounds__(kBnThreads, 2)
bn_bwd_apply_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ y, const unsigned char* __restrict__ mask,
                    const __nv_bfloat16* __restrict__ x, const float* __restrict__ mean, const float* __restrict__ rstd, const P* __restrict__ gamma,
                    const float* __restrict__ sums, __nv_bfloat16* __restrict__ dx, __nv_bfloat16* __restrict__ dres, long long total,
                    int C8, float inv_M) {
  const long long step = (long long)gridDim.x * kBnThreads;
  long long i = (long long)blockIdx.x * kBnThreads + threadIdx.x;
  const int c0 = (int)(i % C8) * 8;
  const int C = C8 * 8;
  // dx = A dz + B (x - mean) + K   with A = gamma rstd, B = -A rstd dgamma / M, K = -A dbeta / M
  float A[8], Bc[8], K[8], mu[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float rs = rstd[c0 + j];
    A[j] = ldp(gamma, c0 + j) * rs;
    Bc[j] = -A[j] * rs * sums[c0 + j] * inv_M;
    K[j] = -A[j] * sums[C + c0 + j] * inv_M;
    mu[j] = mean[c0 + j];
  }
  auto one = [&](long long e) {
    float g[8], v[8], d[8];
    ld8(dy + e * 8, g);
    ld8(x + e * 8, v);
    const unsigned int bits = relu_bits<RELU>(y, mask, (size_t)e * 8);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float dz = (RELU && !((bits >> j) & 1u)) ? 0.f : g[j];
      d[j] = fmaf(A[j], dz, fmaf(Bc[j], v[j] - mu[j], K[j]));
      g[j] = dz;
    }
    st8(dx + e * 8, d);
    if (RES) st8(dres + e * 8, g);
  };
  // four positions per iteration with every load issued first (dy, x and the mask byte of each), then arithmetic and stores
  auto finish = [&](long long e, const uint4& gr, const uint4& xr, unsigned int bits) {
    float g[8], v[8], d[8];
    cvt8(gr, g);
    cvt8(xr, v);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float dz = (RELU && !((bits >> j) & 1u)) ? 0.f : g[j];
      d[j] = fmaf(A[j], dz, fmaf(Bc[j], v[j] - mu[j], K[j]));
      g[j] = dz;
    }
    st8(dx + e * 8, d);
    if (RES) st8(dres + e * 8, g);
  };
  if (!RELU || mask != nullptr) {
    for (; i + 3 * step < total; i += 4 * step) {
      uint4 gr[4], xr[4];
      unsigned int bits[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const long long e = i + u * step;
        gr[u] = ld8raw(dy + e * 8);
        xr[u] = ld8raw(x + e * 8);
        bits[u] = RELU ? (unsigned int)mask[e] : 0xFFu;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) finish(i + u * step, gr[u], xr[u], bits[u]);
    }
  }
  for (; i + step < total; i += 2 * step) {
    one(i);
    one(i + step);
  }
  for (; i < total; i += step) one(i);
}

static bool bn_shape_ok(int C) { return C > 0 && C % 8 == 0 && C <= 16384 && C / 8 <= 64 * kBnGroups; }
static long long gcd_ll(long long a, long long b) { return b == 0 ? a : gcd_ll(b, a % b); }
// Row split of the two reduction passes.  The statistics pass keeps eight 16-byte loads in flight per thread (67 registers, three
// CTAs per SM), so its grid is 3 CTAs per SM: one full wave.  MLT_BN_R / MLT_BN_R_STATS override the multipliers.
static int bn_rows_split(int M, int C, bool stats = false) {
  const int C8 = C / 8;
  const int tiles = ceil_div(C8, kBnGroups);
  const int rpi = kBnThreads / (C8 < kBnGroups ? C8 : kBnGroups);
  static int mult_r = -1, mult_s = -1;
  if (mult_r < 0) { const char* e = getenv("MLT_BN_R"); mult_r = e ? atoi(e) : 4; }
  if (mult_s < 0) { const char* e = getenv("MLT_BN_R_STATS"); mult_s = e ? atoi(e) : 3; }
  const int mult = stats ? mult_s : mult_r;
  int R = (mult * kNumSMs) / tiles;
  const int cap = ceil_div(M, rpi * 8);               // at least 8 iterations per CTA
  if (R > cap) R = cap;
  return R < 1 ? 1 : R;
}
// Grid for the elementwise passes: the stride (grid * 256) must be a multiple of C/8 so channels are loop-invariant.  Three CTAs per
// SM are resident (66-77 registers with the batched loads), so 3 x 148 CTAs is one full wave; measured over the ResNet-50 shapes the
// forward passes take 4.29 ms at 3 against 4.55 ms at 8 (profiles/r2/bn_shapes_*.json), the whole step 24.8 against 25.2 ms.
static int bn_grid(long long items, int C8, int resident_per_sm = 3) {
  long long b = ceil_div64(items, kBnThreads);
  static int env_mult = -1;
  if (env_mult < 0) { const char* e = getenv("MLT_BN_GRID_MULT"); env_mult = e ? atoi(e) : 0; if (env_mult < 0) env_mult = 0; }
  const int mult = env_mult > 0 ? env_mult : resident_per_sm;
  const long long cap = (long long)kNumSMs * mult;
  if (b > cap) b = cap;
  const long long m = C8 / gcd_ll(C8, kBnThreads);       // grid * 256 must be a multiple of C8
  b = (b + m - 1) / m * m;
  return (int)(b < 1 ? m : b);
}

}  // namespace mlt

using namespace mlt;

// scratch (zeroed once by the caller, shared by layers of any width): [tickets 256 B][accumulators, fixed 128 KB window]
// [sums 2*C floats].  Tickets and accumulators are left zero by every call; sums are plain outputs and sit behind the
// fixed window so a narrow layer's sums can never land where a wide layer accumulates.
constexpr int kBnMaxC = 16384;
constexpr long long kBnAccBytes = 2LL * kBnMaxC * 4;
extern "C" long long mlt_bn_scratch_bytes(int M, int C) {
  (void)M;
  return 256 + kBnAccBytes + 2LL * C * 4;
}

static void bn_carve(void* scratch, int M, int C, unsigned int** tickets, float** partial, float** sums) {
  (void)M; (void)C;
  *tickets = (unsigned int*)scratch;
  *partial = (float*)((char*)scratch + 256);
  *sums = (float*)((char*)scratch + 256 + kBnAccBytes);
}

extern "C" int mlt_bn_stats(const void* x, void* mean, void* rstd, void* run_mean, void* run_var, float momentum, float eps,
                            void* scratch, long long scratch_bytes, int M, int C, void* stream) {
  MLT_REQUIRE(M > 0 && bn_shape_ok(C), "mlt_bn_stats: channels (%d) must be a multiple of 8", C);
  MLT_REQUIRE(scratch && scratch_bytes >= mlt_bn_scratch_bytes(M, C), "mlt_bn_stats: scratch too small");
  unsigned int* tk; float *pa, *su;
  bn_carve(scratch, M, C, &tk, &pa, &su);
  dim3 grid(ceil_div(C / 8, kBnGroups), bn_rows_split(M, C, true));
  bn_stats_kernel<<<grid, kBnThreads, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, (float*)mean, (float*)rstd, (float*)run_mean,
                                                                (float*)run_var, momentum, eps, pa, tk, M, C);
  MLT_LAUNCH_CHECK();
  return 0;
}

template <typename P>
static void bn_apply_launch(int g, cudaStream_t st, const void* x, const void* mean, const void* rstd, const void* gamma,
                            const void* beta, const void* residual, void* y, void* mask, long long total, int C8, int relu) {
#define BN_AP(RES, RELU)                                                                                                      \
  bn_apply_kernel<P, RES, RELU><<<g, kBnThreads, 0, st>>>((const __nv_bfloat16*)x, (const float*)mean, (const float*)rstd,        \
      (const P*)gamma, (const P*)beta, (const __nv_bfloat16*)residual, (__nv_bfloat16*)y, (unsigned char*)mask, total, C8)
  if (residual) { if (relu) BN_AP(true, true); else BN_AP(true, false); }
  else { if (relu) BN_AP(false, true); else BN_AP(false, false); }
#undef BN_AP
}

extern "C" int mlt_bn_apply(const void* x, const void* mean, const void* rstd, const void* gamma, const void* beta, int params_bf16,
                            const void* residual, void* y, void* relu_mask, long long M, int C, int relu, void* stream) {
  MLT_REQUIRE(bn_shape_ok(C), "mlt_bn_apply: channels (%d) must be a multiple of 8", C);
  cudaStream_t st = (cudaStream_t)stream;
  const int C8 = C / 8;
  const long long total = M * C8;
  const int g = bn_grid(total, C8);
  if (params_bf16) bn_apply_launch<__nv_bfloat16>(g, st, x, mean, rstd, gamma, beta, residual, y, relu_mask, total, C8, relu);
  else bn_apply_launch<float>(g, st, x, mean, rstd, gamma, beta, residual, y, relu_mask, total, C8, relu);
  MLT_LAUNCH_CHECK();
  return 0;
}

template <typename P>
static void bn_bwd_apply_launch(int g, cudaStream_t st, const void* dy, const void* y, const void* mask, const void* x, const void* mean,
                                const void* rstd, const void* gamma, const float* su, void* dx, void* dres, long long total, int C8,
                                float inv_M, int relu) {
#define BN_BA(RES, RELU)                                                                                                      \
  bn_bwd_apply_kernel<P, RES, RELU><<<g, kBnThreads, 0, st>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)y,                 \
      (const unsigned char*)mask, (const __nv_bfloat16*)x, (const float*)mean, (const float*)rstd, (const P*)gamma, su, (__nv_bfloat16*)dx,                 \
      (__nv_bfloat16*)dres, total, C8, inv_M)
  if (dres) { if (relu) BN_BA(true, true); else BN_BA(true, false); }
  else { if (relu) BN_BA(false, true); else BN_BA(false, false); }
#undef BN_BA
}

extern "C" int mlt_bn_bwd(const void* dy, const void* y, const void* relu_mask, const void* x, const void* mean, const void* rstd, const void* gamma,
                          int params_bf16, void* dx, void* dres, void* dgamma, void* dbeta, int grads_bf16, void* scratch,
                          long long scratch_bytes, int M, int C, int relu, void* stream) {
  MLT_REQUIRE(M > 0 && bn_shape_ok(C), "mlt_bn_bwd: channels (%d) must be a multiple of 8", C);
  MLT_REQUIRE(scratch && scratch_bytes >= mlt_bn_scratch_bytes(M, C), "mlt_bn_bwd: scratch too small");
  unsigned int* tk; float *pa, *su;
  bn_carve(scratch, M, C, &tk, &pa, &su);
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid(ceil_div(C / 8, kBnGroups), bn_rows_split(M, C));
#define BN_RD(G, RELU)                                                                                                        \
  bn_bwd_reduce_kernel<G, RELU><<<grid, kBnThreads, 0, st>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)y,                  \
      (const unsigned char*)relu_mask, (const __nv_bfloat16*)x, (const float*)mean, (const float*)rstd, (G*)dgamma, (G*)dbeta, su, pa, tk, M, C)
  if (grads_bf16) { if (relu) BN_RD(__nv_bfloat16, true); else BN_RD(__nv_bfloat16, false); }
  else { if (relu) BN_RD(float, true); else BN_RD(float, false); }
#undef BN_RD
  MLT_LAUNCH_CHECK();
  const int C8 = C / 8;
  const long long total = (long long)M * C8;
  const int g = bn_grid(total, C8, 2);          // 92-96 registers with the batched loads: two CTAs per SM are resident
  const float inv_M = 1.0f / (float)M;
  if (params_bf16) bn_bwd_apply_launch<__nv_bfloat16>(g, st, dy, y, relu_mask, x, mean, rstd, gamma, su, dx, dres, total, C8, inv_M, relu);
  else bn_bwd_apply_launch<float>(g, st, dy, y, relu_mask, x, mean, rstd, gamma, su, dx, dres, total, C8, inv_M, relu);
  MLT_LAUNCH_CHECK();
  return 0;
}

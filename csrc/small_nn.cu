// Small-shape CUDA-core path for the reference CNN (§8 K7 "tiny-N path", K10).
//
// MLModel's layers are degenerate GEMMs (N = 6 / 16 / 10, K = 75 / 150 / 84): far below one tcgen05 tile, so
// they run as fused fp32 CUDA-core kernels whose job is to minimise launches and memory round trips:
//   conv + bias + ReLU + 2x2 max-pool            -> one kernel (pool arg-max kept as 2 bits for backward)
//   its backward (pool scatter, ReLU mask, wgrad, bias grad, dgrad) -> two kernels
//   linear + bias + activation                   -> one kernel;  dX, dW, db of a linear -> one kernel
// The step is launch-latency bound (~0.13 GFLOP); the roofline that applies is launches x ~2 us, not FLOPs.
// Larger layers go to the tcgen05 GEMM / implicit-GEMM path instead (gemm_tcgen05.cu).
#include "common.cuh"
#include "../include/mlt_b200.h"

namespace mlt {

// =====================================================================================================
// conv2d (stride 1, no padding, NCHW fp32) + bias [+ ReLU] [+ 2x2/2 max-pool]
// grid = (B, co_groups); each CTA stages one input sample and its group's filters in shared memory.
// =====================================================================================================
template <bool RELU, bool POOL>
__global__ void __launch_bounds__(256)
conv2d_direct_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                         float* __restrict__ y, unsigned char* __restrict__ idx, int Cin, int H, int W, int Cout,
                         int KH, int KW, int co_per_cta) {
  extern __shared__ float smem[];
  const int b = blockIdx.x;
  const int co0 = blockIdx.y * co_per_cta;
  const int nco = min(co_per_cta, Cout - co0);
  const int OH = H - KH + 1, OW = W - KW + 1;
  const int PH = POOL ? OH / 2 : OH, PW = POOL ? OW / 2 : OW;
  const int wsz = Cin * KH * KW;
  float* xs = smem;                    // [Cin][H][W]
  float* ws = smem + Cin * H * W;      // [nco][Cin][KH][KW]
  const float* xb = x + (size_t)b * Cin * H * W;
  for (int i = threadIdx.x; i < Cin * H * W; i += blockDim.x) xs[i] = xb[i];
  for (int i = threadIdx.x; i < nco * wsz; i += blockDim.x) ws[i] = w[(size_t)co0 * wsz + i];
  __syncthreads();
  const int per = PH * PW;
  for (int o = threadIdx.x; o < nco * per; o += blockDim.x) {
    const int cl = o / per, r = o - cl * per;
    const int ph = r / PW, pw = r - ph * PW;
    const float* wc = ws + cl * wsz;
    const float bv = bias ? bias[co0 + cl] : 0.f;
    if (POOL) {
      float a00 = bv, a01 = bv, a10 = bv, a11 = bv;
      const int h0 = 2 * ph, w0 = 2 * pw;
      for (int ci = 0; ci < Cin; ++ci) {
        const float* xc = xs + ci * H * W + h0 * W + w0;
        const float* wk = wc + ci * KH * KW;
        for (int kh = 0; kh < KH; ++kh) {
          const float* r0 = xc + kh * W;
          const float* r1 = r0 + W;
          float p0 = r0[0], p1 = r1[0];
          for (int kw = 0; kw < KW; ++kw) {
            const float wv = wk[kh * KW + kw];
            const float q0 = r0[kw + 1], q1 = r1[kw + 1];
            a00 = fmaf(wv, p0, a00); a01 = fmaf(wv, q0, a01);
            a10 = fmaf(wv, p1, a10); a11 = fmaf(wv, q1, a11);
            p0 = q0; p1 = q1;
          }
        }
      }
      float m = a00; int am = 0;                       // first maximum wins, like ATen max_pool2d
      if (a01 > m) { m = a01; am = 1; }
      if (a10 > m) { m = a10; am = 2; }
      if (a11 > m) { m = a11; am = 3; }
      if (RELU) m = fmaxf(m, 0.f);
      const size_t oi = ((size_t)b * Cout + co0 + cl) * per + r;
      y[oi] = m;
      idx[oi] = (unsigned char)am;
    } else {
      float a = bv;
      for (int ci = 0; ci < Cin; ++ci) {
        const float* xc = xs + ci * H * W + ph * W + pw;
        const float* wk = wc + ci * KH * KW;
        for (int kh = 0; kh < KH; ++kh)
          for (int kw = 0; kw < KW; ++kw) a = fmaf(wk[kh * KW + kw], xc[kh * W + kw], a);
      }
      if (RELU) a = fmaxf(a, 0.f);
      y[((size_t)b * Cout + co0 + cl) * per + r] = a;
    }
  }
}

// -----------------------------------------------------------------------------------------------------
// wgrad + bias grad.  grid = (Cout, nchunk): CTA (co, c) walks batch chunk c.  For each sample the upstream
// gradient of channel co is turned into a sparse list (value, input offset) - with pooling only one of the
// four pre-pool positions is live - and each thread owns one filter tap (ci,kh,kw); slot nW is the bias.
// Chunks are combined through scratch: partial[chunk][co][nW+1] + one ticket per co; the last CTA of a
// channel sums the partials in a fixed order (deterministic) and re-zeroes the ticket.
// -----------------------------------------------------------------------------------------------------
template <bool RELU, bool POOL>
__global__ void __launch_bounds__(256)
conv2d_direct_wgrad_kernel(const float* __restrict__ x, const float* __restrict__ dy, const float* __restrict__ y,
                           const unsigned char* __restrict__ idx, float* __restrict__ dw, float* __restrict__ db,
                           float* __restrict__ partial, unsigned int* __restrict__ tickets, int B, int Cin, int H,
                           int W, int Cout, int KH, int KW, int chunk, int accumulate) {
  extern __shared__ float smem[];
  const int co = blockIdx.x, c = blockIdx.y, nchunk = gridDim.y;
  const int OH = H - KH + 1, OW = W - KW + 1;
  const int PH = POOL ? OH / 2 : OH, PW = POOL ? OW / 2 : OW;
  const int per = PH * PW, nW = Cin * KH * KW;
  float* xs = smem;                               // [Cin*H*W]
  float* gs = xs + Cin * H * W;                   // [per] masked upstream gradient
  int* offs = reinterpret_cast<int*>(gs + per);   // [per] offset of the live pre-pool position in the input plane
  __shared__ bool is_last;

  // taps owned by this thread: t, t+256, ... (at most 4 -> up to 1024 taps + bias per output channel)
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  int tap_off[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int t = threadIdx.x + j * 256;
    if (t < nW) {
      const int ci = t / (KH * KW), r = t - ci * KH * KW;
      tap_off[j] = ci * H * W + (r / KW) * W + (r % KW);
    } else {
      tap_off[j] = -1;                            // bias slot (t == nW) or idle
    }
  }
  const int b_lo = c * chunk, b_hi = min(B, b_lo + chunk);
  for (int b = b_lo; b < b_hi; ++b) {
    __syncthreads();
    const float* xb = x + (size_t)b * Cin * H * W;
    for (int i = threadIdx.x; i < Cin * H * W; i += blockDim.x) xs[i] = xb[i];
    const size_t base = ((size_t)b * Cout + co) * per;
    for (int i = threadIdx.x; i < per; i += blockDim.x) {
      float g = dy[base + i];
      if (RELU && !(y[base + i] > 0.f)) g = 0.f;
      const int ph = i / PW, pw = i - ph * PW;
      int off;
      if (POOL) {
        const int am = idx[base + i];
        off = (2 * ph + (am >> 1)) * W + 2 * pw + (am & 1);
      } else {
        off = ph * W + pw;
      }
      gs[i] = g;
      offs[i] = off;
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int t = threadIdx.x + j * 256;
      if (t < nW) {
        const float* xt = xs + tap_off[j];
        float a = acc[j];
#pragma unroll 4
        for (int i = 0; i < per; ++i) a = fmaf(gs[i], xt[offs[i]], a);
        acc[j] = a;
      } else if (t == nW) {
        float a = acc[j];
        for (int i = 0; i < per; ++i) a += gs[i];
        acc[j] = a;
      }
    }
  }
  // publish
  const int slots = nW + 1;
  if (nchunk == 1) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int t = threadIdx.x + j * 256;
      if (t < nW) dw[(size_t)co * nW + t] = (accumulate ? dw[(size_t)co * nW + t] : 0.f) + acc[j];
      else if (t == nW && db) db[co] = (accumulate ? db[co] : 0.f) + acc[j];
    }
    return;
  }
  float* mine = partial + ((size_t)c * Cout + co) * slots;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int t = threadIdx.x + j * 256;
    if (t < slots) mine[t] = acc[j];
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int tk = atomicAdd(&tickets[co], 1u);
    is_last = (tk == (unsigned int)nchunk - 1);
    if (is_last) tickets[co] = 0u;                 // ready for the next launch
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  for (int t = threadIdx.x; t < slots; t += blockDim.x) {
    float s = 0.f;
    for (int k = 0; k < nchunk; ++k) s += __ldcg(partial + ((size_t)k * Cout + co) * slots + t);
    if (t < nW) dw[(size_t)co * nW + t] = (accumulate ? dw[(size_t)co * nW + t] : 0.f) + s;
    else if (db) db[co] = (accumulate ? db[co] : 0.f) + s;
  }
}

// -----------------------------------------------------------------------------------------------------
// dgrad: dx[b,ci,h,w] = sum_{co,kh,kw} dz[b,co,h-kh,w-kw] * w[co,ci,kh,kw].  grid = (B, Cin).  The pre-pool
// gradient plane of the sample is rebuilt in shared memory from (dy, y, idx) with a zero halo of K-1 so the
// tap loops carry no bounds checks; KS > 0 fixes the (square) filter size at compile time and unrolls them.
// -----------------------------------------------------------------------------------------------------
template <bool RELU, bool POOL, int KS>
__global__ void __launch_bounds__(256)
conv2d_direct_dgrad_kernel(const float* __restrict__ dy, const float* __restrict__ y, const unsigned char* __restrict__ idx,
                           const float* __restrict__ w, float* __restrict__ dx, int Cin, int H, int W, int Cout, int KH_,
                           int KW_) {
  extern __shared__ float smem[];
  const int KH = KS > 0 ? KS : KH_, KW = KS > 0 ? KS : KW_;
  const int b = blockIdx.x, ci = blockIdx.y;
  const int OH = H - KH + 1, OW = W - KW + 1;
  const int PH = POOL ? OH / 2 : OH, PW = POOL ? OW / 2 : OW;
  const int per = PH * PW, ksz = KH * KW;
  const int ZH = OH + 2 * (KH - 1), ZW = OW + 2 * (KW - 1), zplane = ZH * ZW;   // haloed plane
  float* dz = smem;                       // [Cout][ZH][ZW]
  float* ws = smem + Cout * zplane;       // [Cout][KH][KW]   (this CTA's input channel only)
  for (int i = threadIdx.x; i < Cout * zplane; i += blockDim.x) dz[i] = 0.f;
  for (int i = threadIdx.x; i < Cout * ksz; i += blockDim.x) {
    const int co = i / ksz, r = i - co * ksz;
    ws[i] = w[((size_t)co * Cin + ci) * ksz + r];
  }
  __syncthreads();
  for (int i = threadIdx.x; i < Cout * per; i += blockDim.x) {
    const int co = i / per, r = i - co * per;
    const size_t gi = ((size_t)b * Cout + co) * per + r;
    float g = dy[gi];
    if (RELU && !(y[gi] > 0.f)) g = 0.f;
    int oh, ow;
    if (POOL) {
      const int ph = r / PW, pw = r - ph * PW, am = idx[gi];
      oh = 2 * ph + (am >> 1);
      ow = 2 * pw + (am & 1);
    } else {
      oh = r / OW;
      ow = r - oh * OW;
    }
    dz[co * zplane + (oh + KH - 1) * ZW + (ow + KW - 1)] = g;
  }
  __syncthreads();
  for (int o = threadIdx.x; o < H * W; o += blockDim.x) {
    const int h = o / W, wq = o - h * W;
    // dz index for tap (kh,kw): row (h - kh) + (KH-1), col (wq - kw) + (KW-1)
    const float* base = dz + (h + KH - 1) * ZW + (wq + KW - 1);
    float a = 0.f;
    for (int co = 0; co < Cout; ++co) {
      const float* dzc = base + co * zplane;
      const float* wk = ws + co * ksz;
      if (KS > 0) {
#pragma unroll
        for (int kh = 0; kh < (KS > 0 ? KS : 1); ++kh)
#pragma unroll
          for (int kw = 0; kw < (KS > 0 ? KS : 1); ++kw) a = fmaf(dzc[-kh * ZW - kw], wk[kh * KS + kw], a);
      } else {
        for (int kh = 0; kh < KH; ++kh)
          for (int kw = 0; kw < KW; ++kw) a = fmaf(dzc[-kh * ZW - kw], wk[kh * KW + kw], a);
      }
    }
    dx[((size_t)b * Cin + ci) * H * W + o] = a;
  }
}

// =====================================================================================================
// small linear: y = act(x W^T + b), fp32.  One warp produces 4 batch rows of one output feature.
// =====================================================================================================
constexpr int kLinRows = 4;

template <typename T>
__global__ void __launch_bounds__(256)
linear_small_fwd_kernel(const T* __restrict__ x, const T* __restrict__ w, const T* __restrict__ bias,
                        T* __restrict__ y, int B, int K, int N, int act) {
  const int lane = threadIdx.x & 31;
  const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int row_tiles = ceil_div(B, kLinRows);
  if (gw >= N * row_tiles) return;
  const int n = gw % N, b0 = (gw / N) * kLinRows;
  const T* wr = w + (size_t)n * K;
  float acc[kLinRows] = {0.f, 0.f, 0.f, 0.f};
  for (int k = lane; k < K; k += 32) {
    const float wv = to_float(wr[k]);
#pragma unroll
    for (int r = 0; r < kLinRows; ++r)
      if (b0 + r < B) acc[r] = fmaf(wv, to_float(x[(size_t)(b0 + r) * K + k]), acc[r]);
  }
#pragma unroll
  for (int r = 0; r < kLinRows; ++r) acc[r] = warp_sum(acc[r]);
  if (lane < kLinRows && b0 + lane < B) {
    float v = acc[0];
#pragma unroll
    for (int r = 1; r < kLinRows; ++r) if (lane == r) v = acc[r];
    v += bias ? to_float(bias[n]) : 0.f;
    y[(size_t)(b0 + lane) * N + n] = from_float<T>(act_fwd_rt(act, v));
  }
}

__device__ __forceinline__ float act_grad_from_y(int act, float yv) {
  switch (act) {
    case ACT_RELU: return yv > 0.f ? 1.f : 0.f;
    case ACT_TANH: return 1.f - yv * yv;
    default: return 1.f;
  }
}

// blocks [0, nbx): dx[b,k] = sum_n dz[b,n] w[n,k]      (skipped when dx == nullptr)
// blocks [nbx, ..): dw[n,k] = sum_b dz[b,n] x[b,k];  db[n] = sum_b dz[b,n]
// T: activation / weight dtype, G: gradient-sink dtype (the flat gradient arena's)
template <typename T, typename G>
__global__ void __launch_bounds__(256)
linear_small_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ y, const T* __restrict__ x,
                        const T* __restrict__ w, T* __restrict__ dx, G* __restrict__ dw,
                        G* __restrict__ db, int B, int K, int N, int act, int nbx, int accumulate) {
  if ((int)blockIdx.x < nbx) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * K) return;
    const int b = i / K, k = i - b * K;
    float a = 0.f;
    for (int n = 0; n < N; ++n) {
      const float dz = to_float(dy[(size_t)b * N + n]) * act_grad_from_y(act, to_float(y[(size_t)b * N + n]));
      a = fmaf(dz, to_float(w[(size_t)n * K + k]), a);
    }
    dx[i] = from_float<T>(a);
  } else {
    const int i = (blockIdx.x - nbx) * blockDim.x + threadIdx.x;
    if (i >= N * K) return;
    const int n = i / K, k = i - n * K;
    float a = 0.f, s = 0.f;
    for (int b = 0; b < B; ++b) {
      const float dz = to_float(dy[(size_t)b * N + n]) * act_grad_from_y(act, to_float(y[(size_t)b * N + n]));
      a = fmaf(dz, to_float(x[(size_t)b * K + k]), a);
      s += dz;
    }
    dw[i] = from_float<G>((accumulate ? to_float(dw[i]) : 0.f) + a);
    if (k == 0 && db) db[n] = from_float<G>((accumulate ? to_float(db[n]) : 0.f) + s);
  }
}

template <typename K>
static int set_smem(K kernel, size_t bytes) {
  if (bytes > 227 * 1024) {
    set_error("direct conv path needs %zu B of shared memory (> 227 KB); use the implicit-GEMM path", bytes);
    return 1;
  }
  if (bytes > 48 * 1024) MLT_CHECK_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  return 0;
}

}  // namespace mlt

using namespace mlt;

#define DISPATCH_RP(relu, pool, ...)                                              \
  if (relu && pool) { constexpr bool R_ = true, P_ = true; __VA_ARGS__; }         \
  else if (relu) { constexpr bool R_ = true, P_ = false; __VA_ARGS__; }           \
  else if (pool) { constexpr bool R_ = false, P_ = true; __VA_ARGS__; }           \
  else { constexpr bool R_ = false, P_ = false; __VA_ARGS__; }

extern "C" int mlt_conv2d_direct_fwd(const void* x, const void* w, const void* bias, void* y, void* pool_idx, int B, int Cin,
                                     int H, int W, int Cout, int KH, int KW, int relu, int pool, void* stream) {
  MLT_REQUIRE(H >= KH && W >= KW, "mlt_conv2d_direct_fwd: kernel larger than input");
  MLT_REQUIRE(!pool || pool_idx != nullptr, "mlt_conv2d_direct_fwd: pool needs an index buffer");
  const int wsz = Cin * KH * KW;
  int co_per = Cout;                                   // split output channels until the grid fills the chip
  while (co_per > 1 && B * ceil_div(Cout, co_per) < kNumSMs) co_per = ceil_div(co_per, 2);
  const size_t smem = ((size_t)Cin * H * W + (size_t)co_per * wsz) * sizeof(float);
  dim3 grid(B, ceil_div(Cout, co_per));
  DISPATCH_RP(relu, pool, {
    auto k = conv2d_direct_fwd_kernel<R_, P_>;
    if (set_smem(k, smem)) return 1;
    k<<<grid, 256, smem, (cudaStream_t)stream>>>((const float*)x, (const float*)w, (const float*)bias, (float*)y,
                                                (unsigned char*)pool_idx, Cin, H, W, Cout, KH, KW, co_per);
  });
  MLT_LAUNCH_CHECK();
  return 0;
}

// scratch layout: [tickets: round_up(Cout*4, 256) bytes][partials: nchunk * Cout * (taps+1) floats]
static void conv_bwd_chunks(int B, int Cout, int* nchunk, int* chunk) {
  int n = (2 * kNumSMs) / (Cout > 0 ? Cout : 1);     // aim for ~2 CTAs per SM
  if (n > B) n = B;
  if (n < 1) n = 1;
  *chunk = ceil_div(B, n);
  *nchunk = ceil_div(B, *chunk);
}
static long long conv_bwd_ticket_bytes(int Cout) { return ((long long)Cout * 4 + 255) / 256 * 256; }

extern "C" long long mlt_conv2d_direct_bwd_scratch_bytes(int B, int Cin, int Cout, int KH, int KW) {
  int nchunk, chunk;
  conv_bwd_chunks(B, Cout, &nchunk, &chunk);
  const long long slots = (long long)Cin * KH * KW + 1;
  return conv_bwd_ticket_bytes(Cout) + (long long)nchunk * Cout * slots * 4;
}

extern "C" int mlt_conv2d_direct_bwd(const void* x, const void* w, const void* dy, const void* y, const void* pool_idx,
                                     void* dx, void* dw, void* db, void* scratch, long long scratch_bytes, int B, int Cin,
                                     int H, int W, int Cout, int KH, int KW, int relu, int pool, int accumulate,
                                     void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  const int nW = Cin * KH * KW;
  MLT_REQUIRE(nW + 1 <= 1024, "mlt_conv2d_direct_bwd: %d filter taps per channel exceeds the direct path (1023)", nW);
  const int OH = H - KH + 1, OW = W - KW + 1;
  const int per = pool ? (OH / 2) * (OW / 2) : OH * OW;
  int nchunk, chunk;
  conv_bwd_chunks(B, Cout, &nchunk, &chunk);
  const long long need = mlt_conv2d_direct_bwd_scratch_bytes(B, Cin, Cout, KH, KW);
  MLT_REQUIRE(nchunk == 1 || (scratch != nullptr && scratch_bytes >= need),
              "mlt_conv2d_direct_bwd: scratch too small (%lld < %lld bytes)", scratch_bytes, need);
  unsigned int* tickets = (unsigned int*)scratch;
  float* partial = scratch ? (float*)((char*)scratch + conv_bwd_ticket_bytes(Cout)) : nullptr;
  {
    const size_t smem = ((size_t)Cin * H * W + 2 * (size_t)per) * sizeof(float);
    dim3 grid(Cout, nchunk);
    DISPATCH_RP(relu, pool, {
      auto k = conv2d_direct_wgrad_kernel<R_, P_>;
      if (set_smem(k, smem)) return 1;
      k<<<grid, 256, smem, st>>>((const float*)x, (const float*)dy, (const float*)y, (const unsigned char*)pool_idx,
                                 (float*)dw, (float*)db, partial, tickets, B, Cin, H, W, Cout, KH, KW, chunk, accumulate);
    });
    MLT_LAUNCH_CHECK();
  }
  if (dx != nullptr) {
    const size_t smem = ((size_t)Cout * (OH + 2 * (KH - 1)) * (OW + 2 * (KW - 1)) + (size_t)Cout * KH * KW) * sizeof(float);
    dim3 grid(B, Cin);
    DISPATCH_RP(relu, pool, {
      if (KH == 5 && KW == 5) {
        auto k = conv2d_direct_dgrad_kernel<R_, P_, 5>;
        if (set_smem(k, smem)) return 1;
        k<<<grid, 256, smem, st>>>((const float*)dy, (const float*)y, (const unsigned char*)pool_idx, (const float*)w,
                                   (float*)dx, Cin, H, W, Cout, KH, KW);
      } else if (KH == 3 && KW == 3) {
        auto k = conv2d_direct_dgrad_kernel<R_, P_, 3>;
        if (set_smem(k, smem)) return 1;
        k<<<grid, 256, smem, st>>>((const float*)dy, (const float*)y, (const unsigned char*)pool_idx, (const float*)w,
                                   (float*)dx, Cin, H, W, Cout, KH, KW);
      } else {
        auto k = conv2d_direct_dgrad_kernel<R_, P_, 0>;
        if (set_smem(k, smem)) return 1;
        k<<<grid, 256, smem, st>>>((const float*)dy, (const float*)y, (const unsigned char*)pool_idx, (const float*)w,
                                   (float*)dx, Cin, H, W, Cout, KH, KW);
      }
    });
    MLT_LAUNCH_CHECK();
  }
  return 0;
}

// dtype: 0 = fp32 tensors, 1 = bf16 tensors (weights, activations); grads_bf16 tells the dtype of dw / db.
extern "C" int mlt_linear_small_fwd(const void* x, const void* w, const void* bias, void* y, int B, int K, int N, int act,
                                    int is_bf16, void* stream) {
  MLT_REQUIRE(act == ACT_NONE || act == ACT_RELU || act == ACT_TANH, "mlt_linear_small_fwd: unsupported activation %d", act);
  const long long warps = (long long)N * ceil_div(B, kLinRows);
  const int grid = (int)ceil_div64(warps * 32, 256);
  cudaStream_t st = (cudaStream_t)stream;
  if (is_bf16)
    linear_small_fwd_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)w,
                                                                 (const __nv_bfloat16*)bias, (__nv_bfloat16*)y, B, K, N, act);
  else
    linear_small_fwd_kernel<float><<<grid, 256, 0, st>>>((const float*)x, (const float*)w, (const float*)bias, (float*)y, B, K, N, act);
  MLT_LAUNCH_CHECK();
  return 0;
}

extern "C" int mlt_linear_small_bwd(const void* dy, const void* y, const void* x, const void* w, void* dx, void* dw, void* db,
                                    int B, int K, int N, int act, int accumulate, int is_bf16, int grads_bf16, void* stream) {
  MLT_REQUIRE(act == ACT_NONE || act == ACT_RELU || act == ACT_TANH, "mlt_linear_small_bwd: unsupported activation %d", act);
  MLT_REQUIRE(is_bf16 || !grads_bf16, "mlt_linear_small_bwd: fp32 tensors with bf16 gradient sinks is not a supported mix");
  const int nbx = dx ? ceil_div(B * K, 256) : 0;
  const int nbw = ceil_div(N * K, 256);
  cudaStream_t st = (cudaStream_t)stream;
#define LSB(T, G)                                                                                                          \
  linear_small_bwd_kernel<T, G><<<nbx + nbw, 256, 0, st>>>((const T*)dy, (const T*)y, (const T*)x, (const T*)w, (T*)dx, (G*)dw, \
                                                           (G*)db, B, K, N, act, nbx, accumulate)
  if (!is_bf16) LSB(float, float);
  else if (grads_bf16) LSB(__nv_bfloat16, __nv_bfloat16);
  else LSB(__nv_bfloat16, float);
#undef LSB
  MLT_LAUNCH_CHECK();
  return 0;
}

// =====================================================================================================
// Allocation-free variant of mlt_conv2d_direct_bwd.
//
// Who needs it: callers that are not allowed to obtain device memory at the point where they run this
// backward pass.  Examples: a plug-in living inside a host application that handed every free byte of the GPU
// to its own memory pool at start-up, so a later cudaMalloc fails or is forbidden by the host's contract;
// and latency-critical loops that cannot tolerate the device-wide synchronisation cudaMalloc/cudaFree imply.
// Such callers still own their inputs and outputs, but have nowhere to get the cross-CTA reduction scratch
// (per-channel tickets + per-chunk partial sums) that the normal entry point expects from them.
//
// Here that scratch is a fixed array in the library's device image.  Consequences worth knowing:
//   * It costs MLT_CONV2D_BWD_NOALLOC_SCRATCH_BYTES (256 KiB) on EVERY GPU that loads this library, from load
//     until the process exits, whether or not this function is ever called.
//   * Problems whose scratch need (mlt_conv2d_direct_bwd_scratch_bytes) exceeds that size are rejected with an
//     error; use mlt_conv2d_direct_bwd with caller-provided scratch for those.
//   * The first call on each device looks up the array's device address (cudaGetSymbolAddress, which may load
//     the module).  Make that first call outside any stream capture.  Later calls on the device only enqueue
//     a small memset and the kernels.
//   * All calls on a device share the one array and it keeps whatever the previous call left behind.  The
//     ticket words, which must be zero when the kernel starts, are cleared on `stream` at the start of every
//     call; the partial sums are fully overwritten before they are read.  Two calls whose GPU work overlaps
//     in time (different streams, or different host threads on the same device) race on the array and produce
//     garbage: serialise them on one stream or with events.
// It performs no cudaMalloc / cudaFree and calls no allocating library routine.
// =====================================================================================================
#include <mutex>

static __device__ __align__(256) unsigned char g_conv2d_bwd_noalloc_scratch[MLT_CONV2D_BWD_NOALLOC_SCRATCH_BYTES];

namespace {
constexpr int kNoAllocMaxDevices = 64;
std::mutex g_noalloc_mu;
void* g_noalloc_scratch_addr[kNoAllocMaxDevices] = {};   // device address of the array, per device ordinal
}  // namespace

extern "C" int mlt_conv2d_direct_bwd_noalloc(const void* x, const void* w, const void* dy, const void* y,
                                             const void* pool_idx, void* dx, void* dw, void* db, int B, int Cin, int H,
                                             int W, int Cout, int KH, int KW, int relu, int pool, int accumulate,
                                             void* stream) {
  MLT_REQUIRE(B > 0 && Cin > 0 && Cout > 0 && KH > 0 && KW > 0, "mlt_conv2d_direct_bwd_noalloc: empty problem");
  const long long need = mlt_conv2d_direct_bwd_scratch_bytes(B, Cin, Cout, KH, KW);
  MLT_REQUIRE(need <= (long long)MLT_CONV2D_BWD_NOALLOC_SCRATCH_BYTES,
              "mlt_conv2d_direct_bwd_noalloc: this problem needs %lld bytes of reduction scratch but the built-in buffer "
              "holds %d; call mlt_conv2d_direct_bwd with caller-provided scratch instead",
              need, (int)MLT_CONV2D_BWD_NOALLOC_SCRATCH_BYTES);
  int dev = -1;
  MLT_CHECK_CUDA(cudaGetDevice(&dev));
  MLT_REQUIRE(dev >= 0 && dev < kNoAllocMaxDevices, "mlt_conv2d_direct_bwd_noalloc: device ordinal %d out of range", dev);
  void* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (g_noalloc_scratch_addr[dev] == nullptr) {
      // In host code the array's name is only a handle; the kernels need its address on the CURRENT device.
      void* addr = nullptr;
      MLT_CHECK_CUDA(cudaGetSymbolAddress(&addr, g_conv2d_bwd_noalloc_scratch));
      MLT_REQUIRE(addr != nullptr, "mlt_conv2d_direct_bwd_noalloc: cudaGetSymbolAddress returned a null address");
      g_noalloc_scratch_addr[dev] = addr;
    }
    scratch = g_noalloc_scratch_addr[dev];
  }
  // The array still holds the previous call's data: tickets must start at zero.
  MLT_CHECK_CUDA(cudaMemsetAsync(scratch, 0, (size_t)conv_bwd_ticket_bytes(Cout), (cudaStream_t)stream));
  return mlt_conv2d_direct_bwd(x, w, dy, y, pool_idx, dx, dw, db, scratch, (long long)MLT_CONV2D_BWD_NOALLOC_SCRATCH_BYTES,
                               B, Cin, H, W, Cout, KH, KW, relu, pool, accumulate, stream);
}

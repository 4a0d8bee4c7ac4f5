// Bandwidth-bound helpers around the GEMMs: activation backward, bias gradient (column sum),
// dtype casts.  Everything is vectorised to 16-byte accesses; the roofline is bytes / HBM bandwidth.
#include "common.cuh"
#include "colreduce.cuh"
#include "../include/mlt_b200.h"

namespace mlt {

__device__ __forceinline__ float act_grad(int act, int from_pre, float s) {
  // s is the pre-activation z (from_pre) or the activation output y
  switch (act) {
    case ACT_RELU: return s > 0.f ? 1.f : 0.f;
    case ACT_TANH: { const float t = from_pre ? tanhf(s) : s; return 1.f - t * t; }
    case ACT_GELU: return gelu_bwd(s);      // needs the pre-activation
    default: return 1.f;
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
act_backward_kernel(const T* __restrict__ dy, const T* __restrict__ saved, T* __restrict__ dz, long long n, int act, int from_pre) {
  constexpr int V = 16 / sizeof(T);
  const long long nv = n / V;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += stride) {
    const uint4 a = reinterpret_cast<const uint4*>(dy)[i];
    const uint4 s = reinterpret_cast<const uint4*>(saved)[i];
    uint4 o;
    const T* ap = reinterpret_cast<const T*>(&a);
    const T* sp = reinterpret_cast<const T*>(&s);
    T* op = reinterpret_cast<T*>(&o);
#pragma unroll
    for (int j = 0; j < V; ++j) op[j] = from_float<T>(to_float(ap[j]) * act_grad(act, from_pre, to_float(sp[j])));
    reinterpret_cast<uint4*>(dz)[i] = o;
  }
  for (long long i = nv * V + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    dz[i] = from_float<T>(to_float(dy[i]) * act_grad(act, from_pre, to_float(saved[i])));
}

// Column sums of x[M][N] (row pitch ld).  grid = (ceil(N/64), R).  A warp reads 64 consecutive columns of a row
// (2 per lane), warps stride over the CTA's rows; CTAs of one column tile combine through scratch
// partial[R][N] + one ticket per column tile; the last one writes the result (deterministic order).
template <typename T, typename O>
__global__ void __launch_bounds__(256)
colsum_kernel(const T* __restrict__ x, O* __restrict__ out, float* __restrict__ partial, unsigned int* __restrict__ tickets,
              int M, int N, long long ld, int accumulate) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c0 = blockIdx.x * 64 + lane * 2;
  const int R = gridDim.y;
  const int rows_per = ceil_div(M, R);
  const int r_lo = blockIdx.y * rows_per, r_hi = min(M, r_lo + rows_per);
  float s0 = 0.f, s1 = 0.f;
  if (c0 < N) {
    const bool pair = c0 + 1 < N;
    for (int r = r_lo + warp; r < r_hi; r += 8) {
      const T* p = x + (size_t)r * ld + c0;
      s0 += to_float(p[0]);
      if (pair) s1 += to_float(p[1]);
    }
  }
  __shared__ float sh[8][64];
  __shared__ bool is_last;
  sh[warp][lane * 2] = s0;
  sh[warp][lane * 2 + 1] = s1;
  __syncthreads();
  if (threadIdx.x < 64) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += sh[w][threadIdx.x];
    const int c = blockIdx.x * 64 + threadIdx.x;
    if (c < N) {
      if (R == 1) out[c] = from_float<O>((accumulate ? to_float(out[c]) : 0.f) + t);
      else partial[(size_t)blockIdx.y * N + c] = t;
    }
  }
  if (R == 1) return;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int tk = atomicAdd(&tickets[blockIdx.x], 1u);
    is_last = tk == (unsigned int)R - 1;
    if (is_last) tickets[blockIdx.x] = 0u;
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  if (threadIdx.x < 64) {
    const int c = blockIdx.x * 64 + threadIdx.x;
    if (c < N) {
      float t = 0.f;
      for (int k = 0; k < R; ++k) t += __ldcg(partial + (size_t)k * N + c);
      out[c] = from_float<O>((accumulate ? to_float(out[c]) : 0.f) + t);
    }
  }
}

// Vectorised variant: every lane owns 16 bytes of a row (8 bf16 / 4 fp32 columns), rows are unrolled 4-deep so
// several loads are in flight.  grid = (ceil(N / (32*V)), R).  Same scratch protocol as colsum_kernel.
template <typename T> struct ColVec;
template <> struct ColVec<__nv_bfloat16> {
  static constexpr int V = 8;
  __device__ static void load(const __nv_bfloat16* p, float (&v)[8]) {
    const uint4 raw = *reinterpret_cast<const uint4*>(p);
    const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[j]));
      v[2 * j] = f.x; v[2 * j + 1] = f.y;
    }
  }
};
template <> struct ColVec<float> {
  static constexpr int V = 4;
  __device__ static void load(const float* p, float (&v)[8]) {
    const float4 a = *reinterpret_cast<const float4*>(p);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = v[5] = v[6] = v[7] = 0.f;
  }
};

template <typename T, typename O>
__global__ void __launch_bounds__(256)
colsum_vec_kernel(const T* __restrict__ x, O* __restrict__ out, float* __restrict__ partial, unsigned int* __restrict__ tickets,
                  int M, int N, long long ld, int accumulate) {
  constexpr int V = ColVec<T>::V, W = 32 * V;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c0 = blockIdx.x * W + lane * V;
  const int R = gridDim.y;
  const int rows_per = ceil_div(M, R);
  const int r_lo = blockIdx.y * rows_per, r_hi = min(M, r_lo + rows_per);
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (c0 < N) {
    int r = r_lo + warp;
    for (; r + 56 < r_hi; r += 64) {          // 8 independent 16-byte loads per lane in flight
      float a[8][8];
#pragma unroll
      for (int u = 0; u < 8; ++u) ColVec<T>::load(x + (size_t)(r + 8 * u) * ld + c0, a[u]);
#pragma unroll
      for (int u = 0; u < 8; ++u)
#pragma unroll
        for (int j = 0; j < V; ++j) acc[j] += a[u][j];
    }
    for (; r < r_hi; r += 8) {
      float a[8];
      ColVec<T>::load(x + (size_t)r * ld + c0, a);
#pragma unroll
      for (int j = 0; j < V; ++j) acc[j] += a[j];
    }
  }
  __shared__ float sh[8][32 * 8];
  __shared__ bool is_last;
#pragma unroll
  for (int j = 0; j < V; ++j) sh[warp][lane * V + j] = acc[j];
  __syncthreads();
  if (threadIdx.x < W) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += sh[w][threadIdx.x];
    const int c = blockIdx.x * W + threadIdx.x;
    if (c < N) {
      if (R == 1) out[c] = from_float<O>((accumulate ? to_float(out[c]) : 0.f) + t);
      else partial[(size_t)blockIdx.y * N + c] = t;
    }
  }
  if (R == 1) return;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int tk = atomicAdd(&tickets[blockIdx.x], 1u);
    is_last = tk == (unsigned int)R - 1;
    if (is_last) tickets[blockIdx.x] = 0u;
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  if (threadIdx.x < W) {
    const int c = blockIdx.x * W + threadIdx.x;
    if (c < N) {
      float t = 0.f;
      for (int k = 0; k < R; ++k) t += __ldcg(partial + (size_t)k * N + c);
      out[c] = from_float<O>((accumulate ? to_float(out[c]) : 0.f) + t);
    }
  }
}

// Bias gradients of the big GEMM layers: contiguous bf16 [M, N] with N a multiple of 8, read as a flat stream of 16-byte
// groups and folded through shared memory + fp32 atomics (colreduce.cuh) - the same scheme as the BatchNorm reductions.
template <typename O>
__global__ void __launch_bounds__(kBnThreads, 3)
colsum_flat_kernel(const __nv_bfloat16* __restrict__ x, O* __restrict__ out, float* __restrict__ acc, unsigned int* __restrict__ tickets,
                   int M, int N, int accumulate) {
  __shared__ float tot[2 * kBnGroups * 8];
  struct {                                       // eight loads in flight per thread (see colreduce.cuh, LOADFIRST)
    const __nv_bfloat16* x;
    __device__ __forceinline__ uint4 load(size_t off) const { return ld8raw(x + off); }
    __device__ __forceinline__ void pin(const uint4& raw) const { asm volatile("" ::"r"(raw.x) : "memory"); }
    __device__ __forceinline__ void use(const uint4& raw, float (&a)[8], float (&b)[8]) const {
      float v[8];
      cvt8(raw, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) a[j] += v[j];
      (void)b;
    }
  } f{x};
  if (!col_reduce_atomic<false, 8, true>(f, acc, tickets, M, N, tot)) return;
  const int g_lo = blockIdx.x * kBnGroups;
  const int nch = min(kBnGroups, (N >> 3) - g_lo) * 8;
  for (int ch = threadIdx.x; ch < nch; ch += kBnThreads) {
    const int c = g_lo * 8 + ch;
    out[c] = from_float<O>((accumulate ? to_float(out[c]) : 0.f) + tot[ch]);
  }
}

template <typename S, typename D>
__global__ void __launch_bounds__(256) cast_kernel(const S* __restrict__ src, D* __restrict__ dst, long long n) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) dst[i] = from_float<D>(to_float(src[i]));
}

static int ew_grid(long long work_items) {
  long long b = ceil_div64(work_items, 256);
  const long long cap = (long long)kNumSMs * 8;
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

static int colsum_splits(int M, int N) {
  const int tiles = ceil_div(N, 64);
  int R = (2 * kNumSMs) / tiles;
  if (R > ceil_div(M, 64)) R = ceil_div(M, 64);
  return R < 1 ? 1 : R;
}

}  // namespace mlt

using namespace mlt;

extern "C" int mlt_act_backward(const void* dy, const void* saved, void* dz, long long n, int act, int from_pre, int is_bf16,
                                void* stream) {
  MLT_REQUIRE(act != ACT_GELU || from_pre, "mlt_act_backward: GELU backward needs the pre-activation");
  if (n <= 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  if (is_bf16)
    act_backward_kernel<__nv_bfloat16><<<ew_grid(n / 8 + 1), 256, 0, st>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)saved,
                                                                         (__nv_bfloat16*)dz, n, act, from_pre);
  else
    act_backward_kernel<float><<<ew_grid(n / 4 + 1), 256, 0, st>>>((const float*)dy, (const float*)saved, (float*)dz, n, act, from_pre);
  MLT_LAUNCH_CHECK();
  return 0;
}

extern "C" long long mlt_colsum_scratch_bytes(int M, int N) {
  // sized for the narrowest column tile (64) so either kernel variant fits
  const int tiles = ceil_div(N, 64);
  return ((long long)tiles * 4 + 255) / 256 * 256 + (long long)(4 * kNumSMs) * N * 4;
}

// out[N] = sum over rows of x[M][N];  x bf16 or fp32, out bf16 or fp32.  scratch: mlt_colsum_scratch_bytes, zeroed once.
extern "C" int mlt_colsum(const void* x, int x_bf16, void* out, int out_bf16, void* scratch, long long scratch_bytes, int M,
                          int N, long long ld, int accumulate, void* stream) {
  MLT_REQUIRE(M > 0 && N > 0, "mlt_colsum: empty problem");
  const int V = x_bf16 ? 8 : 4;
  const bool vec = (N % V == 0) && (ld % V == 0) && (((uintptr_t)x) % 16 == 0);
  const int tile_w = vec ? 32 * V : 64;
  const int tiles = ceil_div(N, tile_w);
  int R = ((vec ? 4 : 2) * kNumSMs) / tiles;
  if (R > ceil_div(M, 64)) R = ceil_div(M, 64);
  if (R < 1) R = 1;
  MLT_REQUIRE(R == 1 || (scratch && scratch_bytes >= mlt_colsum_scratch_bytes(M, N)), "mlt_colsum: scratch too small");
  unsigned int* tickets = (unsigned int*)scratch;
  float* partial = scratch ? (float*)((char*)scratch + ((long long)ceil_div(N, 64) * 4 + 255) / 256 * 256) : nullptr;
  dim3 grid(tiles, R);
  cudaStream_t st = (cudaStream_t)stream;
#define CS(K, T, O) K<T, O><<<grid, 256, 0, st>>>((const T*)x, (O*)out, partial, tickets, M, N, ld, accumulate)
  if (vec) {
    if (x_bf16 && out_bf16) CS(colsum_vec_kernel, __nv_bfloat16, __nv_bfloat16);
    else if (x_bf16) CS(colsum_vec_kernel, __nv_bfloat16, float);
    else if (out_bf16) CS(colsum_vec_kernel, float, __nv_bfloat16);
    else CS(colsum_vec_kernel, float, float);
  } else {
    if (x_bf16 && out_bf16) CS(colsum_kernel, __nv_bfloat16, __nv_bfloat16);
    else if (x_bf16) CS(colsum_kernel, __nv_bfloat16, float);
    else if (out_bf16) CS(colsum_kernel, float, __nv_bfloat16);
    else CS(colsum_kernel, float, float);
  }
#undef CS
  MLT_LAUNCH_CHECK();
  return 0;
}

// Fast path of mlt_colsum for contiguous bf16 inputs with N % 8 == 0.  scratch (zeroed once, left zero by every call):
// [tickets 256 B][N fp32 accumulators].
extern "C" long long mlt_colsum_flat_scratch_bytes(int N) { return 256 + (long long)N * 4; }

extern "C" int mlt_colsum_flat(const void* x, void* out, int out_bf16, void* scratch, long long scratch_bytes, int M, int N,
                               int accumulate, void* stream) {
  MLT_REQUIRE(M > 0 && N > 0 && N % 8 == 0 && N / 8 <= 64 * kBnGroups, "mlt_colsum_flat: N (%d) must be a multiple of 8 up to 16384", N);
  MLT_REQUIRE(scratch && scratch_bytes >= mlt_colsum_flat_scratch_bytes(N), "mlt_colsum_flat: scratch too small");
  MLT_REQUIRE(((uintptr_t)x % 16) == 0, "mlt_colsum_flat: input must be 16-byte aligned");
  const int C8 = N / 8;
  const int tiles = ceil_div(C8, kBnGroups);
  const int rpi = kBnThreads / (C8 < kBnGroups ? C8 : kBnGroups);
  int R = (4 * kNumSMs) / tiles;
  const int cap = ceil_div(M, rpi * 8);
  if (R > cap) R = cap;
  if (R < 1) R = 1;
  unsigned int* tickets = (unsigned int*)scratch;
  float* acc = (float*)((char*)scratch + 256);
  dim3 grid(tiles, R);
  cudaStream_t st = (cudaStream_t)stream;
  if (out_bf16) colsum_flat_kernel<__nv_bfloat16><<<grid, kBnThreads, 0, st>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)out, acc, tickets, M, N, accumulate);
  else colsum_flat_kernel<float><<<grid, kBnThreads, 0, st>>>((const __nv_bfloat16*)x, (float*)out, acc, tickets, M, N, accumulate);
  MLT_LAUNCH_CHECK();
  return 0;
}

extern "C" int mlt_cast(const void* src, int src_bf16, void* dst, int dst_bf16, long long n, void* stream) {
  if (n <= 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  const int g = ew_grid(n);
  if (src_bf16 && !dst_bf16) cast_kernel<__nv_bfloat16, float><<<g, 256, 0, st>>>((const __nv_bfloat16*)src, (float*)dst, n);
  else if (!src_bf16 && dst_bf16) cast_kernel<float, __nv_bfloat16><<<g, 256, 0, st>>>((const float*)src, (__nv_bfloat16*)dst, n);
  else MLT_CHECK_CUDA(cudaMemcpyAsync(dst, src, (size_t)n * (src_bf16 ? 2 : 4), cudaMemcpyDeviceToDevice, st));
  MLT_LAUNCH_CHECK();
  return 0;
}

// ---- fp32 operands on the bf16 tensor cores: three-term split (weak #11 of the round-1 review: no cuBLAS for fp32 layers) ----
// x = hi + lo with hi = bf16(x), lo = bf16(x - hi).  A product of two fp32 matrices is recovered to ~2^-16 relative from bf16
// MMAs with fp32 accumulation as  A_hi B_hi + A_hi B_lo + A_lo B_hi  (the dropped lo*lo term is ~2^-16 of the result) - more
// accurate than TF32 (2^-11) at three bf16 products.  The three products become ONE GEMM by tripling the reduction dimension:
//   [A_hi | A_hi | A_lo] x [B_hi | B_lo | B_hi]^T          (pattern 0 = hi,hi,lo for one operand; 1 = hi,lo,hi for the other)
// This kernel writes such a tripled operand from an fp32 [R, C] matrix: mode 0 concatenates along the columns (out [R, 3C], the
// K-major operand), mode 1 stacks along the rows (out [3R, C], the MN-major operand).
namespace mlt {
__global__ void __launch_bounds__(256)
split3_kernel(const float* __restrict__ x, long long ldx, __nv_bfloat16* __restrict__ out, long long R, long long C, int mode, int pattern) {
  const long long n = R * C;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / C, c = i - r * C;
    const float v = x[r * ldx + c];
    const __nv_bfloat16 hi = __float2bfloat16_rn(v);
    const __nv_bfloat16 lo = __float2bfloat16_rn(v - __bfloat162float(hi));
    const __nv_bfloat16 t0 = hi, t1 = pattern == 0 ? hi : lo, t2 = pattern == 0 ? lo : hi;
    if (mode == 0) {
      __nv_bfloat16* o = out + r * 3 * C + c;
      o[0] = t0; o[C] = t1; o[2 * C] = t2;
    } else {
      out[i] = t0; out[n + i] = t1; out[2 * n + i] = t2;
    }
  }
}
}  // namespace mlt

extern "C" int mlt_split3_bf16(const void* x_f32, long long ldx, void* out_bf16, long long R, long long C, int mode, int pattern, void* stream) {
  MLT_REQUIRE(R > 0 && C > 0 && ldx >= C, "mlt_split3_bf16: bad shape");
  MLT_REQUIRE((mode == 0 || mode == 1) && (pattern == 0 || pattern == 1), "mlt_split3_bf16: mode / pattern are 0 or 1");
  const int g = ew_grid(R * C);
  mlt::split3_kernel<<<g, 256, 0, (cudaStream_t)stream>>>((const float*)x_f32, ldx, (__nv_bfloat16*)out_bf16, R, C, mode, pattern);
  MLT_LAUNCH_CHECK();
  return 0;
}

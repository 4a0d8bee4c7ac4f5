// BERT embedding lookup: word + position + token-type gather-sum (forward) and its scatter / reductions (backward)
// (SURVEY §8 K14).  The LayerNorm that follows is the generic kernel in norm.cu.
#include "common.cuh"
#include "../include/mlt_b200.h"

namespace mlt {

template <typename P>
__global__ void __launch_bounds__(256)
embedding_sum_fwd_kernel(const long long* __restrict__ ids, const P* __restrict__ word, const P* __restrict__ pos,
                         const P* __restrict__ type, __nv_bfloat16* __restrict__ out, int T, int S, int H, long long V) {
  const int lane = threadIdx.x & 31;
  const int t = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (t >= T) return;
  long long id = ids[t];
  if (V > 0) id = id < 0 ? 0 : (id >= V ? V - 1 : id);          // an id outside the table must not read (or, in backward, write) out of bounds
  const P* w = word + (size_t)id * H;
  const P* p = pos + (size_t)(t % S) * H;
  for (int c = lane * 2; c < H; c += 64) {
    const float a = to_float(w[c]) + to_float(p[c]) + to_float(type[c]);
    const float b = to_float(w[c + 1]) + to_float(p[c + 1]) + to_float(type[c + 1]);
    *reinterpret_cast<uint32_t*>(out + (size_t)t * H + c) = pack_bf16(a, b);
  }
}

__device__ __forceinline__ void atomic_add2(float* dst, float a, float b) { atomicAdd(dst, a); atomicAdd(dst + 1, b); }
__device__ __forceinline__ void atomic_add2(__nv_bfloat16* dst, float a, float b) {
  atomicAdd(reinterpret_cast<__nv_bfloat162*>(dst), __floats2bfloat162_rn(a, b));
}

// dword[ids[t]] += dh[t]  (word rows are hit by several tokens: atomics; the caller zeroes dword first)
template <typename G>
__global__ void __launch_bounds__(256)
embedding_word_bwd_kernel(const long long* __restrict__ ids, const __nv_bfloat16* __restrict__ dh, G* __restrict__ dword, int T, int H,
                          long long V) {
  const int lane = threadIdx.x & 31;
  const int t = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (t >= T) return;
  long long id = ids[t];
  id = id < 0 ? 0 : (id >= V ? V - 1 : id);                      // same clamp as the forward pass
  G* dst = dword + (size_t)id * H;
  for (int c = lane * 2; c < H; c += 64) {
    const float2 v = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(dh + (size_t)t * H + c));
    atomic_add2(dst + c, v.x, v.y);
  }
}

// dpos[s] = sum_b dh[b*S + s]  (deterministic: one thread per (s, column pair) walks the batch)
template <typename G>
__global__ void __launch_bounds__(256)
embedding_pos_bwd_kernel(const __nv_bfloat16* __restrict__ dh, G* __restrict__ dpos, int B, int S, int H) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;      // over S * H/2
  const int hp = H / 2;
  if (i >= (long long)S * hp) return;
  const int s = (int)(i / hp), c = (int)(i - (long long)s * hp) * 2;
  float a = 0.f, b = 0.f;
  for (int bb = 0; bb < B; ++bb) {
    const float2 v = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(dh + ((size_t)bb * S + s) * H + c));
    a += v.x; b += v.y;
  }
  dpos[(size_t)s * H + c] = from_float<G>(a);
  dpos[(size_t)s * H + c + 1] = from_float<G>(b);
}

}  // namespace mlt

using namespace mlt;

// V: rows of the word table (ids are clamped into it; 0 = trust the ids)
extern "C" int mlt_embedding_sum_fwd_v(const void* ids_i64, const void* word, const void* pos, const void* type, int params_bf16,
                                       void* out_bf16, int T, int S, int H, long long V, void* stream) {
  MLT_REQUIRE(T > 0 && H % 2 == 0, "mlt_embedding_sum_fwd: bad shape (T=%d H=%d)", T, H);
  cudaStream_t st = (cudaStream_t)stream;
  if (params_bf16)
    embedding_sum_fwd_kernel<__nv_bfloat16><<<ceil_div(T, 8), 256, 0, st>>>((const long long*)ids_i64, (const __nv_bfloat16*)word,
        (const __nv_bfloat16*)pos, (const __nv_bfloat16*)type, (__nv_bfloat16*)out_bf16, T, S, H, V);
  else
    embedding_sum_fwd_kernel<float><<<ceil_div(T, 8), 256, 0, st>>>((const long long*)ids_i64, (const float*)word, (const float*)pos,
                                                                     (const float*)type, (__nv_bfloat16*)out_bf16, T, S, H, V);
  MLT_LAUNCH_CHECK();
  return 0;
}

extern "C" int mlt_embedding_sum_fwd(const void* ids_i64, const void* word, const void* pos, const void* type, int params_bf16,
                                     void* out_bf16, int T, int S, int H, void* stream) {
  return mlt_embedding_sum_fwd_v(ids_i64, word, pos, type, params_bf16, out_bf16, T, S, H, 0, stream);
}

// dword must hold V*H elements; it is zeroed here (cudaMemsetAsync) before the scatter.  dpos rows [0, S) are overwritten.
extern "C" int mlt_embedding_bwd(const void* ids_i64, const void* dh_bf16, void* dword, long long V, void* dpos, int grads_bf16,
                                 int B, int S, int H, void* stream) {
  const int T = B * S;
  MLT_REQUIRE(T > 0 && H % 2 == 0, "mlt_embedding_bwd: bad shape");
  cudaStream_t st = (cudaStream_t)stream;
  MLT_CHECK_CUDA(cudaMemsetAsync(dword, 0, (size_t)V * H * (grads_bf16 ? 2 : 4), st));
  const int pos_blocks = (int)ceil_div64((long long)S * (H / 2), 256);
  if (grads_bf16) {
    embedding_word_bwd_kernel<__nv_bfloat16><<<ceil_div(T, 8), 256, 0, st>>>((const long long*)ids_i64, (const __nv_bfloat16*)dh_bf16,
                                                                           (__nv_bfloat16*)dword, T, H, V);
    embedding_pos_bwd_kernel<__nv_bfloat16><<<pos_blocks, 256, 0, st>>>((const __nv_bfloat16*)dh_bf16, (__nv_bfloat16*)dpos, B, S, H);
  } else {
    embedding_word_bwd_kernel<float><<<ceil_div(T, 8), 256, 0, st>>>((const long long*)ids_i64, (const __nv_bfloat16*)dh_bf16,
                                                                   (float*)dword, T, H, V);
    embedding_pos_bwd_kernel<float><<<pos_blocks, 256, 0, st>>>((const __nv_bfloat16*)dh_bf16, (float*)dpos, B, S, H);
  }
  MLT_LAUNCH_CHECK();
  return 0;
}

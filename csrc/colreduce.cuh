// Two-quantity per-column reduction over the rows of an [M, C] matrix (BatchNorm statistics / gradients, LayerNorm
// dgamma / dbeta).  A lane owns 8 columns (16 bytes of bf16), warps stride over rows, CTAs of a column tile combine
// through scratch partials + a ticket; the last CTA to arrive holds the totals.
#pragma once
#include "common.cuh"

namespace mlt {

constexpr int kBnTile = 256;     // channels per CTA column tile (32 lanes x 8)

// Generic two-quantity column reduction skeleton.  F(row, c0, a[8], b[8]) accumulates the row's contribution.
// partial layout: [R][2][C]; tickets: one per column tile.  Returns true in the CTA that must finalise, with the
// totals of its tile left in shared memory (tot[2][kBnTile]).
template <typename F>
__device__ __forceinline__ bool column_reduce2(F f, float* partial, unsigned int* tickets, int M, int C, float (*tot)[kBnTile]) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c0 = blockIdx.x * kBnTile + lane * 8;
  const int R = gridDim.y;
  const int rows_per = ceil_div(M, R);
  const int r_lo = blockIdx.y * rows_per, r_hi = min(M, r_lo + rows_per);
  float a[8], b[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { a[j] = 0.f; b[j] = 0.f; }
  if (c0 < C)
    for (int r = r_lo + warp; r < r_hi; r += 8) f(r, c0, a, b);
  __shared__ float sh[8][2][kBnTile];
  __shared__ bool is_last;
#pragma unroll
  for (int j = 0; j < 8; ++j) { sh[warp][0][lane * 8 + j] = a[j]; sh[warp][1][lane * 8 + j] = b[j]; }
  __syncthreads();
  const int c = blockIdx.x * kBnTile + threadIdx.x;
  float t0 = 0.f, t1 = 0.f;
#pragma unroll
  for (int w = 0; w < 8; ++w) { t0 += sh[w][0][threadIdx.x]; t1 += sh[w][1][threadIdx.x]; }
  if (R == 1) {
    tot[0][threadIdx.x] = t0; tot[1][threadIdx.x] = t1;
    __syncthreads();
    return true;
  }
  if (c < C) {
    partial[((size_t)blockIdx.y * 2 + 0) * C + c] = t0;
    partial[((size_t)blockIdx.y * 2 + 1) * C + c] = t1;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int tk = atomicAdd(&tickets[blockIdx.x], 1u);
    is_last = tk == (unsigned int)R - 1;
    if (is_last) tickets[blockIdx.x] = 0u;
  }
  __syncthreads();
  if (!is_last) return false;
  __threadfence();
  t0 = t1 = 0.f;
  if (c < C)
    for (int k = 0; k < R; ++k) {
      t0 += __ldcg(partial + ((size_t)k * 2 + 0) * C + c);
      t1 += __ldcg(partial + ((size_t)k * 2 + 1) * C + c);
    }
  tot[0][threadIdx.x] = t0; tot[1][threadIdx.x] = t1;
  __syncthreads();
  return true;
}


}  // namespace mlt

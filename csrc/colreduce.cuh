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


// ---- flat, coalesced, atomics-folded variant (BatchNorm, bias gradients) -------------------------------------------
constexpr int kBnThreads = 256;
constexpr int kBnGroups = 32;         // channel groups (of 8) per CTA column tile: 256 channels, 512-byte row segments

__device__ __forceinline__ void ld8(const __nv_bfloat16* p, float (&v)[8]) {
  const uint4 raw = *reinterpret_cast<const uint4*>(p);
  const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[j]));
    v[2 * j] = f.x; v[2 * j + 1] = f.y;
  }
}
// volatile: the compiler keeps a run of these together instead of sinking each next to its consumer (which, with in-order issue,
// leaves one or two loads in flight per thread)
__device__ __forceinline__ uint4 ld8raw(const __nv_bfloat16* p) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ void cvt8(const uint4& raw, float (&v)[8]) {
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

// One- or two-quantity per-column reduction over the rows of a bf16 [M, C] matrix (C a multiple of 8).  grid = (column tiles, R row slices).  F(offset_of_group, a[8], b[8])
// accumulates one 8-channel group of one row.  A CTA folds its rows through shared memory and adds its per-channel
// partials to acc[2][C] with fp32 atomics (a few hundred thousand per call, spread over 2*C addresses); the last CTA of
// a tile (ticket) picks the totals up into tot[2][cw * 8] (shared memory), re-zeroes acc and returns true.
// LOADFIRST: F has two call forms, f.load(off) -> raw value and f.use(raw, a, b); the U loads of an iteration are issued
// before any of them is consumed (a single-stream pass needs that many 16-byte loads in flight per thread to fill HBM: with
// load-use-load-use order ncu showed bn_stats at 52 % of DRAM peak, 28 % issue-active, long-scoreboard stalls).
template <bool TWO, int U = 4, bool LOADFIRST = false, typename F>
__device__ __forceinline__ bool col_reduce_atomic(F f, float* acc, unsigned int* tickets, int M, int C, float* tot) {
  const int C8 = C >> 3;
  const int g_lo = blockIdx.x * kBnGroups;
  const int cw = min(kBnGroups, C8 - g_lo);            // channel groups in this tile
  const int rpi = kBnThreads / cw;                     // rows per iteration
  const int cg = threadIdx.x % cw, rsub = threadIdx.x / cw;
  const int R = gridDim.y;
  const int rows_per = ceil_div(M, R);
  const int r_lo = blockIdx.y * rows_per, r_hi = min(M, r_lo + rows_per);
  float a[8], b[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { a[j] = 0.f; b[j] = 0.f; }
  if (rsub < rpi) {
    const size_t col = (size_t)(g_lo + cg) * 8;
    int r = r_lo + rsub;
    if constexpr (LOADFIRST) {
      for (; r + (U - 1) * rpi < r_hi; r += U * rpi) {
        decltype(f.load((size_t)0)) raw[U];
#pragma unroll
        for (int u = 0; u < U; ++u) raw[u] = f.load((size_t)(r + u * rpi) * C + col);
#pragma unroll
        for (int u = 0; u < U; ++u) f.pin(raw[u]);       // every load has been issued before the first value is touched
#pragma unroll
        for (int u = 0; u < U; ++u) f.use(raw[u], a, b);
      }
      for (; r < r_hi; r += rpi) f.use(f.load((size_t)r * C + col), a, b);
    } else {
      for (; r + (U - 1) * rpi < r_hi; r += U * rpi) {   // U independent rows in flight
#pragma unroll
        for (int u = 0; u < U; ++u) f((size_t)(r + u * rpi) * C + col, a, b);
      }
      for (; r < r_hi; r += rpi) f((size_t)r * C + col, a, b);
    }
  }
  __shared__ float sh[kBnThreads][17];
  __shared__ bool is_last;
#pragma unroll
  for (int j = 0; j < 8; ++j) { sh[threadIdx.x][j] = a[j]; sh[threadIdx.x][8 + j] = b[j]; }
  __syncthreads();
  const int nch = cw * 8;
  for (int ch = threadIdx.x; ch < nch; ch += kBnThreads) {
    const int g = ch >> 3, j = ch & 7;
    float t0 = 0.f, t1 = 0.f;
    for (int s = 0; s < rpi; ++s) { t0 += sh[s * cw + g][j]; if (TWO) t1 += sh[s * cw + g][8 + j]; }
    if (R == 1) { tot[ch] = t0; tot[nch + ch] = t1; }
    else {
      atomicAdd(acc + g_lo * 8 + ch, t0);
      if (TWO) atomicAdd(acc + C + g_lo * 8 + ch, t1);
    }
  }
  if (R == 1) { __syncthreads(); return true; }
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
  for (int ch = threadIdx.x; ch < nch; ch += kBnThreads) {
    float* p0 = acc + g_lo * 8 + ch;
    float* p1 = acc + C + g_lo * 8 + ch;
    tot[ch] = __ldcg(p0);
    *p0 = 0.f;                                         // ready for the next call on this stream
    if (TWO) { tot[nch + ch] = __ldcg(p1); *p1 = 0.f; }
  }
  __syncthreads();
  return true;
}


}  // namespace mlt

// Fused multi-head attention on tcgen05: softmax(Q K^T / sqrt(d)) V and its backward without the [S, S]
// score matrix ever touching HBM.  Head dim 64, sequence length a multiple of 64 up to 512 (BERT-base: 512), optional
// key-padding mask as an additive per-key bias (HuggingFace's extended attention mask: 0 for real tokens, a large negative
// number for padding), staged once per CTA in shared memory and added to the scaled scores in all three kernels.
//
// All three kernels share one shape: a CTA owns a 128-row tile (queries, or keys in the dK/dV kernel), walks the other
// sequence axis in steps of 64 columns, and runs 320 threads, two CTAs per SM:
//   warp 0      TMA producer: the resident tile(s) once, the streamed 64-row operand tiles through a 2-3 stage ring.
//   warp 1      MMA issuer (one elected thread): score MMAs (M=128, N=64, K=64) into TMEM, then the accumulating
//               output MMAs whose A operand is the bf16 tile the softmax warps wrote into 128-byte swizzled smem.
//   warps 2-9   softmax: TMEM lane quadrant = warp % 4, column half = (warp - 2) / 4, so a thread owns 32 columns of one
//               row per step.  tcgen05.ld the scores, exponentiate, write bf16 to smem, arrive on an mbarrier.
// Two CTAs per SM (<= 113 KB smem, 256 TMEM columns each) keep two such dependency chains in flight per scheduler,
// which is what hides the MMA -> commit -> tcgen05.ld latency of a single chain.
//
// forward   S_j = Q K_j^T (two TMEM score buffers), online softmax with a lazy rescale of the O accumulator in TMEM
//           (only when the running max moved by more than 2^8), O += P_j V_j (V is the MN-major B operand), out = O / l,
//           lse kept for backward.
// backward  delta = rowsum(dO * O) (tiny kernel);  dQ kernel: dP = dO V^T, dS = P (dP - delta) scale, dQ += dS K;
//           dK/dV kernel in the transposed orientation (rows are keys): S^T = K Q^T, dP^T = V dO^T, dV += P^T dO,
//           dK += dS^T Q.  Each output tile is owned by exactly one CTA: no atomics, no cross-CTA reduction.
#include "common.cuh"
#include "ptx.cuh"
#include "../include/mlt_b200.h"
#include <mutex>

namespace mlt {

using namespace ptx;

namespace fa {
constexpr int kThreads = 320;
constexpr int kSoftmaxThreads = 256;
constexpr int kSoftmaxWarps = 8;              // barrier arrivals are one elected lane per warp
constexpr int kTile = 128 * 64 * 2;            // 16 KB: a [128 x 64] bf16 tile, 128-byte rows, SW128
constexpr int kHalf = 64 * 64 * 2;             // 8 KB: 64 rows of it
constexpr int kMaxS = 512;
constexpr uint32_t kTmemCols = 256;
constexpr float kLog2e = 1.4426950408889634f;
// forward: Q | 3 x (K half, V half) | 2 x P | row-max exchange | barriers
constexpr int kFwdStages = 3;
constexpr int kSmemFwd = kTile + kFwdStages * 2 * kHalf + 2 * kTile + 2 * 2 * 128 * 4 + kMaxS * 4 + 256 + 1024;
// dQ: Q | dO | 3 x (K half, V half) | dS | barriers
constexpr int kDqStages = 3;
constexpr int kSmemDq = 2 * kTile + kDqStages * 2 * kHalf + kTile + kMaxS * 4 + 256 + 1024;
// dK/dV: K | V | 2 x (Q half, dO half) | P | dS | lse, delta | barriers
constexpr int kDkvStages = 2;
constexpr int kSmemDkv = 2 * kTile + kDkvStages * 2 * kHalf + 2 * kTile + 2 * kMaxS * 4 + 256 + 1024;
}  // namespace fa

// Dropout on the attention probabilities.  The keep / drop decision of element (sample, head, query i, key j) is one byte of a
// 32-bit counter hash: the word belongs to the 2 x 2 block (i / 2, j / 2), byte (i & 1) * 2 + (j & 1) is the element's, and the
// element is kept when that byte is below `thr` (keep probability thr / 256, survivors scaled by 256 / thr, so the estimate
// stays unbiased at a drop rate quantised to 1/256).  A 2 x 2 block costs a thread one hash per two of its elements in
// either orientation (rows are queries in the forward and dQ kernels, keys in the dK/dV kernel), and all three kernels
// regenerate the same mask from (seed, *step, call) - nothing is stored.
struct FaDrop {
  const long long* step;     // device step counter (read at kernel start: a replayed CUDA graph draws fresh masks), or nullptr
  uint32_t seed_lo, seed_hi, call, thr;
  float scale;               // 256 / thr
};

struct FaParams {
  __nv_bfloat16* out;        // [B, S, H*64]
  float* lse;                // [B, H, S]  (natural-log sum-exp of the scaled scores)
  int S, H;
  long long row_stride;      // of out: H * 64
  float scale_log2;          // (1/sqrt(d)) * log2(e)
  float scale;
  const float* kbias;        // [B, S] additive key bias in natural-log units (added to the scaled scores), or nullptr
  FaDrop drop;
};

struct FaBwdParams {
  const __nv_bfloat16* out;   // forward output  [B, S, H*64]
  const __nv_bfloat16* dout;  // its gradient
  const float* lse;           // [B, H, S]
  float* delta;               // [B, H, S]
  __nv_bfloat16 *dq, *dk, *dv;
  int B, S, H;
  long long row_stride;       // of out / dout: H * 64
  long long ld_dqkv;          // row stride of dq, dk, dv (H * 64, or 3 * H * 64 when they are slices of one packed tensor)
  float scale_log2, scale;
  const float* kbias;         // as in FaParams
  FaDrop drop;
};

// Optional per-CTA timeline (analysis only): 8 x int64 per CTA when a buffer has been installed.
__device__ long long* g_fa_dbg = nullptr;
__device__ __forceinline__ long long gtimer() {
  long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#define FA_DBG(slot)                                                                                          \
  do {                                                                                                        \
    if (dbg != nullptr) dbg[((blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x) * 8 + (slot)] = gtimer(); \
  } while (0)

__device__ __forceinline__ void softmax_bar() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

// All lanes of the warp have done their part (TMEM reads retired, or smem writes + proxy fence): one lane arrives.
__device__ __forceinline__ void warp_arrive(uint64_t* bar, int lane) {
  __syncwarp();
  if (lane == 0) mbar_arrive(bar);
}

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__device__ __forceinline__ uint32_t fa_mix32(uint32_t x) {
  x = (x ^ (x >> 16)) * 0x7feb352du;
  x = (x ^ (x >> 15)) * 0x846ca68bu;
  return x ^ (x >> 16);
}
__device__ __forceinline__ uint2 fa_drop_key(const FaDrop& d) {
  const long long step = d.step != nullptr ? *d.step : 0;
  return make_uint2(fa_mix32(d.seed_lo ^ (d.call * 0x9E3779B9u) ^ (uint32_t)step),
                    fa_mix32(d.seed_hi ^ (d.call * 0x85EBCA6Bu) ^ (uint32_t)((unsigned long long)step >> 32) ^ 0x5bd1e995u));
}
__device__ __forceinline__ uint32_t fa_drop_word(uint2 key, uint32_t idx) {
  uint32_t x = idx ^ key.x;
  x = (x ^ (x >> 16)) * 0x7feb352du;
  x = (x ^ (x >> 15) ^ key.y) * 0x846ca68bu;
  return x ^ (x >> 16);
}
// Keep bits of 32 consecutive keys of one query row: bit c = element (i, j0 + c); idx0 = index of block (i / 2, j0 / 2).
__device__ __forceinline__ uint32_t fa_keep_row(uint2 key, uint32_t idx0, int odd_row, uint32_t thr) {
  uint32_t m = 0;
#pragma unroll
  for (int t = 0; t < 16; ++t) {
    const uint32_t hw = fa_drop_word(key, idx0 + t) >> (odd_row * 16);
    m |= ((hw & 0xFFu) < thr ? 1u : 0u) << (2 * t);
    m |= (((hw >> 8) & 0xFFu) < thr ? 1u : 0u) << (2 * t + 1);
  }
  return m;
}
// Keep bits of 32 consecutive queries of one key: bit c = element (i0 + c, j); idx0 = index of block (i0 / 2, j / 2), and one
// block row further is `stride` indices on.
__device__ __forceinline__ uint32_t fa_keep_col(uint2 key, uint32_t idx0, uint32_t stride, int odd_col, uint32_t thr) {
  uint32_t m = 0;
#pragma unroll
  for (int t = 0; t < 16; ++t) {
    const uint32_t w = fa_drop_word(key, idx0 + t * stride) >> (odd_col * 8);
    m |= ((w & 0xFFu) < thr ? 1u : 0u) << (2 * t);
    m |= (((w >> 16) & 0xFFu) < thr ? 1u : 0u) << (2 * t + 1);
  }
  return m;
}

// 8 consecutive bf16 of row `row` of a K-major SW128 tile: 16-byte chunk index XOR (row & 7).
__device__ __forceinline__ void store_row_chunk(uint8_t* tile_row, int row, int chunk, const float* v) {
  uint4 pk;
  pk.x = pack_bf16(v[0], v[1]); pk.y = pack_bf16(v[2], v[3]); pk.z = pack_bf16(v[4], v[5]); pk.w = pack_bf16(v[6], v[7]);
  *reinterpret_cast<uint4*>(tile_row + ((chunk ^ (row & 7)) << 4)) = pk;
}

// 32 fp32 accumulator columns -> 32 bf16 in global memory, scaled.
__device__ __forceinline__ void store_out32(__nv_bfloat16* g, uint32_t taddr, float mul) {
  uint32_t r[32];
  tmem_ld_32x32b_x32(taddr, r);
  tmem_ld_wait();
#pragma unroll
  for (int i = 0; i < 32; i += 8) {
    uint4 pk;
    pk.x = pack_bf16(__uint_as_float(r[i]) * mul, __uint_as_float(r[i + 1]) * mul);
    pk.y = pack_bf16(__uint_as_float(r[i + 2]) * mul, __uint_as_float(r[i + 3]) * mul);
    pk.z = pack_bf16(__uint_as_float(r[i + 4]) * mul, __uint_as_float(r[i + 5]) * mul);
    pk.w = pack_bf16(__uint_as_float(r[i + 6]) * mul, __uint_as_float(r[i + 7]) * mul);
    *reinterpret_cast<uint4*>(g + i) = pk;
  }
}

// ---- forward ------------------------------------------------------------------------------------------------------
template <bool BIAS, bool DROP>
__global__ void __launch_bounds__(fa::kThreads, 2)
attention_fwd_tcgen05_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                             const __grid_constant__ CUtensorMap tmV, const FaParams p) {
  using namespace fa;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sKV = sQ + kTile;                       // stage s: K half at s * 2 * kHalf, V half right after
  uint8_t* sP = sKV + kFwdStages * 2 * kHalf;      // 2 x [128 q x 64 keys]
  float* s_max = reinterpret_cast<float*>(sP + 2 * kTile);     // [2 buffers][2 halves][128 rows]
  float* s_kb = s_max + 2 * 2 * 128;                            // [S] key bias * log2(e)
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_kb + kMaxS);
  uint64_t* q_full = bars;                         // 1
  uint64_t* kv_full = bars + 1;                    // kFwdStages
  uint64_t* kv_free = kv_full + kFwdStages;        // kFwdStages
  uint64_t* s_full = kv_free + kFwdStages;         // 2
  uint64_t* s_free = s_full + 2;                   // 2
  uint64_t* p_full = s_free + 2;                   // 2
  uint64_t* p_free = p_full + 2;                   // 2
  uint64_t* o_done = p_free + 2;                   // 1
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_done + 1);
  constexpr uint32_t kColS = 0, kColO = 128;       // S buffers at 0 and 64, O at 128

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int qt = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int NH = p.S / 64;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmQ); prefetch_tmap(&tmK); prefetch_tmap(&tmV);
    mbar_init(q_full, 1);
    for (int j = 0; j < kFwdStages; ++j) { mbar_init(&kv_full[j], 1); mbar_init(&kv_free[j], 1); }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&s_full[i], 1); mbar_init(&s_free[i], kSoftmaxWarps);
      mbar_init(&p_full[i], kSoftmaxWarps); mbar_init(&p_free[i], 1);
    }
    mbar_init(o_done, 1);
    fence_mbar_init();
  }
  if (warp == 1) { tmem_alloc(tmem_slot, kTmemCols); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    if (elect_one()) {
      mbar_arrive_expect_tx(q_full, kTile);
      tma_load_4d(sQ, &tmQ, q_full, 0, qt * 128, h, b);
    }
    __syncwarp();
    for (int hs = 0; hs < NH; ++hs) {
      const int st = hs % kFwdStages, n = hs / kFwdStages;
      mbar_wait(&kv_free[st], (n & 1) ^ 1);
      if (elect_one()) {
        mbar_arrive_expect_tx(&kv_full[st], 2 * kHalf);
        tma_load_4d(sKV + st * 2 * kHalf, &tmK, &kv_full[st], 0, hs * 64, h, b);
        tma_load_4d(sKV + st * 2 * kHalf + kHalf, &tmV, &kv_full[st], 0, hs * 64, h, b);
      }
      __syncwarp();
    }
  } else if (warp == 1) {
    // The whole warp walks the loop and waits; one elected lane issues (keeps the operands in uniform registers).
    const uint32_t idesc_s = make_idesc_bf16(128, 64, 0, 0);       // Q, K both K-major
    const uint32_t idesc_o = make_idesc_bf16(128, 64, 0, 1);       // P K-major, V MN-major
    const uint32_t aQ = smem_u32(sQ), aKV = smem_u32(sKV), aP = smem_u32(sP);
    auto issue_pv = [&](int hs) {
      const int u = hs & 1, st = hs % kFwdStages;
      mbar_wait(&p_full[u], (hs >> 1) & 1);
      tc_fence_after();
      const uint32_t aV = aKV + st * 2 * kHalf + kHalf;
      if (elect_one()) {
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          umma_f16(tmem + kColO, make_smem_desc_sw128(aP + u * kTile + kk * 32, 16, 1024),
                   make_smem_desc_sw128(aV + kk * 2048, 8192, 1024), idesc_o, (hs > 0) || (kk > 0));
        umma_commit(&p_free[u]);                  // P buffer reusable, O holds step hs
        umma_commit(&kv_free[st]);
      }
      __syncwarp();
    };
    mbar_wait(q_full, 0);
    for (int hs = 0; hs < NH; ++hs) {
      const int u = hs & 1, st = hs % kFwdStages;
      mbar_wait(&kv_full[st], (hs / kFwdStages) & 1);
      mbar_wait(&s_free[u], ((hs >> 1) & 1) ^ 1);
      tc_fence_after();
      const uint32_t aK = aKV + st * 2 * kHalf;
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
          umma_f16(tmem + kColS + u * 64, make_smem_desc_sw128(aQ + k * 32, 16, 1024),
                   make_smem_desc_sw128(aK + k * 32, 16, 1024), idesc_s, k != 0);
        umma_commit(&s_full[u]);
      }
      __syncwarp();
      if (hs > 0) issue_pv(hs - 1);
    }
    issue_pv(NH - 1);
    if (elect_one()) umma_commit(o_done);
    __syncwarp();
  } else {
    const int quad = warp & 3, hf = (warp - 2) >> 2;
    const int row = quad * 32 + lane;                   // query row within the tile
    const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
    if (BIAS) {                                         // key bias of this sample, pre-multiplied into the exp2 domain
      for (int i = threadIdx.x - 64; i < p.S; i += kSoftmaxThreads) s_kb[i] = p.kbias[(size_t)b * p.S + i] * kLog2e;
      softmax_bar();
    }
    // m is the running reference maximum: of the raw scores without a bias, of the scaled + biased exponents (log2 domain) with one
    const float sl2 = p.scale_log2;
    float m = -INFINITY, l = 0.f;                       // l: this thread's half of the row sum
    uint2 dkey = make_uint2(0u, 0u);
    uint32_t didx = 0;                                  // hash index of this row's first 2 x 2 block
    if (DROP) {
      dkey = fa_drop_key(p.drop);
      didx = ((uint32_t)(b * p.H + h) * (uint32_t)(p.S >> 1) + (uint32_t)((qt * 128 + row) >> 1)) * (uint32_t)(p.S >> 1);
    }
    for (int hs = 0; hs < NH; ++hs) {
      const int u = hs & 1;
      mbar_wait(&s_full[u], (hs >> 1) & 1);
      tc_fence_after();
      uint32_t r[32];
      tmem_ld_32x32b_x32(tmem + lane_base + kColS + u * 64 + hf * 32, r);
      tmem_ld_wait();
      tc_fence_before();
      warp_arrive(&s_free[u], lane);                          // score buffer may be overwritten by S_{hs+2}
      if (BIAS) {
        const float4* kb4 = reinterpret_cast<const float4*>(s_kb + hs * 64 + hf * 32);
#pragma unroll
        for (int i4 = 0; i4 < 8; ++i4) {
          const float4 kb = kb4[i4];
          r[i4 * 4] = __float_as_uint(fmaf(__uint_as_float(r[i4 * 4]), sl2, kb.x));
          r[i4 * 4 + 1] = __float_as_uint(fmaf(__uint_as_float(r[i4 * 4 + 1]), sl2, kb.y));
          r[i4 * 4 + 2] = __float_as_uint(fmaf(__uint_as_float(r[i4 * 4 + 2]), sl2, kb.z));
          r[i4 * 4 + 3] = __float_as_uint(fmaf(__uint_as_float(r[i4 * 4 + 3]), sl2, kb.w));
        }
      }
      float mx0 = __uint_as_float(r[0]), mx1 = __uint_as_float(r[1]), mx2 = __uint_as_float(r[2]), mx3 = __uint_as_float(r[3]);
#pragma unroll
      for (int i = 4; i < 32; i += 4) {
        mx0 = fmaxf(mx0, __uint_as_float(r[i])); mx1 = fmaxf(mx1, __uint_as_float(r[i + 1]));
        mx2 = fmaxf(mx2, __uint_as_float(r[i + 2])); mx3 = fmaxf(mx3, __uint_as_float(r[i + 3]));
      }
      const float mine = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));
      float* xm = s_max + u * 256;
      xm[hf * 128 + row] = mine;
      softmax_bar();
      const float mx = fmaxf(m, fmaxf(mine, xm[(hf ^ 1) * 128 + row]));
      // lazy rescale: keep the old reference max unless some row of this warp moved by more than 2^8
      const float to_exp = BIAS ? 1.0f : sl2;                     // converts a difference of m's into the exp2 domain
      const bool grow = (mx - m) * to_exp > 8.0f;                 // also true on the first step (m = -inf)
      if (__any_sync(0xffffffffu, grow)) {
        const float alpha = fast_exp2((m - mx) * to_exp);         // 0 on the first step
        l *= alpha;
        m = mx;
        if (hs > 0) {
          mbar_wait(&p_free[(hs - 1) & 1], ((hs - 1) >> 1) & 1);  // every PV issued so far has landed in O
          tc_fence_after();
          uint32_t o[32];
          tmem_ld_32x32b_x32(tmem + lane_base + kColO + hf * 32, o);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
          tmem_st_32x32b_x32(tmem + lane_base + kColO + hf * 32, o);
          tmem_st_wait();
        }
      }
      const float off = -m * to_exp;
      float pv[32];
      float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
      for (int i = 0; i < 32; i += 4) {
        if (BIAS) {
          pv[i] = fast_exp2(__uint_as_float(r[i]) + off);         s0 += pv[i];
          pv[i + 1] = fast_exp2(__uint_as_float(r[i + 1]) + off); s1 += pv[i + 1];
          pv[i + 2] = fast_exp2(__uint_as_float(r[i + 2]) + off); s2 += pv[i + 2];
          pv[i + 3] = fast_exp2(__uint_as_float(r[i + 3]) + off); s3 += pv[i + 3];
        } else {
          pv[i] = fast_exp2(fmaf(__uint_as_float(r[i]), sl2, off));         s0 += pv[i];
          pv[i + 1] = fast_exp2(fmaf(__uint_as_float(r[i + 1]), sl2, off)); s1 += pv[i + 1];
          pv[i + 2] = fast_exp2(fmaf(__uint_as_float(r[i + 2]), sl2, off)); s2 += pv[i + 2];
          pv[i + 3] = fast_exp2(fmaf(__uint_as_float(r[i + 3]), sl2, off)); s3 += pv[i + 3];
        }
      }
      l += (s0 + s1) + (s2 + s3);                       // the normaliser is the sum of ALL probabilities, dropped ones included
      if (DROP) {
        const uint32_t keep = fa_keep_row(dkey, didx + (uint32_t)((hs * 64 + hf * 32) >> 1), row & 1, p.drop.thr);
#pragma unroll
        for (int i = 0; i < 32; ++i) pv[i] = ((keep >> i) & 1u) ? pv[i] : 0.f;
      }
      mbar_wait(&p_free[u], ((hs >> 1) & 1) ^ 1);       // PV of step hs-2 no longer reads this P buffer
      uint8_t* prow = sP + u * kTile + row * 128;
#pragma unroll
      for (int g = 0; g < 4; ++g) store_row_chunk(prow, row, hf * 4 + g, pv + g * 8);
      fence_proxy_async();
      tc_fence_before();
      warp_arrive(&p_full[u], lane);
    }
    // epilogue: O / l -> bf16, log-sum-exp
    float* xl = s_max;                                  // reuse: the two halves of the row sum
    softmax_bar();
    xl[hf * 128 + row] = l;
    softmax_bar();
    const float lt = l + xl[(hf ^ 1) * 128 + row];
    mbar_wait(o_done, 0);
    tc_fence_after();
    const int srow = qt * 128 + row;
    if (srow < p.S) {                                   // S % 128 == 64: the last tile's upper half is padding (TMA zero fill)
      __nv_bfloat16* orow = p.out + ((size_t)b * p.S + srow) * p.row_stride + (size_t)h * 64 + hf * 32;
      store_out32(orow, tmem + lane_base + kColO + hf * 32, (DROP ? p.drop.scale : 1.0f) / lt);
      if (hf == 0 && p.lse != nullptr)
        p.lse[((size_t)b * p.H + h) * p.S + srow] = (BIAS ? m * (1.0f / kLog2e) : m * p.scale) + logf(lt);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, kTmemCols);
}

// ---- backward -----------------------------------------------------------------------------------------------------
// delta[b, h, s] = sum_d dO[b, s, h, d] * O[b, s, h, d].  8 lanes per (row, head): fully coalesced 16-byte loads.
__global__ void __launch_bounds__(256) attention_delta_kernel(const FaBwdParams p) {
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long unit = gid >> 3;                       // (b, s, h) flattened, h fastest
  const int sub = (int)(gid & 7);
  const long long total = (long long)p.B * p.S * p.H;
  float acc = 0.f;
  if (unit < total) {
    const uint4 a = reinterpret_cast<const uint4*>(p.out)[unit * 8 + sub];
    const uint4 g = reinterpret_cast<const uint4*>(p.dout)[unit * 8 + sub];
    const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, gw[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 af = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&aw[i]));
      const float2 gf = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&gw[i]));
      acc = fmaf(af.x, gf.x, acc);
      acc = fmaf(af.y, gf.y, acc);
    }
  }
  acc += __shfl_xor_sync(0xffffffffu, acc, 1);
  acc += __shfl_xor_sync(0xffffffffu, acc, 2);
  acc += __shfl_xor_sync(0xffffffffu, acc, 4);
  if (sub == 0 && unit < total) {
    const int hh = (int)(unit % p.H);
    const long long bs = unit / p.H;
    const int s = (int)(bs % p.S);
    const long long b = bs / p.S;
    p.delta[(b * p.H + hh) * p.S + s] = acc;
  }
}

template <bool BIAS, bool DROP>
__global__ void __launch_bounds__(fa::kThreads, 2)
attention_bwd_dq_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                        const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmdO,
                        const FaBwdParams p) {
  using namespace fa;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sdO = sQ + kTile;
  uint8_t* sKV = sdO + kTile;                      // stage s: K half, V half
  uint8_t* sdS = sKV + kDqStages * 2 * kHalf;      // [128 q x 64 keys]
  float* s_kb = reinterpret_cast<float*>(sdS + kTile);          // [S] key bias * log2(e)
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_kb + kMaxS);
  uint64_t* q_full = bars;                         // 1 (Q and dO)
  uint64_t* kv_full = bars + 1;                    // kDqStages
  uint64_t* kv_free = kv_full + kDqStages;         // kDqStages
  uint64_t* st_full = kv_free + kDqStages;         // 1
  uint64_t* st_free = st_full + 1;                 // 1
  uint64_t* ds_full = st_free + 1;                 // 1
  uint64_t* ds_free = ds_full + 1;                 // 1
  uint64_t* done = ds_free + 1;                    // 1
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done + 1);
  constexpr uint32_t kColStage = 0, kColOut = 128; // [S(64) | dP(64)], dQ

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int qt = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int NH = p.S / 64;
  long long* dbg = g_fa_dbg;
  if (threadIdx.x == 0) FA_DBG(0);

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmQ); prefetch_tmap(&tmK); prefetch_tmap(&tmV); prefetch_tmap(&tmdO);
    mbar_init(q_full, 1);
    for (int j = 0; j < kDqStages; ++j) { mbar_init(&kv_full[j], 1); mbar_init(&kv_free[j], 1); }
    mbar_init(st_full, 1); mbar_init(st_free, kSoftmaxWarps); mbar_init(ds_full, kSoftmaxWarps); mbar_init(ds_free, 1);
    mbar_init(done, 1);
    fence_mbar_init();
  }
  if (warp == 1) { tmem_alloc(tmem_slot, kTmemCols); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  if (threadIdx.x == 0) FA_DBG(1);

  if (warp == 0) {
    if (elect_one()) {
      mbar_arrive_expect_tx(q_full, 2 * kTile);
      tma_load_4d(sQ, &tmQ, q_full, 0, qt * 128, h, b);
      tma_load_4d(sdO, &tmdO, q_full, 0, qt * 128, h, b);
    }
    __syncwarp();
    for (int hs = 0; hs < NH; ++hs) {
      const int st = hs % kDqStages, n = hs / kDqStages;
      mbar_wait(&kv_free[st], (n & 1) ^ 1);
      if (elect_one()) {
        mbar_arrive_expect_tx(&kv_full[st], 2 * kHalf);
        tma_load_4d(sKV + st * 2 * kHalf, &tmK, &kv_full[st], 0, hs * 64, h, b);
        tma_load_4d(sKV + st * 2 * kHalf + kHalf, &tmV, &kv_full[st], 0, hs * 64, h, b);
      }
      __syncwarp();
    }
  } else if (warp == 1) {
    const uint32_t idesc_s = make_idesc_bf16(128, 64, 0, 0);
    const uint32_t idesc_o = make_idesc_bf16(128, 64, 0, 1);
    const uint32_t aQ = smem_u32(sQ), adO = smem_u32(sdO), aKV = smem_u32(sKV), adS = smem_u32(sdS);
    auto issue_scores = [&](int hs) {
      const int st = hs % kDqStages;
      const uint32_t aK = aKV + st * 2 * kHalf, aV = aK + kHalf;
      mbar_wait(&kv_full[st], (hs / kDqStages) & 1);
      mbar_wait(st_free, (hs & 1) ^ 1);                 // the softmax warps have read step hs-1 out of TMEM
      tc_fence_after();
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < 4; ++k)           // S = Q K^T
          umma_f16(tmem + kColStage, make_smem_desc_sw128(aQ + k * 32, 16, 1024), make_smem_desc_sw128(aK + k * 32, 16, 1024), idesc_s, k != 0);
#pragma unroll
        for (int k = 0; k < 4; ++k)           // dP = dO V^T
          umma_f16(tmem + kColStage + 64, make_smem_desc_sw128(adO + k * 32, 16, 1024), make_smem_desc_sw128(aV + k * 32, 16, 1024), idesc_s, k != 0);
        umma_commit(st_full);
      }
      __syncwarp();
    };
    mbar_wait(q_full, 0);
    if (lane == 0) FA_DBG(2);
    issue_scores(0);
    for (int hs = 0; hs < NH; ++hs) {
      if (hs + 1 < NH) issue_scores(hs + 1);            // overlaps the softmax of step hs
      const int st = hs % kDqStages;
      const uint32_t aK = aKV + st * 2 * kHalf;
      mbar_wait(ds_full, hs & 1);
      tc_fence_after();
      if (elect_one()) {
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)        // dQ += dS K   (K rows as the MN-major B operand)
          umma_f16(tmem + kColOut, make_smem_desc_sw128(adS + kk * 32, 16, 1024), make_smem_desc_sw128(aK + kk * 2048, 8192, 1024),
                   idesc_o, (hs > 0) || (kk > 0));
        umma_commit(ds_free);
        umma_commit(&kv_free[st]);
      }
      __syncwarp();
    }
    if (elect_one()) umma_commit(done);
    __syncwarp();
  } else {
    const int quad = warp & 3, hf = (warp - 2) >> 2;
    const int row = quad * 32 + lane;
    const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
    const int srow = qt * 128 + row;
    const bool live = srow < p.S;                       // S % 128 == 64: rows past the end are padding
    const size_t soff = ((size_t)b * p.H + h) * p.S + (live ? srow : 0);
    const float lse2 = live ? p.lse[soff] * kLog2e : 0.f;
    const float dsc = live ? p.delta[soff] * p.scale : 0.f;
    uint8_t* drow = sdS + row * 128;
    if (BIAS) {
      for (int i = threadIdx.x - 64; i < p.S; i += kSoftmaxThreads) s_kb[i] = p.kbias[(size_t)b * p.S + i] * kLog2e;
      softmax_bar();
    }
    // dP = (dO V^T) * keep / (1 - p): a dropped element contributes only the -delta term
    const float psc = DROP ? p.scale * p.drop.scale : p.scale;
    uint2 dkey = make_uint2(0u, 0u);
    uint32_t didx = 0;
    if (DROP) {
      dkey = fa_drop_key(p.drop);
      didx = ((uint32_t)(b * p.H + h) * (uint32_t)(p.S >> 1) + (uint32_t)(srow >> 1)) * (uint32_t)(p.S >> 1);
    }
    for (int hs = 0; hs < NH; ++hs) {
      mbar_wait(st_full, hs & 1);
      if (threadIdx.x == 64 && hs == 0) FA_DBG(4);
      tc_fence_after();
      uint32_t rs[32], rp[32];
      tmem_ld_32x32b_x32(tmem + lane_base + kColStage + hf * 32, rs);
      tmem_ld_32x32b_x32(tmem + lane_base + kColStage + 64 + hf * 32, rp);
      tmem_ld_wait();
      tc_fence_before();
      warp_arrive(st_free, lane);
      float ds[32];
      uint32_t keep = 0xFFFFFFFFu;
      if (DROP) keep = fa_keep_row(dkey, didx + (uint32_t)((hs * 64 + hf * 32) >> 1), row & 1, p.drop.thr);
      if (BIAS) {
        const float4* kb4 = reinterpret_cast<const float4*>(s_kb + hs * 64 + hf * 32);
#pragma unroll
        for (int i4 = 0; i4 < 8; ++i4) {
          const float4 kq = kb4[i4];
          const float kv4[4] = {kq.x, kq.y, kq.z, kq.w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int i = i4 * 4 + e;
            const float pr = fast_exp2(fmaf(__uint_as_float(rs[i]), p.scale_log2, kv4[e] - lse2));
            const float g = fmaf(__uint_as_float(rp[i]), psc, -dsc);
            ds[i] = pr * ((!DROP || ((keep >> i) & 1u)) ? g : -dsc);
          }
        }
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const float pr = fast_exp2(fmaf(__uint_as_float(rs[i]), p.scale_log2, -lse2));
          const float g = fmaf(__uint_as_float(rp[i]), psc, -dsc);
          ds[i] = pr * ((!DROP || ((keep >> i) & 1u)) ? g : -dsc);
        }
      }
      mbar_wait(ds_free, (hs & 1) ^ 1);
#pragma unroll
      for (int g = 0; g < 4; ++g) store_row_chunk(drow, row, hf * 4 + g, ds + g * 8);
      fence_proxy_async();
      warp_arrive(ds_full, lane);
    }
    if (threadIdx.x == 64) FA_DBG(5);
    mbar_wait(done, 0);
    if (threadIdx.x == 64) FA_DBG(6);
    tc_fence_after();
    if (live)
      store_out32(p.dq + ((size_t)b * p.S + srow) * p.ld_dqkv + (size_t)h * 64 + hf * 32, tmem + lane_base + kColOut + hf * 32, 1.0f);
    if (threadIdx.x == 64) FA_DBG(3);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, kTmemCols);
  if (threadIdx.x == 32) FA_DBG(7);
}

template <bool BIAS, bool DROP>
__global__ void __launch_bounds__(fa::kThreads, 2)
attention_bwd_dkdv_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                          const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmdO,
                          const FaBwdParams p) {
  using namespace fa;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sK = smem;
  uint8_t* sV = sK + kTile;
  uint8_t* sQdO = sV + kTile;                      // stage s: Q half (64 queries), dO half
  uint8_t* sP = sQdO + kDkvStages * 2 * kHalf;     // [128 keys x 64 queries]
  uint8_t* sdS = sP + kTile;
  float* s_lse2 = reinterpret_cast<float*>(sdS + kTile);       // [S]
  float* s_dsc = s_lse2 + kMaxS;                               // [S]  delta * scale
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_dsc + kMaxS);
  uint64_t* kv_full = bars;                        // 1
  uint64_t* qdo_full = bars + 1;                   // kDkvStages
  uint64_t* qdo_free = qdo_full + kDkvStages;      // kDkvStages
  uint64_t* st_full = qdo_free + kDkvStages;       // 1
  uint64_t* st_free = st_full + 1;                 // 1
  uint64_t* pd_full = st_free + 1;                 // 1
  uint64_t* pd_free = pd_full + 1;                 // 1
  uint64_t* done = pd_free + 1;                    // 1
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done + 1);
  constexpr uint32_t kColStage = 0, kColdV = 128, kColdK = 192;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kt = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int NH = p.S / 64;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmQ); prefetch_tmap(&tmK); prefetch_tmap(&tmV); prefetch_tmap(&tmdO);
    mbar_init(kv_full, 1);
    for (int i = 0; i < kDkvStages; ++i) { mbar_init(&qdo_full[i], 1); mbar_init(&qdo_free[i], 1); }
    mbar_init(st_full, 1); mbar_init(st_free, kSoftmaxWarps); mbar_init(pd_full, kSoftmaxWarps); mbar_init(pd_free, 1);
    mbar_init(done, 1);
    fence_mbar_init();
  }
  if (warp == 1) { tmem_alloc(tmem_slot, kTmemCols); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    if (elect_one()) {
      mbar_arrive_expect_tx(kv_full, 2 * kTile);
      tma_load_4d(sK, &tmK, kv_full, 0, kt * 128, h, b);
      tma_load_4d(sV, &tmV, kv_full, 0, kt * 128, h, b);
    }
    __syncwarp();
    for (int hs = 0; hs < NH; ++hs) {
      const int st = hs % kDkvStages, n = hs / kDkvStages;
      mbar_wait(&qdo_free[st], (n & 1) ^ 1);
      if (elect_one()) {
        mbar_arrive_expect_tx(&qdo_full[st], 2 * kHalf);
        tma_load_4d(sQdO + st * 2 * kHalf, &tmQ, &qdo_full[st], 0, hs * 64, h, b);
        tma_load_4d(sQdO + st * 2 * kHalf + kHalf, &tmdO, &qdo_full[st], 0, hs * 64, h, b);
      }
      __syncwarp();
    }
  } else if (warp == 1) {
    const uint32_t idesc_s = make_idesc_bf16(128, 64, 0, 0);
    const uint32_t idesc_o = make_idesc_bf16(128, 64, 0, 1);
    const uint32_t aK = smem_u32(sK), aV = smem_u32(sV), aQdO = smem_u32(sQdO), aP = smem_u32(sP), adS = smem_u32(sdS);
    auto issue_scores = [&](int hs) {
      const int st = hs % kDkvStages;
      const uint32_t aQ = aQdO + st * 2 * kHalf, adO = aQ + kHalf;
      mbar_wait(&qdo_full[st], (hs / kDkvStages) & 1);
      mbar_wait(st_free, (hs & 1) ^ 1);
      tc_fence_after();
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < 4; ++k)         // S^T = K Q^T
          umma_f16(tmem + kColStage, make_smem_desc_sw128(aK + k * 32, 16, 1024), make_smem_desc_sw128(aQ + k * 32, 16, 1024), idesc_s, k != 0);
#pragma unroll
        for (int k = 0; k < 4; ++k)         // dP^T = V dO^T
          umma_f16(tmem + kColStage + 64, make_smem_desc_sw128(aV + k * 32, 16, 1024), make_smem_desc_sw128(adO + k * 32, 16, 1024), idesc_s, k != 0);
        umma_commit(st_full);
      }
      __syncwarp();
    };
    mbar_wait(kv_full, 0);
    issue_scores(0);
    for (int hs = 0; hs < NH; ++hs) {
      if (hs + 1 < NH) issue_scores(hs + 1);
      const int st = hs % kDkvStages;
      const uint32_t aQ = aQdO + st * 2 * kHalf, adO = aQ + kHalf;
      mbar_wait(pd_full, hs & 1);
      tc_fence_after();
      if (elect_one()) {
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)      // dV += P^T dO   (A: [keys x 64 q] K-major, B: dO rows as MN-major)
          umma_f16(tmem + kColdV, make_smem_desc_sw128(aP + kk * 32, 16, 1024), make_smem_desc_sw128(adO + kk * 2048, 8192, 1024),
                   idesc_o, (hs > 0) || (kk > 0));
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)      // dK += dS^T Q
          umma_f16(tmem + kColdK, make_smem_desc_sw128(adS + kk * 32, 16, 1024), make_smem_desc_sw128(aQ + kk * 2048, 8192, 1024),
                   idesc_o, (hs > 0) || (kk > 0));
        umma_commit(pd_free);
        umma_commit(&qdo_free[st]);
      }
      __syncwarp();
    }
    if (elect_one()) umma_commit(done);
    __syncwarp();
  } else {
    const int quad = warp & 3, hf = (warp - 2) >> 2;
    const int row = quad * 32 + lane;                         // key row within the chunk
    const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
    const size_t sbase = ((size_t)b * p.H + h) * p.S;
    for (int i = threadIdx.x - 64; i < p.S; i += kSoftmaxThreads) {
      s_lse2[i] = p.lse[sbase + i] * kLog2e;
      s_dsc[i] = p.delta[sbase + i] * p.scale;
    }
    softmax_bar();
    uint8_t* prow = sP + row * 128;
    uint8_t* drow = sdS + row * 128;
    const int krow = kt * 128 + row;                          // this thread's key
    const bool live = krow < p.S;
    // rows are keys here: the key bias is one number per thread (a padded-out key row gets a bias that zeroes its P)
    const float kb = live ? (BIAS ? p.kbias[(size_t)b * p.S + krow] * kLog2e : 0.f) : -3.0e4f;
    const float psc = DROP ? p.scale * p.drop.scale : p.scale;
    const uint32_t S2 = (uint32_t)(p.S >> 1);
    uint2 dkey = make_uint2(0u, 0u);
    uint32_t didx = 0;                                        // hash index of block (query 0, this key / 2)
    if (DROP) {
      dkey = fa_drop_key(p.drop);
      didx = (uint32_t)(b * p.H + h) * S2 * S2 + (uint32_t)(krow >> 1);
    }
    for (int hs = 0; hs < NH; ++hs) {
      mbar_wait(st_full, hs & 1);
      tc_fence_after();
      uint32_t rs[32], rp[32];
      tmem_ld_32x32b_x32(tmem + lane_base + kColStage + hf * 32, rs);
      tmem_ld_32x32b_x32(tmem + lane_base + kColStage + 64 + hf * 32, rp);
      tmem_ld_wait();
      tc_fence_before();
      warp_arrive(st_free, lane);
      float pr[32], ds[32];
      const float4* l4 = reinterpret_cast<const float4*>(s_lse2 + hs * 64 + hf * 32);
      const float4* d4 = reinterpret_cast<const float4*>(s_dsc + hs * 64 + hf * 32);
#pragma unroll
      for (int i4 = 0; i4 < 8; ++i4) {
        const float4 lq = l4[i4], dq = d4[i4];
        const float lv[4] = {lq.x, lq.y, lq.z, lq.w}, dv[4] = {dq.x, dq.y, dq.z, dq.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int i = i4 * 4 + e;
          pr[i] = fast_exp2(fmaf(__uint_as_float(rs[i]), p.scale_log2, kb - lv[e]));
          ds[i] = pr[i] * fmaf(__uint_as_float(rp[i]), psc, -dv[e]);
        }
      }
      if (DROP) {      // dropped elements: dS keeps only the -delta term, and P^T (the dV operand) is zero there
        const uint32_t keep = fa_keep_col(dkey, didx + (uint32_t)((hs * 64 + hf * 32) >> 1) * S2, S2, row & 1, p.drop.thr);
        const float4* d4b = reinterpret_cast<const float4*>(s_dsc + hs * 64 + hf * 32);
#pragma unroll
        for (int i4 = 0; i4 < 8; ++i4) {
          const float4 dq = d4b[i4];
          const float dv[4] = {dq.x, dq.y, dq.z, dq.w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int i = i4 * 4 + e;
            if (!((keep >> i) & 1u)) { ds[i] = -pr[i] * dv[e]; pr[i] = 0.f; }
          }
        }
      }
      mbar_wait(pd_free, (hs & 1) ^ 1);
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        store_row_chunk(prow, row, hf * 4 + g, pr + g * 8);
        store_row_chunk(drow, row, hf * 4 + g, ds + g * 8);
      }
      fence_proxy_async();
      warp_arrive(pd_full, lane);
    }
    mbar_wait(done, 0);
    tc_fence_after();
    if (live) {
      const size_t goff = ((size_t)b * p.S + krow) * p.ld_dqkv + (size_t)h * 64 + hf * 32;
      store_out32(p.dv + goff, tmem + lane_base + kColdV + hf * 32, DROP ? p.drop.scale : 1.0f);
      store_out32(p.dk + goff, tmem + lane_base + kColdK + hf * 32, 1.0f);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, kTmemCols);
}

// ---- host ---------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn2)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static int fa_make_map(CUtensorMap* m, const void* base, int S, int H, int B, long long row_stride, int box_rows) {
  static EncodeTiledFn2 enc = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* fp = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      enc = reinterpret_cast<EncodeTiledFn2>(fp);
  });
  MLT_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled is not available from the driver");
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) { MLT_CHECK_CUDA(cudaFree(nullptr)); ctx_bound = true; }
  // [b][s][h][d]: dims innermost first {d, s, h, b}
  cuuint64_t dims[4] = {64, (cuuint64_t)S, (cuuint64_t)H, (cuuint64_t)B};
  cuuint64_t strides[3] = {(cuuint64_t)row_stride * 2, 64 * 2, (cuuint64_t)S * row_stride * 2};
  cuuint32_t box[4] = {64, (cuuint32_t)box_rows, 1, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  MLT_REQUIRE(r == CUDA_SUCCESS, "attention: cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return 0;
}

template <bool BIAS, bool DROP>
static cudaError_t fa_set_attrs_for() {
  cudaError_t e = cudaFuncSetAttribute(attention_fwd_tcgen05_kernel<BIAS, DROP>, cudaFuncAttributeMaxDynamicSharedMemorySize, fa::kSmemFwd);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(attention_bwd_dq_kernel<BIAS, DROP>, cudaFuncAttributeMaxDynamicSharedMemorySize, fa::kSmemDq);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(attention_bwd_dkdv_kernel<BIAS, DROP>, cudaFuncAttributeMaxDynamicSharedMemorySize, fa::kSmemDkv);
  return e;
}

static int fa_set_attrs() {
  static bool attr_set[64] = {};
  int dev = 0;
  MLT_CHECK_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63]) {
    MLT_CHECK_CUDA((fa_set_attrs_for<false, false>()));
    MLT_CHECK_CUDA((fa_set_attrs_for<true, false>()));
    MLT_CHECK_CUDA((fa_set_attrs_for<false, true>()));
    MLT_CHECK_CUDA((fa_set_attrs_for<true, true>()));
    attr_set[dev & 63] = true;
  }
  return 0;
}

static int fa_fill_drop(FaDrop* d, float p_drop, unsigned long long seed, const void* step_ptr, unsigned int call_id) {
  MLT_REQUIRE(p_drop >= 0.f && p_drop < 1.f, "attention: dropout probability %f must be in [0, 1)", (double)p_drop);
  int thr = (int)lrintf((1.0f - p_drop) * 256.0f);
  thr = thr < 1 ? 1 : (thr > 256 ? 256 : thr);
  d->step = (const long long*)step_ptr;
  d->seed_lo = (uint32_t)seed; d->seed_hi = (uint32_t)(seed >> 32); d->call = call_id;
  d->thr = (uint32_t)thr; d->scale = 256.0f / (float)thr;
  return 0;
}

}  // namespace mlt

using namespace mlt;

// Analysis hook: install (or clear with NULL) a device buffer of 8 int64 per CTA that the dQ kernel fills with timestamps.
extern "C" int mlt_attention_set_debug(void* buf) {
  long long* b = (long long*)buf;
  MLT_CHECK_CUDA(cudaMemcpyToSymbol(g_fa_dbg, &b, sizeof(b)));
  return 0;
}

// q, k, v: [B, S, H*64] bf16 with row stride ld_qkv (0 = H*64; 3*H*64 for slices of a packed QKV projection);
// out: [B, S, H*64] bf16;  lse: [B, H, S] fp32 or NULL.  key_bias: NULL, or [B, S] fp32 added to every score of that key after
// the 1/sqrt(d) scaling (HuggingFace's extended attention mask: 0 for tokens, a large negative number for padding).
// p_drop > 0: dropout on the probabilities inside the kernel (FaDrop above); the mask is a function of (seed, *step_ptr, call_id)
// and the backward call must be given the same three.
extern "C" int mlt_attention_fwd_ex(const void* q, const void* k, const void* v, long long ld_qkv, const void* key_bias, void* out,
                                    void* lse, int B, int S, int H, int D, float p_drop, unsigned long long seed, const void* step_ptr,
                                    unsigned int call_id, void* stream) {
  MLT_REQUIRE(D == 64, "mlt_attention_fwd: head dim %d unsupported (64)", D);
  MLT_REQUIRE(S % 64 == 0 && S >= 64 && S <= fa::kMaxS, "mlt_attention_fwd: sequence length %d unsupported (64..512, multiple of 64)", S);
  MLT_REQUIRE(((uintptr_t)q % 16) == 0 && ((uintptr_t)k % 16) == 0 && ((uintptr_t)v % 16) == 0 && ((uintptr_t)out % 16) == 0,
              "mlt_attention_fwd: unaligned tensors");
  const long long rs = (long long)H * D;
  const long long ldq = ld_qkv > 0 ? ld_qkv : rs;
  MLT_REQUIRE(ldq >= rs && ldq % 8 == 0, "mlt_attention_fwd: bad q/k/v row stride %lld", ldq);
  CUtensorMap tq, tk, tv;
  if (fa_make_map(&tq, q, S, H, B, ldq, 128) || fa_make_map(&tk, k, S, H, B, ldq, 64) || fa_make_map(&tv, v, S, H, B, ldq, 64)) return 1;
  FaParams p;
  p.out = (__nv_bfloat16*)out; p.lse = (float*)lse; p.S = S; p.H = H; p.row_stride = rs;
  p.scale = 1.0f / sqrtf((float)D);
  p.scale_log2 = p.scale * fa::kLog2e;
  p.kbias = (const float*)key_bias;
  if (fa_fill_drop(&p.drop, p_drop, seed, step_ptr, call_id)) return 1;
  if (fa_set_attrs()) return 1;
  dim3 grid((S + 127) / 128, H, B);
  cudaStream_t st = (cudaStream_t)stream;
  const bool drop = p_drop > 0.f;
#define MLT_FA_FWD(BIAS, DROP) attention_fwd_tcgen05_kernel<BIAS, DROP><<<grid, fa::kThreads, fa::kSmemFwd, st>>>(tq, tk, tv, p)
  if (key_bias) { if (drop) MLT_FA_FWD(true, true); else MLT_FA_FWD(true, false); }
  else { if (drop) MLT_FA_FWD(false, true); else MLT_FA_FWD(false, false); }
#undef MLT_FA_FWD
  MLT_LAUNCH_CHECK();
  return 0;
}

extern "C" int mlt_attention_fwd_masked(const void* q, const void* k, const void* v, long long ld_qkv, const void* key_bias, void* out,
                                        void* lse, int B, int S, int H, int D, void* stream) {
  return mlt_attention_fwd_ex(q, k, v, ld_qkv, key_bias, out, lse, B, S, H, D, 0.f, 0ull, nullptr, 0u, stream);
}

extern "C" int mlt_attention_fwd(const void* q, const void* k, const void* v, long long ld_qkv, void* out, void* lse, int B, int S,
                                 int H, int D, void* stream) {
  return mlt_attention_fwd_masked(q, k, v, ld_qkv, nullptr, out, lse, B, S, H, D, stream);
}

// Backward of mlt_attention_fwd[_masked].  out / lse are the forward's results, dout the gradient of out; delta is a [B, H, S] fp32
// scratch the call fills.  dq, dk, dv: [B, S, H*64] bf16 with row stride ld_dqkv (0 = H*64).  key_bias as in the forward.
extern "C" int mlt_attention_bwd_ex(const void* q, const void* k, const void* v, long long ld_qkv, const void* key_bias,
                                    const void* out, const void* dout, const void* lse, void* delta, void* dq, void* dk, void* dv,
                                    long long ld_dqkv, int B, int S, int H, int D, float p_drop, unsigned long long seed,
                                    const void* step_ptr, unsigned int call_id, void* stream) {
  MLT_REQUIRE(D == 64, "mlt_attention_bwd: head dim %d unsupported (64)", D);
  MLT_REQUIRE(S % 64 == 0 && S >= 64 && S <= fa::kMaxS, "mlt_attention_bwd: sequence length %d unsupported (64..512, multiple of 64)", S);
  const void* al[] = {q, k, v, out, dout, dq, dk, dv};
  for (const void* a : al) MLT_REQUIRE(((uintptr_t)a % 16) == 0, "mlt_attention_bwd: unaligned tensors");
  const long long rs = (long long)H * D;
  const long long ldq = ld_qkv > 0 ? ld_qkv : rs, ldd = ld_dqkv > 0 ? ld_dqkv : rs;
  MLT_REQUIRE(ldq >= rs && ldq % 8 == 0 && ldd >= rs && ldd % 8 == 0, "mlt_attention_bwd: bad row strides %lld / %lld", ldq, ldd);
  CUtensorMap tq, tk, tv, tdo, tq64, tk64, tv64, tdo64;
  if (fa_make_map(&tq, q, S, H, B, ldq, 128) || fa_make_map(&tk, k, S, H, B, ldq, 128) || fa_make_map(&tv, v, S, H, B, ldq, 128) ||
      fa_make_map(&tdo, dout, S, H, B, rs, 128) || fa_make_map(&tq64, q, S, H, B, ldq, 64) || fa_make_map(&tk64, k, S, H, B, ldq, 64) ||
      fa_make_map(&tv64, v, S, H, B, ldq, 64) || fa_make_map(&tdo64, dout, S, H, B, rs, 64)) return 1;
  FaBwdParams p;
  p.out = (const __nv_bfloat16*)out; p.dout = (const __nv_bfloat16*)dout; p.lse = (const float*)lse; p.delta = (float*)delta;
  p.dq = (__nv_bfloat16*)dq; p.dk = (__nv_bfloat16*)dk; p.dv = (__nv_bfloat16*)dv;
  p.B = B; p.S = S; p.H = H; p.row_stride = rs; p.ld_dqkv = ldd;
  p.scale = 1.0f / sqrtf((float)D);
  p.scale_log2 = p.scale * fa::kLog2e;
  p.kbias = (const float*)key_bias;
  if (fa_fill_drop(&p.drop, p_drop, seed, step_ptr, call_id)) return 1;
  if (fa_set_attrs()) return 1;
  cudaStream_t st = (cudaStream_t)stream;
  const long long dthreads = (long long)B * S * H * 8;
  attention_delta_kernel<<<(unsigned)((dthreads + 255) / 256), 256, 0, st>>>(p);
  MLT_LAUNCH_CHECK();
  dim3 grid((S + 127) / 128, H, B);
  const bool drop = p_drop > 0.f;
#define MLT_FA_BWD(BIAS, DROP)                                                                                       \
  do {                                                                                                               \
    attention_bwd_dq_kernel<BIAS, DROP><<<grid, fa::kThreads, fa::kSmemDq, st>>>(tq, tk64, tv64, tdo, p);            \
    MLT_LAUNCH_CHECK();                                                                                              \
    attention_bwd_dkdv_kernel<BIAS, DROP><<<grid, fa::kThreads, fa::kSmemDkv, st>>>(tq64, tk, tv, tdo64, p);         \
  } while (0)
  if (key_bias) { if (drop) MLT_FA_BWD(true, true); else MLT_FA_BWD(true, false); }
  else { if (drop) MLT_FA_BWD(false, true); else MLT_FA_BWD(false, false); }
#undef MLT_FA_BWD
  MLT_LAUNCH_CHECK();
  return 0;
}

extern "C" int mlt_attention_bwd_masked(const void* q, const void* k, const void* v, long long ld_qkv, const void* key_bias,
                                        const void* out, const void* dout, const void* lse, void* delta, void* dq, void* dk, void* dv,
                                        long long ld_dqkv, int B, int S, int H, int D, void* stream) {
  return mlt_attention_bwd_ex(q, k, v, ld_qkv, key_bias, out, dout, lse, delta, dq, dk, dv, ld_dqkv, B, S, H, D, 0.f, 0ull, nullptr, 0u,
                              stream);
}

extern "C" int mlt_attention_bwd(const void* q, const void* k, const void* v, long long ld_qkv, const void* out, const void* dout,
                                 const void* lse, void* delta, void* dq, void* dk, void* dv, long long ld_dqkv, int B, int S, int H,
                                 int D, void* stream) {
  return mlt_attention_bwd_masked(q, k, v, ld_qkv, nullptr, out, dout, lse, delta, dq, dk, dv, ld_dqkv, B, S, H, D, stream);
}

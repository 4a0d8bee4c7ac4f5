// bf16 GEMM on the 5th-generation tensor cores: D[M,N] = A[M,K] * B[N,K]^T (+ epilogue).
//
// Structure (one persistent CTA per SM, 192 threads, warp-specialised):
//   warp 0      TMA producer: cp.async.bulk.tensor loads of 128-byte-swizzled A/B tiles into a STAGES-deep smem ring,
//               completion counted on `full[s]` mbarriers.
//   warp 1      MMA issuer: one elected thread issues tcgen05.mma (M=128, N=BN, K=16) reading smem through UMMA
//               descriptors, accumulating fp32 in TMEM; tcgen05.commit frees smem slots (`empty[s]`) and publishes the
//               finished accumulator (`tmem_full[a]`).  Two accumulator stages: the epilogue of tile i overlaps the MMAs
//               of tile i+1.  This warp also owns the TMEM allocation.
//   warps 2-5   epilogue: tcgen05.ld the accumulator (each warp its own 32-lane quadrant), apply bias / activation /
//               activation-gradient, convert and store; then release the accumulator stage (`tmem_empty[a]`).
// Operands may be K-major (reduction dim contiguous: activations [M,K], nn.Linear weights [N,K]) or MN-major
// (reduction dim strided: what dgrad and wgrad see), selected per operand through the instruction descriptor's
// major bits and the matching TMA box / smem-descriptor geometry, so backward needs no transposes.
// Replaces the cuBLAS calls behind reference src/model.py:21-23 (fc layers) and every BERT / ResNet-FC GEMM.
#include "common.cuh"
#include "ptx.cuh"
#include "im2col_tma.h"
#include "../include/mlt_b200.h"
#include <mutex>
#include <unordered_map>

namespace mlt {

using namespace ptx;

constexpr int BM = 128;        // tile rows (UMMA M)
constexpr int BK = 64;         // k-block: 64 bf16 = 128 bytes = one swizzle row
constexpr int UMMA_K = 16;
constexpr int kEpiWarps = 8;       // two per TMEM lane quadrant, each takes half of the tile's columns
constexpr int kGemmThreads = 64 + kEpiWarps * 32;

struct GemmParams {
  int M, N, K;
  void* out;
  long long ldo;
  int out_f32;
  const void* bias;
  int bias_f32;
  int act;        // Act
  int aux_mode;   // 0 none | 1 store pre-activation (bf16) to aux | 2 multiply by act'(aux) | 3 store act'(pre) | 4 multiply by aux | 5 add aux
  void* aux;
  long long ld_aux;
  int a_mn, b_mn;
  int num_m, num_n, num_k;
  int nb1, nb2;                  // batch = nb1 * nb2 independent problems (attention: heads x samples)
  long long ob1, ob2;            // output batch strides (elements)
  int splits, kb_per_split;      // split-K: the batch index doubles as the K-slice id, partials go to an fp32 workspace
  // implicit-GEMM convolution: one operand is the patch matrix of an NHWC tensor, read through a TMA im2col map.
  //   1: A = patches [pixels x KH*KW*C], K-major (forward)      2: B = patches as the MN-major operand (weight gradient)
  //   3: data gradient of a stride-1 convolution: A = patches of dy (as 1), B = the weights [Co][tap][C] read as the
  //      MN-major operand with the tap flipped (batch coordinate of a 4-D tiled map)
  int im2col;
  int cv_OW, cv_OH, cv_stride, cv_pad, cv_KW, cv_taps, cv_C;
};

// AUX: the epilogue also writes a second bf16 tensor (pre-activation or activation derivative) through its own slabs and
// TMA stores; those 32 KB come out of the operand ring (3 stages instead of 4 at BN = 256).
// DB: short-K problems are bound by the output stream, not by the MMAs.  Every epilogue warp then gets TWO store slabs and
// alternates between them, so converting / writing the next 4 KB into shared memory overlaps the bulk store that is still
// reading the previous one (one slab serialises them: measured 4.5 TB/s on [802816 x 256] = [.. x 64] x [64 x 256], cuBLAS
// 6.6); the second set of slabs comes out of the operand ring, which a one- or two-k-block problem does not need that deep.
template <int BN, bool AUX, bool DB = false>
struct GemmCfg {
  static constexpr int kABytes = BM * BK * 2;                // 16 KB
  static constexpr int kBBytes = BN * BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = DB ? (BN == 256 ? 3 : (BN == 128 ? 5 : 6)) : (BN == 256 ? (AUX ? 3 : 4) : (BN == 128 ? 6 : 8));
  static constexpr int kTmemCols = 2 * BN;                   // two accumulator stages (power of two >= 32)
  static constexpr int kStagingBytes = kEpiWarps * 4096 * ((AUX || DB) ? 2 : 1);   // per epilogue warp: 32-row x 128-byte slab(s) for TMA stores
  static constexpr int kSmemBytes = kStages * kStageBytes + kStagingBytes + 1024 /*align slack*/ + 256 /*barriers*/;
};

__device__ __forceinline__ void tile_coords(int tile, int num_m, int num_n, int& m_blk, int& n_blk) {
  // groups of 8 M-blocks walk N together so concurrently running CTAs share A and B tiles in L2
  constexpr int G = 8;
  const int per_group = G * num_n;
  const int group = tile / per_group;
  const int first_m = group * G;
  const int gsize = min(G, num_m - first_m);
  const int r = tile - group * per_group;
  m_blk = first_m + r % gsize;
  n_blk = r / gsize;
}

__device__ __forceinline__ float act_grad_from_pre(int act, float z) {
  switch (act) {
    case ACT_RELU: return z > 0.f ? 1.f : 0.f;
    case ACT_GELU: return gelu_bwd(z);
    case ACT_TANH: { const float t = tanhf(z); return 1.f - t * t; }
    default: return 1.f;
  }
}

// One output tile's worth of epilogue work for one warp: rows [store_row0, +32) (this lane: `row`), 32-column chunks
// c_lo .. c_hi-1 of the BN-wide accumulator at `taddr`.  TMEM -> registers -> (bias / activation / aux) -> swizzled smem
// slab -> TMA store.  Waits on the accumulator's full barrier itself so the aux prefetch can be issued before it.
// PAIR (BN == 64, bf16 output): the two warps of a TMEM lane quadrant each take ONE 32-column chunk and fill the two 64-byte
// halves of the same slab rows; they meet on a named barrier and the first of them issues the store - all eight epilogue
// warps work on a 64-column tile instead of four.
__device__ __forceinline__ void pair_bar(int id) { asm volatile("bar.sync %0, 64;" ::"r"(id) : "memory"); }

template <int BN, bool AUX, bool DB>
__device__ __forceinline__ void epilogue_tile(const GemmParams& p, const CUtensorMap* tmOut, const CUtensorMap* tmAux, uint32_t taddr,
                                              int row, int store_row0, int n0, int z1, int z2, uint8_t* slab_base, uint8_t* aux_slab,
                                              int lane, int c_lo, int c_hi, uint64_t* full_bar, uint32_t full_parity, int& buf,
                                              int pair_id, int pair_half) {
  const int chunks_per_slab = p.out_f32 ? 1 : 2;          // a 32-column chunk is 128 B of fp32 or 64 B of bf16
  const bool pair = pair_id > 0;
  const bool aux_in = p.aux_mode == 2 || p.aux_mode == 4 || p.aux_mode == 5;
  // aux operand of the first chunk: requested before the accumulator is ready, then one chunk ahead
  const __nv_bfloat16* aux_row = reinterpret_cast<const __nv_bfloat16*>(p.aux) + (size_t)row * p.ld_aux;
  uint4 pa[4] = {};
  auto fetch_aux = [&](int c) {
    const int nb = n0 + c * 32;
    if (aux_in && row < p.M && nb + 32 <= p.N) {
#pragma unroll
      for (int j = 0; j < 4; ++j) pa[j] = *reinterpret_cast<const uint4*>(aux_row + nb + 8 * j);
    }
  };
  if (c_lo < c_hi) fetch_aux(c_lo);
  mbar_wait(full_bar, full_parity);
  tc_fence_after();
#pragma unroll 1
  for (int c = c_lo; c < c_hi; ++c) {
    uint32_t r[32];
    tmem_ld_32x32b_x32(taddr + c * 32, r);
    uint4 ca[4] = {pa[0], pa[1], pa[2], pa[3]};
    if (c + 1 < c_hi) fetch_aux(c + 1);
    tmem_ld_wait();
    const int nb = n0 + c * 32;
    const int within = pair ? pair_half : (c - c_lo) % chunks_per_slab;
    uint8_t* slab = slab_base + (DB ? buf * 4096 : 0);
    if (pair) {
      // the bulk store that last read this buffer must have drained it: the issuing warp checks, then both may write
      if (pair_half == 0 && lane == 0) tma_store_wait_read<DB ? 1 : 0>();
      pair_bar(pair_id);
    } else if (within == 0) {
      if (lane == 0) tma_store_wait_read<DB ? 1 : 0>();
      __syncwarp();
    }
    float v[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
    const bool live = nb < p.N;                          // whole chunk beyond N: nothing to compute, TMA clips it
    const bool full_chunk = nb + 32 <= p.N;
    if (live && p.bias != nullptr) {
      if (p.bias_f32) {
        const float* bp = reinterpret_cast<const float*>(p.bias) + nb;
#pragma unroll
        for (int j = 0; j < 32; ++j) if (full_chunk || nb + j < p.N) v[j] += __ldg(bp + j);
      } else if (full_chunk && (reinterpret_cast<uintptr_t>(p.bias) & 15) == 0) {
        const uint4* bp = reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(p.bias) + nb);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint4 raw = __ldg(bp + j);
          const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[e]));
            v[8 * j + 2 * e] += f.x; v[8 * j + 2 * e + 1] += f.y;
          }
        }
      } else {
        const __nv_bfloat16* bp = reinterpret_cast<const __nv_bfloat16*>(p.bias) + nb;
#pragma unroll
        for (int j = 0; j < 32; ++j) if (full_chunk || nb + j < p.N) v[j] += __bfloat162float(bp[j]);
      }
    }
    bool act_done = false;
    if ((AUX || (live && row < p.M)) && (p.aux_mode == 1 || p.aux_mode == 3)) {
      // keep for backward, with direct stores: the pre-activation (1) or the activation's derivative there (3)
      __nv_bfloat16* ap = reinterpret_cast<__nv_bfloat16*>(p.aux) + (size_t)row * p.ld_aux + nb;
      float s[32];
      if (p.aux_mode == 1) {
#pragma unroll
        for (int j = 0; j < 32; ++j) s[j] = v[j];
      } else if (p.act == ACT_GELU && !p.out_f32) {      // bf16 result: value and derivative from one tanh.approx
#pragma unroll
        for (int j = 0; j < 32; ++j) { float g, dg; gelu_both(v[j], g, dg); v[j] = g; s[j] = dg; }
        act_done = true;
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) s[j] = act_grad_from_pre(p.act, v[j]);
      }
      if (AUX) {                                         // swizzled slab, stored by TMA together with the output slab
        uint8_t* arow = aux_slab + lane * 128;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint4 pk;
          pk.x = pack_bf16(s[8 * j], s[8 * j + 1]); pk.y = pack_bf16(s[8 * j + 2], s[8 * j + 3]);
          pk.z = pack_bf16(s[8 * j + 4], s[8 * j + 5]); pk.w = pack_bf16(s[8 * j + 6], s[8 * j + 7]);
          *reinterpret_cast<uint4*>(arow + (((within * 4 + j) ^ (lane & 7)) << 4)) = pk;
        }
      } else if (full_chunk) {
#pragma unroll
        for (int j = 0; j < 32; j += 8) {
          uint4 pk;
          pk.x = pack_bf16(s[j], s[j + 1]); pk.y = pack_bf16(s[j + 2], s[j + 3]);
          pk.z = pack_bf16(s[j + 4], s[j + 5]); pk.w = pack_bf16(s[j + 6], s[j + 7]);
          *reinterpret_cast<uint4*>(ap + j) = pk;
        }
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) if (nb + j < p.N) ap[j] = __float2bfloat16_rn(s[j]);
      }
    }
    if (aux_in) {                                        // dX = (dY W) * act'(z): z (2) or act'(z) itself (4) in aux
      if (live && row < p.M) {
        float s[32];
        if (full_chunk) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const uint32_t w[4] = {ca[j].x, ca[j].y, ca[j].z, ca[j].w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[e]));
              s[8 * j + 2 * e] = f.x; s[8 * j + 2 * e + 1] = f.y;
            }
          }
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) s[j] = nb + j < p.N ? __bfloat162float(aux_row[nb + j]) : 0.f;
        }
        if (p.aux_mode == 4) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] *= s[j];
        } else if (p.aux_mode == 5) {                    // D = A B^T + aux (aux may be the output itself: a tile reads what it overwrites)
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] += s[j];
        } else if (p.act == ACT_GELU && !p.out_f32) {
#pragma unroll
          for (int j = 0; j < 32; ++j) { float g, dg; gelu_both(s[j], g, dg); v[j] *= dg; }
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] *= act_grad_from_pre(p.act, s[j]);
        }
      }
    } else if (p.act != ACT_NONE && !act_done) {
      if (p.act == ACT_GELU && !p.out_f32) {
#pragma unroll
        for (int j = 0; j < 32; ++j) { float g, dg; gelu_both(v[j], g, dg); v[j] = g; }
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = act_fwd_rt(p.act, v[j]);
      }
    }
    // registers -> swizzled slab: 16-byte chunk index is XORed with (row & 7), the layout SWIZZLE_128B expects
    uint8_t* srow = slab + lane * 128;
    if (p.out_f32) {
#pragma unroll
      for (int j = 0; j < 8; ++j)
        *reinterpret_cast<float4*>(srow + ((j ^ (lane & 7)) << 4)) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        uint4 pk;
        pk.x = pack_bf16(v[8 * j], v[8 * j + 1]); pk.y = pack_bf16(v[8 * j + 2], v[8 * j + 3]);
        pk.z = pack_bf16(v[8 * j + 4], v[8 * j + 5]); pk.w = pack_bf16(v[8 * j + 6], v[8 * j + 7]);
        *reinterpret_cast<uint4*>(srow + (((within * 4 + j) ^ (lane & 7)) << 4)) = pk;
      }
    }
    if (pair) {
      fence_proxy_async();
      pair_bar(pair_id);                                 // both halves of the slab rows are written
      if (pair_half == 0 && lane == 0) {
        tma_store_4d(tmOut, slab, n0, store_row0, z1, z2);
        tma_store_commit();
      }
      if (DB) buf ^= 1;
    } else if (within == chunks_per_slab - 1) {
      fence_proxy_async();                               // generic-proxy smem writes -> visible to the TMA engine
      __syncwarp();
      if (lane == 0) {
        const int col0 = nb - within * 32;
        tma_store_4d(tmOut, slab, col0, store_row0, z1, z2);
        if (AUX) tma_store_4d(tmAux, aux_slab, col0, store_row0, z1, z2);
        tma_store_commit();
      }
      if (DB) buf ^= 1;
    }
  }
}

// The lean epilogue: no bias, no stored / multiplied aux tensor, activation none or ReLU - every convolution GEMM and every
// data / weight gradient GEMM.  The generic epilogue above re-derives its options per 32-column chunk (~970 instructions per
// warp and 128 x 256 tile, measured; two epilogue warps per scheduler cannot hide that much dependent issue latency, so a
// short-K problem ran at the epilogue's pace: 113 us where the output stream needs 78).  Here a warp's whole share of the tile is
// one straight line: tcgen05.ld, cvt.rn.bf16x2 (or nothing for fp32), st.shared with swizzle offsets computed once, one bulk store
// per 4 KB slab.
__device__ __forceinline__ uint32_t cvt_bf16x2(float lo, float hi) {
  uint32_t d;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  return d;
}
__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

template <int BN, bool DB, bool F32>
__device__ __forceinline__ void epilogue_tile_lean(const CUtensorMap* tmOut, uint32_t taddr, int store_row0, int n0, int z1, int z2,
                                                   uint8_t* slab_base, int lane, int c_lo, int c_hi, uint64_t* full_bar,
                                                   uint32_t full_parity, int& buf, int pair_id, int pair_half, bool relu) {
  const uint32_t srow = smem_u32(slab_base) + lane * 128;
  const uint32_t sw = (uint32_t)(lane & 7);
  mbar_wait(full_bar, full_parity);
  tc_fence_after();
  if (pair_id > 0) {                                     // BN == 64, bf16: this warp owns ONE chunk = one 64-byte half of the slab rows
    uint32_t r[32];
    tmem_ld_32x32b_x32(taddr + c_lo * 32, r);
    if (pair_half == 0 && lane == 0) tma_store_wait_read<DB ? 1 : 0>();
    pair_bar(pair_id);
    tmem_ld_wait();
    const uint32_t base = srow + (DB ? buf * 4096 : 0);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float v[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) { v[e] = __uint_as_float(r[8 * j + e]); if (relu) v[e] = fmaxf(v[e], 0.f); }
      sts128(base + ((((uint32_t)(pair_half * 4 + j)) ^ sw) << 4), cvt_bf16x2(v[0], v[1]), cvt_bf16x2(v[2], v[3]), cvt_bf16x2(v[4], v[5]),
             cvt_bf16x2(v[6], v[7]));
    }
    fence_proxy_async();
    pair_bar(pair_id);
    if (pair_half == 0 && lane == 0) {
      tma_store_4d(tmOut, slab_base + (DB ? buf * 4096 : 0), n0, store_row0, z1, z2);
      tma_store_commit();
    }
    if (DB) buf ^= 1;
    return;
  }
  constexpr int kPerSlab = F32 ? 1 : 2;                  // 32-column chunks per 128-byte slab row
#pragma unroll 1
  for (int c = c_lo; c < c_hi; c += kPerSlab) {
    uint32_t r[32], r2[32];
    tmem_ld_32x32b_x32(taddr + c * 32, r);
    if (!F32) tmem_ld_32x32b_x32(taddr + (c + 1) * 32, r2);
    if (lane == 0) tma_store_wait_read<DB ? 1 : 0>();     // the bulk store that last read this buffer has drained it
    __syncwarp();
    tmem_ld_wait();
    const uint32_t base = srow + (DB ? buf * 4096 : 0);
    if (F32) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        uint32_t w[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) w[e] = relu ? __float_as_uint(fmaxf(__uint_as_float(r[4 * j + e]), 0.f)) : r[4 * j + e];
        sts128(base + (((uint32_t)j ^ sw) << 4), w[0], w[1], w[2], w[3]);
      }
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          v[e] = __uint_as_float(j < 4 ? r[8 * j + e] : r2[8 * (j - 4) + e]);
          if (relu) v[e] = fmaxf(v[e], 0.f);
        }
        sts128(base + (((uint32_t)j ^ sw) << 4), cvt_bf16x2(v[0], v[1]), cvt_bf16x2(v[2], v[3]), cvt_bf16x2(v[4], v[5]), cvt_bf16x2(v[6], v[7]));
      }
    }
    fence_proxy_async();
    __syncwarp();
    if (lane == 0) {
      tma_store_4d(tmOut, slab_base + (DB ? buf * 4096 : 0), n0 + c * 32, store_row0, z1, z2);
      tma_store_commit();
    }
    if (DB) buf ^= 1;
  }
}

template <int BN, bool AUX, bool DB>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                         const __grid_constant__ CUtensorMap tmOut, const __grid_constant__ CUtensorMap tmAux,
                         const GemmParams p) {
  using Cfg = GemmCfg<BN, AUX, DB>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* staging = smem + Cfg::kStages * Cfg::kStageBytes;                 // 1024-byte aligned (stage sizes are)
  uint64_t* full = reinterpret_cast<uint64_t*>(staging + Cfg::kStagingBytes);
  uint64_t* empty = full + Cfg::kStages;
  uint64_t* tmem_full = empty + Cfg::kStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles_per_problem = p.num_m * p.num_n;
  const int num_tiles = tiles_per_problem * p.nb1 * p.nb2;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    prefetch_tmap(&tmOut);
    for (int s = 0; s < Cfg::kStages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(&tmem_full[a], 1); mbar_init(&tmem_empty[a], kEpiWarps); }
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_base_slot, Cfg::kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_base_slot;

  if (warp == 0) {
    // ===================================== TMA producer =====================================
    // The whole warp walks the loop (warp-uniform control flow keeps operands in uniform registers); one elected lane issues.
    {
      int stage = 0; uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int m_blk, n_blk;
        const int z = tile / tiles_per_problem;
        tile_coords(tile - z * tiles_per_problem, p.num_m, p.num_n, m_blk, n_blk);
        const int z1 = p.splits > 1 ? 0 : z % p.nb1, z2 = p.splits > 1 ? 0 : z / p.nb1;
        const int kb_lo = p.splits > 1 ? z * p.kb_per_split : 0;
        const int kb_hi = p.splits > 1 ? min(p.num_k, kb_lo + p.kb_per_split) : p.num_k;
        const int m0 = m_blk * BM, n0 = n_blk * BN;
        for (int kb = kb_lo; kb < kb_hi; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          uint8_t* sA = smem + stage * Cfg::kStageBytes;
          uint8_t* sB = sA + Cfg::kABytes;
          const int k0 = kb * BK;
          if (elect_one()) {
          mbar_arrive_expect_tx(&full[stage], Cfg::kStageBytes);
          if (p.im2col == 1 || p.im2col == 3) {
            // k-block kb = (filter tap, 64-channel block); the 128 rows are consecutive output pixels starting at m0
            const int cblocks = p.cv_C >> 6;
            const int tap = kb / cblocks, c0 = (kb - tap * cblocks) << 6;
            const int fr = tap / p.cv_KW, fs = tap - fr * p.cv_KW;
            const int q0 = m0 % p.cv_OW, t = m0 / p.cv_OW;
            const int p0 = t % p.cv_OH, img = t / p.cv_OH;
            tma_load_im2col_4d(sA, &tmA, &full[stage], c0, q0 * p.cv_stride - p.cv_pad, p0 * p.cv_stride - p.cv_pad, img, fs, fr);
          } else if (!p.a_mn) {
            tma_load_4d(sA, &tmA, &full[stage], k0, m0, z1, z2);                       // box {64 k, 128 rows}
          } else {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j) tma_load_4d(sA + j * 8192, &tmA, &full[stage], m0 + 64 * j, k0, z1, z2);   // box {64 rows, 64 k}
          }
          if (p.im2col == 2) {
            // reduction index = output pixel (64 per k-block), N index = patch column = tap * C + channel
            const int q0 = k0 % p.cv_OW, t = k0 / p.cv_OW;
            const int p0 = t % p.cv_OH, img = t / p.cv_OH;
            const int w0 = q0 * p.cv_stride - p.cv_pad, h0 = p0 * p.cv_stride - p.cv_pad;
#pragma unroll
            for (int j = 0; j < BN / 64; ++j) {
              const int col = n0 + 64 * j;
              int tap = col / p.cv_C;
              const int c0 = col - tap * p.cv_C;
              tap = min(tap, p.cv_taps - 1);                 // columns >= N: any valid tap, the store clips them
              const int fr = tap / p.cv_KW, fs = tap - fr * p.cv_KW;
              tma_load_im2col_4d(sB + j * 8192, &tmB, &full[stage], c0, w0, h0, img, fs, fr);
            }
          } else if (p.im2col == 3) {
            const int cblocks = p.cv_C >> 6;
            const int tap = kb / cblocks, co0 = (kb - tap * cblocks) << 6;
#pragma unroll
            for (int j = 0; j < BN / 64; ++j) tma_load_4d(sB + j * 8192, &tmB, &full[stage], n0 + 64 * j, co0, p.cv_taps - 1 - tap, 0);
          } else if (!p.b_mn) {
            tma_load_4d(sB, &tmB, &full[stage], k0, n0, z1, z2);                       // box {64 k, BN rows}
          } else {
#pragma unroll
            for (int j = 0; j < BN / 64; ++j) tma_load_4d(sB + j * 8192, &tmB, &full[stage], n0 + 64 * j, k0, z1, z2);
          }
          }
          __syncwarp();
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================================== MMA issuer ========================================
    {
      const uint32_t idesc = make_idesc_bf16(BM, BN, p.a_mn, p.b_mn);
      // per-UMMA_K advance of the descriptor start address, and LBO/SBO, for each operand's layout
      const uint32_t a_step = p.a_mn ? (UMMA_K * 128) : (UMMA_K * 2);
      const uint32_t b_step = p.b_mn ? (UMMA_K * 128) : (UMMA_K * 2);
      const uint32_t a_lbo = p.a_mn ? 8192u : 16u, b_lbo = p.b_mn ? 8192u : 16u;
      int stage = 0; uint32_t phase = 0;
      int as = 0; uint32_t aphase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        mbar_wait(&tmem_empty[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        const int zs = tile / tiles_per_problem;
        const int kb_lo = p.splits > 1 ? zs * p.kb_per_split : 0;
        const int kb_hi = p.splits > 1 ? min(p.num_k, kb_lo + p.kb_per_split) : p.num_k;
        for (int kb = kb_lo; kb < kb_hi; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t sA = smem_u32(smem + stage * Cfg::kStageBytes);
          const uint32_t sB = sA + Cfg::kABytes;
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k) {
              const uint64_t adesc = make_smem_desc_sw128(sA + k * a_step, a_lbo, 1024);
              const uint64_t bdesc = make_smem_desc_sw128(sB + k * b_step, b_lbo, 1024);
              umma_f16(d_tmem, adesc, bdesc, idesc, (kb > kb_lo) || (k != 0));
            }
            umma_commit(&empty[stage]);                     // smem slot reusable once these MMAs retire
            if (kb == kb_hi - 1) umma_commit(&tmem_full[as]);
          }
          __syncwarp();
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
        }
        if (++as == 2) { as = 0; aphase ^= 1; }
      }
    }
  } else {
    // ===================================== epilogue ===========================================
    // TMEM -> registers -> (bias / activation) -> 128-byte-swizzled smem slab -> TMA store.  Eight warps: warp w reads
    // TMEM lane quadrant w % 4 (32 rows) and takes the column half (w - 2) / 4 of the tile, through one 4 KB slab
    // (32 rows x 128 bytes = 64 bf16 or 32 fp32 columns); TMA clips rows >= M and columns >= N.
    const int q = warp & 3;                                 // TMEM lane quadrant this warp may read
    const int half = (warp - 2) >> 2;
    // BN == 64 with a bf16 output: the quadrant's two warps share one slab (see epilogue_tile, PAIR)
    const bool pair = BN == 64 && !p.out_f32;
    uint8_t* slab = pair ? staging + q * 4096 * (DB ? 2 : 1) : staging + (warp - 2) * 4096 * (DB ? 2 : 1);
    uint8_t* aux_slab = staging + (kEpiWarps + warp - 2) * 4096;      // only present (and used) when AUX
    constexpr int kChunks = BN / 32;
    const int c_lo = BN >= 128 ? half * (kChunks / 2) : (pair ? half : 0);
    const int c_hi = BN >= 128 ? c_lo + kChunks / 2 : (pair ? half + 1 : (half == 0 ? kChunks : 0));
    int sbuf = 0;
    int as = 0; uint32_t aphase = 0;
    // lean path: nothing but (ReLU and) the store (TMA clips rows >= M and columns >= N, so ragged tiles need no care)
    const bool lean = !AUX && p.bias == nullptr && p.aux_mode == 0 && (p.act == ACT_NONE || p.act == ACT_RELU);
    const bool relu = p.act == ACT_RELU;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      int m_blk, n_blk;
      const int z = tile / tiles_per_problem;
      tile_coords(tile - z * tiles_per_problem, p.num_m, p.num_n, m_blk, n_blk);
      const int z1 = z % p.nb1, z2 = z / p.nb1;
      const int row = m_blk * BM + q * 32 + lane;
      const int n0 = n_blk * BN;
      const uint32_t tacc = tmem_base + as * BN + ((uint32_t)(q * 32) << 16);
      if (lean) {
        if (p.out_f32) epilogue_tile_lean<BN, DB, true>(&tmOut, tacc, m_blk * BM + q * 32, n0, z1, z2, slab, lane, c_lo, c_hi, &tmem_full[as],
                                                        aphase, sbuf, 0, half, relu);
        else epilogue_tile_lean<BN, DB, false>(&tmOut, tacc, m_blk * BM + q * 32, n0, z1, z2, slab, lane, c_lo, c_hi, &tmem_full[as], aphase,
                                               sbuf, pair ? 1 + q : 0, half, relu);
      } else
      epilogue_tile<BN, AUX, DB>(p, &tmOut, &tmAux, tacc, row, m_blk * BM + q * 32, n0, z1, z2,
                                 slab, aux_slab, lane, c_lo, c_hi, &tmem_full[as], aphase, sbuf, pair ? 1 + q : 0, half);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[as]);
      if (++as == 2) { as = 0; aphase ^= 1; }
    }
    if (lane == 0) tma_store_wait_all();                     // smem must outlive the last bulk stores
    __syncwarp();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, Cfg::kTmemCols);
}

// ---- CTA-pair variant (cta_group::2) -------------------------------------------------------------------------
// Two CTAs of a cluster (one TPC) own a 256 x 256 tile.  Each loads its own 128 rows of A and HALF of B (128 of the 256
// columns) - 32 KB per stage instead of 48, so six stages fit and the B operand crosses shared memory once per pair -
// and the leader's MMA warp issues tcgen05.mma.cta_group::2 (M = 256): the hardware reads both CTAs' shared memory and
// writes each CTA's 128 accumulator rows into that CTA's TMEM.  TMA completions of both CTAs are counted on the leader's
// `full` barrier; tcgen05.commit multicasts `empty` / `tmem_full` to both; the peer's epilogue warps arrive on the
// leader's `tmem_empty` through the cluster window.  The epilogue is the single-CTA one, each CTA on its own rows.
struct Gemm2Cfg {
  static constexpr int kABytes = BM * BK * 2;                 // 128 rows of A
  static constexpr int kBBytes = 128 * BK * 2;                // 128 of the tile's 256 B rows
  static constexpr int kStageBytes = kABytes + kBBytes;       // 32 KB
  static constexpr int kStages = 6;
  static constexpr int kTmemCols = 512;                       // two 256-column accumulator stages
  static constexpr int kStagingBytes = kEpiWarps * 4096;
  static constexpr int kSmemBytes = kStages * kStageBytes + kStagingBytes + 1024 + 256;
};

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGemmThreads, 1)
gemm_bf16_2cta_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                      const __grid_constant__ CUtensorMap tmOut, const GemmParams p) {
  using Cfg = Gemm2Cfg;
  constexpr int BN = 256, TM = 256;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* staging = smem + Cfg::kStages * Cfg::kStageBytes;
  uint64_t* full = reinterpret_cast<uint64_t*>(staging + Cfg::kStagingBytes);
  uint64_t* empty = full + Cfg::kStages;
  uint64_t* tmem_full = empty + Cfg::kStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;
  const int tiles_per_problem = p.num_m * p.num_n;
  const int num_tiles = tiles_per_problem * p.nb1 * p.nb2;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    prefetch_tmap(&tmOut);
    for (int s = 0; s < Cfg::kStages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(&tmem_full[a], 1); mbar_init(&tmem_empty[a], 2 * kEpiWarps); }
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc_2cta(tmem_base_slot, Cfg::kTmemCols);
    tmem_relinquish_2cta();
  }
  tc_fence_before();
  cluster_sync_all();                       // barriers of both CTAs initialised and visible, TMEM allocated in both
  tc_fence_after();
  const uint32_t tmem_base = *tmem_base_slot;

  if (warp == 0) {
    // ===================================== TMA producer (both CTAs) =====================================
    int stage = 0; uint32_t phase = 0;
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
      int m_blk, n_blk;
      const int z = tile / tiles_per_problem;
      tile_coords(tile - z * tiles_per_problem, p.num_m, p.num_n, m_blk, n_blk);
      const int z1 = p.splits > 1 ? 0 : z % p.nb1, z2 = p.splits > 1 ? 0 : z / p.nb1;
      const int kb_lo = p.splits > 1 ? z * p.kb_per_split : 0;
      const int kb_hi = p.splits > 1 ? min(p.num_k, kb_lo + p.kb_per_split) : p.num_k;
      const int m0 = m_blk * TM + (int)rank * BM, n0 = n_blk * BN + (int)rank * 128;
      for (int kb = kb_lo; kb < kb_hi; ++kb) {
        mbar_wait(&empty[stage], phase ^ 1);
        uint8_t* sA = smem + stage * Cfg::kStageBytes;
        uint8_t* sB = sA + Cfg::kABytes;
        const int k0 = kb * BK;
        if (elect_one()) {
          if (rank == 0) mbar_arrive_expect_tx(&full[stage], 2 * Cfg::kStageBytes);     // both CTAs' bytes
          if (!p.a_mn) {
            tma_load_4d_2cta(sA, &tmA, &full[stage], k0, m0, z1, z2);                   // box {64 k, 128 rows}
          } else {
#pragma unroll
            for (int j = 0; j < 2; ++j) tma_load_4d_2cta(sA + j * 8192, &tmA, &full[stage], m0 + 64 * j, k0, z1, z2);
          }
          if (!p.b_mn) {
            tma_load_4d_2cta(sB, &tmB, &full[stage], k0, n0, z1, z2);                   // box {64 k, 128 rows}
          } else {
#pragma unroll
            for (int j = 0; j < 2; ++j) tma_load_4d_2cta(sB + j * 8192, &tmB, &full[stage], n0 + 64 * j, k0, z1, z2);
          }
        }
        __syncwarp();
        if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ===================================== MMA issuer (leader CTA only) ===================================
    if (rank == 0) {
      const uint32_t idesc = make_idesc_bf16(TM, BN, p.a_mn, p.b_mn);
      const uint32_t a_step = p.a_mn ? (UMMA_K * 128) : (UMMA_K * 2);
      const uint32_t b_step = p.b_mn ? (UMMA_K * 128) : (UMMA_K * 2);
      const uint32_t a_lbo = p.a_mn ? 8192u : 16u, b_lbo = p.b_mn ? 8192u : 16u;
      int stage = 0; uint32_t phase = 0;
      int as = 0; uint32_t aphase = 0;
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
        mbar_wait(&tmem_empty[as], aphase ^ 1);          // both CTAs' epilogues have drained this accumulator stage
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        const int zs = tile / tiles_per_problem;
        const int kb_lo = p.splits > 1 ? zs * p.kb_per_split : 0;
        const int kb_hi = p.splits > 1 ? min(p.num_k, kb_lo + p.kb_per_split) : p.num_k;
        for (int kb = kb_lo; kb < kb_hi; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t sA = smem_u32(smem + stage * Cfg::kStageBytes);
          const uint32_t sB = sA + Cfg::kABytes;
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k) {
              const uint64_t adesc = make_smem_desc_sw128(sA + k * a_step, a_lbo, 1024);
              const uint64_t bdesc = make_smem_desc_sw128(sB + k * b_step, b_lbo, 1024);
              umma_f16_2cta(d_tmem, adesc, bdesc, idesc, (kb > kb_lo) || (k != 0));
            }
            umma_commit_2cta(&empty[stage], 3);            // the slot is free in BOTH CTAs once these MMAs retire
            if (kb == kb_hi - 1) umma_commit_2cta(&tmem_full[as], 3);
          }
          __syncwarp();
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
        }
        if (++as == 2) { as = 0; aphase ^= 1; }
      }
    }
  } else {
    // ===================================== epilogue (both CTAs, own 128 rows) ==============================
    const int q = warp & 3;
    const int half = (warp - 2) >> 2;
    uint8_t* slab = staging + (warp - 2) * 4096;
    constexpr int kChunks = BN / 32;
    const int c_lo = half * (kChunks / 2), c_hi = c_lo + kChunks / 2;
    int sbuf2 = 0;
    int as = 0; uint32_t aphase = 0;
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
      int m_blk, n_blk;
      const int z = tile / tiles_per_problem;
      tile_coords(tile - z * tiles_per_problem, p.num_m, p.num_n, m_blk, n_blk);
      const int z1 = z % p.nb1, z2 = z / p.nb1;
      const int row0 = m_blk * TM + (int)rank * BM + q * 32;
      epilogue_tile<BN, false, false>(p, &tmOut, &tmOut, tmem_base + as * BN + ((uint32_t)(q * 32) << 16), row0 + lane, row0, n_blk * BN, z1, z2,
                               slab, slab, lane, c_lo, c_hi, &tmem_full[as], aphase, sbuf2, 0, 0);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(&tmem_empty[as], 0);
      if (++as == 2) { as = 0; aphase ^= 1; }
    }
    if (lane == 0) tma_store_wait_all();
    __syncwarp();
  }
  tc_fence_before();
  cluster_sync_all();                       // no CTA leaves while its peer may still touch its smem / TMEM / barriers
  if (warp == 1) tmem_dealloc_2cta(tmem_base, Cfg::kTmemCols);
}

// ---- host side: TMA descriptors ---------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

// bf16 tensor [b2][b1][outer][inner] with strides (elements) ld / s1 / s2, box {box_inner, box_outer, 1, 1}, 128B swizzle.
static int make_tmap(CUtensorMap* out, const void* base, long long inner, long long outer, long long ld, int box_inner,
                     int box_outer, int nb1 = 1, long long s1 = 0, int nb2 = 1, long long s2 = 0, int elem_bytes = 2) {
  EncodeTiledFn enc = get_encode();
  MLT_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled is not available from the driver");
  // Driver-API calls need the primary context bound to THIS host thread; autograd runs backward on its own threads
  // that may not have made a runtime call yet.
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) { MLT_CHECK_CUDA(cudaFree(nullptr)); ctx_bound = true; }
  if (nb1 <= 1) { nb1 = 1; s1 = ld * outer; }
  if (nb2 <= 1) { nb2 = 1; s2 = s1 * nb1; }
  if (s1 % 8 != 0) s1 = (s1 + 7) / 8 * 8;          // only reachable for size-1 batch dims (stride never used)
  if (s2 % 8 != 0) s2 = (s2 + 7) / 8 * 8;
  const int eb = elem_bytes;
  MLT_REQUIRE(((uintptr_t)base % 16) == 0 && (ld * eb) % 16 == 0 && (s1 * eb) % 16 == 0 && (s2 * eb) % 16 == 0,
              "TMA needs a 16-byte aligned base and strides that are multiples of 16 bytes (ld=%lld s1=%lld s2=%lld elements of %d B)",
              ld, s1, s2, eb);
  cuuint64_t dims[4] = {(cuuint64_t)inner, (cuuint64_t)outer, (cuuint64_t)nb1, (cuuint64_t)nb2};
  cuuint64_t strides[3] = {(cuuint64_t)ld * eb, (cuuint64_t)s1 * eb, (cuuint64_t)s2 * eb};
  cuuint32_t box[4] = {(cuuint32_t)box_inner, (cuuint32_t)box_outer, 1, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = enc(out, eb == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  MLT_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d (inner=%lld outer=%lld ld=%lld box=%dx%d nb=%dx%d)",
              (int)r, inner, outer, ld, box_inner, box_outer, nb1, nb2);
  return 0;
}

template <int BN, bool AUX, bool DB = false>
static int launch_gemm_impl(const CUtensorMap& tmA, const CUtensorMap& tmB, GemmParams& p, cudaStream_t st) {
  using Cfg = GemmCfg<BN, AUX, DB>;
  // output map: rows of 128 bytes (64 bf16 / 32 fp32 columns) x 32 rows per store, batch dims as for the operands
  CUtensorMap tmOut;
  if (make_tmap(&tmOut, p.out, p.N, p.M, p.ldo, p.out_f32 ? 32 : 64, 32, p.nb1, p.ob1, p.nb2, p.ob2, p.out_f32 ? 4 : 2)) return 1;
  p.num_m = ceil_div(p.M, BM);
  p.num_n = ceil_div(p.N, BN);
  p.num_k = ceil_div(p.K, BK);
  CUtensorMap tmAux = tmOut;
  if (AUX && make_tmap(&tmAux, p.aux, p.N, p.M, p.ld_aux, 64, 32, p.nb1, 0, p.nb2, 0, 2)) return 1;
  static bool attr_set[64] = {};             // function attributes are per device
  auto kern = gemm_bf16_tcgen05_kernel<BN, AUX, DB>;
  int dev = 0;
  MLT_CHECK_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63]) {
    MLT_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    attr_set[dev & 63] = true;
  }
  long long tiles = (long long)p.num_m * p.num_n * p.nb1 * p.nb2;      // with split-K, nb1 carries the number of K slices
  int grid = tiles > kNumSMs ? kNumSMs : (int)tiles;
  kern<<<grid, kGemmThreads, Cfg::kSmemBytes, st>>>(tmA, tmB, tmOut, tmAux, p);
  MLT_LAUNCH_CHECK();
  return 0;
}

// 0 = never, 1 = whenever the shape allows (default).  MLT_GEMM_2CTA overrides.
static int gemm_2cta_mode() {
  static int mode = -1;
  if (mode < 0) {
    const char* e = getenv("MLT_GEMM_2CTA");
    mode = e ? atoi(e) : 1;
  }
  return mode;
}

static int gemm_2cta_min_k() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("MLT_GEMM_2CTA_MIN_K");
    v = e ? atoi(e) : 512;
  }
  return v;
}

static int launch_gemm_2cta(const CUtensorMap& tmA, const CUtensorMap& tmB, GemmParams& p, cudaStream_t st) {
  using Cfg = Gemm2Cfg;
  CUtensorMap tmOut;
  if (make_tmap(&tmOut, p.out, p.N, p.M, p.ldo, p.out_f32 ? 32 : 64, 32, p.nb1, p.ob1, p.nb2, p.ob2, p.out_f32 ? 4 : 2)) return 1;
  p.num_m = ceil_div(p.M, 256);
  p.num_n = ceil_div(p.N, 256);
  p.num_k = ceil_div(p.K, BK);
  static bool attr_set[64] = {};
  int dev = 0;
  MLT_CHECK_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63]) {
    MLT_CHECK_CUDA(cudaFuncSetAttribute(gemm_bf16_2cta_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    attr_set[dev & 63] = true;
  }
  long long tiles = (long long)p.num_m * p.num_n * p.nb1 * p.nb2;
  int clusters = tiles > kNumSMs / 2 ? kNumSMs / 2 : (int)tiles;
  gemm_bf16_2cta_kernel<<<2 * clusters, kGemmThreads, Cfg::kSmemBytes, st>>>(tmA, tmB, tmOut, p);
  MLT_LAUNCH_CHECK();
  return 0;
}

template <int BN>
static int launch_gemm(const CUtensorMap& tmA, const CUtensorMap& tmB, GemmParams& p, cudaStream_t st) {
  // the slab path for a stored aux tensor: bf16 output (so output and aux slabs cover the same 64 columns), one problem
  const bool aux_slab = BN == 256 && (p.aux_mode == 1 || p.aux_mode == 3) && !p.out_f32 && p.nb1 * p.nb2 == 1 &&
                        ((uintptr_t)p.aux % 16) == 0 && (p.ld_aux * 2) % 16 == 0;
  if (BN == 256 && aux_slab) return launch_gemm_impl<256, true>(tmA, tmB, p, st);
  // short reductions (<= 4 k-blocks per tile): the output stream is the bound -> double-buffered store slabs
  const int kblocks = p.splits > 1 ? p.kb_per_split : ceil_div(p.K, BK);
  static const int db_max_k = getenv("MLT_GEMM_DB_MAX_KBLOCKS") ? atoi(getenv("MLT_GEMM_DB_MAX_KBLOCKS")) : 4;
  if (kblocks <= db_max_k) return launch_gemm_impl<BN, false, true>(tmA, tmB, p, st);
  return launch_gemm_impl<BN, false>(tmA, tmB, p, st);
}

// out[m][n] = sum_s ws[s][m][n]   (split-K epilogue; fp32 partials -> bf16 or fp32 output with row pitch ldo)
template <typename O>
__global__ void __launch_bounds__(256)
splitk_reduce_kernel(const float* __restrict__ ws, O* __restrict__ out, long long ldo, int M, int N, int splits) {
  const long long total4 = (long long)M * N / 4;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long plane = (long long)M * N;
  for (long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x; v < total4; v += stride) {
    const float* p = ws + v * 4;
    float4 a = *reinterpret_cast<const float4*>(p);
    int s = 1;
    for (; s + 3 < splits; s += 4) {                    // four partial planes per round, their loads issued together
      const float4 b0 = *reinterpret_cast<const float4*>(p + (s + 0) * plane);
      const float4 b1 = *reinterpret_cast<const float4*>(p + (s + 1) * plane);
      const float4 b2 = *reinterpret_cast<const float4*>(p + (s + 2) * plane);
      const float4 b3 = *reinterpret_cast<const float4*>(p + (s + 3) * plane);
      a.x += b0.x; a.y += b0.y; a.z += b0.z; a.w += b0.w;      // (same order of additions as one plane at a time)
      a.x += b1.x; a.y += b1.y; a.z += b1.z; a.w += b1.w;
      a.x += b2.x; a.y += b2.y; a.z += b2.z; a.w += b2.w;
      a.x += b3.x; a.y += b3.y; a.z += b3.z; a.w += b3.w;
    }
    for (; s < splits; ++s) {
      const float4 b = *reinterpret_cast<const float4*>(p + s * plane);
      a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
    }
    const long long e = v * 4;
    const int m = (int)(e / N), n = (int)(e - (long long)m * N);
    O* o = out + (size_t)m * ldo + n;                   // N % 4 == 0 and a 16-byte row pitch: 4 outputs never straddle a row
    if constexpr (sizeof(O) == 2) {
      uint2 pk;
      pk.x = pack_bf16(a.x, a.y); pk.y = pack_bf16(a.z, a.w);
      *reinterpret_cast<uint2*>(o) = pk;
    } else {
      *reinterpret_cast<float4*>(o) = a;
    }
  }
}

static int launch_splitk_reduce(const float* ws, void* out, long long ldo, int out_f32, int M, int N, int splits, cudaStream_t st) {
  MLT_REQUIRE(N % 4 == 0, "split-K reduce needs N %% 4 == 0 (N=%d)", N);
  long long blocks = ceil_div64((long long)M * N / 4, 256);
  if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
  if (out_f32) splitk_reduce_kernel<float><<<(int)blocks, 256, 0, st>>>(ws, (float*)out, ldo, M, N, splits);
  else splitk_reduce_kernel<__nv_bfloat16><<<(int)blocks, 256, 0, st>>>(ws, (__nv_bfloat16*)out, ldo, M, N, splits);
  MLT_LAUNCH_CHECK();
  return 0;
}

}  // namespace mlt

using namespace mlt;

// D[M,N] = A * B^T with fp32 accumulation.
//   a_mn = 0: A is [M][K] (row stride lda);   a_mn = 1: A is stored [K][M] (row stride lda)
//   b_mn = 0: B is [N][K] (row stride ldb);   b_mn = 1: B is stored [K][N] (row stride ldb)
extern "C" int mlt_gemm_bf16(const void* A, long long lda, int a_mn, const void* B, long long ldb, int b_mn, void* out,
                             long long ldo, int out_f32, int M, int N, int K, const void* bias, int bias_f32, int act,
                             int aux_mode, void* aux, long long ld_aux, int block_n, void* splitk_ws,
                             long long splitk_ws_bytes, void* stream) {
  MLT_REQUIRE(M > 0 && N > 0 && K > 0, "mlt_gemm_bf16: empty problem (M=%d N=%d K=%d)", M, N, K);
  MLT_REQUIRE(aux_mode == 0 || aux != nullptr, "mlt_gemm_bf16: aux_mode %d needs an aux tensor", aux_mode);
  MLT_REQUIRE(((uintptr_t)out % 16) == 0 && (ldo * (out_f32 ? 4 : 2)) % 16 == 0, "mlt_gemm_bf16: output must be 16-byte aligned with a 16-byte multiple row pitch");
  MLT_REQUIRE(aux_mode == 0 || (((uintptr_t)aux % 16) == 0 && (ld_aux * 2) % 16 == 0), "mlt_gemm_bf16: aux must be 16-byte aligned");
  int BN = block_n;
  if (BN == 0) BN = (N >= 512 || (N % 256 == 0)) ? 256 : (N > 64 ? 128 : 64);
  MLT_REQUIRE(BN == 64 || BN == 128 || BN == 256 || BN == 512, "mlt_gemm_bf16: block_n must be 0, 64, 128, 256 or 512 (= 256 on a CTA pair)");
  // CTA-pair kernel (256 x 256 tiles, cta_group::2): block_n 512 forces it, block_n 0 picks it when the shape fills the pairs
  bool pair = BN == 512;
  if (BN == 512) BN = 256;
  // (few-tile problems go to split-K on single CTAs: twice as many units to spread the K slices over)
  if (block_n == 0 && BN == 256 && gemm_2cta_mode() == 1 && N >= 256 && K >= gemm_2cta_min_k() && ceil_div(M, 256) * ceil_div(N, 256) >= 32 &&
      aux_mode != 1 && aux_mode != 3)
    pair = true;
  if (pair) MLT_REQUIRE(aux_mode != 1 && aux_mode != 3, "mlt_gemm_bf16: the CTA-pair kernel has no stored-aux epilogue");
  const int TM = pair ? 256 : BM;
  CUtensorMap tmA, tmB;
  if (!a_mn) { if (make_tmap(&tmA, A, K, M, lda, BK, BM)) return 1; }
  else       { if (make_tmap(&tmA, A, M, K, lda, 64, BK)) return 1; }
  if (!b_mn) { if (make_tmap(&tmB, B, K, N, ldb, BK, pair ? 128 : BN)) return 1; }
  else       { if (make_tmap(&tmB, B, N, K, ldb, 64, BK)) return 1; }
  GemmParams p;
  p.M = M; p.N = N; p.K = K;
  p.out = out; p.ldo = ldo; p.out_f32 = out_f32;
  p.bias = bias; p.bias_f32 = bias_f32; p.act = act;
  p.aux_mode = aux_mode; p.aux = aux; p.ld_aux = ld_aux;
  p.a_mn = a_mn; p.b_mn = b_mn;
  p.nb1 = p.nb2 = 1; p.ob1 = p.ob2 = 0;
  p.splits = 1; p.kb_per_split = 0;
  p.im2col = 0;
  cudaStream_t st = (cudaStream_t)stream;
  // split-K: weight-gradient shapes (few output tiles, very long K) would otherwise occupy a handful of SMs
  const int tiles = ceil_div(M, TM) * ceil_div(N, BN), num_k = ceil_div(K, BK);
  const int units = pair ? kNumSMs / 2 : kNumSMs;          // CTAs, or CTA pairs
  int splits = 1;
  if (splitk_ws != nullptr && bias == nullptr && act == ACT_NONE && aux_mode == 0 && tiles * 2 <= units && num_k >= 8) {
    splits = units / tiles;
    if (splits > 128) splits = 128;                 // tiny weight gradients (64 x 64 over 800 K pixels): one K slice per SM
    if (splits > num_k / 8) splits = num_k / 8;     // at least 8 k-blocks (512 reduction rows) per slice
    if ((long long)splits * M * N * 4 > splitk_ws_bytes) splits = (int)(splitk_ws_bytes / ((long long)M * N * 4));
  }
  if (splits > 1) {
    const int kbps = ceil_div(num_k, splits);
    splits = ceil_div(num_k, kbps);              // no empty slice
    GemmParams q = p;
    q.out = splitk_ws; q.ldo = N; q.out_f32 = 1;
    q.nb1 = splits; q.nb2 = 1; q.ob1 = (long long)M * N; q.ob2 = 0;
    q.splits = splits; q.kb_per_split = kbps;
    int rc = pair ? launch_gemm_2cta(tmA, tmB, q, st)
                  : (BN == 256 ? launch_gemm<256>(tmA, tmB, q, st) : (BN == 128 ? launch_gemm<128>(tmA, tmB, q, st) : launch_gemm<64>(tmA, tmB, q, st)));
    if (rc) return rc;
    return launch_splitk_reduce((const float*)splitk_ws, out, ldo, out_f32, M, N, splits, st);
  }
  if (pair) return launch_gemm_2cta(tmA, tmB, p, st);
  if (BN == 256) return launch_gemm<256>(tmA, tmB, p, st);
  if (BN == 128) return launch_gemm<128>(tmA, tmB, p, st);
  return launch_gemm<64>(tmA, tmB, p, st);
}

// Batched variant: nb1 * nb2 independent problems that share M, N, K.  Operand z = (z1, z2) starts at
// base + z1 * s1 + z2 * s2 (elements); two batch strides cover attention's [batch, seq, head, dim] layouts without copies.
extern "C" int mlt_gemm_bf16_batched(const void* A, long long lda, long long a_s1, long long a_s2, int a_mn, const void* B,
                                     long long ldb, long long b_s1, long long b_s2, int b_mn, void* out, long long ldo,
                                     long long o_s1, long long o_s2, int out_f32, int M, int N, int K, int nb1, int nb2,
                                     int block_n, void* stream) {
  MLT_REQUIRE(M > 0 && N > 0 && K > 0 && nb1 > 0 && nb2 > 0, "mlt_gemm_bf16_batched: empty problem");
  const int esz = out_f32 ? 4 : 2;
  MLT_REQUIRE(((uintptr_t)out % 16) == 0 && (ldo * esz) % 16 == 0 && (o_s1 * esz) % 16 == 0 && (o_s2 * esz) % 16 == 0,
              "mlt_gemm_bf16_batched: output base and strides must be 16-byte multiples");
  int BN = block_n;
  if (BN == 0) BN = (N >= 512 || (N % 256 == 0)) ? 256 : (N > 64 ? 128 : 64);
  MLT_REQUIRE(BN == 64 || BN == 128 || BN == 256, "mlt_gemm_bf16_batched: block_n must be 0, 64, 128 or 256");
  CUtensorMap tmA, tmB;
  if (!a_mn) { if (make_tmap(&tmA, A, K, M, lda, BK, BM, nb1, a_s1, nb2, a_s2)) return 1; }
  else       { if (make_tmap(&tmA, A, M, K, lda, 64, BK, nb1, a_s1, nb2, a_s2)) return 1; }
  if (!b_mn) { if (make_tmap(&tmB, B, K, N, ldb, BK, BN, nb1, b_s1, nb2, b_s2)) return 1; }
  else       { if (make_tmap(&tmB, B, N, K, ldb, 64, BK, nb1, b_s1, nb2, b_s2)) return 1; }
  GemmParams p;
  p.M = M; p.N = N; p.K = K;
  p.out = out; p.ldo = ldo; p.out_f32 = out_f32;
  p.bias = nullptr; p.bias_f32 = 0; p.act = ACT_NONE;
  p.aux_mode = 0; p.aux = nullptr; p.ld_aux = 0;
  p.a_mn = a_mn; p.b_mn = b_mn;
  p.nb1 = nb1; p.nb2 = nb2; p.ob1 = o_s1; p.ob2 = o_s2;
  p.splits = 1; p.kb_per_split = 0;
  p.im2col = 0;
  cudaStream_t st = (cudaStream_t)stream;
  if (BN == 256) return launch_gemm<256>(tmA, tmB, p, st);
  if (BN == 128) return launch_gemm<128>(tmA, tmB, p, st);
  return launch_gemm<64>(tmA, tmB, p, st);
}

// ---- implicit-GEMM convolution (NHWC bf16, Cin a multiple of 64) -------------------------------------------------
static void conv_params(GemmParams& p, int mode, int H, int W, int C, int KH, int KW, int stride, int pad) {
  p.im2col = mode;
  p.cv_OH = (H + 2 * pad - KH) / stride + 1;
  p.cv_OW = (W + 2 * pad - KW) / stride + 1;
  p.cv_stride = stride; p.cv_pad = pad; p.cv_KW = KW; p.cv_taps = KH * KW; p.cv_C = C;
}

// y[B*OH*OW, Co] = patches(x) * w2^T, the patch matrix never materialised: the A tiles come straight out of x through a
// TMA im2col map (one load per filter tap and 64-channel block).  w2: [Co, KH*KW*C] with row pitch ldw.
extern "C" int mlt_conv_fwd_implicit(const void* x, int B, int H, int W, int C, const void* w2, long long ldw, int Co, int KH, int KW,
                                     int stride, int pad, void* y, long long ldy, void* stream) {
  MLT_REQUIRE(C % 64 == 0 && C > 0, "mlt_conv_fwd_implicit: Cin (%d) must be a multiple of 64", C);
  GemmParams p;
  conv_params(p, 1, H, W, C, KH, KW, stride, pad);
  const long long M = (long long)B * p.cv_OH * p.cv_OW;
  MLT_REQUIRE(M > 0 && M < (1LL << 31), "mlt_conv_fwd_implicit: bad output size");
  const int K = KH * KW * C, N = Co;
  int BN = (N >= 512 || (N % 256 == 0)) ? 256 : (N > 64 ? 128 : 64);
  CUtensorMap tmA, tmB;
  if (make_im2col_tmap(&tmA, x, B, H, W, C, KH, KW, stride, pad, 64, BM)) return 1;
  if (make_tmap(&tmB, w2, K, N, ldw, BK, BN)) return 1;
  p.M = (int)M; p.N = N; p.K = K;
  p.out = y; p.ldo = ldy; p.out_f32 = 0;
  p.bias = nullptr; p.bias_f32 = 0; p.act = ACT_NONE; p.aux_mode = 0; p.aux = nullptr; p.ld_aux = 0;
  p.a_mn = 0; p.b_mn = 0;
  p.nb1 = p.nb2 = 1; p.ob1 = p.ob2 = 0; p.splits = 1; p.kb_per_split = 0;
  cudaStream_t st = (cudaStream_t)stream;
  if (BN == 256) return launch_gemm<256>(tmA, tmB, p, st);
  if (BN == 128) return launch_gemm<128>(tmA, tmB, p, st);
  return launch_gemm<64>(tmA, tmB, p, st);
}

// dW[Co, KH*KW*C] = dy^T * patches(x): dy [pixels, Co] is the MN-major A operand, the patches are the MN-major B operand
// read through a TMA im2col map (64 pixels x 64 channels per load).  Split-K over the pixels as for any weight gradient.
extern "C" int mlt_conv_wgrad_implicit(const void* dy, long long lddy, const void* x, int B, int H, int W, int C, int Co, int KH, int KW,
                                       int stride, int pad, void* dw, long long lddw, int out_f32, void* splitk_ws,
                                       long long splitk_ws_bytes, void* stream) {
  MLT_REQUIRE(C % 64 == 0 && C > 0, "mlt_conv_wgrad_implicit: Cin (%d) must be a multiple of 64", C);
  GemmParams p;
  conv_params(p, 2, H, W, C, KH, KW, stride, pad);
  const long long pixels = (long long)B * p.cv_OH * p.cv_OW;
  MLT_REQUIRE(pixels > 0 && pixels < (1LL << 31), "mlt_conv_wgrad_implicit: bad output size");
  const int M = Co, N = KH * KW * C, K = (int)pixels;
  int BN = (N >= 512 || (N % 256 == 0)) ? 256 : (N > 64 ? 128 : 64);
  CUtensorMap tmA, tmB;
  if (make_tmap(&tmA, dy, M, K, lddy, 64, BK)) return 1;
  if (make_im2col_tmap(&tmB, x, B, H, W, C, KH, KW, stride, pad, 64, 64)) return 1;
  p.M = M; p.N = N; p.K = K;
  p.out = dw; p.ldo = lddw; p.out_f32 = out_f32;
  p.bias = nullptr; p.bias_f32 = 0; p.act = ACT_NONE; p.aux_mode = 0; p.aux = nullptr; p.ld_aux = 0;
  p.a_mn = 1; p.b_mn = 1;
  p.nb1 = p.nb2 = 1; p.ob1 = p.ob2 = 0; p.splits = 1; p.kb_per_split = 0;
  cudaStream_t st = (cudaStream_t)stream;
  const int tiles = ceil_div(M, BM) * ceil_div(N, BN), num_k = ceil_div(K, BK);
  int splits = 1;
  if (splitk_ws != nullptr && tiles * 2 <= kNumSMs && num_k >= 8) {
    splits = kNumSMs / tiles;
    if (splits > 128) splits = 128;                 // tiny weight gradients (64 x 64 over 800 K pixels): one K slice per SM
    if (splits > num_k / 8) splits = num_k / 8;     // at least 8 k-blocks (512 reduction rows) per slice
    if ((long long)splits * M * N * 4 > splitk_ws_bytes) splits = (int)(splitk_ws_bytes / ((long long)M * N * 4));
  }
  if (splits > 1) {
    const int kbps = ceil_div(num_k, splits);
    splits = ceil_div(num_k, kbps);
    GemmParams q = p;
    q.out = splitk_ws; q.ldo = N; q.out_f32 = 1;
    q.nb1 = splits; q.nb2 = 1; q.ob1 = (long long)M * N; q.ob2 = 0;
    q.splits = splits; q.kb_per_split = kbps;
    int rc = BN == 256 ? launch_gemm<256>(tmA, tmB, q, st) : (BN == 128 ? launch_gemm<128>(tmA, tmB, q, st) : launch_gemm<64>(tmA, tmB, q, st));
    if (rc) return rc;
    return launch_splitk_reduce((const float*)splitk_ws, dw, lddw, out_f32, M, N, splits, st);
  }
  if (BN == 256) return launch_gemm<256>(tmA, tmB, p, st);
  if (BN == 128) return launch_gemm<128>(tmA, tmB, p, st);
  return launch_gemm<64>(tmA, tmB, p, st);
}

// dx[B*H*W, C] of a STRIDE-1 convolution as an implicit GEMM: dx = patches(dy) * Wf^T, where the patches of dy use padding
// KH-1-pad and the filter is read flipped.  Replaces the "dy * W -> [pixels, KH*KW*C] -> col2im" round trip.
// dy: [B, OH, OW, Co] bf16 (Co a multiple of 64); w2: [Co, KH*KW, C] contiguous (the channels-last weight).
extern "C" int mlt_conv_dgrad_implicit(const void* dy, int B, int OH, int OW, int Co, const void* w2, int C, int KH, int KW, int pad,
                                       void* dx, long long lddx, void* stream) {
  MLT_REQUIRE(Co % 64 == 0 && Co > 0, "mlt_conv_dgrad_implicit: Cout (%d) must be a multiple of 64", Co);
  MLT_REQUIRE(KH - 1 - pad >= 0 && KW - 1 - pad >= 0, "mlt_conv_dgrad_implicit: padding larger than the filter");
  GemmParams p;
  // patches of dy: "input" OH x OW with Co channels, filter KH x KW, stride 1, padding KH-1-pad  ->  output H x W
  conv_params(p, 3, OH, OW, Co, KH, KW, 1, KH - 1 - pad);
  MLT_REQUIRE(KH == KW, "mlt_conv_dgrad_implicit: square filters only");
  const long long M = (long long)B * p.cv_OH * p.cv_OW;
  MLT_REQUIRE(M > 0 && M < (1LL << 31), "mlt_conv_dgrad_implicit: bad size");
  const int K = KH * KW * Co, N = C;
  int BN = (N >= 512 || (N % 256 == 0)) ? 256 : (N > 64 ? 128 : 64);
  CUtensorMap tmA, tmB;
  if (make_im2col_tmap(&tmA, dy, B, OH, OW, Co, KH, KW, 1, KH - 1 - pad, 64, BM)) return 1;
  // weights as [tap][Co][C]: inner = C, rows = Co with pitch KH*KW*C, batch = tap with stride C
  if (make_tmap(&tmB, w2, C, Co, (long long)KH * KW * C, 64, BK, KH * KW, C)) return 1;
  p.M = (int)M; p.N = N; p.K = K;
  p.out = dx; p.ldo = lddx; p.out_f32 = 0;
  p.bias = nullptr; p.bias_f32 = 0; p.act = ACT_NONE; p.aux_mode = 0; p.aux = nullptr; p.ld_aux = 0;
  p.a_mn = 0; p.b_mn = 1;
  p.nb1 = p.nb2 = 1; p.ob1 = p.ob2 = 0; p.splits = 1; p.kb_per_split = 0;
  cudaStream_t st = (cudaStream_t)stream;
  if (BN == 256) return launch_gemm<256>(tmA, tmB, p, st);
  if (BN == 128) return launch_gemm<128>(tmA, tmB, p, st);
  return launch_gemm<64>(tmA, tmB, p, st);
}

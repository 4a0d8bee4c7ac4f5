// Fused multi-head attention forward on tcgen05: softmax(Q K^T / sqrt(d)) V without the [S, S] score
// matrix ever touching HBM.  Head dim 64, sequence length a multiple of 128 up to 512 (BERT-base: 512), no mask.
//
// One CTA per (128-query tile, head, sample); 192 threads:
//   warp 0     TMA: the Q tile, then all K and V chunks (128 keys each) of this head, every chunk on its own mbarrier.
//   warp 1     MMA: S_j = Q K_j^T (M=128, N=128, K=64) into one of two TMEM score buffers, and O += P_j V_j
//              (M=128, N=64, K=128; V is the MN-major B operand straight from its [key][d] layout).  S_{j+1} is issued
//              before P_j is needed, so tensor cores and the softmax warps overlap.
//   warps 2-5  online softmax, one query row per thread: tcgen05.ld the scores, running max / sum in registers,
//              rescale the O accumulator in TMEM (tcgen05.ld/st) when the max moves, write P_j as bf16 into 128-byte
//              swizzled shared memory (the A operand of the PV MMA).  At the end O / l goes out as bf16 and the
//              log-sum-exp per row is kept for backward.
// TMEM: columns [0,128) and [128,256) score buffers, [256,320) the O accumulator (512 allocated, one CTA per SM).
#include "common.cuh"
#include "ptx.cuh"
#include "../include/mlt_b200.h"
#include <mutex>

namespace mlt {

using namespace ptx;

namespace fa {
constexpr int BQ = 128, BKEY = 128, D = 64;
constexpr int kThreads = 192;
constexpr int kQBytes = BQ * D * 2;            // 16 KB
constexpr int kChunkBytes = BKEY * D * 2;      // 16 KB per K or V chunk
constexpr int kPBytes = BQ * BKEY * 2;         // 32 KB: two [128 x 64] K-major sub-tiles
constexpr int kMaxChunks = 4;                  // S <= 512
constexpr int kSmemBytes = kQBytes + 2 * kMaxChunks * kChunkBytes + kPBytes + 1024 + 256;
constexpr uint32_t kTmemCols = 512;
constexpr uint32_t kColS = 0, kColO = 256;
}  // namespace fa

struct FaParams {
  __nv_bfloat16* out;        // [B, S, H*64]
  float* lse;                // [B, H, S]  (natural-log sum-exp of the scaled scores)
  int S, H;
  long long row_stride;      // H * 64
  float scale_log2;          // (1/sqrt(d)) * log2(e)
  float scale;
};

__global__ void __launch_bounds__(fa::kThreads, 1)
attention_fwd_tcgen05_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                             const __grid_constant__ CUtensorMap tmV, const FaParams p) {
  using namespace fa;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + kQBytes;
  uint8_t* sV = sK + kMaxChunks * kChunkBytes;
  uint8_t* sP = sV + kMaxChunks * kChunkBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + kPBytes);
  uint64_t* q_full = bars;                 // 1
  uint64_t* k_full = bars + 1;             // kMaxChunks
  uint64_t* v_full = k_full + kMaxChunks;  // kMaxChunks
  uint64_t* s_full = v_full + kMaxChunks;  // 2
  uint64_t* s_free = s_full + 2;           // 2
  uint64_t* p_full = s_free + 2;           // 1
  uint64_t* p_free = p_full + 1;           // 1
  uint64_t* o_done = p_free + 1;           // 1
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_done + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int qt = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int NK = p.S / BKEY;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmQ); prefetch_tmap(&tmK); prefetch_tmap(&tmV);
    mbar_init(q_full, 1);
    for (int j = 0; j < kMaxChunks; ++j) { mbar_init(&k_full[j], 1); mbar_init(&v_full[j], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&s_full[i], 1); mbar_init(&s_free[i], 128); }
    mbar_init(p_full, 128);
    mbar_init(p_free, 1);
    mbar_init(o_done, 1);
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      mbar_arrive_expect_tx(q_full, kQBytes);
      tma_load_4d(sQ, &tmQ, q_full, 0, qt * BQ, h, b);
      for (int j = 0; j < NK; ++j) {
        mbar_arrive_expect_tx(&k_full[j], kChunkBytes);
        tma_load_4d(sK + j * kChunkBytes, &tmK, &k_full[j], 0, j * BKEY, h, b);
      }
      for (int j = 0; j < NK; ++j) {
        mbar_arrive_expect_tx(&v_full[j], kChunkBytes);
        tma_load_4d(sV + j * kChunkBytes, &tmV, &v_full[j], 0, j * BKEY, h, b);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc_s = make_idesc_bf16(BQ, BKEY, 0, 0);      // Q, K both K-major
      const uint32_t idesc_o = make_idesc_bf16(BQ, D, 0, 1);         // P K-major, V MN-major
      const uint32_t aQ = smem_u32(sQ), aP = smem_u32(sP);
      auto issue_pv = [&](int j) {
        mbar_wait(p_full, j & 1);
        mbar_wait(&v_full[j], 0);
        tc_fence_after();
        const uint32_t aV = smem_u32(sV + j * kChunkBytes);
#pragma unroll
        for (int kk = 0; kk < BKEY / 16; ++kk) {
          const uint64_t ad = make_smem_desc_sw128(aP + (kk >> 2) * (BQ * 128) + (kk & 3) * 32, 16, 1024);
          const uint64_t bd = make_smem_desc_sw128(aV + kk * 2048, 8192, 1024);
          umma_f16(tmem + kColO, ad, bd, idesc_o, (j > 0) || (kk > 0));
        }
        umma_commit(p_free);                      // P smem reusable, O holds chunk j
      };
      mbar_wait(q_full, 0);
      for (int j = 0; j < NK; ++j) {
        mbar_wait(&k_full[j], 0);
        mbar_wait(&s_free[j & 1], ((j >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t aK = smem_u32(sK + j * kChunkBytes);
#pragma unroll
        for (int k = 0; k < D / 16; ++k) {
          const uint64_t ad = make_smem_desc_sw128(aQ + k * 32, 16, 1024);
          const uint64_t bd = make_smem_desc_sw128(aK + k * 32, 16, 1024);
          umma_f16(tmem + kColS + (j & 1) * BKEY, ad, bd, idesc_s, k != 0);
        }
        umma_commit(&s_full[j & 1]);
        if (j > 0) issue_pv(j - 1);
      }
      issue_pv(NK - 1);
      umma_commit(o_done);
    }
  } else {
    const int q = warp & 3;
    const int row = q * 32 + lane;                      // query row within the tile
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    float m = -INFINITY, l = 0.f;
    for (int j = 0; j < NK; ++j) {
      mbar_wait(&s_full[j & 1], (j >> 1) & 1);
      tc_fence_after();
      float s[BKEY];
#pragma unroll
      for (int c = 0; c < BKEY / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32b_x32(tmem + lane_base + kColS + (j & 1) * BKEY + c * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) s[c * 32 + i] = __uint_as_float(r[i]);
      }
      tc_fence_before();
      mbar_arrive(&s_free[j & 1]);                      // score buffer may be overwritten by S_{j+2}
      float mx = m;
#pragma unroll
      for (int i = 0; i < BKEY; ++i) mx = fmaxf(mx, s[i]);
      const float alpha = exp2f((m - mx) * p.scale_log2);       // m = -inf on the first chunk -> alpha = 0
      float sum = 0.f;
#pragma unroll
      for (int i = 0; i < BKEY; ++i) { s[i] = exp2f((s[i] - mx) * p.scale_log2); sum += s[i]; }
      l = l * alpha + sum;
      m = mx;
      // the PV MMA of the previous chunk must have finished: it reads P from smem and accumulates into O
      mbar_wait(p_free, (j & 1) ^ 1);
      tc_fence_after();
      if (j > 0) {                                      // rescale the running output
#pragma unroll
        for (int c = 0; c < D / 32; ++c) {
          uint32_t r[32];
          tmem_ld_32x32b_x32(tmem + lane_base + kColO + c * 32, r);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
          tmem_st_32x32b_x32(tmem + lane_base + kColO + c * 32, r);
        }
        tmem_st_wait();
      }
      // P_j -> smem, two K-major sub-tiles of 64 keys; 16-byte chunk index XOR (row & 7) = the 128-byte swizzle
#pragma unroll
      for (int t = 0; t < 2; ++t) {
        uint8_t* prow = sP + t * (BQ * 128) + row * 128;
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const float* v = s + t * 64 + c * 8;
          uint4 pk;
          pk.x = pack_bf16(v[0], v[1]); pk.y = pack_bf16(v[2], v[3]); pk.z = pack_bf16(v[4], v[5]); pk.w = pack_bf16(v[6], v[7]);
          *reinterpret_cast<uint4*>(prow + ((c ^ (row & 7)) << 4)) = pk;
        }
      }
      fence_proxy_async();
      tc_fence_before();
      mbar_arrive(p_full);
    }
    // epilogue: O / l -> bf16, log-sum-exp
    mbar_wait(o_done, 0);
    tc_fence_after();
    const float inv_l = 1.0f / l;
    const int srow = qt * BQ + row;
    __nv_bfloat16* orow = p.out + ((size_t)b * p.S + srow) * p.row_stride + (size_t)h * D;
#pragma unroll
    for (int c = 0; c < D / 32; ++c) {
      uint32_t r[32];
      tmem_ld_32x32b_x32(tmem + lane_base + kColO + c * 32, r);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; i += 8) {
        uint4 pk;
        pk.x = pack_bf16(__uint_as_float(r[i]) * inv_l, __uint_as_float(r[i + 1]) * inv_l);
        pk.y = pack_bf16(__uint_as_float(r[i + 2]) * inv_l, __uint_as_float(r[i + 3]) * inv_l);
        pk.z = pack_bf16(__uint_as_float(r[i + 4]) * inv_l, __uint_as_float(r[i + 5]) * inv_l);
        pk.w = pack_bf16(__uint_as_float(r[i + 6]) * inv_l, __uint_as_float(r[i + 7]) * inv_l);
        *reinterpret_cast<uint4*>(orow + c * 32 + i) = pk;
      }
    }
    if (p.lse != nullptr) p.lse[((size_t)b * p.H + h) * p.S + srow] = m * p.scale + logf(l);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, kTmemCols);
}

// ---- host ---------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn2)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static int fa_make_map(CUtensorMap* m, const void* base, int S, int H, int B, long long row_stride) {
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
  cuuint32_t box[4] = {64, 128, 1, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  MLT_REQUIRE(r == CUDA_SUCCESS, "attention: cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return 0;
}

}  // namespace mlt

using namespace mlt;

// q, k, v, out: [B, S, H*64] bf16 (row stride H*64);  lse: [B, H, S] fp32 or NULL.
extern "C" int mlt_attention_fwd(const void* q, const void* k, const void* v, void* out, void* lse, int B, int S, int H, int D,
                                 void* stream) {
  MLT_REQUIRE(D == 64, "mlt_attention_fwd: head dim %d unsupported (64)", D);
  MLT_REQUIRE(S % 128 == 0 && S >= 128 && S <= 128 * fa::kMaxChunks, "mlt_attention_fwd: sequence length %d unsupported (128..512, multiple of 128)", S);
  MLT_REQUIRE(((uintptr_t)q % 16) == 0 && ((uintptr_t)k % 16) == 0 && ((uintptr_t)v % 16) == 0 && ((uintptr_t)out % 16) == 0,
              "mlt_attention_fwd: unaligned tensors");
  const long long rs = (long long)H * D;
  CUtensorMap tq, tk, tv;
  if (fa_make_map(&tq, q, S, H, B, rs) || fa_make_map(&tk, k, S, H, B, rs) || fa_make_map(&tv, v, S, H, B, rs)) return 1;
  FaParams p;
  p.out = (__nv_bfloat16*)out; p.lse = (float*)lse; p.S = S; p.H = H; p.row_stride = rs;
  p.scale = 1.0f / sqrtf((float)D);
  p.scale_log2 = p.scale * 1.4426950408889634f;
  static bool attr_set[64] = {};
  int dev = 0;
  MLT_CHECK_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63]) {
    MLT_CHECK_CUDA(cudaFuncSetAttribute(attention_fwd_tcgen05_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, fa::kSmemBytes));
    attr_set[dev & 63] = true;
  }
  dim3 grid(S / fa::BQ, H, B);
  attention_fwd_tcgen05_kernel<<<grid, fa::kThreads, fa::kSmemBytes, (cudaStream_t)stream>>>(tq, tk, tv, p);
  MLT_LAUNCH_CHECK();
  return 0;
}

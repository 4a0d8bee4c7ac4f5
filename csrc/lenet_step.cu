// Whole training step of the reference model in ONE kernel (reference src/model.py:7-24 + src/trainer.py:181-188).
//
// The reference's MLModel step is ~0.13 GFLOP spread over ~60 library launches plus host syncs: pure launch latency.
// Here forward, loss, accuracy, backward, the cross-GPU gradient all-reduce and the optimizer update are one
// cooperative kernel with a single grid barrier:
//
//   phase A  one CTA per sample, everything in shared memory: conv1+ReLU+pool -> conv2+ReLU+pool -> fc1 -> fc2 -> fc3
//            -> softmax cross-entropy -> dlogits -> fc backward (data) -> conv2 dgrad; the per-sample convolution
//            weight-gradient partials and the small per-sample vectors the fc weight gradients need are written to a
//            global scratch slab (3696 floats per sample, L2 resident).
//   barrier  sense-reversing grid barrier (co-residency guaranteed by the cooperative launch).
//   phase B  every thread owns fixed elements of the flat parameter arena: it reduces that element's gradient over
//            the batch, (multi-GPU) pushes it into every peer's receive slot over NVLink, exchanges per-CTA flags,
//            sums the ranks in a fixed order, and applies the optimizer rule in registers.  CTA 0 also folds the
//            loss / accuracy of the batch into the trainer's device accumulators.
//
// Numerics are fp32 throughout, same formulas as the layer kernels in small_nn.cu / loss.cu (the tested oracle).
#include "p2p_device.cuh"
#include "../include/mlt_b200.h"
#include <cooperative_groups.h>
#include <cstdlib>

namespace mlt {

namespace lenet {
constexpr int CIN = 3, H = 32, C1 = 6, K = 5, O1 = 28, P1 = 14, C2 = 16, O2 = 10, P2 = 5;
constexpr int FLAT = C2 * P2 * P2;          // 400
constexpr int F1 = 120, F2 = 84, NC = 10;
constexpr int XN = CIN * H * H;             // 3072
constexpr int A1N = C1 * P1 * P1;           // 1176
constexpr int W1N = C1 * CIN * K * K;       // 450
constexpr int W2N = C2 * C1 * K * K;        // 2400
constexpr int ZH = O2 + 2 * (K - 1);        // 18: haloed conv2-output gradient plane
// per-sample scratch slab (floats)
constexpr int S_A2 = 0, S_Z1 = 400, S_Z2 = 520, S_DL = 604 /*10 dlogits, loss, correct*/, S_D2 = 616, S_D1 = 700;
constexpr int S_PW1 = 824 /*450 w + 6 b*/, S_PW2 = 1280 /*2400 w + 16 b*/, S_STRIDE = 3696;
// parameter ids in module registration order
enum { P_C1W, P_C1B, P_C2W, P_C2B, P_F1W, P_F1B, P_F2W, P_F2B, P_F3W, P_F3B, NPARAM };
__device__ __constant__ int kParamSize[NPARAM] = {W1N, C1, W2N, C2, F1 * FLAT, F1, F2 * F1, F2, NC * F2, NC};
constexpr int kThreads = 1024;             // 32 warps: the per-sample CTA is latency bound, it needs the warps
constexpr int NW = kThreads / 32;
constexpr int kMaxPerThread = 8;            // arena elements owned per thread in phase B
constexpr int kMaxSlots = 8;                // input ring slots a multi-step launch can walk
}  // namespace lenet

struct LeNetArgs {
  float* master;                 // flat fp32 parameter arena (model == master for fp32 compute)
  float* s1;
  float* s2;
  const Hyper* hyper;
  long long numel;               // padded arena length
  long long off[lenet::NPARAM];  // arena element offset of each parameter
  const float* x;                // [B,3,32,32]
  const long long* y;            // [B]
  int B;
  int train;                     // 0: forward + loss only
  int want_acc;
  float* scratch;                // B * S_STRIDE floats
  unsigned int* sync;            // {count, sense}, zero-initialised once
  float* acc3;                   // {loss_sum, metric_sum, steps} or nullptr
  float* loss_out;               // scalar or nullptr
  P2PArgs comm;                  // world == 1: no communication
  // ---- several steps in ONE launch (cluster kernel only): removes the ~10 us launch / drain gap between steps --------
  int nsteps;                    // 0: a single step with the fields above;  >= 1: the multi-step protocol below
  int nslots;                    // ring mode: step i reads xs[i % nslots] / ys[i % nslots]
  const float* xs[lenet::kMaxSlots];
  const long long* ys[lenet::kMaxSlots];
  long long pool_n;              // pool mode (> 0): step i reads xs[0] + (i % pool_n) * B * XN (and ys[0] likewise)
  const int* batch_sizes;        // per-step batch size (device), or nullptr: B
  const Hyper* hyper_table;      // per-step hyper rows (device), or nullptr: hyper
  float* loss_base;              // loss of step i -> loss_base[i] (device or pinned host), or nullptr
  const unsigned int* ready;     // step i may start once *ready >= i + 1 (written by the copy stream), or nullptr
  unsigned int* progress;        // *progress = i + 1 after step i (pinned host), or nullptr
};

__device__ __forceinline__ unsigned int ld_acquire_gpu(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Bulk asynchronous copy global -> shared (the TMA engine, no tensor map): one instruction per contiguous block.
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned int bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"((unsigned int)__cvta_generic_to_shared(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gmem_src, unsigned int bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   (unsigned int)__cvta_generic_to_shared(smem_dst)),
               "l"(gmem_src), "r"(bytes), "r"((unsigned int)__cvta_generic_to_shared(bar))
               : "memory");
}
__device__ __forceinline__ void mbar_wait_parity(unsigned long long* bar, unsigned int parity) {
  unsigned int ok = 0;
  while (!ok) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.b32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"((unsigned int)__cvta_generic_to_shared(bar)), "r"(parity)
        : "memory");
  }
}

__device__ __forceinline__ unsigned int ld_acquire_sys_u32(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ void grid_barrier(unsigned int* sync, unsigned int nblocks, unsigned int sense0) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned int old = atomicAdd(&sync[0], 1u);
    if (old == nblocks - 1) {
      sync[0] = 0u;
      __threadfence();
      atomicExch(&sync[1], sense0 + 1u);
    } else {
      while (ld_acquire_gpu(&sync[1]) == sense0) {
      }
    }
    __threadfence();
  }
  __syncthreads();
}

__device__ __forceinline__ float dot4(const float4& a, const float4& b) { return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w; }

// Phase B (after the grid barrier): batch loss / accuracy into the device accumulators, then every thread reduces the
// gradient of the arena elements it owns over the batch, exchanges it with the peers (multi-GPU) and applies the rule.
template <typename R>
__device__ __forceinline__ void lenet_phase_b(const LeNetArgs& a, float inv_B, int cur_B, const Hyper* cur_hyper, float* cur_loss) {
  using namespace lenet;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (blockIdx.x == 0 && warp == 0) {            // batch loss / accuracy -> device accumulators
    float l = 0.f, c = 0.f;
    for (int b = lane; b < cur_B; b += 32) {
      l += __ldcg(a.scratch + (size_t)b * S_STRIDE + S_DL + 10);
      c += __ldcg(a.scratch + (size_t)b * S_STRIDE + S_DL + 11);
    }
    l = warp_sum(l);
    c = warp_sum(c);
    if (lane == 0) {
      const float loss = l * inv_B;
      if (cur_loss) *cur_loss = loss;
      if (a.acc3) {
        a.acc3[0] += loss;
        if (a.want_acc) a.acc3[1] += c * inv_B;
        a.acc3[2] += 1.0f;
      }
    }
  }
  if (!a.train) return;

  const Hyper h = *cur_hyper;
  const int per_cta = (int)ceil_div64(a.numel, (long long)gridDim.x * kThreads) * kThreads;   // elements owned by a CTA
  const long long e0 = (long long)blockIdx.x * per_cta;
  const int per_thread = per_cta / kThreads;
  float g[kMaxPerThread];
  const int B = cur_B;
#pragma unroll
  for (int r = 0; r < kMaxPerThread; ++r) {
    g[r] = 0.f;
    if (r >= per_thread) continue;
    const long long e = e0 + (long long)r * kThreads + tid;
    if (e >= a.numel) continue;
    // which parameter does arena element e belong to?
    int pid = -1; int i = 0;
#pragma unroll
    for (int p = 0; p < NPARAM; ++p) {
      const long long d = e - a.off[p];
      if (d >= 0 && d < kParamSize[p]) { pid = p; i = (int)d; }
    }
    if (pid < 0) continue;                        // alignment padding between parameters
    const float* sc = a.scratch;
    float s = 0.f;
    switch (pid) {
      case P_F1W: { const int n = i / FLAT, k = i - n * FLAT;
        for (int b = 0; b < B; ++b) s = fmaf(__ldcg(sc + (size_t)b * S_STRIDE + S_D1 + n), __ldcg(sc + (size_t)b * S_STRIDE + S_A2 + k), s); break; }
      case P_F2W: { const int n = i / F1, k = i - n * F1;
        for (int b = 0; b < B; ++b) s = fmaf(__ldcg(sc + (size_t)b * S_STRIDE + S_D2 + n), __ldcg(sc + (size_t)b * S_STRIDE + S_Z1 + k), s); break; }
      case P_F3W: { const int n = i / F2, k = i - n * F2;
        for (int b = 0; b < B; ++b) s = fmaf(__ldcg(sc + (size_t)b * S_STRIDE + S_DL + n), __ldcg(sc + (size_t)b * S_STRIDE + S_Z2 + k), s); break; }
      case P_F1B: for (int b = 0; b < B; ++b) s += __ldcg(sc + (size_t)b * S_STRIDE + S_D1 + i); break;
      case P_F2B: for (int b = 0; b < B; ++b) s += __ldcg(sc + (size_t)b * S_STRIDE + S_D2 + i); break;
      case P_F3B: for (int b = 0; b < B; ++b) s += __ldcg(sc + (size_t)b * S_STRIDE + S_DL + i); break;
      case P_C2W: for (int b = 0; b < B; ++b) s += __ldcg(sc + (size_t)b * S_STRIDE + S_PW2 + i); break;
      case P_C2B: for (int b = 0; b < B; ++b) s += __ldcg(sc + (size_t)b * S_STRIDE + S_PW2 + W2N + i); break;
      case P_C1W: for (int b = 0; b < B; ++b) s += __ldcg(sc + (size_t)b * S_STRIDE + S_PW1 + i); break;
      case P_C1B: for (int b = 0; b < B; ++b) s += __ldcg(sc + (size_t)b * S_STRIDE + S_PW1 + W1N + i); break;
    }
    g[r] = s;
  }

  const P2PArgs& cm = a.comm;
  unsigned int parity = 0;
  if (cm.world > 1) {
    // one-shot push of this CTA's gradient elements into every peer's receive slot, then a per-CTA flag exchange
    const int me = cm.rank;
    parity = (flags_of(cm, me)->tick[0][blockIdx.x] + 1) & 1u;
    const long long slot_off = cm.recv_off + ((long long)parity * cm.world + me) * cm.recv_slot_bytes;
    if (cm.mc_flags) {
      // ONE store per element through the multicast mapping: the switch writes it into slot [parity][me] of every rank
      // (mine too), instead of world-1 unicast stores per element
      float* dst = reinterpret_cast<float*>(cm.mc_base + slot_off);
#pragma unroll
      for (int r = 0; r < kMaxPerThread; ++r) {
        const long long e = e0 + (long long)r * kThreads + tid;
        if (r < per_thread && e < a.numel) mc_st_f32(dst + e, g[r]);
      }
    } else {
      for (int p = 0; p < cm.world; ++p) {
        if (p == me) continue;
        float* dst = reinterpret_cast<float*>(cm.base[p] + slot_off);
#pragma unroll
        for (int r = 0; r < kMaxPerThread; ++r) {
          const long long e = e0 + (long long)r * kThreads + tid;
          if (r < per_thread && e < a.numel) dst[e] = g[r];
        }
      }
    }
    if (!cta_exchange(cm, me, 0, blockIdx.x)) return;        // a lost peer: weights and optimizer state stay untouched
  }
#pragma unroll
  for (int r = 0; r < kMaxPerThread; ++r) {
    if (r >= per_thread) continue;
    const long long e = e0 + (long long)r * kThreads + tid;
    if (e >= a.numel) continue;
    float gs = g[r];
    if (cm.world > 1) {
      gs = 0.f;
      const char* recv = cm.base[cm.rank] + cm.recv_off + (long long)parity * cm.world * cm.recv_slot_bytes;
      for (int p = 0; p < cm.world; ++p)        // fixed rank order: every replica computes identical bits
        gs += (p == cm.rank) ? g[r] : __ldcg(reinterpret_cast<const float*>(recv + p * cm.recv_slot_bytes) + e);
    }
    float pv = a.master[e], s1 = a.s1[e], s2 = 0.f;
    if (R::kStates > 1) s2 = a.s2[e];
    R::apply(pv, gs * h.grad_scale, s1, s2, h);
    a.master[e] = pv;
    a.s1[e] = s1;
    if (R::kStates > 1) a.s2[e] = s2;
  }
}

template <typename R>
__global__ void __launch_bounds__(lenet::kThreads, 1)
lenet_step_kernel(const LeNetArgs a) {
  using namespace lenet;
  extern __shared__ __align__(16) float sm[];
  float* xs = sm;                       // 3072
  float* w1s = xs + XN;                 // 456
  float* a1s = w1s + 456;               // 1176
  float* w2s = a1s + A1N;               // 2416
  float* a2s = w2s + 2416;              // 400
  float* z1s = a2s + FLAT;              // 120
  float* z2s = z1s + F1;                // 84 (+4 pad)
  float* lgs = z2s + 88;                // 16
  float* dls = lgs + 16;                // 16
  float* d2s = dls + 16;                // 84 (+4)
  float* d1s = d2s + 88;                // 120
  float* g2p = d1s + F1;                // 400
  float* g1p = g2p + FLAT;              // 1176
  float* dz2 = g1p + A1N;               // 16*18*18 = 5184 used by dgrad; 6400 reserved (reduction scratch)
  float* dgp = dz2 + 6400;              // 4 x 1176 partial sums of the conv2 dgrad
  int* offs1 = reinterpret_cast<int*>(dgp + 4 * A1N);        // 1176
  int* offs2 = offs1 + A1N;                                  // 400
  unsigned char* idx1 = reinterpret_cast<unsigned char*>(offs2 + FLAT);   // 1176
  unsigned char* idx2 = idx1 + A1N;                                       // 400

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  __shared__ unsigned int s_sense;
  if (tid == 0) s_sense = ld_acquire_gpu(&a.sync[1]);

  const float* Wc1 = a.master + a.off[P_C1W];
  const float* Bc1 = a.master + a.off[P_C1B];
  const float* Wc2 = a.master + a.off[P_C2W];
  const float* Bc2 = a.master + a.off[P_C2B];
  const float* Wf1 = a.master + a.off[P_F1W];
  const float* Bf1 = a.master + a.off[P_F1B];
  const float* Wf2 = a.master + a.off[P_F2W];
  const float* Bf2 = a.master + a.off[P_F2B];
  const float* Wf3 = a.master + a.off[P_F3W];
  const float* Bf3 = a.master + a.off[P_F3B];

  if ((int)blockIdx.x < a.B) {        // this CTA has at least one sample: stage the conv filters once
    for (int i = tid; i < W1N; i += kThreads) w1s[i] = Wc1[i];
    if (tid < C1) w1s[W1N + tid] = Bc1[tid];
    for (int i = tid; i < W2N; i += kThreads) w2s[i] = Wc2[i];
    if (tid < C2) w2s[W2N + tid] = Bc2[tid];
  }
  const float inv_B = 1.0f / (float)a.B;

  // =========================================== phase A ===========================================
  for (int b = blockIdx.x; b < a.B; b += gridDim.x) {
    float* slab = a.scratch + (size_t)b * S_STRIDE;
    __syncthreads();
    {   // input sample -> smem
      const float4* src = reinterpret_cast<const float4*>(a.x + (size_t)b * XN);
      float4* dst = reinterpret_cast<float4*>(xs);
      for (int i = tid; i < XN / 4; i += kThreads) dst[i] = src[i];
    }
    __syncthreads();
    // ---- conv1 + bias + ReLU + 2x2 max-pool: [3,32,32] -> [6,14,14]
    for (int o = tid; o < A1N; o += kThreads) {
      const int co = o / (P1 * P1), r = o - co * (P1 * P1), ph = r / P1, pw = r - ph * P1;
      const float bv = w1s[W1N + co];
      float a00 = bv, a01 = bv, a10 = bv, a11 = bv;
#pragma unroll
      for (int ci = 0; ci < CIN; ++ci) {
        const float* xc = xs + ci * H * H + (2 * ph) * H + 2 * pw;
        const float* wk = w1s + (co * CIN + ci) * K * K;
#pragma unroll
        for (int kh = 0; kh < K; ++kh) {
          const float* r0 = xc + kh * H;
          const float* r1 = r0 + H;
          float p0 = r0[0], p1 = r1[0];
#pragma unroll
          for (int kw = 0; kw < K; ++kw) {
            const float wv = wk[kh * K + kw];
            const float q0 = r0[kw + 1], q1 = r1[kw + 1];
            a00 = fmaf(wv, p0, a00); a01 = fmaf(wv, q0, a01);
            a10 = fmaf(wv, p1, a10); a11 = fmaf(wv, q1, a11);
            p0 = q0; p1 = q1;
          }
        }
      }
      float m = a00; int am = 0;
      if (a01 > m) { m = a01; am = 1; }
      if (a10 > m) { m = a10; am = 2; }
      if (a11 > m) { m = a11; am = 3; }
      a1s[o] = fmaxf(m, 0.f);
      idx1[o] = (unsigned char)am;
      offs1[o] = (2 * ph + (am >> 1)) * H + 2 * pw + (am & 1);     // live pre-pool position in the 32x32 input plane
    }
    __syncthreads();
    // ---- conv2 + bias + ReLU + pool: [6,14,14] -> [16,5,5]
    for (int o = tid; o < FLAT; o += kThreads) {
      const int co = o / (P2 * P2), r = o - co * (P2 * P2), ph = r / P2, pw = r - ph * P2;
      const float bv = w2s[W2N + co];
      float a00 = bv, a01 = bv, a10 = bv, a11 = bv;
      for (int ci = 0; ci < C1; ++ci) {
        const float* xc = a1s + ci * P1 * P1 + (2 * ph) * P1 + 2 * pw;
        const float* wk = w2s + (co * C1 + ci) * K * K;
#pragma unroll
        for (int kh = 0; kh < K; ++kh) {
          const float* r0 = xc + kh * P1;
          const float* r1 = r0 + P1;
          float p0 = r0[0], p1 = r1[0];
#pragma unroll
          for (int kw = 0; kw < K; ++kw) {
            const float wv = wk[kh * K + kw];
            const float q0 = r0[kw + 1], q1 = r1[kw + 1];
            a00 = fmaf(wv, p0, a00); a01 = fmaf(wv, q0, a01);
            a10 = fmaf(wv, p1, a10); a11 = fmaf(wv, q1, a11);
            p0 = q0; p1 = q1;
          }
        }
      }
      float m = a00; int am = 0;
      if (a01 > m) { m = a01; am = 1; }
      if (a10 > m) { m = a10; am = 2; }
      if (a11 > m) { m = a11; am = 3; }
      m = fmaxf(m, 0.f);
      a2s[o] = m;
      slab[S_A2 + o] = m;
      idx2[o] = (unsigned char)am;
      offs2[o] = (2 * ph + (am >> 1)) * P1 + 2 * pw + (am & 1);    // position in the 14x14 conv2 input plane
    }
    __syncthreads();
    // ---- fc1 + ReLU: one warp per neuron, weight rows streamed as float4 from L2
    {
      const float4* xv = reinterpret_cast<const float4*>(a2s);
      for (int n0 = warp; n0 < F1; n0 += NW * 3) {
        float acc[3] = {0.f, 0.f, 0.f};
#pragma unroll
        for (int u = 0; u < 3; ++u) {
          const int n = n0 + NW * u;
          if (n < F1) {
            const float4* wr = reinterpret_cast<const float4*>(Wf1 + (size_t)n * FLAT);
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              const int c4 = lane + 32 * c;
              if (c4 < FLAT / 4) acc[u] += dot4(wr[c4], xv[c4]);
            }
          }
        }
#pragma unroll
        for (int u = 0; u < 3; ++u) {
          const float s = warp_sum(acc[u]);
          const int n = n0 + NW * u;
          if (lane == 0 && n < F1) {
            const float v = fmaxf(s + Bf1[n], 0.f);
            z1s[n] = v;
            slab[S_Z1 + n] = v;
          }
        }
      }
    }
    __syncthreads();
    // ---- fc2 + ReLU
    {
      const float4* xv = reinterpret_cast<const float4*>(z1s);
      for (int n0 = warp; n0 < F2; n0 += NW * 4) {
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int n = n0 + NW * u;
          if (n < F2 && lane < F1 / 4) acc[u] = dot4(reinterpret_cast<const float4*>(Wf2 + (size_t)n * F1)[lane], xv[lane]);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const float s = warp_sum(acc[u]);
          const int n = n0 + NW * u;
          if (lane == 0 && n < F2) {
            const float v = fmaxf(s + Bf2[n], 0.f);
            z2s[n] = v;
            slab[S_Z2 + n] = v;
          }
        }
      }
    }
    __syncthreads();
    // ---- fc3 (logits)
    {
      const float4* xv = reinterpret_cast<const float4*>(z2s);
      for (int n = warp; n < NC; n += NW) {
        float acc = 0.f;
        if (lane < F2 / 4) acc = dot4(reinterpret_cast<const float4*>(Wf3 + (size_t)n * F2)[lane], xv[lane]);
        acc = warp_sum(acc);
        if (lane == 0) lgs[n] = acc + Bf3[n];
      }
    }
    __syncthreads();
    // ---- softmax cross-entropy, accuracy, dlogits
    if (warp == 0) {
      const int tgt = (int)a.y[b];
      const float v = lane < NC ? lgs[lane] : -INFINITY;
      float m = v; int am = lane < NC ? lane : 0x7fffffff;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float om = __shfl_xor_sync(0xffffffffu, m, o);
        const int oa = __shfl_xor_sync(0xffffffffu, am, o);
        if (om > m || (om == m && oa < am)) { m = om; am = oa; }
      }
      const float e = lane < NC ? expf(v - m) : 0.f;
      const float s = warp_sum(e);
      if (lane < NC) {
        float p = e / s;
        if (lane == tgt) p -= 1.0f;
        p *= inv_B;
        dls[lane] = p;
        slab[S_DL + lane] = p;
      }
      if (lane == 0) {
        const float xt = (tgt >= 0 && tgt < NC) ? lgs[tgt] : (logf(s) + m);
        slab[S_DL + 10] = logf(s) + m - xt;                 // per-sample loss
        slab[S_DL + 11] = (am == tgt) ? 1.f : 0.f;
      }
    }
    __syncthreads();
    if (!a.train) continue;
    // ---- backward through fc3 / fc2 (data gradients)
    if (tid < F2) {
      float s = 0.f;
#pragma unroll
      for (int n = 0; n < NC; ++n) s = fmaf(dls[n], Wf3[n * F2 + tid], s);
      s = z2s[tid] > 0.f ? s : 0.f;
      d2s[tid] = s;
      slab[S_D2 + tid] = s;
    }
    __syncthreads();
    if (tid < 8 * F1) {                     // eight slices of the 84-long reduction
      const int k = tid % F1, part = tid / F1;
      float s = 0.f;
      const int n_hi = min(F2, part * 11 + 11);
#pragma unroll 11
      for (int n = part * 11; n < n_hi; ++n) s = fmaf(d2s[n], Wf2[n * F1 + k], s);
      dz2[tid] = s;                         // dz2 doubles as reduction scratch here
    }
    __syncthreads();
    if (tid < F1) {
      float s = 0.f;
#pragma unroll
      for (int part = 0; part < 8; ++part) s += dz2[part * F1 + tid];
      s = z1s[tid] > 0.f ? s : 0.f;
      d1s[tid] = s;
      slab[S_D1 + tid] = s;
    }
    __syncthreads();
    // ---- backward through fc1: g2p[k] = sum_n d1[n] W1[n][k]; warps = 16 row groups x 2 column halves, each
    //      scales and accumulates its rows' float4s; partials are combined through shared memory
    {
      const int rg = warp & 15, half = warp >> 4;
      float4 acc[2];
      acc[0] = acc[1] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
      for (int n = rg; n < F1; n += 16) {
        const float dv = d1s[n];
        const float4* wr = reinterpret_cast<const float4*>(Wf1 + (size_t)n * FLAT);
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          const int c4 = lane + 32 * (2 * half + c);
          if (c4 < FLAT / 4) {
            const float4 w = wr[c4];
            acc[c].x = fmaf(dv, w.x, acc[c].x); acc[c].y = fmaf(dv, w.y, acc[c].y);
            acc[c].z = fmaf(dv, w.z, acc[c].z); acc[c].w = fmaf(dv, w.w, acc[c].w);
          }
        }
      }
      float4* part = reinterpret_cast<float4*>(dz2) + rg * (FLAT / 4);      // [16][100] float4
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        const int c4 = lane + 32 * (2 * half + c);
        if (c4 < FLAT / 4) part[c4] = acc[c];
      }
    }
    __syncthreads();
    for (int k = tid; k < FLAT; k += kThreads) {
      float s = 0.f;
#pragma unroll
      for (int w = 0; w < 16; ++w) s += dz2[w * FLAT + k];
      g2p[k] = a2s[k] > 0.f ? s : 0.f;      // through ReLU (pooled output > 0 <=> the live pre-pool value > 0)
    }
    __syncthreads();
    // ---- conv2: per-sample weight / bias gradient partials, and the haloed dense gradient plane for dgrad
    for (int i = tid; i < C2 * ZH * ZH; i += kThreads) dz2[i] = 0.f;
    __syncthreads();
    for (int t = tid; t < W2N + C2; t += kThreads) {
      float s = 0.f;
      if (t < W2N) {
        const int co = t / (C1 * K * K), r = t - co * (C1 * K * K), ci = r / (K * K), q = r - ci * (K * K);
        const float* xt = a1s + ci * P1 * P1 + (q / K) * P1 + (q % K);
        const float* g = g2p + co * (P2 * P2);
        const int* of = offs2 + co * (P2 * P2);
#pragma unroll
        for (int p = 0; p < P2 * P2; ++p) s = fmaf(g[p], xt[of[p]], s);
      } else {
        const float* g = g2p + (t - W2N) * (P2 * P2);
#pragma unroll
        for (int p = 0; p < P2 * P2; ++p) s += g[p];
      }
      slab[S_PW2 + t] = s;
    }
    for (int o = tid; o < FLAT; o += kThreads) {
      const int co = o / (P2 * P2);
      const int pos = offs2[o];                      // oh * 14 + ow  with oh, ow < 10
      const int oh = pos / P1, ow = pos - oh * P1;
      dz2[co * ZH * ZH + (oh + K - 1) * ZH + (ow + K - 1)] = g2p[o];
    }
    __syncthreads();
    // ---- conv2 dgrad -> gradient w.r.t. the pooled conv1 output, through its ReLU.
    //      Work item = (group of 4 output channels, ci, row h, strip of 7 columns): per (co, kh) it loads 11 gradient
    //      values + 5 taps for 35 FMAs (shared memory is the bound here); the 4 channel groups are summed afterwards.
    for (int it = tid; it < 4 * C1 * P1 * 2; it += kThreads) {
      const int cog = it / (C1 * P1 * 2), r = it - cog * (C1 * P1 * 2);
      const int ci = r / (P1 * 2), r2 = r - ci * (P1 * 2), h = r2 >> 1, w0 = (r2 & 1) * 7;
      float acc[7] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      for (int co = cog * 4; co < cog * 4 + 4; ++co) {
        const float* wk = w2s + (co * C1 + ci) * K * K;
#pragma unroll
        for (int kh = 0; kh < K; ++kh) {
          // output (h, w) needs dz[h - kh + 4][w - kw + 4]; for w = w0..w0+6, kw = 0..4 that is columns w0 .. w0 + 10
          const float* row = dz2 + co * ZH * ZH + (h - kh + K - 1) * ZH + w0;
          float d[11];
#pragma unroll
          for (int j = 0; j < 11; ++j) d[j] = row[j];
#pragma unroll
          for (int kw = 0; kw < K; ++kw) {
            const float wv = wk[kh * K + kw];
#pragma unroll
            for (int j = 0; j < 7; ++j) acc[j] = fmaf(d[j + (K - 1) - kw], wv, acc[j]);
          }
        }
      }
      float* dst = dgp + cog * A1N + ci * P1 * P1 + h * P1 + w0;
#pragma unroll
      for (int j = 0; j < 7; ++j) dst[j] = acc[j];
    }
    __syncthreads();
    for (int o = tid; o < A1N; o += kThreads) {
      const float s = dgp[o] + dgp[A1N + o] + dgp[2 * A1N + o] + dgp[3 * A1N + o];
      g1p[o] = a1s[o] > 0.f ? s : 0.f;
    }
    __syncthreads();
    // ---- conv1: per-sample weight / bias gradient partials
    for (int t = tid; t < W1N + C1; t += kThreads) {
      float s = 0.f;
      if (t < W1N) {
        const int co = t / (CIN * K * K), r = t - co * (CIN * K * K), ci = r / (K * K), q = r - ci * (K * K);
        const float* xt = xs + ci * H * H + (q / K) * H + (q % K);
        const float* g = g1p + co * (P1 * P1);
        const int* of = offs1 + co * (P1 * P1);
#pragma unroll 4
        for (int p = 0; p < P1 * P1; ++p) s = fmaf(g[p], xt[of[p]], s);
      } else {
        const float* g = g1p + (t - W1N) * (P1 * P1);
        for (int p = 0; p < P1 * P1; ++p) s += g[p];
      }
      slab[S_PW1 + t] = s;
    }
  }

  // =========================================== barrier ===========================================
  grid_barrier(a.sync, gridDim.x, s_sense);

  lenet_phase_b<R>(a, inv_B, a.B, a.hyper, a.loss_out);
}

// =====================================================================================================
// Cluster variant of phase A: a sample is carried by a cluster of 4 CTAs (4 SMs) instead of one.  Every stage's
// outputs are split four ways; each CTA writes its share into the shared memory of all four (distributed shared
// memory), then the cluster synchronises.  Stages whose work is tiny (fc3, loss, d2, d1) run redundantly in every CTA.
// With 32 samples and 148 SMs this puts 128 SMs to work in phase A instead of 32.
// =====================================================================================================
namespace lenet {
constexpr int CS = 4;                       // CTAs per sample
}

// Optional per-CTA timeline (analysis only): 16 x int64 per CTA of %globaltimer stamps when a buffer has been installed.
__device__ long long* g_lenet_dbg = nullptr;
// Optional per-STEP stamps (analysis only, bench/epoch_profile.py): CTA 0 appends {step start, step end} in %globaltimer
// nanoseconds to g_lenet_stamps (capacity g_lenet_stamp_cap pairs); [0] of the counter is the number of steps stamped so far.
__device__ long long* g_lenet_stamps = nullptr;
__device__ unsigned int g_lenet_stamp_cap = 0;
__device__ unsigned int g_lenet_stamp_n = 0;
__device__ __forceinline__ long long ln_gtimer() {
  long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#ifndef MLT_LENET_TIMELINE
#define LN_DBG(slot) do { (void)ln_dbg; } while (0)
#else
#define LN_DBG(slot)                                                                           \
  do {                                                                                         \
    if (ln_dbg != nullptr && threadIdx.x == 0) {                                               \
      long long t_;                                                                            \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                                   \
      ln_dbg[blockIdx.x * 16 + (slot)] = t_;                                                   \
    }                                                                                          \
  } while (0)
#endif

// MULTI = false: one step per launch, weights read where they are used (the cold-L2 friendly form).
// MULTI = true : a.nsteps steps per launch (ring / pool inputs, in-kernel barriers between steps) with the fully-connected
//               weights staged in shared memory by bulk copies while the convolutions run.
template <typename R, bool MULTI>
__global__ void __launch_bounds__(lenet::kThreads, 1)
lenet_step_cluster_kernel(const LeNetArgs a) {
  using namespace lenet;
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  const int cr = (int)cluster.block_rank();
  extern __shared__ __align__(16) float sm[];
  float* xs = sm;                       // 3072
  float* w1s = xs + XN;                 // 456
  float* a1s = w1s + 456;               // 1176
  float* w2s = a1s + A1N;               // 2416
  float* a2s = w2s + 2416;              // 400
  float* z1s = a2s + FLAT;              // 120
  float* z2s = z1s + F1;                // 88
  float* lgs = z2s + 88;                // 16
  float* dls = lgs + 16;                // 16
  float* d2s = dls + 16;                // 88
  float* d1s = d2s + 88;                // 120
  float* g2p = d1s + F1;                // 400
  float* g1p = g2p + FLAT;              // 1176
  float* dz2 = g1p + A1N;               // 6400: haloed plane (5184) / reduction scratch
  float* dgp = dz2 + 6400;              // 4 x 1176: this CTA's per-output-channel dgrad partials
  float* g2x = dgp + 4 * A1N;           // 4 x 400 : fc1-backward partials of the 4 CTAs
  float* g1x = g2x + CS * FLAT;         // 4 x 1176: conv2-dgrad partials of the 4 CTAs
  float* w1x = g1x + CS * A1N;          // 4 x 128 : conv1 wgrad position-split partials
  int* offs1 = reinterpret_cast<int*>(w1x + 4 * 128);        // 1176
  int* offs2 = offs1 + A1N;                                  // 400
  // fully-connected weights staged once per step with cp.async while the convolutions run: this CTA's 30 rows of fc1
  // (forward and backward use the same rows), all of fc2 / fc3 and the three bias vectors - the fc stages are chains of
  // tiny dependent products, and an L2 round trip per stage was most of their time
  float* wf1s = reinterpret_cast<float*>(offs2 + FLAT);      // 30 x 400
  float* wf2s = wf1s + 30 * FLAT;                            // 84 x 120
  float* wf3s = wf2s + F2 * F1;                              // 10 x 84
  float* bfs = wf3s + NC * F2;                               // 120 + 84 + 10 (+ pad)
  float* rsm[CS];
#pragma unroll
  for (int r = 0; r < CS; ++r) rsm[r] = cluster.map_shared_rank(sm, r);
  auto bcast_f = [&](float* local, float v) {
    const int off = (int)(local - sm);
#pragma unroll
    for (int r = 0; r < CS; ++r) rsm[r][off] = v;
  };
  auto bcast_i = [&](int* local, int v) {
    const int off = (int)(reinterpret_cast<float*>(local) - sm);
#pragma unroll
    for (int r = 0; r < CS; ++r) reinterpret_cast<int*>(rsm[r])[off] = v;
  };

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
#ifdef MLT_LENET_TIMELINE
  long long* ln_dbg = g_lenet_dbg;
#else
  long long* ln_dbg = nullptr;
#endif
  LN_DBG(0);
  __shared__ unsigned int s_sense;
  if (tid == 0) s_sense = ld_acquire_gpu(&a.sync[1]);

  const float* Wc1 = a.master + a.off[P_C1W];
  const float* Bc1 = a.master + a.off[P_C1B];
  const float* Wc2 = a.master + a.off[P_C2W];
  const float* Bc2 = a.master + a.off[P_C2B];
  const float* Wf1 = a.master + a.off[P_F1W];
  const float* Bf1 = a.master + a.off[P_F1B];
  const float* Wf2 = a.master + a.off[P_F2W];
  const float* Bf2 = a.master + a.off[P_F2B];
  const float* Wf3 = a.master + a.off[P_F3W];
  const float* Bf3 = a.master + a.off[P_F3B];

  const int cluster_id = blockIdx.x / CS, nclusters = gridDim.x / CS;
  auto share = [&](int n, int& lo, int& hi) { const int per = (n + CS - 1) / CS; lo = min(n, cr * per); hi = min(n, lo + per); };
  const int nsteps = MULTI ? a.nsteps : 1;
  constexpr bool multi = MULTI;
  __shared__ int s_abort;
  __shared__ __align__(8) unsigned long long wbar;
  if (MULTI) {
    if (tid == 0) {
      s_abort = 0;
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"((unsigned int)__cvta_generic_to_shared(&wbar)) : "memory");
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
  }
  int f1_lo, f1_hi;
  share(F1, f1_lo, f1_hi);
  // where the fc stages read their weights: staged copies (MULTI) or the arena itself
  const float* pW1 = MULTI ? wf1s : Wf1 + (size_t)f1_lo * FLAT;      // this CTA's rows of fc1
  const float* pW2 = MULTI ? wf2s : Wf2;
  const float* pW3 = MULTI ? wf3s : Wf3;
  const float* pB1 = MULTI ? bfs : Bf1;
  const float* pB2 = MULTI ? bfs + F1 : Bf2;
  const float* pB3 = MULTI ? bfs + F1 + F2 : Bf3;

  unsigned int wcount = 0;                // bulk weight copies this CTA has issued (steps with fewer samples than clusters skip)
  for (int step = 0; step < nsteps; ++step) {
  // ---- this step's inputs (multi-step launches walk a ring of slots or a device-resident pool) -------------------
  const float* cur_x = a.x;
  const long long* cur_y = a.y;
  int cur_B = a.B;
  const Hyper* cur_hyper = a.hyper;
  float* cur_loss = a.loss_out;
  if (multi) {
    if (a.ready != nullptr) {                         // wait until the copy stream has landed this step's batch
      if (tid == 0) {
        const long long t0 = clock64();
        while ((int)(ld_acquire_sys_u32(a.ready) - (unsigned int)(step + 1)) < 0) {
          if (clock64() - t0 > (1LL << 33)) { s_abort = 1; break; }      // ~4 s: the producer died
        }
      }
      __syncthreads();
      if (s_abort) {
        if (blockIdx.x == 0 && tid == 0 && a.progress) { *a.progress = 0xFFFFFFFFu; __threadfence_system(); }
        return;                                        // every CTA polls the same word: all give up together
      }
    }
    if (a.batch_sizes) cur_B = a.batch_sizes[step];
    if (a.pool_n > 0) {
      const long long slot = step % a.pool_n;
      cur_x = a.xs[0] + slot * (long long)a.B * XN;
      cur_y = a.ys[0] + slot * (long long)a.B;
    } else {
      cur_x = a.xs[step % a.nslots];
      cur_y = a.ys[step % a.nslots];
    }
    if (a.hyper_table) cur_hyper = a.hyper_table + step;
    if (a.loss_base) cur_loss = a.loss_base + step;
  }
  unsigned int stamp_at = 0xFFFFFFFFu;
  if (blockIdx.x == 0 && tid == 0 && g_lenet_stamps != nullptr) {
    stamp_at = atomicAdd(&g_lenet_stamp_n, 1u);
    if (stamp_at < g_lenet_stamp_cap) g_lenet_stamps[2 * stamp_at] = ln_gtimer(); else stamp_at = 0xFFFFFFFFu;
  }
  if (cluster_id < cur_B) {
    for (int i = tid; i < W1N; i += kThreads) w1s[i] = Wc1[i];
    if (tid < C1) w1s[W1N + tid] = Bc1[tid];
    for (int i = tid; i < W2N; i += kThreads) w2s[i] = Wc2[i];
    if (tid < C2) w2s[W2N + tid] = Bc2[tid];
    if (MULTI) {
      ++wcount;
      if (tid == 0) {                    // three bulk copies (TMA, no tensor map), completion counted on wbar
        const unsigned int b1 = (unsigned int)(f1_hi - f1_lo) * FLAT * 4, b2 = F2 * F1 * 4, b3 = NC * F2 * 4;
        mbar_expect_tx(&wbar, b1 + b2 + b3);
        bulk_g2s(wf1s, Wf1 + (size_t)f1_lo * FLAT, b1, &wbar);
        bulk_g2s(wf2s, Wf2, b2, &wbar);
        bulk_g2s(wf3s, Wf3, b3, &wbar);
      }
      if (tid < F1) bfs[tid] = Bf1[tid];
      else if (tid < F1 + F2) bfs[tid] = Bf2[tid - F1];
      else if (tid < F1 + F2 + NC) bfs[tid] = Bf3[tid - F1 - F2];
    }
  }
  const float inv_B = 1.0f / (float)cur_B;

  for (int b = cluster_id; b < cur_B; b += nclusters) {
    float* slab = a.scratch + (size_t)b * S_STRIDE;
    cluster.sync();                    // previous sample fully consumed everywhere before anything is overwritten
    LN_DBG(1);
    {
      const float4* src = reinterpret_cast<const float4*>(cur_x + (size_t)b * XN);
      float4* dst = reinterpret_cast<float4*>(xs);
      for (int i = tid; i < XN / 4; i += kThreads) dst[i] = src[i];
    }
    __syncthreads();
    int lo, hi;
    // ---- conv1 + ReLU + pool: my quarter of the 1176 outputs, broadcast to the cluster
    share(A1N, lo, hi);
    for (int o = lo + tid; o < hi; o += kThreads) {
      const int co = o / (P1 * P1), r = o - co * (P1 * P1), ph = r / P1, pw = r - ph * P1;
      const float bv = w1s[W1N + co];
      float a00 = bv, a01 = bv, a10 = bv, a11 = bv;
#pragma unroll
      for (int ci = 0; ci < CIN; ++ci) {
        const float* xc = xs + ci * H * H + (2 * ph) * H + 2 * pw;
        const float* wk = w1s + (co * CIN + ci) * K * K;
#pragma unroll
        for (int kh = 0; kh < K; ++kh) {
          const float* r0 = xc + kh * H;
          const float* r1 = r0 + H;
          float p0 = r0[0], p1 = r1[0];
#pragma unroll
          for (int kw = 0; kw < K; ++kw) {
            const float wv = wk[kh * K + kw];
            const float q0 = r0[kw + 1], q1 = r1[kw + 1];
            a00 = fmaf(wv, p0, a00); a01 = fmaf(wv, q0, a01);
            a10 = fmaf(wv, p1, a10); a11 = fmaf(wv, q1, a11);
            p0 = q0; p1 = q1;
          }
        }
      }
      float m = a00; int am = 0;
      if (a01 > m) { m = a01; am = 1; }
      if (a10 > m) { m = a10; am = 2; }
      if (a11 > m) { m = a11; am = 3; }
      bcast_f(a1s + o, fmaxf(m, 0.f));
      bcast_i(offs1 + o, (2 * ph + (am >> 1)) * H + 2 * pw + (am & 1));
    }
    cluster.sync();
    LN_DBG(2);
    // ---- conv2 + ReLU + pool: my 100 of the 400 outputs
    share(FLAT, lo, hi);
    for (int o = lo + tid; o < hi; o += kThreads) {
      const int co = o / (P2 * P2), r = o - co * (P2 * P2), ph = r / P2, pw = r - ph * P2;
      const float bv = w2s[W2N + co];
      float a00 = bv, a01 = bv, a10 = bv, a11 = bv;
      for (int ci = 0; ci < C1; ++ci) {
        const float* xc = a1s + ci * P1 * P1 + (2 * ph) * P1 + 2 * pw;
        const float* wk = w2s + (co * C1 + ci) * K * K;
#pragma unroll
        for (int kh = 0; kh < K; ++kh) {
          const float* r0 = xc + kh * P1;
          const float* r1 = r0 + P1;
          float p0 = r0[0], p1 = r1[0];
#pragma unroll
          for (int kw = 0; kw < K; ++kw) {
            const float wv = wk[kh * K + kw];
            const float q0 = r0[kw + 1], q1 = r1[kw + 1];
            a00 = fmaf(wv, p0, a00); a01 = fmaf(wv, q0, a01);
            a10 = fmaf(wv, p1, a10); a11 = fmaf(wv, q1, a11);
            p0 = q0; p1 = q1;
          }
        }
      }
      float m = a00; int am = 0;
      if (a01 > m) { m = a01; am = 1; }
      if (a10 > m) { m = a10; am = 2; }
      if (a11 > m) { m = a11; am = 3; }
      m = fmaxf(m, 0.f);
      bcast_f(a2s + o, m);
      bcast_i(offs2 + o, (2 * ph + (am >> 1)) * P1 + 2 * pw + (am & 1));
      slab[S_A2 + o] = m;
    }
    if (MULTI) mbar_wait_parity(&wbar, (wcount - 1u) & 1u);   // the fc weights of this step have landed
    cluster.sync();
    LN_DBG(3);
    // ---- fc1 + ReLU: my 30 neurons, one warp each
    share(F1, lo, hi);
    {
      const float4* xv = reinterpret_cast<const float4*>(a2s);
      const int n = lo + warp;
      if (n < hi) {
        const float4* wr = reinterpret_cast<const float4*>(pW1 + (size_t)(n - f1_lo) * FLAT);
        float acc = 0.f;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const int c4 = lane + 32 * c;
          if (c4 < FLAT / 4) acc += dot4(wr[c4], xv[c4]);
        }
        acc = warp_sum(acc);
        if (lane == 0) {
          const float v = fmaxf(acc + pB1[n], 0.f);
          bcast_f(z1s + n, v);
          slab[S_Z1 + n] = v;
        }
      }
    }
    cluster.sync();
    LN_DBG(4);
    // ---- fc2 + ReLU: my 21 neurons
    share(F2, lo, hi);
    {
      const float4* xv = reinterpret_cast<const float4*>(z1s);
      const int n = lo + warp;
      if (n < hi) {
        float acc = 0.f;
        if (lane < F1 / 4) acc = dot4(reinterpret_cast<const float4*>(pW2 + (size_t)n * F1)[lane], xv[lane]);
        acc = warp_sum(acc);
        if (lane == 0) {
          const float v = fmaxf(acc + pB2[n], 0.f);
          bcast_f(z2s + n, v);
          slab[S_Z2 + n] = v;
        }
      }
    }
    cluster.sync();
    LN_DBG(5);
    // ---- fc3, loss, dlogits, d2, d1: tiny, every CTA computes them for itself (no exchange needed)
    {
      const float4* xv = reinterpret_cast<const float4*>(z2s);
      if (warp < NC) {
        float acc = 0.f;
        if (lane < F2 / 4) acc = dot4(reinterpret_cast<const float4*>(pW3 + (size_t)warp * F2)[lane], xv[lane]);
        acc = warp_sum(acc);
        if (lane == 0) lgs[warp] = acc + pB3[warp];
      }
    }
    __syncthreads();
    if (warp == 0) {
      const int tgt = (int)cur_y[b];
      const float v = lane < NC ? lgs[lane] : -INFINITY;
      float m = v; int am = lane < NC ? lane : 0x7fffffff;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float om = __shfl_xor_sync(0xffffffffu, m, o);
        const int oa = __shfl_xor_sync(0xffffffffu, am, o);
        if (om > m || (om == m && oa < am)) { m = om; am = oa; }
      }
      const float e = lane < NC ? expf(v - m) : 0.f;
      const float s = warp_sum(e);
      if (lane < NC) {
        float p = e / s;
        if (lane == tgt) p -= 1.0f;
        p *= inv_B;
        dls[lane] = p;
        if (cr == 0) slab[S_DL + lane] = p;
      }
      if (lane == 0 && cr == 0) {
        const float xt = (tgt >= 0 && tgt < NC) ? lgs[tgt] : (logf(s) + m);
        slab[S_DL + 10] = logf(s) + m - xt;
        slab[S_DL + 11] = (am == tgt) ? 1.f : 0.f;
      }
    }
    __syncthreads();
    if (!a.train) continue;
    if (tid < F2) {
      float s = 0.f;
#pragma unroll
      for (int n = 0; n < NC; ++n) s = fmaf(dls[n], pW3[n * F2 + tid], s);
      s = z2s[tid] > 0.f ? s : 0.f;
      d2s[tid] = s;
      if (cr == 0) slab[S_D2 + tid] = s;
    }
    __syncthreads();
    if (tid < 8 * F1) {
      const int k = tid % F1, part = tid / F1;
      float s = 0.f;
      const int n_hi = min(F2, part * 11 + 11);
#pragma unroll 11
      for (int n = part * 11; n < n_hi; ++n) s = fmaf(d2s[n], pW2[n * F1 + k], s);
      dz2[tid] = s;
    }
    __syncthreads();
    if (tid < F1) {
      float s = 0.f;
#pragma unroll
      for (int part = 0; part < 8; ++part) s += dz2[part * F1 + tid];
      s = z1s[tid] > 0.f ? s : 0.f;
      d1s[tid] = s;
      if (cr == 0) slab[S_D1 + tid] = s;
    }
    __syncthreads();
    // ---- backward through fc1: my 30 rows of W1, partial sums exchanged across the cluster
    share(F1, lo, hi);
    {
      const int rg = warp & 15, half = warp >> 4;
      float4 acc[2];
      acc[0] = acc[1] = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int n = lo + rg; n < hi; n += 16) {
        const float dv = d1s[n];
        const float4* wr = reinterpret_cast<const float4*>(pW1 + (size_t)(n - f1_lo) * FLAT);
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          const int c4 = lane + 32 * (2 * half + c);
          if (c4 < FLAT / 4) {
            const float4 w = wr[c4];
            acc[c].x = fmaf(dv, w.x, acc[c].x); acc[c].y = fmaf(dv, w.y, acc[c].y);
            acc[c].z = fmaf(dv, w.z, acc[c].z); acc[c].w = fmaf(dv, w.w, acc[c].w);
          }
        }
      }
      float4* part = reinterpret_cast<float4*>(dz2) + rg * (FLAT / 4);
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        const int c4 = lane + 32 * (2 * half + c);
        if (c4 < FLAT / 4) part[c4] = acc[c];
      }
    }
    __syncthreads();
    for (int k = tid; k < FLAT; k += kThreads) {
      float s = 0.f;
#pragma unroll
      for (int w = 0; w < 16; ++w) s += dz2[w * FLAT + k];
      bcast_f(g2x + cr * FLAT + k, s);
    }
    cluster.sync();
    LN_DBG(6);
    for (int k = tid; k < FLAT; k += kThreads) {
      const float s = g2x[k] + g2x[FLAT + k] + g2x[2 * FLAT + k] + g2x[3 * FLAT + k];
      g2p[k] = a2s[k] > 0.f ? s : 0.f;
    }
    for (int i = tid; i < C2 * ZH * ZH; i += kThreads) dz2[i] = 0.f;
    __syncthreads();
    // ---- conv2 weight / bias gradient partials: my 604 of the 2416 slots
    share(W2N + C2, lo, hi);
    for (int t = lo + tid; t < hi; t += kThreads) {
      float s = 0.f;
      if (t < W2N) {
        const int co = t / (C1 * K * K), r = t - co * (C1 * K * K), ci = r / (K * K), q = r - ci * (K * K);
        const float* xt = a1s + ci * P1 * P1 + (q / K) * P1 + (q % K);
        const float* g = g2p + co * (P2 * P2);
        const int* of = offs2 + co * (P2 * P2);
#pragma unroll
        for (int p = 0; p < P2 * P2; ++p) s = fmaf(g[p], xt[of[p]], s);
      } else {
        const float* g = g2p + (t - W2N) * (P2 * P2);
#pragma unroll
        for (int p = 0; p < P2 * P2; ++p) s += g[p];
      }
      slab[S_PW2 + t] = s;
    }
    for (int o = tid; o < FLAT; o += kThreads) {
      const int co = o / (P2 * P2);
      const int pos = offs2[o];
      const int oh = pos / P1, ow = pos - oh * P1;
      dz2[co * ZH * ZH + (oh + K - 1) * ZH + (ow + K - 1)] = g2p[o];
    }
    __syncthreads();
    // ---- conv2 dgrad: this CTA takes output channels 4*cr .. 4*cr+3, one work item per (co, ci, row, 7-column strip)
    for (int it = tid; it < 4 * C1 * P1 * 2; it += kThreads) {
      const int cl = it / (C1 * P1 * 2), r = it - cl * (C1 * P1 * 2);
      const int ci = r / (P1 * 2), r2 = r - ci * (P1 * 2), h = r2 >> 1, w0 = (r2 & 1) * 7;
      const int co = cr * 4 + cl;
      float acc[7] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      const float* wk = w2s + (co * C1 + ci) * K * K;
#pragma unroll
      for (int kh = 0; kh < K; ++kh) {
        const float* row = dz2 + co * ZH * ZH + (h - kh + K - 1) * ZH + w0;
        float d[11];
#pragma unroll
        for (int j = 0; j < 11; ++j) d[j] = row[j];
#pragma unroll
        for (int kw = 0; kw < K; ++kw) {
          const float wv = wk[kh * K + kw];
#pragma unroll
          for (int j = 0; j < 7; ++j) acc[j] = fmaf(d[j + (K - 1) - kw], wv, acc[j]);
        }
      }
      float* dst = dgp + cl * A1N + ci * P1 * P1 + h * P1 + w0;
#pragma unroll
      for (int j = 0; j < 7; ++j) dst[j] = acc[j];
    }
    __syncthreads();
    for (int o = tid; o < A1N; o += kThreads)
      bcast_f(g1x + cr * A1N + o, dgp[o] + dgp[A1N + o] + dgp[2 * A1N + o] + dgp[3 * A1N + o]);
    cluster.sync();
    LN_DBG(7);
    for (int o = tid; o < A1N; o += kThreads) {
      const float s = g1x[o] + g1x[A1N + o] + g1x[2 * A1N + o] + g1x[3 * A1N + o];
      g1p[o] = a1s[o] > 0.f ? s : 0.f;
    }
    __syncthreads();
    // ---- conv1 weight / bias gradient partials: my 114 of the 456 slots, the 196 positions split 4 ways
    share(W1N + C1, lo, hi);
    {
      const int nslot = hi - lo;
      for (int it = tid; it < nslot * 4; it += kThreads) {
        const int sl = it >> 2, part = it & 3, t = lo + sl;
        const int p_lo = part * 49, p_hi = p_lo + 49;
        float s = 0.f;
        if (t < W1N) {
          const int co = t / (CIN * K * K), r = t - co * (CIN * K * K), ci = r / (K * K), q = r - ci * (K * K);
          const float* xt = xs + ci * H * H + (q / K) * H + (q % K);
          const float* g = g1p + co * (P1 * P1);
          const int* of = offs1 + co * (P1 * P1);
#pragma unroll 7
          for (int p = p_lo; p < p_hi; ++p) s = fmaf(g[p], xt[of[p]], s);
        } else {
          const float* g = g1p + (t - W1N) * (P1 * P1);
          for (int p = p_lo; p < p_hi; ++p) s += g[p];
        }
        w1x[part * 128 + sl] = s;
      }
      __syncthreads();
      for (int sl = tid; sl < nslot; sl += kThreads)
        slab[S_PW1 + lo + sl] = w1x[sl] + w1x[128 + sl] + w1x[256 + sl] + w1x[384 + sl];
    }
  }
  cluster.sync();        // nobody leaves while a sibling may still write into its shared memory
  LN_DBG(8);

  grid_barrier(a.sync, gridDim.x, s_sense);
  LN_DBG(9);
  lenet_phase_b<R>(a, inv_B, cur_B, cur_hyper, cur_loss);
  LN_DBG(10);
  if (multi) {
    // the updated weights (and the loss) must be visible everywhere before the next step reads them
    if (tid == 0) s_sense += 1;
    grid_barrier(a.sync, gridDim.x, s_sense + 0u);
    if (tid == 0) s_sense += 1;
    // (no system fence: the host only uses the word to know the step's INPUT slot has been read; losses are read after the run)
    if (blockIdx.x == 0 && tid == 0 && a.progress) *reinterpret_cast<volatile unsigned int*>(a.progress) = (unsigned int)(step + 1);
  }
  if (stamp_at != 0xFFFFFFFFu) g_lenet_stamps[2 * stamp_at + 1] = ln_gtimer();
  }  // step loop
}

namespace lenet {
constexpr size_t kClusterSmemBytes =
    (size_t)(XN + 456 + A1N + 2416 + FLAT + F1 + 88 + 16 + 16 + 88 + F1 + FLAT + A1N + 6400 + 4 * A1N + CS * FLAT + CS * A1N + 4 * 128) * 4 +
    (size_t)(A1N + FLAT) * 4 + 64;
constexpr size_t kClusterSmemBytesMulti = kClusterSmemBytes + (size_t)(30 * FLAT + F2 * F1 + NC * F2 + 256) * 4;
}

namespace lenet {
constexpr size_t kSmemBytes = (size_t)(XN + 456 + A1N + 2416 + FLAT + F1 + 88 + 16 + 16 + 88 + F1 + FLAT + A1N + 6400 + 4 * A1N) * 4 +
                              (size_t)(A1N + FLAT) * 4 + (A1N + FLAT) + 64;
}

}  // namespace mlt

using namespace mlt;

extern "C" long long mlt_lenet_scratch_bytes(int B) { return (long long)B * lenet::S_STRIDE * 4 + 256; }

// One fused training (train=1) or evaluation (train=0) step of the reference MLModel.  fp32; model == master.
// param_off: arena element offsets of {conv1.w, conv1.b, conv2.w, conv2.b, fc1.w, fc1.b, fc2.w, fc2.b, fc3.w, fc3.b}.
// scratch: mlt_lenet_scratch_bytes(B) bytes, zero-initialised once (its first 256 bytes hold the grid-barrier words).
// ctx: NULL or world == 1 for single-GPU; otherwise gradients are all-reduced in-kernel through the symmetric arena.
template <typename R, bool MULTI>
static int lenet_launch_cluster(LeNetArgs& a, long long numel, cudaStream_t st, int dev) {
  using namespace lenet;
  auto kern = lenet_step_cluster_kernel<R, MULTI>;
  constexpr size_t kSmem = MULTI ? kClusterSmemBytesMulti : kClusterSmemBytes;
  static int grid_cl[64] = {};
  cudaLaunchConfig_t cfg{};
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = kSmem;
  cfg.stream = st;
  cudaLaunchAttribute attrs[2];
  attrs[0].id = cudaLaunchAttributeClusterDimension;
  attrs[0].val.clusterDim.x = CS; attrs[0].val.clusterDim.y = 1; attrs[0].val.clusterDim.z = 1;
  attrs[1].id = cudaLaunchAttributeCooperative;
  attrs[1].val.cooperative = 1;
  cfg.attrs = attrs;
  if (grid_cl[dev & 63] == 0) {
    MLT_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem));
    int sms = 0, nclus = 0;
    MLT_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    cfg.gridDim = dim3((sms / CS) * CS);
    cfg.numAttrs = 1;
    MLT_CHECK_CUDA(cudaOccupancyMaxActiveClusters(&nclus, kern, &cfg));
    MLT_REQUIRE(nclus >= 1, "mlt_lenet_step: no 4-CTA cluster fits on this device");
    grid_cl[dev & 63] = nclus * CS;
  }
  const int grid = grid_cl[dev & 63];
  const long long min_grid = ceil_div64(numel, (long long)kThreads * kMaxPerThread);
  MLT_REQUIRE(grid >= min_grid && grid <= kMaxCtas - 1, "mlt_lenet_step: grid %d cannot cover %lld parameters", grid, numel);
  cfg.gridDim = dim3(grid);
  static bool coop_ok = true;                     // cooperative + cluster launch; fall back to a plain cluster launch
  cfg.numAttrs = coop_ok ? 2 : 1;                 // (all CTAs are co-resident by construction: grid <= active clusters)
  cudaError_t le = cudaLaunchKernelEx(&cfg, kern, a);
  if (le != cudaSuccess && coop_ok) {
    (void)cudaGetLastError();
    coop_ok = false;
    cfg.numAttrs = 1;
    le = cudaLaunchKernelEx(&cfg, kern, a);
  }
  MLT_CHECK_CUDA(le);
  return 0;
}

// Launch the step kernel for the arguments in `a` (single step, or a.nsteps steps in one launch: cluster kernel only).
static int lenet_launch(LeNetArgs& a, int rule, long long numel, void* stream) {
  using namespace lenet;
  cudaStream_t st = (cudaStream_t)stream;
  int dev = 0;
  MLT_CHECK_CUDA(cudaGetDevice(&dev));
  int rc = 0;
  static const bool use_cluster = []{ const char* e = getenv("MLT_LENET_CLUSTER"); return !(e && e[0] == '0'); }();
  MLT_REQUIRE(a.nsteps < 1 || use_cluster, "multi-step launches need the cluster kernel (MLT_LENET_CLUSTER=0 is set)");
  MLT_DISPATCH_RULE(rule, R, {
    if (use_cluster) return a.nsteps >= 1 ? lenet_launch_cluster<R, true>(a, numel, st, dev) : lenet_launch_cluster<R, false>(a, numel, st, dev);
    MLT_REQUIRE(a.nsteps < 1, "multi-step launches need the cluster kernel");
    auto kern = lenet_step_kernel<R>;
    static int grid_cache[64] = {};
    if (grid_cache[dev & 63] == 0) {
      MLT_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes));
      int per_sm = 0, sms = 0;
      MLT_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kThreads, kSmemBytes));
      MLT_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
      MLT_REQUIRE(per_sm >= 1, "mlt_lenet_step: kernel does not fit on an SM");
      grid_cache[dev & 63] = sms;           // one CTA per SM: all co-resident, phase B spreads over the whole chip
    }
    int grid = grid_cache[dev & 63];
    const long long min_grid = ceil_div64(numel, (long long)kThreads * kMaxPerThread);
    MLT_REQUIRE(grid >= min_grid && grid <= kMaxCtas - 1, "mlt_lenet_step: grid %d cannot cover %lld parameters", grid, numel);
    void* params[] = {&a};
    MLT_CHECK_CUDA(cudaLaunchCooperativeKernel((void*)kern, dim3(grid), dim3(kThreads), params, kSmemBytes, st));
  });
  return rc;
}


extern "C" int mlt_lenet_step(void* master, void* s1, void* s2, const void* hyper, long long numel, const long long* param_off,
                              int rule, const void* x, const void* y_i64, int B, int train, int want_acc, void* acc3,
                              void* loss_out, void* scratch, long long scratch_bytes, const mlt_p2p_ctx* ctx, void* stream) {
  using namespace lenet;
  MLT_REQUIRE(B > 0, "mlt_lenet_step: empty batch");
  MLT_REQUIRE(scratch != nullptr && scratch_bytes >= mlt_lenet_scratch_bytes(B), "mlt_lenet_step: scratch too small");
  LeNetArgs a;
  a.master = (float*)master; a.s1 = (float*)s1; a.s2 = (float*)s2; a.hyper = (const Hyper*)hyper; a.numel = numel;
  for (int i = 0; i < NPARAM; ++i) a.off[i] = param_off[i];
  a.x = (const float*)x; a.y = (const long long*)y_i64; a.B = B; a.train = train; a.want_acc = want_acc;
  a.sync = (unsigned int*)scratch;
  a.scratch = (float*)((char*)scratch + 256);
  a.acc3 = (float*)acc3; a.loss_out = (float*)loss_out;
  a.nsteps = 0; a.nslots = 0; a.pool_n = 0; a.batch_sizes = nullptr; a.hyper_table = nullptr; a.loss_base = nullptr;
  a.ready = nullptr; a.progress = nullptr;
  for (int i = 0; i < kMaxSlots; ++i) { a.xs[i] = nullptr; a.ys[i] = nullptr; }
  a.comm.world = 1; a.comm.rank = 0; a.comm.mc_flags = 0; a.comm.recv_bucket_off = 0; a.comm.mc_base = nullptr;
  if (ctx != nullptr && ctx->world > 1) {
    MLT_REQUIRE(numel * 4 <= ctx->recv_slot_bytes, "mlt_lenet_step: parameter arena exceeds the one-shot receive slot");
    for (int i = 0; i < kMaxRanks; ++i) a.comm.base[i] = (char*)ctx->base[i];
    a.comm.mc_base = (char*)ctx->mc_base;
    a.comm.world = ctx->world; a.comm.rank = ctx->rank;
    a.comm.flags_off = ctx->flags_off; a.comm.recv_off = ctx->recv_off; a.comm.recv_slot_bytes = ctx->recv_slot_bytes;
    a.comm.mc_flags = (ctx->mc_flags != 0 && ctx->mc_base != nullptr) ? 1 : 0;
  }
  return lenet_launch(a, rule, numel, stream);
}

// `nsteps` consecutive steps in ONE launch (no launch / drain gap between steps).  Inputs: either a ring of `nslots`
// device buffers (xs[i], ys[i]; step i uses slot i % nslots once *ready >= i + 1, *ready being advanced by whoever fills the
// slots, e.g. a copy stream) or, with pool_n > 0, a device-resident pool of pool_n batches of B samples at xs[0] / ys[0].
// batch_sizes / hyper_table / loss_base: per-step device tables (each may be NULL: B, hyper, no per-step loss).
// progress: optional word (pinned host memory) that receives i + 1 after step i, 0xFFFFFFFF if the kernel gave up waiting.
extern "C" int mlt_lenet_steps(void* master, void* s1, void* s2, const void* hyper, long long numel, const long long* param_off,
                               int rule, const void* const* xs, const void* const* ys, int nslots, long long pool_n, int B,
                               const int* batch_sizes, const void* hyper_table, void* loss_base, const void* ready, void* progress,
                               int nsteps, int train, int want_acc, void* acc3, void* scratch, long long scratch_bytes,
                               const mlt_p2p_ctx* ctx, void* stream) {
  using namespace lenet;
  MLT_REQUIRE(B > 0 && nsteps > 0, "mlt_lenet_steps: empty run");
  MLT_REQUIRE(nslots >= 1 && nslots <= kMaxSlots, "mlt_lenet_steps: 1..%d input slots", kMaxSlots);
  MLT_REQUIRE(scratch != nullptr && scratch_bytes >= mlt_lenet_scratch_bytes(B), "mlt_lenet_steps: scratch too small");
  LeNetArgs a;
  a.master = (float*)master; a.s1 = (float*)s1; a.s2 = (float*)s2; a.hyper = (const Hyper*)hyper; a.numel = numel;
  for (int i = 0; i < NPARAM; ++i) a.off[i] = param_off[i];
  a.x = (const float*)xs[0]; a.y = (const long long*)ys[0]; a.B = B; a.train = train; a.want_acc = want_acc;
  a.sync = (unsigned int*)scratch;
  a.scratch = (float*)((char*)scratch + 256);
  a.acc3 = (float*)acc3; a.loss_out = nullptr;
  a.nsteps = nsteps;
  a.nslots = nslots; a.pool_n = pool_n;
  for (int i = 0; i < kMaxSlots; ++i) { a.xs[i] = i < nslots ? (const float*)xs[i] : nullptr; a.ys[i] = i < nslots ? (const long long*)ys[i] : nullptr; }
  a.batch_sizes = batch_sizes; a.hyper_table = (const Hyper*)hyper_table; a.loss_base = (float*)loss_base;
  a.ready = (const unsigned int*)ready; a.progress = (unsigned int*)progress;
  a.comm.world = 1; a.comm.rank = 0; a.comm.mc_flags = 0; a.comm.recv_bucket_off = 0; a.comm.mc_base = nullptr;
  if (ctx != nullptr && ctx->world > 1) {
    MLT_REQUIRE(numel * 4 <= ctx->recv_slot_bytes, "mlt_lenet_steps: parameter arena exceeds the one-shot receive slot");
    for (int i = 0; i < kMaxRanks; ++i) a.comm.base[i] = (char*)ctx->base[i];
    a.comm.mc_base = (char*)ctx->mc_base;
    a.comm.world = ctx->world; a.comm.rank = ctx->rank;
    a.comm.flags_off = ctx->flags_off; a.comm.recv_off = ctx->recv_off; a.comm.recv_slot_bytes = ctx->recv_slot_bytes;
    a.comm.mc_flags = (ctx->mc_flags != 0 && ctx->mc_base != nullptr) ? 1 : 0;
  }
  return lenet_launch(a, rule, numel, stream);
}

// Analysis hook: device buffer of `pairs` x 2 int64 that CTA 0 of the step kernel fills with {start, end} %globaltimer stamps, one
// pair per training step in launch order (NULL clears it and resets the count).
extern "C" int mlt_lenet_set_stamps(void* buf, int pairs) {
  long long* b = (long long*)buf;
  unsigned int cap = buf ? (unsigned int)pairs : 0u, zero = 0u;
  MLT_CHECK_CUDA(cudaMemcpyToSymbol(mlt::g_lenet_stamps, &b, sizeof(b)));
  MLT_CHECK_CUDA(cudaMemcpyToSymbol(mlt::g_lenet_stamp_cap, &cap, sizeof(cap)));
  MLT_CHECK_CUDA(cudaMemcpyToSymbol(mlt::g_lenet_stamp_n, &zero, sizeof(zero)));
  return 0;
}

// Analysis hook: device buffer of 16 int64 per CTA that the cluster kernel fills with %globaltimer stamps (NULL clears it).
extern "C" int mlt_lenet_set_debug(void* buf) {
  long long* b = (long long*)buf;
  MLT_CHECK_CUDA(cudaMemcpyToSymbol(mlt::g_lenet_dbg, &b, sizeof(b)));
  return 0;
}

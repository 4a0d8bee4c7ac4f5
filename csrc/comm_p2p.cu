// Fused gradient all-reduce + optimizer update over NVLink peer memory.
// Replaces DDP's Reducer -> ncclAllReduce -> copy-back -> torch.optim chain behind reference
// src/trainer.py:98,187-188 (and the dead _average_gradients at :152-158, whose semantics - SUM then divide by
// world - are kept: the 1/world factor rides in hyper.grad_scale).
//
// Every rank maps every peer's symmetric arena (symm.cpp), so kernels address remote HBM with plain
// ld/st.global over NVLink; the multicast mapping adds in-switch reduction (multimem.ld_reduce) and
// broadcast (multimem.st).  Three variants, chosen per bucket size:
//   one-shot   push: every CTA stores its chunk of local gradients into each peer's receive slot, raises a flag,
//              waits for the peers' flags, reduces the N copies from LOCAL memory and updates the whole bucket
//              redundantly.  One NVLink hop of latency; receive slots are double-buffered by step parity so no
//              trailing barrier is needed.
//   two-shot   reduce-scatter + all-gather: rank r pulls slice r of every peer's gradients, reduces, updates its
//              1/N slice of master/state (ZeRO-1-like), then pushes the updated compute-dtype weights to all peers.
//   NVLS       same ownership, but the reduction is one multimem.ld_reduce per vector (done in the NVSwitch) and the
//              weight broadcast one multimem.st.
// All reductions accumulate in fp32 in a fixed rank order, so every replica computes bit-identical weights.
//
// Synchronisation: per-CTA flags in the arena, written with st.release.sys and polled with ld.acquire.sys.  Flag
// values are tickets that only grow (one counter per CTA lives beside the flags), so there is no reset phase and
// CUDA-graph replays keep working.  A bounded spin turns a lost peer into a recorded diagnostic, not a hang.
//
// Testing on ONE GPU: gridDim.y emulates the ranks (rank = blockIdx.y, cooperative launch so that all CTAs are
// co-resident); on real multi-GPU runs gridDim.y == 1 and rank comes from the context.
#include "p2p_device.cuh"
#include "../include/mlt_b200.h"

namespace mlt {

// Apply rule R to kElems consecutive elements starting at element index i of this rank's bucket-local state.
// `g` holds the summed gradients.  Writes master/state and returns the new weights in `pnew`.
template <typename R, int E>
__device__ __forceinline__ void update_elems(const P2PArgs& a, int slot, long long i, float (&g)[8], const Hyper& h, float (&pnew)[8]) {
  float* master = a.master[slot];
#pragma unroll
  for (int j = 0; j < E; j += 4) {
    float4 p = *reinterpret_cast<const float4*>(master + i + j);
    float4 s = *reinterpret_cast<const float4*>(a.s1[slot] + i + j);
    float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
    if (R::kStates > 1) t = *reinterpret_cast<const float4*>(a.s2[slot] + i + j);
    const float4 gg = make_float4(g[j] * h.grad_scale, g[j + 1] * h.grad_scale, g[j + 2] * h.grad_scale, g[j + 3] * h.grad_scale);
    update4<R>(p, gg, s, t, h);
    *reinterpret_cast<float4*>(master + i + j) = p;
    *reinterpret_cast<float4*>(a.s1[slot] + i + j) = s;
    if (R::kStates > 1) *reinterpret_cast<float4*>(a.s2[slot] + i + j) = t;
    pnew[j] = p.x; pnew[j + 1] = p.y; pnew[j + 2] = p.z; pnew[j + 3] = p.w;
  }
}

template <typename G, int E>
__device__ __forceinline__ uint4 pack_weights(const float (&p)[8]) {
  uint4 o;
  if (E == 4) {
    o.x = __float_as_uint(p[0]); o.y = __float_as_uint(p[1]); o.z = __float_as_uint(p[2]); o.w = __float_as_uint(p[3]);
  } else {
    o.x = pack_bf16(p[0], p[1]); o.y = pack_bf16(p[2], p[3]); o.z = pack_bf16(p[4], p[5]); o.w = pack_bf16(p[6], p[7]);
  }
  return o;
}

// =====================================================================================================
// K1: one-shot push all-reduce + update.  grid = (nblocks, emulated_ranks).
// =====================================================================================================
template <typename R, typename G>
__global__ void __launch_bounds__(256)
allreduce_oneshot_update_kernel(const P2PArgs a) {
  constexpr int E = Vec<G>::kElems;
  const int slot = blockIdx.y, me = a.rank + slot, cta = blockIdx.x;
  const Hyper h = *a.hyper[slot];
  const long long nvec = a.n / E;
  const long long per_cta = ceil_div64(nvec, gridDim.x);
  const long long v_lo = (long long)cta * per_cta, v_hi = min(nvec, v_lo + per_cta);
  // step parity for the receive slots comes from the ticket this CTA is about to use
  const unsigned int parity = (flags_of(a, me)->tick[0][cta] + 1) & 1u;
  const char* my_grad = a.base[me] + a.grad_off;
  // 1. push my chunk into slot [parity][me] of every peer
  for (int p = 0; p < a.world; ++p) {
    if (p == me) continue;
    char* dst = a.base[p] + a.recv_off + ((long long)parity * a.world + me) * a.recv_slot_bytes + a.recv_bucket_off;
    for (long long v = v_lo + threadIdx.x; v < v_hi; v += blockDim.x)
      st_peer(dst + v * 16, *reinterpret_cast<const uint4*>(my_grad + v * 16));
  }
  // 2. publish + wait for the same chunk from every peer (a lost peer: leave weights and state untouched)
  if (!cta_exchange(a, me, 0, cta)) return;
  // 3. reduce in rank order from local memory, update, write weights
  const char* recv = a.base[me] + a.recv_off + (long long)parity * a.world * a.recv_slot_bytes + a.recv_bucket_off;
  char* model = a.base[me] + a.model_off;
  for (long long v = v_lo + threadIdx.x; v < v_hi; v += blockDim.x) {
    float g[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int p = 0; p < a.world; ++p) {
      const uint4 raw = (p == me) ? *reinterpret_cast<const uint4*>(my_grad + v * 16)
                                  : *reinterpret_cast<const uint4*>(recv + p * a.recv_slot_bytes + v * 16);
      Vec<G>::add(g, raw);
    }
    float pnew[8];
    update_elems<R, E>(a, slot, v * E, g, h, pnew);
    if (!a.model_is_master) *reinterpret_cast<uint4*>(model + v * 16) = pack_weights<G, E>(pnew);
  }
}

// =====================================================================================================
// K2 / K3: reduce-scatter (pull, or multimem.ld_reduce) + update of the owned slice + all-gather of weights
// (push, or multimem.st).  grid = (nblocks, emulated_ranks).  USE_MC selects the NVLS data path.
// =====================================================================================================
template <typename G> __device__ __forceinline__ void mc_ld_reduce(const void* mc, float (&g)[8]);
template <> __device__ __forceinline__ void mc_ld_reduce<float>(const void* mc, float (&g)[8]) {
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(g[0]), "=f"(g[1]), "=f"(g[2]), "=f"(g[3]) : "l"(mc) : "memory");
}
template <> __device__ __forceinline__ void mc_ld_reduce<__nv_bfloat16>(const void* mc, float (&g)[8]) {
  uint4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(mc) : "memory");
  for (int j = 0; j < 8; ++j) g[j] = 0.f;
  Vec<__nv_bfloat16>::add(g, r);
}
__device__ __forceinline__ void mc_st(void* mc, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
               ::"l"(mc), "f"(__uint_as_float(v.x)), "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)), "f"(__uint_as_float(v.w))
               : "memory");
}

template <typename R, typename G, bool USE_MC>
__global__ void __launch_bounds__(256)
allreduce_twoshot_update_kernel(const P2PArgs a) {
  constexpr int E = Vec<G>::kElems;
  const int slot = blockIdx.y, me = a.rank + slot, cta = blockIdx.x;
  const Hyper h = *a.hyper[slot];
  const long long nvec = a.n / E;
  const long long per_rank = ceil_div64(nvec, a.world);
  const long long r_lo = min(nvec, (long long)me * per_rank), r_hi = min(nvec, r_lo + per_rank);     // my slice
  const long long per_cta = ceil_div64(r_hi - r_lo, gridDim.x);
  const long long v_lo = min(r_hi, r_lo + (long long)cta * per_cta), v_hi = min(r_hi, v_lo + per_cta);
  // 1. everyone's gradients are final (each rank's kernel runs after its backward on the same stream)
  if (!cta_exchange(a, me, 0, cta)) return;
  // 2. reduce my slice, update, scatter the new weights
  for (long long v = v_lo + threadIdx.x; v < v_hi; v += blockDim.x) {
    float g[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (USE_MC) {
      mc_ld_reduce<G>(a.mc_base + a.grad_off + v * 16, g);
    } else {
      for (int p = 0; p < a.world; ++p) {
        const char* src = a.base[p] + a.grad_off + v * 16;
        Vec<G>::add(g, p == me ? *reinterpret_cast<const uint4*>(src) : ld_peer(src));
      }
    }
    float pnew[8];
    update_elems<R, E>(a, slot, v * E, g, h, pnew);
    const uint4 w = pack_weights<G, E>(pnew);
    if (USE_MC) {
      mc_st(a.mc_base + a.model_off + v * 16, w);               // in-switch broadcast, includes myself
    } else {
      for (int p = 0; p < a.world; ++p) {
        if (p == me) { if (!a.model_is_master) *reinterpret_cast<uint4*>(a.base[me] + a.model_off + v * 16) = w; }
        else st_peer(a.base[p] + a.model_off + v * 16, w);
      }
    }
  }
  // 3. all slices have landed everywhere, and nobody still reads my gradients
  cta_exchange(a, me, 1, cta);
}

// =====================================================================================================
// K5: broadcast (init parameter sync, BN buffer sync): root stores into every peer, then a flag exchange.
// K6: small-vector all-reduce (loss / metric sums) and barrier: push into slot [me] of every peer, exchange, sum.
// =====================================================================================================
// The root must not store before every peer's EARLIER work on this stream has finished (a peer's own initialisation
// copy into the same range, enqueued before its broadcast kernel, would otherwise land after - and over - the root's
// data): a first exchange says "my stream has reached the broadcast", only then does the root push.
__global__ void __launch_bounds__(256)
bcast_store_kernel(const P2PArgs a, long long off, long long nbytes, int root) {
  const int slot = blockIdx.y, me = a.rank + slot, cta = blockIdx.x;
  if (!cta_exchange(a, me, 0, cta)) return;
  if (me == root) {
    const long long nvec = nbytes / 16;
    const char* src = a.base[me] + off;
    for (int p = 0; p < a.world; ++p) {
      if (p == me) continue;
      char* dst = a.base[p] + off;
      for (long long v = (long long)cta * blockDim.x + threadIdx.x; v < nvec; v += (long long)gridDim.x * blockDim.x)
        st_peer(dst + v * 16, *reinterpret_cast<const uint4*>(src + v * 16));
    }
  }
  cta_exchange(a, me, 1, cta);
}

// vals: per emulated rank, k floats in/out.  Uses the first k*4 bytes of every one-shot receive slot pair region
// reserved for scalars (scalar_off), double-buffered by ticket parity like the one-shot kernel.
__global__ void __launch_bounds__(32)
reduce_scalars_kernel(const P2PArgs a, long long scalar_off, float* const* vals, int k) {
  const int slot = blockIdx.y, me = a.rank + slot;
  const unsigned int parity = (flags_of(a, me)->tick[0][kMaxCtas - 1] + 1) & 1u;
  float* v = vals[slot];
  for (int p = 0; p < a.world; ++p) {
    float* dst = reinterpret_cast<float*>(a.base[p] + scalar_off) + ((long long)parity * a.world + me) * 32;
    for (int j = threadIdx.x; j < k; j += 32) dst[j] = v[j];
  }
  if (!cta_exchange(a, me, 0, kMaxCtas - 1)) return;
  const float* mine = reinterpret_cast<const float*>(a.base[me] + scalar_off) + (long long)parity * a.world * 32;
  for (int j = threadIdx.x; j < k; j += 32) {
    float s = 0.f;
    for (int p = 0; p < a.world; ++p) s += mine[p * 32 + j];
    v[j] = s;
  }
}

}  // namespace mlt

using namespace mlt;

// Host-visible context mirrors P2PArgs' symmetric part (see include/mlt_b200.h).
extern "C" long long mlt_p2p_flag_block_bytes(void) { return (long long)sizeof(FlagBlock); }

static int fill_args(P2PArgs& a, const mlt_p2p_ctx* c, int ranks_here) {
  MLT_REQUIRE(c->world >= 1 && c->world <= kMaxRanks, "p2p: world size %d out of range (1..%d)", c->world, kMaxRanks);
  MLT_REQUIRE(ranks_here >= 1 && c->rank + ranks_here <= c->world, "p2p: bad emulated rank count %d", ranks_here);
  for (int i = 0; i < kMaxRanks; ++i) a.base[i] = (char*)c->base[i];
  a.mc_base = (char*)c->mc_base;
  a.world = c->world;
  a.rank = c->rank;
  a.flags_off = c->flags_off;
  a.recv_off = c->recv_off;
  a.recv_slot_bytes = c->recv_slot_bytes;
  a.mc_flags = (c->mc_flags != 0 && c->mc_base != nullptr && ranks_here == 1) ? 1 : 0;
  a.recv_bucket_off = 0;
  return 0;
}

template <typename K>
static int launch_maybe_coop(K kernel, dim3 grid, int threads, P2PArgs& a, int ranks_here, cudaStream_t st) {
  if (ranks_here > 1) {       // single-GPU emulation: all CTAs must be co-resident because they wait for each other
    void* params[] = {&a};
    MLT_CHECK_CUDA(cudaLaunchCooperativeKernel((void*)kernel, grid, dim3(threads), params, 0, st));
  } else {
    kernel<<<grid, threads, 0, st>>>(a);
  }
  MLT_LAUNCH_CHECK();
  return 0;
}

// algo: 0 one-shot, 1 two-shot, 2 NVLS.  ranks_here > 1 only for the single-GPU emulation used by the tests.
extern "C" int mlt_p2p_allreduce_update(const mlt_p2p_ctx* ctx, int algo, int rule, int grad_is_bf16, long long grad_off,
                                        long long model_off, int model_is_master, void* const* master, void* const* s1,
                                        void* const* s2, void* const* hyper, long long n, int nblocks, int ranks_here,
                                        void* stream) {
  return mlt_p2p_allreduce_update_at(ctx, algo, rule, grad_is_bf16, grad_off, model_off, model_is_master, master, s1, s2, hyper, n,
                                     nblocks, ranks_here, 0, stream);
}

extern "C" int mlt_p2p_allreduce_update_at(const mlt_p2p_ctx* ctx, int algo, int rule, int grad_is_bf16, long long grad_off,
                                           long long model_off, int model_is_master, void* const* master, void* const* s1,
                                           void* const* s2, void* const* hyper, long long n, int nblocks, int ranks_here,
                                           long long recv_bucket_off, void* stream) {
  P2PArgs a;
  if (fill_args(a, ctx, ranks_here)) return 1;
  MLT_REQUIRE(recv_bucket_off >= 0 && recv_bucket_off % 16 == 0, "p2p: receive-slot offset must be a non-negative multiple of 16");
  a.recv_bucket_off = recv_bucket_off;
  MLT_REQUIRE(n > 0 && n % 8 == 0, "p2p: bucket length %lld must be a positive multiple of 8", n);
  MLT_REQUIRE(grad_off % 16 == 0 && model_off % 16 == 0, "p2p: arena offsets must be 16-byte aligned");
  MLT_REQUIRE(algo != 2 || ctx->mc_base != nullptr, "p2p: NVLS variant requested but no multicast mapping exists");
  const long long bytes = n * (grad_is_bf16 ? 2 : 4);
  MLT_REQUIRE(algo != 0 || recv_bucket_off + bytes <= ctx->recv_slot_bytes,
              "p2p: one-shot bucket of %lld bytes at slot offset %lld exceeds the %lld-byte receive slot", bytes, recv_bucket_off,
              ctx->recv_slot_bytes);
  a.grad_off = grad_off; a.model_off = model_off; a.n = n; a.model_is_master = model_is_master;
  for (int i = 0; i < ranks_here; ++i) {
    a.master[i] = (float*)master[i]; a.s1[i] = (float*)s1[i]; a.s2[i] = s2 ? (float*)s2[i] : nullptr;
    a.hyper[i] = (const Hyper*)hyper[i];
  }
  if (nblocks <= 0) {
    const long long vecs = n / (grad_is_bf16 ? 8 : 4) / (algo == 0 ? 1 : ctx->world);
    nblocks = (int)ceil_div64(vecs, 256 * 4);        // ~4 vectors per thread
    const int cap = algo == 0 ? 64 : 128;            // comm kernels share the GPU with backward: leave SMs for it
    if (nblocks > cap) nblocks = cap;
    if (nblocks < 1) nblocks = 1;
  }
  MLT_REQUIRE(nblocks <= kMaxCtas - 1, "p2p: at most %d CTAs per collective", kMaxCtas - 1);
  dim3 grid(nblocks, ranks_here);
  cudaStream_t st = (cudaStream_t)stream;
  MLT_DISPATCH_RULE(rule, R, {
    if (algo == 0) {
      if (grad_is_bf16) return launch_maybe_coop(allreduce_oneshot_update_kernel<R, __nv_bfloat16>, grid, 256, a, ranks_here, st);
      return launch_maybe_coop(allreduce_oneshot_update_kernel<R, float>, grid, 256, a, ranks_here, st);
    } else if (algo == 1) {
      if (grad_is_bf16) return launch_maybe_coop(allreduce_twoshot_update_kernel<R, __nv_bfloat16, false>, grid, 256, a, ranks_here, st);
      return launch_maybe_coop(allreduce_twoshot_update_kernel<R, float, false>, grid, 256, a, ranks_here, st);
    } else {
      if (grad_is_bf16) return launch_maybe_coop(allreduce_twoshot_update_kernel<R, __nv_bfloat16, true>, grid, 256, a, ranks_here, st);
      return launch_maybe_coop(allreduce_twoshot_update_kernel<R, float, true>, grid, 256, a, ranks_here, st);
    }
  });
  return 0;
}

extern "C" int mlt_p2p_broadcast(const mlt_p2p_ctx* ctx, long long off, long long nbytes, int root, int ranks_here, void* stream) {
  P2PArgs a;
  if (fill_args(a, ctx, ranks_here)) return 1;
  MLT_REQUIRE(off % 16 == 0 && nbytes % 16 == 0, "p2p broadcast: offset and size must be multiples of 16 bytes");
  int nblocks = (int)ceil_div64(nbytes / 16, 256 * 8);
  if (nblocks > 64) nblocks = 64;
  if (nblocks < 1) nblocks = 1;
  dim3 grid(nblocks, ranks_here);
  cudaStream_t st = (cudaStream_t)stream;
  if (ranks_here > 1) {
    void* params[] = {&a, &off, &nbytes, &root};
    MLT_CHECK_CUDA(cudaLaunchCooperativeKernel((void*)bcast_store_kernel, grid, dim3(256), params, 0, st));
  } else {
    bcast_store_kernel<<<grid, 256, 0, st>>>(a, off, nbytes, root);
  }
  MLT_LAUNCH_CHECK();
  return 0;
}

// vals_dev: device array of `ranks_here` pointers, each to k (<= 32) floats summed across ranks in place.
extern "C" int mlt_p2p_reduce_scalars(const mlt_p2p_ctx* ctx, long long scalar_off, void* vals_dev, int k, int ranks_here,
                                      void* stream) {
  P2PArgs a;
  if (fill_args(a, ctx, ranks_here)) return 1;
  MLT_REQUIRE(k >= 1 && k <= 32, "p2p reduce_scalars: k=%d out of range (1..32)", k);
  dim3 grid(1, ranks_here);
  cudaStream_t st = (cudaStream_t)stream;
  float* const* vals = (float* const*)vals_dev;
  if (ranks_here > 1) {
    void* params[] = {&a, &scalar_off, &vals, &k};
    MLT_CHECK_CUDA(cudaLaunchCooperativeKernel((void*)reduce_scalars_kernel, grid, dim3(32), params, 0, st));
  } else {
    reduce_scalars_kernel<<<grid, 32, 0, st>>>(a, scalar_off, vals, k);
  }
  MLT_LAUNCH_CHECK();
  return 0;
}

static int p2p_eval_error(const mlt_p2p_ctx* ctx, const unsigned int* e, char* fb);

// The epoch-end form for host values: pin (pinned host memory, >= 40 x 4 bytes) holds k floats on entry; on return pin[0..k) holds
// their sums over all ranks and the timeout diagnostic of this rank has been read in the same round trip - ONE stream
// synchronisation instead of a pageable upload, a clone, a read-back and a separate error read.  vals_dev_table / vals_dev: the
// device pointer table and the k-float device buffer it points to (as for mlt_p2p_reduce_scalars).  Returns like
// mlt_p2p_check_error.
extern "C" int mlt_p2p_reduce_scalars_host(const mlt_p2p_ctx* ctx, long long scalar_off, void* vals_dev_table, void* vals_dev, void* pin,
                                           int k, void* stream) {
  MLT_REQUIRE(k >= 1 && k <= 32, "p2p reduce_scalars: k=%d out of range (1..32)", k);
  cudaStream_t st = (cudaStream_t)stream;
  float* hp = (float*)pin;
  char* fb = (char*)ctx->base[ctx->rank] + ctx->flags_off + offsetof(FlagBlock, err);
  MLT_CHECK_CUDA(cudaMemcpyAsync(vals_dev, hp, (size_t)k * sizeof(float), cudaMemcpyHostToDevice, st));
  if (mlt_p2p_reduce_scalars(ctx, scalar_off, vals_dev_table, k, 1, stream)) return 1;
  MLT_CHECK_CUDA(cudaMemcpyAsync(hp, vals_dev, (size_t)k * sizeof(float), cudaMemcpyDeviceToHost, st));
  MLT_CHECK_CUDA(cudaMemcpyAsync(hp + 32, fb, 8 * sizeof(unsigned int), cudaMemcpyDeviceToHost, st));
  MLT_CHECK_CUDA(cudaStreamSynchronize(st));
  return p2p_eval_error(ctx, reinterpret_cast<const unsigned int*>(hp + 32), fb);
}

// Reads (and clears) the timeout diagnostic of the local arena.  out6 = {code, rank, peer, cta, expected, seen}.
extern "C" int mlt_p2p_check_error(const mlt_p2p_ctx* ctx, unsigned int* out6_host) {
  unsigned int e[8];
  char* fb = (char*)ctx->base[ctx->rank] + ctx->flags_off + offsetof(FlagBlock, err);
  MLT_CHECK_CUDA(cudaMemcpy(e, fb, sizeof(e), cudaMemcpyDeviceToHost));
  for (int i = 0; i < 6; ++i) out6_host[i] = e[i];
  return p2p_eval_error(ctx, e, fb);
}

static int p2p_eval_error(const mlt_p2p_ctx* ctx, const unsigned int* e, char* fb) {
  if (e[0] == 1) {
    MLT_CHECK_CUDA(cudaMemset(fb, 0, 8 * sizeof(unsigned int)));
    set_error("p2p: rank %u timed out waiting for peer %u (cta %u, kind %u): expected ticket %u, saw %u; the update was skipped",
              e[1], e[2], e[3], e[6], e[4], e[5]);
    return 2;
  }
  if (e[0] != 0) {
    MLT_CHECK_CUDA(cudaMemset(fb, 0, 8 * sizeof(unsigned int)));
    set_error("p2p: rank %d was stopped by a peer that timed out waiting in a collective; the update was skipped", ctx->rank);
    return 3;
  }
  return 0;
}

// Wait limit of every flag wait on this rank, in seconds (<= 0 restores the default of ~20 s).
extern "C" int mlt_p2p_set_timeout(const mlt_p2p_ctx* ctx, double seconds) {
  int dev = 0, khz = 0;
  MLT_CHECK_CUDA(cudaGetDevice(&dev));
  MLT_CHECK_CUDA(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev));
  const unsigned long long cycles = seconds > 0 ? (unsigned long long)(seconds * 1e3 * (double)khz) : 0ull;
  char* fb = (char*)ctx->base[ctx->rank] + ctx->flags_off + offsetof(FlagBlock, spin_limit);
  MLT_CHECK_CUDA(cudaMemcpy(fb, &cycles, sizeof(cycles), cudaMemcpyHostToDevice));
  return 0;
}

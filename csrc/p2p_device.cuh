// Device-side building blocks of the NVLink communication kernels: arena flag block, sys-scope flag
// exchange between same-index CTAs of all ranks, peer load/store helpers.  Shared by comm_p2p.cu and lenet_step.cu.
#pragma once
#include "optim_rules.cuh"

namespace mlt {

constexpr int kMaxRanks = 8;
constexpr int kMaxCtas = 512;            // flag rows per kind
constexpr long long kSpinTimeoutCycles = 40LL * 1000 * 1000 * 1000;   // default wait limit: ~20 s at 2 GHz (mlt_p2p_set_timeout overrides)

// Flag block layout (uint32 units), identical on every rank:
//   sig[kind][cta][src_rank]   written by peers        kind 0 = "data pushed / grads ready", 1 = "done"
//   tick[kind][cta]            local ticket counters
//   err[8]                     first timeout diagnostic {code, rank, peer, cta, expected, seen, kind}.  err[0] is STICKY: once it
//                              is non-zero (1 = this rank timed out waiting, 2 = a peer told us it did) every later collective
//                              on this rank returns without waiting and without touching weights or optimizer state, until the
//                              host has read and cleared it (mlt_p2p_check_error) - a lost peer stops training, it never
//                              trains on stale data.
//   spin_limit                 wait limit in clock64() cycles (0 = kSpinTimeoutCycles)
struct FlagBlock {
  unsigned int sig[2][kMaxCtas][kMaxRanks];
  unsigned int tick[2][kMaxCtas];
  unsigned int err[8];
  unsigned long long spin_limit;
};

struct P2PArgs {
  char* base[kMaxRanks];     // arena base of every rank, as mapped in this process
  char* mc_base;             // multicast mapping (nullptr when NVLS is unavailable)
  int world;
  int rank;                  // rank of gridDim.y index 0
  long long flags_off;       // FlagBlock offset in the arena
  long long recv_off;        // one-shot receive slots: [parity][src][recv_slot_bytes]
  long long recv_slot_bytes;
  long long grad_off;        // bucket gradient slice offset in the arena (bytes)
  long long model_off;       // bucket compute-dtype weight slice offset in the arena (bytes)
  // per emulated rank (index blockIdx.y): local, non-symmetric optimizer state of the bucket slice
  float* master[kMaxRanks];
  float* s1[kMaxRanks];
  float* s2[kMaxRanks];
  const Hyper* hyper[kMaxRanks];
  long long n;               // elements in the bucket (multiple of 128 by arena construction)
  int model_is_master;       // fp32 compute: the arena weight slice IS the master (no separate shadow)
  int mc_flags;              // flags (and the LeNet step's gradient push) go through the multicast mapping: one store reaches all
  long long recv_bucket_off; // one-shot: this bucket's byte offset inside every receive slot
};

__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_relaxed_sys(unsigned int* p, unsigned int v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ FlagBlock* flags_of(const P2PArgs& a, int r) {
  return reinterpret_cast<FlagBlock*>(a.base[r] + a.flags_off);
}
// One store through the multicast mapping lands at the same arena offset on EVERY rank (the NVSwitch replicates it).
__device__ __forceinline__ void mc_st_u32(void* mc, unsigned int v) {
  asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(mc), "f"(__uint_as_float(v)) : "memory");
}
__device__ __forceinline__ void mc_st_f32(void* mc, float v) {
  asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(mc), "f"(v) : "memory");
}

__device__ __forceinline__ unsigned int ld_volatile_u32(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Signal every peer's sig[kind][cta][me] with a fresh ticket, then wait until every peer's ticket reached it.
// Must be called by all threads of the CTA (contains __syncthreads); data written before the call by any
// thread of this CTA is visible to peers that observe the flag, and vice versa.
// Returns false (to every thread of the CTA) when the rank is poisoned: a wait ran past the limit here or on a peer.  The
// caller must then skip its update and return; nothing after a failed exchange may be trusted.
__device__ __forceinline__ bool cta_exchange(const P2PArgs& a, int me, int kind, int cta, bool do_signal = true, bool do_wait = true,
                                             unsigned int* ticket_io = nullptr) {
  __shared__ unsigned int s_ticket;
  __shared__ int s_bad;
  __syncthreads();
  if (threadIdx.x == 0) {
    FlagBlock* mine = flags_of(a, me);
    s_bad = ld_volatile_u32(&mine->err[0]) != 0u;
    unsigned int t;
    if (do_signal) {
      t = mine->tick[kind][cta] + 1;
      mine->tick[kind][cta] = t;
      // ONE system-scope fence orders every data store of this CTA (the bar.sync above made them this thread's
      // business) before the flags; the flags themselves are relaxed stores.  A release store per peer would pay
      // one NVLink round trip each, serially (measured: ~5 us per peer).
      __threadfence_system();
      if (a.mc_flags) {      // one store, replicated by the switch (it also lands in my own, unused, sig[..][me])
        mc_st_u32(&reinterpret_cast<FlagBlock*>(a.mc_base + a.flags_off)->sig[kind][cta][me], t);
      } else {
        for (int p = 0; p < a.world; ++p)
          if (p != me) st_relaxed_sys(&flags_of(a, p)->sig[kind][cta][me], t);
      }
    } else {
      t = *ticket_io;
    }
    s_ticket = t;
  }
  __syncthreads();
  const unsigned int t = s_ticket;
  if (ticket_io && threadIdx.x == 0) *ticket_io = t;
  if (do_wait) {
    if (threadIdx.x < a.world && (int)threadIdx.x != me && !s_bad) {
      FlagBlock* mine = flags_of(a, me);
      const unsigned int* f = &mine->sig[kind][cta][threadIdx.x];
      const long long t0 = clock64();
      const unsigned long long lim = mine->spin_limit;
      const long long limit = lim ? (long long)lim : kSpinTimeoutCycles;
      unsigned int seen, spins = 0;
      while ((int)((seen = ld_acquire_sys(f)) - t) < 0) {
        if ((++spins & 63u) != 0u) continue;
        if (ld_volatile_u32(&mine->err[0]) != 0u) { s_bad = 1; break; }          // poisoned meanwhile (by another CTA or a peer)
        if (clock64() - t0 > limit) {
          unsigned int* e = mine->err;
          if (atomicCAS(&e[0], 0u, 1u) == 0u) { e[1] = me; e[2] = threadIdx.x; e[3] = cta; e[4] = t; e[5] = seen; e[6] = kind; }
          for (int p = 0; p < a.world; ++p)                                        // tell the others: they stop at once instead of
            if (p != me) atomicCAS_system(&flags_of(a, p)->err[0], 0u, 2u);       // running into their own limit
          s_bad = 1;
          break;
        }
      }
    }
    __syncthreads();
  }
  return s_bad == 0;
}

// ---- vector helpers: 8 (bf16) or 4 (fp32) elements per 16 bytes -------------------------------------------------
template <typename G> struct Vec;
template <> struct Vec<float> {
  static constexpr int kElems = 4;
  __device__ static void add(float (&acc)[8], const uint4& raw) {
    acc[0] += __uint_as_float(raw.x); acc[1] += __uint_as_float(raw.y); acc[2] += __uint_as_float(raw.z); acc[3] += __uint_as_float(raw.w);
  }
};
template <> struct Vec<__nv_bfloat16> {
  static constexpr int kElems = 8;
  __device__ static void add(float (&acc)[8], const uint4& raw) {
    const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[j]));
      acc[2 * j] += f.x; acc[2 * j + 1] += f.y;
    }
  }
};

__device__ __forceinline__ uint4 ld_peer(const void* p) {      // peer HBM: bypass L1, nothing to keep
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_peer(void* p, const uint4& v) {
  asm volatile("st.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

}  // namespace mlt

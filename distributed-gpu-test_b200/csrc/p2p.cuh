// Device-side view of the peer-memory exchange set up by b200ds_ctx_enable_p2p (p2p.cu), shared by
// the kernels that publish into it and read from it.
#pragma once
#include "ctx.cuh"
#include <string.h>

// p2p.cu: a flag wait that timed out in an earlier launch left a note in the context's mailbox (not exported)
extern "C" int p2p_check_mailbox(const b200ds_ctx* ctx, const char* who);

namespace b200ds {

struct P2PDev {
  int world, rank;
  int64_t cap;                              // floats per half of an exchange buffer
  float* xchg[B200DS_P2P_MAX_WORLD];        // every rank's exchange buffer (own entry = local memory): [2 halves][cap], pulled from
  float* push[B200DS_P2P_MAX_WORLD];        // the part of it the peers store into: [2 halves][world slots][cap], slot r written by rank r
  unsigned long long* flags[B200DS_P2P_MAX_WORLD];   // every rank's flag row: flags[p][r] = last epoch rank r published to p
  unsigned long long* epoch;                // local: epochs completed
  unsigned int* tickets;                    // local: [0] copy-phase arrivals, [1] exit arrivals
  unsigned long long timeout_ns;            // bound of one flag wait
  unsigned long long* mailbox;              // pinned host words a timed-out wait reports into before it traps
};

__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// peer data: never through this SM's L1 (the line may be a stale copy from two epochs ago)
__device__ __forceinline__ float4 ld_peer(const float4* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ float ld_peer1(const float* p) {
  float v;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}

// Wait until *flag >= e.  Ranks drift apart by far more than a spin count covers (first-call module load, a graph
// capture, rank 0 writing a checkpoint), so the bound is wall time on %globaltimer, generous and configurable.  A
// wait that still runs out says which rank it waited for in the pinned mailbox (readable by the host after the
// trap) and traps: a hang would take the box down with it.
__device__ __forceinline__ void p2p_wait_flag(const unsigned long long* flag, unsigned long long e, int peer, const struct P2PDev& d);

// words of the local flag allocation: flag row (MAX_WORLD x u64) | epoch (u64) | tickets (2 x u32)
inline P2PDev make_p2p_dev(const b200ds_ctx* ctx) {
  P2PDev d;
  memset(&d, 0, sizeof(d));
  d.world = ctx->world; d.rank = ctx->rank; d.cap = ctx->p2p_cap;
  for (int w = 0; w < ctx->world; ++w) {
    d.xchg[w] = ctx->p2p_xchg[w];
    d.push[w] = ctx->p2p_xchg[w] ? ctx->p2p_xchg[w] + 2 * ctx->p2p_cap : nullptr;
    d.flags[w] = ctx->p2p_flags[w];
  }
  unsigned long long* words = ctx->p2p_flags[ctx->rank];
  if (words) {
    d.epoch = words + B200DS_P2P_MAX_WORLD;
    d.tickets = reinterpret_cast<unsigned int*>(words + B200DS_P2P_MAX_WORLD + 1);
  }
  d.timeout_ns = ctx->p2p_timeout_ns;
  d.mailbox = ctx->p2p_mailbox_dev;
  return d;
}

__device__ __forceinline__ void p2p_wait_flag(const unsigned long long* flag, unsigned long long e, int peer, const P2PDev& d) {
  if (ld_acquire_sys(flag) >= e) return;
  const unsigned long long t0 = global_timer_ns();
  unsigned int polls = 0;
  while (ld_acquire_sys(flag) < e) {
    if ((++polls & 1023u) == 0 && global_timer_ns() - t0 > d.timeout_ns) {
      if (d.mailbox) {
        d.mailbox[1] = e;
        d.mailbox[0] = 1ull + (unsigned long long)peer;
        __threadfence_system();
      }
      __trap();
    }
  }
}

}  // namespace b200ds

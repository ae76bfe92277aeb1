// Peer-memory region layout and packet helpers shared by csrc/peer.cu (stand-alone all-reduce) and csrc/norm.cu
// (BatchNorm statistic kernels with the exchange fused in).  See peer.cu for the protocol.
#pragma once
#include "common.cuh"

namespace gsb {

constexpr int kPeerMaxWorld = 8;
constexpr int kPeerSlots = 4;
constexpr int kPeerMaxN = 4096;  // doubles per call
constexpr size_t kPeerHeaderBytes = 1024;
constexpr size_t kPeerRegionBytes = kPeerHeaderBytes + (size_t)kPeerSlots * kPeerMaxWorld * kPeerMaxN * 16;

struct PeerHeader {
  unsigned long long counter;  // calls issued by the owner
  unsigned long long error;    // != 0: sequence number of a call whose wait timed out (read by gsb_peer_error)
  unsigned int done;           // multi-CTA callers: CTAs finished in the current call (the last one bumps `counter`)
};

// Wait limit of one exchange in nanoseconds (GSB_PEER_TIMEOUT_S, default 1800 s; 0 = wait for ever like a blocked NCCL
// kernel).  Ranks legitimately drift apart by seconds to minutes in real training (rank 0 validates, checkpoints or logs,
// a data loader stalls), so the limit is a last-resort watchdog: on expiry the kernel records the call's sequence number
// in its region's error word, keeps its LOCAL value and returns; the host sees it through gsb_peer_error() and raises an
// ordinary RuntimeError.  Nothing traps, the CUDA context stays usable.
long long peer_timeout_ns();   // peer.cu
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

struct PeerRegions {
  void* r[kPeerMaxWorld];
};

__device__ __forceinline__ void st_packet(void* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ uint4 ld_packet(const void* p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}


// Sum `v` (this rank's value of element `elem`) over the ranks: packet into every region, poll own region, add in rank
// order.  `tag` / `slot` come from the call's sequence number (PeerHeader.counter + 1).  Returns false on time-out
// (v keeps the local value).
__device__ __forceinline__ bool peer_exchange_sum(double& v, int elem, const PeerRegions& regs, int rank, int world, uint32_t tag,
                                                  int slot, long long timeout_ns) {
  const unsigned long long u = (unsigned long long)__double_as_longlong(v);
  const size_t my_off = kPeerHeaderBytes + ((size_t)slot * kPeerMaxWorld + rank) * kPeerMaxN * 16 + (size_t)elem * 16;
  for (int p = 0; p < world; ++p)
    st_packet(static_cast<char*>(regs.r[(rank + p) % world]) + my_off, (uint32_t)u, tag, (uint32_t)(u >> 32), tag);
  const char* base = static_cast<const char*>(regs.r[rank]) + kPeerHeaderBytes + (size_t)slot * kPeerMaxWorld * kPeerMaxN * 16;
  const unsigned long long t0 = global_ns();
  double a = 0.0;
  for (int q = 0; q < world; ++q) {
    const char* src = base + ((size_t)q * kPeerMaxN + elem) * 16;
    uint4 pk = ld_packet(src);
    unsigned int spins = 0;
    while (pk.y != tag || pk.w != tag) {
      if (timeout_ns > 0 && (++spins & 1023u) == 0 && global_ns() - t0 > (unsigned long long)timeout_ns) return false;
      pk = ld_packet(src);
    }
    a += __longlong_as_double((long long)(((unsigned long long)pk.z << 32) | (unsigned long long)pk.x));
  }
  v = a;
  return true;
}

}  // namespace gsb

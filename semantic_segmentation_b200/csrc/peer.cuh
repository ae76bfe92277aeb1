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
  unsigned long long error;    // != 0: a wait timed out
  unsigned int done;           // multi-CTA callers: CTAs finished in the current call (the last one bumps `counter`)
};

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
// order.  `tag` / `slot` come from the call's sequence number (PeerHeader.counter + 1).  Returns false on time-out.
__device__ __forceinline__ bool peer_exchange_sum(double& v, int elem, const PeerRegions& regs, int rank, int world, uint32_t tag,
                                                  int slot) {
  const unsigned long long u = (unsigned long long)__double_as_longlong(v);
  const size_t my_off = kPeerHeaderBytes + ((size_t)slot * kPeerMaxWorld + rank) * kPeerMaxN * 16 + (size_t)elem * 16;
  for (int p = 0; p < world; ++p)
    st_packet(static_cast<char*>(regs.r[(rank + p) % world]) + my_off, (uint32_t)u, tag, (uint32_t)(u >> 32), tag);
  const char* base = static_cast<const char*>(regs.r[rank]) + kPeerHeaderBytes + (size_t)slot * kPeerMaxWorld * kPeerMaxN * 16;
  const long long t0 = clock64();
  double a = 0.0;
  for (int q = 0; q < world; ++q) {
    const char* src = base + ((size_t)q * kPeerMaxN + elem) * 16;
    uint4 pk = ld_packet(src);
    while (pk.y != tag || pk.w != tag) {
      if (clock64() - t0 > 4000000000ll) return false;  // ~2 s: a peer is gone
      pk = ld_packet(src);
    }
    a += __longlong_as_double((long long)(((unsigned long long)pk.z << 32) | (unsigned long long)pk.x));
  }
  v = a;
  return true;
}

}  // namespace gsb

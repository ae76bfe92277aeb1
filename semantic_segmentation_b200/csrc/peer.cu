// One-shot all-reduce of small fp64 vectors over NVLink peer memory: the sync-BN exchange steps
// (forward [sum x, sum x^2], backward [sum dy, sum dy*xhat], <= 16 KB per layer) and the four CE / Dice sums.
//
// 36 of these sit on the critical path of one U-Net step (one before every BatchNorm apply, forward and backward),
// so what matters is latency, not bandwidth.  Every rank owns a cudaMalloc'ed REGION that its peers map with CUDA
// IPC.  One 1024-thread CTA per rank:
//   1. takes the next sequence number from a counter in its own region (the same call order on every rank, so the
//      numbers agree; a device-side counter keeps CUDA-graph replays correct);
//   2. each thread stores its elements into slot (seq % kSlots, rank) of EVERY region as self-validating 16-byte
//      packets {lo32, seq, hi32, seq} (remote stores over NVLink / NVSwitch) -- no fence, no separate flag: an
//      8-byte half is valid exactly when its tag equals seq (the LL idea of NCCL, sized for doubles);
//   3. the same thread polls the `world` packets of its elements in its own region until the tags match, sums them
//      in rank order -- every rank adds the same numbers in the same order, so the result is bit-identical on all
//      ranks -- and writes the sum over its input.
// One NVLink one-way latency per call instead of store + system fence + flag + acquire.  A slot is reused after
// kSlots calls; a peer can be at most one call ahead (call seq + 2 needs this rank's packets of call seq + 1, which
// are sent only after call seq has been read), so kSlots = 4 leaves margin and a stale tag never equals seq.  A rank
// that waits longer than GSB_PEER_TIMEOUT_S (default 30 minutes; 0 = no limit) records the call in its region's error word,
// keeps its local values and returns -- no trap; the host polls gsb_peer_error().  Every rank must issue the same sequence
// of exchanges (train-mode BatchNorm layers, seg_loss calls): a rank that skips one leaves the others waiting.
#include <stdlib.h>

#include "common.cuh"
#include "peer.cuh"

namespace gsb {

long long peer_timeout_ns() {
  static long long ns = -1;
  if (ns < 0) {
    const char* e = getenv("GSB_PEER_TIMEOUT_S");
    double sec = e != nullptr && e[0] != 0 ? atof(e) : 1800.0;
    if (sec < 0) sec = 0;
    ns = (long long)(sec * 1e9);
  }
  return ns;
}

constexpr int kPeerThreads = 1024;
constexpr int kPeerPerThread = kPeerMaxN / kPeerThreads;

__global__ void __launch_bounds__(kPeerThreads, 1) peer_allreduce_f64_kernel(double* __restrict__ buf, int n, PeerRegions regs,
                                                                              int rank, int world, long long timeout_ns) {
  GSB_PDL_ENTRY();
  __shared__ unsigned long long s_seq;
  PeerHeader* mine = reinterpret_cast<PeerHeader*>(regs.r[rank]);
  if (threadIdx.x == 0) s_seq = ++mine->counter;
  __syncthreads();
  const uint32_t tag = (uint32_t)s_seq;  // never 0 in the first 2^32 calls; stale slots hold tag - kSlots
  const int slot = (int)(s_seq % kPeerSlots);
  const size_t my_off = kPeerHeaderBytes + ((size_t)slot * kPeerMaxWorld + rank) * kPeerMaxN * 16;
  // 2. my elements into every region (own region included: the sum below treats all ranks alike)
#pragma unroll
  for (int j = 0; j < kPeerPerThread; ++j) {
    const int i = threadIdx.x + j * kPeerThreads;
    if (i < n) {
      const unsigned long long u = (unsigned long long)__double_as_longlong(buf[i]);
      for (int p = 0; p < world; ++p) {
        char* dst = static_cast<char*>(regs.r[(rank + p) % world]) + my_off + (size_t)i * 16;
        st_packet(dst, (uint32_t)u, tag, (uint32_t)(u >> 32), tag);
      }
    }
  }
  // 3. poll the packets of my elements, sum in rank order
  const char* base = reinterpret_cast<const char*>(mine) + kPeerHeaderBytes + (size_t)slot * kPeerMaxWorld * kPeerMaxN * 16;
  const unsigned long long t0 = global_ns();
#pragma unroll
  for (int j = 0; j < kPeerPerThread; ++j) {
    const int i = threadIdx.x + j * kPeerThreads;
    if (i < n) {
      double a = 0.0;
      bool ok = true;
      for (int q = 0; q < world && ok; ++q) {
        const char* src = base + ((size_t)q * kPeerMaxN + i) * 16;
        uint4 v = ld_packet(src);
        unsigned int spins = 0;
        while (v.y != tag || v.w != tag) {
          if (timeout_ns > 0 && (++spins & 1023u) == 0 && global_ns() - t0 > (unsigned long long)timeout_ns) {
            ok = false;   // watchdog: record the call, keep the local value, return normally (host: gsb_peer_error)
            break;
          }
          v = ld_packet(src);
        }
        a += __longlong_as_double((long long)(((unsigned long long)v.z << 32) | (unsigned long long)v.x));
      }
      if (ok) buf[i] = a;
      else mine->error = s_seq;
    }
  }
}

}  // namespace gsb

using namespace gsb;

extern "C" int64_t gsb_peer_region_bytes(void) { return (int64_t)kPeerRegionBytes; }
extern "C" int gsb_peer_max_elems(void) { return kPeerMaxN; }
extern "C" int gsb_peer_handle_bytes(void) { return (int)sizeof(cudaIpcMemHandle_t); }

// Allocate and zero this rank's region on the current device; writes its CUDA IPC handle (gsb_peer_handle_bytes() bytes).
extern "C" int gsb_peer_create(void** region, void* handle_out) {
  GSB_REQUIRE(region && handle_out, "peer_create: null pointer");
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, kPeerRegionBytes);
  GSB_REQUIRE(e == cudaSuccess, "peer_create: cudaMalloc failed: %s", cudaGetErrorString(e));
  e = cudaMemset(p, 0, kPeerRegionBytes);
  GSB_REQUIRE(e == cudaSuccess, "peer_create: cudaMemset failed: %s", cudaGetErrorString(e));
  cudaIpcMemHandle_t h;
  e = cudaIpcGetMemHandle(&h, p);
  GSB_REQUIRE(e == cudaSuccess, "peer_create: cudaIpcGetMemHandle failed: %s", cudaGetErrorString(e));
  memcpy(handle_out, &h, sizeof(h));
  e = cudaDeviceSynchronize();
  GSB_REQUIRE(e == cudaSuccess, "peer_create: sync failed: %s", cudaGetErrorString(e));
  *region = p;
  return 0;
}

// Map a peer's region (handle from gsb_peer_create in another process of this node) into this process.
extern "C" int gsb_peer_open(const void* handle, void** peer_region) {
  GSB_REQUIRE(handle && peer_region, "peer_open: null pointer");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  void* p = nullptr;
  cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
  GSB_REQUIRE(e == cudaSuccess, "peer_open: cudaIpcOpenMemHandle failed: %s", cudaGetErrorString(e));
  *peer_region = p;
  return 0;
}

extern "C" int gsb_peer_close(void* peer_region) {
  if (peer_region == nullptr) return 0;
  cudaError_t e = cudaIpcCloseMemHandle(peer_region);
  GSB_REQUIRE(e == cudaSuccess, "peer_close: %s", cudaGetErrorString(e));
  return 0;
}

// Error word of this rank's OWN region: 0, or the sequence number of an exchange that gave up waiting (synchronises the
// device: call it at a step / epoch boundary, not per layer).
extern "C" int gsb_peer_error(void* region, unsigned long long* seq_out) {
  GSB_REQUIRE(region && seq_out, "peer_error: null pointer");
  PeerHeader h;
  cudaError_t e = cudaMemcpy(&h, region, sizeof(h), cudaMemcpyDeviceToHost);
  GSB_REQUIRE(e == cudaSuccess, "peer_error: %s", cudaGetErrorString(e));
  *seq_out = h.error;
  return 0;
}

extern "C" int gsb_peer_destroy(void* region) {
  if (region == nullptr) return 0;
  cudaError_t e = cudaFree(region);
  GSB_REQUIRE(e == cudaSuccess, "peer_destroy: %s", cudaGetErrorString(e));
  return 0;
}

// In-place sum of buf[0..n) (fp64, device) over the `world` ranks whose regions are listed in rank order
// (regions[rank] = this rank's own region from gsb_peer_create, the others from gsb_peer_open).
extern "C" int gsb_peer_allreduce_f64(double* buf, int n, void* const* regions, int rank, int world, gsb_stream s) {
  GSB_REQUIRE(buf && regions, "peer_allreduce: null pointer");
  GSB_REQUIRE(world >= 1 && world <= kPeerMaxWorld && rank >= 0 && rank < world, "peer_allreduce: world %d rank %d", world, rank);
  GSB_REQUIRE(n >= 0 && n <= kPeerMaxN, "peer_allreduce: n = %d exceeds %d", n, kPeerMaxN);
  if (n == 0) return 0;
  PeerRegions regs;
  memset(&regs, 0, sizeof(regs));
  for (int i = 0; i < world; ++i) {
    GSB_REQUIRE(regions[i] != nullptr, "peer_allreduce: region %d is null", i);
    regs.r[i] = regions[i];
  }
  gsb::launch(peer_allreduce_f64_kernel, gsb::LaunchCfg(1, 1024, 0, to_stream(s)), buf, n, regs, rank, world, peer_timeout_ns());
  return check_launch("peer_allreduce_f64");
}

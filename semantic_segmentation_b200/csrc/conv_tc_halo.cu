// Halo-stationary tcgen05 implicit-GEMM convolution (fprop / dgrad) for sm_100a.
//
// Why a second kernel: the shifted-window kernel of conv_tc.cu re-loads the full 128-pixel activation box for
// every filter tap, i.e. 24 KB of TMA traffic per 128x64x64 MMA block.  ncu (profiles/r01_*) shows the tensor
// pipe 14-29 % busy with the L2->SM operand feed as the limiter.  Here one CTA owns a 16x16 output-pixel tile
// (two M=128 accumulators, 8x16 pixels each) and, per 64-channel block, loads the input window ONCE PER
// HORIZONTAL TAP OFFSET with a vertical halo:  box = 64 ch x 16 px x (16 + ndy - 1) rows.  All vertical taps are
// then plain row shifts of that box: +dy rows = +dy*16*128 B, a multiple of the 1024 B swizzle atom, so the UMMA
// shared-memory descriptor of tap (dx, dy) / M-tile m is simply
//      start = slot + m*1024 + dy*2048,  SBO = 2048 (next pixel row),  SWIZZLE_128B, K-major
// and stays canonical (no base-offset games).  A 3x3 convolution therefore moves 3 boxes of 36 KB per 64-channel
// block for 256 output pixels instead of 18 boxes of 16 KB, and every weight tile (BN x 64) is used by two
// M-tiles: 2.4x less operand traffic per FLOP at BN=128.
//
// Roles (192 threads): warp 0 = TMA producer (activation ring + weight ring, separate mbarrier pairs),
// warp 1 = tcgen05.mma issuer (accumulators double-buffered in TMEM: 2 buffers x 2 M-tiles x BN columns),
// warps 2-5 = epilogue (tcgen05.ld -> bias/act -> bf16 -> swizzled smem -> TMA store, fused BN sum / sum-sq).
//
// Covers every stride-1 window structure whose taps, for a fixed (parity map, dx), have consecutive dy:
// 3x3 s1 (unet/unet_parts.py:16,19 fwd + dgrad), 4x4 s1/s2 and the parity phases of transposed convs
// (models_pix2pix/networks.py:582,647,655; architecture_pix2pix/operations.py:14-39).
#include <stdlib.h>

#include <cstdio>
#include "tc_common.cuh"

namespace gsb {

constexpr int kHaloMaxGroups = 16;
constexpr int kHaloMaxDy = 4;
struct HaloPlan {
  int ngroups;
  int8_t map[kHaloMaxGroups], dx[kHaloMaxGroups], dy0[kHaloMaxGroups], ndy[kHaloMaxGroups];
  uint8_t wtap[kHaloMaxGroups][kHaloMaxDy];
};
struct HaloParams {
  int tiles_x, tiles_y;  // (8*mt)x16 tiles per image
  int mt, tw;            // M-tiles (8x16 pixels each) per CTA tile: 2 (16-wide tile) or 1 (8-wide); tw = 8*mt
  int W, H, N;           // extents of the tiled (output) space
  int cin, cout, n_tiles, num_tiles;
  int nA, nB, a_bytes;   // ring depths and activation box bytes (pw x boxh x 128 B)
  int a_stride;          // activation slot stride (a_bytes rounded up to 1024)
  int tb;                // vertical taps per weight slot (1, or ndy for the short N=64 MMAs: fewer barrier round trips)
  int pw, full;          // box width in pixels; full = 1: ONE box per 64-channel block serves every (dx, dy) tap
  int act;
  const float* bias;
  const float* scale;    // optional per-channel factor applied before the bias (inference: BatchNorm folded into the epilogue)
  float* stats;          // [2*spatial_tiles][2][cout] or null
};

// multicast variants (thread-block cluster of CL CTAs working on CL spatial tiles with the same weights)
__device__ __forceinline__ void tma_load_2d_mc(const void* map, uint64_t* bar, void* dst, int c0, int c1, uint16_t mask) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%3, %4}], [%2], %5;"
      ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "h"(mask)
      : "memory");
}
__device__ __forceinline__ void umma_commit_mc(uint64_t* bar, uint16_t mask) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"(mask) : "memory");
}
// ---- CTA-pair forms (tcgen05 cta_group::2): the two CTAs of a cluster issue ONE M = 256 MMA per step.  CTA r supplies rows
// [128 r, 128 r + 128) of A from its own shared memory and N/2 columns of B, so every CTA reads only half of each weight tile
// from shared memory (the N = 64 / 128 tiles of the one-CTA form are bound by shared-memory operand bandwidth: 192 / 128 bytes
// per clock for 128 available).  The MMA is issued by the leader (cluster rank 0); TMA loads of both CTAs signal the LEADER's
// barriers (peer bit of the shared::cluster address cleared), commits are multicast to both CTAs.
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;
__device__ __forceinline__ void mbar_expect_tx_cluster(uint32_t bar_addr, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cluster.b64 _, [%0], %1;" ::"r"(bar_addr), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t bar_addr) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(bar_addr) : "memory");
}
__device__ __forceinline__ void tma_load_4d_pair(const void* map, uint32_t bar_addr, void* dst, int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(bar_addr), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d_pair(const void* map, uint32_t bar_addr, void* dst, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(bar_addr), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void umma_bf16_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit_pair_mc(uint64_t* bar, uint16_t mask) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"(mask) : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t* holder, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(holder)), "r"(cols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t addr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}

__device__ __forceinline__ void sts_128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void sts_f32(uint32_t addr, float v) {
  asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ float lds_f32(uint32_t addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t lds_32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .b32 rx;\n\t.reg .pred px;\n\t"
      "elect.sync rx|px, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, px;\n\t}"
      : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// BN = output channels per tile (64 / 128 double-buffered accumulators, 256 single-buffered);
// CL = cluster size: the CL CTAs of a cluster own CL different 16x16 tiles with the SAME weight tile, each
// loads 1/CL of every weight slot and multicasts it to all of them (L2 weight reads / CL).
// PAIR (needs CL == 2): the cta_group::2 form described above; B_SLOT = bytes of one weight tap in THIS CTA's shared memory.
template <int BN, int CL, bool PAIR = false>
__global__ void __launch_bounds__(320, 1)
conv_tc_halo_kernel(const __grid_constant__ TcMaps maps, const __grid_constant__ HaloPlan plan, const HaloParams p) {
  GSB_PDL_ENTRY();
  static_assert(!PAIR || CL == 2, "the CTA-pair MMA needs a cluster of two");
  constexpr int B_BYTES = BN * 128;
  constexpr int B_SLOT = PAIR ? B_BYTES / 2 : B_BYTES;
  constexpr int OUT_COLS = BN > 128 ? 128 : BN;
  constexpr int OUT_BYTES = 128 * OUT_COLS * 2;
  constexpr int RED_BYTES = 8 * 2 * OUT_COLS * 4;
  constexpr int NBUF = BN > 128 ? 1 : 2;
  constexpr int TMEM_COLS = NBUF * 2 * BN;
  constexpr uint16_t kMask = (uint16_t)((1u << CL) - 1u);
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = smem;
  uint8_t* sB = sA + p.nA * p.a_stride;
  uint8_t* s_out = sB + p.nB * p.tb * B_SLOT;
  float* s_red = reinterpret_cast<float*>(s_out + OUT_BYTES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_out + OUT_BYTES + RED_BYTES);
  float* s_acc = reinterpret_cast<float*>(bars + 32);  // [2][cout] running BN sums of this CTA (only when p.stats)
  uint64_t* fullA = bars;
  uint64_t* emptyA = bars + 4;
  uint64_t* fullB = bars + 8;
  uint64_t* emptyB = bars + 16;
  uint64_t* tfull = bars + 24;
  uint64_t* tempty = bars + 26;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(bars + 28);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 4; ++i) {
      mbar_init(&fullA[i], PAIR ? 2 : 1);   // PAIR: the leader's barrier collects the loads of both CTAs
      mbar_init(&emptyA[i], 1);
    }
    for (int i = 0; i < 8; ++i) {
      mbar_init(&fullB[i], PAIR ? 2 : 1);
      mbar_init(&emptyB[i], PAIR ? 1 : CL);  // one-CTA form: every CTA of the cluster reads the slot its peers multicast into
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull[i], 1);
      mbar_init(&tempty[i], PAIR ? 512 : 256);   // PAIR: the epilogue threads of both CTAs release the leader's MMA warp
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    prefetch_tmap(&maps.w);
    prefetch_tmap(&maps.out);
    prefetch_tmap(&maps.in[0]);
  }
  __syncthreads();
  if (warp == 1) {
    if (PAIR) tmem_alloc_pair(tmem_holder, TMEM_COLS); else tmem_alloc(tmem_holder, TMEM_COLS);
  }
  tc_fence_before();
  if (CL > 1) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;
  const int kblocks = p.cin >> 6;
  const int per_img = p.tiles_x * p.tiles_y;
  const int spatial = per_img * p.N;
  const int rank = CL > 1 ? (int)(blockIdx.x % CL) : 0;
  const int cluster_id = (int)(blockIdx.x / CL), nclusters = (int)(gridDim.x / CL);
  const int num_ctiles = ((spatial + CL - 1) / CL) * p.n_tiles;  // cluster tiles: CL spatial tiles x one weight tile

  if (warp == 0) {
    if (lane == 0) {  // ---------------- TMA producer ----------------
      int sa = 0, sb = 0;
      uint32_t pa = 0, pb = 0;
      for (int tile = cluster_id; tile < num_ctiles; tile += nclusters) {
        const int nt = tile % p.n_tiles, st = (tile / p.n_tiles) * CL + rank;
        const bool ghost = st >= spatial;  // odd tail: out-of-range image index -> zero-filled loads, no stores
        const int n = ghost ? p.N : st / per_img, r = ghost ? 0 : st % per_img;
        const int x0 = (r % p.tiles_x) * p.tw, y0 = (r / p.tiles_x) * 16;
        for (int kb = 0; kb < kblocks; ++kb) {
          for (int g = 0; g < plan.ngroups; ++g) {
            if (!p.full || g == 0) {
              mbar_wait(&emptyA[sa], pa ^ 1u);
              if (PAIR) {
                const uint32_t fb = smem_u32(&fullA[sa]) & kPeerBitMask;   // the leader's barrier
                mbar_expect_tx_cluster(fb, (uint32_t)p.a_bytes);
                tma_load_4d_pair(&maps.in[plan.map[g]], fb, sA + sa * p.a_stride, kb * 64, x0 + plan.dx[g], y0 + plan.dy0[g], n);
              } else {
                mbar_expect_tx(&fullA[sa], (uint32_t)p.a_bytes);
                tma_load_4d(&maps.in[plan.map[g]], &fullA[sa], sA + sa * p.a_stride, kb * 64, x0 + plan.dx[g], y0 + plan.dy0[g], n);
              }
              if (++sa == p.nA) { sa = 0; pa ^= 1u; }
            }
            const int nd = plan.ndy[g];
            for (int j0 = 0; j0 < nd; j0 += p.tb) {
              const int nt_ = nd - j0 < p.tb ? nd - j0 : p.tb;
              mbar_wait(&emptyB[sb], pb ^ 1u);
              const uint32_t fbB = smem_u32(&fullB[sb]) & kPeerBitMask;
              if (PAIR) mbar_expect_tx_cluster(fbB, (uint32_t)(nt_ * B_SLOT));
              else mbar_expect_tx(&fullB[sb], (uint32_t)(nt_ * B_BYTES));
              for (int t = 0; t < nt_; ++t) {
                const int wrow = (int)plan.wtap[g][j0 + t] * p.cout + nt * BN;
                uint8_t* dst = sB + (sb * p.tb + t) * B_SLOT;
                if (PAIR)      // this CTA keeps ITS half of the weight rows only: the pair MMA reads both halves in place
                  tma_load_2d_pair(&maps.w, fbB, dst, kb * 64, wrow + rank * (BN / 2));
                else if (CL > 1)
                  tma_load_2d_mc(&maps.w, &fullB[sb], dst + rank * (B_BYTES / CL), kb * 64, wrow + rank * (BN / CL), kMask);
                else
                  tma_load_2d(&maps.w, &fullB[sb], dst, kb * 64, wrow);
              }
              if (++sb == p.nB) { sb = 0; pb ^= 1u; }
            }
          }
        }
      }
    }
  } else if (warp == 1 && PAIR && rank != 0) {
    // the peer's MMA warp has nothing to issue: the leader's instructions drive both tensor cores
  } else if (warp == 1) {
    // ---------------- MMA issuer ----------------
    // The WHOLE warp runs the loop (warp-uniform control flow and operands, so descriptors and barrier addresses stay
    // in uniform registers); one elected lane issues.  With `if (lane == 0)` around the loop every tcgen05.mma cost
    // ~20 dependent per-thread instructions (~120 cycles, ncu profiles/r01e) against 32-128 cycles of tensor time.
    constexpr uint32_t idesc = umma_idesc(PAIR ? 256 : 128, BN, 0, 0);
    int sa = 0, sb = 0;
    uint32_t pa = 0, pb = 0, acc = 0, acc_phase = 0;
    const uint32_t row_bytes = (uint32_t)p.pw * 128u;
    const uint64_t desc_hi_a = umma_desc(0, 16, row_bytes), desc_hi_b = umma_desc(0, 16, 1024);
    const uint32_t sA_u = smem_u32(sA), sB_u = smem_u32(sB);
    for (int tile = cluster_id; tile < num_ctiles; tile += nclusters) {
      mbar_wait(&tempty[acc], acc_phase ^ 1u);
      tc_fence_after();
      const uint32_t d0 = tmem_base + acc * (2 * BN);
      uint32_t accum = 0;
      for (int kb = 0; kb < kblocks; ++kb) {
        for (int g = 0; g < plan.ngroups; ++g) {
          if (!p.full || g == 0) {
            mbar_wait(&fullA[sa], pa);
            tc_fence_after();
          }
          // full mode: horizontal taps are pixel (= 128 B row) shifts inside the one box; the swizzle XOR acts on
          // absolute shared-memory address bits, so a row-shifted start stays consistent with what TMA wrote
          const uint32_t a_base = sA_u + sa * p.a_stride + (p.full ? (uint32_t)(plan.dx[g] - plan.dx[0]) * 128u : 0u);
          const int nd = plan.ndy[g];
          for (int j0 = 0; j0 < nd; j0 += p.tb) {
            const int nt_ = nd - j0 < p.tb ? nd - j0 : p.tb;
            mbar_wait(&fullB[sb], pb);
            tc_fence_after();
            const uint64_t bdesc0 = desc_hi_b | (uint64_t)(((sB_u + sb * p.tb * B_SLOT) & 0x3FFFFu) >> 4);
            const uint64_t adesc0 = desc_hi_a | (uint64_t)(((a_base + j0 * row_bytes) & 0x3FFFFu) >> 4);
            if (elect_one()) {
              for (int t = 0; t < nt_; ++t) {
                const uint64_t adesc = adesc0 + (uint64_t)(t * (row_bytes >> 4)), bdesc = bdesc0 + (uint64_t)(t * (B_SLOT >> 4));
#pragma unroll
                for (int m = 0; m < 2; ++m) {
                  if (m < p.mt) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                      if (PAIR)
                        umma_bf16_pair(d0 + m * BN, adesc + (uint64_t)(64 * m + 2 * k), bdesc + (uint64_t)(2 * k), idesc, accum | (uint32_t)(t | k));
                      else
                        umma_bf16(d0 + m * BN, adesc + (uint64_t)(64 * m + 2 * k), bdesc + (uint64_t)(2 * k), idesc, accum | (uint32_t)(t | k));
                    }
                  }
                }
              }
              if (PAIR) umma_commit_pair_mc(&emptyB[sb], kMask);
              else if (CL > 1) umma_commit_mc(&emptyB[sb], kMask);
              else umma_commit(&emptyB[sb]);
            }
            __syncwarp();
            accum = 1;
            if (++sb == p.nB) { sb = 0; pb ^= 1u; }
          }
          if (!p.full || g == plan.ngroups - 1) {
            if (elect_one()) {
              if (PAIR) umma_commit_pair_mc(&emptyA[sa], kMask); else umma_commit(&emptyA[sa]);
            }
            __syncwarp();
            if (++sa == p.nA) { sa = 0; pa ^= 1u; }
          }
        }
      }
      if (elect_one()) {
        if (PAIR) umma_commit_pair_mc(&tfull[acc], kMask); else umma_commit(&tfull[acc]);
      }
      __syncwarp();
      if (++acc == NBUF) { acc = 0; acc_phase ^= 1u; }
    }
  } else {  // ---------------- epilogue warps 2..9 ----------------
    // Eight warps: two per TMEM lane quarter, each taking half of the columns of a staging pass.  With four warps
    // (one per scheduler, no ILP) the epilogue ran at IPC ~0.24 and the MMA warp waited on `tempty` for the
    // small-K layers (ncu profiles/r01e: 10.1 k cycles per tile against 2.3 k cycles of MMA at K = 576).
    const int q = warp & 3;              // TMEM lane quarter this warp may access
    const int half = (warp - 2) >> 2;    // column half of the staging pass
    const int row = q * 32 + lane;       // accumulator row: pixel (row % 8, row / 8) of the 8x16 M-tile
    const int et = threadIdx.x - 64;     // 0..255
    constexpr int CPW = OUT_COLS / 64;   // 32-column chunks per warp and staging pass
    const float slope = p.act == GSB_ACT_RELU ? 0.f : (p.act == GSB_ACT_LRELU ? 0.2f : 1.f);
    const bool is_tanh = p.act == GSB_ACT_TANH;
    const bool simple = p.bias == nullptr && p.scale == nullptr && p.act == GSB_ACT_NONE;
    const uint32_t so = smem_u32(s_out), sr = smem_u32(s_red), sacc = smem_u32(s_acc);
    uint32_t acc = 0, acc_phase = 0;
    if (p.stats != nullptr) {
      for (int j = et; j < 2 * p.cout; j += 256) sts_f32(sacc + 4 * j, 0.f);
      named_bar_sync(2, 256);
    }
    for (int tile = cluster_id; tile < num_ctiles; tile += nclusters) {
      const int nt = tile % p.n_tiles, st = (tile / p.n_tiles) * CL + rank;
      const bool ghost = st >= spatial;
      const int n = ghost ? p.N : st / per_img, r = ghost ? 0 : st % per_img;
      const int x0 = (r % p.tiles_x) * p.tw, y0 = (r / p.tiles_x) * 16;
      const bool plain = simple && !ghost && (x0 + p.tw <= p.W) && (y0 + 16 <= p.H);  // no bias / act / masking
      mbar_wait(&tfull[acc], acc_phase);
      tc_fence_after();
#pragma unroll 1
      const int n_mh = p.mt * (BN / OUT_COLS);
      for (int mh = 0; mh < n_mh; ++mh) {
        const int m = mh / (BN / OUT_COLS), h = mh % (BN / OUT_COLS);  // M-tile, 128-column half
        const bool valid = !ghost && (x0 + m * 8 + (row & 7) < p.W) && (y0 + (row >> 3) < p.H);
        if (et == 0) tma_store_wait_read();  // the previous store has finished reading s_out
        named_bar_sync(1, 256);
#pragma unroll
        for (int cc = 0; cc < CPW; ++cc) {
          const int ch = half * CPW + cc;
          uint32_t v[32];
          tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + acc * (2 * BN) + m * BN + h * OUT_COLS + ch * 32, v);
          const uint32_t slab = so + (ch >> 1) * (128 * 128) + row * 128;
          if (plain) {
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              uint32_t pk[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                __nv_bfloat162 hh = __floats2bfloat162_rn(__uint_as_float(v[g * 8 + 2 * e]), __uint_as_float(v[g * 8 + 2 * e + 1]));
                pk[e] = *reinterpret_cast<uint32_t*>(&hh);
              }
              const int c16 = (ch & 1) * 4 + g;  // 16-byte chunk inside the 128-byte row
              sts_128(slab + ((c16 ^ (row & 7)) << 4), pk[0], pk[1], pk[2], pk[3]);
            }
          } else {
            const int cbase = nt * BN + h * OUT_COLS + ch * 32;
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              uint32_t pk[4];
              // per-channel factor / offset of these 8 columns: four 16-byte loads (warp-uniform addresses) instead of 32
              // scalar ones -- the scalar form made the inference epilogue of the 64-column kernel 27 % slower than the
              // training one (profiles/r02zz_eval_launch_summary.txt)
              float sc8[8], bi8[8];
              if (p.scale != nullptr) {
                const float4 a = __ldg(reinterpret_cast<const float4*>(p.scale + cbase + g * 8));
                const float4 b = __ldg(reinterpret_cast<const float4*>(p.scale + cbase + g * 8 + 4));
                sc8[0] = a.x; sc8[1] = a.y; sc8[2] = a.z; sc8[3] = a.w; sc8[4] = b.x; sc8[5] = b.y; sc8[6] = b.z; sc8[7] = b.w;
              }
              if (p.bias != nullptr) {
                const float4 a = __ldg(reinterpret_cast<const float4*>(p.bias + cbase + g * 8));
                const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + cbase + g * 8 + 4));
                bi8[0] = a.x; bi8[1] = a.y; bi8[2] = a.z; bi8[3] = a.w; bi8[4] = b.x; bi8[5] = b.y; bi8[6] = b.z; bi8[7] = b.w;
              }
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                float f0 = __uint_as_float(v[g * 8 + 2 * e]), f1 = __uint_as_float(v[g * 8 + 2 * e + 1]);
                if (p.scale != nullptr) {
                  f0 *= sc8[2 * e];
                  f1 *= sc8[2 * e + 1];
                }
                if (p.bias != nullptr) {
                  f0 += bi8[2 * e];
                  f1 += bi8[2 * e + 1];
                }
                if (is_tanh) {
                  f0 = tanhf(f0);
                  f1 = tanhf(f1);
                } else {
                  f0 = f0 > 0.f ? f0 : f0 * slope;
                  f1 = f1 > 0.f ? f1 : f1 * slope;
                }
                __nv_bfloat162 hh = __floats2bfloat162_rn(f0, f1);
                pk[e] = valid ? *reinterpret_cast<uint32_t*>(&hh) : 0u;
              }
              const int c16 = (ch & 1) * 4 + g;
              sts_128(slab + ((c16 ^ (row & 7)) << 4), pk[0], pk[1], pk[2], pk[3]);
            }
          }
        }
        if (mh == n_mh - 1) {
          tc_fence_before();
          // accumulators drained: the MMA warp (PAIR: the leader's) may reuse this TMEM buffer
          if (PAIR) mbar_arrive_cluster(smem_u32(&tempty[acc]) & kPeerBitMask); else mbar_arrive(&tempty[acc]);
        }
        fence_proxy_async();
        named_bar_sync(1, 256);
        if (et == 0 && !ghost) {
#pragma unroll
          for (int s = 0; s < OUT_COLS / 64; ++s)
            tma_store_4d(&maps.out, s_out + s * (128 * 128), nt * BN + h * OUT_COLS + s * 64, x0 + m * 8, y0, n);
          tma_store_commit();
        }
        if (p.stats != nullptr && !ghost) {
          // per-channel sum / sum-of-squares of the bf16 values just staged (masked rows hold zeros):
          // thread = (channel pair w, row group rg of 16 rows)
          const int w = et & 31, rg = et >> 5;
#pragma unroll
          for (int s = 0; s < OUT_COLS / 64; ++s) {
            float s0a = 0.f, s0b = 0.f, s1a = 0.f, s1b = 0.f;
            const uint32_t slab = so + s * (128 * 128);
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const int rr = rg * 16 + i;
              const uint32_t u = lds_32(slab + rr * 128 + (((w >> 2) ^ (rr & 7)) << 4) + (w & 3) * 4);
              const float fx = __uint_as_float(u << 16), fy = __uint_as_float(u & 0xffff0000u);
              s0a += fx; s0b += fy;
              s1a = fmaf(fx, fx, s1a); s1b = fmaf(fy, fy, s1b);
            }
            const uint32_t rd = sr + rg * (2 * OUT_COLS) * 4;
            sts_f32(rd + (s * 64 + 2 * w) * 4, s0a);
            sts_f32(rd + (s * 64 + 2 * w + 1) * 4, s0b);
            sts_f32(rd + (OUT_COLS + s * 64 + 2 * w) * 4, s1a);
            sts_f32(rd + (OUT_COLS + s * 64 + 2 * w + 1) * 4, s1b);
          }
          named_bar_sync(2, 256);
          if (et < 2 * OUT_COLS) {  // (which, channel) -> always the same thread: no race on s_acc
            const int j = et;
            float a = 0.f;
#pragma unroll
            for (int g8 = 0; g8 < 8; ++g8) a += lds_f32(sr + (g8 * 2 * OUT_COLS + j) * 4);
            const int which = j / OUT_COLS, c = j % OUT_COLS;
            const uint32_t dst = sacc + (which * p.cout + nt * BN + h * OUT_COLS + c) * 4;
            sts_f32(dst, lds_f32(dst) + a);
          }
          named_bar_sync(2, 256);
        }
      }
      if (++acc == NBUF) { acc = 0; acc_phase ^= 1u; }
    }
    if (p.stats != nullptr)  // one statistics row per CTA (CTAs without tiles write zeros)
      for (int j = et; j < 2 * p.cout; j += 256) p.stats[(int64_t)blockIdx.x * 2 * p.cout + j] = lds_f32(sacc + 4 * j);
    if (et == 0) tma_store_wait_read();
  }
  tc_fence_before();
  if (CL > 1) cluster_sync_all(); else __syncthreads();  // no CTA may exit while a peer still multicasts into it
  if (warp == 1) {
    tc_fence_after();
    if (PAIR) tmem_dealloc_pair(tmem_base, TMEM_COLS); else tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------------------
// Weight gradient, halo-stationary:  dWp[tap][co][ci] = sum_pix dY[pix][co] * X[pix @ tap][ci]
//
// GEMM with K = pixels, so both operands are MN-major in the TMA boxes (64 channels x 128 B rows of pixels).
// One CTA owns: an M-tile = two 64-channel blocks of the SHIFTED tensor (each tied to a (parity map, dx) group),
// an N-tile = BN channels of the TILED tensor, and a range of 8x16-pixel boxes (split-K).  The shifted boxes are
// loaded with a vertical halo (8 x (16+ndy-1) pixels); the ndy vertical taps are K-row shifts of +1024 B and
// accumulate into ndy separate TMEM accumulators, so one pass over the pixels produces all vertical taps:
// 3x less operand traffic than one (tap, block) pair per CTA.  Epilogue: split s stores its fp32 partial sums into slice
// s of the caller's workspace with plain stores (every slice is written completely); the slices are added in slice
// order by the un-pack kernel, so the result does not depend on the order in which CTAs finish (was: red.global.add).
// ------------------------------------------------------------------------------------------------------
struct WgHaloParams {
  int tiles_x, tiles_y;  // 8x16 pixel boxes per image
  int boxes, boxes_per_split;
  int a_blocks, a_cblocks;  // (group, 64-channel) blocks of the shifted side; channel blocks per group
  int n_tiles, ndy;
  int a_is_in;              // 1: shifted side is the forward input (rows = ci), 0: it is dY (rows = co)
  int cin, cout;
  int a_block_bytes, nstages;
  float* dwp;            // slice workspace [split][taps][cout][cin]
  int64_t slice_elems;
};

template <int BN>
__global__ void __launch_bounds__(192, 1)
conv_tc_wgrad_halo_kernel(const __grid_constant__ TcMaps maps, const __grid_constant__ HaloPlan plan, const WgHaloParams p) {
  GSB_PDL_ENTRY();
  constexpr int B_BYTES = (BN / 64) * 16384;
  constexpr int TMEM_COLS = BN == 128 ? 512 : 256;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int stage_bytes = 2 * p.a_block_bytes + B_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + p.nstages * stage_bytes);
  uint64_t* full = bars;
  uint64_t* empty = bars + 8;
  uint64_t* tfull = bars + 16;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(bars + 17);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 8; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    mbar_init(tfull, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    prefetch_tmap(&maps.w);
    prefetch_tmap(&maps.in[0]);
  }
  __syncthreads();
  if (warp == 1) tmem_alloc(tmem_holder, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;

  const int mt = blockIdx.x / p.n_tiles, nt = blockIdx.x % p.n_tiles;
  const int blk0 = mt * 2, blk1 = (mt * 2 + 1 < p.a_blocks) ? mt * 2 + 1 : mt * 2;
  const int box_begin = blockIdx.y * p.boxes_per_split;
  const int box_end = min(box_begin + p.boxes_per_split, p.boxes);
  const int nsteps = box_end - box_begin;
  const int per_img = p.tiles_x * p.tiles_y;

  if (warp == 0) {
    if (lane == 0 && nsteps > 0) {
      const int g0 = blk0 / p.a_cblocks, c0 = (blk0 % p.a_cblocks) * 64;
      const int g1 = blk1 / p.a_cblocks, c1 = (blk1 % p.a_cblocks) * 64;
      const void* m0 = &maps.in[plan.map[g0]];
      const void* m1 = &maps.in[plan.map[g1]];
      const int dx0 = plan.dx[g0], dy0 = plan.dy0[g0], dx1 = plan.dx[g1], dy1 = plan.dy0[g1];
      int stage = 0;
      uint32_t phase = 0;
      for (int b = box_begin; b < box_end; ++b) {
        const int n = b / per_img, r = b % per_img;
        const int x0 = (r % p.tiles_x) * 8, y0 = (r / p.tiles_x) * 16;
        mbar_wait(&empty[stage], phase ^ 1u);
        mbar_expect_tx(&full[stage], (uint32_t)stage_bytes);
        uint8_t* sa = smem + stage * stage_bytes;
        tma_load_4d(m0, &full[stage], sa, c0, x0 + dx0, y0 + dy0, n);
        tma_load_4d(m1, &full[stage], sa + p.a_block_bytes, c1, x0 + dx1, y0 + dy1, n);
#pragma unroll
        for (int j = 0; j < BN / 64; ++j)
          tma_load_4d(&maps.w, &full[stage], sa + 2 * p.a_block_bytes + j * 16384, nt * BN + j * 64, x0, y0, n);
        if (++stage == p.nstages) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp == 1) {
    if (nsteps > 0) {  // whole warp runs the loop (uniform operands), one elected lane issues
      constexpr uint32_t idesc = umma_idesc(128, BN, 1, 1);
      int stage = 0;
      uint32_t phase = 0;
      const uint64_t hi_a = umma_desc(0, (uint32_t)p.a_block_bytes, 1024), hi_b = umma_desc(0, 16384, 1024);
      const uint32_t s0 = smem_u32(smem);
      for (int s = 0; s < nsteps; ++s) {
        mbar_wait(&full[stage], phase);
        tc_fence_after();
        const uint32_t sa = s0 + stage * stage_bytes;
        const uint64_t adesc = hi_a | (uint64_t)((sa & 0x3FFFFu) >> 4);
        const uint64_t bdesc = hi_b | (uint64_t)(((sa + 2 * p.a_block_bytes) & 0x3FFFFu) >> 4);
        if (elect_one()) {
          for (int j = 0; j < p.ndy; ++j) {
#pragma unroll
            for (int k = 0; k < 8; ++k)  // 16 pixels (two 8-pixel rows) per MMA; vertical tap j = +j rows
              umma_bf16(tmem_base + j * BN, adesc + (uint64_t)(64 * j + 128 * k), bdesc + (uint64_t)(128 * k), idesc,
                        (uint32_t)(s | k));
          }
          umma_commit(&empty[stage]);
        }
        __syncwarp();
        if (++stage == p.nstages) { stage = 0; phase ^= 1u; }
      }
      if (elect_one()) umma_commit(tfull);
      __syncwarp();
    }
  } else if (nsteps > 0) {
    const int q = warp & 3, row = q * 32 + lane;
    const int blk = mt * 2 + (row >> 6);
    const bool rvalid = blk < p.a_blocks;
    const int g = rvalid ? blk / p.a_cblocks : 0;
    const int a_ch = rvalid ? (blk % p.a_cblocks) * 64 + (row & 63) : 0;
    mbar_wait(tfull, 0);
    tc_fence_after();
    float* __restrict__ slice = p.dwp + (int64_t)blockIdx.y * p.slice_elems;
#pragma unroll 1
    for (int j = 0; j < p.ndy; ++j) {
      const int tap = (int)plan.wtap[g][j];
#pragma unroll 1
      for (int ch = 0; ch < BN / 32; ++ch) {
        uint32_t v[32];
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + j * BN + ch * 32, v);
        if (rvalid && !p.a_is_in) {   // transposed convolutions: 128 contiguous bytes per thread (see conv_tc_wgrad_kernel)
          float4* dst = reinterpret_cast<float4*>(slice + ((int64_t)tap * p.cout + a_ch) * p.cin + nt * BN + ch * 32);
#pragma unroll
          for (int i = 0; i < 8; ++i)
            dst[i] = make_float4(__uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1]), __uint_as_float(v[4 * i + 2]),
                                 __uint_as_float(v[4 * i + 3]));
        } else if (rvalid) {
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const int b_ch = nt * BN + ch * 32 + i;
            slice[((int64_t)tap * p.cout + b_ch) * p.cin + a_ch] = __uint_as_float(v[i]);
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------------------
// Host side
// ------------------------------------------------------------------------------------------------------
static bool halo_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("GSB_TC_HALO");
    v = (e != nullptr && e[0] == '0') ? 0 : 1;
  }
  return v != 0;
}

// Group the taps by (parity map, dx); inside a group dy must be consecutive.  Returns false when the structure
// does not fit (caller falls back to the shifted-window kernel).
static bool build_halo_plan(const TcTaps& taps, HaloPlan& plan, int& max_ndy, bool& uniform) {
  memset(&plan, 0, sizeof(plan));
  int dys[kHaloMaxGroups][kMaxTaps], wts[kHaloMaxGroups][kMaxTaps], cnt[kHaloMaxGroups];
  for (int t = 0; t < taps.ntaps; ++t) {
    int g = -1;
    for (int i = 0; i < plan.ngroups; ++i)
      if (plan.map[i] == taps.map[t] && plan.dx[i] == taps.dx[t]) g = i;
    if (g < 0) {
      if (plan.ngroups == kHaloMaxGroups) return false;
      g = plan.ngroups++;
      plan.map[g] = taps.map[t];
      plan.dx[g] = taps.dx[t];
      cnt[g] = 0;
    }
    dys[g][cnt[g]] = taps.dy[t];
    wts[g][cnt[g]] = taps.wtap[t];
    cnt[g]++;
  }
  max_ndy = 1;
  uniform = true;
  for (int g = 0; g < plan.ngroups; ++g) {
    if (cnt[g] > kHaloMaxDy) return false;
    for (int a = 0; a < cnt[g]; ++a)  // insertion sort by dy
      for (int b = a + 1; b < cnt[g]; ++b)
        if (dys[g][b] < dys[g][a]) {
          int t0 = dys[g][a]; dys[g][a] = dys[g][b]; dys[g][b] = t0;
          t0 = wts[g][a]; wts[g][a] = wts[g][b]; wts[g][b] = t0;
        }
    for (int a = 1; a < cnt[g]; ++a)
      if (dys[g][a] != dys[g][a - 1] + 1) return false;
    plan.dy0[g] = (int8_t)dys[g][0];
    plan.ndy[g] = (int8_t)cnt[g];
    for (int a = 0; a < cnt[g]; ++a) plan.wtap[g][a] = (uint8_t)wts[g][a];
    if (cnt[g] > max_ndy) max_ndy = cnt[g];
    if (cnt[g] != cnt[0]) uniform = false;
  }
  // order the groups by (map, dx): the full-halo mode wants dx ascending from the box origin
  for (int a = 0; a < plan.ngroups; ++a)
    for (int b = a + 1; b < plan.ngroups; ++b)
      if (plan.map[b] < plan.map[a] || (plan.map[b] == plan.map[a] && plan.dx[b] < plan.dx[a])) {
        int8_t t8;
        t8 = plan.map[a]; plan.map[a] = plan.map[b]; plan.map[b] = t8;
        t8 = plan.dx[a]; plan.dx[a] = plan.dx[b]; plan.dx[b] = t8;
        t8 = plan.dy0[a]; plan.dy0[a] = plan.dy0[b]; plan.dy0[b] = t8;
        t8 = plan.ndy[a]; plan.ndy[a] = plan.ndy[b]; plan.ndy[b] = t8;
        for (int j = 0; j < kHaloMaxDy; ++j) {
          uint8_t u8 = plan.wtap[a][j]; plan.wtap[a][j] = plan.wtap[b][j]; plan.wtap[b][j] = u8;
        }
      }
  return plan.ngroups > 0;
}

static bool env_flag(const char* name, bool dflt) {
  const char* e = getenv(name);
  if (e == nullptr || e[0] == 0) return dflt;
  return e[0] != '0';
}
static bool cluster_enabled() {
  static int v = -1;
  if (v < 0) v = env_flag("GSB_TC_CLUSTER", true) ? 1 : 0;
  return v != 0;
}
static bool fullhalo_enabled() {
  static int v = -1;
  if (v < 0) v = env_flag("GSB_TC_FULLHALO", true) ? 1 : 0;
  return v != 0;
}
// GSB_TC_PAIR: 0 = one-CTA MMAs everywhere, 1 = cta_group::2 (CTA pair) for the 64- and 128-column tiles (default),
// 2 = also for the 256-column tiles
static int pair_mode() {   // read on every call (tests and A/B runs switch it inside one process)
  const char* e = getenv("GSB_TC_PAIR");
  return (e != nullptr && e[0] >= '0' && e[0] <= '2') ? e[0] - '0' : 1;
}
static bool bn256_enabled() {
  static int v = -1;
  if (v < 0) v = env_flag("GSB_TC_BN256", true) ? 1 : 0;
  return v != 0;
}

bool conv_halo_applicable(const SubView& outv) { return halo_enabled() && outv.W >= 16 && outv.H >= 16; }
int conv_halo_m_tiles(int w, int h, int n) {  // upper bound on the statistics rows (one per CTA)
  (void)w; (void)h; (void)n;
  return num_sms();
}

// Returns 1 when the launch was issued, 0 when the tap structure does not fit (caller falls back), -1 on error.
int conv_halo_fprop(const SubView& outv, const SubView* shifted, int nmaps, const TcTaps& taps, const void* wp, int taps_total,
                    int cin, int cout, const float* bias, int act, float* stats, int* stats_rows, cudaStream_t st,
                    const float* scale) {
  if (!conv_halo_applicable(outv)) return 0;
  HaloPlan plan;
  int max_ndy;
  bool uniform;
  if (!build_halo_plan(taps, plan, max_ndy, uniform)) return 0;
  const int boxh = 16 + max_ndy - 1;
  // full-halo mode: one parity map, every group spans the same rows, dx consecutive -> a single box per channel block
  int full = fullhalo_enabled() && uniform && nmaps == 1 && plan.ngroups > 1 ? 1 : 0;
  for (int g = 0; full && g < plan.ngroups; ++g)
    if (plan.dy0[g] != plan.dy0[0] || plan.dx[g] != plan.dx[0] + g) full = 0;
  if (16 + plan.ngroups - 1 > 24) full = 0;
  // Tile shape (BN output channels x mt M-tiles of 8x16 pixels): minimise  tiles-per-CTA x time-per-tile.  Per-SM rates
  // are the measured ones (profiles/README.md: 1500 / 1100 / 870 TFLOP/s over 148 SMs at BN = 256 / 128 / 64); narrow
  // tiles (mt = 1) re-read the weight tile twice as often.  Large launches keep the widest tile; launches with fewer
  // tiles than SMs (deep layers at 8 images per GPU) trade per-tile efficiency for occupancy.
  int BN = 64, mt = 2;
  {
    const double k_flop = 2.0 * 128.0 * (double)cin * (double)taps.ntaps;
    double best = 1e30;
    const int sms = num_sms_avail();
    for (int cand = 0; cand < 6; ++cand) {
      const int bn = cand / 2 == 0 ? 256 : (cand / 2 == 1 ? 128 : 64), m = cand % 2 == 0 ? 2 : 1;
      if (cout % bn != 0) continue;
      if (bn == 256 && !(cin >= 128 && bn256_enabled())) continue;
      if (m == 1 && !env_flag("GSB_TC_NARROW", true)) continue;
      const double rate = (bn == 256 ? 10.1e6 : (bn == 128 ? 7.4e6 : 5.9e6)) * (m == 1 ? 0.9 : 1.0);  // FLOP per us per SM
      const long long sp = (long long)(((int)outv.W + 8 * m - 1) / (8 * m)) * (((int)outv.H + 15) / 16) * (long long)outv.N;
      const long long tiles = ((sp + 1) / 2) * 2 * (cout / bn);
      const long long per_cta = (tiles + sms - 1) / sms;
      const double t = (double)per_cta * (k_flop * m * bn / rate + 1.5) + 4.0;
      if (t < best * 0.97) { best = t; BN = bn; mt = m; }
    }
    // GSB_TC_TILE="<BN>,<mt>" pins the shape (tests sweep every variant; ignored when it does not divide Cout)
    const char* e = getenv("GSB_TC_TILE");
    int fb = 0, fm = 0;
    if (e != nullptr && sscanf(e, "%d,%d", &fb, &fm) == 2 && (fb == 64 || fb == 128 || fb == 256) && (fm == 1 || fm == 2) &&
        cout % fb == 0) {
      BN = fb;
      mt = fm;
    }
  }
  const int tw = 8 * mt;
  const int pw = full ? tw + plan.ngroups - 1 : tw;
  TcMaps maps;
  memset(&maps, 0, sizeof(maps));
  for (int i = 0; i < nmaps; ++i)
    if (make_act_map(&maps.in[i], shifted[i], pw, boxh, 1)) return -1;
  for (int i = nmaps; i < 4; ++i) maps.in[i] = maps.in[0];
  if (make_act_map(&maps.out, outv, 8, 16, 1)) return -1;
  const int spatial = (((int)outv.W + tw - 1) / tw) * (((int)outv.H + 15) / 16) * (int)outv.N;
  const int CL = (cluster_enabled() && spatial >= 2) ? 2 : 1;
  const bool pair = CL == 2 && (pair_mode() == 2 || (pair_mode() == 1 && BN <= 128));
  if (make_w_map(&maps.w, wp, (int64_t)taps_total * cout, cin, BN / CL)) return -1;
  HaloParams p;
  p.mt = mt; p.tw = tw;
  p.tiles_x = ((int)outv.W + tw - 1) / tw;
  p.tiles_y = ((int)outv.H + 15) / 16;
  p.W = (int)outv.W; p.H = (int)outv.H; p.N = (int)outv.N;
  p.cin = cin; p.cout = cout;
  p.n_tiles = cout / BN;
  p.num_tiles = ((spatial + CL - 1) / CL) * p.n_tiles;  // cluster tiles
  p.pw = pw;
  p.full = full;
  p.a_bytes = p.pw * boxh * 128;
  p.a_stride = (p.a_bytes + 1023) & ~1023;
  const int b_bytes = pair ? BN * 64 : BN * 128;   // CTA pair: every CTA holds half of the weight rows
  const int out_cols = BN > 128 ? 128 : BN;
  const int fixed = 1024 + 128 * out_cols * 2 + 8 * 2 * out_cols * 4 + 256 + (stats ? 2 * cout * 4 : 0);
  p.tb = (BN == 64 && uniform) ? max_ndy : 1;
  p.nB = BN == 64 ? (p.tb > 1 ? 3 : 6) : (BN == 128 ? 4 : 3);
  int nA = (232448 - fixed - p.nB * p.tb * b_bytes) / p.a_stride;
  if (nA < 2 && p.tb > 1) {  // tall windows (4 vertical taps): back to one tap per weight slot
    p.tb = 1;
    p.nB = 6;
    nA = (232448 - fixed - p.nB * p.tb * b_bytes) / p.a_stride;
  }
  while (nA < 2 && p.nB > 2) {
    --p.nB;
    nA = (232448 - fixed - p.nB * p.tb * b_bytes) / p.a_stride;
  }
  p.nA = nA > 4 ? 4 : nA;
  GSB_REQUIRE(p.nA >= 2, "conv_halo: shared memory budget");
  p.act = act; p.bias = bias; p.scale = scale; p.stats = stats;
  const int smem = fixed + p.nA * p.a_stride + p.nB * p.tb * b_bytes;
  const int max_clusters = num_sms_avail() / CL;
  const int grid = (p.num_tiles < max_clusters ? p.num_tiles : max_clusters) * CL;
  if (stats_rows) *stats_rows = grid;  // one row of running sums per CTA
  void (*kern)(TcMaps, HaloPlan, HaloParams) = nullptr;
  if (pair) kern = BN == 256 ? conv_tc_halo_kernel<256, 2, true> : (BN == 128 ? conv_tc_halo_kernel<128, 2, true> : conv_tc_halo_kernel<64, 2, true>);
  else if (CL == 2) kern = BN == 256 ? conv_tc_halo_kernel<256, 2> : (BN == 128 ? conv_tc_halo_kernel<128, 2> : conv_tc_halo_kernel<64, 2>);
  else kern = BN == 256 ? conv_tc_halo_kernel<256, 1> : (BN == 128 ? conv_tc_halo_kernel<128, 1> : conv_tc_halo_kernel<64, 1>);
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  GSB_REQUIRE(e == cudaSuccess, "cudaFuncSetAttribute(smem=%d) failed: %s", smem, cudaGetErrorString(e));
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(320);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CL;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;   // PDL, see common.cuh
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  e = cudaLaunchKernelEx(&cfg, kern, maps, plan, p);
  GSB_REQUIRE(e == cudaSuccess, "conv_tc_halo launch failed: %s", cudaGetErrorString(e));
  if (check_launch("conv_tc_halo launch")) return -1;
  return 1;
}

// Weight gradient through the halo kernel.  Returns 1 when launched (or planned), 0 when not applicable, -1 on error.
// ws = slice workspace [max_slices][taps_total * cout * cin]; *used = slices written (each completely).  ws == nullptr:
// plan only (*used = the split the launch would take).
int conv_halo_wgrad(const SubView& tiled, const SubView* shifted, int nmaps, const TcTaps& taps, int shifted_c, int a_is_in,
                    int cin, int cout, int taps_total, float* ws, int max_slices, int* used, cudaStream_t st) {
  if (!halo_enabled() || tiled.W < 8 || tiled.H < 16) return 0;
  HaloPlan plan;
  int max_ndy;
  bool uniform;
  if (!build_halo_plan(taps, plan, max_ndy, uniform) || !uniform) return 0;
  if (taps.ntaps != taps_total) return 0;   // a skipped tap would leave its slice entries unwritten
  const int BN = (tiled.C % 128 == 0) ? 128 : 64;
  if (max_ndy * BN > (BN == 128 ? 512 : 256)) return 0;
  const int boxh = 16 + max_ndy - 1;
  TcMaps maps;
  memset(&maps, 0, sizeof(maps));
  if (ws != nullptr) {
    for (int i = 0; i < nmaps; ++i)
      if (make_act_map(&maps.in[i], shifted[i], 8, boxh, 1)) return -1;
    for (int i = nmaps; i < 4; ++i) maps.in[i] = maps.in[0];
    if (make_act_map(&maps.w, tiled, 8, 16, 1)) return -1;
    maps.out = maps.w;
  }
  WgHaloParams p;
  p.tiles_x = ((int)tiled.W + 7) / 8;
  p.tiles_y = ((int)tiled.H + 15) / 16;
  p.boxes = p.tiles_x * p.tiles_y * (int)tiled.N;
  p.a_cblocks = shifted_c / 64;
  p.a_blocks = plan.ngroups * p.a_cblocks;
  p.n_tiles = (int)tiled.C / BN;
  p.ndy = max_ndy;
  p.a_is_in = a_is_in;
  p.cin = cin; p.cout = cout;
  p.a_block_bytes = 8 * boxh * 128;
  p.dwp = ws;
  p.slice_elems = (int64_t)taps_total * cout * cin;
  const int stage_bytes = 2 * p.a_block_bytes + (BN / 64) * 16384;
  int nst = (232448 - 1024 - 256) / stage_bytes;
  p.nstages = nst > 8 ? 8 : nst;
  GSB_REQUIRE(p.nstages >= 2, "conv_halo_wgrad: shared memory budget");
  const int m_tiles = (p.a_blocks + 1) / 2;
  const int base = m_tiles * p.n_tiles;
  // split-K over the pixel boxes: fill whole waves of SMs, keep >= 2 boxes per CTA
  const int sms = num_sms_avail();
  // GSB_WGRAD_ONE_WAVE: 0 = off, 1 = every layer, 2 = only the 128-column kernel (A/B runs)
  static const int one_wave_mode = []() { const char* e = getenv("GSB_WGRAD_ONE_WAVE"); return (e && e[0]) ? atoi(e) : 1; }();
  const bool one_wave = one_wave_mode == 1 || (one_wave_mode == 2 && BN == 128);
  int max_split = p.boxes / 2 > 0 ? p.boxes / 2 : 1;
  if (ws != nullptr && max_split > max_slices) max_split = max_slices;
  int best = 1;
  double best_eff = 0.0;
  for (int sp = 1; sp <= max_split && sp <= 4 * sms; ++sp) {
    const long long ctas = (long long)base * sp;
    const long long waves = (ctas + sms - 1) / sms;
    const double eff = (double)ctas / (double)(waves * sms);
    if (eff > best_eff + 1e-9) { best_eff = eff; best = sp; }
    // the FIRST split that fills a wave: one wave of CTAs with twice the pixel range costs the same MMA time as two
    // waves, writes half the partial-sum slices (3 * 128 * BN fp32 per CTA) and halves what the fold kernel reads
    if (eff >= 0.95 && (one_wave || ctas >= sms)) { best = sp; break; }
  }
  p.boxes_per_split = (p.boxes + best - 1) / best;
  const int split = (p.boxes + p.boxes_per_split - 1) / p.boxes_per_split;
  GSB_REQUIRE(split <= 65535, "conv_halo_wgrad: split too large");
  *used = split;
  if (ws == nullptr) return 1;
  GSB_REQUIRE(split <= max_slices, "conv_halo_wgrad: %d slices needed, workspace holds %d", split, max_slices);
  const int smem = 1024 + 256 + p.nstages * stage_bytes;
  cudaError_t e;
  if (BN == 128) {
    e = cudaFuncSetAttribute(conv_tc_wgrad_halo_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    GSB_REQUIRE(e == cudaSuccess, "cudaFuncSetAttribute(smem=%d) failed: %s", smem, cudaGetErrorString(e));
    gsb::launch(conv_tc_wgrad_halo_kernel<128>, gsb::LaunchCfg(dim3(base, split), 192, smem, st), maps, plan, p);
  } else {
    e = cudaFuncSetAttribute(conv_tc_wgrad_halo_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    GSB_REQUIRE(e == cudaSuccess, "cudaFuncSetAttribute(smem=%d) failed: %s", smem, cudaGetErrorString(e));
    gsb::launch(conv_tc_wgrad_halo_kernel<64>, gsb::LaunchCfg(dim3(base, split), 192, smem, st), maps, plan, p);
  }
  if (check_launch("conv_tc_wgrad_halo launch")) return -1;
  return 1;
}

}  // namespace gsb

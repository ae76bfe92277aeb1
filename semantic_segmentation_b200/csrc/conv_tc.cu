// tcgen05 / TMEM / TMA implicit-GEMM convolution kernels for sm_100a (bf16 in, fp32 accumulate).
//
// Replaces the cuDNN convolutions behind nn.Conv2d / nn.ConvTranspose2d on the hot layers of
// unet/unet_parts.py:16,19 (3x3 s1 p1), models_pix2pix/networks.py:582,647,655 (4x4 s2/s1 p1) and
// architecture_pix2pix/operations.py:14-39 (merged 8x8 s2 p3 transposed conv), forward, data-gradient
// and weight-gradient.
//
// Formulation ("shifted-window implicit GEMM"):
//   out[pix][co] = sum_tap sum_ci  A_tap[pix][ci] * Wp[tap][co][ci]
// The M tile is a (tw x th x tn) box of 128 output pixels.  For every filter tap the producer issues ONE
// 4-D TMA load of the same box shifted by the tap offset (dx, dy) from the NHWC activation: out-of-bounds
// rows/columns are zero-filled by the TMA unit, which is exactly the convolution padding.  Stride-2
// convolutions use the four parity sub-views of the input (doubled strides), stride-2 transposed
// convolutions the four parity sub-views of the output, so every tap is again a plain shifted box.
// The box lands in shared memory as 128 rows x 64 bf16 (128 B, SWIZZLE_128B) = the canonical K-major UMMA
// operand; weights are pre-packed [tap][cout][cin] (K-major) and fetched with a 2-D TMA.  A single thread
// issues tcgen05.mma (M=128, N=BN, K=16) into a double-buffered TMEM accumulator; four epilogue warps read
// TMEM (tcgen05.ld 32x32b), add bias / activation, round to bf16, stage the tile in swizzled shared memory,
// TMA-store it, and reduce the per-channel BatchNorm sum / sum-of-squares of the stored values.
//
// Weight gradient: dWp[tap][co][ci] = sum_pix dY[pix][co] * X[pix@tap][ci] is a GEMM whose K dimension is
// the pixel index, i.e. both operands are MN-major in the very same TMA boxes (64 pixels x 64 channels).
// Two (tap, 64-channel) blocks of the shifted operand form the M=128 rows, so Cin=64 layers still fill the
// tensor core; split-K over pixel boxes, fp32 red.global.add into the packed gradient.
#include <stdlib.h>

#include "tc_common.cuh"

namespace gsb {

// ------------------------------------------------------------------------------------------------------
// Kernel parameters
// ------------------------------------------------------------------------------------------------------
struct TcParams {
  int tw, th, tn;                 // pixel box (tw*th*tn = 128 fprop / 64 wgrad)
  int tiles_x, tiles_y, tiles_z;  // boxes covering (W, H, N) of the tiled space
  int W, H, N;                    // extents of the tiled space (validity masks)
  int cin, cout;
  int n_tiles, num_tiles;
  int act;
  const float* bias;
  float* stats;                   // [m_tiles][2][cout] or null
  int ks_split;                   // > 1: split-K -- a tile's K steps are spread over ks_split CTAs, part i stores into slice i
  float* ws;                      // [ks_split][N*H*W of the tiled space][cout] fp32 partial sums (split-K only)
  int64_t ws_slice;               // elements per slice
};

template <int BN>
struct FpropCfg {
  static constexpr int A_BYTES = 128 * 128;
  static constexpr int B_BYTES = BN * 128;
  static constexpr int STAGE = A_BYTES + B_BYTES;
  static constexpr int STAGES = BN == 64 ? 6 : (BN == 128 ? 5 : 3);
  static constexpr int OUT_BYTES = 128 * BN * 2;
  static constexpr int RED_BYTES = 4 * 2 * BN * 4;
  static constexpr int SMEM = STAGES * STAGE + OUT_BYTES + RED_BYTES + 256 + 1024;  // + barriers + alignment slack
  static constexpr int TMEM_COLS = 2 * BN;  // 128 / 256 / 512
};

template <int BN>
__global__ void __launch_bounds__(192, 1)
conv_tc_fprop_kernel(const __grid_constant__ TcMaps maps, const __grid_constant__ TcTaps taps, const TcParams p) {
  GSB_PDL_ENTRY();
  using Cfg = FpropCfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* s_out = smem + Cfg::STAGES * Cfg::STAGE;
  float* s_red = reinterpret_cast<float*>(s_out + Cfg::OUT_BYTES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_out + Cfg::OUT_BYTES + Cfg::RED_BYTES);
  uint64_t* full = bars;
  uint64_t* empty = bars + Cfg::STAGES;
  uint64_t* tfull = bars + 2 * Cfg::STAGES;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int i = 0; i < Cfg::STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull[i], 1);
      mbar_init(&tempty[i], 128);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    prefetch_tmap(&maps.w);
    prefetch_tmap(&maps.out);
    prefetch_tmap(&maps.in[0]);
  }
  __syncthreads();
  if (warp == 1) tmem_alloc(tmem_holder, Cfg::TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;
  const int kblocks = p.cin >> 6;
  const int ksteps = taps.ntaps * kblocks;

  if (warp == 0) {
    if (lane == 0) {  // ---------------- TMA producer ----------------
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
        const int part = tile % p.ks_split, tl = tile / p.ks_split;
        const int ks0 = (int)((long long)part * ksteps / p.ks_split), ks1 = (int)((long long)(part + 1) * ksteps / p.ks_split);
        const int nt = tl % p.n_tiles, mt = tl / p.n_tiles;
        const int tx = mt % p.tiles_x, r = mt / p.tiles_x;
        const int x0 = tx * p.tw, y0 = (r % p.tiles_y) * p.th, n0 = (r / p.tiles_y) * p.tn;
        for (int ks = ks0; ks < ks1; ++ks) {
          const int t = ks / kblocks, kb = ks - t * kblocks;
          const void* amap = &maps.in[taps.map[t]];
          const int xx = x0 + taps.dx[t], yy = y0 + taps.dy[t];
          const int wrow = (int)taps.wtap[t] * p.cout + nt * BN;
          mbar_wait(&empty[stage], phase ^ 1u);
          mbar_expect_tx(&full[stage], Cfg::STAGE);
          uint8_t* sa = smem + stage * Cfg::STAGE;
          tma_load_4d(amap, &full[stage], sa, kb * 64, xx, yy, n0);
          tma_load_2d(&maps.w, &full[stage], sa + Cfg::A_BYTES, kb * 64, wrow);
          if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {  // ---------------- MMA issuer ----------------
      constexpr uint32_t idesc = umma_idesc(128, BN, 0, 0);
      int stage = 0;
      uint32_t phase = 0, acc = 0, acc_phase = 0;
      for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
        mbar_wait(&tempty[acc], acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        const int part = tile % p.ks_split;
        const int nks = (int)((long long)(part + 1) * ksteps / p.ks_split) - (int)((long long)part * ksteps / p.ks_split);
        for (int ks = 0; ks < nks; ++ks) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * Cfg::STAGE);
          const uint64_t adesc = umma_desc(sa, 16, 1024);
          const uint64_t bdesc = umma_desc(sa + Cfg::A_BYTES, 16, 1024);
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (ks | k) != 0 ? 1u : 0u);
          umma_commit(&empty[stage]);
          if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1u; }
        }
        umma_commit(&tfull[acc]);
        acc ^= 1u;
        if (acc == 0) acc_phase ^= 1u;
      }
    }
  } else {  // ---------------- epilogue warps 2..5 ----------------
    const int q = warp & 3;            // TMEM lane quarter this warp may access
    const int row = q * 32 + lane;     // accumulator row == pixel index inside the box
    const int et = threadIdx.x - 64;   // 0..127
    uint32_t acc = 0, acc_phase = 0;
    for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
      const int tl = tile / p.ks_split;
      const int nt = tl % p.n_tiles, mt = tl / p.n_tiles;
      const int tx = mt % p.tiles_x, r = mt / p.tiles_x;
      const int x0 = tx * p.tw, y0 = (r % p.tiles_y) * p.th, n0 = (r / p.tiles_y) * p.tn;
      const int px = row % p.tw, py = (row / p.tw) % p.th, pn = row / (p.tw * p.th);
      const bool valid = (x0 + px < p.W) && (y0 + py < p.H) && (n0 + pn < p.N);
      mbar_wait(&tfull[acc], acc_phase);
      tc_fence_after();
      if (p.ks_split > 1) {
        // split-K: this CTA holds a partial sum over its K range -> plain stores into its own slice of the workspace; the
        // slices are added in slice order, then bias, activation, rounding and the store to the output view, in
        // splitk_finalize_kernel (deterministic: no atomics)
        float* wrow = p.ws + (int64_t)(tile % p.ks_split) * p.ws_slice +
                      (((int64_t)(n0 + pn) * p.H + (y0 + py)) * p.W + (x0 + px)) * p.cout + nt * BN;
#pragma unroll 1
        for (int ch = 0; ch < BN / 32; ++ch) {
          uint32_t v[32];
          tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + acc * BN + ch * 32, v);
          if (valid) {
#pragma unroll
            for (int i = 0; i < 8; ++i)
              *reinterpret_cast<float4*>(wrow + ch * 32 + 4 * i) =
                  make_float4(__uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1]), __uint_as_float(v[4 * i + 2]),
                              __uint_as_float(v[4 * i + 3]));
          }
        }
        tc_fence_before();
        mbar_arrive(&tempty[acc]);
        acc ^= 1u;
        if (acc == 0) acc_phase ^= 1u;
        continue;
      }
      if (et == 0) tma_store_wait_read();  // previous tile's store has finished reading s_out
      named_bar_sync(1, 128);
#pragma unroll 1
      for (int ch = 0; ch < BN / 32; ++ch) {
        uint32_t v[32];
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + acc * BN + ch * 32, v);
        const int cbase = nt * BN + ch * 32;
        uint8_t* slab = s_out + (ch >> 1) * (128 * 128) + row * 128;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint32_t pk[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            float f0 = __uint_as_float(v[g * 8 + 2 * e]), f1 = __uint_as_float(v[g * 8 + 2 * e + 1]);
            if (p.bias != nullptr) {
              f0 += __ldg(p.bias + cbase + g * 8 + 2 * e);
              f1 += __ldg(p.bias + cbase + g * 8 + 2 * e + 1);
            }
            f0 = valid ? act_fwd(f0, p.act) : 0.f;
            f1 = valid ? act_fwd(f1, p.act) : 0.f;
            __nv_bfloat162 h = __floats2bfloat162_rn(f0, f1);
            pk[e] = *reinterpret_cast<uint32_t*>(&h);
          }
          const int c16 = (ch & 1) * 4 + g;  // 16-byte chunk inside the 128-byte row
          *reinterpret_cast<uint4*>(slab + ((c16 ^ (row & 7)) << 4)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
        }
      }
      tc_fence_before();
      mbar_arrive(&tempty[acc]);  // accumulator drained: the MMA warp may start the tile after next
      fence_proxy_async();
      named_bar_sync(1, 128);
      if (et == 0) {
#pragma unroll
        for (int s = 0; s < BN / 64; ++s) tma_store_4d(&maps.out, s_out + s * (128 * 128), nt * BN + s * 64, x0, y0, n0);
        tma_store_commit();
      }
      if (p.stats != nullptr) {
        // per-channel sum / sum-of-squares of the bf16 values just staged (invalid rows hold zeros)
        const int w = et & 31, rg = et >> 5;
#pragma unroll
        for (int s = 0; s < BN / 64; ++s) {
          float s0a = 0.f, s0b = 0.f, s1a = 0.f, s1b = 0.f;
          const uint8_t* slab = s_out + s * (128 * 128);
#pragma unroll 8
          for (int i = 0; i < 32; ++i) {
            const int rr = rg * 32 + i;
            uint32_t u = *reinterpret_cast<const uint32_t*>(slab + rr * 128 + (((w >> 2) ^ (rr & 7)) << 4) + (w & 3) * 4);
            float2 f = __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&u));
            s0a += f.x; s0b += f.y;
            s1a = fmaf(f.x, f.x, s1a); s1b = fmaf(f.y, f.y, s1b);
          }
          float* rd = s_red + rg * (2 * BN);
          rd[s * 64 + 2 * w] = s0a;
          rd[s * 64 + 2 * w + 1] = s0b;
          rd[BN + s * 64 + 2 * w] = s1a;
          rd[BN + s * 64 + 2 * w + 1] = s1b;
        }
        named_bar_sync(2, 128);
        for (int j = et; j < 2 * BN; j += 128) {
          float a = s_red[j] + s_red[2 * BN + j] + s_red[4 * BN + j] + s_red[6 * BN + j];
          const int which = j / BN, c = j % BN;
          p.stats[((int64_t)mt * 2 + which) * p.cout + nt * BN + c] = a;
        }
        named_bar_sync(2, 128);
      }
      acc ^= 1u;
      if (acc == 0) acc_phase ^= 1u;
    }
    if (et == 0) tma_store_wait_read();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}

// Split-K epilogue: sum over the slices ws[slice][pixel][cout] (fp32, slice order) -> (+bias) -> activation -> bf16 ->
// output view (8 channels per thread).
__global__ void __launch_bounds__(256) splitk_finalize_kernel(const float* __restrict__ ws, int nslices, int64_t slice_elems,
                                                             int64_t pixels, int W, int H, int cout,
                                                             const float* __restrict__ bias, int act, __nv_bfloat16* __restrict__ out,
                                                             int64_t sw, int64_t sh, int64_t sn) {
  GSB_PDL_ENTRY();
  const int G = cout >> 3;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < pixels * G; i += (int64_t)gridDim.x * blockDim.x) {
    const int g = (int)(i % G);
    const int64_t pix = i / G;
    const int x = (int)(pix % W), y = (int)((pix / W) % H);
    const int64_t n = pix / ((int64_t)W * H);
    float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int s = 0; s < nslices; ++s) {
      const float* src = ws + (int64_t)s * slice_elems + pix * cout + g * 8;
      const float4 a = *reinterpret_cast<const float4*>(src), b = *reinterpret_cast<const float4*>(src + 4);
      v[0] += a.x; v[1] += a.y; v[2] += a.z; v[3] += a.w;
      v[4] += b.x; v[5] += b.y; v[6] += b.z; v[7] += b.w;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = act_fwd(v[j] + (bias ? __ldg(bias + g * 8 + j) : 0.f), act);
    Vec8<__nv_bfloat16>::store(out + n * sn + y * sh + x * sw + g * 8, v);
  }
}

// ------------------------------------------------------------------------------------------------------
// Weight gradient
// ------------------------------------------------------------------------------------------------------
struct WgParams {
  int tw, th, tn;                 // 64-pixel box
  int tiles_x, tiles_y, tiles_z;
  int boxes, boxes_per_split;
  int a_blocks, a_cblocks;        // shifted-side (tap, 64-channel) blocks; channel blocks per tap
  int n_tiles;                    // unshifted-side channel tiles of BN
  int a_is_in;                    // 1: shifted side is the forward input (rows = ci), 0: it is dY (rows = co)
  int cin, cout;
  float* dwp;                     // slice workspace [split][taps][cout][cin]: split s stores into slice s (no atomics)
  int64_t slice_elems;
};

template <int BN>
struct WgradCfg {
  static constexpr int A_BYTES = 2 * 64 * 128;        // two 64-pixel x 64-channel boxes
  static constexpr int B_BYTES = (BN / 64) * 64 * 128;
  static constexpr int STAGE = A_BYTES + B_BYTES;
  static constexpr int STAGES = BN == 64 ? 8 : (BN == 128 ? 6 : 4);
  static constexpr int SMEM = STAGES * STAGE + 256 + 1024;
  static constexpr int TMEM_COLS = BN < 32 ? 32 : BN;
};

template <int BN>
__global__ void __launch_bounds__(192, 1)
conv_tc_wgrad_kernel(const __grid_constant__ TcMaps maps, const __grid_constant__ TcTaps taps, const WgParams p) {
  GSB_PDL_ENTRY();
  using Cfg = WgradCfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::STAGES * Cfg::STAGE);
  uint64_t* full = bars;
  uint64_t* empty = bars + Cfg::STAGES;
  uint64_t* tfull = bars + 2 * Cfg::STAGES;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(tfull + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int i = 0; i < Cfg::STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    mbar_init(tfull, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    prefetch_tmap(&maps.w);
    prefetch_tmap(&maps.in[0]);
  }
  __syncthreads();
  if (warp == 1) tmem_alloc(tmem_holder, Cfg::TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;

  const int mt = blockIdx.x / p.n_tiles, nt = blockIdx.x % p.n_tiles;
  const int blk0 = mt * 2, blk1 = (mt * 2 + 1 < p.a_blocks) ? mt * 2 + 1 : mt * 2;
  const int box_begin = blockIdx.y * p.boxes_per_split;
  const int box_end = min(box_begin + p.boxes_per_split, p.boxes);
  const int nsteps = box_end - box_begin;

  if (warp == 0) {
    if (lane == 0 && nsteps > 0) {
      const int t0 = blk0 / p.a_cblocks, c0 = (blk0 % p.a_cblocks) * 64;
      const int t1 = blk1 / p.a_cblocks, c1 = (blk1 % p.a_cblocks) * 64;
      const void* m0 = &maps.in[taps.map[t0]];
      const void* m1 = &maps.in[taps.map[t1]];
      int stage = 0;
      uint32_t phase = 0;
      for (int b = box_begin; b < box_end; ++b) {
        const int tx = b % p.tiles_x, r = b / p.tiles_x;
        const int x0 = tx * p.tw, y0 = (r % p.tiles_y) * p.th, n0 = (r / p.tiles_y) * p.tn;
        mbar_wait(&empty[stage], phase ^ 1u);
        mbar_expect_tx(&full[stage], Cfg::STAGE);
        uint8_t* sa = smem + stage * Cfg::STAGE;
        tma_load_4d(m0, &full[stage], sa, c0, x0 + taps.dx[t0], y0 + taps.dy[t0], n0);
        tma_load_4d(m1, &full[stage], sa + 8192, c1, x0 + taps.dx[t1], y0 + taps.dy[t1], n0);
#pragma unroll
        for (int j = 0; j < BN / 64; ++j)
          tma_load_4d(&maps.w, &full[stage], sa + Cfg::A_BYTES + j * 8192, nt * BN + j * 64, x0, y0, n0);
        if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && nsteps > 0) {
      constexpr uint32_t idesc = umma_idesc(128, BN, 1, 1);
      int stage = 0;
      uint32_t phase = 0;
      for (int s = 0; s < nsteps; ++s) {
        mbar_wait(&full[stage], phase);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + stage * Cfg::STAGE);
        const uint64_t adesc = umma_desc(sa, 8192, 1024);
        const uint64_t bdesc = umma_desc(sa + Cfg::A_BYTES, 8192, 1024);
#pragma unroll
        for (int k = 0; k < 4; ++k)  // 16 pixels (K) per MMA = 2048 B further down the box
          umma_bf16(tmem_base, adesc + (uint64_t)(128 * k), bdesc + (uint64_t)(128 * k), idesc, (s | k) != 0 ? 1u : 0u);
        umma_commit(&empty[stage]);
        if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1u; }
      }
      umma_commit(tfull);
    }
  } else if (nsteps > 0) {
    const int q = warp & 3, row = q * 32 + lane;
    const int blk = mt * 2 + (row >> 6);
    const bool rvalid = blk < p.a_blocks;
    const int tap = rvalid ? (int)taps.wtap[blk / p.a_cblocks] : 0;
    const int a_ch = rvalid ? (blk % p.a_cblocks) * 64 + (row & 63) : 0;
    mbar_wait(tfull, 0);
    tc_fence_after();
    float* __restrict__ slice = p.dwp + (int64_t)blockIdx.y * p.slice_elems;
#pragma unroll 1
    for (int ch = 0; ch < BN / 32; ++ch) {
      uint32_t v[32];
      tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + ch * 32, v);
      if (rvalid && !p.a_is_in) {
        // transposed convolutions: the thread's 32 accumulator columns are 128 contiguous bytes of its own row
        // [tap][a_ch][nt * BN + 32 ch ..]: eight 16-byte stores (scalar stores wrote 4 bytes of every 32-byte sector per
        // instruction: the deep generator levels ran at 0.7 TB/s of slice writes, profiles/r02p_p2p_launch_summary.txt)
        float4* dst = reinterpret_cast<float4*>(slice + ((int64_t)tap * p.cout + a_ch) * p.cin + nt * BN + ch * 32);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          dst[j] = make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]), __uint_as_float(v[4 * j + 2]),
                               __uint_as_float(v[4 * j + 3]));
      } else if (rvalid) {
#pragma unroll
        for (int j = 0; j < 32; ++j) {   // lanes = consecutive a_ch: coalesced 128-byte rows per column
          const int b_ch = nt * BN + ch * 32 + j;
          slice[((int64_t)tap * p.cout + b_ch) * p.cin + a_ch] = __uint_as_float(v[j]);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------------------
// Host side: tensor maps, tap tables, launch
// ------------------------------------------------------------------------------------------------------
bool conv_halo_applicable(const SubView& outv);
int conv_halo_m_tiles(int w, int h, int n);
int conv_halo_fprop(const SubView& outv, const SubView* shifted, int nmaps, const TcTaps& taps, const void* wp, int taps_total,
                    int cin, int cout, const float* bias, int act, float* stats, int* stats_rows, cudaStream_t st,
                    const float* scale = nullptr);
int conv_halo_wgrad(const SubView& tiled, const SubView* shifted, int nmaps, const TcTaps& taps, int shifted_c, int a_is_in,
                    int cin, int cout, int taps_total, float* ws, int max_slices, int* used, cudaStream_t st);

static int pow2ceil(int v) {
  int p = 1;
  while (p < v) p <<= 1;
  return p;
}
static void pick_box(int W, int H, int pixels, int& tw, int& th, int& tn) {
  tw = pow2ceil(W);
  if (tw > 16) tw = 16;
  if (tw > pixels) tw = pixels;
  th = pow2ceil(H);
  if (th > pixels / tw) th = pixels / tw;
  tn = pixels / (tw * th);
}

int conv_tc_m_tiles(const gsb_view* out) {  // upper bound over both fprop kernels (the launch reports the rows it wrote)
  int tw, th, tn;
  pick_box(out->w, out->h, 128, tw, th, tn);
  int a = ((out->w + tw - 1) / tw) * ((out->h + th - 1) / th) * ((out->n + tn - 1) / tn);
  int b = conv_halo_m_tiles(out->w, out->h, out->n);
  return a > b ? a : b;
}

static bool tc_view_ok(const gsb_view* v) {
  return v->dtype == GSB_BF16 && v->sc == 1 && v->c % 64 == 0 && v->sw % 8 == 0 && v->sh % 8 == 0 && v->sn % 8 == 0 &&
         ((uintptr_t)v->ptr % 16 == 0);
}

bool conv_tc_supported(const gsb_view* in, const gsb_view* out, const gsb_conv_geom* g) {
  if (!tc_view_ok(in) || !tc_view_ok(out)) return false;
  if (g->stride != 1 && g->stride != 2) return false;
  if (g->kh * g->kw > kMaxTaps) return false;
  if (g->kh > 16 || g->kw > 16) return false;
  return true;
}

// Tap tables.  Tiled space = `tiled` view; shifted operand = `shifted` view.
//   rel = +1: shifted coordinate = tiled*stride - pad + k   (conv forward / convT wgrad)
//   rel = -1: shifted coordinate = (tiled + pad - k)/stride (convT-gather, sub-sampled per output parity)
struct TapPlan {
  TcTaps taps;
  SubView shifted[4];
  int nmaps;
};

template <typename Kern, typename P>
static int launch_tc(Kern kern, int grid_x, int grid_y, int smem, const TcMaps& maps, const TcTaps& taps, const P& prm,
                     cudaStream_t st) {
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  GSB_REQUIRE(e == cudaSuccess, "cudaFuncSetAttribute(smem=%d) failed: %s", smem, cudaGetErrorString(e));
  gsb::launch(kern, gsb::LaunchCfg(dim3(grid_x, grid_y), 192, smem, st), maps, taps, prm);
  return check_launch("conv_tc launch");
}


// One fprop-style launch over a tiled output (sub)view.
static int fprop_launch(const SubView& outv, const SubView* shifted, int nmaps, const TcTaps& taps, const void* wp, int taps_total,
                        int cin, int cout, const float* bias, int act, float* stats, int* stats_rows, float* scratch,
                        size_t scratch_bytes, cudaStream_t st, const float* scale = nullptr) {
  {  // halo-stationary kernel first (16x16 tiles, operand reuse across vertical taps); falls through when not applicable
    int r = conv_halo_fprop(outv, shifted, nmaps, taps, wp, taps_total, cin, cout, bias, act, stats, stats_rows, st, scale);
    if (r != 0) return r < 0 ? -1 : 0;
  }
  if (scale != nullptr) return 1;   // the per-channel factor exists in the halo kernel's epilogue only: not applicable
  TcMaps maps;
  memset(&maps, 0, sizeof(maps));
  int tw, th, tn;
  pick_box((int)outv.W, (int)outv.H, 128, tw, th, tn);
  for (int i = 0; i < nmaps; ++i)
    if (make_act_map(&maps.in[i], shifted[i], tw, th, tn)) return -1;
  for (int i = nmaps; i < 4; ++i) maps.in[i] = maps.in[0];
  if (make_act_map(&maps.out, outv, tw, th, tn)) return -1;
  const int BN = (cout % 128 == 0) ? 128 : 64;
  if (make_w_map(&maps.w, wp, (int64_t)taps_total * cout, cin, BN)) return -1;
  TcParams p;
  p.tw = tw; p.th = th; p.tn = tn;
  p.tiles_x = ((int)outv.W + tw - 1) / tw;
  p.tiles_y = ((int)outv.H + th - 1) / th;
  p.tiles_z = ((int)outv.N + tn - 1) / tn;
  p.W = (int)outv.W; p.H = (int)outv.H; p.N = (int)outv.N;
  p.cin = cin; p.cout = cout;
  p.n_tiles = cout / BN;
  p.num_tiles = p.tiles_x * p.tiles_y * p.tiles_z * p.n_tiles;
  p.act = act; p.bias = bias; p.stats = stats;
  p.ks_split = 1; p.ws = nullptr; p.ws_slice = 0;
  if (stats_rows) *stats_rows = p.tiles_x * p.tiles_y * p.tiles_z;
  // Deep generator / discriminator levels: a handful of tiles would stream the whole weight tensor through 4..32 SMs
  // (2.3 ms per Pix2Pix step, ncu profiles/r01o).  Split the K loop over enough CTAs to fill the GPU instead.
  const int ksteps = taps.ntaps * (cin / 64);
  const int64_t pixels = outv.N * outv.H * outv.W;
  const int sms = num_sms_avail();
  static const bool splitk_on = []() { const char* e = getenv("GSB_TC_SPLITK"); return !(e && e[0] == '0'); }();
  const size_t slice_bytes = (size_t)pixels * cout * sizeof(float);
  if (splitk_on && scratch != nullptr && p.num_tiles * 2 <= sms && ksteps >= 8 && 2 * slice_bytes <= scratch_bytes &&
      cout % 8 == 0 && ((uintptr_t)scratch % 16) == 0) {
    // the caller's scratch holds one fp32 partial-sum slice per K part (every valid pixel row of a slice is written by
    // exactly one CTA, so no clearing and no atomics: the finalize kernel adds the slices in slice order)
    int split = sms / p.num_tiles;
    if (split > ksteps / 4) split = ksteps / 4;
    if ((size_t)split * slice_bytes > scratch_bytes) split = (int)(scratch_bytes / slice_bytes);
    if (split > 1) {
      p.ks_split = split;
      p.ws = scratch;
      p.ws_slice = pixels * cout;
      p.num_tiles *= split;
      p.stats = nullptr;
      if (stats_rows) *stats_rows = 0;   // no fused statistics on this path: the caller reduces the stored output
    }
  }
  int grid = p.num_tiles < sms ? p.num_tiles : sms;
  int rc = BN == 128 ? launch_tc(conv_tc_fprop_kernel<128>, grid, 1, FpropCfg<128>::SMEM, maps, taps, p, st)
                     : launch_tc(conv_tc_fprop_kernel<64>, grid, 1, FpropCfg<64>::SMEM, maps, taps, p, st);
  if (rc != 0 || p.ks_split == 1) return rc;
  int64_t items = pixels * (cout / 8);
  int fgrid = (int)((items + 255) / 256);
  if (fgrid > num_sms() * 8) fgrid = num_sms() * 8;
  gsb::launch(splitk_finalize_kernel, gsb::LaunchCfg(fgrid, 256, 0, st), p.ws, p.ks_split, p.ws_slice, pixels, (int)outv.W, (int)outv.H, cout, bias, act,
                                                (__nv_bfloat16*)const_cast<void*>(outv.ptr), outv.sw, outv.sh, outv.sn);
  return check_launch("conv_tc split-K finalize");
}

static int floor_div(int a, int b) { return (a >= 0) ? a / b : -((-a + b - 1) / b); }

int conv_gather_tc(const gsb_view* in, const void* wp, const float* bias, const gsb_view* out, const gsb_conv_geom* g, int act,
                   float* stats, int* stats_rows, float* scratch, size_t scratch_bytes, cudaStream_t st, const float* scale) {
  const int s = g->stride, taps_total = g->kh * g->kw;
  if (scale != nullptr && g->mode != GSB_MODE_CONV) return 1;   // (returns 1 = "scaled epilogue not applicable")
  GSB_REQUIRE(stats == nullptr || !(g->mode == GSB_MODE_CONVT && s > 1), "conv_tc: fused BN statistics unsupported for strided transposed conv");
  if (g->mode == GSB_MODE_CONV) {
    // iy = oy*s - pad + ky = s*(oy + a) + par
    TcTaps taps;
    memset(&taps, 0, sizeof(taps));
    SubView subs[4];
    int map_of[4] = {-1, -1, -1, -1}, nmaps = 0;
    for (int ky = 0; ky < g->kh; ++ky)
      for (int kx = 0; kx < g->kw; ++kx) {
        int dyv = ky - g->pad, dxv = kx - g->pad;
        int ay = floor_div(dyv, s), ax = floor_div(dxv, s);
        int py = dyv - ay * s, px = dxv - ax * s;
        if (py >= in->h || px >= in->w) continue;  // empty parity sub-view: tap never hits
        int key = py * s + px;
        if (map_of[key] < 0) {
          map_of[key] = nmaps;
          subs[nmaps++] = subview(in, s, py, px);
        }
        int t = taps.ntaps++;
        taps.dy[t] = (int8_t)ay; taps.dx[t] = (int8_t)ax; taps.map[t] = (int8_t)map_of[key];
        taps.wtap[t] = (uint8_t)(ky * g->kw + kx);
      }
    GSB_REQUIRE(taps.ntaps > 0, "conv_tc: no taps");
    SubView ov = subview(out, 1, 0, 0);
    return fprop_launch(ov, subs, nmaps, taps, wp, taps_total, in->c, out->c, bias, act, stats, stats_rows, scratch, scratch_bytes, st,
                        scale);
  }
  // transposed-conv gather: one launch per output parity
  for (int py = 0; py < s; ++py)
    for (int px = 0; px < s; ++px) {
      if (py >= out->h || px >= out->w) continue;
      TcTaps taps;
      memset(&taps, 0, sizeof(taps));
      for (int ky = 0; ky < g->kh; ++ky) {
        if ((py + g->pad - ky) % s != 0) continue;
        for (int kx = 0; kx < g->kw; ++kx) {
          if ((px + g->pad - kx) % s != 0) continue;
          int t = taps.ntaps++;
          taps.dy[t] = (int8_t)((py + g->pad - ky) / s);
          taps.dx[t] = (int8_t)((px + g->pad - kx) / s);
          taps.map[t] = 0;
          taps.wtap[t] = (uint8_t)(ky * g->kw + kx);
        }
      }
      SubView ov = subview(out, s, py, px);
      SubView iv = subview(in, 1, 0, 0);
      if (taps.ntaps == 0) continue;  // (cannot happen for k >= s)
      if (fprop_launch(ov, &iv, 1, taps, wp, taps_total, in->c, out->c, bias, act, stats, s == 1 ? stats_rows : nullptr, scratch,
                       scratch_bytes, st))
        return -1;
    }
  return 0;
}

// ws == nullptr: plan only (*used = number of slices the launch would write).
int conv_wgrad_tc(const gsb_view* in, const gsb_view* dout, float* ws, int max_slices, int* used, const gsb_conv_geom* g,
                  cudaStream_t st) {
  const int s = g->stride;
  const int taps_total = g->kh * g->kw;
  const bool conv = g->mode == GSB_MODE_CONV;
  // driver (tiled, unshifted) side: dY for conv, X for transposed conv; shifted side is the other one
  const gsb_view* tiled = conv ? dout : in;
  const gsb_view* shv = conv ? in : dout;
  TcTaps taps;
  memset(&taps, 0, sizeof(taps));
  SubView subs[4];
  int map_of[4] = {-1, -1, -1, -1}, nmaps = 0;
  for (int ky = 0; ky < g->kh; ++ky)
    for (int kx = 0; kx < g->kw; ++kx) {
      int dyv = ky - g->pad, dxv = kx - g->pad;  // shifted = tiled*s + (k - pad) for both modes
      int ay = floor_div(dyv, s), ax = floor_div(dxv, s);
      int py = dyv - ay * s, px = dxv - ax * s;
      if (py >= shv->h || px >= shv->w) continue;
      int key = py * s + px;
      if (map_of[key] < 0) {
        map_of[key] = nmaps;
        subs[nmaps++] = subview(shv, s, py, px);
      }
      int t = taps.ntaps++;
      taps.dy[t] = (int8_t)ay; taps.dx[t] = (int8_t)ax; taps.map[t] = (int8_t)map_of[key];
      taps.wtap[t] = (uint8_t)(ky * g->kw + kx);
    }
  GSB_REQUIRE(taps.ntaps > 0, "conv_wgrad_tc: no taps");
  if (getenv("GSB_TC_HALO_WGRAD") == nullptr || getenv("GSB_TC_HALO_WGRAD")[0] != '0') {
    int r = conv_halo_wgrad(subview(tiled, 1, 0, 0), subs, nmaps, taps, shv->c, conv ? 1 : 0, in->c, dout->c, taps_total, ws,
                            max_slices, used, st);
    if (r != 0) return r < 0 ? -1 : 0;
  }
  TcMaps maps;
  memset(&maps, 0, sizeof(maps));
  int tw, th, tn;
  pick_box(tiled->w, tiled->h, 64, tw, th, tn);
  if (ws != nullptr) {
    for (int i = 0; i < nmaps; ++i)
      if (make_act_map(&maps.in[i], subs[i], tw, th, tn)) return -1;
    for (int i = nmaps; i < 4; ++i) maps.in[i] = maps.in[0];
    SubView tv = subview(tiled, 1, 0, 0);
    if (make_act_map(&maps.w, tv, tw, th, tn)) return -1;
    maps.out = maps.w;
  }
  const int b_ch = tiled->c;
  const int BN = (b_ch % 128 == 0) ? 128 : 64;
  WgParams p;
  p.tw = tw; p.th = th; p.tn = tn;
  p.tiles_x = (tiled->w + tw - 1) / tw;
  p.tiles_y = (tiled->h + th - 1) / th;
  p.tiles_z = (tiled->n + tn - 1) / tn;
  p.boxes = p.tiles_x * p.tiles_y * p.tiles_z;
  p.a_cblocks = shv->c / 64;
  p.a_blocks = taps.ntaps * p.a_cblocks;
  p.n_tiles = b_ch / BN;
  p.a_is_in = conv ? 1 : 0;
  p.cin = in->c; p.cout = dout->c;
  p.dwp = ws;
  p.slice_elems = (int64_t)taps_total * dout->c * in->c;
  const int m_tiles = (p.a_blocks + 1) / 2;
  const int64_t base = (int64_t)m_tiles * p.n_tiles;
  int64_t want = ((int64_t)num_sms_avail() * 2 + base - 1) / base;  // ~2 waves
  int64_t max_split = (p.boxes + 15) / 16;                    // >= 16 k-steps per CTA
  if (max_split < 1) max_split = 1;
  if (ws != nullptr && max_split > max_slices) max_split = max_slices;
  int64_t split = want < 1 ? 1 : (want > max_split ? max_split : want);
  p.boxes_per_split = (int)((p.boxes + split - 1) / split);
  split = (p.boxes + p.boxes_per_split - 1) / p.boxes_per_split;
  GSB_REQUIRE(split <= 65535, "conv_wgrad_tc: split too large");
  *used = (int)split;
  if (ws == nullptr) return 0;
  GSB_REQUIRE(split <= max_slices, "conv_wgrad_tc: %d slices needed, workspace holds %d", (int)split, max_slices);
  if (taps.ntaps != taps_total) {
    // taps whose parity sub-view is empty never hit (inputs narrower than the stride): their entries stay zero
    cudaError_t e = cudaMemsetAsync(ws, 0, (size_t)split * p.slice_elems * sizeof(float), st);
    GSB_REQUIRE(e == cudaSuccess, "conv_wgrad_tc: memset failed: %s", cudaGetErrorString(e));
  }
  if (BN == 128) return launch_tc(conv_tc_wgrad_kernel<128>, (int)base, (int)split, WgradCfg<128>::SMEM, maps, taps, p, st);
  return launch_tc(conv_tc_wgrad_kernel<64>, (int)base, (int)split, WgradCfg<64>::SMEM, maps, taps, p, st);
}

}  // namespace gsb

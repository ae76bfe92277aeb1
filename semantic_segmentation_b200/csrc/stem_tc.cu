// Network-input ("stem") convolutions on the tensor cores: Conv2d(1 -> 64, 3x3 s1 p1) of the U-Net (unet_parts.py:16 via
// unet_model.py:14) and Conv2d(1 -> 64, 4x4 s2 p1) of the Pix2Pix generator (networks.py:582), forward and weight gradient,
// reading the fp32 NCHW network input directly.
//
// The SIMT stem kernels (edge.cu) spend 576 FMAs per pixel on 9 input values and run at 0.17 - 0.24 of the HBM roofline
// (issue-bound).  Here a CTA turns 128 consecutive output pixels into an im2col tile [128 pixels][64 columns] in shared
// memory (bf16, the SWIZZLE_128B layout TMA would write) and lets tcgen05.mma do the arithmetic; the threads only gather,
// convert and store, so the kernels are bound by the bf16 output / gradient stream.
//
// fp32 fidelity: the input is fp32 and stays fp32 in effect.  Every value is split x = xh + xl (xh = bf16(x), xl = bf16(x - xh),
// |x - xh - xl| <= 2^-17 |x|), the weights likewise, and the K dimension carries the three significant products:
//     forward   columns [xh(T) | xl(T) | xh(T)] . rows [wh(T) | wh(T) | wl(T)]   (T = cin*k*k <= 21; 3T <= 63 of K = 64)
//     wgrad     columns [xh(T) | xl(T)] against the bf16 gradient (exact), dW[co][t] = D[t][co] + D[T + t][co]
// with fp32 accumulation in TMEM: 1e-5 relative, far below the bf16 rounding of the stored output.
#include <stdlib.h>

#include "tc_common.cuh"

namespace gsb {

constexpr int kStemTcRowBytes = 128;   // one im2col row = 64 bf16 columns

// 16-byte chunk j of row r in a SWIZZLE_128B tile (8-row x 128-byte atoms, chunk index XOR row-in-atom)
__device__ __forceinline__ uint32_t swz_off(int r, int j) {
  return (uint32_t)((r >> 3) * 1024 + (r & 7) * 128 + ((j ^ (r & 7)) << 4));
}
__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}

// hi / lo bf16 split of T values into COPIES * T columns (copy 0 = hi, 1 = lo, 2 = hi), zero padded to 64 columns; writes
// the first NCH 16-byte chunks of row r at their swizzled positions.
template <int T, int COPIES, int NCH>
__device__ __forceinline__ void write_split_row(uint8_t* tile, int r, const float (&v)[T]) {
  float e[NCH * 8];
#pragma unroll
  for (int i = 0; i < NCH * 8; ++i) e[i] = 0.f;
#pragma unroll
  for (int t = 0; t < T; ++t) {
    const float hi = __bfloat162float(__float2bfloat16_rn(v[t]));
    e[t] = hi;
    e[T + t] = v[t] - hi;          // rounded to bf16 by the pack below
    if (COPIES == 3) e[2 * T + t] = hi;
  }
#pragma unroll
  for (int j = 0; j < NCH; ++j) {
    uint4 q;
    q.x = pack_bf16(e[8 * j + 0], e[8 * j + 1]);
    q.y = pack_bf16(e[8 * j + 2], e[8 * j + 3]);
    q.z = pack_bf16(e[8 * j + 4], e[8 * j + 5]);
    q.w = pack_bf16(e[8 * j + 6], e[8 * j + 7]);
    *reinterpret_cast<uint4*>(tile + swz_off(r, j)) = q;
  }
}

// The T input values under the receptive field of output pixel p (zero outside the image / past the end)
template <int T, int K, int S>
__device__ __forceinline__ void gather_taps(const float* __restrict__ x, int H, int W, int OH, int OW, int pad, int64_t p,
                                            bool valid, float (&v)[T]) {
  constexpr int CIN = T / (K * K);
  const int ox = (int)(p % OW);
  const int64_t r = p / OW;
  const int oy = (int)(r % OH);
  const int64_t n = r / OH;
#pragma unroll
  for (int ci = 0; ci < CIN; ++ci)
#pragma unroll
    for (int ky = 0; ky < K; ++ky) {
      const int iy = oy * S - pad + ky;
      const bool yok = valid && iy >= 0 && iy < H;
      const float* row = x + ((n * CIN + ci) * H + (yok ? iy : 0)) * (int64_t)W;
#pragma unroll
      for (int kx = 0; kx < K; ++kx) {
        const int ix = ox * S - pad + kx;
        v[(ci * K + ky) * K + kx] = (yok && ix >= 0 && ix < W) ? __ldg(row + ix) : 0.f;
      }
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// forward: y[p][co] = sum_t x[p @ t] w[co][t], bf16 NHWC out, optional BatchNorm statistics of the STORED values
// ---------------------------------------------------------------------------------------------------------------------
template <int T, int K, int S>
__global__ void __launch_bounds__(128, 6) stem_tc_fwd_kernel(const __grid_constant__ CUtensorMap ymap, int use_tma,
                                                         const float* __restrict__ x, int H, int W, const float* __restrict__ w,
                                                         __nv_bfloat16* __restrict__ y, int64_t yld, int OH, int OW, int pad,
                                                         float* __restrict__ stats, int64_t P, int num_tiles,
                                                         const float* __restrict__ aff_scale, const float* __restrict__ aff_shift,
                                                         int aff_relu) {
  GSB_PDL_ENTRY();
  constexpr int KS = (3 * T + 15) / 16;       // K = 16 steps that hold non-zero columns
  constexpr int NCH = 2 * KS;                 // 16-byte chunks per row that the MMAs read
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = smem;                         // [128 pixels][64 columns] bf16, swizzled; re-used as the output staging tile
  uint8_t* sB = smem + 16384;                 // [64 co][64 columns]
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 16384 + 8192);
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(bar + 1);
  float* red = reinterpret_cast<float*>(bar + 2);   // [4 row groups][2][64]; with an affine epilogue: [scale 64 | shift 64]
  const int tid = threadIdx.x, warp = tid >> 5;
  if (tid == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (warp == 0) tmem_alloc(tmem_holder, 64);
  if (tid < 64) {   // weight rows [wh | wh | wl]
    float wv[T];
#pragma unroll
    for (int t = 0; t < T; ++t) wv[t] = __ldg(w + tid * T + t);
    float e[NCH * 8];
#pragma unroll
    for (int i = 0; i < NCH * 8; ++i) e[i] = 0.f;
#pragma unroll
    for (int t = 0; t < T; ++t) {
      const float hi = __bfloat162float(__float2bfloat16_rn(wv[t]));
      e[t] = hi;
      e[T + t] = hi;
      e[2 * T + t] = wv[t] - hi;
    }
#pragma unroll
    for (int j = 0; j < NCH; ++j) {
      uint4 q;
      q.x = pack_bf16(e[8 * j + 0], e[8 * j + 1]);
      q.y = pack_bf16(e[8 * j + 2], e[8 * j + 3]);
      q.z = pack_bf16(e[8 * j + 4], e[8 * j + 5]);
      q.w = pack_bf16(e[8 * j + 6], e[8 * j + 7]);
      *reinterpret_cast<uint4*>(sB + swz_off(tid, j)) = q;
    }
  }
  // inference: out = act(acc * scale[co] + shift[co]) (BatchNorm of the running statistics folded in; no statistics then)
  const bool affine = aff_scale != nullptr;
  if (affine && tid < 64) {
    red[tid] = __ldg(aff_scale + tid);
    red[64 + tid] = aff_shift ? __ldg(aff_shift + tid) : 0.f;
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;
  constexpr uint32_t idesc = umma_idesc(128, 64, 0, 0);
  const uint64_t adesc = umma_desc(smem_u32(sA), 16, 1024), bdesc = umma_desc(smem_u32(sB), 16, 1024);
  uint32_t phase = 0;
  const int cp = tid & 31, rg = tid >> 5;     // statistics: channel pair (2 cp, 2 cp + 1), rows [32 rg, 32 rg + 32)
  float2 s0 = make_float2(0.f, 0.f), s1 = make_float2(0.f, 0.f);
  // software pipeline: the input values of the NEXT tile are gathered (global loads in flight) while the tensor core, the
  // epilogue and the statistics pass work on the current one -- the serial chain gather -> MMA -> store per tile was the limit
  float v[T];
  int64_t p = (int64_t)blockIdx.x * 128 + tid;
  bool valid = p < P;
  if ((int)blockIdx.x < num_tiles) gather_taps<T, K, S>(x, H, W, OH, OW, pad, p, valid, v);
  for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
    write_split_row<T, 3, NCH>(sA, tid, v);
    fence_proxy_async();            // generic-proxy writes -> visible to the tensor core's async-proxy reads
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) umma_bf16(tmem_base, adesc + 2 * ks, bdesc + 2 * ks, idesc, ks > 0 ? 1u : 0u);
      umma_commit(bar);
    }
    const int64_t p_next = (int64_t)(tile + gridDim.x) * 128 + tid;
    const bool valid_next = p_next < P;
    if (tile + (int)gridDim.x < num_tiles) gather_taps<T, K, S>(x, H, W, OH, OW, pad, p_next, valid_next, v);
    mbar_wait(bar, phase);
    phase ^= 1u;
    tc_fence_after();
    uint32_t a0[32], a1[32];
    tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16), a0);
    tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + 32, a1);
    if (affine) {
      const float4* r4 = reinterpret_cast<const float4*>(red);   // 16-byte broadcast reads: [scale 64 | shift 64]
#pragma unroll
      for (int c4 = 0; c4 < 8; ++c4) {
        const float4 s0 = r4[c4], s1 = r4[8 + c4], h0 = r4[16 + c4], h1 = r4[24 + c4];
        const float sa[4] = {s0.x, s0.y, s0.z, s0.w}, sb[4] = {s1.x, s1.y, s1.z, s1.w};
        const float ha[4] = {h0.x, h0.y, h0.z, h0.w}, hb[4] = {h1.x, h1.y, h1.z, h1.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          float f0 = fmaf(__uint_as_float(a0[4 * c4 + k]), sa[k], ha[k]);
          float f1 = fmaf(__uint_as_float(a1[4 * c4 + k]), sb[k], hb[k]);
          if (aff_relu) { f0 = fmaxf(f0, 0.f); f1 = fmaxf(f1, 0.f); }
          a0[4 * c4 + k] = __float_as_uint(f0);
          a1[4 * c4 + k] = __float_as_uint(f1);
        }
      }
    }
    uint4 q[8];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      q[j].x = pack_bf16(__uint_as_float(a0[8 * j + 0]), __uint_as_float(a0[8 * j + 1]));
      q[j].y = pack_bf16(__uint_as_float(a0[8 * j + 2]), __uint_as_float(a0[8 * j + 3]));
      q[j].z = pack_bf16(__uint_as_float(a0[8 * j + 4]), __uint_as_float(a0[8 * j + 5]));
      q[j].w = pack_bf16(__uint_as_float(a0[8 * j + 6]), __uint_as_float(a0[8 * j + 7]));
      q[4 + j].x = pack_bf16(__uint_as_float(a1[8 * j + 0]), __uint_as_float(a1[8 * j + 1]));
      q[4 + j].y = pack_bf16(__uint_as_float(a1[8 * j + 2]), __uint_as_float(a1[8 * j + 3]));
      q[4 + j].z = pack_bf16(__uint_as_float(a1[8 * j + 4]), __uint_as_float(a1[8 * j + 5]));
      q[4 + j].w = pack_bf16(__uint_as_float(a1[8 * j + 6]), __uint_as_float(a1[8 * j + 7]));
    }
    if (valid && !use_tma) {   // strided output view: one 128-byte row per thread
      uint4* dst = reinterpret_cast<uint4*>(y + p * yld);
#pragma unroll
      for (int j = 0; j < 8; ++j) dst[j] = q[j];
    }
    if (stats != nullptr || use_tma) {
      // stage the stored tile (zero rows past the end) in the A buffer -- the MMAs are done with it -- in the 128-byte
      // swizzle: that is the layout a TMA store reads (dense output: ONE bulk store of 16 KB per tile instead of 8
      // row-strided 16-byte stores per thread, which kept the load / store unit queue full: ncu profiles/r02k, lg_throttle
      // 3.9 of 11 stall cycles per issue), and it keeps the column sums below conflict free
#pragma unroll
      for (int j = 0; j < 8; ++j)
        *reinterpret_cast<uint4*>(sA + tid * 128 + ((j ^ (tid & 7)) << 4)) = valid ? q[j] : make_uint4(0u, 0u, 0u, 0u);
      if (use_tma) fence_proxy_async();
      __syncthreads();
      if (use_tma && tid == 0) {
        tma_store_2d(&ymap, sA, 0, tile * 128);     // rows past P are clipped by the tensor map
        tma_store_commit();
      }
    }
    if (stats != nullptr) {
      // thread (cp, rg) walks 32 rows of one 32-bit word column of the staged tile
      const uint32_t* s32 = reinterpret_cast<const uint32_t*>(sA);
#pragma unroll 8
      for (int i = 0; i < 32; ++i) {
        const int rr = rg * 32 + i;
        const uint32_t wv = s32[rr * 32 + ((((cp >> 2) ^ (rr & 7)) << 2) | (cp & 3))];
        const float2 f = make_float2(__uint_as_float(wv << 16), __uint_as_float(wv & 0xffff0000u));
        s0 = __fadd2_rn(s0, f);
        s1 = __ffma2_rn(f, f, s1);
      }
    }
    if (use_tma && tid == 0) tma_store_wait_read();   // the bulk store has read the staging tile
    tc_fence_before();
    __syncthreads();                // accumulator and staging tile are free for the next tile
    p = p_next;
    valid = valid_next;
  }
  if (stats != nullptr) {
    red[(rg * 2 + 0) * 64 + 2 * cp] = s0.x;
    red[(rg * 2 + 0) * 64 + 2 * cp + 1] = s0.y;
    red[(rg * 2 + 1) * 64 + 2 * cp] = s1.x;
    red[(rg * 2 + 1) * 64 + 2 * cp + 1] = s1.y;
    __syncthreads();
    {   // 128 threads = [2][64]: the four row groups are added in order (deterministic)
      const int which = tid >> 6, c = tid & 63;
      float t = 0.f;
#pragma unroll
      for (int g = 0; g < 4; ++g) t += red[(g * 2 + which) * 64 + c];
      stats[(int64_t)blockIdx.x * 128 + which * 64 + c] = t;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 64);
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// weight gradient: dw[co][t] = sum_p dY[p][co] x[p @ t].  D[m = im2col column][n = co] accumulates in TMEM over all the
// tiles of a CTA (M = 128: the tile's 64 zero columns of a second block pad it); one partial row per CTA, added in CTA
// order by gsb::sum_slices.  The gradient tile arrives by TMA, the im2col tile is built by the threads.
// ---------------------------------------------------------------------------------------------------------------------
template <int T, int K, int S>
__global__ void __launch_bounds__(128) stem_tc_wgrad_kernel(const __grid_constant__ CUtensorMap dymap, const float* __restrict__ x,
                                                           int H, int W, int OH, int OW, int pad, float* __restrict__ partial,
                                                           int64_t P, int num_tiles) {
  GSB_PDL_ENTRY();
  constexpr int NCH = (2 * T + 7) / 8;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = smem;                         // [2 blocks][128 pixels][64 columns]: block 1 and the unused chunks stay zero
  uint8_t* sB = smem + 32768;                 // gradient tile [128 pixels][64 co], TMA, SWIZZLE_128B
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + 32768 + 16384);
  uint64_t* done = full + 1;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(full + 2);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    mbar_init(full, 1);
    mbar_init(done, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    prefetch_tmap(&dymap);
  }
  for (int i = tid; i < 32768 / 16; i += 128) reinterpret_cast<uint4*>(sA)[i] = make_uint4(0u, 0u, 0u, 0u);
  __syncthreads();
  if (warp == 0) tmem_alloc(tmem_holder, 64);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;
  constexpr uint32_t idesc = umma_idesc(128, 64, 1, 1);
  const uint64_t adesc = umma_desc(smem_u32(sA), 16384, 1024), bdesc = umma_desc(smem_u32(sB), 16384, 1024);
  int it = 0;
  for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
    if (it > 0) mbar_wait(done, (uint32_t)((it - 1) & 1));   // the previous tile's MMAs have read both buffers
    if (tid == 0) {
      mbar_expect_tx(full, 16384);
      tma_load_2d(&dymap, full, sB, 0, tile * 128);          // rows past the end are zero filled
    }
    const int64_t p = (int64_t)tile * 128 + tid;
    float v[T];
    gather_taps<T, K, S>(x, H, W, OH, OW, pad, p, p < P, v);
    write_split_row<T, 2, NCH>(sA, tid, v);
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      mbar_wait(full, (uint32_t)(it & 1));
      tc_fence_after();
#pragma unroll
      for (int ks = 0; ks < 8; ++ks)   // 16 pixels (K) per MMA = 2048 bytes further down both tiles
        umma_bf16(tmem_base, adesc + (uint64_t)(128 * ks), bdesc + (uint64_t)(128 * ks), idesc, (it | ks) != 0 ? 1u : 0u);
      umma_commit(done);
    }
  }
  if (it > 0) mbar_wait(done, (uint32_t)((it - 1) & 1));
  tc_fence_after();
  if (warp == 0) {   // lanes = im2col columns m (hi taps 0..T-1, lo taps T..2T-1), registers = co
    uint32_t d0[32], d1[32];
    tmem_ld32(tmem_base, d0);
    tmem_ld32(tmem_base + 32, d1);
    float* row = partial + (int64_t)blockIdx.x * 64 * T;
#pragma unroll
    for (int c = 0; c < 32; ++c) {
      float a = __uint_as_float(d0[c]), b = __uint_as_float(d1[c]);
      if (it == 0) a = b = 0.f;
      const float a2 = __shfl_down_sync(0xffffffffu, a, T), b2 = __shfl_down_sync(0xffffffffu, b, T);
      if (lane < T) {
        row[c * T + lane] = a + a2;
        row[(32 + c) * T + lane] = b + b2;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 64);
  }
}

static bool stem_tc_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("GSB_STEM_TC");
    v = (e != nullptr && e[0] == '0') ? 0 : 1;
  }
  return v != 0;
}

// Shapes the tensor-core stem covers: bf16 NHWC 64-channel side, cin * k * k in {9 (3x3 s1), 16 (4x4 s2)}, cin = 1.
bool stem_tc_applicable(int cin, int cout, int k, int stride, int dtype, int64_t pixels) {
  if (!stem_tc_enabled() || !gsb_tc_available() || dtype != GSB_BF16 || cout != 64 || cin != 1) return false;
  if (pixels <= 0 || pixels >= ((int64_t)1 << 37)) return false;
  return (k == 3 && stride == 1) || (k == 4 && stride == 2);
}

// Returns the number of statistic rows written (= CTAs) through *rows; 0 on success.
int stem_tc_fwd(const float* x, int h, int w, const float* weight, const gsb_view* y, int k, int stride, int pad, float* stats,
                int max_rows, int* rows, cudaStream_t st, const float* aff_scale, const float* aff_shift, int aff_relu) {
  const int64_t P = num_pixels(y);
  const int num_tiles = (int)((P + 127) / 128);
  int grid = num_sms() * 8;
  if (grid > num_tiles) grid = num_tiles;
  if (stats != nullptr && grid > max_rows) grid = max_rows;
  const int smem = 1024 + 16384 + 8192 + 64 + 4 * 2 * 64 * (int)sizeof(float);
  __nv_bfloat16* yp = reinterpret_cast<__nv_bfloat16*>(y->ptr);
  static const bool tma_on = []() { const char* e = getenv("GSB_STEM_TMA_STORE"); return !(e && e[0] == '0'); }();
  const int use_tma = (tma_on && y->sw == 64 && ((uintptr_t)y->ptr & 15) == 0 && P < ((int64_t)1 << 31)) ? 1 : 0;
  CUtensorMap ymap;
  memset(&ymap, 0, sizeof(ymap));
  if (use_tma && make_w_map(&ymap, y->ptr, P, 64, 128)) return -1;
  if (k == 3)
    gsb::launch(stem_tc_fwd_kernel<9, 3, 1>, gsb::LaunchCfg(grid, 128, smem, st), ymap, use_tma, x, h, w, weight, yp, (int64_t)y->sw,
                y->h, y->w, pad, stats, P, num_tiles, aff_scale, aff_shift, aff_relu);
  else
    gsb::launch(stem_tc_fwd_kernel<16, 4, 2>, gsb::LaunchCfg(grid, 128, smem, st), ymap, use_tma, x, h, w, weight, yp, (int64_t)y->sw,
                y->h, y->w, pad, stats, P, num_tiles, aff_scale, aff_shift, aff_relu);
  if (rows) *rows = stats != nullptr ? grid : 0;
  return 0;
}

// partial: [*rows][64 * cin * k * k] fp32 rows (one per CTA), to be added in row order by the caller.
int stem_tc_wgrad(const float* x, int h, int w, const gsb_view* dy, int k, int stride, int pad, float* partial, int max_rows,
                  int* rows, cudaStream_t st) {
  (void)stride;
  const int64_t P = num_pixels(dy);
  const int num_tiles = (int)((P + 127) / 128);
  int grid = num_sms() * 4;
  if (grid > num_tiles) grid = num_tiles;
  if (grid > max_rows) grid = max_rows;
  GSB_REQUIRE(grid >= 1, "stem_tc_wgrad: workspace holds no row");
  CUtensorMap map;
  if (make_w_map(&map, dy->ptr, P, 64, 128)) return -1;
  const int smem = 1024 + 32768 + 16384 + 64;
  cudaError_t e;
  if (k == 3) {
    e = cudaFuncSetAttribute(stem_tc_wgrad_kernel<9, 3, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    GSB_REQUIRE(e == cudaSuccess, "stem_tc_wgrad: smem attribute: %s", cudaGetErrorString(e));
    gsb::launch(stem_tc_wgrad_kernel<9, 3, 1>, gsb::LaunchCfg(grid, 128, smem, st), map, x, h, w, dy->h, dy->w, pad, partial, P, num_tiles);
  } else {
    e = cudaFuncSetAttribute(stem_tc_wgrad_kernel<16, 4, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    GSB_REQUIRE(e == cudaSuccess, "stem_tc_wgrad: smem attribute: %s", cudaGetErrorString(e));
    gsb::launch(stem_tc_wgrad_kernel<16, 4, 2>, gsb::LaunchCfg(grid, 128, smem, st), map, x, h, w, dy->h, dy->w, pad, partial, P, num_tiles);
  }
  *rows = grid;
  return 0;
}

}  // namespace gsb

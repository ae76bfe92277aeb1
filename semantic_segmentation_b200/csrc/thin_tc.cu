// The generator's outermost up-cell on the tensor cores: ConvTranspose2d(CW -> 1, 8x8 s2 p3) -- the merged 4x4 / 6x6 / 8x8 cell of
// networks.py:590-592 with operations.py:14-39 -- forward, data gradient and weight gradient.
//
// One side of this layer has a single channel, so none of its three passes is a GEMM over channels; the SIMT kernels
// (thin_conv.cu) spend 64 taps x CW channels of fp32 FMAs per pixel behind scalar index arithmetic and take 365 / 180 / 420 us
// per step at 8 x 256 x 256 (profiles/r02n_p2p_launch_summary.txt) for 2.1 GFLOP each.  Seen from the WIDE tensor
// X[p][ci] (p = input pixel of the transposed convolution) all three are GEMMs with the 64 taps as a matrix dimension:
//     forward    Z[p][t]   = sum_ci X[p][ci] * Wp[t][ci]      (M = pixels, N = 64 taps, K = CW)   + col2im: out[q] = sum Z[p][t]
//     dgrad      dX[p][ci] = sum_t  A[p][t]  * Wp[t][ci]      (M = pixels, N = CW,      K = 64 taps), A = im2col of dOut
//     wgrad      dW[t][ci] = sum_p  A[p][t]  * X[p][ci]       (M = 64 taps, N = CW,     K = pixels)
// with q = 2 p - pad + k the output pixel that tap k = (ky, kx) of input pixel p lands on.  A (the im2col tile of the
// one-channel fp32 tensor, rounded to bf16 like every other gradient / activation of the bf16 mode) is built by the threads
// in the SWIZZLE_128B layout TMA would write, exactly as the stem kernels do (stem_tc.cu); X and dX move by TMA.
#include <stdlib.h>

#include "tc_common.cuh"

namespace gsb {

namespace {

constexpr int kTaps = 64, kK = 8, kS = 2;

__device__ __forceinline__ uint32_t tswz(int r, int j) {   // 16-byte chunk j of row r in a SWIZZLE_128B tile
  return (uint32_t)((r >> 3) * 1024 + (r & 7) * 128 + ((j ^ (r & 7)) << 4));
}
__device__ __forceinline__ uint32_t tpack(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}

struct ThinSide {      // the one-channel fp32 tensor [N][OH][OW] (any strides)
  const float* ptr;
  int64_t sn, sh, sw;
  int OH, OW;
};

// Row `row` of the im2col tile: the 64 values d[n][2 py - pad + ky][2 px - pad + kx] of wide pixel p (zero outside the
// tensor / past the end), bf16, chunk ky = the eight kx taps of kernel row ky.
__device__ __forceinline__ void write_im2col_row(uint8_t* tile, int row, const ThinSide& d, int H, int W, int pad, int64_t p,
                                                 bool valid) {
  const int px = (int)(p % W);
  const int64_t r = p / W;
  const int py = (int)(r % H);
  const int64_t n = r / H;
  const int qx0 = px * kS - pad;
#pragma unroll
  for (int ky = 0; ky < kK; ++ky) {
    const int qy = py * kS - pad + ky;
    const bool yok = valid && qy >= 0 && qy < d.OH;
    const float* rowp = d.ptr + n * d.sn + (int64_t)(yok ? qy : 0) * d.sh;
    float v[kK];
#pragma unroll
    for (int kx = 0; kx < kK; ++kx) {
      const int qx = qx0 + kx;
      v[kx] = (yok && qx >= 0 && qx < d.OW) ? __ldg(rowp + (int64_t)qx * d.sw) : 0.f;
    }
    *reinterpret_cast<uint4*>(tile + tswz(row, ky)) =
        make_uint4(tpack(v[0], v[1]), tpack(v[2], v[3]), tpack(v[4], v[5]), tpack(v[6], v[7]));
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// dgrad: dX[p][ci] = sum_t A[p][t] * Wp[t][ci].  A K-major (built here), B = Wp[t][ci] is [K][N]: MN-major blocks of 64 ci.
// ---------------------------------------------------------------------------------------------------------------------
template <int CW>
__global__ void __launch_bounds__(128) thin_tc_dgrad_kernel(const __grid_constant__ CUtensorMap dxmap, ThinSide d, int H, int W,
                                                            int pad, const __nv_bfloat16* __restrict__ wp, int64_t P,
                                                            int num_tiles) {
  GSB_PDL_ENTRY();
  constexpr int NB = CW / 64;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = smem;                          // [128 pixels][64 taps]
  uint8_t* sB = smem + 16384;                  // NB blocks [64 taps][64 ci]
  uint8_t* sO = sB + NB * 8192;                // NB blocks [128 pixels][64 ci]: output staging for the TMA store
  uint64_t* bar = reinterpret_cast<uint64_t*>(sO + NB * 16384);
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(bar + 1);
  const int tid = threadIdx.x, warp = tid >> 5;
  if (tid == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    prefetch_tmap(&dxmap);
  }
  __syncthreads();
  if (warp == 0) tmem_alloc(tmem_holder, CW);
  for (int i = tid; i < kTaps * (CW / 8); i += 128) {   // weights: row t, chunk c (8 channels) -> block c / 8
    const int t = i / (CW / 8), c = i % (CW / 8);
    const uint4 q = __ldg(reinterpret_cast<const uint4*>(wp + (int64_t)t * CW + c * 8));
    *reinterpret_cast<uint4*>(sB + (c >> 3) * 8192 + tswz(t, c & 7)) = q;
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;
  constexpr uint32_t idesc = umma_idesc(128, CW, 0, 1);
  const uint64_t adesc = umma_desc(smem_u32(sA), 16, 1024), bdesc = umma_desc(smem_u32(sB), 8192, 1024);
  uint32_t phase = 0;
  for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
    const int64_t p = (int64_t)tile * 128 + tid;
    write_im2col_row(sA, tid, d, H, W, pad, p, p < P);
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
#pragma unroll
      for (int ks = 0; ks < 4; ++ks)   // 16 taps per MMA: +32 B along the K-major A rows, +16 rows (2048 B) down B
        umma_bf16(tmem_base, adesc + 2 * ks, bdesc + 128 * ks, idesc, ks > 0 ? 1u : 0u);
      umma_commit(bar);
      tma_store_wait_read();           // the previous tile's bulk stores have read the staging buffers
    }
    mbar_wait(bar, phase);
    phase ^= 1u;
    tc_fence_after();
    __syncthreads();                   // (thread 0 has passed the store wait)
#pragma unroll
    for (int c = 0; c < CW / 32; ++c) {
      uint32_t a[32];
      tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + c * 32, a);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const uint4 q = make_uint4(tpack(__uint_as_float(a[8 * j + 0]), __uint_as_float(a[8 * j + 1])),
                                   tpack(__uint_as_float(a[8 * j + 2]), __uint_as_float(a[8 * j + 3])),
                                   tpack(__uint_as_float(a[8 * j + 4]), __uint_as_float(a[8 * j + 5])),
                                   tpack(__uint_as_float(a[8 * j + 6]), __uint_as_float(a[8 * j + 7])));
        *reinterpret_cast<uint4*>(sO + (c >> 1) * 16384 + tswz(tid, (c & 1) * 4 + j)) = q;
      }
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();                   // staging complete; the accumulator and the A tile are free
    if (tid == 0) {
#pragma unroll
      for (int b = 0; b < NB; ++b) tma_store_2d(&dxmap, sO + b * 16384, b * 64, tile * 128);   // rows past P are clipped
      tma_store_commit();
    }
  }
  if (tid == 0) tma_store_wait_read();
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, CW);
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// wgrad: dW[t][ci] = sum_p A[p][t] * X[p][ci].  K = pixels: A (im2col, built here) and B (X tile, TMA) are both MN-major.
// M = 128 with the second 64-column block left zero (as the stem weight gradient does); every CTA accumulates its tiles
// in TMEM and STORES one partial-sum slice [64 taps][CW] (added in slice order by the caller: no atomics).
// ---------------------------------------------------------------------------------------------------------------------
template <int CW>
__global__ void __launch_bounds__(128) thin_tc_wgrad_kernel(const __grid_constant__ CUtensorMap xmap, ThinSide d, int H, int W,
                                                            int pad, float* __restrict__ slices, int64_t P, int num_tiles) {
  GSB_PDL_ENTRY();
  constexpr int NB = CW / 64;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = smem;                          // [2 blocks][128 pixels][64 columns]; block 1 stays zero
  uint8_t* sB = smem + 32768;                  // NB blocks [128 pixels][64 ci], TMA
  uint64_t* full = reinterpret_cast<uint64_t*>(sB + NB * 16384);
  uint64_t* done = full + 1;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(full + 2);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    mbar_init(full, 1);
    mbar_init(done, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    prefetch_tmap(&xmap);
  }
  for (int i = tid; i < 16384 / 16; i += 128) reinterpret_cast<uint4*>(sA + 16384)[i] = make_uint4(0u, 0u, 0u, 0u);
  __syncthreads();
  if (warp == 0) tmem_alloc(tmem_holder, CW);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;
  constexpr uint32_t idesc = umma_idesc(128, CW, 1, 1);
  const uint64_t adesc = umma_desc(smem_u32(sA), 16384, 1024), bdesc = umma_desc(smem_u32(sB), 16384, 1024);
  int it = 0;
  for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
    if (it > 0) mbar_wait(done, (uint32_t)((it - 1) & 1));   // the previous tile's MMAs have read both buffers
    if (tid == 0) {
      mbar_expect_tx(full, NB * 16384);
#pragma unroll
      for (int b = 0; b < NB; ++b) tma_load_2d(&xmap, full, sB + b * 16384, b * 64, tile * 128);   // rows past P: zeros
    }
    const int64_t p = (int64_t)tile * 128 + tid;
    write_im2col_row(sA, tid, d, H, W, pad, p, p < P);
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
  if (warp < 2) {   // TMEM lanes 0..63 = taps, columns = ci
    float* row = slices + ((int64_t)blockIdx.x * kTaps + warp * 32 + lane) * CW;
#pragma unroll
    for (int c = 0; c < CW / 32; ++c) {
      uint32_t a[32];
      tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + c * 32, a);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float4 v = make_float4(__uint_as_float(a[4 * j]), __uint_as_float(a[4 * j + 1]), __uint_as_float(a[4 * j + 2]),
                               __uint_as_float(a[4 * j + 3]));
        if (it == 0) v = make_float4(0.f, 0.f, 0.f, 0.f);   // a CTA without tiles still owns (and clears) its slice
        *reinterpret_cast<float4*>(row + c * 32 + 4 * j) = v;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, CW);
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// forward, stage 1: Z[p][t] = sum_ci X[p][ci] * Wp[t][ci] (fp32, caller scratch [P][64]).  A = X tile (TMA, K-major, CW/64
// atoms along K), B = Wp rows t (K-major).  Stage 2 (col2im) adds the 16 Z entries that land on each output pixel.
// ---------------------------------------------------------------------------------------------------------------------
template <int CW>
__global__ void __launch_bounds__(128) thin_tc_fwd_z_kernel(const __grid_constant__ CUtensorMap xmap,
                                                            const __nv_bfloat16* __restrict__ wp, float* __restrict__ z,
                                                            int64_t P, int num_tiles) {
  GSB_PDL_ENTRY();
  constexpr int NB = CW / 64;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sX = smem;                          // 2 stages x NB atoms [128 pixels][64 ci]
  uint8_t* sW = smem + 2 * NB * 16384;         // NB atoms [64 taps][64 ci]
  uint64_t* full = reinterpret_cast<uint64_t*>(sW + NB * 8192);   // [2]
  uint64_t* done = full + 2;                                       // [2]
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(full + 4);
  const int tid = threadIdx.x, warp = tid >> 5;
  if (tid == 0) {
    mbar_init(&full[0], 1); mbar_init(&full[1], 1);
    mbar_init(&done[0], 1); mbar_init(&done[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    prefetch_tmap(&xmap);
  }
  __syncthreads();
  if (warp == 0) tmem_alloc(tmem_holder, 128);   // two accumulators of 64 columns
  for (int i = tid; i < kTaps * (CW / 8); i += 128) {   // weights: row t, chunk c (8 channels) -> atom c / 8
    const int t = i / (CW / 8), c = i % (CW / 8);
    const uint4 q = __ldg(reinterpret_cast<const uint4*>(wp + (int64_t)t * CW + c * 8));
    *reinterpret_cast<uint4*>(sW + (c >> 3) * 8192 + tswz(t, c & 7)) = q;
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;
  constexpr uint32_t idesc = umma_idesc(128, 64, 0, 0);
  const int my_tiles = blockIdx.x < num_tiles ? (num_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
  auto load_tile = [&](int it) {   // thread 0: X tile `it` of this CTA into stage it & 1
    const int tile = blockIdx.x + it * gridDim.x, st = it & 1;
    mbar_expect_tx(&full[st], NB * 16384);
#pragma unroll
    for (int b = 0; b < NB; ++b) tma_load_2d(&xmap, &full[st], sX + (st * NB + b) * 16384, b * 64, tile * 128);
  };
  if (tid == 0 && my_tiles > 0) load_tile(0);
  for (int it = 0; it < my_tiles; ++it) {
    const int st = it & 1;
    if (tid == 0) {
      // stage st ^ 1 was read by the MMAs of tile it - 1: wait for them before the next load overwrites it
      if (it >= 1) mbar_wait(&done[st ^ 1], (uint32_t)(((it - 1) >> 1) & 1));
      if (it + 1 < my_tiles) load_tile(it + 1);
      mbar_wait(&full[st], (uint32_t)((it >> 1) & 1));
      tc_fence_after();
#pragma unroll
      for (int b = 0; b < NB; ++b) {
        const uint64_t adesc = umma_desc(smem_u32(sX + (st * NB + b) * 16384), 16, 1024);
        const uint64_t bdesc = umma_desc(smem_u32(sW + b * 8192), 16, 1024);
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) umma_bf16(tmem_base + st * 64, adesc + 2 * ks, bdesc + 2 * ks, idesc, (b | ks) != 0 ? 1u : 0u);
      }
      umma_commit(&done[st]);
    }
    if (it >= 1) {   // epilogue of tile it - 1 (accumulator st ^ 1) while the tensor core works on tile it
      const int pt = it - 1, ps = pt & 1;
      mbar_wait(&done[ps], (uint32_t)((pt >> 1) & 1));
      tc_fence_after();
      const int64_t p = (int64_t)(blockIdx.x + pt * gridDim.x) * 128 + tid;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t a[32];
        tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + ps * 64 + c * 32, a);
        if (p < P) {
          float4* dst = reinterpret_cast<float4*>(z + p * kTaps + c * 32);
#pragma unroll
          for (int j = 0; j < 8; ++j)
            dst[j] = make_float4(__uint_as_float(a[4 * j]), __uint_as_float(a[4 * j + 1]), __uint_as_float(a[4 * j + 2]),
                                 __uint_as_float(a[4 * j + 3]));
        }
      }
      tc_fence_before();
    }
    __syncthreads();   // accumulator st ^ 1 is drained before thread 0 issues tile it + 1 into it
  }
  if (my_tiles > 0) {
    const int pt = my_tiles - 1, ps = pt & 1;
    mbar_wait(&done[ps], (uint32_t)((pt >> 1) & 1));
    tc_fence_after();
    const int64_t p = (int64_t)(blockIdx.x + pt * gridDim.x) * 128 + tid;
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      uint32_t a[32];
      tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + ps * 64 + c * 32, a);
      if (p < P) {
        float4* dst = reinterpret_cast<float4*>(z + p * kTaps + c * 32);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          dst[j] = make_float4(__uint_as_float(a[4 * j]), __uint_as_float(a[4 * j + 1]), __uint_as_float(a[4 * j + 2]),
                               __uint_as_float(a[4 * j + 3]));
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 128);
  }
}

// forward, stage 2: out[n][qy][qx] = act(bias + sum over the taps (ky, kx) with qy + pad - ky, qx + pad - kx even and the
// input pixel inside the tensor of Z[p][ky * 8 + kx]); ascending (ky, kx) order -> deterministic.
__global__ void __launch_bounds__(256) thin_tc_col2im_kernel(const float* __restrict__ z, int N, int H, int W, int pad,
                                                             const float* __restrict__ bias, int act, float* __restrict__ out,
                                                             int64_t osn, int64_t osh, int64_t osw, int OH, int OW) {
  GSB_PDL_ENTRY();
  const int64_t total = (int64_t)N * OH * OW;
  const float b0 = bias ? __ldg(bias) : 0.f;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (int64_t)gridDim.x * 256) {
    const int qx = (int)(i % OW);
    const int64_t r = i / OW;
    const int qy = (int)(r % OH);
    const int64_t n = r / OH;
    float acc = 0.f;
    const int ky0 = (qy + pad) & 1, kx0 = (qx + pad) & 1;
#pragma unroll
    for (int a = 0; a < 4; ++a) {
      const int ky = ky0 + 2 * a, py = (qy + pad - ky) >> 1;
      if (py < 0 || py >= H) continue;
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const int kx = kx0 + 2 * b, px = (qx + pad - kx) >> 1;
        if (px < 0 || px >= W) continue;
        acc += __ldg(z + ((n * H + py) * (int64_t)W + px) * kTaps + ky * kK + kx);
      }
    }
    out[n * osn + (int64_t)qy * osh + (int64_t)qx * osw] = act_fwd(acc + b0, act);
  }
}

}  // namespace

static bool thin_tc_enabled() {   // read per call: tests and A/B runs switch it inside one process
  const char* e = getenv("GSB_THIN_TC");
  return !(e != nullptr && e[0] == '0');   // on by default (GPU tests: test_thin_tc_outermost_cell, the Pix2Pix suites)
}
static bool thin_side_ok(const gsb_view* v) { return v->c == 1 && v->dtype == GSB_F32 && v->ptr != nullptr; }
static bool wide_side_ok(const gsb_view* v) {   // dense pixel-linear bf16 NHWC with 64 or 128 channels
  return (v->c == 64 || v->c == 128) && v->dtype == GSB_BF16 && v->sc == 1 && v->sw == v->c && v->sh == (int64_t)v->w * v->c &&
         v->sn == (int64_t)v->h * v->w * v->c && ((uintptr_t)v->ptr & 15) == 0 && num_pixels(v) < ((int64_t)1 << 31);
}
static bool thin_tc_geom_ok(const gsb_view* wide, const gsb_view* thin, const gsb_conv_geom* g) {
  return g->kh == kK && g->kw == kK && g->stride == kS && g->pad >= 0 && g->pad < kK && thin->n == wide->n &&
         thin->h == (wide->h - 1) * kS - 2 * g->pad + kK && thin->w == (wide->w - 1) * kS - 2 * g->pad + kK;
}
static ThinSide thin_side(const gsb_view* v) {
  ThinSide t;
  t.ptr = reinterpret_cast<const float*>(v->ptr);
  t.sn = v->sn; t.sh = v->sh; t.sw = v->sw;
  t.OH = v->h; t.OW = v->w;
  return t;
}
static int tile_grid(int num_tiles, int per_sm) {
  const int cap = num_sms() * per_sm;
  return num_tiles < cap ? num_tiles : cap;
}
template <typename K>
static int set_smem(K kern, int smem) {
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  GSB_REQUIRE(e == cudaSuccess, "thin_tc: shared memory attribute (%d bytes): %s", smem, cudaGetErrorString(e));
  return 0;
}

// gsb_conv_gather hook: 1 = launched, 0 = not this layer shape (the caller goes on), -1 = error.
//   mode CONVT, in = wide, out = thin  : the layer's forward (needs scratch >= pixels(in) * 64 floats for Z)
//   mode CONV,  in = thin, out = wide  : the layer's data gradient
int thin_tc_gather_try(const gsb_view* in, const void* wp, const float* bias, const gsb_view* out, const gsb_conv_geom* g, int act,
                       void* scratch, size_t scratch_bytes, cudaStream_t st) {
  if (!thin_tc_enabled()) return 0;
  if (g->mode == GSB_MODE_CONVT && wide_side_ok(in) && thin_side_ok(out) && thin_tc_geom_ok(in, out, g)) {
    const int64_t P = num_pixels(in);
    if (scratch == nullptr || scratch_bytes < (size_t)P * kTaps * sizeof(float) || ((uintptr_t)scratch & 15) != 0) return 0;
    const int num_tiles = (int)((P + 127) / 128), CW = in->c, NB = CW / 64;
    CUtensorMap xmap;
    if (make_w_map(&xmap, in->ptr, P, CW, 128)) return -1;
    float* z = reinterpret_cast<float*>(scratch);
    const int smem = 1024 + 2 * NB * 16384 + NB * 8192 + 64;
    const __nv_bfloat16* w16 = reinterpret_cast<const __nv_bfloat16*>(wp);
    if (CW == 128) {
      if (set_smem(thin_tc_fwd_z_kernel<128>, smem)) return -1;
      gsb::launch(thin_tc_fwd_z_kernel<128>, gsb::LaunchCfg(tile_grid(num_tiles, 2), 128, smem, st), xmap, w16, z, P, num_tiles);
    } else {
      if (set_smem(thin_tc_fwd_z_kernel<64>, smem)) return -1;
      gsb::launch(thin_tc_fwd_z_kernel<64>, gsb::LaunchCfg(tile_grid(num_tiles, 4), 128, smem, st), xmap, w16, z, P, num_tiles);
    }
    const int64_t Q = num_pixels(out);
    int64_t blocks = (Q + 255) / 256;
    if (blocks > (int64_t)num_sms() * 16) blocks = (int64_t)num_sms() * 16;
    gsb::launch(thin_tc_col2im_kernel, gsb::LaunchCfg((int)blocks, 256, 0, st), (const float*)z, in->n, in->h, in->w, g->pad, bias, act,
                reinterpret_cast<float*>(out->ptr), out->sn, out->sh, out->sw, out->h, out->w);
    return check_launch("thin_tc_fwd", 2) ? -1 : 1;
  }
  if (g->mode == GSB_MODE_CONV && thin_side_ok(in) && wide_side_ok(out) && thin_tc_geom_ok(out, in, g) && bias == nullptr &&
      act == GSB_ACT_NONE) {
    const int64_t P = num_pixels(out);
    const int num_tiles = (int)((P + 127) / 128), CW = out->c, NB = CW / 64;
    CUtensorMap dxmap;
    if (make_w_map(&dxmap, out->ptr, P, CW, 128)) return -1;
    const int smem = 1024 + 16384 + NB * 8192 + NB * 16384 + 64;
    const __nv_bfloat16* w16 = reinterpret_cast<const __nv_bfloat16*>(wp);
    if (CW == 128) {
      if (set_smem(thin_tc_dgrad_kernel<128>, smem)) return -1;
      gsb::launch(thin_tc_dgrad_kernel<128>, gsb::LaunchCfg(tile_grid(num_tiles, 3), 128, smem, st), dxmap, thin_side(in), out->h, out->w,
                  g->pad, w16, P, num_tiles);
    } else {
      if (set_smem(thin_tc_dgrad_kernel<64>, smem)) return -1;
      gsb::launch(thin_tc_dgrad_kernel<64>, gsb::LaunchCfg(tile_grid(num_tiles, 4), 128, smem, st), dxmap, thin_side(in), out->h, out->w,
                  g->pad, w16, P, num_tiles);
    }
    return check_launch("thin_tc_dgrad") ? -1 : 1;
  }
  return 0;
}

// gsb_conv_wgrad hook (forward layer mode CONVT: in = wide, dout = thin).  ws == nullptr: plan only.
int thin_tc_wgrad_try(const gsb_view* in, const gsb_view* dout, float* ws, int max_slices, int* used, const gsb_conv_geom* g,
                      cudaStream_t st) {
  if (!thin_tc_enabled()) return 0;
  if (!(g->mode == GSB_MODE_CONVT && wide_side_ok(in) && thin_side_ok(dout) && thin_tc_geom_ok(in, dout, g))) return 0;
  const int64_t P = num_pixels(in);
  const int num_tiles = (int)((P + 127) / 128), CW = in->c, NB = CW / 64;
  int grid = tile_grid(num_tiles, 2);
  if (ws != nullptr && grid > max_slices) grid = max_slices;
  *used = grid;
  if (ws == nullptr) return 1;
  GSB_REQUIRE(grid >= 1, "thin_tc_wgrad: no workspace slices");
  CUtensorMap xmap;
  if (make_w_map(&xmap, in->ptr, P, CW, 128)) return -1;
  const int smem = 1024 + 32768 + NB * 16384 + 64;
  if (CW == 128) {
    if (set_smem(thin_tc_wgrad_kernel<128>, smem)) return -1;
    gsb::launch(thin_tc_wgrad_kernel<128>, gsb::LaunchCfg(grid, 128, smem, st), xmap, thin_side(dout), in->h, in->w, g->pad, ws, P, num_tiles);
  } else {
    if (set_smem(thin_tc_wgrad_kernel<64>, smem)) return -1;
    gsb::launch(thin_tc_wgrad_kernel<64>, gsb::LaunchCfg(grid, 128, smem, st), xmap, thin_side(dout), in->h, in->w, g->pad, ws, P, num_tiles);
  }
  return check_launch("thin_tc_wgrad") ? -1 : 1;
}

}  // namespace gsb

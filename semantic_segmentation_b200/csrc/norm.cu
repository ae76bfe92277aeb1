// BatchNorm2d statistics / apply / backward fused with activation and dropout.
// Replaces nn.BatchNorm2d + nn.ReLU / nn.LeakyReLU(0.2) / nn.Dropout(0.5) of
// unet/unet_parts.py:17-21 and models_pix2pix/networks.py:583-607,648-657.
//
// All kernels are HBM-bound: every thread owns 8 consecutive channels (one 128-bit access in
// bf16) of a pixel and a fixed channel group, so per-channel parameters live in registers.
// Algorithmic bytes: reduce = 1 read; act = 1 read + 1 (or 2) writes; bwd_reduce = 2 (3) reads;
// bwd_apply = 2 (3) reads + 1 write -- at the storage dtype.
#include <stdlib.h>
#include <type_traits>
#include "common.cuh"
#include "peer.cuh"

namespace gsb {

constexpr int kThreads = 256;
constexpr int kUnroll = 2;     // pixels in flight per thread (memory-level parallelism)
#ifndef GSB_BN_BWD_UNROLL
#define GSB_BN_BWD_UNROLL 1
#endif
constexpr int kUnrollBwd = GSB_BN_BWD_UNROLL;  // backward kernels read two tensors: fewer registers, more resident warps

struct PixView {  // pixel-linear view
  void* ptr;
  int64_t ld;
};

template <typename T>
__device__ __forceinline__ void load8(const PixView& v, int64_t p, int c0, float (&f)[8]) {
  Vec8<T>::load(reinterpret_cast<const T*>(v.ptr) + p * v.ld + c0, f);
}
template <typename T>
__device__ __forceinline__ void store8(const PixView& v, int64_t p, int c0, const float (&f)[8]) {
  Vec8<T>::store(reinterpret_cast<T*>(v.ptr) + p * v.ld + c0, f);
}

// per-image mode: move a pixel-linear view to image `img` (P pixels per image), once, outside the pixel loops
template <typename T>
__device__ __forceinline__ void advance_view(PixView& v, int64_t pixels) {
  if (v.ptr != nullptr) v.ptr = reinterpret_cast<T*>(v.ptr) + pixels * v.ld;
}

__device__ __forceinline__ void load_param8(const float* p, int c0, float (&f)[8], float dflt) {
  if (p == nullptr) {
#pragma unroll
    for (int i = 0; i < 8; ++i) f[i] = dflt;
  } else {
    float4 a = __ldg(reinterpret_cast<const float4*>(p + c0));
    float4 b = __ldg(reinterpret_cast<const float4*>(p + c0 + 4));
    f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w;
    f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
  }
}

// keep bits for the 8 channels of (pixel p, channel group c0); logical index = p*C + c0
// (explicit masks are indexed locally; Philox masks by the GLOBAL element index = local index + drop_offset, so that a
// batch sharded over N ranks draws the masks of the one-rank run: drop_offset = first global sample * H * W * C)
__device__ __forceinline__ uint32_t keep_bits(uint64_t seed, int64_t drop_offset, const uint8_t* mask, int64_t p, int C, int c0) {
  if (mask != nullptr) {
    const uint8_t* m = mask + p * C + c0;
    uint2 r = *reinterpret_cast<const uint2*>(m);
    uint32_t bits = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      bits |= ((r.x >> (8 * i)) & 0xFFu ? 1u : 0u) << i;
      bits |= ((r.y >> (8 * i)) & 0xFFu ? 1u : 0u) << (4 + i);
    }
    return bits;
  }
  if (seed != 0) return dropout_keep8(seed, (uint64_t)drop_offset + (uint64_t)p * (uint64_t)C + (uint64_t)c0);
  return 0xFFu;
}

// ---------------------------------------------------------------------------------------
// Per-channel reductions.  MODE 0: sum x, sum x^2.   MODE 1: sum dy, sum dy*xhat.
// Block partials -> partial[block][2][C] fp32; second kernel reduces blocks in fp64.
// ---------------------------------------------------------------------------------------
// FAST = the U-Net layer shape (act1 = ReLU, no second branch, no dropout): the activation switches, the dropout
// mask and the second gradient stream fold away at compile time (3-4x fewer instructions per 16-byte item; the
// generic kernels are issue-bound, ncu profiles/r01b).
template <typename T, int MODE, bool FAST, bool PERIMG = false>
__global__ void __launch_bounds__(kThreads) bn_reduce_kernel(
    PixView x, PixView dA1, PixView dA2, int64_t P, int C, const float* __restrict__ scale,
    const float* __restrict__ shift, const float* __restrict__ center, const float* __restrict__ mean,
    const float* __restrict__ invstd, int act1,
    int act2, uint64_t drop_seed, int64_t drop_offset, const uint8_t* __restrict__ drop_mask, float* __restrict__ partial) {
  GSB_PDL_ENTRY();
  constexpr int KU = MODE == 1 ? kUnrollBwd : kUnroll;
  // fp32 validation mode: fp64 accumulation, and every block writes TWO partial rows (hi, lo = the fp64 sum split into two
  // floats) which the fp64 second stage adds back together.  sum x^2 - (sum x)^2 / n cancels badly when a channel's mean
  // dominates its spread (BatchNorm over the 8..32 samples of the generator's deep levels): fp32 partial sums put the
  // whole network 1e-4 away from the reference there, fp64 ones 1e-6 (torch's CPU kernel is two-pass).
  constexpr bool WIDE = std::is_same<T, float>::value;
  using Acc = typename std::conditional<WIDE, double, float>::type;
  extern __shared__ double red_raw[];
  Acc* red = reinterpret_cast<Acc*>(red_raw);  // [lanes][2][C]
  const int G = C >> 3;
  const int lanes = kThreads / G;
  const int g = threadIdx.x % G, lane = threadIdx.x / G;
  const int c0 = g * 8;
  // per-image statistics (InstanceNorm, gridDim.y = images): P = pixels of ONE image, this block row works on image
  // blockIdx.y, the per-channel parameter arrays are [images][C] and a partial row is [2][images * C]
  // (PERIMG is a template parameter: the whole-batch kernels must not pay for the indexing -- an earlier run-time form cost
  // the bf16 bn_act kernel 20 % of its bandwidth through a different load schedule)
  const int img = PERIMG ? (int)blockIdx.y : 0, NC = PERIMG ? (int)gridDim.y * C : C, poff = img * C;
  const int64_t pbase = (int64_t)img * P;
  if (PERIMG && img != 0) {
    advance_view<T>(x, pbase);
    advance_view<T>(dA1, pbase);
    advance_view<T>(dA2, pbase);
  }
  Acc s0[8], s1[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) s0[i] = s1[i] = (Acc)0;
  float sc[8], sh[8], mu[8], is[8], ce[8];
  if (MODE == 1) {
    load_param8(scale ? scale + poff : nullptr, c0, sc, 1.f);
    load_param8(shift ? shift + poff : nullptr, c0, sh, 0.f);
    load_param8((FAST || !center) ? nullptr : center + poff, c0, ce, 0.f);
    load_param8(mean ? mean + poff : nullptr, c0, mu, 0.f);
    load_param8(invstd ? invstd + poff : nullptr, c0, is, 1.f);
  }
  const bool has_drop = FAST ? false : ((drop_seed != 0) || (drop_mask != nullptr));
  const bool two = FAST ? false : (dA2.ptr != nullptr);
  if (FAST) { act1 = GSB_ACT_RELU; act2 = GSB_ACT_NONE; }
  const int64_t step = (int64_t)gridDim.x * lanes;
  for (int64_t p0 = (int64_t)blockIdx.x * lanes + lane; p0 < P; p0 += step * KU) {
    float v[KU][8], a[KU][8], b[KU][8];
#pragma unroll
    for (int u = 0; u < KU; ++u)
      if (p0 + u * step < P) {
        load8<T>(x, p0 + u * step, c0, v[u]);
        if (MODE == 1) {
          load8<T>(dA1, p0 + u * step, c0, a[u]);
          if (two) load8<T>(dA2, p0 + u * step, c0, b[u]);
        }
      }
#pragma unroll
    for (int u = 0; u < KU; ++u) {
      const int64_t p = p0 + u * step;
      if (p >= P) break;
      if (MODE == 0) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          s0[i] += (Acc)v[u][i];
          s1[i] += (Acc)v[u][i] * (Acc)v[u][i];
        }
      } else {
        uint32_t keep = has_drop ? keep_bits(drop_seed, drop_offset, drop_mask, pbase + p, C, c0) : 0xFFu;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          float z = FAST ? fmaf(v[u][i], sc[i], sh[i]) : fmaf(v[u][i] - ce[i], sc[i], sh[i]);
          float dm = has_drop ? (((keep >> i) & 1u) ? 2.f : 0.f) : 1.f;
          float dy = a[u][i] * dm * act_grad(z * dm, act1);
          if (two) dy += b[u][i] * act_grad(z, act2);
          float xh = (v[u][i] - mu[i]) * is[i];
          s0[i] += (Acc)dy;
          s1[i] += (Acc)dy * (Acc)xh;
        }
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    red[(lane * 2 + 0) * C + c0 + i] = s0[i];
    red[(lane * 2 + 1) * C + c0 + i] = s1[i];
  }
  __syncthreads();
  for (int j = threadIdx.x; j < 2 * C; j += kThreads) {
    Acc acc = (Acc)0;
    for (int l = 0; l < lanes; ++l) acc += red[l * 2 * C + j];
    const int col = j < C ? poff + j : NC + poff + (j - C);
    if (WIDE) {
      const float hi = (float)acc;
      partial[((int64_t)blockIdx.x * 2) * 2 * NC + col] = hi;
      partial[((int64_t)blockIdx.x * 2 + 1) * 2 * NC + col] = (float)((double)acc - (double)hi);
    } else {
      partial[(int64_t)blockIdx.x * 2 * NC + col] = (float)acc;
    }
  }
}

// Backward statistics of the U-Net layer shape (bf16, ReLU, one gradient stream, no dropout): sum dy, sum dy*xhat with
// dy = dA * [scale*y + shift > 0].  The generic kernel held 76 registers (3 blocks / SM) with ONE 16-byte load pair in
// flight per thread and sat at 69 % of HBM bandwidth stalled on the long scoreboard (ncu profiles/r01n).  Here the
// loads stay packed (uint4) until they are consumed, KU pixel rows are in flight per thread, and the normalisation is
// applied after the loop:  sum dy*xhat = invstd * (sum dy*y - mean * sum dy)  (mean / invstd leave the loop; the
// per-thread sums cover ~100 pixels of bf16 data, so the subtraction costs nothing measurable in accuracy).
template <int KU>
__global__ void __launch_bounds__(kThreads, 4) bn_bwd_reduce_relu_bf16_kernel(
    const __nv_bfloat16* __restrict__ x, int64_t xld, const __nv_bfloat16* __restrict__ dA, int64_t dld, int64_t P, int C,
    const float* __restrict__ scale, const float* __restrict__ shift, const float* __restrict__ mean,
    const float* __restrict__ invstd, float* __restrict__ partial) {
  GSB_PDL_ENTRY();
  extern __shared__ float red[];  // [lanes][2][C]
  const int G = C >> 3;
  const int lanes = kThreads / G;
  const int g = threadIdx.x % G, lane = threadIdx.x / G;
  const int c0 = g * 8;
  float s0[8], s1[8], sc[8], sh[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) s0[i] = s1[i] = 0.f;
  load_param8(scale, c0, sc, 1.f);
  load_param8(shift, c0, sh, 0.f);
  const int64_t step = (int64_t)gridDim.x * lanes;
  const __nv_bfloat16* xp = x + c0;
  const __nv_bfloat16* dp = dA + c0;
  for (int64_t p0 = (int64_t)blockIdx.x * lanes + lane; p0 < P; p0 += step * KU) {
    uint4 rv[KU], ra[KU];
#pragma unroll
    for (int u = 0; u < KU; ++u) {
      const int64_t p = p0 + u * step;
      if (p < P) {
        rv[u] = *reinterpret_cast<const uint4*>(xp + p * xld);
        ra[u] = *reinterpret_cast<const uint4*>(dp + p * dld);
      } else {
        rv[u] = make_uint4(0, 0, 0, 0);
        ra[u] = make_uint4(0, 0, 0, 0);  // dA = 0 contributes nothing
      }
    }
#pragma unroll
    for (int u = 0; u < KU; ++u) {
      const uint32_t* wv = reinterpret_cast<const uint32_t*>(&rv[u]);
      const uint32_t* wa = reinterpret_cast<const uint32_t*>(&ra[u]);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float v0 = __uint_as_float(wv[i] << 16), v1 = __uint_as_float(wv[i] & 0xffff0000u);
        const float a0 = __uint_as_float(wa[i] << 16), a1 = __uint_as_float(wa[i] & 0xffff0000u);
        const float d0 = fmaf(v0, sc[2 * i], sh[2 * i]) > 0.f ? a0 : 0.f;
        const float d1 = fmaf(v1, sc[2 * i + 1], sh[2 * i + 1]) > 0.f ? a1 : 0.f;
        s0[2 * i] += d0;
        s0[2 * i + 1] += d1;
        s1[2 * i] = fmaf(d0, v0, s1[2 * i]);
        s1[2 * i + 1] = fmaf(d1, v1, s1[2 * i + 1]);
      }
    }
  }
  {
    float mu[8], is[8];
    load_param8(mean, c0, mu, 0.f);
    load_param8(invstd, c0, is, 1.f);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      red[(lane * 2 + 0) * C + c0 + i] = s0[i];
      red[(lane * 2 + 1) * C + c0 + i] = is[i] * fmaf(-mu[i], s0[i], s1[i]);
    }
  }
  __syncthreads();
  for (int j = threadIdx.x; j < 2 * C; j += kThreads) {
    float acc = 0.f;
    for (int l = 0; l < lanes; ++l) acc += red[l * 2 * C + j];
    partial[(int64_t)blockIdx.x * 2 * C + j] = acc;
  }
}

// The same statistics with the MaxPool2d(2) backward folded in (the encoder's Down -> DoubleConv seam in the backward
// direction): the gradient of the block's output is  dA = dskip + route(dpool)  where dskip already sits in the concat
// gradient (written by the decoder's data gradient) and route() sends each pooled gradient to the FIRST maximum of its
// 2x2 window.  A thread owns one window of 8 channels: it recomputes the stored activations a = bf16(relu(scale*y+shift))
// (exactly what bn_act wrote, so the argmax is the one maxpool2_bwd finds), adds the routed gradient, WRITES dA back
// (rounded to bf16 as maxpool2_bwd's accumulate does: bn_bwd_apply reads it next) and accumulates sum dy, sum dy*y.
// Replaces maxpool2_bwd (reads a, dpool, dskip; writes dskip) + this kernel's two reads by one pass: 2 tensor reads
// less per encoder level.  H, W even, kThreads % (C/8) == 0.
__global__ void __launch_bounds__(kThreads, 2) bn_bwd_reduce_relu_pool_bf16_kernel(
    const __nv_bfloat16* __restrict__ x, int64_t xld, __nv_bfloat16* __restrict__ dA, int64_t dld,
    const __nv_bfloat16* __restrict__ dpool, int64_t pld, int N, int H, int W, int C, const float* __restrict__ scale,
    const float* __restrict__ shift, const float* __restrict__ mean, const float* __restrict__ invstd,
    float* __restrict__ partial) {
  GSB_PDL_ENTRY();
  extern __shared__ float red[];  // [lanes][2][C]
  const int G = C >> 3;
  const int lanes = kThreads / G;
  const int g = threadIdx.x % G, lane = threadIdx.x / G;
  const int c0 = g * 8;
  float s0[8], s1[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) s0[i] = s1[i] = 0.f;
  const unsigned OW = (unsigned)W >> 1, OH = (unsigned)H >> 1;
  const int64_t windows = (int64_t)N * OH * OW;
  const int64_t step = (int64_t)gridDim.x * lanes;
  for (int64_t wi = (int64_t)blockIdx.x * lanes + lane; wi < windows; wi += step) {
    unsigned ox, oy, n;
    if ((wi >> 32) == 0) {
      const unsigned u = (unsigned)wi, row = u / OW;
      ox = u - row * OW;
      n = row / OH;
      oy = row - n * OH;
    } else {
      const int64_t row = wi / OW;
      ox = (unsigned)(wi - row * OW);
      n = (unsigned)(row / OH);
      oy = (unsigned)(row - (int64_t)n * OH);
    }
    const int64_t p00 = ((int64_t)n * H + 2 * oy) * W + 2 * ox;
    uint4 rv[4], ra[4];
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int64_t p = p00 + (t >> 1) * (int64_t)W + (t & 1);
      rv[t] = *reinterpret_cast<const uint4*>(x + p * xld + c0);
      ra[t] = *reinterpret_cast<const uint4*>(dA + p * dld + c0);
    }
    const uint4 rp = *reinterpret_cast<const uint4*>(dpool + wi * pld + c0);
    const uint32_t wp[4] = {rp.x, rp.y, rp.z, rp.w};
    // the routed gradient replaces the loaded one IN PLACE (ra[t] word i), so no second set of 16 registers is live:
    // (3 blocks per SM at 80 registers + spills measured the same 0.61 of the HBM roofline as 2 blocks: r02z_full_new.raw.csv)
    uint32_t aw[4][4];
#pragma unroll
    for (int t = 0; t < 4; ++t) { aw[t][0] = ra[t].x; aw[t][1] = ra[t].y; aw[t][2] = ra[t].z; aw[t][3] = ra[t].w; }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      // scale / shift of this channel pair from L1 (16 registers less than keeping all eight pairs)
      const float2 sc2 = __ldg(reinterpret_cast<const float2*>(scale + c0 + 2 * i));
      const float2 sh2 = __ldg(reinterpret_cast<const float2*>(shift + c0 + 2 * i));
      float v0[4], v1[4], a0[4], a1[4];
      int arg0 = 0, arg1 = 0;
      float m0 = 0.f, m1 = 0.f;
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const uint32_t wv = reinterpret_cast<const uint32_t*>(&rv[t])[i];
        v0[t] = __uint_as_float(wv << 16);
        v1[t] = __uint_as_float(wv & 0xffff0000u);
        // stored activation (bf16, round to nearest even) as bn_act_kernel<bf16, true> writes it
        const float z0 = fmaf(v0[t], sc2.x, sh2.x), z1 = fmaf(v1[t], sc2.y, sh2.y);
        a0[t] = __bfloat162float(__float2bfloat16_rn(z0 > 0.f ? z0 : 0.f));
        a1[t] = __bfloat162float(__float2bfloat16_rn(z1 > 0.f ? z1 : 0.f));
        if (t == 0) { m0 = a0[0]; m1 = a1[0]; }
        else {
          if (a0[t] > m0) { m0 = a0[t]; arg0 = t; }   // first maximum wins (maxpool2_bwd_kernel; no NaN after the ReLU)
          if (a1[t] > m1) { m1 = a1[t]; arg1 = t; }
        }
      }
      const float g0 = __uint_as_float(wp[i] << 16), g1 = __uint_as_float(wp[i] & 0xffff0000u);
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const uint32_t wa = aw[t][i];
        float d0 = __uint_as_float(wa << 16), d1 = __uint_as_float(wa & 0xffff0000u);
        d0 += arg0 == t ? g0 : 0.f;
        d1 += arg1 == t ? g1 : 0.f;
        const __nv_bfloat162 h = __floats2bfloat162_rn(d0, d1);
        aw[t][i] = *reinterpret_cast<const uint32_t*>(&h);
        d0 = __low2float(h);
        d1 = __high2float(h);
        const float e0 = a0[t] > 0.f ? d0 : 0.f, e1 = a1[t] > 0.f ? d1 : 0.f;   // z > 0 <=> stored a > 0
        s0[2 * i] += e0;
        s0[2 * i + 1] += e1;
        s1[2 * i] = fmaf(e0, v0[t], s1[2 * i]);
        s1[2 * i + 1] = fmaf(e1, v1[t], s1[2 * i + 1]);
      }
    }
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int64_t p = p00 + (t >> 1) * (int64_t)W + (t & 1);
      *reinterpret_cast<uint4*>(dA + p * dld + c0) = make_uint4(aw[t][0], aw[t][1], aw[t][2], aw[t][3]);
    }
  }
  {
    float mu[8], is[8];
    load_param8(mean, c0, mu, 0.f);
    load_param8(invstd, c0, is, 1.f);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      red[(lane * 2 + 0) * C + c0 + i] = s0[i];
      red[(lane * 2 + 1) * C + c0 + i] = is[i] * fmaf(-mu[i], s0[i], s1[i]);
    }
  }
  __syncthreads();
  for (int j = threadIdx.x; j < 2 * C; j += kThreads) {
    float acc = 0.f;
    for (int l = 0; l < lanes; ++l) acc += red[l * 2 * C + j];
    partial[(int64_t)blockIdx.x * 2 * C + j] = acc;
  }
}

// partial[rows][n] fp32 -> sums[n] fp64.  Block = 32 columns x 32 row lanes: coalesced 128-byte row segments,
// four independent loads in flight per thread, fp64 accumulation, fixed summation order (deterministic).
__global__ void __launch_bounds__(1024) reduce_partials_kernel(const float* __restrict__ partial, int rows, int n,
                                                              double* __restrict__ sums) {
  GSB_PDL_ENTRY();
  __shared__ double red[32][33];
  const int c = threadIdx.x & 31, l = threadIdx.x >> 5;
  const int col = blockIdx.x * 32 + c;
  double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
  if (col < n) {
    int r = l;
    for (; r + 96 < rows; r += 128) {
      const float v0 = partial[(int64_t)r * n + col], v1 = partial[(int64_t)(r + 32) * n + col];
      const float v2 = partial[(int64_t)(r + 64) * n + col], v3 = partial[(int64_t)(r + 96) * n + col];
      a0 += (double)v0; a1 += (double)v1; a2 += (double)v2; a3 += (double)v3;
    }
    for (; r < rows; r += 32) a0 += (double)partial[(int64_t)r * n + col];
  }
  red[l][c] = (a0 + a1) + (a2 + a3);
  __syncthreads();
  if (l == 0 && col < n) {
    double t = 0.0;
#pragma unroll
    for (int i = 0; i < 32; ++i) t += red[i][c];
    sums[col] = t;
  }
}

__global__ void bn_finalize_kernel(const double* __restrict__ sums, double count, const float* __restrict__ gamma,
                                   const float* __restrict__ beta, float* running_mean, float* running_var,
                                   int64_t* nbt, float momentum, float eps, int C, float* scale, float* shift,
                                   float* mean_out, float* invstd_out, int centered) {
  GSB_PDL_ENTRY();
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c == 0 && nbt != nullptr && sums != nullptr) *nbt += 1;
  if (c >= C) return;
  float g = gamma ? gamma[c] : 1.f, b = beta ? beta[c] : 0.f;
  float m, is;
  if (sums != nullptr) {
    double mu = sums[c] / count;
    double var = sums[C + c] / count - mu * mu;
    if (var < 0.0) var = 0.0;
    m = (float)mu;
    is = (float)(1.0 / sqrt(var + (double)eps));
    if (running_mean) running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * m;
    if (running_var) {
      double unbiased = count > 1.0 ? var * count / (count - 1.0) : var;
      running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)unbiased;
    }
  } else {
    m = running_mean[c];
    is = 1.f / sqrtf(running_var[c] + eps);
  }
  float sc = g * is;
  scale[c] = sc;
  // centered (fp32 validation mode): the consumers evaluate (x - mean) * scale + beta like torch does; the folded form
  // x * scale + (beta - mean * scale) loses |mean| / std digits to cancellation (channels of the generator's deep levels
  // are normalised over 8..32 samples), which put the fp32 mode 1e-4 away from the reference and flipped ReLU branches
  shift[c] = centered ? b : b - m * sc;
  if (mean_out) mean_out[c] = m;
  if (invstd_out) invstd_out[c] = is;
}

template <typename T, bool FAST, bool PERIMG = false>
__global__ void __launch_bounds__(kThreads) bn_act_kernel(PixView x, int64_t P, int C, const float* __restrict__ scale,
                                                         const float* __restrict__ shift, const float* __restrict__ center,
                                                         int act1, PixView out1,
                                                         int act2, PixView out2, uint64_t drop_seed, int64_t drop_offset,
                                                         const uint8_t* __restrict__ drop_mask) {
  GSB_PDL_ENTRY();
  const int G = C >> 3;
  const int lanes = kThreads / G;
  const int g = threadIdx.x % G, lane = threadIdx.x / G;
  const int c0 = g * 8;
  const int poff = PERIMG ? (int)blockIdx.y * C : 0;   // per-image parameters (InstanceNorm): gridDim.y = images, P per image
  const int64_t pbase = PERIMG ? (int64_t)blockIdx.y * P : 0;
  if (PERIMG && blockIdx.y != 0) {
    advance_view<T>(x, pbase);
    advance_view<T>(out1, pbase);
    advance_view<T>(out2, pbase);
  }
  float sc[8], sh[8], ce[8];
  load_param8(scale ? scale + poff : nullptr, c0, sc, 1.f);
  load_param8(shift ? shift + poff : nullptr, c0, sh, 0.f);
  load_param8((FAST || !center) ? nullptr : center + poff, c0, ce, 0.f);   // z = (x - center) * scale + shift
  const bool has_drop = FAST ? false : ((drop_seed != 0) || (drop_mask != nullptr));
  const bool has_out2 = FAST ? false : (out2.ptr != nullptr);
  if (FAST) { act1 = GSB_ACT_RELU; act2 = GSB_ACT_NONE; }
  const int64_t step = (int64_t)gridDim.x * lanes;
  for (int64_t p0 = (int64_t)blockIdx.x * lanes + lane; p0 < P; p0 += step * kUnroll) {
    float v[kUnroll][8];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u)
      if (p0 + u * step < P) load8<T>(x, p0 + u * step, c0, v[u]);
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const int64_t p = p0 + u * step;
      if (p >= P) break;
      float o1[8], o2[8];
      uint32_t keep = has_drop ? keep_bits(drop_seed, drop_offset, drop_mask, pbase + p, C, c0) : 0xFFu;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float z = FAST ? fmaf(v[u][i], sc[i], sh[i]) : fmaf(v[u][i] - ce[i], sc[i], sh[i]);
        float dm = has_drop ? (((keep >> i) & 1u) ? 2.f : 0.f) : 1.f;
        o1[i] = act_fwd(z * dm, act1);
        if (has_out2) o2[i] = act_fwd(z, act2);
      }
      store8<T>(out1, p, c0, o1);
      if (has_out2) store8<T>(out2, p, c0, o2);
    }
  }
}

// BatchNorm apply + ReLU + MaxPool2d(2) in one pass (the encoder's `DoubleConv -> Down` seam, unet_parts.py:21,34): a thread
// owns one 2x2 window of 8 channels, writes the four activations (the skip tensor, possibly a channel slice of a concat
// buffer) and their maximum (the pooled tensor), so the pooling pass does not read the activation back (1 read + 1/4
// write of the tensor saved).  Same arithmetic as bn_act_kernel<T, true> followed by maxpool2_fwd_kernel: rounding to
// the storage type is monotonic, so round(max) == max(round) and the pooled values are bit-identical.  H, W even.
template <typename T>
__global__ void __launch_bounds__(kThreads) bn_act_pool_kernel(PixView x, int N, int H, int W, int C, const float* __restrict__ scale,
                                                              const float* __restrict__ shift, PixView out1, PixView pooled) {
  GSB_PDL_ENTRY();
  const unsigned G = (unsigned)C >> 3, OW = (unsigned)W >> 1, OH = (unsigned)H >> 1;
  const uint64_t total = (uint64_t)N * OH * OW * G;
  const bool fixed_g = (kThreads % G) == 0;   // the grid stride is a multiple of G: a thread keeps its channel group
  float sc[8], sh[8];
  if (fixed_g) {
    load_param8(scale, (int)(threadIdx.x % G) * 8, sc, 1.f);
    load_param8(shift, (int)(threadIdx.x % G) * 8, sh, 0.f);
  }
  for (uint64_t i = (uint64_t)blockIdx.x * kThreads + threadIdx.x; i < total; i += (uint64_t)gridDim.x * kThreads) {
    unsigned g, ox, oy, n;
    if ((i >> 32) == 0) {
      const unsigned u = (unsigned)i, q = u / G;
      g = u - q * G;
      const unsigned row = q / OW;
      ox = q - row * OW;
      n = row / OH;
      oy = row - n * OH;
    } else {
      const uint64_t q = i / G;
      g = (unsigned)(i - q * G);
      const uint64_t row = q / OW;
      ox = (unsigned)(q - row * OW);
      n = (unsigned)(row / OH);
      oy = (unsigned)(row - (uint64_t)n * OH);
    }
    const int c0 = (int)g * 8;
    if (!fixed_g) {
      load_param8(scale, c0, sc, 1.f);
      load_param8(shift, c0, sh, 0.f);
    }
    const int64_t p00 = ((int64_t)n * H + 2 * oy) * W + 2 * ox;
    float v[4][8];
#pragma unroll
    for (int t = 0; t < 4; ++t) load8<T>(x, p00 + (t >> 1) * (int64_t)W + (t & 1), c0, v[t]);
    float m[8];
#pragma unroll
    for (int t = 0; t < 4; ++t) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float z = fmaf(v[t][k], sc[k], sh[k]);
        v[t][k] = z > 0.f ? z : 0.f;
        m[k] = t == 0 ? v[t][k] : fmaxf(m[k], v[t][k]);
      }
      store8<T>(out1, p00 + (t >> 1) * (int64_t)W + (t & 1), c0, v[t]);
    }
    store8<T>(pooled, ((int64_t)n * OH + oy) * OW + ox, c0, m);
  }
}

template <typename T, bool FAST, bool PERIMG = false>
__global__ void __launch_bounds__(kThreads) bn_bwd_apply_kernel(
    PixView x, PixView dA1, PixView dA2, int64_t P, int C, const float* __restrict__ scale,
    const float* __restrict__ shift, const float* __restrict__ center, const float* __restrict__ mean,
    const float* __restrict__ invstd, int act1,
    int act2, uint64_t drop_seed, int64_t drop_offset, const uint8_t* __restrict__ drop_mask, const double* __restrict__ sums,
    double inv_count, PixView dx) {
  GSB_PDL_ENTRY();
  const int G = C >> 3;
  const int lanes = kThreads / G;
  const int g = threadIdx.x % G, lane = threadIdx.x / G;
  const int c0 = g * 8;
  const int NC = PERIMG ? (int)gridDim.y * C : C, poff = PERIMG ? (int)blockIdx.y * C : 0;   // per-image parameters / sums
  const int64_t pbase = PERIMG ? (int64_t)blockIdx.y * P : 0;
  if (PERIMG && blockIdx.y != 0) {
    advance_view<T>(x, pbase);
    advance_view<T>(dA1, pbase);
    advance_view<T>(dA2, pbase);
    advance_view<T>(dx, pbase);
  }
  float sc[8], sh[8], mu[8], is[8], m0[8], m1[8], ce[8];
  load_param8(scale ? scale + poff : nullptr, c0, sc, 1.f);
  load_param8(shift ? shift + poff : nullptr, c0, sh, 0.f);
  load_param8((FAST || !center) ? nullptr : center + poff, c0, ce, 0.f);
  const bool has_bn = FAST ? true : (mean != nullptr);
  load_param8(mean ? mean + poff : nullptr, c0, mu, 0.f);
  load_param8(invstd ? invstd + poff : nullptr, c0, is, 1.f);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    m0[i] = has_bn ? (float)(sums[poff + c0 + i] * inv_count) : 0.f;
    m1[i] = has_bn ? (float)(sums[NC + poff + c0 + i] * inv_count) : 0.f;
  }
  const bool has_drop = FAST ? false : ((drop_seed != 0) || (drop_mask != nullptr));
  const bool two = FAST ? false : (dA2.ptr != nullptr);
  if (FAST) { act1 = GSB_ACT_RELU; act2 = GSB_ACT_NONE; }
  const int64_t step = (int64_t)gridDim.x * lanes;
  if (FAST) {
    // dx = sc*(dy - m0 - xhat*m1) = sc*dy + A*x + B with A = -sc*m1*invstd, B = -sc*m0 - A*mean: four parameter
    // vectors in registers instead of six, three FMAs per element
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float A = -sc[i] * m1[i] * is[i];
      m1[i] = A;
      m0[i] = -sc[i] * m0[i] - A * mu[i];
    }
    for (int64_t p0 = (int64_t)blockIdx.x * lanes + lane; p0 < P; p0 += step * kUnrollBwd) {
      float v[kUnrollBwd][8], a[kUnrollBwd][8];
#pragma unroll
      for (int u = 0; u < kUnrollBwd; ++u)
        if (p0 + u * step < P) {
          load8<T>(x, p0 + u * step, c0, v[u]);
          load8<T>(dA1, p0 + u * step, c0, a[u]);
        }
#pragma unroll
      for (int u = 0; u < kUnrollBwd; ++u) {
        const int64_t p = p0 + u * step;
        if (p >= P) break;
        float o[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float z = fmaf(v[u][i], sc[i], sh[i]);
          const float dy = z > 0.f ? a[u][i] : 0.f;
          o[i] = fmaf(sc[i], dy, fmaf(m1[i], v[u][i], m0[i]));
        }
        store8<T>(dx, p, c0, o);
      }
    }
    return;
  }
  for (int64_t p0 = (int64_t)blockIdx.x * lanes + lane; p0 < P; p0 += step * kUnrollBwd) {
    float v[kUnrollBwd][8], a[kUnrollBwd][8], b[kUnrollBwd][8];
#pragma unroll
    for (int u = 0; u < kUnrollBwd; ++u)
      if (p0 + u * step < P) {
        load8<T>(x, p0 + u * step, c0, v[u]);
        load8<T>(dA1, p0 + u * step, c0, a[u]);
        if (two) load8<T>(dA2, p0 + u * step, c0, b[u]);
      }
#pragma unroll
    for (int u = 0; u < kUnrollBwd; ++u) {
      const int64_t p = p0 + u * step;
      if (p >= P) break;
      float o[8];
      uint32_t keep = has_drop ? keep_bits(drop_seed, drop_offset, drop_mask, pbase + p, C, c0) : 0xFFu;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float z = fmaf(v[u][i] - ce[i], sc[i], sh[i]);
        float dm = has_drop ? (((keep >> i) & 1u) ? 2.f : 0.f) : 1.f;
        float dy = a[u][i] * dm * act_grad(z * dm, act1);
        if (two) dy += b[u][i] * act_grad(z, act2);
        if (has_bn) {
          float xh = (v[u][i] - mu[i]) * is[i];
          o[i] = sc[i] * (dy - m0[i] - xh * m1[i]);
        } else {
          o[i] = dy * sc[i];
        }
      }
      store8<T>(dx, p, c0, o);
    }
  }
}

__global__ void bn_param_grad_kernel(const double* __restrict__ sums, int C, float* dgamma, float* dbeta, int accumulate = 1) {
  GSB_PDL_ENTRY();
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  if (dbeta) dbeta[c] = (accumulate ? dbeta[c] : 0.f) + (float)sums[c];
  if (dgamma) dgamma[c] = (accumulate ? dgamma[c] : 0.f) + (float)sums[C + c];
}

// Statistic rows -> per-channel totals -> (cross-rank exchange) -> consumer, in ONE launch (was reduce_partials +
// all-reduce + bn_finalize forward, reduce_partials + bn_param_grad + all-reduce backward: 5 launches per BatchNorm layer
// and direction on the critical path of the step).  Block b owns channels [16b, 16b+16): 32 columns (16 x sum, 16 x
// second sum) x 32 row lanes, fp64, fixed order.  world > 1: the 32 totals are exchanged through the peer regions
// (csrc/peer.cu protocol, element index = column); every block reads the call's sequence number, the last block to
// finish bumps the counter.
//   MODE 0 (forward):  totals = [sum x, sum x^2] -> scale / shift / mean / invstd, running statistics, num_batches_tracked
//   MODE 1 (backward): totals = [sum dy, sum dy*xhat] -> dbeta / dgamma from the LOCAL totals (the gradient all-reduce
//                      adds the other ranks), global totals -> sums_out for bn_bwd_apply
struct BnFusedFwd {
  double count;
  const float* gamma;
  const float* beta;
  float* running_mean;
  float* running_var;
  int64_t* nbt;
  float momentum, eps;
  float* scale;
  float* shift;
  float* mean;
  float* invstd;
  int centered;   // shift = beta, consumers evaluate (x - mean) * scale + beta (fp32 validation mode)
};
struct BnFusedBwd {
  float* dgamma;
  float* dbeta;
  int accumulate;
  double* sums;
};

template <int MODE>
__global__ void __launch_bounds__(1024) bn_stats_fused_kernel(const float* __restrict__ partial, int rows, int C, PeerRegions regs,
                                                             int rank, int world, BnFusedFwd f, BnFusedBwd b,
                                                             long long timeout_ns) {
  GSB_PDL_ENTRY();
  __shared__ double red[32][33];
  __shared__ double tot[32];
  const int c = threadIdx.x & 31, l = threadIdx.x >> 5;
  const int ch = blockIdx.x * 16 + (c & 15);
  const int col = (c < 16 ? 0 : C) + ch;
  const bool ok = ch < C;
  const int n = 2 * C;
  double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
  if (ok) {
    int r = l;
    for (; r + 96 < rows; r += 128) {
      const float v0 = partial[(int64_t)r * n + col], v1 = partial[(int64_t)(r + 32) * n + col];
      const float v2 = partial[(int64_t)(r + 64) * n + col], v3 = partial[(int64_t)(r + 96) * n + col];
      a0 += (double)v0; a1 += (double)v1; a2 += (double)v2; a3 += (double)v3;
    }
    for (; r < rows; r += 32) a0 += (double)partial[(int64_t)r * n + col];
  }
  red[l][c] = (a0 + a1) + (a2 + a3);
  __syncthreads();
  PeerHeader* hdr = world > 1 ? reinterpret_cast<PeerHeader*>(regs.r[rank]) : nullptr;
  unsigned long long seq = 0;
  if (l == 0) {
    double t = 0.0;
#pragma unroll
    for (int i = 0; i < 32; ++i) t += red[i][c];
    if (MODE == 1 && ok) {
      float* dst = c < 16 ? b.dbeta : b.dgamma;
      if (dst != nullptr) dst[ch] = (b.accumulate ? dst[ch] : 0.f) + (float)t;
    }
    if (world > 1) {
      seq = *reinterpret_cast<volatile unsigned long long*>(&hdr->counter) + 1ull;
      if (ok && !peer_exchange_sum(t, col, regs, rank, world, (uint32_t)seq, (int)(seq % kPeerSlots), timeout_ns))
        hdr->error = seq;   // watchdog expired: local statistics are used, the host raises after gsb_peer_error()
    }
    tot[c] = t;
    if (MODE == 1 && ok && b.sums != nullptr) b.sums[col] = t;
  }
  __syncthreads();
  if (MODE == 0) {
    if (threadIdx.x < 16 && blockIdx.x * 16 + threadIdx.x < C) {
      const int cc = blockIdx.x * 16 + threadIdx.x;
      const float g = f.gamma ? f.gamma[cc] : 1.f, be = f.beta ? f.beta[cc] : 0.f;
      const double mu = tot[threadIdx.x] / f.count;
      double var = tot[16 + threadIdx.x] / f.count - mu * mu;
      if (var < 0.0) var = 0.0;
      const float m = (float)mu;
      const float is = (float)(1.0 / sqrt(var + (double)f.eps));
      if (f.running_mean) f.running_mean[cc] = (1.f - f.momentum) * f.running_mean[cc] + f.momentum * m;
      if (f.running_var) {
        const double unbiased = f.count > 1.0 ? var * f.count / (f.count - 1.0) : var;
        f.running_var[cc] = (1.f - f.momentum) * f.running_var[cc] + f.momentum * (float)unbiased;
      }
      const float sc = g * is;
      f.scale[cc] = sc;
      f.shift[cc] = f.centered ? be : be - m * sc;
      if (f.mean) f.mean[cc] = m;
      if (f.invstd) f.invstd[cc] = is;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0 && f.nbt != nullptr) *f.nbt += 1;
  }
  if (world > 1 && threadIdx.x == 0) {  // lane 0 of warp 0 holds seq; all of this block's packets have been consumed
    __threadfence();
    const unsigned int prev = atomicAdd(&hdr->done, 1u);
    if (prev == gridDim.x - 1) {
      hdr->done = 0;
      hdr->counter = seq;
    }
  }
}

static bool relu_stats_fast_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("GSB_BN_BWD_FAST2");
    v = (e != nullptr && e[0] == '0') ? 0 : 1;
  }
  return v != 0;
}

static int grid_for(int64_t P, int lanes) {
  int64_t want = (P + lanes - 1) / lanes;
  int64_t cap = (int64_t)num_sms() * 8;
  return (int)(want < cap ? (want > 0 ? want : 1) : cap);
}

static int check_vec_view(const gsb_view* v, const char* name) {
  GSB_REQUIRE(v != nullptr && v->ptr != nullptr, "%s: null view", name);
  GSB_REQUIRE(vec8_ok(v), "%s: view must be pixel-linear NHWC with c%%8==0, ld%%8==0, 16B-aligned", name);
  GSB_REQUIRE(v->c >= 8 && kThreads % (v->c / 8) == 0, "%s: channel count %d unsupported (c/8 must divide 256)",
              name, v->c);
  return 0;
}
static PixView pv(const gsb_view* v) { return v ? PixView{v->ptr, v->sw} : PixView{nullptr, 0}; }

}  // namespace gsb

using namespace gsb;

extern "C" size_t gsb_bn_workspace_bytes(const gsb_view* x) {
  return (size_t)num_sms() * 8 * 2 * 2 * (size_t)x->c * sizeof(float);   // fp32 views write two rows (hi, lo) per block
}

// per_image != 0: statistics per (image, channel) -- InstanceNorm -- in the same launches: sums[2][n * c], blocks (x, image)
static int per_image_grid(int64_t P_img, int lanes, int n) {
  int g = grid_for(P_img, lanes);
  const int cap = (num_sms() * 8) / (n > 0 ? n : 1);     // so that rows x 2 x n x c still fits gsb_bn_workspace_bytes
  if (g > cap) g = cap;
  return g < 1 ? 1 : g;
}

extern "C" int gsb_bn_reduce(const gsb_view* x, int per_image, double* sums, void* workspace, size_t workspace_bytes,
                             gsb_stream s) {
  if (check_vec_view(x, "bn_reduce.x")) return -1;
  const int C = x->c, lanes = kThreads / (C / 8);
  const int nimg = per_image ? x->n : 1;
  const int64_t P = num_pixels(x) / nimg;
  int grid = per_image ? per_image_grid(P, lanes, nimg) : grid_for(P, lanes);
  const int wide = x->dtype == GSB_BF16 ? 1 : 2;
  GSB_REQUIRE(workspace_bytes >= (size_t)grid * wide * 2 * C * nimg * sizeof(float), "bn_reduce: workspace too small");
  size_t smem = (size_t)lanes * 2 * C * sizeof(float) * wide;
  float* partial = reinterpret_cast<float*>(workspace);
  PixView nul{nullptr, 0};
#define GSB_BN_REDUCE0(TT, PI)                                                                                              \
  do {                                                                                                                      \
    if (smem > 48 * 1024) cudaFuncSetAttribute(bn_reduce_kernel<TT, 0, false, PI>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
    gsb::launch(bn_reduce_kernel<TT, 0, false, PI>, gsb::LaunchCfg(dim3(grid, nimg), kThreads, smem, to_stream(s)), pv(x), nul, nul, P, C, \
                nullptr, nullptr, nullptr, nullptr, nullptr, 0, 0, 0, 0, nullptr, partial);                                  \
  } while (0)
  if (x->dtype == GSB_BF16) {
    if (per_image) GSB_BN_REDUCE0(__nv_bfloat16, true); else GSB_BN_REDUCE0(__nv_bfloat16, false);
  } else {
    if (per_image) GSB_BN_REDUCE0(float, true); else GSB_BN_REDUCE0(float, false);
  }
#undef GSB_BN_REDUCE0
  int n = 2 * C * nimg;
  gsb::launch(reduce_partials_kernel, gsb::LaunchCfg((n + 31) / 32, 1024, 0, to_stream(s)), partial, grid * wide, n, sums);
  return check_launch("bn_reduce", 2);
}

extern "C" int gsb_bn_reduce_partials(const float* partial, int rows, int c, double* sums, gsb_stream s) {
  int n = 2 * c;
  gsb::launch(reduce_partials_kernel, gsb::LaunchCfg((n + 31) / 32, 1024, 0, to_stream(s)), partial, rows, n, sums);
  return check_launch("bn_reduce_partials");
}

extern "C" int gsb_bn_finalize(const double* sums, double count, const float* gamma, const float* beta,
                               float* running_mean, float* running_var, int64_t* num_batches_tracked, float momentum,
                               float eps, int c, float* scale, float* shift, float* mean, float* invstd, int centered,
                               gsb_stream s) {
  GSB_REQUIRE(scale && shift, "bn_finalize: scale/shift required");
  GSB_REQUIRE(!centered || mean != nullptr, "bn_finalize: the centered form needs the mean output");
  GSB_REQUIRE(sums != nullptr || (running_mean && running_var), "bn_finalize: eval mode needs running stats");
  gsb::launch(bn_finalize_kernel, gsb::LaunchCfg((c + 127) / 128, 128, 0, to_stream(s)), sums, count, gamma, beta, running_mean, running_var,
                                                               num_batches_tracked, momentum, eps, c, scale, shift,
                                                               mean, invstd, centered);
  return check_launch("bn_finalize");
}

extern "C" int gsb_bn_act(const gsb_view* x, const float* scale, const float* shift, const float* center, int act1,
                          const gsb_view* out1,
                          int act2, const gsb_view* out2, uint64_t drop_seed, int64_t drop_offset, const uint8_t* drop_mask,
                          int per_image, gsb_stream s) {
  if (check_vec_view(x, "bn_act.x") || check_vec_view(out1, "bn_act.out1")) return -1;
  if (out2 && out2->ptr && check_vec_view(out2, "bn_act.out2")) return -1;
  GSB_REQUIRE(out1->dtype == x->dtype && out1->c == x->c && num_pixels(out1) == num_pixels(x), "bn_act: out1 mismatch");
  if (out2 && out2->ptr)
    GSB_REQUIRE(out2->dtype == x->dtype && out2->c == x->c && num_pixels(out2) == num_pixels(x), "bn_act: out2 mismatch");
  const int C = x->c, lanes = kThreads / (C / 8);
  const int nimg = per_image ? x->n : 1;     // per_image: scale / shift / center are [n][c] (InstanceNorm)
  const int64_t P = num_pixels(x) / nimg;
  dim3 grid(grid_for(P, lanes * 4), nimg);
  PixView o2 = (out2 && out2->ptr) ? pv(out2) : PixView{nullptr, 0};
  const bool fast = act1 == GSB_ACT_RELU && o2.ptr == nullptr && drop_seed == 0 && drop_mask == nullptr && scale && shift &&
                    center == nullptr && !per_image;
  if (x->dtype == GSB_BF16) {
    if (fast)
      gsb::launch(bn_act_kernel<__nv_bfloat16, true>, gsb::LaunchCfg(grid, kThreads, 0, to_stream(s)), pv(x), P, C, scale, shift, center, act1, pv(out1), act2, o2,
                                                                              drop_seed, drop_offset, drop_mask);
    else if (per_image)
      gsb::launch(bn_act_kernel<__nv_bfloat16, false, true>, gsb::LaunchCfg(grid, kThreads, 0, to_stream(s)), pv(x), P, C, scale, shift, center, act1,
                  pv(out1), act2, o2, drop_seed, drop_offset, drop_mask);
    else
      gsb::launch(bn_act_kernel<__nv_bfloat16, false>, gsb::LaunchCfg(grid, kThreads, 0, to_stream(s)), pv(x), P, C, scale, shift, center, act1, pv(out1), act2,
                                                                               o2, drop_seed, drop_offset, drop_mask);
  } else if (per_image) {
    gsb::launch(bn_act_kernel<float, false, true>, gsb::LaunchCfg(grid, kThreads, 0, to_stream(s)), pv(x), P, C, scale, shift, center, act1, pv(out1),
                act2, o2, drop_seed, drop_offset, drop_mask);
  } else {
    gsb::launch(bn_act_kernel<float, false>, gsb::LaunchCfg(grid, kThreads, 0, to_stream(s)), pv(x), P, C, scale, shift, center, act1, pv(out1), act2, o2,
                                                                     drop_seed, drop_offset, drop_mask);
  }
  return check_launch("bn_act");
}

extern "C" int gsb_bn_act_pool(const gsb_view* x, const float* scale, const float* shift, const gsb_view* out1,
                               const gsb_view* pooled, gsb_stream s) {
  if (check_vec_view(x, "bn_act_pool.x") || check_vec_view(out1, "bn_act_pool.out1") || check_vec_view(pooled, "bn_act_pool.pooled"))
    return -1;
  GSB_REQUIRE(scale && shift, "bn_act_pool: scale / shift required");
  GSB_REQUIRE(out1->dtype == x->dtype && out1->c == x->c && out1->n == x->n && out1->h == x->h && out1->w == x->w,
              "bn_act_pool: out1 mismatch");
  GSB_REQUIRE(x->h % 2 == 0 && x->w % 2 == 0 && x->h > 0 && x->w > 0, "bn_act_pool: even height and width required");
  GSB_REQUIRE(pooled->dtype == x->dtype && pooled->c == x->c && pooled->n == x->n && pooled->h == x->h / 2 && pooled->w == x->w / 2,
              "bn_act_pool: pooled mismatch");
  const int64_t total = (int64_t)x->n * (x->h / 2) * (x->w / 2) * (x->c / 8);
  int64_t blocks = (total + kThreads - 1) / kThreads;
  const int64_t cap = (int64_t)num_sms() * 16;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  if (x->dtype == GSB_BF16)
    gsb::launch(bn_act_pool_kernel<__nv_bfloat16>, gsb::LaunchCfg((int)blocks, kThreads, 0, to_stream(s)), pv(x), x->n, x->h, x->w, x->c,
                scale, shift, pv(out1), pv(pooled));
  else
    gsb::launch(bn_act_pool_kernel<float>, gsb::LaunchCfg((int)blocks, kThreads, 0, to_stream(s)), pv(x), x->n, x->h, x->w, x->c, scale,
                shift, pv(out1), pv(pooled));
  return check_launch("bn_act_pool");
}

// Explicit Dropout(0.5) keep mask (uint8, dense NHWC, local indexing) holding exactly the bits the fused Philox path draws
// for seed = (seed_k + *calls) mod 2^62 + 1 + layer_term -- the per-call seed rule of the generator program
// (gan_program.py) with the CALL COUNT IN DEVICE MEMORY, so that a captured CUDA graph of the training step draws a new
// mask at every replay (a host-side count would freeze the capture step's masks into the graph).
namespace gsb {
__global__ void __launch_bounds__(256) dropout_mask_kernel(uint8_t* __restrict__ mask, int64_t numel, uint64_t seed_k,
                                                          const int64_t* __restrict__ calls, uint64_t layer_term,
                                                          int64_t drop_offset) {
  GSB_PDL_ENTRY();
  const uint64_t seed = ((seed_k + (uint64_t)*calls) & ((1ull << 62) - 1)) + 1ull + layer_term;
  const int64_t groups = numel >> 3;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < groups; i += (int64_t)gridDim.x * 256) {
    const uint32_t bits = dropout_keep8(seed, (uint64_t)drop_offset + (uint64_t)i * 8ull);
    uint2 r;
    r.x = ((bits >> 0) & 1u) | (((bits >> 1) & 1u) << 8) | (((bits >> 2) & 1u) << 16) | (((bits >> 3) & 1u) << 24);
    r.y = ((bits >> 4) & 1u) | (((bits >> 5) & 1u) << 8) | (((bits >> 6) & 1u) << 16) | (((bits >> 7) & 1u) << 24);
    *reinterpret_cast<uint2*>(mask + i * 8) = r;
  }
}
}  // namespace gsb

extern "C" int gsb_dropout_mask(uint8_t* mask, int64_t numel, uint64_t seed_k, const int64_t* calls, uint64_t layer_term,
                                int64_t drop_offset, gsb_stream s) {
  GSB_REQUIRE(mask && calls && numel > 0 && numel % 8 == 0 && ((uintptr_t)mask & 7) == 0 && drop_offset % 8 == 0,
              "dropout_mask: bad argument (numel and drop_offset must be multiples of 8)");
  int64_t blocks = (numel / 8 + 255) / 256;
  if (blocks > (int64_t)num_sms() * 8) blocks = (int64_t)num_sms() * 8;
  gsb::launch(gsb::dropout_mask_kernel, gsb::LaunchCfg((int)blocks, 256, 0, to_stream(s)), mask, numel, seed_k, calls, layer_term, drop_offset);
  return check_launch("dropout_mask");
}

static int check_bwd_views(const gsb_view* x, const gsb_view* dA1, const gsb_view* dA2) {
  if (check_vec_view(x, "bn_bwd.x") || check_vec_view(dA1, "bn_bwd.dA1")) return -1;
  GSB_REQUIRE(dA1->dtype == x->dtype && dA1->c == x->c && num_pixels(dA1) == num_pixels(x), "bn_bwd: dA1 mismatch");
  if (dA2 && dA2->ptr) {
    if (check_vec_view(dA2, "bn_bwd.dA2")) return -1;
    GSB_REQUIRE(dA2->dtype == x->dtype && dA2->c == x->c && num_pixels(dA2) == num_pixels(x), "bn_bwd: dA2 mismatch");
  }
  return 0;
}

// Stage 1 of the backward statistics: per-block rows of [sum dy, sum dy*xhat] into `workspace`; returns the row count.
static int bn_bwd_partial_launch(const gsb_view* x, const gsb_view* dA1, const gsb_view* dA2, const float* scale,
                                 const float* shift, const float* center, const float* mean, const float* invstd, int act1, int act2,
                                 uint64_t drop_seed, int64_t drop_offset, const uint8_t* drop_mask, void* workspace, size_t workspace_bytes,
                                 int* rows_out, int per_image, gsb_stream s) {
  if (check_bwd_views(x, dA1, dA2)) return -1;
  const int C = x->c, lanes = kThreads / (C / 8);
  const int nimg = per_image ? x->n : 1;
  const int64_t P = num_pixels(x) / nimg;
  const int gx = per_image ? per_image_grid(P, lanes, nimg) : grid_for(P, lanes);
  dim3 grid(gx, nimg);
  const int wide = x->dtype == GSB_BF16 ? 1 : 2;
  GSB_REQUIRE(workspace_bytes >= (size_t)gx * wide * 2 * C * nimg * sizeof(float), "bn_bwd_reduce: workspace too small");
  size_t smem = (size_t)lanes * 2 * C * sizeof(float) * wide;
  if (smem > 48 * 1024) {
    cudaFuncSetAttribute(bn_reduce_kernel<float, 1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(bn_reduce_kernel<float, 1, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  }
  float* partial = reinterpret_cast<float*>(workspace);
  PixView d2 = (dA2 && dA2->ptr) ? pv(dA2) : PixView{nullptr, 0};
  const bool fast = act1 == GSB_ACT_RELU && d2.ptr == nullptr && drop_seed == 0 && drop_mask == nullptr && scale && shift &&
                    mean && invstd && center == nullptr && !per_image;
  if (x->dtype == GSB_BF16) {
    if (fast && relu_stats_fast_enabled())
      gsb::launch(bn_bwd_reduce_relu_bf16_kernel<2>, gsb::LaunchCfg(grid, kThreads, smem, to_stream(s)), 
          (const __nv_bfloat16*)x->ptr, x->sw, (const __nv_bfloat16*)dA1->ptr, dA1->sw, P, C, scale, shift, mean, invstd, partial);
    else if (fast)
      gsb::launch(bn_reduce_kernel<__nv_bfloat16, 1, true>, gsb::LaunchCfg(grid, kThreads, smem, to_stream(s)), 
          pv(x), pv(dA1), d2, P, C, scale, shift, center, mean, invstd, act1, act2, drop_seed, drop_offset, drop_mask, partial);
    else if (per_image)
      gsb::launch(bn_reduce_kernel<__nv_bfloat16, 1, false, true>, gsb::LaunchCfg(grid, kThreads, smem, to_stream(s)), 
          pv(x), pv(dA1), d2, P, C, scale, shift, center, mean, invstd, act1, act2, drop_seed, drop_offset, drop_mask, partial);
    else
      gsb::launch(bn_reduce_kernel<__nv_bfloat16, 1, false>, gsb::LaunchCfg(grid, kThreads, smem, to_stream(s)), 
          pv(x), pv(dA1), d2, P, C, scale, shift, center, mean, invstd, act1, act2, drop_seed, drop_offset, drop_mask, partial);
  } else if (per_image) {
    gsb::launch(bn_reduce_kernel<float, 1, false, true>, gsb::LaunchCfg(grid, kThreads, smem, to_stream(s)), pv(x), pv(dA1), d2, P, C, scale, shift, center, mean,
                invstd, act1, act2, drop_seed, drop_offset, drop_mask, partial);
  } else {
    gsb::launch(bn_reduce_kernel<float, 1, false>, gsb::LaunchCfg(grid, kThreads, smem, to_stream(s)), pv(x), pv(dA1), d2, P, C, scale, shift, center, mean,
                                                                             invstd, act1, act2, drop_seed, drop_offset, drop_mask, partial);
  }
  *rows_out = gx * wide;
  return 0;
}

extern "C" int gsb_bn_bwd_reduce(const gsb_view* x, const gsb_view* dA1, const gsb_view* dA2, const float* scale,
                                 const float* shift, const float* center, const float* mean, const float* invstd, int act1, int act2,
                                 uint64_t drop_seed, int64_t drop_offset, const uint8_t* drop_mask, double* sums, void* workspace,
                                 size_t workspace_bytes, gsb_stream s) {
  int grid = 0;
  if (bn_bwd_partial_launch(x, dA1, dA2, scale, shift, center, mean, invstd, act1, act2, drop_seed, drop_offset, drop_mask, workspace,
                            workspace_bytes, &grid, 0, s))
    return -1;
  int n = 2 * x->c;
  gsb::launch(reduce_partials_kernel, gsb::LaunchCfg((n + 31) / 32, 1024, 0, to_stream(s)), reinterpret_cast<float*>(workspace), grid, n, sums);
  return check_launch("bn_bwd_reduce", 2);
}

extern "C" int gsb_bn_bwd_partial(const gsb_view* x, const gsb_view* dA1, const gsb_view* dA2, const float* scale,
                                  const float* shift, const float* center, const float* mean, const float* invstd, int act1, int act2,
                                  uint64_t drop_seed, int64_t drop_offset, const uint8_t* drop_mask, void* workspace, size_t workspace_bytes,
                                  int* rows, int per_image, gsb_stream s) {
  GSB_REQUIRE(rows != nullptr, "bn_bwd_partial: rows is null");
  if (bn_bwd_partial_launch(x, dA1, dA2, scale, shift, center, mean, invstd, act1, act2, drop_seed, drop_offset, drop_mask, workspace,
                            workspace_bytes, rows, per_image, s))
    return -1;
  return check_launch("bn_bwd_partial");
}

// Stage 1 of the backward statistics with the MaxPool2d(2) backward folded in: dA (in / out) += route(dpool), rows of
// [sum dy, sum dy*xhat] into `workspace`.  bf16, BatchNorm + ReLU, even h and w (the caller falls back to
// gsb_maxpool2_bwd + gsb_bn_bwd_partial otherwise).
extern "C" int gsb_bn_bwd_partial_pool(const gsb_view* x, const gsb_view* dA, const gsb_view* dpool, const float* scale,
                                       const float* shift, const float* mean, const float* invstd, void* workspace,
                                       size_t workspace_bytes, int* rows, gsb_stream s) {
  GSB_REQUIRE(rows != nullptr && workspace != nullptr, "bn_bwd_partial_pool: null argument");
  if (check_vec_view(x, "bn_bwd_pool.x") || check_vec_view(dA, "bn_bwd_pool.dA") || check_vec_view(dpool, "bn_bwd_pool.dpool")) return -1;
  GSB_REQUIRE(x->dtype == GSB_BF16 && dA->dtype == GSB_BF16 && dpool->dtype == GSB_BF16, "bn_bwd_partial_pool: bf16 only");
  GSB_REQUIRE(scale && shift && mean && invstd, "bn_bwd_partial_pool: BatchNorm parameters required");
  GSB_REQUIRE(dA->n == x->n && dA->h == x->h && dA->w == x->w && dA->c == x->c, "bn_bwd_partial_pool: dA mismatch");
  GSB_REQUIRE(x->h % 2 == 0 && x->w % 2 == 0 && x->h > 0 && x->w > 0, "bn_bwd_partial_pool: even height and width required");
  GSB_REQUIRE(dpool->n == x->n && dpool->h == x->h / 2 && dpool->w == x->w / 2 && dpool->c == x->c, "bn_bwd_partial_pool: dpool mismatch");
  const int C = x->c, lanes = kThreads / (C / 8);
  const int64_t windows = num_pixels(dpool);
  const int gx = grid_for(windows, lanes);
  GSB_REQUIRE(workspace_bytes >= (size_t)gx * 2 * C * sizeof(float), "bn_bwd_partial_pool: workspace too small");
  const size_t smem = (size_t)lanes * 2 * C * sizeof(float);
  gsb::launch(bn_bwd_reduce_relu_pool_bf16_kernel, gsb::LaunchCfg(gx, kThreads, smem, to_stream(s)), (const __nv_bfloat16*)x->ptr, x->sw,
              (__nv_bfloat16*)dA->ptr, dA->sw, (const __nv_bfloat16*)dpool->ptr, dpool->sw, x->n, x->h, x->w, C, scale, shift, mean,
              invstd, reinterpret_cast<float*>(workspace));
  *rows = gx;
  return check_launch("bn_bwd_partial_pool");
}

static int fill_regions(PeerRegions& regs, void* const* regions, int rank, int world, int c, const char* who) {
  memset(&regs, 0, sizeof(regs));
  GSB_REQUIRE(world >= 1 && world <= kPeerMaxWorld && rank >= 0 && rank < world, "%s: world %d rank %d", who, world, rank);
  if (world > 1) {
    GSB_REQUIRE(regions != nullptr, "%s: peer regions required for world > 1", who);
    GSB_REQUIRE(2 * c <= kPeerMaxN, "%s: %d channels exceed the peer slot size", who, c);
    for (int i = 0; i < world; ++i) {
      GSB_REQUIRE(regions[i] != nullptr, "%s: region %d is null", who, i);
      regs.r[i] = regions[i];
    }
  }
  return 0;
}

// Forward: statistic rows [rows, 2c] (conv epilogue / stem / bn_reduce stage 1) -> scale, shift, mean, invstd and the
// running-statistics update, with the cross-rank sum of [sum x, sum x^2] fused in when world > 1 (`count` = GLOBAL count).
extern "C" int gsb_bn_stats_finalize(const float* partial, int rows, int c, double count, void* const* regions, int rank,
                                     int world, const float* gamma, const float* beta, float* running_mean, float* running_var,
                                     int64_t* num_batches_tracked, float momentum, float eps, float* scale, float* shift,
                                     float* mean, float* invstd, int centered, gsb_stream s) {
  GSB_REQUIRE(partial && scale && shift && rows > 0 && c > 0, "bn_stats_finalize: bad arguments");
  GSB_REQUIRE(!centered || mean != nullptr, "bn_stats_finalize: the centered form needs the mean output");
  PeerRegions regs;
  if (fill_regions(regs, regions, rank, world, c, "bn_stats_finalize")) return -1;
  BnFusedFwd f{count, gamma, beta, running_mean, running_var, num_batches_tracked, momentum, eps, scale, shift, mean, invstd, centered};
  BnFusedBwd b{nullptr, nullptr, 0, nullptr};
  gsb::launch(bn_stats_fused_kernel<0>, gsb::LaunchCfg((c + 15) / 16, 1024, 0, to_stream(s)), partial, rows, c, regs, rank, world, f, b,
              peer_timeout_ns());
  return check_launch("bn_stats_finalize");
}

// Backward: rows of [sum dy, sum dy*xhat] -> dbeta / dgamma from the local totals, global totals -> sums[2c] (fp64).
extern "C" int gsb_bn_bwd_stats(const float* partial, int rows, int c, void* const* regions, int rank, int world, float* dgamma,
                                float* dbeta, int accumulate, double* sums, gsb_stream s) {
  GSB_REQUIRE(partial && sums && rows > 0 && c > 0, "bn_bwd_stats: bad arguments");
  PeerRegions regs;
  if (fill_regions(regs, regions, rank, world, c, "bn_bwd_stats")) return -1;
  BnFusedFwd f{};
  BnFusedBwd b{dgamma, dbeta, accumulate, sums};
  gsb::launch(bn_stats_fused_kernel<1>, gsb::LaunchCfg((c + 15) / 16, 1024, 0, to_stream(s)), partial, rows, c, regs, rank, world, f, b,
              peer_timeout_ns());
  return check_launch("bn_bwd_stats");
}

extern "C" int gsb_bn_bwd_apply(const gsb_view* x, const gsb_view* dA1, const gsb_view* dA2, const float* scale,
                                const float* shift, const float* center, const float* mean, const float* invstd, int act1, int act2,
                                uint64_t drop_seed, int64_t drop_offset, const uint8_t* drop_mask, const double* sums, double count,
                                const gsb_view* dx, float* dgamma, float* dbeta, int per_image, gsb_stream s) {
  if (check_bwd_views(x, dA1, dA2) || check_vec_view(dx, "bn_bwd.dx")) return -1;
  GSB_REQUIRE(dx->dtype == x->dtype && dx->c == x->c && num_pixels(dx) == num_pixels(x), "bn_bwd_apply: dx mismatch");
  GSB_REQUIRE(mean == nullptr || sums != nullptr, "bn_bwd_apply: sums required with BN");
  const int C = x->c, lanes = kThreads / (C / 8);
  const int nimg = per_image ? x->n : 1;     // per_image: parameters [n][c], sums [2][n * c], count = pixels per image
  const int64_t P = num_pixels(x) / nimg;
  dim3 grid(grid_for(P, lanes * 4), nimg);
  PixView d2 = (dA2 && dA2->ptr) ? pv(dA2) : PixView{nullptr, 0};
  double inv = count > 0 ? 1.0 / count : 0.0;
  const bool fast = act1 == GSB_ACT_RELU && d2.ptr == nullptr && drop_seed == 0 && drop_mask == nullptr && scale && shift &&
                    mean && invstd && center == nullptr && !per_image;
  GSB_REQUIRE(!per_image || (dgamma == nullptr && dbeta == nullptr), "bn_bwd_apply: per-image statistics have no affine parameters");
  if (x->dtype == GSB_BF16) {
    if (fast)
      gsb::launch(bn_bwd_apply_kernel<__nv_bfloat16, true>, gsb::LaunchCfg(grid, kThreads, 0, to_stream(s)), 
          pv(x), pv(dA1), d2, P, C, scale, shift, center, mean, invstd, act1, act2, drop_seed, drop_offset, drop_mask, sums, inv, pv(dx));
    else if (per_image)
      gsb::launch(bn_bwd_apply_kernel<__nv_bfloat16, false, true>, gsb::LaunchCfg(grid, kThreads, 0, to_stream(s)), 
          pv(x), pv(dA1), d2, P, C, scale, shift, center, mean, invstd, act1, act2, drop_seed, drop_offset, drop_mask, sums, inv, pv(dx));
    else
      gsb::launch(bn_bwd_apply_kernel<__nv_bfloat16, false>, gsb::LaunchCfg(grid, kThreads, 0, to_stream(s)), 
          pv(x), pv(dA1), d2, P, C, scale, shift, center, mean, invstd, act1, act2, drop_seed, drop_offset, drop_mask, sums, inv, pv(dx));
  } else if (per_image) {
    gsb::launch(bn_bwd_apply_kernel<float, false, true>, gsb::LaunchCfg(grid, kThreads, 0, to_stream(s)), pv(x), pv(dA1), d2, P, C, scale, shift, center, mean,
                invstd, act1, act2, drop_seed, drop_offset, drop_mask, sums, inv, pv(dx));
  } else {
    gsb::launch(bn_bwd_apply_kernel<float, false>, gsb::LaunchCfg(grid, kThreads, 0, to_stream(s)), pv(x), pv(dA1), d2, P, C, scale, shift, center, mean, invstd,
                                                                           act1, act2, drop_seed, drop_offset, drop_mask, sums, inv, pv(dx));
  }
  if (mean != nullptr && (dgamma || dbeta))
    gsb::launch(bn_param_grad_kernel, gsb::LaunchCfg((C + 127) / 128, 128, 0, to_stream(s)), sums, C, dgamma, dbeta, 1);
  return check_launch("bn_bwd_apply", (mean != nullptr && (dgamma || dbeta)) ? 2 : 1);
}

extern "C" int gsb_bn_param_grad(const double* sums, int c, float* dgamma, float* dbeta, int accumulate, gsb_stream s) {
  GSB_REQUIRE(sums != nullptr, "bn_param_grad: null sums");
  gsb::launch(bn_param_grad_kernel, gsb::LaunchCfg((c + 127) / 128, 128, 0, to_stream(s)), sums, c, dgamma, dbeta, accumulate);
  return check_launch("bn_param_grad");
}

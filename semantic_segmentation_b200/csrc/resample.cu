// MaxPool2d(2) and bilinear x2 upsample (+pad, +concat-slice write): bandwidth kernels.
// Replaces unet/unet_parts.py:34 (nn.MaxPool2d(2)) and :50,57-67 (nn.Upsample(scale_factor=2,
// 'bilinear', align_corners=True) -> F.pad -> torch.cat).  128-bit NHWC accesses; the upsample
// writes straight into the upper channel half of the concat buffer, so no cat copy exists.
// Algorithmic bytes: pool fwd = 1 read + 1/4 write; upsample fwd = 1/4 read + 1 write.
#include <stdlib.h>
#include "common.cuh"

namespace gsb {

// (group, x, y, n) of flat item i = ((n*OH + oy)*OW + ox)*G + g; 32-bit divisions whenever the index fits
struct ItemIdx {
  int g, ox, oy, n;
  int64_t p;
};
__device__ __forceinline__ ItemIdx decode_item(int64_t i, int G, int OW, int OH) {
  ItemIdx r;
  if ((i >> 32) == 0) {
    const unsigned u = (unsigned)i;
    const unsigned p = u / (unsigned)G;
    r.g = (int)(u - p * (unsigned)G);
    const unsigned row = p / (unsigned)OW;
    r.ox = (int)(p - row * (unsigned)OW);
    const unsigned n = row / (unsigned)OH;
    r.oy = (int)(row - n * (unsigned)OH);
    r.n = (int)n;
    r.p = (int64_t)p;
  } else {
    r.g = (int)(i % G);
    r.p = i / G;
    r.ox = (int)(r.p % OW);
    const int64_t row = r.p / OW;
    r.oy = (int)(row % OH);
    r.n = (int)(row / OH);
  }
  return r;
}

template <typename T>
__global__ void __launch_bounds__(256) maxpool2_fwd_kernel(const T* __restrict__ x, int64_t xld, int N, int H, int W, int C,
                                                          T* __restrict__ y, int64_t yld) {
  GSB_PDL_ENTRY();
  const int G = C >> 3, OH = H >> 1, OW = W >> 1;
  const int64_t total = (int64_t)N * OH * OW * G;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const ItemIdx it = decode_item(i, G, OW, OH);
    const int g = it.g, ox = it.ox, oy = it.oy, n = it.n;
    const int64_t p = it.p;
    const T* base = x + (((int64_t)n * H + 2 * oy) * W + 2 * ox) * xld + g * 8;
    float m[8], v[8];
    Vec8<T>::load(base, m);
#pragma unroll
    for (int t = 1; t < 4; ++t) {
      Vec8<T>::load(base + ((t >> 1) * (int64_t)W + (t & 1)) * xld, v);
#pragma unroll
      for (int k = 0; k < 8; ++k)
        if (v[k] > m[k] || v[k] != v[k]) m[k] = v[k];
    }
    Vec8<T>::store(y + p * yld + g * 8, m);
  }
}

template <typename T>
__global__ void __launch_bounds__(256) maxpool2_bwd_kernel(const T* __restrict__ x, int64_t xld, int N, int H, int W, int C,
                                                          const T* __restrict__ dy, int64_t dyld, T* __restrict__ dx,
                                                          int64_t dxld, int accumulate) {
  GSB_PDL_ENTRY();
  const int G = C >> 3, OH = H >> 1, OW = W >> 1;
  const int WH = (H + 1) >> 1, WW = (W + 1) >> 1;  // windows incl. the ragged last row/col
  const int64_t total = (int64_t)N * WH * WW * G;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const ItemIdx it = decode_item(i, G, WW, WH);
    const int g = it.g, ox = it.ox, oy = it.oy, n = it.n;
    const int64_t pix0 = ((int64_t)n * H + 2 * oy) * W + 2 * ox;
    if (oy >= OH || ox >= OW) {  // not a pooling window: gradient is zero there
      if (!accumulate) {
        float z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int t = 0; t < 4; ++t) {
          int yy = 2 * oy + (t >> 1), xx = 2 * ox + (t & 1);
          if (yy < H && xx < W) Vec8<T>::store(dx + (pix0 + (t >> 1) * (int64_t)W + (t & 1)) * dxld + g * 8, z);
        }
      }
      continue;
    }
    float v[4][8], gr[8];
#pragma unroll
    for (int t = 0; t < 4; ++t) Vec8<T>::load(x + (pix0 + (t >> 1) * (int64_t)W + (t & 1)) * xld + g * 8, v[t]);
    Vec8<T>::load(dy + (((int64_t)n * OH + oy) * OW + ox) * dyld + g * 8, gr);
    int arg[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float m = v[0][k];
      int a = 0;
#pragma unroll
      for (int t = 1; t < 4; ++t)
        if (v[t][k] > m || v[t][k] != v[t][k]) { m = v[t][k]; a = t; }
      arg[k] = a;
    }
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      T* d = dx + (pix0 + (t >> 1) * (int64_t)W + (t & 1)) * dxld + g * 8;
      float o[8];
      if (accumulate) Vec8<T>::load(d, o);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        float c = arg[k] == t ? gr[k] : 0.f;
        o[k] = accumulate ? o[k] + c : c;
      }
      Vec8<T>::store(d, o);
    }
  }
}

struct Lerp {
  int i0, i1;
  float l0, l1;
};
// PyTorch align_corners=True source index (ATen UpSample.h: area_pixel_compute_scale /
// guard_index_and_lambda), all in fp32.
__device__ __forceinline__ Lerp lerp_for(int o, int in, int out) {
  float scale = out > 1 ? (float)(in - 1) / (float)(out - 1) : 0.f;
  float real = __fmul_rn(scale, (float)o);  // rounded product as ATen's CPU path (no contraction into the subtraction below)
  int i0 = (int)real;
  if (i0 > in - 1) i0 = in - 1;
  float l1 = real - (float)i0;
  l1 = fminf(fmaxf(l1, 0.f), 1.f);
  Lerp L;
  L.i0 = i0;
  L.i1 = i0 + ((i0 + 1 > in - 1) ? 0 : 1);
  L.l1 = l1;
  L.l0 = 1.f - l1;
  return L;
}

template <typename T>
__global__ void __launch_bounds__(256) upsample2_fwd_kernel(const T* __restrict__ x, int64_t xld, int N, int h, int w, int C,
                                                           T* __restrict__ out, int64_t old, int OH, int OW, int pad_top,
                                                           int pad_left) {
  GSB_PDL_ENTRY();
  const int G = C >> 3;
  const int64_t total = (int64_t)N * OH * OW * G;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const ItemIdx it = decode_item(i, G, OW, OH);
    const int g = it.g, ox = it.ox, oy = it.oy, n = it.n;
    const int64_t p = it.p;
    int uy = oy - pad_top, ux = ox - pad_left;
    float o[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (uy >= 0 && uy < 2 * h && ux >= 0 && ux < 2 * w) {
      Lerp ly = lerp_for(uy, h, 2 * h), lx = lerp_for(ux, w, 2 * w);
      const T* b = x + (int64_t)n * h * w * xld + g * 8;
      float v00[8], v01[8], v10[8], v11[8];
      Vec8<T>::load(b + ((int64_t)ly.i0 * w + lx.i0) * xld, v00);
      Vec8<T>::load(b + ((int64_t)ly.i0 * w + lx.i1) * xld, v01);
      Vec8<T>::load(b + ((int64_t)ly.i1 * w + lx.i0) * xld, v10);
      Vec8<T>::load(b + ((int64_t)ly.i1 * w + lx.i1) * xld, v11);
#pragma unroll
      for (int k = 0; k < 8; ++k)
        o[k] = ly.l0 * (lx.l0 * v00[k] + lx.l1 * v01[k]) + ly.l1 * (lx.l0 * v10[k] + lx.l1 * v11[k]);
    }
    Vec8<T>::store(out + p * old + g * 8, o);
  }
}

// dx[iy][ix] = sum over upsampled pixels whose lerp touches (iy,ix): exact transpose of the forward.
template <typename T>
__global__ void __launch_bounds__(256) upsample2_bwd_kernel(const T* __restrict__ dout, int64_t dld, int N, int OH, int OW,
                                                           int C, T* __restrict__ dx, int64_t xld, int h, int w, int pad_top,
                                                           int pad_left) {
  GSB_PDL_ENTRY();
  const int G = C >> 3;
  const int64_t total = (int64_t)N * h * w * G;
  const float sy = 2 * h > 1 ? (float)(h - 1) / (float)(2 * h - 1) : 0.f;
  const float sx = 2 * w > 1 ? (float)(w - 1) / (float)(2 * w - 1) : 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const ItemIdx it = decode_item(i, G, w, h);
    const int g = it.g, ix = it.ox, iy = it.oy, n = it.n;
    const int64_t p = it.p;
    int ylo = 0, yhi = 2 * h - 1, xlo = 0, xhi = 2 * w - 1;
    if (sy > 0.f) {
      ylo = max(0, (int)floorf((float)(iy - 1) / sy) - 1);
      yhi = min(2 * h - 1, (int)ceilf((float)(iy + 1) / sy) + 1);
    }
    if (sx > 0.f) {
      xlo = max(0, (int)floorf((float)(ix - 1) / sx) - 1);
      xhi = min(2 * w - 1, (int)ceilf((float)(ix + 1) / sx) + 1);
    }
    // separable: the (<= 4) contributing output columns and their weights are found once, not once per row
    constexpr int kMaxC = 6;
    int cx[kMaxC];
    float cw[kMaxC];
    int ncx = 0;
    for (int ux = xlo; ux <= xhi && ncx < kMaxC; ++ux) {
      Lerp lx = lerp_for(ux, w, 2 * w);
      float wx = (lx.i0 == ix ? lx.l0 : 0.f) + (lx.i1 == ix ? lx.l1 : 0.f);
      int ox = ux + pad_left;
      if (wx != 0.f && ox >= 0 && ox < OW) {
        cx[ncx] = ox;
        cw[ncx] = wx;
        ++ncx;
      }
    }
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int uy = ylo; uy <= yhi; ++uy) {
      Lerp ly = lerp_for(uy, h, 2 * h);
      float wy = (ly.i0 == iy ? ly.l0 : 0.f) + (ly.i1 == iy ? ly.l1 : 0.f);
      if (wy == 0.f) continue;
      int oy = uy + pad_top;
      if (oy < 0 || oy >= OH) continue;
      const T* rowp = dout + ((int64_t)n * OH + oy) * OW * dld + g * 8;
#pragma unroll
      for (int c = 0; c < kMaxC; ++c) {
        if (c < ncx) {
          float v[8];
          Vec8<T>::load(rowp + (int64_t)cx[c] * dld, v);
          float wgt = wy * cw[c];
#pragma unroll
          for (int k = 0; k < 8; ++k) acc[k] = fmaf(wgt, v[k], acc[k]);
        }
      }
    }
    Vec8<T>::store(dx + p * xld + g * 8, acc);
  }
}

// ---------------------------------------------------------------------------------------------------------
// Row-blocked bilinear kernels: one block per output (fwd) / input (bwd) row.  The horizontal lerp table is built
// once per block in shared memory, the vertical lerp once per row, the item -> (x, channel group) split is a
// 32-bit division: ~2x fewer instructions per 16-byte item than the flat kernels above, which are issue-bound
// (profiles/r01j: 0.66 + 0.92 ms per step against ~0.4 ms of HBM time).  Same arithmetic, same summation order.
// ---------------------------------------------------------------------------------------------------------
constexpr int kMaxUpW = 2048;

template <typename T>
__global__ void __launch_bounds__(256) upsample2_fwd_rows_kernel(const T* __restrict__ x, int64_t xld, int N, int h, int w, int C,
                                                                T* __restrict__ out, int64_t old, int OH, int OW, int pad_top,
                                                                int pad_left) {
  GSB_PDL_ENTRY();
  __shared__ int s_i0[kMaxUpW], s_i1[kMaxUpW];
  __shared__ float s_l1[kMaxUpW];
  for (int u = threadIdx.x; u < 2 * w; u += blockDim.x) {
    Lerp l = lerp_for(u, w, 2 * w);
    s_i0[u] = l.i0; s_i1[u] = l.i1; s_l1[u] = l.l1;
  }
  __syncthreads();
  const unsigned G = (unsigned)C >> 3, items = (unsigned)OW * G;
  for (int row = blockIdx.x; row < N * OH; row += gridDim.x) {
    const int n = row / OH, oy = row - n * OH;
    const int uy = oy - pad_top;
    const bool rowvalid = uy >= 0 && uy < 2 * h;
    Lerp ly = lerp_for(rowvalid ? uy : 0, h, 2 * h);
    const T* r0 = x + ((int64_t)n * h + ly.i0) * w * xld;
    const T* r1 = x + ((int64_t)n * h + ly.i1) * w * xld;
    T* orow = out + (int64_t)row * OW * old;
    for (unsigned it = threadIdx.x; it < items; it += blockDim.x) {
      const unsigned ox = it / G, g = it - ox * G;
      const int ux = (int)ox - pad_left;
      float o[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      if (rowvalid && ux >= 0 && ux < 2 * w) {
        const int i0 = s_i0[ux], i1 = s_i1[ux];
        const float lx1 = s_l1[ux], lx0 = 1.f - lx1;
        float v00[8], v01[8], v10[8], v11[8];
        Vec8<T>::load(r0 + (int64_t)i0 * xld + g * 8, v00);
        Vec8<T>::load(r0 + (int64_t)i1 * xld + g * 8, v01);
        Vec8<T>::load(r1 + (int64_t)i0 * xld + g * 8, v10);
        Vec8<T>::load(r1 + (int64_t)i1 * xld + g * 8, v11);
        // packed fp32x2 arithmetic (FMUL2 / FFMA2): the kernel is issue-bound (ncu profiles/r01n: 77 % issue active)
        const float2 a0 = make_float2(lx0, lx0), a1 = make_float2(lx1, lx1);
        const float2 b0 = make_float2(ly.l0, ly.l0), b1 = make_float2(ly.l1, ly.l1);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float2 top = __ffma2_rn(a1, make_float2(v01[2 * k], v01[2 * k + 1]),
                                        __fmul2_rn(a0, make_float2(v00[2 * k], v00[2 * k + 1])));
          const float2 bot = __ffma2_rn(a1, make_float2(v11[2 * k], v11[2 * k + 1]),
                                        __fmul2_rn(a0, make_float2(v10[2 * k], v10[2 * k + 1])));
          const float2 r = __ffma2_rn(b1, bot, __fmul2_rn(b0, top));
          o[2 * k] = r.x;
          o[2 * k + 1] = r.y;
        }
      }
      Vec8<T>::store(orow + (int64_t)ox * old + g * 8, o);
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(256) upsample2_bwd_rows_kernel(const T* __restrict__ dout, int64_t dld, int N, int OH, int OW,
                                                                int C, T* __restrict__ dx, int64_t xld, int h, int w,
                                                                int pad_top, int pad_left) {
  GSB_PDL_ENTRY();
  __shared__ int s_i0[kMaxUpW], s_i1[kMaxUpW];
  __shared__ float s_l1[kMaxUpW];
  __shared__ int s_oy[8];
  __shared__ float s_wy[8];
  for (int u = threadIdx.x; u < 2 * w; u += blockDim.x) {
    Lerp l = lerp_for(u, w, 2 * w);
    s_i0[u] = l.i0; s_i1[u] = l.i1; s_l1[u] = l.l1;
  }
  const unsigned G = (unsigned)C >> 3, items = (unsigned)w * G;
  for (int row = blockIdx.x; row < N * h; row += gridDim.x) {
    const int n = row / h, iy = row - n * h;
    __syncthreads();  // table ready / previous row's candidates consumed
    if (threadIdx.x < 6) {  // contributing output rows uy in [2iy-2, 2iy+3], ascending (same order as the flat kernel)
      const int uy = 2 * iy - 2 + (int)threadIdx.x;
      float wy = 0.f;
      int oy = -1;
      if (uy >= 0 && uy < 2 * h) {
        Lerp ly = lerp_for(uy, h, 2 * h);
        wy = (ly.i0 == iy ? ly.l0 : 0.f) + (ly.i1 == iy ? ly.l1 : 0.f);
        oy = uy + pad_top;
        if (oy < 0 || oy >= OH) wy = 0.f;
      }
      s_oy[threadIdx.x] = oy;
      s_wy[threadIdx.x] = wy;
    }
    __syncthreads();
    for (unsigned it = threadIdx.x; it < items; it += blockDim.x) {
      const unsigned ix = it / G, g = it - ix * G;
      int cx[6];
      float cw[6];
#pragma unroll
      for (int c = 0; c < 6; ++c) {
        const int ux = 2 * (int)ix - 2 + c;
        float wx = 0.f;
        int ox = 0;
        if (ux >= 0 && ux < 2 * w) {
          const float l1 = s_l1[ux];
          wx = (s_i0[ux] == (int)ix ? 1.f - l1 : 0.f) + (s_i1[ux] == (int)ix ? l1 : 0.f);
          ox = ux + pad_left;
          if (ox < 0 || ox >= OW) { wx = 0.f; ox = 0; }
        }
        cx[c] = ox;
        cw[c] = wx;
      }
      float2 acc2[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) acc2[k] = make_float2(0.f, 0.f);
#pragma unroll
      for (int r = 0; r < 6; ++r) {
        const float wy = s_wy[r];
        if (wy == 0.f) continue;
        const T* rowp = dout + ((int64_t)n * OH + s_oy[r]) * OW * dld + g * 8;
#pragma unroll
        for (int c = 0; c < 6; ++c) {
          if (cw[c] != 0.f) {
            float v[8];
            Vec8<T>::load(rowp + (int64_t)cx[c] * dld, v);
            const float wgt = wy * cw[c];
            const float2 w2 = make_float2(wgt, wgt);
#pragma unroll
            for (int k = 0; k < 4; ++k) acc2[k] = __ffma2_rn(w2, make_float2(v[2 * k], v[2 * k + 1]), acc2[k]);
          }
        }
      }
      float acc[8];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        acc[2 * k] = acc2[k].x;
        acc[2 * k + 1] = acc2[k].y;
      }
      Vec8<T>::store(dx + ((int64_t)row * w + ix) * xld + g * 8, acc);
    }
  }
}

// Backward, window form: for an input column ix the contributing output columns are a run of at most five, so the
// block prologue stores (first column, five weights) per ix and the row loop issues five UNCONDITIONAL, independent
// 16-byte loads per contributing output row.  The kernel above hides its loads behind per-thread weight tests (ncu
// profiles/r01n: issue 48 %, long-scoreboard bound, 1.5 TB/s).  Same products, same ascending summation order.
// Why five: the sources that touch ix lie in the open interval ((ix-1)/s, (ix+1)/s) with s = (w-1)/(2w-1), of length
// 4 + 2/(w-1) <= 5 for w >= 3 (and there are only 2w <= 4 columns for w <= 2).  A wider run traps (fails loudly).
constexpr int kUpWin = 5;
template <typename T>
__global__ void __launch_bounds__(256) upsample2_bwd_win_kernel(const T* __restrict__ dout, int64_t dld, int N, int OH, int OW, int C,
                                                               T* __restrict__ dx, int64_t xld, int h, int w, int pad_top,
                                                               int pad_left) {
  GSB_PDL_ENTRY();
  __shared__ int s_c0[kMaxUpW / 2];
  __shared__ float s_cw[kMaxUpW / 2][kUpWin];
  __shared__ int s_oy[8];
  __shared__ float s_wy[8];
  __shared__ int s_bad;
  if (threadIdx.x == 0) s_bad = 0;
  __syncthreads();
  for (int ix = threadIdx.x; ix < w; ix += blockDim.x) {
    int lo = 1 << 30, hi = -1;
    for (int c = 0; c < 6; ++c) {  // candidates ux in [2ix-2, 2ix+3]
      const int ux = 2 * ix - 2 + c;
      if (ux < 0 || ux >= 2 * w) continue;
      const Lerp l = lerp_for(ux, w, 2 * w);
      const float wx = (l.i0 == ix ? l.l0 : 0.f) + (l.i1 == ix ? l.l1 : 0.f);
      const int ox = ux + pad_left;
      if (wx != 0.f && ox >= 0 && ox < OW) {
        lo = min(lo, ox);
        hi = max(hi, ox);
      }
    }
    int ws = hi < 0 ? 0 : lo;
    if (ws > OW - kUpWin) ws = OW - kUpWin;
    if (ws < 0) ws = 0;
    if (hi >= 0 && (lo < ws || hi > ws + kUpWin - 1)) s_bad = 1;
    s_c0[ix] = ws;
    for (int k = 0; k < kUpWin; ++k) {
      const int ox = ws + k, ux = ox - pad_left;
      float wx = 0.f;
      if (ox < OW && ux >= 0 && ux < 2 * w) {
        const Lerp l = lerp_for(ux, w, 2 * w);
        wx = (l.i0 == ix ? l.l0 : 0.f) + (l.i1 == ix ? l.l1 : 0.f);
      }
      s_cw[ix][k] = wx;
    }
  }
  __syncthreads();
  if (s_bad) __trap();
  const unsigned G = (unsigned)C >> 3, items = (unsigned)w * G;
  const int lastcol = OW - 1;
  for (int row = blockIdx.x; row < N * h; row += gridDim.x) {
    const int n = row / h, iy = row - n * h;
    __syncthreads();
    if (threadIdx.x < 6) {
      const int uy = 2 * iy - 2 + (int)threadIdx.x;
      float wy = 0.f;
      int oy = 0;
      if (uy >= 0 && uy < 2 * h) {
        const Lerp ly = lerp_for(uy, h, 2 * h);
        wy = (ly.i0 == iy ? ly.l0 : 0.f) + (ly.i1 == iy ? ly.l1 : 0.f);
        oy = uy + pad_top;
        if (oy < 0 || oy >= OH) { wy = 0.f; oy = 0; }
      }
      s_oy[threadIdx.x] = oy;
      s_wy[threadIdx.x] = wy;
    }
    __syncthreads();
    for (unsigned it = threadIdx.x; it < items; it += blockDim.x) {
      const unsigned ix = it / G, g = it - ix * G;
      const int c0 = s_c0[ix];
      float cw[kUpWin];
#pragma unroll
      for (int k = 0; k < kUpWin; ++k) cw[k] = s_cw[ix][k];
      float2 acc2[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) acc2[k] = make_float2(0.f, 0.f);
#pragma unroll
      for (int r = 0; r < 6; ++r) {
        const float wy = s_wy[r];
        if (wy == 0.f) continue;  // block-uniform
        const T* rowp = dout + ((int64_t)n * OH + s_oy[r]) * OW * dld + g * 8;
        float v[kUpWin][8];
#pragma unroll
        for (int k = 0; k < kUpWin; ++k) Vec8<T>::load(rowp + (int64_t)min(c0 + k, lastcol) * dld, v[k]);
#pragma unroll
        for (int k = 0; k < kUpWin; ++k) {
          const float wgt = wy * cw[k];
          const float2 w2 = make_float2(wgt, wgt);
#pragma unroll
          for (int j = 0; j < 4; ++j) acc2[j] = __ffma2_rn(w2, make_float2(v[k][2 * j], v[k][2 * j + 1]), acc2[j]);
        }
      }
      float acc[8];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        acc[2 * k] = acc2[k].x;
        acc[2 * k + 1] = acc2[k].y;
      }
      Vec8<T>::store(dx + ((int64_t)row * w + ix) * xld + g * 8, acc);
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// Block form of the align_corners x2 lerp (un-padded case OH = 2h, OW = 2w, the one every even-sized level takes).
// With scale = (in-1)/(2in-1) < 1/2 the source of output u = 2b+1 lies in [b, b+1/2) and that of u = 2b+2 in
// (b+1/2, b+1): both outputs of block b = (u+1)/2 - 1 interpolate between the SAME two inputs b and b+1 (clamped to
// the tensor at b = -1 and b = in-1, where only one of the two outputs exists).  So a 2x2 output block reads one 2x2
// input block: 4 loads for 4 outputs (the row kernels above issue 4 loads per output and sit at ~0.4 of the HBM
// roofline, issue / L1 bound), and the transpose is separable with a 4-tap window: input i receives from outputs
// 2i-1 .. 2i+2.  The weight on the upper tap is clamp(scale*u - b, 0, 1), which equals lerp_for's l1 whenever
// (int)(scale*u) == b; when fp32 rounding puts scale*u an ulp across an integer (only possible at u = 2in-1) the
// result differs from lerp_for's by <= ulp * |x[b+1] - x[b]|.
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float up_blk_w1(int u, int in) {
  const float scale = (float)(in - 1) / (float)(2 * in - 1);
  const int b = ((u + 1) >> 1) - 1;
  return fminf(fmaxf(__fmul_rn(scale, (float)u) - (float)b, 0.f), 1.f);  // rounded product, as lerp_for / ATen
}
// weight with which output u feeds input i (the transpose of the above); 0 outside the tensor
__device__ __forceinline__ float up_blk_wgt(int u, int i, int in) {
  if (u < 0 || u >= 2 * in) return 0.f;
  const int b = ((u + 1) >> 1) - 1;
  const float w1 = up_blk_w1(u, in), w0 = 1.f - w1;
  const int ia = max(b, 0), ib = min(b + 1, in - 1);
  return (ia == i ? w0 : 0.f) + (ib == i ? w1 : 0.f);
}

template <typename T>
__global__ void __launch_bounds__(256) upsample2_fwd_blk_kernel(const T* __restrict__ x, int64_t xld, int N, int h, int w, int C,
                                                               T* __restrict__ out, int64_t old) {
  GSB_PDL_ENTRY();
  // one flat grid-stride loop over (image, row block, column block, channel group): no per-row tail (a row of the
  // production shape has 129 * 8 = 1032 items, 4.03 per thread of a row-owning block)
  __shared__ float s_wx[kMaxUpW], s_wy[kMaxUpW];
  for (int v = threadIdx.x; v < 2 * w; v += blockDim.x) s_wx[v] = up_blk_w1(v, w);
  for (int v = threadIdx.x; v < 2 * h; v += blockDim.x) s_wy[v] = up_blk_w1(v, h);
  __syncthreads();
  const unsigned G = (unsigned)C >> 3, per_row = (unsigned)(w + 1) * G;
  const unsigned H1 = (unsigned)h + 1;
  const int OW = 2 * w;
  const uint64_t total = (uint64_t)N * H1 * per_row;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (uint64_t)gridDim.x * blockDim.x) {
    unsigned rb, it;
    if ((i >> 32) == 0) {
      rb = (unsigned)i / per_row;
      it = (unsigned)i - rb * per_row;
    } else {
      rb = (unsigned)(i / per_row);
      it = (unsigned)(i - (uint64_t)rb * per_row);
    }
    const unsigned n = rb / H1;
    const int bi = (int)(rb - n * H1) - 1;
    const int u0 = 2 * bi + 1, u1 = u0 + 1;
    const bool hasr0 = bi >= 0, hasr1 = bi < h - 1;
    const float wy0 = s_wy[hasr0 ? u0 : 0], wy1 = s_wy[hasr1 ? u1 : 0];   // rows outside the tensor are not stored
    const T* r0 = x + ((int64_t)n * h + max(bi, 0)) * w * xld;
    const T* r1 = x + ((int64_t)n * h + min(bi + 1, h - 1)) * w * xld;
    T* o0 = out + ((int64_t)n * 2 * h + u0) * OW * old;
    T* o1 = o0 + (int64_t)OW * old;
    const float2 y00 = make_float2(1.f - wy0, 1.f - wy0), y01 = make_float2(wy0, wy0);
    const float2 y10 = make_float2(1.f - wy1, 1.f - wy1), y11 = make_float2(wy1, wy1);
    const unsigned bjp = it / G, g = it - bjp * G;
    const int bj = (int)bjp - 1;
    const int ca = max(bj, 0), cb = min(bj + 1, w - 1);
    const int v0 = 2 * bj + 1, v1 = v0 + 1;
    const bool hasc0 = bj >= 0, hasc1 = bj < w - 1;
    float v00[8], v01[8], v10[8], v11[8];
    Vec8<T>::load(r0 + (int64_t)ca * xld + g * 8, v00);
    Vec8<T>::load(r0 + (int64_t)cb * xld + g * 8, v01);
    Vec8<T>::load(r1 + (int64_t)ca * xld + g * 8, v10);
    Vec8<T>::load(r1 + (int64_t)cb * xld + g * 8, v11);
    const float wx0 = s_wx[hasc0 ? v0 : 0], wx1 = s_wx[hasc1 ? v1 : 0];
    const float2 a00 = make_float2(1.f - wx0, 1.f - wx0), a01 = make_float2(wx0, wx0);
    const float2 a10 = make_float2(1.f - wx1, 1.f - wx1), a11 = make_float2(wx1, wx1);
    float o00[8], o01[8], o10[8], o11[8];  // [row][col]
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float2 p00 = make_float2(v00[2 * k], v00[2 * k + 1]), p01 = make_float2(v01[2 * k], v01[2 * k + 1]);
      const float2 p10 = make_float2(v10[2 * k], v10[2 * k + 1]), p11 = make_float2(v11[2 * k], v11[2 * k + 1]);
      // same association as the row kernel: ly0 * (lx0 * v00 + lx1 * v01) + ly1 * (lx0 * v10 + lx1 * v11)
      const float2 top0 = __ffma2_rn(a01, p01, __fmul2_rn(a00, p00)), bot0 = __ffma2_rn(a01, p11, __fmul2_rn(a00, p10));
      const float2 top1 = __ffma2_rn(a11, p01, __fmul2_rn(a10, p00)), bot1 = __ffma2_rn(a11, p11, __fmul2_rn(a10, p10));
      const float2 q00 = __ffma2_rn(y01, bot0, __fmul2_rn(y00, top0)), q01 = __ffma2_rn(y01, bot1, __fmul2_rn(y00, top1));
      const float2 q10 = __ffma2_rn(y11, bot0, __fmul2_rn(y10, top0)), q11 = __ffma2_rn(y11, bot1, __fmul2_rn(y10, top1));
      o00[2 * k] = q00.x; o00[2 * k + 1] = q00.y;
      o01[2 * k] = q01.x; o01[2 * k + 1] = q01.y;
      o10[2 * k] = q10.x; o10[2 * k + 1] = q10.y;
      o11[2 * k] = q11.x; o11[2 * k + 1] = q11.y;
    }
    if (hasr0) {
      if (hasc0) Vec8<T>::store(o0 + (int64_t)v0 * old + g * 8, o00);
      if (hasc1) Vec8<T>::store(o0 + (int64_t)v1 * old + g * 8, o01);
    }
    if (hasr1) {
      if (hasc0) Vec8<T>::store(o1 + (int64_t)v0 * old + g * 8, o10);
      if (hasc1) Vec8<T>::store(o1 + (int64_t)v1 * old + g * 8, o11);
    }
  }
}
// Backward, separable: a block owns one input row and a run of JC input columns.  Phase 1 folds the (<= 4) contributing
// output rows into one fp32 row tile in shared memory (4 coalesced 16-byte loads per output column and channel group),
// phase 2 folds the 4 contributing columns of that tile per input column: 8 global loads per dx vector where the window
// kernel issues 20.  Tile layout [column][half][group][4 floats] so that both phases move 16-byte words conflict-free.
template <typename T> struct RawVec8;
template <> struct RawVec8<__nv_bfloat16> {
  using type = uint4;
  static __device__ __forceinline__ uint4 load(const __nv_bfloat16* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }
  static __device__ __forceinline__ void unpack(const uint4& r, float (&v)[8]) {
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      v[2 * i] = __uint_as_float(w[i] << 16);
      v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
  }
};
template <> struct RawVec8<float> {
  struct type { float4 a, b; };
  static __device__ __forceinline__ type load(const float* p) {
    type t;
    t.a = __ldg(reinterpret_cast<const float4*>(p));
    t.b = __ldg(reinterpret_cast<const float4*>(p + 4));
    return t;
  }
  static __device__ __forceinline__ void unpack(const type& r, float (&v)[8]) {
    v[0] = r.a.x; v[1] = r.a.y; v[2] = r.a.z; v[3] = r.a.w;
    v[4] = r.b.x; v[5] = r.b.y; v[6] = r.b.z; v[7] = r.b.w;
  }
};
constexpr int kSepMaxJC = 512;
template <typename T>
__global__ void __launch_bounds__(256) upsample2_bwd_sep_kernel(const T* __restrict__ dout, int64_t dld, int N, int C,
                                                               T* __restrict__ dx, int64_t xld, int h, int w, int JC) {
  GSB_PDL_ENTRY();
  extern __shared__ __align__(16) float s_t[];
  __shared__ float s_cw[kSepMaxJC][4];
  const int j0 = blockIdx.y * JC, jn = min(JC, w - j0);
  const int vbase = 2 * j0 - 1, ncol = 2 * jn + 2, OW = 2 * w;
  for (int i = threadIdx.x; i < jn * 4; i += blockDim.x) {
    const int ix = i >> 2, k = i & 3;
    s_cw[ix][k] = up_blk_wgt(2 * (j0 + ix) - 1 + k, j0 + ix, w);
  }
  const unsigned G = (unsigned)C >> 3, items1 = (unsigned)ncol * G, items2 = (unsigned)jn * G;
  const int halfC = C >> 1;
  for (int row = blockIdx.x; row < N * h; row += gridDim.x) {
    const int n = row / h, iy = row - n * h;
    float wy[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) wy[r] = up_blk_wgt(2 * iy - 1 + r, iy, h);
    const T* base = dout + ((int64_t)n * 2 * h + (2 * iy - 1)) * OW * dld;
    const int64_t rstride = (int64_t)OW * dld;
    int64_t roff[4];   // rows outside the tensor (block-uniform) weigh 0: they are read from the nearest row inside it
#pragma unroll
    for (int r = 0; r < 4; ++r) roff[r] = (int64_t)min(max(r, iy == 0 ? 1 : 0), iy == h - 1 ? 2 : 3) * rstride;
    __syncthreads();  // the previous row's phase 2 (and the weight table) is done
    // batches of U items per thread: the 4U row loads of a batch are independent and issued together (one memory round
    // trip per batch; one item at a time left the block latency-bound: 5 dependent round trips per row at 1040 items)
    constexpr int U = sizeof(T) == 2 ? 5 : 2;
    using Raw = typename RawVec8<T>::type;
    for (unsigned b0 = 0; b0 < items1; b0 += 256 * U) {
      Raw raw[U][4];
      bool ok[U];
      unsigned cs[U], gs[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const unsigned it = b0 + u * 256 + threadIdx.x;
        const unsigned c = it / G, g = it - c * G;
        cs[u] = c; gs[u] = g;
        const int v = vbase + (int)c;
        ok[u] = it < items1 && v >= 0 && v < OW;
        const T* p = base + (int64_t)(ok[u] ? v : 0) * dld + g * 8;
#pragma unroll
        for (int r = 0; r < 4; ++r)
          if (ok[u]) raw[u][r] = RawVec8<T>::load(p + roff[r]);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const unsigned it = b0 + u * 256 + threadIdx.x;
        if (it >= items1) continue;
        const unsigned c = cs[u], g = gs[u];
        float2 acc[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) acc[k] = make_float2(0.f, 0.f);
        if (ok[u]) {
#pragma unroll
          for (int r = 0; r < 4; ++r) {
            float t[8];
            RawVec8<T>::unpack(raw[u][r], t);
            const float2 w2 = make_float2(wy[r], wy[r]);
#pragma unroll
            for (int k = 0; k < 4; ++k) acc[k] = __ffma2_rn(w2, make_float2(t[2 * k], t[2 * k + 1]), acc[k]);
          }
        }
        float* d = s_t + (size_t)c * C + g * 4;
        *reinterpret_cast<float4*>(d) = make_float4(acc[0].x, acc[0].y, acc[1].x, acc[1].y);
        *reinterpret_cast<float4*>(d + halfC) = make_float4(acc[2].x, acc[2].y, acc[3].x, acc[3].y);
      }
    }
    __syncthreads();
    for (unsigned it = threadIdx.x; it < items2; it += blockDim.x) {
      const unsigned ix = it / G, g = it - ix * G;
      float2 acc[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) acc[k] = make_float2(0.f, 0.f);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float* sp = s_t + (size_t)(2 * ix + k) * C + g * 4;
        const float4 lo = *reinterpret_cast<const float4*>(sp), hi = *reinterpret_cast<const float4*>(sp + halfC);
        const float cw = s_cw[ix][k];
        const float2 w2 = make_float2(cw, cw);
        acc[0] = __ffma2_rn(w2, make_float2(lo.x, lo.y), acc[0]);
        acc[1] = __ffma2_rn(w2, make_float2(lo.z, lo.w), acc[1]);
        acc[2] = __ffma2_rn(w2, make_float2(hi.x, hi.y), acc[2]);
        acc[3] = __ffma2_rn(w2, make_float2(hi.z, hi.w), acc[3]);
      }
      float o[8];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        o[2 * k] = acc[k].x;
        o[2 * k + 1] = acc[k].y;
      }
      Vec8<T>::store(dx + ((int64_t)row * w + j0 + ix) * xld + g * 8, o);
    }
  }
}
static int ew_grid(int64_t total) {
  int64_t b = (total + 255) / 256;
  int64_t cap = (int64_t)num_sms() * 16;
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}
static int chk(const gsb_view* v, const char* nm) {
  GSB_REQUIRE(v && v->ptr, "%s: null view", nm);
  GSB_REQUIRE(vec8_ok(v), "%s: view must be pixel-linear NHWC, c%%8==0, 16B aligned", nm);
  return 0;
}

}  // namespace gsb

using namespace gsb;

#define DISPATCH_T(dtype, ...)                 \
  if ((dtype) == GSB_BF16) {                   \
    using T = __nv_bfloat16;                   \
    __VA_ARGS__;                               \
  } else {                                     \
    using T = float;                           \
    __VA_ARGS__;                               \
  }

extern "C" int gsb_maxpool2_fwd(const gsb_view* x, const gsb_view* y, gsb_stream s) {
  if (chk(x, "maxpool.x") || chk(y, "maxpool.y")) return -1;
  GSB_REQUIRE(y->n == x->n && y->h == x->h / 2 && y->w == x->w / 2 && y->c == x->c && y->dtype == x->dtype,
              "maxpool2_fwd: shape mismatch");
  int64_t total = num_pixels(y) * (x->c / 8);
  if (total == 0) return 0;
  DISPATCH_T(x->dtype, gsb::launch(maxpool2_fwd_kernel<T>, gsb::LaunchCfg(ew_grid(total), 256, 0, to_stream(s)), 
                           (const T*)x->ptr, x->sw, x->n, x->h, x->w, x->c, (T*)y->ptr, y->sw));
  return check_launch("maxpool2_fwd");
}

extern "C" int gsb_maxpool2_bwd(const gsb_view* x, const gsb_view* dy, const gsb_view* dx, int accumulate, gsb_stream s) {
  if (chk(x, "maxpool_bwd.x") || chk(dy, "maxpool_bwd.dy") || chk(dx, "maxpool_bwd.dx")) return -1;
  GSB_REQUIRE(dy->n == x->n && dy->h == x->h / 2 && dy->w == x->w / 2 && dy->c == x->c && dy->dtype == x->dtype,
              "maxpool2_bwd: dy mismatch");
  GSB_REQUIRE(dx->n == x->n && dx->h == x->h && dx->w == x->w && dx->c == x->c && dx->dtype == x->dtype,
              "maxpool2_bwd: dx mismatch");
  int64_t total = (int64_t)x->n * ((x->h + 1) / 2) * ((x->w + 1) / 2) * (x->c / 8);
  if (total == 0) return 0;
  DISPATCH_T(x->dtype, gsb::launch(maxpool2_bwd_kernel<T>, gsb::LaunchCfg(ew_grid(total), 256, 0, to_stream(s)), 
                           (const T*)x->ptr, x->sw, x->n, x->h, x->w, x->c, (const T*)dy->ptr, dy->sw, (T*)dx->ptr, dx->sw,
                           accumulate));
  return check_launch("maxpool2_bwd");
}

extern "C" int gsb_upsample2_fwd(const gsb_view* x, const gsb_view* out, int pad_top, int pad_left, gsb_stream s) {
  if (chk(x, "upsample.x") || chk(out, "upsample.out")) return -1;
  GSB_REQUIRE(out->n == x->n && out->c == x->c && out->dtype == x->dtype, "upsample2_fwd: mismatch");
  int64_t total = num_pixels(out) * (x->c / 8);
  if (total == 0) return 0;
  static const bool blk = []() { const char* e = getenv("GSB_UP_BLK"); return !(e && e[0] == '0'); }();
  if (blk && pad_top == 0 && pad_left == 0 && out->h == 2 * x->h && out->w == 2 * x->w && 2 * x->w <= kMaxUpW &&
      (int64_t)x->n * (x->h + 1) < ((int64_t)1 << 30)) {
    const int grid = ew_grid((int64_t)x->n * (x->h + 1) * (x->w + 1) * (x->c / 8));
    DISPATCH_T(x->dtype, gsb::launch(upsample2_fwd_blk_kernel<T>, gsb::LaunchCfg(grid, 256, 0, to_stream(s)),
                             (const T*)x->ptr, x->sw, x->n, x->h, x->w, x->c, (T*)out->ptr, out->sw));
    return check_launch("upsample2_fwd");
  }
  if (2 * x->w <= kMaxUpW && (int64_t)out->n * out->h < ((int64_t)1 << 30)) {
    const int rows = out->n * out->h;
    const int grid = rows < num_sms() * 16 ? rows : num_sms() * 16;
    DISPATCH_T(x->dtype, gsb::launch(upsample2_fwd_rows_kernel<T>, gsb::LaunchCfg(grid, 256, 0, to_stream(s)), 
                             (const T*)x->ptr, x->sw, x->n, x->h, x->w, x->c, (T*)out->ptr, out->sw, out->h, out->w, pad_top,
                             pad_left));
    return check_launch("upsample2_fwd");
  }
  DISPATCH_T(x->dtype, gsb::launch(upsample2_fwd_kernel<T>, gsb::LaunchCfg(ew_grid(total), 256, 0, to_stream(s)), 
                           (const T*)x->ptr, x->sw, x->n, x->h, x->w, x->c, (T*)out->ptr, out->sw, out->h, out->w, pad_top,
                           pad_left));
  return check_launch("upsample2_fwd");
}

extern "C" int gsb_upsample2_bwd(const gsb_view* dout, const gsb_view* dx, int pad_top, int pad_left, gsb_stream s) {
  if (chk(dout, "upsample_bwd.dout") || chk(dx, "upsample_bwd.dx")) return -1;
  GSB_REQUIRE(dout->n == dx->n && dout->c == dx->c && dout->dtype == dx->dtype, "upsample2_bwd: mismatch");
  int64_t total = num_pixels(dx) * (dx->c / 8);
  if (total == 0) return 0;
  static const bool blk = []() { const char* e = getenv("GSB_UP_BLK"); return !(e && e[0] == '0'); }();
  if (blk && pad_top == 0 && pad_left == 0 && dout->h == 2 * dx->h && dout->w == 2 * dx->w && dx->c <= 1024 &&
      (int64_t)dx->n * dx->h < ((int64_t)1 << 30)) {
    int jc = 4096 / dx->c;
    if (jc < 1) jc = 1;
    if (jc > dx->w) jc = dx->w;
    if (jc > kSepMaxJC) jc = kSepMaxJC;
    const int chunks = (dx->w + jc - 1) / jc;
    const int rows = dx->n * dx->h;
    int gx = num_sms() * 16 / chunks;
    if (gx < 1) gx = 1;
    if (gx > rows) gx = rows;
    const size_t smem = (size_t)(2 * jc + 2) * dx->c * sizeof(float);
    DISPATCH_T(dx->dtype, gsb::launch(upsample2_bwd_sep_kernel<T>, gsb::LaunchCfg(dim3(gx, chunks), 256, smem, to_stream(s)),
                              (const T*)dout->ptr, dout->sw, dout->n, dout->c, (T*)dx->ptr, dx->sw, dx->h, dx->w, jc));
    return check_launch("upsample2_bwd");
  }
  if (2 * dx->w <= kMaxUpW && (int64_t)dx->n * dx->h < ((int64_t)1 << 30)) {
    const int rows = dx->n * dx->h;
    const int grid = rows < num_sms() * 16 ? rows : num_sms() * 16;
    static const bool win = []() { const char* e = getenv("GSB_UP_BWD_WIN"); return !(e && e[0] == '0'); }();
    if (win) {
      DISPATCH_T(dx->dtype, gsb::launch(upsample2_bwd_win_kernel<T>, gsb::LaunchCfg(grid, 256, 0, to_stream(s)), 
                                (const T*)dout->ptr, dout->sw, dout->n, dout->h, dout->w, dout->c, (T*)dx->ptr, dx->sw, dx->h,
                                dx->w, pad_top, pad_left));
    } else {
      DISPATCH_T(dx->dtype, gsb::launch(upsample2_bwd_rows_kernel<T>, gsb::LaunchCfg(grid, 256, 0, to_stream(s)), 
                                (const T*)dout->ptr, dout->sw, dout->n, dout->h, dout->w, dout->c, (T*)dx->ptr, dx->sw, dx->h,
                                dx->w, pad_top, pad_left));
    }
    return check_launch("upsample2_bwd");
  }
  DISPATCH_T(dx->dtype, gsb::launch(upsample2_bwd_kernel<T>, gsb::LaunchCfg(ew_grid(total), 256, 0, to_stream(s)), 
                            (const T*)dout->ptr, dout->sw, dout->n, dout->h, dout->w, dout->c, (T*)dx->ptr, dx->sw, dx->h,
                            dx->w, pad_top, pad_left));
  return check_launch("upsample2_bwd");
}

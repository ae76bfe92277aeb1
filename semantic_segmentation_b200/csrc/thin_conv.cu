// "Thin" convolutions: one side of the layer has 1..4 channels, so neither a 128-wide tcgen05 tile nor the 64x64
// SIMT implicit-GEMM tile fits, and the shape-generic kernel of conv_direct.cu spends milliseconds on them
// (profiles: 8.8 ms of the 27 ms Pix2Pix step in the generator's outermost 128 -> 1 merged 8x8 transposed conv alone).
//   * thin_gather : many channels -> <= 4 channels   (models_pix2pix/networks.py:590-592 outermost up-cell + Tanh,
//                   :660 PatchGAN output conv, data-gradient of :640 towards the 2..4 input channels)
//   * thin_scatter: <= 4 channels -> many channels   (data-gradients of the two layers above)
//   * thin_wgrad  : weight gradient when dout has <= 4 channels
// All three are bandwidth / FMA-issue kernels with the small operand's weights in shared memory and 128-bit accesses
// on the wide NHWC tensor; the narrow tensor is addressed through its element strides (NCHW fp32 at the API edge).
#include <stdlib.h>

#include "common.cuh"

namespace gsb {

struct TView {
  void* ptr;
  int dtype;
  int N, H, W, C;
  int64_t sn, sh, sw, sc;
};
static TView tv(const gsb_view* v) { return TView{v->ptr, v->dtype, v->n, v->h, v->w, v->c, v->sn, v->sh, v->sw, v->sc}; }

// first valid tap index and step along one axis for destination coordinate o (gather relation of conv_direct.cu)
__device__ __forceinline__ void tap_range(int mode, int o, int stride, int pad, int& k0, int& kstep) {
  if (mode == GSB_MODE_CONV) {
    k0 = 0;
    kstep = 1;
  } else {
    k0 = (o + pad) % stride;
    kstep = stride;
  }
}
__device__ __forceinline__ bool tap_src(int mode, int o, int k, int stride, int pad, int limit, int& i) {
  if (mode == GSB_MODE_CONV) {
    i = o * stride - pad + k;
  } else {
    const int t = o + pad - k;  // divisible by stride by construction of tap_range
    if (t < 0) return false;
    i = t / stride;
  }
  return i >= 0 && i < limit;
}

__device__ __forceinline__ float wload(const void* wp, int wdtype, int64_t i) {
  return wdtype == GSB_BF16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(wp)[i]) : reinterpret_cast<const float*>(wp)[i];
}

// ---------------------------------------------------------------------------------------------------------
// thin_gather: out[p][co] = act(bias[co] + sum_taps sum_ci in[p@tap][ci] * wp[tap][co][ci]),  CO = out.C <= 4.
// L lanes per output pixel, lane l owns the 8-channel slices l, l+L, ...; shuffle reduction.
// ---------------------------------------------------------------------------------------------------------
template <typename T, int CO>
__global__ void __launch_bounds__(256) thin_gather_kernel(TView in, const void* __restrict__ wp, int wdtype,
                                                         const float* __restrict__ bias, TView out, int mode, int KH, int KW,
                                                         int stride, int pad, int act, int L) {
  GSB_PDL_ENTRY();
  extern __shared__ float sw[];  // [taps][CO][Cin]
  const int Cin = in.C, taps = KH * KW;
  for (int i = threadIdx.x; i < taps * CO * Cin; i += blockDim.x) {
    const int ci = i % Cin, co = (i / Cin) % CO, t = i / (Cin * CO);
    sw[i] = co < out.C ? wload(wp, wdtype, ((int64_t)t * out.C + co) * Cin + ci) : 0.f;
  }
  __syncthreads();
  const int slices = Cin >> 3;
  const int l = threadIdx.x % L;
  const unsigned P = (unsigned)out.N * out.H * out.W;
  const unsigned ppb = blockDim.x / L;  // pixels per block iteration
  const T* ip = reinterpret_cast<const T*>(in.ptr);
  // uniform trip count per warp (shuffles below): iterate on the block's first pixel
  for (unsigned p0 = blockIdx.x * ppb; p0 < P; p0 += gridDim.x * ppb) {
    const unsigned p = p0 + threadIdx.x / L;
    const bool valid = p < P;
    float acc[CO];
#pragma unroll
    for (int c = 0; c < CO; ++c) acc[c] = 0.f;
    int ox = 0, oy = 0, n = 0;
    if (valid) {
      ox = (int)(p % (unsigned)out.W);
      const unsigned r = p / (unsigned)out.W;
      oy = (int)(r % (unsigned)out.H);
      n = (int)(r / (unsigned)out.H);
      int ky0, kys, kx0, kxs;
      tap_range(mode, oy, stride, pad, ky0, kys);
      tap_range(mode, ox, stride, pad, kx0, kxs);
      for (int ky = ky0; ky < KH; ky += kys) {
        int iy;
        if (!tap_src(mode, oy, ky, stride, pad, in.H, iy)) continue;
        for (int kx = kx0; kx < KW; kx += kxs) {
          int ix;
          if (!tap_src(mode, ox, kx, stride, pad, in.W, ix)) continue;
          const T* src = ip + (int64_t)n * in.sn + (int64_t)iy * in.sh + (int64_t)ix * in.sw;
          const float* wt = sw + (ky * KW + kx) * CO * Cin;
          for (int s = l; s < slices; s += L) {
            float v[8];
            Vec8<T>::load(src + s * 8, v);
#pragma unroll
            for (int c = 0; c < CO; ++c) {
              const float4 w0 = *reinterpret_cast<const float4*>(wt + c * Cin + s * 8);
              const float4 w1 = *reinterpret_cast<const float4*>(wt + c * Cin + s * 8 + 4);
              acc[c] = fmaf(v[0], w0.x, acc[c]); acc[c] = fmaf(v[1], w0.y, acc[c]);
              acc[c] = fmaf(v[2], w0.z, acc[c]); acc[c] = fmaf(v[3], w0.w, acc[c]);
              acc[c] = fmaf(v[4], w1.x, acc[c]); acc[c] = fmaf(v[5], w1.y, acc[c]);
              acc[c] = fmaf(v[6], w1.z, acc[c]); acc[c] = fmaf(v[7], w1.w, acc[c]);
            }
          }
        }
      }
    }
#pragma unroll
    for (int c = 0; c < CO; ++c)
      for (int o = L >> 1; o > 0; o >>= 1) acc[c] += __shfl_xor_sync(0xffffffffu, acc[c], o);
    if (valid && l == 0) {
#pragma unroll
      for (int c = 0; c < CO; ++c)
        if (c < out.C) {
          const float z = act_fwd(acc[c] + (bias ? bias[c] : 0.f), act);
          st_elem(out.ptr, (int64_t)n * out.sn + (int64_t)oy * out.sh + (int64_t)ox * out.sw + (int64_t)c * out.sc, out.dtype, z);
        }
    }
  }
}

// Register-blocked form: one lane group computes PXB output pixels of one row that share their tap set (neighbours for a
// convolution, pixels `stride` apart -- the same sub-pixel phase -- for a transposed convolution), so every weight vector
// read from shared memory feeds PXB * 8 FMAs instead of 8.  The one-pixel form above issues two LDS.128 per 8 FMAs and is
// bound by shared-memory bandwidth (generator output cell 128 -> 1, merged 8x8: 170 us per launch against ~10 us of HBM
// time, profiles/r01r).
template <typename T, int CO, int PXB>
__global__ void __launch_bounds__(256) thin_gather_px_kernel(TView in, const void* __restrict__ wp, int wdtype,
                                                            const float* __restrict__ bias, TView out, int mode, int KH, int KW,
                                                            int stride, int pad, int act, int L) {
  GSB_PDL_ENTRY();
  extern __shared__ float sw[];  // [taps][CO][Cin]
  const int Cin = in.C, taps = KH * KW;
  for (int i = threadIdx.x; i < taps * CO * Cin; i += blockDim.x) {
    const int ci = i % Cin, co = (i / Cin) % CO, t = i / (Cin * CO);
    sw[i] = co < out.C ? wload(wp, wdtype, ((int64_t)t * out.C + co) * Cin + ci) : 0.f;
  }
  __syncthreads();
  const int slices = Cin >> 3;
  const int l = threadIdx.x % L;
  const bool conv = mode == GSB_MODE_CONV;
  const int spacing = conv ? 1 : stride;   // distance between the output pixels of one item
  const int xstep = conv ? stride : 1;     // distance between the input pixels they read for one tap
  const unsigned wq = ((unsigned)out.W + spacing - 1) / spacing, XB = (wq + PXB - 1) / PXB;
  const unsigned items = (unsigned)out.N * out.H * spacing * XB;
  const unsigned ipb = blockDim.x / L;
  const T* ip = reinterpret_cast<const T*>(in.ptr);
  for (unsigned it0 = blockIdx.x * ipb; it0 < items; it0 += gridDim.x * ipb) {   // uniform trip count (shuffles below)
    const unsigned it = it0 + threadIdx.x / L;
    const bool valid = it < items;
    float acc[PXB][CO];
#pragma unroll
    for (int j = 0; j < PXB; ++j)
#pragma unroll
      for (int c = 0; c < CO; ++c) acc[j][c] = 0.f;
    int ox0 = 0, oy = 0, n = 0;
    if (valid) {
      const unsigned xb = it % XB;
      unsigned r = it / XB;
      const int ph = (int)(r % (unsigned)spacing);
      r /= (unsigned)spacing;
      oy = (int)(r % (unsigned)out.H);
      n = (int)(r / (unsigned)out.H);
      ox0 = ph + spacing * (int)(xb * PXB);
      int ky0, kys, kx0, kxs;
      tap_range(mode, oy, stride, pad, ky0, kys);
      tap_range(mode, ox0, stride, pad, kx0, kxs);     // depends on ox0 only through its phase: shared by the PXB pixels
      for (int ky = ky0; ky < KH; ky += kys) {
        int iy;
        if (!tap_src(mode, oy, ky, stride, pad, in.H, iy)) continue;
        const T* srow = ip + (int64_t)n * in.sn + (int64_t)iy * in.sh;
        for (int kx = kx0; kx < KW; kx += kxs) {
          // input column of pixel 0 (may lie outside; pixel j reads ixb + j * xstep)
          const int ixb = conv ? ox0 * stride - pad + kx : (ox0 + pad - kx) / stride;   // exact: kx = (ox0 + pad) mod stride
          bool ok[PXB];
          bool any = false;
#pragma unroll
          for (int j = 0; j < PXB; ++j) {
            const int ix = ixb + j * xstep;
            ok[j] = (ox0 + j * spacing < out.W) && ix >= 0 && ix < in.W;
            any |= ok[j];
          }
          if (!any) continue;
          const float* wt = sw + (ky * KW + kx) * CO * Cin;
          for (int sidx = l; sidx < slices; sidx += L) {
            float4 w0[CO], w1[CO];
#pragma unroll
            for (int c = 0; c < CO; ++c) {
              w0[c] = *reinterpret_cast<const float4*>(wt + c * Cin + sidx * 8);
              w1[c] = *reinterpret_cast<const float4*>(wt + c * Cin + sidx * 8 + 4);
            }
#pragma unroll
            for (int j = 0; j < PXB; ++j) {
              if (!ok[j]) continue;
              float v[8];
              Vec8<T>::load(srow + (int64_t)(ixb + j * xstep) * in.sw + sidx * 8, v);
#pragma unroll
              for (int c = 0; c < CO; ++c) {
                float a = acc[j][c];
                a = fmaf(v[0], w0[c].x, a); a = fmaf(v[1], w0[c].y, a); a = fmaf(v[2], w0[c].z, a); a = fmaf(v[3], w0[c].w, a);
                a = fmaf(v[4], w1[c].x, a); a = fmaf(v[5], w1[c].y, a); a = fmaf(v[6], w1[c].z, a); a = fmaf(v[7], w1[c].w, a);
                acc[j][c] = a;
              }
            }
          }
        }
      }
    }
#pragma unroll
    for (int j = 0; j < PXB; ++j)
#pragma unroll
      for (int c = 0; c < CO; ++c)
        for (int o = L >> 1; o > 0; o >>= 1) acc[j][c] += __shfl_xor_sync(0xffffffffu, acc[j][c], o);
    if (valid && l == 0) {
#pragma unroll
      for (int j = 0; j < PXB; ++j) {
        const int ox = ox0 + j * spacing;
        if (ox >= out.W) continue;
#pragma unroll
        for (int c = 0; c < CO; ++c)
          if (c < out.C) {
            const float z = act_fwd(acc[j][c] + (bias ? bias[c] : 0.f), act);
            st_elem(out.ptr, (int64_t)n * out.sn + (int64_t)oy * out.sh + (int64_t)ox * out.sw + (int64_t)c * out.sc, out.dtype, z);
          }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// thin_scatter: out[p][c] = act(bias[c] + sum_taps sum_{ci < CI} in[p@tap][ci] * wp[tap][c][ci]),  CI = in.C <= 4,
// out NHWC with c % 8 == 0.  Thread = (pixel, 8-channel group).
// ---------------------------------------------------------------------------------------------------------
template <typename T, int CI>
__global__ void __launch_bounds__(256) thin_scatter_kernel(TView in, const void* __restrict__ wp, int wdtype,
                                                          const float* __restrict__ bias, TView out, int mode, int KH, int KW,
                                                          int stride, int pad, int act) {
  GSB_PDL_ENTRY();
  extern __shared__ float sw[];  // [taps][CI][Cout]
  const int Cout = out.C, taps = KH * KW;
  for (int i = threadIdx.x; i < taps * CI * Cout; i += blockDim.x) {
    const int c = i % Cout, ci = (i / Cout) % CI, t = i / (Cout * CI);
    sw[i] = ci < in.C ? wload(wp, wdtype, ((int64_t)t * Cout + c) * in.C + ci) : 0.f;
  }
  __syncthreads();
  const unsigned G = (unsigned)Cout >> 3;
  const unsigned total = (unsigned)out.N * out.H * out.W * G;
  T* op = reinterpret_cast<T*>(out.ptr);
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const unsigned p = i / G, g = i - p * G;
    const int ox = (int)(p % (unsigned)out.W);
    const unsigned r = p / (unsigned)out.W;
    const int oy = (int)(r % (unsigned)out.H), n = (int)(r / (unsigned)out.H);
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = bias ? bias[g * 8 + j] : 0.f;
    int ky0, kys, kx0, kxs;
    tap_range(mode, oy, stride, pad, ky0, kys);
    tap_range(mode, ox, stride, pad, kx0, kxs);
    for (int ky = ky0; ky < KH; ky += kys) {
      int iy;
      if (!tap_src(mode, oy, ky, stride, pad, in.H, iy)) continue;
      for (int kx = kx0; kx < KW; kx += kxs) {
        int ix;
        if (!tap_src(mode, ox, kx, stride, pad, in.W, ix)) continue;
        const int64_t base = (int64_t)n * in.sn + (int64_t)iy * in.sh + (int64_t)ix * in.sw;
        const float* wt = sw + (ky * KW + kx) * CI * Cout + g * 8;
#pragma unroll
        for (int ci = 0; ci < CI; ++ci) {
          if (ci < in.C) {
            const float v = ld_elem(in.ptr, base + (int64_t)ci * in.sc, in.dtype);
            const float4 w0 = *reinterpret_cast<const float4*>(wt + ci * Cout);
            const float4 w1 = *reinterpret_cast<const float4*>(wt + ci * Cout + 4);
            acc[0] = fmaf(v, w0.x, acc[0]); acc[1] = fmaf(v, w0.y, acc[1]);
            acc[2] = fmaf(v, w0.z, acc[2]); acc[3] = fmaf(v, w0.w, acc[3]);
            acc[4] = fmaf(v, w1.x, acc[4]); acc[5] = fmaf(v, w1.y, acc[5]);
            acc[6] = fmaf(v, w1.z, acc[6]); acc[7] = fmaf(v, w1.w, acc[7]);
          }
        }
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = act_fwd(acc[j], act);
    Vec8<T>::store(op + (int64_t)n * out.sn + (int64_t)oy * out.sh + (int64_t)ox * out.sw + g * 8, acc);
  }
}

// Register-blocked thin_scatter (same idea as thin_gather_px_kernel): thread = (PXB output pixels of one row that share
// their tap set, 8-channel group); a weight vector read from shared memory feeds PXB * 8 FMAs.
template <typename T, int CI, int PXB>
__global__ void __launch_bounds__(256) thin_scatter_px_kernel(TView in, const void* __restrict__ wp, int wdtype,
                                                             const float* __restrict__ bias, TView out, int mode, int KH, int KW,
                                                             int stride, int pad, int act) {
  GSB_PDL_ENTRY();
  extern __shared__ float sw[];  // [taps][CI][Cout]
  const int Cout = out.C, taps = KH * KW;
  for (int i = threadIdx.x; i < taps * CI * Cout; i += blockDim.x) {
    const int c = i % Cout, ci = (i / Cout) % CI, t = i / (Cout * CI);
    sw[i] = ci < in.C ? wload(wp, wdtype, ((int64_t)t * Cout + c) * in.C + ci) : 0.f;
  }
  __syncthreads();
  const bool conv = mode == GSB_MODE_CONV;
  const int spacing = conv ? 1 : stride, xstep = conv ? stride : 1;
  const unsigned G = (unsigned)Cout >> 3;
  const unsigned wq = ((unsigned)out.W + spacing - 1) / spacing, XB = (wq + PXB - 1) / PXB;
  const unsigned total = (unsigned)out.N * out.H * spacing * XB * G;
  T* op = reinterpret_cast<T*>(out.ptr);
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const unsigned it = i / G, g = i - it * G;
    const unsigned xb = it % XB;
    unsigned r = it / XB;
    const int ph = (int)(r % (unsigned)spacing);
    r /= (unsigned)spacing;
    const int oy = (int)(r % (unsigned)out.H), n = (int)(r / (unsigned)out.H);
    const int ox0 = ph + spacing * (int)(xb * PXB);
    if (ox0 >= out.W) continue;
    float acc[PXB][8];
#pragma unroll
    for (int j = 0; j < PXB; ++j)
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[j][e] = bias ? bias[g * 8 + e] : 0.f;
    int ky0, kys, kx0, kxs;
    tap_range(mode, oy, stride, pad, ky0, kys);
    tap_range(mode, ox0, stride, pad, kx0, kxs);
    for (int ky = ky0; ky < KH; ky += kys) {
      int iy;
      if (!tap_src(mode, oy, ky, stride, pad, in.H, iy)) continue;
      const int64_t rbase = (int64_t)n * in.sn + (int64_t)iy * in.sh;
      for (int kx = kx0; kx < KW; kx += kxs) {
        const int ixb = conv ? ox0 * stride - pad + kx : (ox0 + pad - kx) / stride;   // exact for the transposed form
        const float* wt = sw + (ky * KW + kx) * CI * Cout + g * 8;
#pragma unroll
        for (int ci = 0; ci < CI; ++ci) {
          if (ci < in.C) {
            float v[PXB];
            bool any = false;
#pragma unroll
            for (int j = 0; j < PXB; ++j) {
              const int ix = ixb + j * xstep;
              const bool ok = (ox0 + j * spacing < out.W) && ix >= 0 && ix < in.W;
              v[j] = ok ? ld_elem(in.ptr, rbase + (int64_t)ix * in.sw + (int64_t)ci * in.sc, in.dtype) : 0.f;
              any |= ok;
            }
            if (!any) continue;
            const float4 w0 = *reinterpret_cast<const float4*>(wt + ci * Cout);
            const float4 w1 = *reinterpret_cast<const float4*>(wt + ci * Cout + 4);
#pragma unroll
            for (int j = 0; j < PXB; ++j) {
              acc[j][0] = fmaf(v[j], w0.x, acc[j][0]); acc[j][1] = fmaf(v[j], w0.y, acc[j][1]);
              acc[j][2] = fmaf(v[j], w0.z, acc[j][2]); acc[j][3] = fmaf(v[j], w0.w, acc[j][3]);
              acc[j][4] = fmaf(v[j], w1.x, acc[j][4]); acc[j][5] = fmaf(v[j], w1.y, acc[j][5]);
              acc[j][6] = fmaf(v[j], w1.z, acc[j][6]); acc[j][7] = fmaf(v[j], w1.w, acc[j][7]);
            }
          }
        }
      }
    }
#pragma unroll
    for (int j = 0; j < PXB; ++j) {
      const int ox = ox0 + j * spacing;
      if (ox >= out.W) continue;
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[j][e] = act_fwd(acc[j][e], act);
      Vec8<T>::store(op + (int64_t)n * out.sn + (int64_t)oy * out.sh + (int64_t)ox * out.sw + g * 8, acc[j]);
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// thin_wgrad: dwp[tap][co][ci] += sum over pixels of dout[q][co] * in[p][ci],  CO = dout.C <= 4 (dwp is [taps][CO][Cin]).
// Thread = (pixel lane over the WIDE tensor `in`, 8-channel group); TP taps of one kernel row band per pass
// (blockIdx.y = pass) accumulate in registers; block reduction in shared memory, one atomic per output and block.
// Relation (forward layer): CONV  q = (p + pad - k) / stride ;  CONVT  q = p * stride - pad + k.
// ---------------------------------------------------------------------------------------------------------
// KWT > 0: kernel width known at compile time (4 or 8: the PatchGAN head and the merged 8x8 up-convolution) -- the tap ->
// (ky, kx) split of the unrolled tap loop is then constant and the per-tap index arithmetic collapses to two adds.
// Accumulation in packed fp32x2 (FFMA2): the kernel is bound by the FMA pipe (128 ci x 64 taps per pixel, 516 us per launch
// for the generator's output layer, ncu profiles/r01o).
template <typename T, int TP, int KWT>
__global__ void __launch_bounds__(256) thin_wgrad_kernel(TView in, TView dout, float* __restrict__ dwp, int64_t slice_elems,
                                                        int mode, int KH, int KW, int stride, int pad, int co) {
  GSB_PDL_ENTRY();
  extern __shared__ float sacc[];  // [TP][Cin]
  const int Cin = in.C, taps = KH * KW;
  const int tap0 = blockIdx.y * TP;
  for (int i = threadIdx.x; i < TP * Cin; i += blockDim.x) sacc[i] = 0.f;
  __syncthreads();
  const unsigned G = (unsigned)Cin >> 3, lanes = blockDim.x / G;
  const unsigned g = threadIdx.x % G, lane = threadIdx.x / G;
  float2 acc[TP][4];
#pragma unroll
  for (int t = 0; t < TP; ++t)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[t][j] = make_float2(0.f, 0.f);
  const unsigned P = (unsigned)in.N * in.H * in.W;
  const T* ip = reinterpret_cast<const T*>(in.ptr);
  const int ky0 = KWT > 0 ? tap0 / KWT : 0;   // TP is a multiple of KWT: a pass covers whole kernel rows
  for (unsigned p = blockIdx.x * lanes + lane; p < P; p += gridDim.x * lanes) {
    const int ix = (int)(p % (unsigned)in.W);
    const unsigned r = p / (unsigned)in.W;
    const int iy = (int)(r % (unsigned)in.H), n = (int)(r / (unsigned)in.H);
    float v[8];
    Vec8<T>::load(ip + (int64_t)n * in.sn + (int64_t)iy * in.sh + (int64_t)ix * in.sw + g * 8, v);
    float2 v2[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) v2[j] = make_float2(v[2 * j], v[2 * j + 1]);
    const int64_t dbase = (int64_t)n * dout.sn + (int64_t)co * dout.sc;
#pragma unroll
    for (int t = 0; t < TP; ++t) {
      const int tap = tap0 + t;
      if (tap < taps) {
        const int ky = KWT > 0 ? ky0 + t / KWT : tap / KW;
        const int kx = KWT > 0 ? t % KWT : tap - ky * KW;
        int qy, qx;
        bool ok;
        if (mode == GSB_MODE_CONV) {
          const int ty = iy + pad - ky, tx = ix + pad - kx;
          qy = ty / stride; qx = tx / stride;
          ok = ty >= 0 && tx >= 0 && qy * stride == ty && qx * stride == tx && qy < dout.H && qx < dout.W;
        } else {
          qy = iy * stride - pad + ky; qx = ix * stride - pad + kx;
          ok = qy >= 0 && qx >= 0 && qy < dout.H && qx < dout.W;
        }
        if (ok) {
          const float d = ld_elem(dout.ptr, dbase + (int64_t)qy * dout.sh + (int64_t)qx * dout.sw, dout.dtype);
          const float2 d2 = make_float2(d, d);
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[t][j] = __ffma2_rn(d2, v2[j], acc[t][j]);
        }
      }
    }
  }
  // deterministic reduction: the pixel lanes add their registers into shared memory one lane after the other, and the
  // CTA STORES its sums into its own slice (blockIdx.x) of the workspace; the caller adds the slices in slice order
  for (unsigned l = 0; l < lanes; ++l) {
    if (lane == l) {
#pragma unroll
      for (int t = 0; t < TP; ++t)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          sacc[t * Cin + g * 8 + 2 * j] += acc[t][j].x;
          sacc[t * Cin + g * 8 + 2 * j + 1] += acc[t][j].y;
        }
    }
    __syncthreads();
  }
  float* __restrict__ slice = dwp + (int64_t)blockIdx.x * slice_elems;
  for (int i = threadIdx.x; i < TP * Cin; i += blockDim.x) {
    const int t = i / Cin, ci = i - t * Cin;
    if (tap0 + t < taps) slice[((int64_t)(tap0 + t) * dout.C + co) * Cin + ci] = sacc[i];
  }
}

static bool wide_ok(const gsb_view* v) {
  return v->sc == 1 && v->c % 8 == 0 && v->c >= 8 && v->sw % 8 == 0 && v->sh % 8 == 0 && v->sn % 8 == 0 &&
         ((uintptr_t)v->ptr % 16 == 0) && (v->dtype == GSB_BF16 || v->dtype == GSB_F32);
}
static int pow2_floor(int v) {
  int p = 1;
  while (p * 2 <= v) p *= 2;
  return p;
}
static int grid_sm(int64_t want, int per_sm) {
  int64_t cap = (int64_t)num_sms() * per_sm;
  return (int)(want < 1 ? 1 : (want > cap ? cap : want));
}

// Returns 1 when launched, 0 when the shape is not thin (caller continues with the generic kernel), -1 on error.
int thin_gather_try(const gsb_view* in, const void* wp, int wdtype, const float* bias, const gsb_view* out,
                    const gsb_conv_geom* g, int act, int accumulate, cudaStream_t st) {
  if (accumulate || g->kh * g->kw > 64) return 0;
  const int64_t P = num_pixels(out);
  if (P <= 0 || P >= ((int64_t)1 << 28)) return 0;
  const int taps = g->kh * g->kw;
  if (out->c <= 4 && wide_ok(in)) {
    const int CO = out->c <= 1 ? 1 : (out->c <= 2 ? 2 : 4);
    const size_t smem = (size_t)taps * CO * in->c * sizeof(float);
    if (smem > 96 * 1024) return 0;
    int L = pow2_floor(in->c / 8);
    if (L > 16) L = 16;
    const int ppb = 256 / L;
    TView vi = tv(in), vo = tv(out);
    static const bool px_on = []() { const char* e = getenv("GSB_THIN_PX"); return !(e && e[0] == '0'); }();
    constexpr int PXB = 4;
    const int spacing = g->mode == GSB_MODE_CONV ? 1 : g->stride;
    const int64_t wq = (out->w + spacing - 1) / spacing, xb = (wq + PXB - 1) / PXB;
    const int64_t items = (int64_t)out->n * out->h * spacing * xb;
    // small outputs (the PatchGAN head: 8 x 30 x 30 pixels = 120 CTAs of the blocked kernel, 81 us of latency): one pixel per
    // lane group fills the GPU four times better
    if (px_on && (items + ppb - 1) / ppb >= num_sms()) {
      // register-blocked kernel: 4 pixels of one row and phase per lane group
      const int pgrid = grid_sm((items + ppb - 1) / ppb, 8);
#define GSB_THIN_GP(T, C)                                                                                             \
  do {                                                                                                                \
    cudaFuncSetAttribute(thin_gather_px_kernel<T, C, PXB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);   \
    gsb::launch(thin_gather_px_kernel<T, C, PXB>, gsb::LaunchCfg(pgrid, 256, smem, st), vi, wp, wdtype, bias, vo, g->mode, g->kh, g->kw,  \
                g->stride, g->pad, act, L);                                                                           \
  } while (0)
      if (in->dtype == GSB_BF16) {
        if (CO == 1) GSB_THIN_GP(__nv_bfloat16, 1); else if (CO == 2) GSB_THIN_GP(__nv_bfloat16, 2); else GSB_THIN_GP(__nv_bfloat16, 4);
      } else {
        if (CO == 1) GSB_THIN_GP(float, 1); else if (CO == 2) GSB_THIN_GP(float, 2); else GSB_THIN_GP(float, 4);
      }
#undef GSB_THIN_GP
      return check_launch("thin_gather_px") ? -1 : 1;
    }
    const int grid = grid_sm((P + ppb - 1) / ppb, 8);
#define GSB_THIN_G(T, C)                                                                                              \
  do {                                                                                                                \
    cudaFuncSetAttribute(thin_gather_kernel<T, C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);           \
    gsb::launch(thin_gather_kernel<T, C>, gsb::LaunchCfg(grid, 256, smem, st), vi, wp, wdtype, bias, vo, g->mode, g->kh, g->kw, g->stride,     \
                                                      g->pad, act, L);                                                \
  } while (0)
    if (in->dtype == GSB_BF16) {
      if (CO == 1) GSB_THIN_G(__nv_bfloat16, 1); else if (CO == 2) GSB_THIN_G(__nv_bfloat16, 2); else GSB_THIN_G(__nv_bfloat16, 4);
    } else {
      if (CO == 1) GSB_THIN_G(float, 1); else if (CO == 2) GSB_THIN_G(float, 2); else GSB_THIN_G(float, 4);
    }
#undef GSB_THIN_G
    return check_launch("thin_gather") ? -1 : 1;
  }
  if (in->c <= 4 && wide_ok(out)) {
    const int CI = in->c <= 1 ? 1 : (in->c <= 2 ? 2 : 4);
    const size_t smem = (size_t)taps * CI * out->c * sizeof(float);
    if (smem > 96 * 1024) return 0;
    const int64_t total = P * (out->c / 8);
    if (total >= ((int64_t)1 << 31)) return 0;
    TView vi = tv(in), vo = tv(out);
    static const bool spx_on = []() { const char* e = getenv("GSB_THIN_SPX"); return !(e && e[0] == '0'); }();   // on by default (validated r01s)
    if (spx_on) {
      constexpr int PXB = 4;
      const int spacing = g->mode == GSB_MODE_CONV ? 1 : g->stride;
      const int64_t wq = (out->w + spacing - 1) / spacing, xb = (wq + PXB - 1) / PXB;
      const int64_t ptotal = (int64_t)out->n * out->h * spacing * xb * (out->c / 8);
      const int pgrid = grid_sm((ptotal + 255) / 256, 8);
#define GSB_THIN_SP(T, C)                                                                                             \
  do {                                                                                                                \
    cudaFuncSetAttribute(thin_scatter_px_kernel<T, C, PXB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);  \
    gsb::launch(thin_scatter_px_kernel<T, C, PXB>, gsb::LaunchCfg(pgrid, 256, smem, st), vi, wp, wdtype, bias, vo, g->mode, g->kh, g->kw, \
                g->stride, g->pad, act);                                                                              \
  } while (0)
      if (out->dtype == GSB_BF16) {
        if (CI == 1) GSB_THIN_SP(__nv_bfloat16, 1); else if (CI == 2) GSB_THIN_SP(__nv_bfloat16, 2); else GSB_THIN_SP(__nv_bfloat16, 4);
      } else {
        if (CI == 1) GSB_THIN_SP(float, 1); else if (CI == 2) GSB_THIN_SP(float, 2); else GSB_THIN_SP(float, 4);
      }
#undef GSB_THIN_SP
      return check_launch("thin_scatter_px") ? -1 : 1;
    }
    const int grid = grid_sm((total + 255) / 256, 8);
#define GSB_THIN_S(T, C)                                                                                              \
  do {                                                                                                                \
    cudaFuncSetAttribute(thin_scatter_kernel<T, C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);          \
    gsb::launch(thin_scatter_kernel<T, C>, gsb::LaunchCfg(grid, 256, smem, st), vi, wp, wdtype, bias, vo, g->mode, g->kh, g->kw, g->stride,    \
                                                       g->pad, act);                                                  \
  } while (0)
    if (out->dtype == GSB_BF16) {
      if (CI == 1) GSB_THIN_S(__nv_bfloat16, 1); else if (CI == 2) GSB_THIN_S(__nv_bfloat16, 2); else GSB_THIN_S(__nv_bfloat16, 4);
    } else {
      if (CI == 1) GSB_THIN_S(float, 1); else if (CI == 2) GSB_THIN_S(float, 2); else GSB_THIN_S(float, 4);
    }
#undef GSB_THIN_S
    return check_launch("thin_scatter") ? -1 : 1;
  }
  return 0;
}

// ws == nullptr: plan only.  *used = slices written (one per CTA column, each completely).
int thin_wgrad_try(const gsb_view* in, const gsb_view* dout, float* dwp, int max_slices, int* used, const gsb_conv_geom* g,
                   cudaStream_t st) {
  if (!(dout->c <= 4 && wide_ok(in)) || in->c > 1024 || 256 % (in->c / 8) != 0) return 0;
  const int64_t P = num_pixels(in);
  if (P <= 0 || P >= ((int64_t)1 << 31)) return 0;
  // taps per pass: 8 (one row of the merged 8x8 up-convolution: 64 accumulators, two blocks per SM) or 16
  const int TPr = (g->kw == 8 && in->dtype == GSB_BF16) ? 8 : 16;
  const int taps = g->kh * g->kw, passes = (taps + TPr - 1) / TPr;
  const int G = in->c / 8, lanes = 256 / G;
  int grid = grid_sm((P + lanes * 16 - 1) / (lanes * 16), 2);
  if (grid < num_sms()) grid = grid_sm((P + lanes * 4 - 1) / (lanes * 4), 2);   // small inputs: shorter pixel loops, more CTAs
  if (dwp != nullptr && grid > max_slices) grid = max_slices;
  *used = grid;
  if (dwp == nullptr) return 1;
  GSB_REQUIRE(grid >= 1, "thin_wgrad: no workspace slices");
  const int64_t slice_elems = (int64_t)taps * dout->c * in->c;
  const size_t smem = (size_t)TPr * in->c * sizeof(float);
  TView vi = tv(in), vd = tv(dout);
  for (int co = 0; co < dout->c; ++co) {
#define GSB_THIN_WG(TT, TPV, KWT)                                                                                          \
  gsb::launch(thin_wgrad_kernel<TT, TPV, KWT>, gsb::LaunchCfg(dim3(grid, passes), 256, smem, st), vi, vd, dwp, slice_elems, g->mode, g->kh, g->kw, g->stride, g->pad, co)
    if (in->dtype == GSB_BF16) {
      if (g->kw == 8) GSB_THIN_WG(__nv_bfloat16, 8, 8);
      else if (g->kw == 4) GSB_THIN_WG(__nv_bfloat16, 16, 4);
      else GSB_THIN_WG(__nv_bfloat16, 16, 0);
    } else {
      GSB_THIN_WG(float, 16, 0);
    }
#undef GSB_THIN_WG
  }
  return check_launch("thin_wgrad", dout->c) ? -1 : 1;
}

}  // namespace gsb

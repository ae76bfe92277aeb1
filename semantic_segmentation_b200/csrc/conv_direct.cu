// SIMT implicit-GEMM convolution: the shape-generic path of libgenseg_b200.
//
// One gather primitive covers nn.Conv2d / nn.ConvTranspose2d forward and data-gradient
// (unet/unet_parts.py:16,19,53,74; models_pix2pix/networks.py:582,640-660;
// architecture_pix2pix/operations.py:14-39), with arbitrary element strides so NCHW fp32
// tensors at the API edge (Cin = 1..4, Cout = 1..3) and channel slices of NHWC concat buffers
// are read/written in place.  It is the production kernel for the layers a 128-wide
// tcgen05 tile cannot fill (first/last convs) and the fp32 validation mode for all layers;
// the hot 3x3/4x4/8x8 layers run in conv_tc.cu.
#include "common.cuh"

namespace gsb {

struct DView {
  void* ptr;
  int dtype;
  int N, H, W, C;
  int64_t sn, sh, sw, sc;
};
static DView dv(const gsb_view* v) { return DView{v->ptr, v->dtype, v->n, v->h, v->w, v->c, v->sn, v->sh, v->sw, v->sc}; }

constexpr int BK = 16;

// source coordinate of tap k for destination coordinate o; returns false when the tap does not hit
__device__ __forceinline__ bool src_coord(int mode, int o, int k, int stride, int pad, int limit, int& i) {
  if (mode == GSB_MODE_CONV) {
    i = o * stride - pad + k;
    return i >= 0 && i < limit;
  }
  int t = o + pad - k;
  if (t < 0) return false;
  i = t / stride;
  return (t - i * stride) == 0 && i < limit;
}

template <int BM, int BN, int TM, int TN>
__global__ void __launch_bounds__(256) conv_gather_kernel(DView in, const void* __restrict__ wp, int wdtype,
                                                          const float* __restrict__ bias, DView out, int mode, int KH,
                                                          int KW, int stride, int pad, int act, int accumulate,
                                                          int sub, int py, int px) {
  GSB_PDL_ENTRY();
  static_assert((BM / TM) * (BN / TN) == 256, "tile/thread mismatch");
  constexpr int AR = BM / 16;                    // A rows loaded per thread
  constexpr int BR = (BN + 15) / 16;             // B rows loaded per thread
  __shared__ float As[BK][BM + 4];
  __shared__ float Bs[BK][BN + 4];
  const int tid = threadIdx.x;
  const int tx = tid % (BN / TN), ty = tid / (BN / TN);
  const int QH = (out.H - py + sub - 1) / sub, QW = (out.W - px + sub - 1) / sub;
  const int64_t M = (int64_t)out.N * QH * QW;
  const int64_t m0 = (int64_t)blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  const int Cin = in.C, Cout = out.C;
  const int kl = tid % 16, rl = tid / 16;

  int an[AR], aoy[AR], aox[AR];
#pragma unroll
  for (int i = 0; i < AR; ++i) {
    int64_t m = m0 + rl + i * 16;
    if (m < M) {
      int qx = (int)(m % QW);
      int64_t r = m / QW;
      int qy = (int)(r % QH);
      an[i] = (int)(r / QH);
      aoy[i] = qy * sub + py;
      aox[i] = qx * sub + px;
    } else {
      an[i] = -1;
      aoy[i] = aox[i] = 0;
    }
  }
  // two-level accumulation (this is the kernel of the fp32 validation mode): fp32 FMAs inside a BK-channel block, fp64
  // across blocks and taps, so a K = 8192 reduction is as accurate as torch's blocked CPU kernel (a plain fp32 running sum
  // was 3x further from the fp64 reference and flipped 3x more ReLU branches in whole-network gradient comparisons)
  double acc[TM][TN];
#pragma unroll
  for (int a = 0; a < TM; ++a)
#pragma unroll
    for (int b = 0; b < TN; ++b) acc[a][b] = 0.0;

  // transposed conv with sub == stride: only taps with k == (o + pad) mod stride contribute
  int ky0 = 0, kx0 = 0, kstep = 1;
  if (mode == GSB_MODE_CONVT && sub == stride && stride > 1) {
    ky0 = (py + pad) % stride;
    kx0 = (px + pad) % stride;
    kstep = stride;
  }
  for (int ky = ky0; ky < KH; ky += kstep) {
    for (int kx = kx0; kx < KW; kx += kstep) {
      const int tap = ky * KW + kx;
      int64_t aoff[AR];
      bool aval[AR];
#pragma unroll
      for (int i = 0; i < AR; ++i) {
        int iy = 0, ix = 0;
        bool v = an[i] >= 0 && src_coord(mode, aoy[i], ky, stride, pad, in.H, iy) &&
                 src_coord(mode, aox[i], kx, stride, pad, in.W, ix);
        aval[i] = v;
        aoff[i] = v ? (int64_t)an[i] * in.sn + (int64_t)iy * in.sh + (int64_t)ix * in.sw : 0;
      }
      for (int c0 = 0; c0 < Cin; c0 += BK) {
        const int ci = c0 + kl;
#pragma unroll
        for (int i = 0; i < AR; ++i) {
          float v = 0.f;
          if (aval[i] && ci < Cin) v = ld_elem(in.ptr, aoff[i] + (int64_t)ci * in.sc, in.dtype);
          As[kl][rl + i * 16] = v;
        }
#pragma unroll
        for (int j = 0; j < BR; ++j) {
          int nl = rl + j * 16;
          if (nl < BN) {
            int co = n0 + nl;
            float v = 0.f;
            if (co < Cout && ci < Cin) v = ld_elem(wp, ((int64_t)tap * Cout + co) * Cin + ci, wdtype);
            Bs[kl][nl] = v;
          }
        }
        __syncthreads();
        float part[TM][TN];
#pragma unroll
        for (int i = 0; i < TM; ++i)
#pragma unroll
          for (int j = 0; j < TN; ++j) part[i][j] = 0.f;
#pragma unroll
        for (int k = 0; k < BK; ++k) {
          float a[TM], b[TN];
#pragma unroll
          for (int i = 0; i < TM; ++i) a[i] = As[k][ty * TM + i];
#pragma unroll
          for (int j = 0; j < TN; ++j) b[j] = Bs[k][tx * TN + j];
#pragma unroll
          for (int i = 0; i < TM; ++i)
#pragma unroll
            for (int j = 0; j < TN; ++j) part[i][j] = fmaf(a[i], b[j], part[i][j]);
        }
#pragma unroll
        for (int i = 0; i < TM; ++i)
#pragma unroll
          for (int j = 0; j < TN; ++j) acc[i][j] += (double)part[i][j];
        __syncthreads();
      }
    }
  }
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    int64_t m = m0 + ty * TM + i;
    if (m >= M) continue;
    int qx = (int)(m % QW);
    int64_t r = m / QW;
    int qy = (int)(r % QH);
    int n = (int)(r / QH);
    int64_t obase = (int64_t)n * out.sn + (int64_t)(qy * sub + py) * out.sh + (int64_t)(qx * sub + px) * out.sw;
#pragma unroll
    for (int j = 0; j < TN; ++j) {
      int co = n0 + tx * TN + j;
      if (co >= Cout) continue;
      float v = (float)acc[i][j];
      if (bias != nullptr) v += bias[co];
      int64_t o = obase + (int64_t)co * out.sc;
      if (accumulate) v += ld_elem(out.ptr, o, out.dtype);
      st_elem(out.ptr, o, out.dtype, act_fwd(v, act));
    }
  }
}

// slice[blockIdx.z][tap][co][ci] = sum over this split's driver pixels of dout[.][co] * in[.][ci]  (plain stores: every
// slice is written completely, the caller adds the slices in slice order -> deterministic, no atomics).
// Driver = the lower-resolution side: output pixels for CONV, input pixels for CONVT.
template <int BM, int BN, int TM, int TN>
__global__ void __launch_bounds__(256) conv_wgrad_kernel(DView in, DView dout, float* __restrict__ dwp, int64_t slice_elems,
                                                         int mode, int KH, int KW, int stride, int pad, int64_t chunk,
                                                         int tiles_n) {
  GSB_PDL_ENTRY();
  static_assert((BM / TM) * (BN / TN) == 256, "tile/thread mismatch");
  __shared__ float As[BK][BM + 4];  // dout  [k][co]
  __shared__ float Bs[BK][BN + 4];  // in    [k][ci]
  const int tid = threadIdx.x;
  const int tx = tid % (BN / TN), ty = tid / (BN / TN);
  const int Cin = in.C, Cout = dout.C;
  const int co0 = (blockIdx.x / tiles_n) * BM, ci0 = (blockIdx.x % tiles_n) * BN;
  const int tap = blockIdx.y, ky = tap / KW, kx = tap % KW;
  const bool conv = mode == GSB_MODE_CONV;
  const int DH = conv ? dout.H : in.H, DW = conv ? dout.W : in.W, DN = in.N;
  const int64_t P = (int64_t)DN * DH * DW;
  const int64_t p_begin = (int64_t)blockIdx.z * chunk;
  const int64_t p_end = p_begin + chunk < P ? p_begin + chunk : P;
  // Two-level accumulation: fp32 FMAs inside a BK-pixel block, fp64 across blocks.  A weight gradient behind BatchNorm is a
  // sum of thousands of products that nearly cancel (sum_p dY[p] = 0 per channel), so a plain fp32 running sum left the
  // fp32 validation mode 1e-3 away from the fp64 reference where torch's blocked CPU kernel is at 1e-5.
  double acc[TM][TN];
#pragma unroll
  for (int a = 0; a < TM; ++a)
#pragma unroll
    for (int b = 0; b < TN; ++b) acc[a][b] = 0.0;

  __shared__ int64_t s_in[BK], s_out[BK];
  for (int64_t pb = p_begin; pb < p_end; pb += BK) {
    if (tid < BK) {
      int64_t p = pb + tid;
      int64_t oi = -1, oo = -1;
      if (p < p_end) {
        int x = (int)(p % DW);
        int64_t r = p / DW;
        int y = (int)(r % DH);
        int n = (int)(r / DH);
        int iy, ix, oy, ox;
        bool ok;
        if (conv) {
          oy = y; ox = x;
          iy = oy * stride - pad + ky;
          ix = ox * stride - pad + kx;
          ok = iy >= 0 && iy < in.H && ix >= 0 && ix < in.W;
        } else {
          iy = y; ix = x;
          oy = iy * stride - pad + ky;
          ox = ix * stride - pad + kx;
          ok = oy >= 0 && oy < dout.H && ox >= 0 && ox < dout.W;
        }
        if (ok) {
          oi = (int64_t)n * in.sn + (int64_t)iy * in.sh + (int64_t)ix * in.sw;
          oo = (int64_t)n * dout.sn + (int64_t)oy * dout.sh + (int64_t)ox * dout.sw;
        }
      }
      s_in[tid] = oi;
      s_out[tid] = oo;
    }
    __syncthreads();
    for (int idx = tid; idx < BK * BM; idx += 256) {
      int k = idx / BM, c = idx % BM;
      int co = co0 + c;
      float v = 0.f;
      if (s_out[k] >= 0 && co < Cout) v = ld_elem(dout.ptr, s_out[k] + (int64_t)co * dout.sc, dout.dtype);
      As[k][c] = v;
    }
    for (int idx = tid; idx < BK * BN; idx += 256) {
      int k = idx / BN, c = idx % BN;
      int ci = ci0 + c;
      float v = 0.f;
      if (s_in[k] >= 0 && ci < Cin) v = ld_elem(in.ptr, s_in[k] + (int64_t)ci * in.sc, in.dtype);
      Bs[k][c] = v;
    }
    __syncthreads();
    float part[TM][TN];
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
      for (int j = 0; j < TN; ++j) part[i][j] = 0.f;
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      float a[TM], b[TN];
#pragma unroll
      for (int i = 0; i < TM; ++i) a[i] = As[k][ty * TM + i];
#pragma unroll
      for (int j = 0; j < TN; ++j) b[j] = Bs[k][tx * TN + j];
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) part[i][j] = fmaf(a[i], b[j], part[i][j]);
    }
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
      for (int j = 0; j < TN; ++j) acc[i][j] += (double)part[i][j];
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    int co = co0 + ty * TM + i;
    if (co >= Cout) continue;
#pragma unroll
    for (int j = 0; j < TN; ++j) {
      int ci = ci0 + tx * TN + j;
      if (ci >= Cin) continue;
      dwp[(int64_t)blockIdx.z * slice_elems + ((int64_t)tap * Cout + co) * Cin + ci] = (float)acc[i][j];
    }
  }
}

// One CTA per channel; fixed summation order (thread-strided fp64 partials, shuffle tree, 8 warp sums in index order):
// deterministic, and the result is STORED (the earlier form added partial sums of several CTAs with atomics).
__global__ void bias_grad_kernel(DView d, float* __restrict__ db) {
  GSB_PDL_ENTRY();
  const int c = blockIdx.x;
  const int64_t P = (int64_t)d.N * d.H * d.W;
  double acc = 0.0;   // bias gradients are cancellation-heavy sums: accumulate in fp64
  for (int64_t p = threadIdx.x; p < P; p += blockDim.x) {
    int x = (int)(p % d.W);
    int64_t r = p / d.W;
    int y = (int)(r % d.H);
    int n = (int)(r / d.H);
    acc += ld_elem(d.ptr, (int64_t)n * d.sn + (int64_t)y * d.sh + (int64_t)x * d.sw + (int64_t)c * d.sc, d.dtype);
  }
  __shared__ double red[8];
  acc = warp_sum_d(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    double v = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.0;
    v = warp_sum_d(v);
    if (threadIdx.x == 0) db[c] = (float)v;
  }
}

// Any other view (the 1..3-channel fp32 NCHW network outputs): grid = (pixel chunks, channels); a CTA adds a strided
// pixel range of one channel in fp64 and writes TWO partial rows (hi + lo floats); gsb::sum_slices adds the rows in row
// order.  (One CTA per channel took 1.3 ms for the generator's 8 x 1 x 256 x 256 output: profiles/r02m_p2p_launch_summary.)
__global__ void __launch_bounds__(256) bias_grad_rows_kernel(DView d, float* __restrict__ partial) {
  GSB_PDL_ENTRY();
  const int c = blockIdx.y, C = gridDim.y;
  const int64_t P = (int64_t)d.N * d.H * d.W;
  double acc = 0.0;
  for (int64_t p = (int64_t)blockIdx.x * 256 + threadIdx.x; p < P; p += (int64_t)gridDim.x * 256) {
    int x, y, n;
    if ((p >> 31) == 0) {
      const unsigned u = (unsigned)p, r = u / (unsigned)d.W;
      x = (int)(u - r * (unsigned)d.W);
      n = (int)(r / (unsigned)d.H);
      y = (int)(r - (unsigned)n * (unsigned)d.H);
    } else {
      x = (int)(p % d.W);
      const int64_t r = p / d.W;
      y = (int)(r % d.H);
      n = (int)(r / d.H);
    }
    acc += ld_elem(d.ptr, (int64_t)n * d.sn + (int64_t)y * d.sh + (int64_t)x * d.sw + (int64_t)c * d.sc, d.dtype);
  }
  __shared__ double red[8];
  acc = warp_sum_d(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    double v = threadIdx.x < 8 ? red[threadIdx.x] : 0.0;
    v = warp_sum_d(v);
    if (threadIdx.x == 0) {
      const float hi = (float)v;
      partial[((int64_t)blockIdx.x * 2) * C + c] = hi;
      partial[((int64_t)blockIdx.x * 2 + 1) * C + c] = (float)(v - (double)hi);
    }
  }
}

// NHWC views with C % 8 == 0: 16-byte loads, every CTA covers a pixel range and writes TWO partial rows (its fp64 sums
// split into hi + lo floats); gsb::sum_slices adds the rows in row order.
template <typename T>
__global__ void __launch_bounds__(256) bias_grad_vec_kernel(const T* __restrict__ d, int64_t ld, int64_t P, int C,
                                                           float* __restrict__ partial) {
  GSB_PDL_ENTRY();
  __shared__ double red[256][8];
  const int G = C >> 3, lanes = 256 / G;
  const int g = threadIdx.x % G, lane = threadIdx.x / G;
  double acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.0;
  for (int64_t p = (int64_t)blockIdx.x * lanes + lane; p < P; p += (int64_t)gridDim.x * lanes) {
    float v[8];
    Vec8<T>::load(d + p * ld + g * 8, v);
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] += (double)v[j];
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) red[threadIdx.x][j] = acc[j];
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += 256) {
    double t = 0.0;
    for (int l = 0; l < lanes; ++l) t += red[l * G + (c >> 3)][c & 7];
    const float hi = (float)t;
    partial[((int64_t)blockIdx.x * 2) * C + c] = hi;
    partial[((int64_t)blockIdx.x * 2 + 1) * C + c] = (float)(t - (double)hi);
  }
}

int thin_gather_try(const gsb_view* in, const void* wp, int wdtype, const float* bias, const gsb_view* out,
                    const gsb_conv_geom* g, int act, int accumulate, cudaStream_t st);
int thin_wgrad_try(const gsb_view* in, const gsb_view* dout, float* ws, int max_slices, int* used, const gsb_conv_geom* g,
                   cudaStream_t st);

int conv_gather_direct(const gsb_view* in, const void* wp, int wdtype, const float* bias, const gsb_view* out,
                       const gsb_conv_geom* g, int act, int accumulate, cudaStream_t st) {
  {  // layers with 1..4 channels on one side: dedicated bandwidth kernels (thin_conv.cu)
    int r = thin_gather_try(in, wp, wdtype, bias, out, g, act, accumulate, st);
    if (r != 0) return r < 0 ? -1 : 0;
  }
  DView vi = dv(in), vo = dv(out);
  int sub = (g->mode == GSB_MODE_CONVT && g->stride > 1) ? g->stride : 1;
  for (int py = 0; py < sub; ++py)
    for (int px = 0; px < sub; ++px) {
      int QH = (out->h - py + sub - 1) / sub, QW = (out->w - px + sub - 1) / sub;
      int64_t M = (int64_t)out->n * QH * QW;
      if (M <= 0) continue;
      if (out->c <= 8) {
        dim3 grid((unsigned)((M + 127) / 128), (unsigned)((out->c + 7) / 8));
        gsb::launch(conv_gather_kernel<128, 8, 4, 1>, gsb::LaunchCfg(grid, 256, 0, st), vi, wp, wdtype, bias, vo, g->mode, g->kh, g->kw, g->stride,
                                                               g->pad, act, accumulate, sub, py, px);
      } else {
        dim3 grid((unsigned)((M + 63) / 64), (unsigned)((out->c + 63) / 64));
        gsb::launch(conv_gather_kernel<64, 64, 4, 4>, gsb::LaunchCfg(grid, 256, 0, st), vi, wp, wdtype, bias, vo, g->mode, g->kh, g->kw, g->stride,
                                                               g->pad, act, accumulate, sub, py, px);
      }
    }
  return check_launch("conv_gather(direct)", sub * sub);
}

// ws == nullptr: plan only (*used = number of slices the launch would write).
int conv_wgrad_direct(const gsb_view* in, const gsb_view* dout, float* ws, int max_slices, int* used, const gsb_conv_geom* g,
                      cudaStream_t st) {
  {
    int r = thin_wgrad_try(in, dout, ws, max_slices, used, g, st);
    if (r != 0) return r < 0 ? -1 : 0;
  }
  const int64_t slice_elems = (int64_t)g->kh * g->kw * dout->c * in->c;
  int rc = 0;
  DView vi = dv(in), vd = dv(dout);
  const bool conv = g->mode == GSB_MODE_CONV;
  int64_t P = conv ? num_pixels(dout) : num_pixels(in);
  int taps = g->kh * g->kw;
  auto launch = [&](auto kernel, int BM, int BN) {
    int tiles_m = (dout->c + BM - 1) / BM, tiles_n = (in->c + BN - 1) / BN;
    int64_t base = (int64_t)tiles_m * tiles_n * taps;
    int64_t want = ((int64_t)num_sms() * 4 + base - 1) / base;
    int64_t max_split = (P + 255) / 256;
    if (ws != nullptr && max_split > max_slices) max_split = max_slices;
    if (max_split > 256) max_split = 256;
    int64_t split = want < 1 ? 1 : (want > max_split ? max_split : want);
    int64_t chunk = ((P + split - 1) / split + BK - 1) / BK * BK;
    split = (P + chunk - 1) / chunk;
    if (split < 1) split = 1;
    *used = (int)split;
    if (ws == nullptr) return;
    if (split > max_slices) { rc = -1; return; }
    dim3 grid((unsigned)(tiles_m * tiles_n), (unsigned)taps, (unsigned)split);
    gsb::launch(kernel, gsb::LaunchCfg(grid, 256, 0, st), vi, vd, ws, slice_elems, g->mode, g->kh, g->kw, g->stride, g->pad, chunk, tiles_n);
  };
  if (dout->c <= 8)
    launch(conv_wgrad_kernel<8, 128, 1, 4>, 8, 128);
  else if (in->c <= 8)
    launch(conv_wgrad_kernel<64, 8, 2, 1>, 64, 8);
  else
    launch(conv_wgrad_kernel<64, 64, 4, 4>, 64, 64);
  GSB_REQUIRE(rc == 0, "conv_wgrad(direct): %d slices needed, workspace holds %d", *used, max_slices);
  if (ws == nullptr) return 0;
  return check_launch("conv_wgrad(direct)");
}

}  // namespace gsb

using namespace gsb;

extern "C" size_t gsb_bias_grad_workspace_bytes(const gsb_view* dout) {
  return (size_t)num_sms() * 4 * 2 * (size_t)(dout ? dout->c : 0) * sizeof(float);
}

// db[c] = sum over pixels of dout[., c]  (stored, not accumulated; deterministic)
extern "C" int gsb_bias_grad(const gsb_view* dout, float* db, void* workspace, size_t workspace_bytes, gsb_stream s) {
  GSB_REQUIRE(dout && dout->ptr && db, "bias_grad: null argument");
  const int C = dout->c;
  const int64_t P = num_pixels(dout);
  if (vec8_ok(dout) && C >= 8 && 256 % (C / 8) == 0 && C / 8 <= 256 && workspace != nullptr) {
    const int lanes = 256 / (C / 8);
    int64_t want = (P + (int64_t)lanes * 8 - 1) / ((int64_t)lanes * 8);
    int64_t cap = (int64_t)(workspace_bytes / (2 * (size_t)C * sizeof(float)));
    if (cap > (int64_t)num_sms() * 4) cap = (int64_t)num_sms() * 4;
    int grid = (int)(want < 1 ? 1 : (want > cap ? cap : want));
    if (grid >= 1 && cap >= 1) {
      float* partial = reinterpret_cast<float*>(workspace);
      if (dout->dtype == GSB_BF16)
        gsb::launch(bias_grad_vec_kernel<__nv_bfloat16>, gsb::LaunchCfg(grid, 256, 0, to_stream(s)), (const __nv_bfloat16*)dout->ptr,
                    (int64_t)dout->sw, P, C, partial);
      else
        gsb::launch(bias_grad_vec_kernel<float>, gsb::LaunchCfg(grid, 256, 0, to_stream(s)), (const float*)dout->ptr, (int64_t)dout->sw, P, C,
                    partial);
      gsb::sum_slices(partial, 2 * grid, C, C, db, 0, to_stream(s));
      return check_launch("bias_grad", 2);
    }
  }
  if (workspace != nullptr && C >= 1 && C <= 65535 && P >= 4096) {
    int64_t want = (P + 2047) / 2048;
    int64_t cap = (int64_t)(workspace_bytes / (2 * (size_t)C * sizeof(float)));
    if (cap > (int64_t)num_sms() * 4) cap = (int64_t)num_sms() * 4;
    const int gx = (int)(want > cap ? cap : want);
    if (gx >= 1) {
      float* partial = reinterpret_cast<float*>(workspace);
      gsb::launch(bias_grad_rows_kernel, gsb::LaunchCfg(dim3((unsigned)gx, (unsigned)C), 256, 0, to_stream(s)), dv(dout), partial);
      gsb::sum_slices(partial, 2 * gx, C, C, db, 0, to_stream(s));
      return check_launch("bias_grad", 2);
    }
  }
  dim3 grid((unsigned)C, 1u);
  gsb::launch(bias_grad_kernel, gsb::LaunchCfg(grid, 256, 0, to_stream(s)), dv(dout), db);
  return check_launch("bias_grad");
}

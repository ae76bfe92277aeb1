// Bandwidth kernels for the network-edge convolutions, where one side has 1..4 channels and a 128-wide
// tensor-core tile cannot be filled:
//   * "stem": Conv2d(Cin<=4 -> Cout, k x k, stride, pad) reading the fp32 NCHW network input and writing the
//     NHWC activation (unet/unet_model.py:14 via unet_parts.py:16; networks.py:582 outermost, :640), forward
//     and weight gradient;
//   * "head": the 1x1 OutConv (unet_parts.py:71-77) K -> C<=8 classes, reading the NHWC activation and writing
//     fp32 NCHW logits; backward fuses dX, dW and db in one pass over (x, dlogits).
// All four are HBM-bound; algorithmic bytes: stem fwd = N*H*W*(4 Cin + Cout*esz); stem wgrad = one read of dY
// per 16-tap pass + x; head fwd = N*H*W*(K*esz + 4C); head bwd = N*H*W*(2 K*esz + 4C).
#include <stdlib.h>

#include "common.cuh"

namespace gsb {

constexpr int kStemMaxTaps = 64;   // k*k*Cin

// ---------------------------------------------------------------------------------------------------------
// stem forward: thread = (output pixel, 8-channel group)
// ---------------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) stem_fwd_kernel(const float* __restrict__ x, int N, int Cin, int H, int W,
                                                      const float* __restrict__ w /*[Cout][Cin][k][k]*/,
                                                      const float* __restrict__ bias, T* __restrict__ y, int64_t yld, int OH,
                                                      int OW, int Cout, int k, int stride, int pad, int act) {
  GSB_PDL_ENTRY();
  extern __shared__ float sw[];  // [taps][Cout]
  const int taps = Cin * k * k;
  for (int i = threadIdx.x; i < taps * Cout; i += blockDim.x) {
    int t = i / Cout, co = i % Cout;
    sw[i] = w[(int64_t)co * taps + t];
  }
  __syncthreads();
  const int G = Cout >> 3;
  const int64_t total = (int64_t)N * OH * OW * G;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int g = (int)(i % G);
    const int64_t p = i / G;
    const int ox = (int)(p % OW);
    const int64_t r = p / OW;
    const int oy = (int)(r % OH), n = (int)(r / OH);
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = bias ? bias[g * 8 + j] : 0.f;
    int t = 0;
    for (int ci = 0; ci < Cin; ++ci) {
      const float* xp = x + ((int64_t)n * Cin + ci) * H * W;
      for (int ky = 0; ky < k; ++ky) {
        const int iy = oy * stride - pad + ky;
        for (int kx = 0; kx < k; ++kx, ++t) {
          const int ix = ox * stride - pad + kx;
          float v = (iy >= 0 && iy < H && ix >= 0 && ix < W) ? __ldg(xp + (int64_t)iy * W + ix) : 0.f;
          const float4 w0 = *reinterpret_cast<const float4*>(sw + t * Cout + g * 8);
          const float4 w1 = *reinterpret_cast<const float4*>(sw + t * Cout + g * 8 + 4);
          acc[0] = fmaf(v, w0.x, acc[0]); acc[1] = fmaf(v, w0.y, acc[1]);
          acc[2] = fmaf(v, w0.z, acc[2]); acc[3] = fmaf(v, w0.w, acc[3]);
          acc[4] = fmaf(v, w1.x, acc[4]); acc[5] = fmaf(v, w1.y, acc[5]);
          acc[6] = fmaf(v, w1.z, acc[6]); acc[7] = fmaf(v, w1.w, acc[7]);
        }
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = act_fwd(acc[j], act);
    Vec8<T>::store(y + p * yld + g * 8, acc);
  }
}

// ---------------------------------------------------------------------------------------------------------
// stem weight gradient: dw[co][tap] = sum_pix dY[pix][co] * x[pix@tap]; TP taps per pass kept in registers.
// Deterministic reduction (all weight-gradient kernels of this file): the pixel lanes of a CTA add their registers into
// shared memory one lane after the other, the CTA STORES its sums into row blockIdx.x of a partial buffer, and
// gsb::sum_slices adds the rows in row order.  No atomics: two runs give the same bits.
// ---------------------------------------------------------------------------------------------------------
template <typename T, int TP>
__global__ void __launch_bounds__(256) stem_wgrad_kernel(const float* __restrict__ x, int N, int Cin, int H, int W,
                                                        const T* __restrict__ dy, int64_t dld, int OH, int OW, int Cout,
                                                        int k, int stride, int pad, int tap0, float* __restrict__ dw) {
  GSB_PDL_ENTRY();
  extern __shared__ float sacc[];  // [G][TP][8]
  __shared__ int s_ci[TP], s_ky[TP], s_kx[TP];
  const int taps = Cin * k * k;
  if (threadIdx.x < TP) {
    const int tap = tap0 + threadIdx.x;
    s_ci[threadIdx.x] = tap / (k * k);
    s_ky[threadIdx.x] = (tap % (k * k)) / k;
    s_kx[threadIdx.x] = tap % k;
  }
  const int G = Cout >> 3;
  const int lanes = blockDim.x / G;
  const int g = threadIdx.x % G, lane = threadIdx.x / G;
  for (int i = threadIdx.x; i < G * TP * 8; i += blockDim.x) sacc[i] = 0.f;
  __syncthreads();
  float acc[TP][8];
#pragma unroll
  for (int t = 0; t < TP; ++t)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[t][j] = 0.f;
  const int64_t P = (int64_t)N * OH * OW;
  for (int64_t p = (int64_t)blockIdx.x * lanes + lane; p < P; p += (int64_t)gridDim.x * lanes) {
    const int ox = (int)(p % OW);
    const int64_t r = p / OW;
    const int oy = (int)(r % OH), n = (int)(r / OH);
    float d[8];
    Vec8<T>::load(dy + p * dld + g * 8, d);
#pragma unroll
    for (int t = 0; t < TP; ++t) {
      const int tap = tap0 + t;
      if (tap < taps) {
        const int ci = s_ci[t];
        const int iy = oy * stride - pad + s_ky[t], ix = ox * stride - pad + s_kx[t];
        float v = (iy >= 0 && iy < H && ix >= 0 && ix < W) ? __ldg(x + (((int64_t)n * Cin + ci) * H + iy) * W + ix) : 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[t][j] = fmaf(v, d[j], acc[t][j]);
      }
    }
  }
  for (int l = 0; l < lanes; ++l) {
    if (lane == l) {
#pragma unroll
      for (int t = 0; t < TP; ++t)
#pragma unroll
        for (int j = 0; j < 8; ++j) sacc[(g * TP + t) * 8 + j] += acc[t][j];
    }
    __syncthreads();
  }
  float* __restrict__ row = dw + (int64_t)blockIdx.x * Cout * taps;
  for (int i = threadIdx.x; i < G * TP * 8; i += blockDim.x) {
    const int j = i % 8, t = (i / 8) % TP, gg = i / (8 * TP);
    const int tap = tap0 + t;
    if (tap < taps) row[(int64_t)(gg * 8 + j) * taps + tap] = sacc[i];
  }
}

// ---------------------------------------------------------------------------------------------------------
// Fast stem kernels (K x K, stride S known at compile time): every thread owns PX consecutive output pixels of
// one 8-channel group, so the input row segment is loaded once per (ci, ky) and reused by all K horizontal taps,
// the weight vectors are read once per tap for PX pixels, and the index arithmetic (32-bit) is paid once per PX
// pixels.  The generic kernels above are issue-bound (~480 instructions per 16-byte store, ncu profiles/r01b).
// The forward optionally accumulates the BatchNorm sum / sum-of-squares of the STORED (rounded) values.
// ---------------------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ float stored_value(float v) { return v; }
template <>
__device__ __forceinline__ float stored_value<__nv_bfloat16>(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }

// NOACT: no activation follows inside this kernel (the U-Net stem feeds BatchNorm) -> the per-element activation switch
// folds away.  Accumulation and statistics use the packed fp32x2 pipe (FFMA2 / FADD2, sm_100): the kernel is
// instruction-bound (ncu profiles/r01n: 57 % issue at 25 % occupancy, ALU pipe above the FMA pipe), and the packed
// forms halve its arithmetic instruction count.
template <typename T, int PX, int K, int S, bool NOACT>
__global__ void __launch_bounds__(256, 2) stem_fwd_fast_kernel(const float* __restrict__ x, int N, int Cin, int H, int W,
                                                           const float* __restrict__ w, const float* __restrict__ bias,
                                                           T* __restrict__ y, int64_t yld, int OH, int OW, int Cout, int pad,
                                                           int act, float* __restrict__ stats) {
  GSB_PDL_ENTRY();
  extern __shared__ float sw[];  // [taps][Cout] | red[2][Cout]
  const int taps = Cin * K * K;
  float* red = sw + taps * Cout;
  for (int i = threadIdx.x; i < taps * Cout; i += blockDim.x) {
    int t = i / Cout, co = i % Cout;
    sw[i] = w[(int64_t)co * taps + t];
  }
  if (stats != nullptr)
    for (int i = threadIdx.x; i < 2 * Cout; i += blockDim.x) red[i] = 0.f;
  __syncthreads();
  constexpr int SPAN = (PX - 1) * S + K;
  const unsigned G = (unsigned)Cout >> 3, XQ = (unsigned)(OW + PX - 1) / PX;
  const unsigned total = (unsigned)N * (unsigned)OH * XQ * G;
  const unsigned g = threadIdx.x % G;  // fixed per thread: the grid stride is a multiple of G
  float2 s0[4], s1[4], b2[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    s0[j] = s1[j] = make_float2(0.f, 0.f);
    b2[j] = bias ? make_float2(bias[g * 8 + 2 * j], bias[g * 8 + 2 * j + 1]) : make_float2(0.f, 0.f);
  }
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const unsigned q = i / G;
    const unsigned xq = q % XQ, row = q / XQ;
    const int oy = (int)(row % (unsigned)OH), n = (int)(row / (unsigned)OH);
    const int ox0 = (int)xq * PX;
    float2 acc[PX][4];
#pragma unroll
    for (int px = 0; px < PX; ++px)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[px][j] = b2[j];
    const int ixb = ox0 * S - pad;
    const bool x_inside = ixb >= 0 && ixb + SPAN <= W;
    for (int ci = 0; ci < Cin; ++ci) {
#pragma unroll
      for (int ky = 0; ky < K; ++ky) {
        const int iy = oy * S - pad + ky;
        if (iy < 0 || iy >= H) continue;
        const float* xr = x + (((int64_t)n * Cin + ci) * H + iy) * W;
        float xv[SPAN];
        if (x_inside) {
#pragma unroll
          for (int e = 0; e < SPAN; ++e) xv[e] = __ldg(xr + ixb + e);
        } else {
#pragma unroll
          for (int e = 0; e < SPAN; ++e) {
            const int ix = ixb + e;
            xv[e] = (ix >= 0 && ix < W) ? __ldg(xr + ix) : 0.f;
          }
        }
        const float* wr = sw + ((ci * K + ky) * K) * Cout + g * 8;
#pragma unroll
        for (int kx = 0; kx < K; ++kx) {
          const float4 w0 = *reinterpret_cast<const float4*>(wr + kx * Cout);
          const float4 w1 = *reinterpret_cast<const float4*>(wr + kx * Cout + 4);
          const float2 wa = make_float2(w0.x, w0.y), wb = make_float2(w0.z, w0.w);
          const float2 wc = make_float2(w1.x, w1.y), wd = make_float2(w1.z, w1.w);
#pragma unroll
          for (int px = 0; px < PX; ++px) {
            const float2 v = make_float2(xv[px * S + kx], xv[px * S + kx]);
            acc[px][0] = __ffma2_rn(v, wa, acc[px][0]);
            acc[px][1] = __ffma2_rn(v, wb, acc[px][1]);
            acc[px][2] = __ffma2_rn(v, wc, acc[px][2]);
            acc[px][3] = __ffma2_rn(v, wd, acc[px][3]);
          }
        }
      }
    }
    const int64_t pbase = ((int64_t)n * OH + oy) * OW + ox0;
#pragma unroll
    for (int px = 0; px < PX; ++px) {
      if (ox0 + px < OW) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          o[2 * j] = NOACT ? acc[px][j].x : act_fwd(acc[px][j].x, act);
          o[2 * j + 1] = NOACT ? acc[px][j].y : act_fwd(acc[px][j].y, act);
        }
        Vec8<T>::store(y + (pbase + px) * yld + g * 8, o);
        if (stats != nullptr) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float2 r = make_float2(stored_value<T>(o[2 * j]), stored_value<T>(o[2 * j + 1]));
            s0[j] = __fadd2_rn(s0[j], r);
            s1[j] = __ffma2_rn(r, r, s1[j]);
          }
        }
      }
    }
  }
  if (stats != nullptr) {
    // threads sharing a channel group add in a fixed order (deterministic statistics rows)
    const unsigned lanes = blockDim.x / G, lane = threadIdx.x / G;
    for (unsigned l = 0; l < lanes; ++l) {
      if (lane == l) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          red[g * 8 + 2 * j] += s0[j].x;
          red[g * 8 + 2 * j + 1] += s0[j].y;
          red[Cout + g * 8 + 2 * j] += s1[j].x;
          red[Cout + g * 8 + 2 * j + 1] += s1[j].y;
        }
      }
      __syncthreads();
    }
    for (int i = threadIdx.x; i < 2 * Cout; i += blockDim.x) stats[(int64_t)blockIdx.x * 2 * Cout + i] = red[i];
  }
}

// dw[co][ci][ky][kx] += sum_pix dY[pix][co] * x[pix @ (ky,kx)]; blockIdx.y = ci; KYN*K taps of 8 channels in registers:
// KYN = K for the 3x3 stem; the 4x4 stride-2 first convolutions of the Pix2Pix generator / discriminator take KYN = 2
// kernel rows per CTA (blockIdx.z = which pair), 64 accumulator registers instead of 128
template <typename T, int PX, int K, int S, int KYN>
__global__ void __launch_bounds__(256, 2) stem_wgrad_fast_kernel(const float* __restrict__ x, int N, int Cin, int H, int W,
                                                             const T* __restrict__ dy, int64_t dld, int OH, int OW, int Cout,
                                                             int pad, float* __restrict__ dw) {
  GSB_PDL_ENTRY();
  extern __shared__ float sacc[];  // [G][KYN*K][8]
  constexpr int SPAN = (PX - 1) * S + K;
  constexpr int KT = KYN * K;      // taps of this CTA
  const int ci = blockIdx.y, ky0 = blockIdx.z * KYN;
  const unsigned G = (unsigned)Cout >> 3, XQ = (unsigned)(OW + PX - 1) / PX;
  const unsigned lanes = blockDim.x / G;
  const unsigned g = threadIdx.x % G, lane = threadIdx.x / G;
  for (int i = threadIdx.x; i < (int)G * KT * 8; i += blockDim.x) sacc[i] = 0.f;
  __syncthreads();
  float2 acc[KT][4];  // packed fp32x2 accumulators (FFMA2): half the arithmetic instructions of the scalar form
#pragma unroll
  for (int t = 0; t < KT; ++t)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[t][j] = make_float2(0.f, 0.f);
  const unsigned items = (unsigned)N * (unsigned)OH * XQ;
  for (unsigned q = blockIdx.x * lanes + lane; q < items; q += gridDim.x * lanes) {
    const unsigned xq = q % XQ, row = q / XQ;
    const int oy = (int)(row % (unsigned)OH), n = (int)(row / (unsigned)OH);
    const int ox0 = (int)xq * PX;
    const int64_t pbase = ((int64_t)n * OH + oy) * OW + ox0;
    float2 d[PX][4];
#pragma unroll
    for (int px = 0; px < PX; ++px) {
      float t8[8];
      if (ox0 + px < OW) {
        Vec8<T>::load(dy + (pbase + px) * dld + g * 8, t8);
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) t8[j] = 0.f;
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) d[px][j] = make_float2(t8[2 * j], t8[2 * j + 1]);
    }
    const int ixb = ox0 * S - pad;
    const bool x_inside = ixb >= 0 && ixb + SPAN <= W;
#pragma unroll
    for (int ky = 0; ky < KYN; ++ky) {
      const int iy = oy * S - pad + ky0 + ky;
      if (iy < 0 || iy >= H) continue;
      const float* xr = x + (((int64_t)n * Cin + ci) * H + iy) * W;
      float xv[SPAN];
      if (x_inside) {
#pragma unroll
        for (int e = 0; e < SPAN; ++e) xv[e] = __ldg(xr + ixb + e);
      } else {
#pragma unroll
        for (int e = 0; e < SPAN; ++e) {
          const int ix = ixb + e;
          xv[e] = (ix >= 0 && ix < W) ? __ldg(xr + ix) : 0.f;
        }
      }
#pragma unroll
      for (int kx = 0; kx < K; ++kx)
#pragma unroll
        for (int px = 0; px < PX; ++px) {
          const float2 v = make_float2(xv[px * S + kx], xv[px * S + kx]);
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[ky * K + kx][j] = __ffma2_rn(v, d[px][j], acc[ky * K + kx][j]);
        }
    }
  }
  for (unsigned l = 0; l < lanes; ++l) {
    if (lane == l) {
#pragma unroll
      for (int t = 0; t < KT; ++t)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          sacc[(g * KT + t) * 8 + 2 * j] += acc[t][j].x;
          sacc[(g * KT + t) * 8 + 2 * j + 1] += acc[t][j].y;
        }
    }
    __syncthreads();
  }
  const int taps = Cin * K * K;
  float* __restrict__ row = dw + (int64_t)blockIdx.x * Cout * taps;   // partial row of this CTA column
  for (int i = threadIdx.x; i < (int)G * KT * 8; i += blockDim.x) {
    const int j = i % 8, t = (i / 8) % KT, gg = i / (8 * KT);
    row[(int64_t)(gg * 8 + j) * taps + ci * K * K + ky0 * K + t] = sacc[i];
  }
}

// ---------------------------------------------------------------------------------------------------------
// head forward: 8 lanes per pixel, each lane a contiguous 1/8 of the K channels; 3 shuffles per class
// ---------------------------------------------------------------------------------------------------------
constexpr int kHeadMaxC = 8;

// CT = compile-time bound on the class count (2 or 8): weights of the lane's 8 channels live in registers when
// K == 64 (OutConv of the U-Net), pixel arithmetic is 32-bit.
template <typename T, int CT>
__global__ void __launch_bounds__(256, 4) head_fwd_kernel(const T* __restrict__ x, int64_t xld, unsigned P, unsigned HW, int K, int C,
                                                      const float* __restrict__ w /*[C][K]*/, const float* __restrict__ bias,
                                                      float* __restrict__ out /*[N][C][HW]*/) {
  GSB_PDL_ENTRY();
  extern __shared__ float sw[];  // [C][K]
  for (int i = threadIdx.x; i < C * K; i += blockDim.x) sw[i] = w[i];
  __syncthreads();
  const int sub = threadIdx.x & 7;
  const int chunks = K >> 6;  // K multiple of 64: each lane handles 8 channels per 64-channel chunk
  float wr[CT][8];
#pragma unroll
  for (int c = 0; c < CT; ++c)
#pragma unroll
    for (int j = 0; j < 8; ++j) wr[c][j] = c < C ? sw[c * K + sub * 8 + j] : 0.f;
  const unsigned stride_p = (gridDim.x * blockDim.x) >> 3;
  // warp-uniform trip count (the shuffles below need all 32 lanes): iterate on the warp's first pixel.  Two pixels
  // (pw, pw + stride_p) are in flight per thread: with one the kernel sat at 2.7 TB/s waiting on its single 16-byte load
  // (ncu profiles/r01n: issue 57 %, long-scoreboard stalls).
  for (unsigned pw = (blockIdx.x * blockDim.x + (threadIdx.x & ~31u)) >> 3; pw < P; pw += 2 * stride_p) {
    unsigned pp[2];
    bool valid[2];
    float v[2][8], acc[2][CT];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      pp[u] = pw + u * stride_p + ((threadIdx.x & 31) >> 3);
      valid[u] = pp[u] < P;
      if (valid[u]) Vec8<T>::load(x + (int64_t)pp[u] * xld + sub * 8, v[u]);
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
#pragma unroll
      for (int c = 0; c < CT; ++c) acc[u][c] = 0.f;
      if (valid[u]) {
#pragma unroll
        for (int c = 0; c < CT; ++c)
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[u][c] = fmaf(v[u][j], wr[c][j], acc[u][c]);
        for (int ch = 1; ch < chunks; ++ch) {
          const int k0 = ch * 64 + sub * 8;
          float t[8];
          Vec8<T>::load(x + (int64_t)pp[u] * xld + k0, t);
#pragma unroll
          for (int c = 0; c < CT; ++c)
            if (c < C) {
#pragma unroll
              for (int j = 0; j < 8; ++j) acc[u][c] = fmaf(t[j], sw[c * K + k0 + j], acc[u][c]);
            }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
#pragma unroll
      for (int c = 0; c < CT; ++c) {
        acc[u][c] += __shfl_xor_sync(0xffffffffu, acc[u][c], 1);
        acc[u][c] += __shfl_xor_sync(0xffffffffu, acc[u][c], 2);
        acc[u][c] += __shfl_xor_sync(0xffffffffu, acc[u][c], 4);
      }
      if (valid[u] && sub == 0) {
        const unsigned n = pp[u] / HW, r = pp[u] - n * HW;
#pragma unroll
        for (int c = 0; c < CT; ++c)
          if (c < C) out[((int64_t)n * C + c) * HW + r] = acc[u][c] + (bias ? bias[c] : 0.f);
      }
    }
  }
}

// OutConv of the U-Net (K = 64 channels -> C <= 2 classes): TWO lanes per pixel, 32 channels each -- four independent 16-byte
// loads in flight per thread, the two class accumulators packed in one fp32x2 register (FFMA2), one shuffle per class.
// The 8-lane kernel above spends 6 shuffles + 6 adds per 16-byte load and ran at 0.45 of the HBM roofline (issue-bound).
template <typename T>
__global__ void __launch_bounds__(256, 3) head_fwd_k64_kernel(const T* __restrict__ x, int64_t xld, unsigned P, unsigned HW, int C,
                                                          const float* __restrict__ w /*[C][64]*/, const float* __restrict__ bias,
                                                          float* __restrict__ out /*[N][C][HW]*/) {
  GSB_PDL_ENTRY();
  const int sub = threadIdx.x & 1, lane = threadIdx.x & 31;
  float2 wr[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) wr[i] = make_float2(__ldg(w + sub * 32 + i), C > 1 ? __ldg(w + 64 + sub * 32 + i) : 0.f);
  const float b0 = bias ? bias[0] : 0.f, b1 = (bias && C > 1) ? bias[1] : 0.f;
  const unsigned stride = (gridDim.x * blockDim.x) >> 1;
  for (unsigned pw = (blockIdx.x * blockDim.x + (threadIdx.x & ~31u)) >> 1; pw < P; pw += stride) {   // warp-uniform trip count
    const unsigned p = pw + (lane >> 1);
    const bool valid = p < P;
    float v[4][8];
    if (valid) {
      const T* src = x + (int64_t)p * xld + sub * 32;
#pragma unroll
      for (int j = 0; j < 4; ++j) Vec8<T>::load(src + j * 8, v[j]);
    }
    float2 acc = make_float2(0.f, 0.f);
    if (valid) {
#pragma unroll
      for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int i = 0; i < 8; ++i) acc = __ffma2_rn(make_float2(v[j][i], v[j][i]), wr[j * 8 + i], acc);
    }
    acc.x += __shfl_xor_sync(0xffffffffu, acc.x, 1);
    acc.y += __shfl_xor_sync(0xffffffffu, acc.y, 1);
    if (valid && sub == 0) {
      const unsigned n = p / HW, r = p - n * HW;
      out[((int64_t)n * C) * HW + r] = acc.x + b0;
      if (C > 1) out[((int64_t)n * C + 1) * HW + r] = acc.y + b1;
    }
  }
}

// head backward: dx[p][k] = sum_c d[p][c] w[c][k];  dw[c][k] += sum_p d[p][c] x[p][k];  db[c] += sum_p d[p][c]
template <typename T, int CT>
__global__ void __launch_bounds__(256, 4) head_bwd_kernel(const T* __restrict__ x, int64_t xld, unsigned P, unsigned HW, int K, int C,
                                                      const float* __restrict__ w, const float* __restrict__ dout,
                                                      T* __restrict__ dx, int64_t dxld, float* __restrict__ dw,
                                                      float* __restrict__ db) {
  GSB_PDL_ENTRY();
  extern __shared__ float sm[];  // sw[C][K] | sdw[C][K] | sdb[C]
  float* sw = sm;
  float* sdw = sm + C * K;
  float* sdb = sdw + C * K;
  for (int i = threadIdx.x; i < C * K; i += blockDim.x) {
    sw[i] = w[i];
    sdw[i] = 0.f;
  }
  if (threadIdx.x < C) sdb[threadIdx.x] = 0.f;
  __syncthreads();
  const unsigned G = (unsigned)K >> 3;  // 8-channel groups
  const unsigned lanes = blockDim.x / G;
  const unsigned g = threadIdx.x % G, lane = threadIdx.x / G;
  float aw[CT][8], ab[CT], wr[CT][8];
#pragma unroll
  for (int c = 0; c < CT; ++c) {
    ab[c] = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      aw[c][j] = 0.f;
      wr[c][j] = c < C ? sw[c * K + g * 8 + j] : 0.f;
    }
  }
  // two pixels in flight per thread (one left the kernel latency-bound at 3.3 TB/s, ncu profiles/r01n)
  const unsigned pstep = gridDim.x * lanes;
  for (unsigned p0 = blockIdx.x * lanes + lane; p0 < P; p0 += 2 * pstep) {
    float d[2][CT], v[2][8];
    bool valid[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const unsigned p = p0 + u * pstep;
      valid[u] = p < P;
      const unsigned n = valid[u] ? p / HW : 0, r = valid[u] ? p - n * HW : 0;
#pragma unroll
      for (int c = 0; c < CT; ++c) d[u][c] = (valid[u] && c < C) ? __ldg(dout + ((int64_t)n * C + c) * HW + r) : 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) v[u][j] = 0.f;
      if (valid[u] && dw != nullptr) Vec8<T>::load(x + (int64_t)p * xld + g * 8, v[u]);
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      float o[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = 0.f;
#pragma unroll
      for (int c = 0; c < CT; ++c) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          o[j] = fmaf(d[u][c], wr[c][j], o[j]);
          if (dw != nullptr) aw[c][j] = fmaf(d[u][c], v[u][j], aw[c][j]);
        }
        ab[c] += d[u][c];
      }
      if (valid[u] && dx != nullptr) Vec8<T>::store(dx + (int64_t)(p0 + u * pstep) * dxld + g * 8, o);
    }
  }
  // deterministic: lanes add in lane order, the CTA stores row blockIdx.x = [dw (C*K) | db (C)] of the partial buffer
  for (unsigned l = 0; l < lanes; ++l) {
    if (lane == l) {
      if (dw != nullptr) {
#pragma unroll
        for (int c = 0; c < CT; ++c)
          if (c < C) {
#pragma unroll
            for (int j = 0; j < 8; ++j) sdw[c * K + g * 8 + j] += aw[c][j];
          }
      }
      if (db != nullptr && g == 0) {
#pragma unroll
        for (int c = 0; c < CT; ++c)
          if (c < C) sdb[c] += ab[c];
      }
    }
    __syncthreads();
  }
  const int rowlen = C * K + kHeadMaxC;
  if (dw != nullptr)
    for (int i = threadIdx.x; i < C * K; i += blockDim.x) dw[(int64_t)blockIdx.x * rowlen + i] = sdw[i];
  if (db != nullptr && threadIdx.x < C) db[(int64_t)blockIdx.x * rowlen + threadIdx.x] = sdb[threadIdx.x];
}

// tensor-core stem (stem_tc.cu)
bool stem_tc_applicable(int cin, int cout, int k, int stride, int dtype, int64_t pixels);
int stem_tc_fwd(const float* x, int h, int w, const float* weight, const gsb_view* y, int k, int stride, int pad, float* stats,
                int max_rows, int* rows, cudaStream_t st, const float* aff_scale = nullptr, const float* aff_shift = nullptr,
                int aff_relu = 0);
int stem_tc_wgrad(const float* x, int h, int w, const gsb_view* dy, int k, int stride, int pad, float* partial, int max_rows,
                  int* rows, cudaStream_t st);

static int grid_cap(int64_t want, int per_sm) {
  int64_t cap = (int64_t)num_sms() * per_sm;
  return (int)(want < 1 ? 1 : (want > cap ? cap : want));
}

}  // namespace gsb

using namespace gsb;

#define EDGE_DISPATCH_T(dtype, ...) \
  if ((dtype) == GSB_BF16) {        \
    using T = __nv_bfloat16;        \
    __VA_ARGS__;                    \
  } else {                          \
    using T = float;                \
    __VA_ARGS__;                    \
  }

static int stem_check(const float* x, const gsb_view* y, int cin, int k) {
  GSB_REQUIRE(x && y && y->ptr, "stem: null argument");
  GSB_REQUIRE(vec8_ok(y), "stem: output must be a pixel-linear NHWC view with c%%8==0");
  GSB_REQUIRE(cin * k * k <= kStemMaxTaps && y->c <= 256 && 256 % (y->c / 8) == 0, "stem: unsupported shape (cin=%d k=%d cout=%d)", cin, k, y->c);
  return 0;
}

extern "C" int gsb_stem_stats_rows(void) { return num_sms() * 8; }

// Inference form of the stem: out = act(conv(x) * scale[co] + shift[co]) in one pass (eval-mode Conv -> BatchNorm -> ReLU of
// unet_parts.py:16-18 on the 1-channel network input).  0 = launched, 1 = not applicable (the caller runs gsb_stem_conv_fwd +
// gsb_bn_act), -1 = error.  act: GSB_ACT_NONE or GSB_ACT_RELU; tensor-core stem shapes only.
extern "C" int gsb_stem_conv_fwd_affine(const float* x, int n, int cin, int h, int w, const float* weight, const float* scale,
                                        const float* shift, const gsb_view* y, int k, int stride, int pad, int act, gsb_stream s) {
  if (stem_check(x, y, cin, k)) return -1;
  GSB_REQUIRE(weight && scale && y->n == n && y->h == (h + 2 * pad - k) / stride + 1 && y->w == (w + 2 * pad - k) / stride + 1,
              "stem_conv_fwd_affine: shape mismatch");
  if ((act != GSB_ACT_NONE && act != GSB_ACT_RELU) || !stem_tc_applicable(cin, y->c, k, stride, y->dtype, num_pixels(y))) return 1;
  int rows = 0;
  if (stem_tc_fwd(x, h, w, weight, y, k, stride, pad, nullptr, gsb_stem_stats_rows(), &rows, to_stream(s), scale, shift,
                  act == GSB_ACT_RELU ? 1 : 0))
    return -1;
  return check_launch("stem_conv_fwd_affine");
}

extern "C" int gsb_stem_conv_fwd(const float* x, int n, int cin, int h, int w, const float* weight, const float* bias,
                                 const gsb_view* y, int k, int stride, int pad, int act, float* stats_partial, int* stats_rows,
                                 gsb_stream s) {
  if (stem_check(x, y, cin, k)) return -1;
  GSB_REQUIRE(weight && y->n == n && y->h == (h + 2 * pad - k) / stride + 1 && y->w == (w + 2 * pad - k) / stride + 1,
              "stem_conv_fwd: shape mismatch");
  const int cout = y->c;
  int64_t total = num_pixels(y) * (cout / 8);
  if (stats_rows) *stats_rows = 0;
  if (bias == nullptr && act == GSB_ACT_NONE && stem_tc_applicable(cin, cout, k, stride, y->dtype, num_pixels(y))) {
    // im2col tile in shared memory + tcgen05.mma (hi / lo bf16 split of the fp32 input and weights): bound by the output stream
    int rows = 0;
    if (stem_tc_fwd(x, h, w, weight, y, k, stride, pad, stats_partial, gsb_stem_stats_rows(), &rows, to_stream(s))) return -1;
    if (stats_partial && stats_rows) *stats_rows = rows;
    return check_launch("stem_conv_fwd(tc)");
  }
  const bool k3 = (k == 3 && stride == 1), k4 = (k == 4 && stride == 2);
  if ((k3 || k4) && total < (int64_t)1 << 31) {
    constexpr int PX = 4;
    const int64_t items = (int64_t)y->n * y->h * ((y->w + PX - 1) / PX) * (cout / 8);
    int grid = grid_cap((items + 255) / 256, 8);
    size_t smem = ((size_t)cin * k * k * cout + 2 * cout) * sizeof(float);
    if (k3 && act == GSB_ACT_NONE) {
      EDGE_DISPATCH_T(y->dtype, gsb::launch(stem_fwd_fast_kernel<T, PX, 3, 1, true>, gsb::LaunchCfg(grid, 256, smem, to_stream(s)), 
                                    x, n, cin, h, w, weight, bias, (T*)y->ptr, y->sw, y->h, y->w, cout, pad, act, stats_partial));
    } else if (k3) {
      EDGE_DISPATCH_T(y->dtype, gsb::launch(stem_fwd_fast_kernel<T, PX, 3, 1, false>, gsb::LaunchCfg(grid, 256, smem, to_stream(s)), 
                                    x, n, cin, h, w, weight, bias, (T*)y->ptr, y->sw, y->h, y->w, cout, pad, act, stats_partial));
    } else {
      EDGE_DISPATCH_T(y->dtype, gsb::launch(stem_fwd_fast_kernel<T, PX, 4, 2, false>, gsb::LaunchCfg(grid, 256, smem, to_stream(s)), 
                                    x, n, cin, h, w, weight, bias, (T*)y->ptr, y->sw, y->h, y->w, cout, pad, act, stats_partial));
    }
    if (stats_partial && stats_rows) *stats_rows = grid;
    return check_launch("stem_conv_fwd");
  }
  size_t smem = (size_t)cin * k * k * cout * sizeof(float);
  int grid = grid_cap((total + 255) / 256, 16);
  EDGE_DISPATCH_T(y->dtype, gsb::launch(stem_fwd_kernel<T>, gsb::LaunchCfg(grid, 256, smem, to_stream(s)), x, n, cin, h, w, weight, bias, (T*)y->ptr, y->sw,
                                                                                   y->h, y->w, cout, k, stride, pad, act));
  return check_launch("stem_conv_fwd");
}

// Partial rows of stem_conv_wgrad (row = cout * cin*k*k floats) and head_conv_bwd (row = c * x.c + 8 floats): the workspace
// should hold gsb_edge_rows() rows; with fewer the launch uses fewer CTA columns.
extern "C" int gsb_edge_rows(void) { return num_sms() * 4; }

extern "C" int gsb_stem_conv_wgrad(const float* x, int n, int cin, int h, int w, const gsb_view* dy, int k, int stride, int pad,
                                   float* dweight, void* workspace, size_t workspace_bytes, gsb_stream s) {
  if (stem_check(x, dy, cin, k)) return -1;
  GSB_REQUIRE(dweight && dy->n == n && workspace, "stem_conv_wgrad: bad argument");
  const int cout = dy->c, taps = cin * k * k;
  constexpr int TP = 9;
  const int G = cout / 8, lanes = 256 / G;
  const size_t row_bytes = (size_t)cout * taps * sizeof(float);
  const int64_t max_rows = (int64_t)(workspace_bytes / row_bytes);
  GSB_REQUIRE(max_rows >= 1, "stem_conv_wgrad: workspace too small");
  float* partial = reinterpret_cast<float*>(workspace);
  auto rows_for = [&](int64_t want, int per_sm) {
    int g = grid_cap(want, per_sm);
    return (int)(g > max_rows ? max_rows : g);
  };
  if (dy->sw == cout && ((uintptr_t)dy->ptr % 128) == 0 && stem_tc_applicable(cin, cout, k, stride, dy->dtype, num_pixels(dy))) {
    int rows = 0;
    if (stem_tc_wgrad(x, h, w, dy, k, stride, pad, partial, (int)(max_rows > (1 << 20) ? (1 << 20) : max_rows), &rows, to_stream(s)))
      return -1;
    gsb::sum_slices(partial, rows, (int64_t)cout * taps, (int64_t)cout * taps, dweight, 0, to_stream(s));
    return check_launch("stem_conv_wgrad(tc)", 2);
  }
  if (k == 3 && stride == 1 && num_pixels(dy) < (int64_t)1 << 31) {
    constexpr int PX = 4;
    const int64_t items = (int64_t)dy->n * dy->h * ((dy->w + PX - 1) / PX);
    int gx = rows_for((items + lanes - 1) / lanes, 4);
    size_t smem = (size_t)G * 9 * 8 * sizeof(float);
    EDGE_DISPATCH_T(dy->dtype, gsb::launch(stem_wgrad_fast_kernel<T, PX, 3, 1, 3>, gsb::LaunchCfg(dim3(gx, cin), 256, smem, to_stream(s)), 
                                   x, n, cin, h, w, (const T*)dy->ptr, dy->sw, dy->h, dy->w, cout, pad, partial));
    gsb::sum_slices(partial, gx, (int64_t)cout * taps, (int64_t)cout * taps, dweight, 0, to_stream(s));
    return check_launch("stem_conv_wgrad", 2);
  }
  if (k == 4 && stride == 2 && num_pixels(dy) < (int64_t)1 << 31 && getenv("GSB_STEM_WGRAD_FAST42") == nullptr) {
    // first convolutions of the Pix2Pix generator (1 -> 64) and discriminator (2 -> 64): one launch instead of
    // cin * 16 / 9 launches of the generic kernel (0.75 ms per Pix2Pix step, profiles/r01r)
    constexpr int PX = 2;
    const int64_t items = (int64_t)dy->n * dy->h * ((dy->w + PX - 1) / PX);
    int gx = rows_for((items + lanes - 1) / lanes, 2);
    size_t smem = (size_t)G * 8 * 8 * sizeof(float);
    EDGE_DISPATCH_T(dy->dtype, gsb::launch(stem_wgrad_fast_kernel<T, PX, 4, 2, 2>, gsb::LaunchCfg(dim3(gx, cin, 2), 256, smem, to_stream(s)), 
                                   x, n, cin, h, w, (const T*)dy->ptr, dy->sw, dy->h, dy->w, cout, pad, partial));
    gsb::sum_slices(partial, gx, (int64_t)cout * taps, (int64_t)cout * taps, dweight, 0, to_stream(s));
    return check_launch("stem_conv_wgrad", 2);
  }
  int grid = rows_for((num_pixels(dy) + lanes - 1) / lanes, 4);
  size_t smem = (size_t)G * TP * 8 * sizeof(float);
  int launches = 0;
  for (int tap0 = 0; tap0 < taps; tap0 += TP, ++launches) {
    EDGE_DISPATCH_T(dy->dtype, gsb::launch(stem_wgrad_kernel<T, TP>, gsb::LaunchCfg(grid, 256, smem, to_stream(s)), 
                                   x, n, cin, h, w, (const T*)dy->ptr, dy->sw, dy->h, dy->w, cout, k, stride, pad, tap0, partial));
  }
  gsb::sum_slices(partial, grid, (int64_t)cout * taps, (int64_t)cout * taps, dweight, 0, to_stream(s));
  return check_launch("stem_conv_wgrad", launches + 1);
}

static int head_check(const gsb_view* x, int c) {
  GSB_REQUIRE(x && x->ptr, "head: null view");
  GSB_REQUIRE(vec8_ok(x) && x->c % 64 == 0 && x->c <= 256, "head: input must be pixel-linear NHWC with c in {64,128,192,256}");
  GSB_REQUIRE(c >= 1 && c <= kHeadMaxC, "head: n_classes %d unsupported by the fused 1x1 kernel (max %d)", c, kHeadMaxC);
  return 0;
}

extern "C" int gsb_head_conv_fwd(const gsb_view* x, const float* weight, const float* bias, int c, float* logits, gsb_stream s) {
  if (head_check(x, c)) return -1;
  GSB_REQUIRE(weight && logits, "head_conv_fwd: null pointer");
  const int64_t P = num_pixels(x), HW = (int64_t)x->h * x->w;
  int grid = grid_cap((P * 8 + 255) / 256, 16);
  size_t smem = (size_t)c * x->c * sizeof(float);
  GSB_REQUIRE(P < ((int64_t)1 << 28), "head_conv_fwd: too many pixels");
  if (c <= 2 && x->c == 64) {
    int g2 = grid_cap((P * 2 + 255) / 256, 12);
    EDGE_DISPATCH_T(x->dtype, gsb::launch(head_fwd_k64_kernel<T>, gsb::LaunchCfg(g2, 256, 0, to_stream(s)), (const T*)x->ptr, x->sw, (unsigned)P,
                                          (unsigned)HW, c, weight, bias, logits));
    return check_launch("head_conv_fwd");
  }
  if (c <= 2) {
    EDGE_DISPATCH_T(x->dtype, gsb::launch(head_fwd_kernel<T, 2>, gsb::LaunchCfg(grid, 256, smem, to_stream(s)), (const T*)x->ptr, x->sw, (unsigned)P, (unsigned)HW,
                                                                                        x->c, c, weight, bias, logits));
  } else {
    EDGE_DISPATCH_T(x->dtype, gsb::launch(head_fwd_kernel<T, 8>, gsb::LaunchCfg(grid, 256, smem, to_stream(s)), (const T*)x->ptr, x->sw, (unsigned)P, (unsigned)HW,
                                                                                        x->c, c, weight, bias, logits));
  }
  return check_launch("head_conv_fwd");
}

extern "C" int gsb_head_conv_bwd(const gsb_view* x, const float* weight, int c, const float* dlogits, const gsb_view* dx,
                                 float* dweight, float* dbias, void* workspace, size_t workspace_bytes, gsb_stream s) {
  if (head_check(x, c)) return -1;
  GSB_REQUIRE(weight && dlogits, "head_conv_bwd: null pointer");
  GSB_REQUIRE(workspace || (!dweight && !dbias), "head_conv_bwd: workspace required for the parameter gradients");
  if (dx && dx->ptr)
    GSB_REQUIRE(vec8_ok(dx) && dx->c == x->c && num_pixels(dx) == num_pixels(x) && dx->dtype == x->dtype, "head_conv_bwd: dx mismatch");
  const int64_t P = num_pixels(x), HW = (int64_t)x->h * x->w;
  const int K = x->c, G = K / 8, lanes = 256 / G;
  int grid = grid_cap((P + lanes - 1) / lanes, 4);
  const int rowlen = c * K + kHeadMaxC;
  float* partial = reinterpret_cast<float*>(workspace);
  if (dweight || dbias) {
    const int64_t max_rows = (int64_t)(workspace_bytes / ((size_t)rowlen * sizeof(float)));
    GSB_REQUIRE(max_rows >= 1, "head_conv_bwd: workspace too small");
    if (grid > max_rows) grid = (int)max_rows;
  }
  float* pdw = dweight ? partial : nullptr;
  float* pdb = dbias ? partial + (size_t)c * K : nullptr;
  size_t smem = (size_t)(2 * c * K + c) * sizeof(float);
  GSB_REQUIRE(P < ((int64_t)1 << 31), "head_conv_bwd: too many pixels");
  if (c <= 2) {
    EDGE_DISPATCH_T(x->dtype, gsb::launch(head_bwd_kernel<T, 2>, gsb::LaunchCfg(grid, 256, smem, to_stream(s)), 
                                  (const T*)x->ptr, x->sw, (unsigned)P, (unsigned)HW, K, c, weight, dlogits,
                                  (dx && dx->ptr) ? (T*)dx->ptr : nullptr, dx ? dx->sw : 0, pdw, pdb));
  } else {
    EDGE_DISPATCH_T(x->dtype, gsb::launch(head_bwd_kernel<T, 8>, gsb::LaunchCfg(grid, 256, smem, to_stream(s)), 
                                  (const T*)x->ptr, x->sw, (unsigned)P, (unsigned)HW, K, c, weight, dlogits,
                                  (dx && dx->ptr) ? (T*)dx->ptr : nullptr, dx ? dx->sw : 0, pdw, pdb));
  }
  int launches = 1;
  if (dweight) { gsb::sum_slices(pdw, grid, rowlen, (int64_t)c * K, dweight, 0, to_stream(s)); ++launches; }
  if (dbias) { gsb::sum_slices(pdb, grid, rowlen, c, dbias, 0, to_stream(s)); ++launches; }
  return check_launch("head_conv_bwd", launches);
}

// Device-side image chain of the GenSeg end-to-end step, replacing the reference's host round trip:
//
//   running_files/train_end2end_jsrt.py:186-200
//     fake_mask  = where(augment(mask) > 0.1, 1, 0)                 (imgaug affine chain on the CPU, :100-112,:187-190)
//     fake_image = (G(fake_mask) - min) / (max - min)               (min / max over the WHOLE batch tensor, :197)
//     fake_image = fake_image.mul(255).add_(0.5).clamp_(0,255).to(uint8)                                    (:199)
//     fake_image = adjust_gamma(equalize(fake_image), 0.5) / 255.0  (torchvision, per image and channel,    :200)
//
// Integer / LUT arithmetic, HBM-bound and tiny (N x 64 K pixels): three launches
//   minmax_kernel        ordered-key atomicMin / atomicMax over the batch tensor
//   quant_hist_kernel    fp32 -> uint8 with the reference's operation order (no FMA contraction) + 256-bin histogram per plane
//   equalize_lut_kernel  per plane: cumulative histogram -> torchvision's equalisation LUT -> gamma LUT -> k / 255.0f
// plus mask_affine_kernel: inverse-mapped bilinear warp (zero outside) of the masks and the > 0.1 threshold.
// Bit-exact against torchvision on the CPU (tests/test_gpu_augment.py); the gamma LUT (256 bytes) is supplied by the
// host so that any gamma follows torch's own pow.
#include "common.cuh"

namespace gsb {

__device__ __forceinline__ uint32_t f32_key(float f) {  // monotonic float -> unsigned
  const uint32_t b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float key_f32(uint32_t k) {
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

__global__ void minmax_init_kernel(uint32_t* keys) {
  GSB_PDL_ENTRY();
  keys[0] = 0xffffffffu;  // min
  keys[1] = 0u;           // max
}

__global__ void __launch_bounds__(256) minmax_kernel(const float* __restrict__ x, int64_t n, uint32_t* __restrict__ keys) {
  GSB_PDL_ENTRY();
  uint32_t lo = 0xffffffffu, hi = 0u;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const uint32_t k = f32_key(x[i]);
    lo = min(lo, k);
    hi = max(hi, k);
  }
  for (int o = 16; o > 0; o >>= 1) {
    lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
  }
  __shared__ uint32_t s_lo[8], s_hi[8];
  if ((threadIdx.x & 31) == 0) {
    s_lo[threadIdx.x >> 5] = lo;
    s_hi[threadIdx.x >> 5] = hi;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; ++w) {
      lo = min(lo, s_lo[w]);
      hi = max(hi, s_hi[w]);
    }
    atomicMin(&keys[0], lo);
    atomicMax(&keys[1], hi);
  }
}

// q = uint8(clamp(((x - mn) / (mx - mn)) * 255 + 0.5, 0, 255)), exactly the reference's fp32 operation sequence.
// keys == nullptr: x is already in [0, 1] scale-free form and only mul / add / clamp / truncate apply.
__global__ void __launch_bounds__(256) quant_hist_kernel(const float* __restrict__ x, const uint32_t* __restrict__ keys,
                                                        uint8_t* __restrict__ q, uint32_t* __restrict__ hist, int plane) {
  GSB_PDL_ENTRY();
  __shared__ uint32_t s_hist[256];
  s_hist[threadIdx.x] = 0;
  __syncthreads();
  float mn = 0.f, den = 1.f;
  if (keys != nullptr) {
    mn = key_f32(keys[0]);
    den = __fsub_rn(key_f32(keys[1]), mn);
  }
  const int64_t base = (int64_t)blockIdx.y * plane;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < plane; i += gridDim.x * blockDim.x) {
    float t = x[base + i];
    if (keys != nullptr) t = __fdiv_rn(__fsub_rn(t, mn), den);
    t = __fadd_rn(__fmul_rn(t, 255.f), 0.5f);
    t = fminf(fmaxf(t, 0.f), 255.f);  // NaN (max == min) -> 0, like .to(uint8) of a clamped NaN is unspecified in the reference
    const uint32_t v = (uint32_t)t;
    q[base + i] = (uint8_t)v;
    atomicAdd(&s_hist[v], 1u);
  }
  __syncthreads();
  if (s_hist[threadIdx.x] != 0) atomicAdd(&hist[blockIdx.y * 256 + threadIdx.x], s_hist[threadIdx.x]);
}

__global__ void __launch_bounds__(256) hist_u8_kernel(const uint8_t* __restrict__ q, uint32_t* __restrict__ hist, int plane) {
  GSB_PDL_ENTRY();
  __shared__ uint32_t s_hist[256];
  s_hist[threadIdx.x] = 0;
  __syncthreads();
  const int64_t base = (int64_t)blockIdx.y * plane;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < plane; i += gridDim.x * blockDim.x)
    atomicAdd(&s_hist[q[base + i]], 1u);
  __syncthreads();
  if (s_hist[threadIdx.x] != 0) atomicAdd(&hist[blockIdx.y * 256 + threadIdx.x], s_hist[threadIdx.x]);
}

// torchvision.transforms.functional.equalize (tensor path, _scale_channel) followed by a 256-entry LUT (gamma):
//   step = floor((sum(hist) - last non-zero bin) / 255);  step == 0 -> identity
//   lut[v] = clamp(floor((cumsum[v-1] + step / 2) / step), 0, 255),  lut[0] = 0
// out_f (optional) = post[lut[q]] / 255.0f, out_u8 (optional) = post[lut[q]].
__global__ void __launch_bounds__(256) equalize_lut_kernel(const uint8_t* __restrict__ q, const uint32_t* __restrict__ hist,
                                                          const uint8_t* __restrict__ post, float* __restrict__ out_f,
                                                          uint8_t* __restrict__ out_u8, int plane, int do_equalize) {
  GSB_PDL_ENTRY();
  __shared__ uint32_t s_cum[256];
  __shared__ float s_lut_f[256];
  __shared__ uint8_t s_lut[256];
  __shared__ int s_last;
  const int t = threadIdx.x;
  const uint32_t h = hist[blockIdx.y * 256 + t];
  s_cum[t] = h;
  if (t == 0) s_last = -1;
  __syncthreads();
  if (h != 0) atomicMax(&s_last, t);
  for (int o = 1; o < 256; o <<= 1) {  // inclusive scan
    const uint32_t v = t >= o ? s_cum[t - o] : 0u;
    __syncthreads();
    s_cum[t] += v;
    __syncthreads();
  }
  const uint32_t total = s_cum[255];
  const uint32_t last = s_last >= 0 ? hist[blockIdx.y * 256 + s_last] : 0u;
  const uint32_t step = (total - last) / 255u;
  uint32_t l = (uint32_t)t;
  if (do_equalize && step != 0) {
    l = t == 0 ? 0u : (s_cum[t - 1] + step / 2u) / step;
    l = l > 255u ? 255u : l;
  }
  const uint8_t pv = post != nullptr ? post[l] : (uint8_t)l;
  s_lut[t] = pv;
  s_lut_f[t] = __fdiv_rn((float)pv, 255.f);
  __syncthreads();
  const int64_t base = (int64_t)blockIdx.y * plane;
  for (int i = blockIdx.x * blockDim.x + t; i < plane; i += gridDim.x * blockDim.x) {
    const uint8_t v = q[base + i];
    if (out_f != nullptr) out_f[base + i] = s_lut_f[v];
    if (out_u8 != nullptr) out_u8[base + i] = s_lut[v];
  }
}

// out[n, y, x] = (bilinear(mask[n], M_n * (x, y, 1)) > thr) ? 1 : 0, zero outside the image.  M_n (2x3, row-major) maps
// OUTPUT pixel coordinates to INPUT pixel coordinates (the host composes flip / crop-pad / scale / translate / rotate /
// shear in the drawn order and inverts).  thr < 0: the interpolated value itself is written.
__global__ void __launch_bounds__(256) mask_affine_kernel(const float* __restrict__ in, const float* __restrict__ mats,
                                                         float* __restrict__ out, int h, int w, float thr) {
  GSB_PDL_ENTRY();
  const int n = blockIdx.y;
  const float* m = mats + n * 6;
  const float m0 = m[0], m1 = m[1], m2 = m[2], m3 = m[3], m4 = m[4], m5 = m[5];
  const float* src = in + (int64_t)n * h * w;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < h * w; i += gridDim.x * blockDim.x) {
    const float xo = (float)(i % w), yo = (float)(i / w);
    const float xs = __fadd_rn(__fadd_rn(__fmul_rn(m0, xo), __fmul_rn(m1, yo)), m2);
    const float ys = __fadd_rn(__fadd_rn(__fmul_rn(m3, xo), __fmul_rn(m4, yo)), m5);
    const float xf = floorf(xs), yf = floorf(ys);
    const int x0 = (int)xf, y0 = (int)yf;
    const float ax = __fsub_rn(xs, xf), ay = __fsub_rn(ys, yf);
    auto at = [&](int yy, int xx) -> float { return (yy >= 0 && yy < h && xx >= 0 && xx < w) ? src[yy * w + xx] : 0.f; };
    const float v00 = at(y0, x0), v01 = at(y0, x0 + 1), v10 = at(y0 + 1, x0), v11 = at(y0 + 1, x0 + 1);
    const float top = __fadd_rn(__fmul_rn(v00, __fsub_rn(1.f, ax)), __fmul_rn(v01, ax));
    const float bot = __fadd_rn(__fmul_rn(v10, __fsub_rn(1.f, ax)), __fmul_rn(v11, ax));
    const float v = __fadd_rn(__fmul_rn(top, __fsub_rn(1.f, ay)), __fmul_rn(bot, ay));
    out[(int64_t)n * h * w + i] = thr < 0.f ? v : (v > thr ? 1.f : 0.f);
  }
}

static int plane_blocks(int plane, int planes) {
  int b = (plane + 256 * 8 - 1) / (256 * 8);
  const int cap = (num_sms() * 8 + planes - 1) / planes;
  if (b > cap) b = cap;
  return b < 1 ? 1 : b;
}

}  // namespace gsb

using namespace gsb;

// Ordered keys of (min, max) over x[0..n): keys[0..1] (device, 8 bytes), decoded inside gsb_quantize_hist.
extern "C" int gsb_minmax_f32(const float* x, int64_t n, uint32_t* keys, gsb_stream s) {
  GSB_REQUIRE(x && keys && n > 0, "minmax_f32: bad arguments");
  cudaStream_t st = to_stream(s);
  gsb::launch(minmax_init_kernel, gsb::LaunchCfg(1, 1, 0, st), keys);
  int64_t b = (n + 256 * 8 - 1) / (256 * 8);
  const int64_t cap = (int64_t)num_sms() * 8;
  gsb::launch(minmax_kernel, gsb::LaunchCfg((int)(b > cap ? cap : b), 256, 0, st), x, n, keys);
  return check_launch("minmax_f32", 2);
}

// x: [planes, plane] fp32 -> q uint8 (same shape) and hist [planes, 256] (zeroed here).  keys = output of gsb_minmax_f32
// over the whole tensor (batch min-max normalisation) or NULL (x already in [0,1]).
extern "C" int gsb_quantize_hist(const float* x, const uint32_t* keys, uint8_t* q, uint32_t* hist, int planes, int plane,
                                 gsb_stream s) {
  GSB_REQUIRE(x && q && hist && planes > 0 && plane > 0 && planes <= 65535, "quantize_hist: bad arguments");
  cudaStream_t st = to_stream(s);
  cudaError_t e = cudaMemsetAsync(hist, 0, (size_t)planes * 256 * sizeof(uint32_t), st);
  GSB_REQUIRE(e == cudaSuccess, "quantize_hist: memset failed: %s", cudaGetErrorString(e));
  gsb::launch(quant_hist_kernel, gsb::LaunchCfg(dim3(plane_blocks(plane, planes), planes), 256, 0, st), x, keys, q, hist, plane);
  return check_launch("quantize_hist");
}

// Histogram-equalise (do_equalize != 0) each uint8 plane like torchvision.transforms.functional.equalize, then map
// through `post` (256-entry device LUT, e.g. adjust_gamma; NULL = identity).  hist == NULL: computed here into ws_hist
// ([planes, 256] uint32 workspace).  Writes out_f = value / 255.0f and / or out_u8 (either may be NULL).
extern "C" int gsb_equalize_lut(const uint8_t* q, const uint32_t* hist, uint32_t* ws_hist, const uint8_t* post, float* out_f,
                                uint8_t* out_u8, int planes, int plane, int do_equalize, gsb_stream s) {
  GSB_REQUIRE(q && (hist || ws_hist) && (out_f || out_u8) && planes > 0 && plane > 0 && planes <= 65535,
              "equalize_lut: bad arguments");
  cudaStream_t st = to_stream(s);
  const int blocks = plane_blocks(plane, planes);
  if (hist == nullptr) {
    cudaError_t e = cudaMemsetAsync(ws_hist, 0, (size_t)planes * 256 * sizeof(uint32_t), st);
    GSB_REQUIRE(e == cudaSuccess, "equalize_lut: memset failed: %s", cudaGetErrorString(e));
    gsb::launch(hist_u8_kernel, gsb::LaunchCfg(dim3(blocks, planes), 256, 0, st), q, ws_hist, plane);
    hist = ws_hist;
  }
  gsb::launch(equalize_lut_kernel, gsb::LaunchCfg(dim3(blocks, planes), 256, 0, st), q, hist, post, out_f, out_u8, plane, do_equalize);
  return check_launch("equalize_lut");
}

// Affine warp of n single-channel fp32 masks [n, h, w] (mats: [n, 6] output->input pixel maps, device) + threshold.
extern "C" int gsb_mask_affine(const float* in, const float* mats, float* out, int n, int h, int w, float threshold,
                               gsb_stream s) {
  GSB_REQUIRE(in && mats && out && n > 0 && h > 0 && w > 0 && n <= 65535, "mask_affine: bad arguments");
  gsb::launch(mask_affine_kernel, gsb::LaunchCfg(dim3(plane_blocks(h * w, n), n), 256, 0, to_stream(s)), in, mats, out, h, w, threshold);
  return check_launch("mask_affine");
}

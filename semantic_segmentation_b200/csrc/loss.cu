// Fused segmentation / GAN losses and the integer Dice evaluation.
//
// seg loss = CrossEntropy (c>1) or BCE-with-logits (c==1) + Dice loss with the reference's
// GLOBAL-batch ratio (util/dice_score.py:5-28 with reduce_batch_first=True; training
// composition running_files/train_end2end_jsrt.py:136,182-183).  One pass over the fp32 NCHW
// logits + int64 labels produces the four sums; a second pass writes dlogits.  Eval
// (unet/evaluate.py:29-40) is an argmax / sigmoid>0.5 pass with exact integer counts.
// All HBM-bound: algorithmic bytes = logits (4c B/pixel) + labels (8 B/pixel) [+ dlogits].
#include <mutex>
#include "common.cuh"

namespace gsb {

constexpr int kMaxC = 16;

__device__ __forceinline__ double block_sum_d(double v, double* sh) {
  v = warp_sum_d(v);
  int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) sh[w] = v;
  __syncthreads();
  double r = 0.0;
  if (w == 0) {
    r = l < (blockDim.x >> 5) ? sh[l] : 0.0;
    r = warp_sum_d(r);
  }
  return r;  // valid in thread 0
}

// CT > 0: compile-time class count (the per-class arrays stay in registers: with a run-time bound they were indexed
// dynamically, i.e. lived in local memory); CT == 0: any C <= kMaxC.
template <int CT>
__global__ void __launch_bounds__(256) seg_loss_reduce_kernel(const float* __restrict__ logits,
                                                             const int64_t* __restrict__ labels, int64_t P, int64_t HW,
                                                             int C, double* __restrict__ partial) {
  GSB_PDL_ENTRY();
  __shared__ double sh[8];
  const int Cc = CT > 0 ? CT : C;
  constexpr int CB = CT > 0 ? CT : kMaxC;
  float ce = 0.f, inter = 0.f, sp = 0.f, st = 0.f;
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < P; p += (int64_t)gridDim.x * blockDim.x) {
    int64_t n = p / HW, r = p - n * HW;
    const float* z = logits + n * Cc * HW + r;
    int t = (int)labels[p];
    if (Cc == 1) {
      float x = z[0];
      float pr = 1.f / (1.f + expf(-x));
      ce += fmaxf(x, 0.f) - x * (float)t + log1pf(expf(-fabsf(x)));
      inter += pr * (float)t;
      sp += pr;
      st += (float)t;
    } else {
      float v[kMaxC], m = -INFINITY;
#pragma unroll
      for (int c = 0; c < CB; ++c) {
        if (CT == 0 && c >= Cc) break;
        v[c] = z[(int64_t)c * HW];
        m = fmaxf(m, v[c]);
      }
      float se = 0.f, zt = 0.f, et = 0.f;
#pragma unroll
      for (int c = 0; c < CB; ++c) {
        if (CT == 0 && c >= Cc) break;
        float e = expf(v[c] - m);
        se += e;
        if (c == t) { zt = v[c]; et = e; }
      }
      float inv = 1.f / se;
      ce += (m + logf(se)) - zt;
      inter += et * inv;
      float s = 0.f;
#pragma unroll
      for (int c = 0; c < CB; ++c) {
        if (CT == 0 && c >= Cc) break;
        s += expf(v[c] - m) * inv;
      }
      sp += s;
      st += 1.f;
    }
  }
  double r0 = block_sum_d((double)ce, sh);
  double r1 = block_sum_d((double)inter, sh);
  double r2 = block_sum_d((double)sp, sh);
  double r3 = block_sum_d((double)st, sh);
  if (threadIdx.x == 0) {
    partial[blockIdx.x * 4 + 0] = r0;
    partial[blockIdx.x * 4 + 1] = r1;
    partial[blockIdx.x * 4 + 2] = r2;
    partial[blockIdx.x * 4 + 3] = r3;
  }
}

__global__ void sum_rows4_kernel(const double* __restrict__ partial, int rows, double* __restrict__ sums) {
  GSB_PDL_ENTRY();
  __shared__ double sh[8];
  for (int j = 0; j < 4; ++j) {
    double a = 0.0;
    for (int r = threadIdx.x; r < rows; r += blockDim.x) a += partial[r * 4 + j];
    a = block_sum_d(a, sh);
    if (threadIdx.x == 0) sums[j] = a;
  }
}

__device__ __forceinline__ void dice_terms(const double* sums, double& dice, double& a, double& b) {
  // dice = (2I+eps)/(S+eps); d dice / d p = (2 t (S+eps) - (2I+eps)) / (S+eps)^2 = a*t - b
  const double eps = 1e-6;
  double I2 = 2.0 * sums[1], S = sums[2] + sums[3];
  if (S == 0.0) {  // empty-set rule, util/dice_score.py:14
    dice = (I2 + eps) / (I2 + eps);
    a = b = 0.0;
    return;
  }
  dice = (I2 + eps) / (S + eps);
  a = 2.0 / (S + eps);
  b = (I2 + eps) / ((S + eps) * (S + eps));
}

__global__ void seg_loss_finalize_kernel(const double* __restrict__ sums, double count, float* out) {
  GSB_PDL_ENTRY();
  double dice, a, b;
  dice_terms(sums, dice, a, b);
  double ce = sums[0] / count;
  out[1] = (float)ce;
  out[2] = (float)(1.0 - dice);
  out[0] = (float)ce + (float)(1.0 - dice);
}

template <int CT>
__global__ void __launch_bounds__(256) seg_loss_bwd_kernel(const float* __restrict__ logits, const int64_t* __restrict__ labels,
                                                          int64_t P, int64_t HW, int C, const double* __restrict__ sums,
                                                          double count, const float* __restrict__ gscale,
                                                          float* __restrict__ dlogits) {
  GSB_PDL_ENTRY();
  const int Cc = CT > 0 ? CT : C;
  constexpr int CB = CT > 0 ? CT : kMaxC;
  double dice, ad, bd;
  dice_terms(sums, dice, ad, bd);
  const float gs = gscale ? *gscale : 1.f;
  const float a = (float)ad, b = (float)bd, invn = (float)(1.0 / count);
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < P; p += (int64_t)gridDim.x * blockDim.x) {
    int64_t n = p / HW, r = p - n * HW;
    const float* z = logits + n * Cc * HW + r;
    float* dz = dlogits + n * Cc * HW + r;
    int t = (int)labels[p];
    if (Cc == 1) {
      float x = z[0];
      float pr = 1.f / (1.f + expf(-x));
      float g = -(a * (float)t - b);  // d(1-dice)/dp
      dz[0] = gs * ((pr - (float)t) * invn + g * pr * (1.f - pr));
    } else {
      float v[kMaxC], m = -INFINITY;
#pragma unroll
      for (int c = 0; c < CB; ++c) {
        if (CT == 0 && c >= Cc) break;
        v[c] = z[(int64_t)c * HW];
        m = fmaxf(m, v[c]);
      }
      float se = 0.f;
#pragma unroll
      for (int c = 0; c < CB; ++c) {
        if (CT == 0 && c >= Cc) break;
        v[c] = expf(v[c] - m);
        se += v[c];
      }
      float inv = 1.f / se, dot = 0.f;
#pragma unroll
      for (int c = 0; c < CB; ++c) {
        if (CT == 0 && c >= Cc) break;
        v[c] *= inv;
        float g = -(a * (c == t ? 1.f : 0.f) - b);
        dot += v[c] * g;
      }
#pragma unroll
      for (int c = 0; c < CB; ++c) {
        if (CT == 0 && c >= Cc) break;
        float g = -(a * (c == t ? 1.f : 0.f) - b);
        dz[(int64_t)c * HW] = gs * ((v[c] - (c == t ? 1.f : 0.f)) * invn + v[c] * (g - dot));
      }
    }
  }
}

// grid = (chunks, n).  counts[n][K][3] (K = max(C-1,1)) int64: inter, pred, true.
__global__ void __launch_bounds__(256) eval_counts_kernel(const float* __restrict__ logits, const int64_t* __restrict__ labels,
                                                         int64_t HW, int C, uint8_t* __restrict__ mask,
                                                         unsigned long long* __restrict__ counts) {
  GSB_PDL_ENTRY();
  __shared__ unsigned int sc[kMaxC * 3];
  const int K = C > 1 ? C - 1 : 1;
  for (int i = threadIdx.x; i < K * 3; i += blockDim.x) sc[i] = 0;
  __syncthreads();
  const int n = blockIdx.y;
  const int64_t iters = (HW + (int64_t)gridDim.x * blockDim.x - 1) / ((int64_t)gridDim.x * blockDim.x);
  for (int64_t it = 0; it < iters; ++it) {
    int64_t r = (it * gridDim.x + blockIdx.x) * blockDim.x + threadIdx.x;
    bool valid = r < HW;
    int pred = -1, tru = -1;
    if (valid) {
      const float* z = logits + (int64_t)n * C * HW + r;
      tru = (int)labels[(int64_t)n * HW + r];
      if (C == 1) {
        float x = z[0];
        float sg = 1.f / (1.f + expf(-x));
        pred = sg > 0.5f ? 1 : 0;
      } else {
        float best = z[0];
        pred = 0;
        for (int c = 1; c < C; ++c) {
          float v = z[(int64_t)c * HW];
          if (v > best || (v != v && !(best != best))) { best = v; pred = c; }
        }
      }
      if (mask) mask[(int64_t)n * HW + r] = (uint8_t)pred;
    }
    for (int k = 1; k <= K; ++k) {
      unsigned bp = __ballot_sync(0xffffffffu, valid && pred == k);
      unsigned bt = __ballot_sync(0xffffffffu, valid && tru == k);
      if ((threadIdx.x & 31) == 0) {
        unsigned ci = __popc(bp & bt), cp = __popc(bp), ct = __popc(bt);
        if (ci) atomicAdd(&sc[(k - 1) * 3 + 0], ci);
        if (cp) atomicAdd(&sc[(k - 1) * 3 + 1], cp);
        if (ct) atomicAdd(&sc[(k - 1) * 3 + 2], ct);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < K * 3; i += blockDim.x)
    if (sc[i]) atomicAdd(&counts[(int64_t)n * K * 3 + i], (unsigned long long)sc[i]);
}

// partial[blockIdx.x][slice][3]: per-CTA sums, added in row order by sum_rows_kernel (deterministic, no atomics)
__global__ void __launch_bounds__(256) dice_sums_kernel(const float* __restrict__ a, const float* __restrict__ b, int64_t inner,
                                                       double* __restrict__ partial) {
  GSB_PDL_ENTRY();
  __shared__ double sh[8];
  const int64_t s = blockIdx.y;
  const float* pa = a + s * inner;
  const float* pb = b + s * inner;
  float ab = 0.f, sa = 0.f, sb = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < inner; i += (int64_t)gridDim.x * blockDim.x) {
    float x = pa[i], y = pb[i];
    ab = fmaf(x, y, ab);
    sa += x;
    sb += y;
  }
  double r0 = block_sum_d((double)ab, sh), r1 = block_sum_d((double)sa, sh), r2 = block_sum_d((double)sb, sh);
  if (threadIdx.x == 0) {
    double* row = partial + ((int64_t)blockIdx.x * gridDim.y + s) * 3;
    row[0] = r0; row[1] = r1; row[2] = r2;
  }
}

// sums[j] = (sum over rows of partial[r][j]) / divisor, fixed order (thread-strided, shuffle tree, warp sums in index order);
// out_f (optional) receives the fp32 value of column 0
__global__ void __launch_bounds__(256) sum_rows_kernel(const double* __restrict__ partial, int rows, int ncols, double divisor,
                                                      double* __restrict__ sums, float* __restrict__ out_f) {
  GSB_PDL_ENTRY();
  __shared__ double sh[8];
  for (int j = 0; j < ncols; ++j) {
    double a = 0.0;
    for (int r = threadIdx.x; r < rows; r += blockDim.x) a += partial[(int64_t)r * ncols + j];
    a = block_sum_d(a, sh);
    if (threadIdx.x == 0) {
      if (sums) sums[j] = a / divisor;
      if (out_f && j == 0) out_f[0] = (float)(a / divisor);
    }
  }
}

// mode 0 vanilla (BCE with logits), 1 lsgan (MSE), 2 wgangp.  acc[blockIdx.x] = this CTA's sum of the per-element loss.
__global__ void __launch_bounds__(256) gan_loss_kernel(const float* __restrict__ pred, int64_t n, int mode, float label,
                                                      double* acc, const float* __restrict__ gscale, float* __restrict__ dpred) {
  GSB_PDL_ENTRY();
  __shared__ double sh[8];
  float l = 0.f;
  const float gs = (gscale ? *gscale : 1.f) / (float)n;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float x = pred[i], g;
    if (mode == 0) {
      l += fmaxf(x, 0.f) - x * label + log1pf(expf(-fabsf(x)));
      g = 1.f / (1.f + expf(-x)) - label;
    } else if (mode == 1) {
      float d = x - label;
      l += d * d;
      g = 2.f * d;
    } else {
      float sgn = label > 0.5f ? -1.f : 1.f;
      l += sgn * x;
      g = sgn;
    }
    if (dpred) dpred[i] = gs * g;
  }
  if (acc) {
    double r = block_sum_d((double)l, sh);
    if (threadIdx.x == 0) acc[blockIdx.x] = r;
  }
}

__global__ void __launch_bounds__(256) l1_loss_kernel(const float* __restrict__ a, const float* __restrict__ b, int64_t n,
                                                     double* acc, const float* __restrict__ gscale, float* __restrict__ da) {
  GSB_PDL_ENTRY();
  __shared__ double sh[8];
  float l = 0.f;
  const float gs = (gscale ? *gscale : 1.f) / (float)n;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float d = a[i] - b[i];
    l += fabsf(d);
    if (da) da[i] = d > 0.f ? gs : (d < 0.f ? -gs : 0.f);
  }
  if (acc) {
    double r = block_sum_d((double)l, sh);
    if (threadIdx.x == 0) acc[blockIdx.x] = r;
  }
}

__global__ void __launch_bounds__(256) act_bwd_kernel(const float* __restrict__ out, const float* __restrict__ dout, int act,
                                                     float* __restrict__ dx, int64_t n) {
  GSB_PDL_ENTRY();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float o = out[i], g = dout[i], d;
    switch (act) {
      case GSB_ACT_RELU: d = o > 0.f ? 1.f : 0.f; break;
      case GSB_ACT_LRELU: d = o > 0.f ? 1.f : 0.2f; break;
      case GSB_ACT_TANH: d = 1.f - o * o; break;
      default: d = 1.f;
    }
    dx[i] = g * d;
  }
}

static int grid1d(int64_t total, int per_sm = 8) {
  int64_t b = (total + 255) / 256;
  int64_t cap = (int64_t)num_sms() * per_sm;
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

}  // namespace gsb

using namespace gsb;

extern "C" size_t gsb_seg_loss_workspace_bytes(int n, int c, int h, int w) {
  (void)n; (void)c; (void)h; (void)w;
  return (size_t)num_sms() * 8 * 4 * sizeof(double);
}

extern "C" int gsb_seg_loss_reduce(const float* logits, const int64_t* labels, int n, int c, int h, int w, double* sums,
                                   void* workspace, size_t workspace_bytes, gsb_stream s) {
  GSB_REQUIRE(logits && labels && sums && workspace, "seg_loss_reduce: null pointer");
  GSB_REQUIRE(c >= 1 && c <= kMaxC, "seg_loss: n_classes %d unsupported (max %d)", c, kMaxC);
  int64_t HW = (int64_t)h * w, P = (int64_t)n * HW;
  int grid = grid1d(P);
  GSB_REQUIRE(workspace_bytes >= (size_t)grid * 4 * sizeof(double), "seg_loss_reduce: workspace too small");
  double* partial = reinterpret_cast<double*>(workspace);
  if (c == 2) gsb::launch(seg_loss_reduce_kernel<2>, gsb::LaunchCfg(grid, 256, 0, to_stream(s)), logits, labels, P, HW, c, partial);
  else if (c == 1) gsb::launch(seg_loss_reduce_kernel<1>, gsb::LaunchCfg(grid, 256, 0, to_stream(s)), logits, labels, P, HW, c, partial);
  else gsb::launch(seg_loss_reduce_kernel<0>, gsb::LaunchCfg(grid, 256, 0, to_stream(s)), logits, labels, P, HW, c, partial);
  gsb::launch(sum_rows4_kernel, gsb::LaunchCfg(1, 256, 0, to_stream(s)), partial, grid, sums);
  return check_launch("seg_loss_reduce", 2);
}

extern "C" int gsb_seg_loss_finalize(const double* sums, double count, float* out, gsb_stream s) {
  GSB_REQUIRE(sums && out && count > 0, "seg_loss_finalize: bad argument");
  gsb::launch(seg_loss_finalize_kernel, gsb::LaunchCfg(1, 1, 0, to_stream(s)), sums, count, out);
  return check_launch("seg_loss_finalize");
}

extern "C" int gsb_seg_loss_bwd(const float* logits, const int64_t* labels, int n, int c, int h, int w, const double* sums,
                                double count, const float* gscale, float* dlogits, gsb_stream s) {
  GSB_REQUIRE(logits && labels && sums && dlogits && count > 0, "seg_loss_bwd: bad argument");
  GSB_REQUIRE(c >= 1 && c <= kMaxC, "seg_loss: n_classes %d unsupported (max %d)", c, kMaxC);
  int64_t HW = (int64_t)h * w, P = (int64_t)n * HW;
  const gsb::LaunchCfg lc(grid1d(P, 16), 256, 0, to_stream(s));
  if (c == 2) gsb::launch(seg_loss_bwd_kernel<2>, lc, logits, labels, P, HW, c, sums, count, gscale, dlogits);
  else if (c == 1) gsb::launch(seg_loss_bwd_kernel<1>, lc, logits, labels, P, HW, c, sums, count, gscale, dlogits);
  else gsb::launch(seg_loss_bwd_kernel<0>, lc, logits, labels, P, HW, c, sums, count, gscale, dlogits);
  return check_launch("seg_loss_bwd");
}

extern "C" int gsb_eval_counts(const float* logits, const int64_t* labels, int n, int c, int h, int w, uint8_t* mask,
                               int64_t* counts, gsb_stream s) {
  GSB_REQUIRE(logits && labels && counts, "eval_counts: null pointer");
  GSB_REQUIRE(c >= 1 && c <= kMaxC, "eval_counts: n_classes %d unsupported (max %d)", c, kMaxC);
  int K = c > 1 ? c - 1 : 1;
  cudaMemsetAsync(counts, 0, (size_t)n * K * 3 * sizeof(int64_t), to_stream(s));
  int64_t HW = (int64_t)h * w;
  int gx = (int)((HW + 256 * 8 - 1) / (256 * 8));
  if (gx < 1) gx = 1;
  dim3 grid((unsigned)gx, (unsigned)n);
  gsb::launch(eval_counts_kernel, gsb::LaunchCfg(grid, 256, 0, to_stream(s)), logits, labels, HW, c, mask, reinterpret_cast<unsigned long long*>(counts));
  return check_launch("eval_counts");
}

// workspace: fp64 partial rows [gx][slices][3]; gx = min(ceil(inner / 4096), 1024, what the workspace holds)
extern "C" int gsb_dice_sums(const float* a, const float* b, int64_t slices, int64_t inner, double* sums, void* workspace,
                             size_t workspace_bytes, gsb_stream s) {
  GSB_REQUIRE(a && b && sums && workspace && slices > 0 && slices < 65536, "dice_sums: bad argument");
  const size_t row_bytes = (size_t)slices * 3 * sizeof(double);
  GSB_REQUIRE(workspace_bytes >= row_bytes, "dice_sums: workspace too small");
  int64_t gx = (inner + 256 * 16 - 1) / (256 * 16);
  if (gx < 1) gx = 1;
  if (gx > 1024) gx = 1024;
  if ((size_t)gx * row_bytes > workspace_bytes) gx = (int64_t)(workspace_bytes / row_bytes);
  double* partial = reinterpret_cast<double*>(workspace);
  dim3 grid((unsigned)gx, (unsigned)slices);
  gsb::launch(dice_sums_kernel, gsb::LaunchCfg(grid, 256, 0, to_stream(s)), a, b, inner, partial);
  gsb::launch(sum_rows_kernel, gsb::LaunchCfg(1, 256, 0, to_stream(s)), (const double*)partial, (int)gx, (int)(slices * 3), 1.0, sums,
              (float*)nullptr);
  return check_launch("dice_sums", 2);
}

// ---------------------------------------------------------------------------------------------------------
// gsb_dice_sums_noalloc: gsb_dice_sums for callers that cannot allocate (no counterpart in the reference).
// The fp64 partial rows live in a file-scope device array of 96 KiB that every GPU
// loading the library holds until exit.  In HOST code `g_dice_noalloc_scratch` is only a stand-in for the symbol, not
// a device address: the address is looked up once per device with cudaGetSymbolAddress (that device current) and it
// is that pointer the kernels receive.  All calls share the array: calls that overlap on the GPU (two streams) race.
// Nothing has to start at zero: every row the summation reads is stored by the launch in front of it.
// ---------------------------------------------------------------------------------------------------------
constexpr int kDiceNoAllocDoubles = 12288;      // 96 KiB
constexpr int kDiceNoAllocMaxSlices = 64;       // -> at least 64 partial rows
static __device__ __align__(256) double g_dice_noalloc_scratch[kDiceNoAllocDoubles];

static double* dice_noalloc_scratch_ptr() {   // null + gsb_last_error on failure
  constexpr int kMaxDev = 64;
  static double* ptrs[kMaxDev] = {};
  static std::mutex mu;
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDev) {
    gsb::set_error("dice_sums_noalloc: cudaGetDevice failed or device index %d out of range (%s)", dev, cudaGetErrorString(e));
    return nullptr;
  }
  std::lock_guard<std::mutex> lock(mu);
  if (ptrs[dev] == nullptr) {
    double* p = nullptr;
    e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_dice_noalloc_scratch);   // the current device's copy
    if (e != cudaSuccess || p == nullptr) {
      gsb::set_error("dice_sums_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
      return nullptr;
    }
    ptrs[dev] = p;
  }
  return ptrs[dev];
}

extern "C" size_t gsb_dice_sums_noalloc_max_slices(void) { return (size_t)kDiceNoAllocMaxSlices; }

extern "C" int gsb_dice_sums_noalloc(const float* a, const float* b, int64_t slices, int64_t inner, double* sums, gsb_stream s) {
  GSB_REQUIRE(a && b && sums && inner > 0, "dice_sums_noalloc: bad argument");
  GSB_REQUIRE(slices > 0 && slices <= kDiceNoAllocMaxSlices,
              "dice_sums_noalloc: %lld slices exceed the built-in scratch (max %d slices, %d KiB); use gsb_dice_sums with a "
              "caller workspace", (long long)slices, kDiceNoAllocMaxSlices, (int)(kDiceNoAllocDoubles * sizeof(double) / 1024));
  double* partial = dice_noalloc_scratch_ptr();
  if (partial == nullptr) return -1;
  int64_t gx = (inner + 256 * 16 - 1) / (256 * 16);   // same row count rule as gsb_dice_sums
  if (gx < 1) gx = 1;
  if (gx > 1024) gx = 1024;
  const int64_t cap = kDiceNoAllocDoubles / (slices * 3);
  if (gx > cap) gx = cap;
  dim3 grid((unsigned)gx, (unsigned)slices);
  gsb::launch(dice_sums_kernel, gsb::LaunchCfg(grid, 256, 0, to_stream(s)), a, b, inner, partial);
  gsb::launch(sum_rows_kernel, gsb::LaunchCfg(1, 256, 0, to_stream(s)), (const double*)partial, (int)gx, (int)(slices * 3), 1.0, sums,
              (float*)nullptr);
  return check_launch("dice_sums_noalloc", 2);
}

// workspace (forward value only): one double per CTA; the launch uses min(grid, workspace_bytes / 8) CTAs
extern "C" int gsb_gan_loss(const float* pred, int64_t numel, int mode, float label, float* loss, const float* gscale,
                            float* dpred, void* workspace, size_t workspace_bytes, gsb_stream s) {
  GSB_REQUIRE(pred && numel > 0 && mode >= 0 && mode <= 2, "gan_loss: bad argument");
  GSB_REQUIRE(loss == nullptr || (workspace != nullptr && workspace_bytes >= sizeof(double)),
              "gan_loss: workspace (>= 8 bytes) required for the forward value");
  double* acc = loss ? reinterpret_cast<double*>(workspace) : nullptr;
  int grid = grid1d(numel, 4);
  if (acc && (size_t)grid * sizeof(double) > workspace_bytes) grid = (int)(workspace_bytes / sizeof(double));
  gsb::launch(gan_loss_kernel, gsb::LaunchCfg(grid, 256, 0, to_stream(s)), pred, numel, mode, label, acc, gscale, dpred);
  if (acc)
    gsb::launch(sum_rows_kernel, gsb::LaunchCfg(1, 256, 0, to_stream(s)), (const double*)acc, grid, 1, (double)numel, (double*)nullptr, loss);
  return check_launch("gan_loss", acc ? 2 : 1);
}

extern "C" int gsb_l1_loss(const float* a, const float* b, int64_t numel, float* loss, const float* gscale, float* da,
                           void* workspace, size_t workspace_bytes, gsb_stream s) {
  GSB_REQUIRE(a && b && numel > 0, "l1_loss: bad argument");
  GSB_REQUIRE(loss == nullptr || (workspace != nullptr && workspace_bytes >= sizeof(double)),
              "l1_loss: workspace (>= 8 bytes) required for the forward value");
  double* acc = loss ? reinterpret_cast<double*>(workspace) : nullptr;
  int grid = grid1d(numel, 8);
  if (acc && (size_t)grid * sizeof(double) > workspace_bytes) grid = (int)(workspace_bytes / sizeof(double));
  gsb::launch(l1_loss_kernel, gsb::LaunchCfg(grid, 256, 0, to_stream(s)), a, b, numel, acc, gscale, da);
  if (acc)
    gsb::launch(sum_rows_kernel, gsb::LaunchCfg(1, 256, 0, to_stream(s)), (const double*)acc, grid, 1, (double)numel, (double*)nullptr, loss);
  return check_launch("l1_loss", acc ? 2 : 1);
}

extern "C" int gsb_act_bwd(const float* out, const float* dout, int act, float* dx, int64_t numel, gsb_stream s) {
  GSB_REQUIRE(out && dout && dx, "act_bwd: null pointer");
  if (numel == 0) return 0;
  gsb::launch(act_bwd_kernel, gsb::LaunchCfg(grid1d(numel, 8), 256, 0, to_stream(s)), out, dout, act, dx, numel);
  return check_launch("act_bwd");
}

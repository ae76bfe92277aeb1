// Fused flat-buffer optimisers: torch.optim.RMSprop(momentum=0.9, weight_decay=1e-8)
// of running_files/train_end2end_jsrt.py:71-72 and torch.optim.Adam(betas=(0.5, 0.999)) of
// models_pix2pix/pix2pix_model.py:75-78, over one contiguous fp32 parameter range per launch
// (parameters and gradients live in flat buffers, so one kernel replaces ~110 foreach launches
// and the gradient all-reduce averaging folds in as grad_scale).  HBM-bound:
// RMSprop = 4 reads + 3 writes, Adam = 4 reads + 3 writes of 4 B per parameter.
#include "common.cuh"

namespace gsb {

__global__ void __launch_bounds__(256) rmsprop_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                     float* __restrict__ sq, float* __restrict__ buf, int64_t n, float lr,
                                                     float alpha, float eps, float wd, float momentum, float gscale) {
  GSB_PDL_ENTRY();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float w = p[i];
    float gr = g[i] * gscale;
    if (wd != 0.f) gr = fmaf(wd, w, gr);
    float s = alpha * sq[i] + (1.f - alpha) * gr * gr;
    sq[i] = s;
    float avg = sqrtf(s) + eps;
    if (momentum > 0.f) {
      float b = momentum * buf[i] + gr / avg;
      buf[i] = b;
      p[i] = w - lr * b;
    } else {
      p[i] = w - lr * gr / avg;
    }
  }
}

__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                                  float* __restrict__ v, int64_t n, float step_size, float beta1, float beta2,
                                                  float eps, float wd, float inv_sqrt_bc2, float gscale) {
  GSB_PDL_ENTRY();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float w = p[i];
    float gr = g[i] * gscale;
    if (wd != 0.f) gr = fmaf(wd, w, gr);
    float mi = beta1 * m[i] + (1.f - beta1) * gr;
    float vi = beta2 * v[i] + (1.f - beta2) * gr * gr;
    m[i] = mi;
    v[i] = vi;
    float denom = sqrtf(vi) * inv_sqrt_bc2 + eps;
    p[i] = w - step_size * (mi / denom);
  }
}

static int grid1d(int64_t total) {
  int64_t b = (total + 255) / 256;
  int64_t cap = (int64_t)num_sms() * 16;
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

}  // namespace gsb

using namespace gsb;

extern "C" int gsb_rmsprop_step(float* p, const float* g, float* square_avg, float* mom_buf, int64_t numel, float lr,
                                float alpha, float eps, float weight_decay, float momentum, float grad_scale, gsb_stream s) {
  GSB_REQUIRE(p && g && square_avg && (momentum == 0.f || mom_buf), "rmsprop_step: null pointer");
  if (numel == 0) return 0;
  gsb::launch(rmsprop_kernel, gsb::LaunchCfg(grid1d(numel), 256, 0, to_stream(s)), p, g, square_avg, mom_buf, numel, lr, alpha, eps, weight_decay,
                                                          momentum, grad_scale);
  return check_launch("rmsprop_step");
}

extern "C" int gsb_adam_step(float* p, const float* g, float* exp_avg, float* exp_avg_sq, int64_t numel, float lr, float beta1,
                             float beta2, float eps, float weight_decay, int64_t step, float grad_scale, gsb_stream s) {
  GSB_REQUIRE(p && g && exp_avg && exp_avg_sq && step >= 1, "adam_step: bad argument");
  if (numel == 0) return 0;
  double bc1 = 1.0 - pow((double)beta1, (double)step), bc2 = 1.0 - pow((double)beta2, (double)step);
  gsb::launch(adam_kernel, gsb::LaunchCfg(grid1d(numel), 256, 0, to_stream(s)), p, g, exp_avg, exp_avg_sq, numel, (float)(lr / bc1), beta1, beta2, eps,
                                                       weight_decay, (float)(1.0 / sqrt(bc2)), grad_scale);
  return check_launch("adam_step");
}

// ---------------------------------------------------------------------------------------------------------
// Multi-tensor variants: every parameter tensor of a network in ONE launch (block -> (tensor, chunk) table in the
// kernel parameters) instead of one 5 us launch per tensor (56 per U-Net step).
// ---------------------------------------------------------------------------------------------------------
namespace gsb {
constexpr int kMaxOptJobs = 48;
constexpr int kOptPerBlock = 256 * 8;
struct OptJobs {
  float* p[kMaxOptJobs];
  const float* g[kMaxOptJobs];
  float* s1[kMaxOptJobs];
  float* s2[kMaxOptJobs];
  int64_t n[kMaxOptJobs];
  int blk0[kMaxOptJobs + 1];
  int count;
};
__device__ __forceinline__ int opt_find_job(const int* blk0, int count, int b) {
  int lo = 0, hi = count - 1;
  while (lo < hi) {
    int mid = (lo + hi + 1) >> 1;
    if (blk0[mid] <= b) lo = mid; else hi = mid - 1;
  }
  return lo;
}

__global__ void __launch_bounds__(256) rmsprop_multi_kernel(const __grid_constant__ OptJobs J, float lr, float alpha, float eps,
                                                           float wd, float momentum, float gscale) {
  GSB_PDL_ENTRY();
  const int j = opt_find_job(J.blk0, J.count, blockIdx.x);
  float* __restrict__ p = J.p[j];
  const float* __restrict__ g = J.g[j];
  float* __restrict__ sq = J.s1[j];
  float* __restrict__ buf = J.s2[j];
  const int64_t n = J.n[j], base = (int64_t)(blockIdx.x - J.blk0[j]) * kOptPerBlock;
#pragma unroll
  for (int u = 0; u < 8; ++u) {
    const int64_t i = base + u * 256 + threadIdx.x;
    if (i < n) {
      float w = p[i];
      float gr = g[i] * gscale;
      if (wd != 0.f) gr = fmaf(wd, w, gr);
      float s = alpha * sq[i] + (1.f - alpha) * gr * gr;
      sq[i] = s;
      float avg = sqrtf(s) + eps;
      if (momentum > 0.f) {
        float b = momentum * buf[i] + gr / avg;
        buf[i] = b;
        p[i] = w - lr * b;
      } else {
        p[i] = w - lr * gr / avg;
      }
    }
  }
}

// DEV: the step count lives in device memory (`step_dev`, already incremented for this step) and the bias corrections are
// formed here -- a captured CUDA graph of the training step then advances like the eager loop (with a host-side count the
// corrections of the capture step would be frozen into the graph).  step_size carries lr in that case.
template <bool DEV>
__global__ void __launch_bounds__(256) adam_multi_kernel(const __grid_constant__ OptJobs J, float step_size, float beta1,
                                                        float beta2, float eps, float wd, float inv_sqrt_bc2, float gscale,
                                                        const int64_t* __restrict__ step_dev) {
  GSB_PDL_ENTRY();
  if (DEV) {
    __shared__ float s_bc[2];
    if (threadIdx.x == 0) {
      const double t = (double)*step_dev;
      const double bc1 = 1.0 - pow((double)beta1, t), bc2 = 1.0 - pow((double)beta2, t);
      s_bc[0] = (float)((double)step_size / bc1);
      s_bc[1] = (float)(1.0 / sqrt(bc2));
    }
    __syncthreads();
    step_size = s_bc[0];
    inv_sqrt_bc2 = s_bc[1];
  }
  const int j = opt_find_job(J.blk0, J.count, blockIdx.x);
  float* __restrict__ p = J.p[j];
  const float* __restrict__ g = J.g[j];
  float* __restrict__ m = J.s1[j];
  float* __restrict__ v = J.s2[j];
  const int64_t n = J.n[j], base = (int64_t)(blockIdx.x - J.blk0[j]) * kOptPerBlock;
#pragma unroll
  for (int u = 0; u < 8; ++u) {
    const int64_t i = base + u * 256 + threadIdx.x;
    if (i < n) {
      float w = p[i];
      float gr = g[i] * gscale;
      if (wd != 0.f) gr = fmaf(wd, w, gr);
      float mi = beta1 * m[i] + (1.f - beta1) * gr;
      float vi = beta2 * v[i] + (1.f - beta2) * gr * gr;
      m[i] = mi;
      v[i] = vi;
      float denom = sqrtf(vi) * inv_sqrt_bc2 + eps;
      p[i] = w - step_size * (mi / denom);
    }
  }
}

template <typename F>
static int run_opt_multi(int n, float* const* p, const float* const* g, float* const* s1, float* const* s2, const int64_t* numel,
                         bool need_s2, const char* what, F launch) {
  int launches = 0;
  for (int begin = 0; begin < n; begin += kMaxOptJobs) {
    OptJobs J;
    memset(&J, 0, sizeof(J));
    J.count = n - begin < kMaxOptJobs ? n - begin : kMaxOptJobs;
    int blocks = 0;
    for (int k = 0; k < J.count; ++k) {
      const int i = begin + k;
      GSB_REQUIRE(p[i] && g[i] && s1[i] && (!need_s2 || s2[i]) && numel[i] >= 0, "%s: bad tensor %d", what, i);
      J.p[k] = p[i]; J.g[k] = g[i]; J.s1[k] = s1[i]; J.s2[k] = s2 ? s2[i] : nullptr;
      J.n[k] = numel[i];
      J.blk0[k] = blocks;
      blocks += (int)((numel[i] + kOptPerBlock - 1) / kOptPerBlock);
    }
    J.blk0[J.count] = blocks;
    if (blocks == 0) continue;
    launch(J, blocks);
    ++launches;
  }
  return check_launch(what, launches > 0 ? launches : 1);
}
}  // namespace gsb

extern "C" int gsb_rmsprop_step_multi(int n, float* const* p, const float* const* g, float* const* square_avg,
                                      float* const* mom_buf, const int64_t* numel, float lr, float alpha, float eps,
                                      float weight_decay, float momentum, float grad_scale, gsb_stream s) {
  if (n <= 0) return 0;
  GSB_REQUIRE(p && g && square_avg && numel && (momentum == 0.f || mom_buf), "rmsprop_step_multi: null argument");
  cudaStream_t st = gsb::to_stream(s);
  return gsb::run_opt_multi(n, p, g, square_avg, mom_buf, numel, momentum != 0.f, "rmsprop_step_multi",
                            [&](const gsb::OptJobs& J, int blocks) {
                              gsb::launch(gsb::rmsprop_multi_kernel, gsb::LaunchCfg(blocks, 256, 0, st), J, lr, alpha, eps, weight_decay, momentum, grad_scale);
                            });
}

extern "C" int gsb_adam_step_multi(int n, float* const* p, const float* const* g, float* const* exp_avg, float* const* exp_avg_sq,
                                   const int64_t* numel, float lr, float beta1, float beta2, float eps, float weight_decay,
                                   int64_t step, float grad_scale, gsb_stream s) {
  if (n <= 0) return 0;
  GSB_REQUIRE(p && g && exp_avg && exp_avg_sq && numel && step >= 1, "adam_step_multi: bad argument");
  double bc1 = 1.0 - pow((double)beta1, (double)step), bc2 = 1.0 - pow((double)beta2, (double)step);
  cudaStream_t st = gsb::to_stream(s);
  const float step_size = (float)(lr / bc1), isb = (float)(1.0 / sqrt(bc2));
  return gsb::run_opt_multi(n, p, g, exp_avg, exp_avg_sq, numel, true, "adam_step_multi", [&](const gsb::OptJobs& J, int blocks) {
    gsb::launch(gsb::adam_multi_kernel<false>, gsb::LaunchCfg(blocks, 256, 0, st), J, step_size, beta1, beta2, eps, weight_decay, isb,
                grad_scale, (const int64_t*)nullptr);
  });
}

// Adam with the step count in device memory (CUDA-graph capturable training steps): *step_dev = the count of THIS step
// (>= 1; the caller advances it with gsb_counter_add before the call, once per optimiser step).
extern "C" int gsb_adam_step_multi_dev(int n, float* const* p, const float* const* g, float* const* exp_avg, float* const* exp_avg_sq,
                                       const int64_t* numel, float lr, float beta1, float beta2, float eps, float weight_decay,
                                       const int64_t* step_dev, float grad_scale, gsb_stream s) {
  if (n <= 0) return 0;
  GSB_REQUIRE(p && g && exp_avg && exp_avg_sq && numel && step_dev, "adam_step_multi_dev: bad argument");
  cudaStream_t st = gsb::to_stream(s);
  return gsb::run_opt_multi(n, p, g, exp_avg, exp_avg_sq, numel, true, "adam_step_multi_dev", [&](const gsb::OptJobs& J, int blocks) {
    gsb::launch(gsb::adam_multi_kernel<true>, gsb::LaunchCfg(blocks, 256, 0, st), J, lr, beta1, beta2, eps, weight_decay, 1.f, grad_scale,
                step_dev);
  });
}

namespace gsb {
__global__ void counter_add_kernel(int64_t* ctr, int64_t delta) {
  GSB_PDL_ENTRY();
  if (threadIdx.x == 0 && blockIdx.x == 0) *ctr += delta;
}
}  // namespace gsb

// *ctr += delta on the stream (device-side step / call counters of captured training steps)
extern "C" int gsb_counter_add(int64_t* ctr, int64_t delta, gsb_stream s) {
  GSB_REQUIRE(ctr != nullptr, "counter_add: null counter");
  gsb::launch(gsb::counter_add_kernel, gsb::LaunchCfg(1, 32, 0, gsb::to_stream(s)), ctr, delta);
  return check_launch("counter_add");
}

// Shared device/host helpers for libgenseg_b200 (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/genseg_b200.h"

namespace gsb {

// ---- errors ---------------------------------------------------------------------------
void set_error(const char* fmt, ...);
int check_launch(const char* what, int n = 1);  // n = kernels launched by this entry point

#define GSB_REQUIRE(cond, ...)          \
  do {                                  \
    if (!(cond)) {                      \
      gsb::set_error(__VA_ARGS__);      \
      return -1;                        \
    }                                   \
  } while (0)

static inline cudaStream_t to_stream(gsb_stream s) { return reinterpret_cast<cudaStream_t>(s); }

// ---- launches: programmatic dependent launch (PDL) ------------------------------------------------------------------
// Every kernel of the library is launched through gsb::launch with cudaLaunchAttributeProgrammaticStreamSerialization and
// begins with GSB_PDL_ENTRY(): `griddepcontrol.launch_dependents` lets the NEXT kernel of the stream (or of the captured
// graph) be scheduled as soon as every CTA of this one has started, so its launch latency, block scheduling and prologue
// (barrier init, TMEM allocation, descriptor prefetch) overlap this kernel's tail; `griddepcontrol.wait` then blocks until
// the PREVIOUS kernel has completed and its memory is visible.  Rule: no global-memory access before the wait.
// Measured (r01r, CUDA-graph replay of the U-Net step, 170 launches): 4.253 -> 4.200 ms at 8 images per GPU, 22.14 -> 22.54 ms
// at 64 -- within run-to-run noise: graph replay already hides the launch bubbles.  The attribute is therefore OFF unless
// GSB_PDL=1 (the device instructions are no-ops without it); the launch helper stays as the single launch path.
bool pdl_enabled();
struct LaunchCfg {
  dim3 grid, block;
  size_t smem;
  cudaStream_t stream;
  LaunchCfg(dim3 g, dim3 b, size_t s = 0, cudaStream_t st = 0) : grid(g), block(b), smem(s), stream(st) {}
};
template <typename... KArgs, typename... Args>
static inline cudaError_t launch(void (*kern)(KArgs...), const LaunchCfg& lc, Args&&... args) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = lc.grid;
  cfg.blockDim = lc.block;
  cfg.dynamicSmemBytes = lc.smem;
  cfg.stream = lc.stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<Args&&>(args)...);
}
#define GSB_PDL_ENTRY()                                          \
  do {                                                           \
    asm volatile("griddepcontrol.launch_dependents;");           \
    asm volatile("griddepcontrol.wait;" ::: "memory");           \
  } while (0)
// dst[i] (+)= sum over s < nslices of ws[s * slice_stride + i], i < n, in slice order with fp64 accumulation (base.cu)
int sum_slices(const float* ws, int nslices, int64_t slice_stride, int64_t n, float* dst, int accumulate, cudaStream_t st);
int num_sms();        // SMs of the current device
int num_sms_avail();  // SMs the persistent tensor-core kernels size their grids for (total minus gsb_set_sm_reserve)

// A view is "pixel-linear" when pixel p = (n*h + y)*w + x lives at ptr + p*sw (+ c).
static inline bool pixel_linear(const gsb_view* v) {
  return v->sc == 1 && v->sh == (int64_t)v->w * v->sw && v->sn == (int64_t)v->h * v->sh;
}
static inline bool vec8_ok(const gsb_view* v) {
  size_t esz = v->dtype == GSB_BF16 ? 2 : 4;
  return pixel_linear(v) && v->c % 8 == 0 && v->sw % 8 == 0 && ((uintptr_t)v->ptr % 16 == 0) && (esz > 0);
}
static inline int64_t num_pixels(const gsb_view* v) { return (int64_t)v->n * v->h * v->w; }

// ---- element access -------------------------------------------------------------------
__device__ __forceinline__ float ld_elem(const void* p, int64_t i, int dtype) {
  return dtype == GSB_BF16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p)[i])
                           : reinterpret_cast<const float*>(p)[i];
}
__device__ __forceinline__ void st_elem(void* p, int64_t i, int dtype, float v) {
  if (dtype == GSB_BF16)
    reinterpret_cast<__nv_bfloat16*>(p)[i] = __float2bfloat16_rn(v);
  else
    reinterpret_cast<float*>(p)[i] = v;
}

// 8 consecutive channels <-> 8 floats (128-bit accesses for bf16, 2x128-bit for fp32).
template <typename T>
struct Vec8;
template <>
struct Vec8<__nv_bfloat16> {
  static __device__ __forceinline__ void load(const __nv_bfloat16* p, float (&v)[8]) {
    uint4 r = *reinterpret_cast<const uint4*>(p);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float2 f = __bfloat1622float2(h[i]);
      v[2 * i] = f.x;
      v[2 * i + 1] = f.y;
    }
  }
  static __device__ __forceinline__ void store(__nv_bfloat16* p, const float (&v)[8]) {
    uint4 r;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
    *reinterpret_cast<uint4*>(p) = r;
  }
};
template <>
struct Vec8<float> {
  static __device__ __forceinline__ void load(const float* p, float (&v)[8]) {
    float4 a = *reinterpret_cast<const float4*>(p);
    float4 b = *reinterpret_cast<const float4*>(p + 4);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
    v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
  static __device__ __forceinline__ void store(float* p, const float (&v)[8]) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
  }
};

// ---- activations ----------------------------------------------------------------------
__device__ __forceinline__ float act_fwd(float z, int kind) {
  switch (kind) {
    case GSB_ACT_RELU: return z > 0.f ? z : 0.f;
    case GSB_ACT_LRELU: return z > 0.f ? z : 0.2f * z;
    case GSB_ACT_TANH: return tanhf(z);
    default: return z;
  }
}
// derivative w.r.t. the pre-activation z (PyTorch conventions: relu'(0)=0, lrelu'(0)=slope)
__device__ __forceinline__ float act_grad(float z, int kind) {
  switch (kind) {
    case GSB_ACT_RELU: return z > 0.f ? 1.f : 0.f;
    case GSB_ACT_LRELU: return z > 0.f ? 1.f : 0.2f;
    case GSB_ACT_TANH: {
      float t = tanhf(z);
      return 1.f - t * t;
    }
    default: return 1.f;
  }
}

// ---- Philox4x32-10 (stateless dropout masks: recomputed in backward) --------------------
__device__ __forceinline__ uint4 philox4x32(uint4 ctr, uint2 key) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
    uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += W0;
    key.y += W1;
  }
  return ctr;
}
// keep-mask bit (p_keep = 0.5) for 8 consecutive logical elements starting at idx (idx % 8 == 0)
__device__ __forceinline__ uint32_t dropout_keep8(uint64_t seed, uint64_t idx) {
  uint4 c = make_uint4((uint32_t)(idx >> 3), (uint32_t)(idx >> 35), 0u, 0u);
  uint4 r = philox4x32(c, make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
  return r.x & 0xFFu;
}

// ---- reductions -----------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

}  // namespace gsb

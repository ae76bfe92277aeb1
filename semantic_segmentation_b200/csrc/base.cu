// Error plumbing, version and device queries of libgenseg_b200.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"

namespace gsb {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static unsigned long long g_launches = 0;
unsigned long long launch_count() { return g_launches; }

int check_launch(const char* what, int n) {
  g_launches += (unsigned long long)n;
  cudaError_t e = cudaPeekAtLastError();
  if (e != cudaSuccess) {
    cudaGetLastError();
    set_error("%s: CUDA error: %s", what, cudaGetErrorString(e));
    return -3;
  }
  return 0;
}

bool pdl_enabled() {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("GSB_PDL");
    on = (e && e[0] == '1') ? 1 : 0;   // off by default: measured neutral (profiles/README.md, r01r)
  }
  return on == 1;
}

int num_sms() {  // of the CURRENT device (cached per device ordinal)
  static int sms[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (sms[dev] == 0) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
    sms[dev] = v;
  }
  return sms[dev];
}

// SMs the persistent kernels size their grids for: all of them, minus the ones set aside for a collective that runs
// concurrently on another stream (gsb_set_sm_reserve; see parallel.py -- the gradient all-reduce overlapped with the
// encoder backward).  A persistent grid that does not fit beside the collective's CTAs would run its tail as a second wave.
static int g_sm_reserve = 0;
int num_sms_avail() {
  const int n = num_sms() - g_sm_reserve;
  return n < 8 ? 8 : n;
}

// dst[i] (+)= sum_s ws[s][i], s ascending, fp64 accumulation: the deterministic second stage of every split reduction of
// the library (weight-gradient slices, per-CTA partial rows of the edge kernels); no atomics anywhere.
__global__ void __launch_bounds__(256) sum_slices_kernel(const float* __restrict__ ws, int nslices, int64_t slice_stride, int64_t n,
                                                        float* __restrict__ dst, int accumulate, int vec) {
  GSB_PDL_ENTRY();
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  if (vec) {
    const int64_t n4 = n >> 2;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
      double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
#pragma unroll 4
      for (int s = 0; s < nslices; ++s) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(ws + (int64_t)s * slice_stride) + i);
        a0 += (double)v.x; a1 += (double)v.y; a2 += (double)v.z; a3 += (double)v.w;
      }
      float4* d = reinterpret_cast<float4*>(dst) + i;
      if (accumulate) {
        const float4 o = *d;
        a0 += (double)o.x; a1 += (double)o.y; a2 += (double)o.z; a3 += (double)o.w;
      }
      *d = make_float4((float)a0, (float)a1, (float)a2, (float)a3);
    }
  } else {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
      double a = 0.0;
#pragma unroll 4
      for (int s = 0; s < nslices; ++s) a += (double)__ldg(ws + (int64_t)s * slice_stride + i);
      if (accumulate) a += (double)dst[i];
      dst[i] = (float)a;
    }
  }
}

// Many slices, few columns (per-CTA partial rows of the edge kernels: 592 rows of 576 floats): one thread per column would
// walk hundreds of dependent loads (130 us for 1.4 MB).  Here a block owns 8 columns x 32 slice lanes: lane l adds slices
// l, l + 32, ... (fp64), the 32 lane sums are added in lane order.  Fixed mapping -> deterministic.
__global__ void __launch_bounds__(256) sum_slices_wide_kernel(const float* __restrict__ ws, int nslices, int64_t slice_stride,
                                                             int64_t n, float* __restrict__ dst, int accumulate) {
  GSB_PDL_ENTRY();
  __shared__ double red[32][9];
  const int c = threadIdx.x & 7, l = threadIdx.x >> 3;
  const int64_t col = (int64_t)blockIdx.x * 8 + c;
  double a = 0.0;
  if (col < n)
    for (int s = l; s < nslices; s += 32) a += (double)__ldg(ws + (int64_t)s * slice_stride + col);
  red[l][c] = a;
  __syncthreads();
  if (l == 0 && col < n) {
    double t = 0.0;
#pragma unroll
    for (int i = 0; i < 32; ++i) t += red[i][c];
    if (accumulate) t += (double)dst[col];
    dst[col] = (float)t;
  }
}

int sum_slices(const float* ws, int nslices, int64_t slice_stride, int64_t n, float* dst, int accumulate, cudaStream_t st) {
  if (nslices >= 16 && n <= ((int64_t)1 << 22)) {
    launch(sum_slices_wide_kernel, LaunchCfg((unsigned)((n + 7) / 8), 256, 0, st), ws, nslices, slice_stride, n, dst, accumulate);
    return 0;
  }
  const int vec = (n % 4 == 0) && (slice_stride % 4 == 0) && ((uintptr_t)ws % 16 == 0) && ((uintptr_t)dst % 16 == 0);
  const int64_t items = vec ? n / 4 : n;
  int64_t blocks = (items + 255) / 256;
  const int64_t cap = (int64_t)num_sms() * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  launch(sum_slices_kernel, LaunchCfg((unsigned)blocks, 256, 0, st), ws, nslices, slice_stride, n, dst, accumulate, vec);
  return 0;
}

}  // namespace gsb

extern "C" int gsb_sum_slices(const float* ws, int nslices, int64_t n, float* dst, int accumulate, gsb_stream s) {
  GSB_REQUIRE(ws && dst && nslices >= 1 && n >= 1, "sum_slices: bad argument");
  gsb::sum_slices(ws, nslices, n, n, dst, accumulate, gsb::to_stream(s));
  return gsb::check_launch("sum_slices");
}

extern "C" unsigned long long gsb_launch_count(void) { return gsb::launch_count(); }
extern "C" int gsb_version(void) { return 100; }
extern "C" const char* gsb_last_error(void) { return gsb::g_err; }
extern "C" int gsb_tc_available(void) {
  int dev = 0, major = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) return 0;
  return major == 10 ? 1 : 0;
}

extern "C" int gsb_set_sm_reserve(int n) {
  const int prev = gsb::g_sm_reserve;
  gsb::g_sm_reserve = n < 0 ? 0 : (n > 64 ? 64 : n);
  return prev;
}

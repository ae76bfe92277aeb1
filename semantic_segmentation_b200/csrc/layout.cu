// Layout conversion and weight packing kernels (all bandwidth-trivial: weights are <= 1.1 GB,
// re-packed every forward because Betty's DARTS step perturbs p.data in place -- SURVEY fact 10).
#include "common.cuh"

namespace gsb {

struct CView {
  void* ptr;
  int dtype;
  int N, H, W, C;
  int64_t sn, sh, sw, sc;
};
static CView cv(const gsb_view* v) { return CView{v->ptr, v->dtype, v->n, v->h, v->w, v->c, v->sn, v->sh, v->sw, v->sc}; }

__global__ void copy_view_kernel(CView src, CView dst, int c_fastest) {
  const int64_t total = (int64_t)dst.N * dst.H * dst.W * dst.C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int n, y, x, c;
    int64_t r = i;
    if (c_fastest) {
      c = (int)(r % dst.C); r /= dst.C;
      x = (int)(r % dst.W); r /= dst.W;
      y = (int)(r % dst.H); n = (int)(r / dst.H);
    } else {
      x = (int)(r % dst.W); r /= dst.W;
      y = (int)(r % dst.H); r /= dst.H;
      c = (int)(r % dst.C); n = (int)(r / dst.C);
    }
    float v = ld_elem(src.ptr, (int64_t)n * src.sn + (int64_t)y * src.sh + (int64_t)x * src.sw + (int64_t)c * src.sc, src.dtype);
    st_elem(dst.ptr, (int64_t)n * dst.sn + (int64_t)y * dst.sh + (int64_t)x * dst.sw + (int64_t)c * dst.sc, dst.dtype, v);
  }
}

__global__ void pack_weight_kernel(const float* __restrict__ src, int d0, int d1, int taps, int rows_from_d0, void* dst,
                                   int dst_dtype) {
  const int rows = rows_from_d0 ? d0 : d1, cols = rows_from_d0 ? d1 : d0;
  const int64_t total = (int64_t)taps * rows * cols;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int c = (int)(i % cols);
    int64_t r2 = i / cols;
    int r = (int)(r2 % rows);
    int tap = (int)(r2 / rows);
    int i0 = rows_from_d0 ? r : c, i1 = rows_from_d0 ? c : r;
    st_elem(dst, i, dst_dtype, src[((int64_t)i0 * d1 + i1) * taps + tap]);
  }
}

__global__ void unpack_wgrad_kernel(const float* __restrict__ src, int d0, int d1, int taps, int rows_from_d0,
                                    float* __restrict__ dst, int accumulate) {
  const int rows = rows_from_d0 ? d0 : d1, cols = rows_from_d0 ? d1 : d0;
  const int64_t total = (int64_t)taps * d0 * d1;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int tap = (int)(i % taps);
    int64_t r2 = i / taps;
    int i1 = (int)(r2 % d1);
    int i0 = (int)(r2 / d1);
    int r = rows_from_d0 ? i0 : i1, c = rows_from_d0 ? i1 : i0;
    float v = src[((int64_t)tap * rows + r) * cols + c];
    dst[i] = accumulate ? dst[i] + v : v;
  }
}

__device__ __forceinline__ void softmax3(const float* alpha, float (&w)[3]) {
  float m = fmaxf(alpha[0], fmaxf(alpha[1], alpha[2]));
  float e0 = expf(alpha[0] - m), e1 = expf(alpha[1] - m), e2 = expf(alpha[2] - m);
  float inv = 1.f / (e0 + e1 + e2);
  w[0] = e0 * inv; w[1] = e1 * inv; w[2] = e2 * inv;
}

__global__ void merge_upconv_kernel(const float* __restrict__ w4, const float* __restrict__ w6, const float* __restrict__ w8,
                                    const float* b4, const float* b6, const float* b8, const float* __restrict__ alpha,
                                    int64_t pairs, int cout, float* __restrict__ wm, float* bm) {
  float w[3];
  softmax3(alpha, w);
  const int64_t total = pairs * 64;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int kx = (int)(i & 7), ky = (int)((i >> 3) & 7);
    int64_t pr = i >> 6;
    float v = w[2] * w8[i];
    if (ky >= 1 && ky < 7 && kx >= 1 && kx < 7) v += w[1] * w6[pr * 36 + (ky - 1) * 6 + (kx - 1)];
    if (ky >= 2 && ky < 6 && kx >= 2 && kx < 6) v += w[0] * w4[pr * 16 + (ky - 2) * 4 + (kx - 2)];
    wm[i] = v;
  }
  if (bm != nullptr && blockIdx.x == 0)
    for (int c = threadIdx.x; c < cout; c += blockDim.x) bm[c] = w[0] * b4[c] + w[1] * b6[c] + w[2] * b8[c];
}

// dots[3] += <crop_i(dWm), W_i> (+ <db, b_i>);  dW_i = w_i * crop_i(dWm)
__global__ void unmerge_upconv_kernel(const float* __restrict__ dwm, const float* dbm, const float* __restrict__ w4,
                                      const float* __restrict__ w6, const float* __restrict__ w8, const float* b4,
                                      const float* b6, const float* b8, const float* __restrict__ alpha, int64_t pairs,
                                      int cout, float* __restrict__ dw4, float* __restrict__ dw6, float* __restrict__ dw8,
                                      float* db4, float* db6, float* db8, double* dots) {
  float w[3];
  softmax3(alpha, w);
  const int64_t total = pairs * 64;
  float d0 = 0.f, d1 = 0.f, d2 = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int kx = (int)(i & 7), ky = (int)((i >> 3) & 7);
    int64_t pr = i >> 6;
    float g = dwm[i];
    dw8[i] = w[2] * g;
    d2 = fmaf(g, w8[i], d2);
    if (ky >= 1 && ky < 7 && kx >= 1 && kx < 7) {
      int64_t j = pr * 36 + (ky - 1) * 6 + (kx - 1);
      dw6[j] = w[1] * g;
      d1 = fmaf(g, w6[j], d1);
    }
    if (ky >= 2 && ky < 6 && kx >= 2 && kx < 6) {
      int64_t j = pr * 16 + (ky - 2) * 4 + (kx - 2);
      dw4[j] = w[0] * g;
      d0 = fmaf(g, w4[j], d0);
    }
  }
  if (dbm != nullptr && blockIdx.x == 0) {
    for (int c = threadIdx.x; c < cout; c += blockDim.x) {
      float g = dbm[c];
      db4[c] = w[0] * g; db6[c] = w[1] * g; db8[c] = w[2] * g;
      d0 = fmaf(g, b4[c], d0); d1 = fmaf(g, b6[c], d1); d2 = fmaf(g, b8[c], d2);
    }
  }
  d0 = warp_sum(d0); d1 = warp_sum(d1); d2 = warp_sum(d2);
  if ((threadIdx.x & 31) == 0) {
    atomicAdd(dots + 0, (double)d0);
    atomicAdd(dots + 1, (double)d1);
    atomicAdd(dots + 2, (double)d2);
  }
}

__global__ void dalpha_kernel(const double* __restrict__ dots, const float* __restrict__ alpha, float* dalpha) {
  float w[3];
  softmax3(alpha, w);
  double mean = w[0] * dots[0] + w[1] * dots[1] + w[2] * dots[2];
  for (int j = 0; j < 3; ++j) dalpha[j] += (float)(w[j] * (dots[j] - mean));
}

static int ew_grid(int64_t total) {
  int64_t b = (total + 255) / 256;
  int64_t cap = (int64_t)num_sms() * 16;
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

}  // namespace gsb

using namespace gsb;

extern "C" int gsb_copy_view(const gsb_view* src, const gsb_view* dst, gsb_stream s) {
  GSB_REQUIRE(src && dst && src->ptr && dst->ptr, "copy_view: null view");
  GSB_REQUIRE(src->n == dst->n && src->h == dst->h && src->w == dst->w && src->c == dst->c, "copy_view: shape mismatch");
  int64_t total = (int64_t)dst->n * dst->h * dst->w * dst->c;
  if (total == 0) return 0;
  copy_view_kernel<<<ew_grid(total), 256, 0, to_stream(s)>>>(cv(src), cv(dst), dst->sc == 1 ? 1 : 0);
  return check_launch("copy_view");
}

extern "C" int gsb_pack_weight(const float* src, int d0, int d1, int kh, int kw, int rows_from_d0, void* dst, int dst_dtype,
                               gsb_stream s) {
  GSB_REQUIRE(src && dst, "pack_weight: null pointer");
  int64_t total = (int64_t)d0 * d1 * kh * kw;
  pack_weight_kernel<<<ew_grid(total), 256, 0, to_stream(s)>>>(src, d0, d1, kh * kw, rows_from_d0, dst, dst_dtype);
  return check_launch("pack_weight");
}

extern "C" int gsb_unpack_wgrad(const float* src, int d0, int d1, int kh, int kw, int rows_from_d0, float* dst,
                                int accumulate, gsb_stream s) {
  GSB_REQUIRE(src && dst, "unpack_wgrad: null pointer");
  int64_t total = (int64_t)d0 * d1 * kh * kw;
  unpack_wgrad_kernel<<<ew_grid(total), 256, 0, to_stream(s)>>>(src, d0, d1, kh * kw, rows_from_d0, dst, accumulate);
  return check_launch("unpack_wgrad");
}

extern "C" int gsb_merge_upconv(const float* w4, const float* w6, const float* w8, const float* b4, const float* b6,
                                const float* b8, const float* alpha, int cin, int cout, float* w_merged, float* b_merged,
                                gsb_stream s) {
  GSB_REQUIRE(w4 && w6 && w8 && alpha && w_merged, "merge_upconv: null pointer");
  GSB_REQUIRE(b_merged == nullptr || (b4 && b6 && b8), "merge_upconv: biases missing");
  int64_t pairs = (int64_t)cin * cout;
  merge_upconv_kernel<<<ew_grid(pairs * 64), 256, 0, to_stream(s)>>>(w4, w6, w8, b4, b6, b8, alpha, pairs, cout, w_merged,
                                                                    b_merged);
  return check_launch("merge_upconv");
}

extern "C" int gsb_unmerge_upconv(const float* dw_merged, const float* db_merged, const float* w4, const float* w6,
                                  const float* w8, const float* b4, const float* b6, const float* b8, const float* alpha,
                                  int cin, int cout, float* dw4, float* dw6, float* dw8, float* db4, float* db6, float* db8,
                                  float* dalpha, void* workspace, size_t workspace_bytes, gsb_stream s) {
  GSB_REQUIRE(dw_merged && w4 && w6 && w8 && alpha && dw4 && dw6 && dw8 && dalpha, "unmerge_upconv: null pointer");
  GSB_REQUIRE(workspace && workspace_bytes >= 3 * sizeof(double), "unmerge_upconv: workspace too small");
  GSB_REQUIRE(db_merged == nullptr || (b4 && b6 && b8 && db4 && db6 && db8), "unmerge_upconv: biases missing");
  double* dots = reinterpret_cast<double*>(workspace);
  cudaMemsetAsync(dots, 0, 3 * sizeof(double), to_stream(s));
  int64_t pairs = (int64_t)cin * cout;
  unmerge_upconv_kernel<<<ew_grid(pairs * 64), 256, 0, to_stream(s)>>>(dw_merged, db_merged, w4, w6, w8, b4, b6, b8, alpha,
                                                                      pairs, cout, dw4, dw6, dw8, db4, db6, db8, dots);
  dalpha_kernel<<<1, 1, 0, to_stream(s)>>>(dots, alpha, dalpha);
  return check_launch("unmerge_upconv", 2);
}

// ---------------------------------------------------------------------------------------------------------
// Batched ("multi-tensor") weight pack / weight-gradient unpack: one launch for every layer of a network instead of
// one 5-8 us latency-bound launch per layer (34 + 17 launches per U-Net step; visible at 8 images per GPU).
// ---------------------------------------------------------------------------------------------------------
namespace gsb {
constexpr int kMaxLayoutJobs = 64;
struct LayoutJobs {
  const void* src[kMaxLayoutJobs];
  void* dst[kMaxLayoutJobs];
  int d0[kMaxLayoutJobs], d1[kMaxLayoutJobs], taps[kMaxLayoutJobs];
  int blk0[kMaxLayoutJobs + 1];
  uint8_t rows_from_d0[kMaxLayoutJobs];
  int count, dst_dtype, accumulate;
};
constexpr int kLayoutPerBlock = 256 * 8;

__device__ __forceinline__ int find_job(const int* blk0, int count, int b) {
  int lo = 0, hi = count - 1;
  while (lo < hi) {
    int mid = (lo + hi + 1) >> 1;
    if (blk0[mid] <= b) lo = mid; else hi = mid - 1;
  }
  return lo;
}

__global__ void __launch_bounds__(256) pack_weight_multi_kernel(const __grid_constant__ LayoutJobs J) {
  const int j = find_job(J.blk0, J.count, blockIdx.x);
  const int d0 = J.d0[j], d1 = J.d1[j], taps = J.taps[j], rfd = J.rows_from_d0[j];
  const int rows = rfd ? d0 : d1, cols = rfd ? d1 : d0;
  const int64_t total = (int64_t)taps * rows * cols;
  const float* src = reinterpret_cast<const float*>(J.src[j]);
  const int64_t base = (int64_t)(blockIdx.x - J.blk0[j]) * kLayoutPerBlock;
#pragma unroll
  for (int u = 0; u < 8; ++u) {
    const int64_t i = base + u * 256 + threadIdx.x;
    if (i < total) {
      int c = (int)(i % cols);
      int64_t r2 = i / cols;
      int r = (int)(r2 % rows);
      int tap = (int)(r2 / rows);
      int i0 = rfd ? r : c, i1 = rfd ? c : r;
      st_elem(J.dst[j], i, J.dst_dtype, src[((int64_t)i0 * d1 + i1) * taps + tap]);
    }
  }
}

__global__ void __launch_bounds__(256) unpack_wgrad_multi_kernel(const __grid_constant__ LayoutJobs J) {
  const int j = find_job(J.blk0, J.count, blockIdx.x);
  const int d0 = J.d0[j], d1 = J.d1[j], taps = J.taps[j], rfd = J.rows_from_d0[j];
  const int rows = rfd ? d0 : d1, cols = rfd ? d1 : d0;
  const int64_t total = (int64_t)taps * d0 * d1;
  const float* src = reinterpret_cast<const float*>(J.src[j]);
  float* dst = reinterpret_cast<float*>(J.dst[j]);
  const int64_t base = (int64_t)(blockIdx.x - J.blk0[j]) * kLayoutPerBlock;
#pragma unroll
  for (int u = 0; u < 8; ++u) {
    const int64_t i = base + u * 256 + threadIdx.x;
    if (i < total) {
      int tap = (int)(i % taps);
      int64_t r2 = i / taps;
      int i1 = (int)(r2 % d1);
      int i0 = (int)(r2 / d1);
      int r = rfd ? i0 : i1, c = rfd ? i1 : i0;
      float v = src[((int64_t)tap * rows + r) * cols + c];
      dst[i] = J.accumulate ? dst[i] + v : v;
    }
  }
}

static int run_layout_multi(bool pack, int n, const void* const* src, void* const* dst, const int* d0, const int* d1,
                            const int* taps, const int* rows_from_d0, int dst_dtype, int accumulate, cudaStream_t st) {
  int launches = 0;
  for (int begin = 0; begin < n; begin += kMaxLayoutJobs) {
    LayoutJobs J;
    memset(&J, 0, sizeof(J));
    J.count = n - begin < kMaxLayoutJobs ? n - begin : kMaxLayoutJobs;
    J.dst_dtype = dst_dtype;
    J.accumulate = accumulate;
    int blocks = 0;
    for (int k = 0; k < J.count; ++k) {
      const int i = begin + k;
      GSB_REQUIRE(src[i] && dst[i] && d0[i] > 0 && d1[i] > 0 && taps[i] > 0, "layout_multi: bad job %d", i);
      J.src[k] = src[i]; J.dst[k] = dst[i];
      J.d0[k] = d0[i]; J.d1[k] = d1[i]; J.taps[k] = taps[i];
      J.rows_from_d0[k] = (uint8_t)(rows_from_d0[i] != 0);
      J.blk0[k] = blocks;
      const int64_t total = (int64_t)d0[i] * d1[i] * taps[i];
      blocks += (int)((total + kLayoutPerBlock - 1) / kLayoutPerBlock);
    }
    J.blk0[J.count] = blocks;
    if (blocks == 0) continue;
    if (pack) pack_weight_multi_kernel<<<blocks, 256, 0, st>>>(J);
    else unpack_wgrad_multi_kernel<<<blocks, 256, 0, st>>>(J);
    ++launches;
  }
  return check_launch(pack ? "pack_weight_multi" : "unpack_wgrad_multi", launches > 0 ? launches : 1);
}
}  // namespace gsb

extern "C" int gsb_pack_weight_multi(int n, const void* const* src, void* const* dst, const int* d0, const int* d1,
                                     const int* taps, const int* rows_from_d0, int dst_dtype, gsb_stream s) {
  if (n <= 0) return 0;
  GSB_REQUIRE(src && dst && d0 && d1 && taps && rows_from_d0, "pack_weight_multi: null argument");
  return gsb::run_layout_multi(true, n, src, dst, d0, d1, taps, rows_from_d0, dst_dtype, 0, gsb::to_stream(s));
}

extern "C" int gsb_unpack_wgrad_multi(int n, const void* const* src, void* const* dst, const int* d0, const int* d1,
                                      const int* taps, const int* rows_from_d0, int accumulate, gsb_stream s) {
  if (n <= 0) return 0;
  GSB_REQUIRE(src && dst && d0 && d1 && taps && rows_from_d0, "unpack_wgrad_multi: null argument");
  return gsb::run_layout_multi(false, n, src, dst, d0, d1, taps, rows_from_d0, GSB_F32, accumulate, gsb::to_stream(s));
}

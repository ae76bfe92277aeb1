// Layout conversion and weight packing kernels (all bandwidth-trivial: weights are <= 1.1 GB,
// re-packed every forward because Betty's DARTS step perturbs p.data in place -- SURVEY fact 10).
#include "common.cuh"

namespace gsb {

static int run_layout_multi(bool pack, int n, const void* const* src, void* const* dst, const int* d0, const int* d1,
                            const int* taps, const int* rows_from_d0, const int* nslices, int dst_dtype, int accumulate,
                            cudaStream_t st);


struct CView {
  void* ptr;
  int dtype;
  int N, H, W, C;
  int64_t sn, sh, sw, sc;
};
static CView cv(const gsb_view* v) { return CView{v->ptr, v->dtype, v->n, v->h, v->w, v->c, v->sn, v->sh, v->sw, v->sc}; }

__global__ void copy_view_kernel(CView src, CView dst, int c_fastest) {
  GSB_PDL_ENTRY();
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
  GSB_PDL_ENTRY();
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
  GSB_PDL_ENTRY();
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
  GSB_PDL_ENTRY();
  float w[3];
  softmax3(alpha, w);
  const int64_t rows_total = pairs * 8;   // one thread = one kernel row (8 taps) of one (ci, co) pair
  for (int64_t rix = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; rix < rows_total; rix += (int64_t)gridDim.x * blockDim.x) {
    const int ky = (int)(rix & 7);
    const int64_t pr = rix >> 3, i = rix * 8;
    const float4 wa = *reinterpret_cast<const float4*>(w8 + i), wb = *reinterpret_cast<const float4*>(w8 + i + 4);
    float v[8] = {w[2] * wa.x, w[2] * wa.y, w[2] * wa.z, w[2] * wa.w, w[2] * wb.x, w[2] * wb.y, w[2] * wb.z, w[2] * wb.w};
    if (ky >= 1 && ky < 7) {
      const int64_t j = pr * 36 + (ky - 1) * 6;
      const float2 u0 = *reinterpret_cast<const float2*>(w6 + j), u1 = *reinterpret_cast<const float2*>(w6 + j + 2),
                   u2 = *reinterpret_cast<const float2*>(w6 + j + 4);
      v[1] += w[1] * u0.x; v[2] += w[1] * u0.y; v[3] += w[1] * u1.x; v[4] += w[1] * u1.y; v[5] += w[1] * u2.x; v[6] += w[1] * u2.y;
    }
    if (ky >= 2 && ky < 6) {
      const float4 u = *reinterpret_cast<const float4*>(w4 + pr * 16 + (ky - 2) * 4);
      v[2] += w[0] * u.x; v[3] += w[0] * u.y; v[4] += w[0] * u.z; v[5] += w[0] * u.w;
    }
    *reinterpret_cast<float4*>(wm + i) = make_float4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<float4*>(wm + i + 4) = make_float4(v[4], v[5], v[6], v[7]);
  }
  if (bm != nullptr && blockIdx.x == 0)
    for (int c = threadIdx.x; c < cout; c += blockDim.x) bm[c] = w[0] * b4[c] + w[1] * b6[c] + w[2] * b8[c];
}

// dots[blockIdx.x][3] = this CTA's part of <crop_i(dWm), W_i> (+ <db, b_i>);  dW_i = w_i * crop_i(dWm).
// The per-CTA rows are added in row order by dalpha_kernel (deterministic, no atomics).
__global__ void unmerge_upconv_kernel(const float* __restrict__ dwm, const float* dbm, const float* __restrict__ w4,
                                      const float* __restrict__ w6, const float* __restrict__ w8, const float* b4,
                                      const float* b6, const float* b8, const float* __restrict__ alpha, int64_t pairs,
                                      int cout, float* __restrict__ dw4, float* __restrict__ dw6, float* __restrict__ dw8,
                                      float* db4, float* db6, float* db8, double* dots) {
  GSB_PDL_ENTRY();
  float w[3];
  softmax3(alpha, w);
  // one thread = one kernel row (8 taps) of one (ci, co) pair: 2 x 16-byte loads per stream instead of one element per
  // iteration (the element-wise form sat at ~1.5 TB/s with a single load in flight per thread)
  const int64_t rows_total = pairs * 8;
  float d0 = 0.f, d1 = 0.f, d2 = 0.f;
  for (int64_t rix = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; rix < rows_total; rix += (int64_t)gridDim.x * blockDim.x) {
    const int ky = (int)(rix & 7);
    const int64_t pr = rix >> 3, i = rix * 8;
    const float4 ga = *reinterpret_cast<const float4*>(dwm + i), gb = *reinterpret_cast<const float4*>(dwm + i + 4);
    const float4 wa = *reinterpret_cast<const float4*>(w8 + i), wb = *reinterpret_cast<const float4*>(w8 + i + 4);
    const float g[8] = {ga.x, ga.y, ga.z, ga.w, gb.x, gb.y, gb.z, gb.w};
    *reinterpret_cast<float4*>(dw8 + i) = make_float4(w[2] * ga.x, w[2] * ga.y, w[2] * ga.z, w[2] * ga.w);
    *reinterpret_cast<float4*>(dw8 + i + 4) = make_float4(w[2] * gb.x, w[2] * gb.y, w[2] * gb.z, w[2] * gb.w);
    d2 = fmaf(ga.x, wa.x, d2); d2 = fmaf(ga.y, wa.y, d2); d2 = fmaf(ga.z, wa.z, d2); d2 = fmaf(ga.w, wa.w, d2);
    d2 = fmaf(gb.x, wb.x, d2); d2 = fmaf(gb.y, wb.y, d2); d2 = fmaf(gb.z, wb.z, d2); d2 = fmaf(gb.w, wb.w, d2);
    if (ky >= 1 && ky < 7) {
      const int64_t j = pr * 36 + (ky - 1) * 6;        // 6 taps kx = 1..6, 8-byte aligned
      const float2 u0 = *reinterpret_cast<const float2*>(w6 + j), u1 = *reinterpret_cast<const float2*>(w6 + j + 2),
                   u2 = *reinterpret_cast<const float2*>(w6 + j + 4);
      *reinterpret_cast<float2*>(dw6 + j) = make_float2(w[1] * g[1], w[1] * g[2]);
      *reinterpret_cast<float2*>(dw6 + j + 2) = make_float2(w[1] * g[3], w[1] * g[4]);
      *reinterpret_cast<float2*>(dw6 + j + 4) = make_float2(w[1] * g[5], w[1] * g[6]);
      d1 = fmaf(g[1], u0.x, d1); d1 = fmaf(g[2], u0.y, d1); d1 = fmaf(g[3], u1.x, d1);
      d1 = fmaf(g[4], u1.y, d1); d1 = fmaf(g[5], u2.x, d1); d1 = fmaf(g[6], u2.y, d1);
    }
    if (ky >= 2 && ky < 6) {
      const int64_t j = pr * 16 + (ky - 2) * 4;        // 4 taps kx = 2..5, 16-byte aligned
      const float4 u = *reinterpret_cast<const float4*>(w4 + j);
      *reinterpret_cast<float4*>(dw4 + j) = make_float4(w[0] * g[2], w[0] * g[3], w[0] * g[4], w[0] * g[5]);
      d0 = fmaf(g[2], u.x, d0); d0 = fmaf(g[3], u.y, d0); d0 = fmaf(g[4], u.z, d0); d0 = fmaf(g[5], u.w, d0);
    }
  }
  if (dbm != nullptr && blockIdx.x == 0) {
    for (int c = threadIdx.x; c < cout; c += blockDim.x) {
      float g = dbm[c];
      db4[c] = w[0] * g; db6[c] = w[1] * g; db8[c] = w[2] * g;
      d0 = fmaf(g, b4[c], d0); d1 = fmaf(g, b6[c], d1); d2 = fmaf(g, b8[c], d2);
    }
  }
  __shared__ double sred[8][3];
  d0 = warp_sum(d0); d1 = warp_sum(d1); d2 = warp_sum(d2);
  if ((threadIdx.x & 31) == 0) {
    sred[threadIdx.x >> 5][0] = (double)d0;
    sred[threadIdx.x >> 5][1] = (double)d1;
    sred[threadIdx.x >> 5][2] = (double)d2;
  }
  __syncthreads();
  if (threadIdx.x < 3) {
    double t = 0.0;
    for (int w_ = 0; w_ < (int)(blockDim.x >> 5); ++w_) t += sred[w_][threadIdx.x];
    dots[(int64_t)blockIdx.x * 3 + threadIdx.x] = t;
  }
}

__global__ void __launch_bounds__(256) dalpha_kernel(const double* __restrict__ dots, int rows, const float* __restrict__ alpha,
                                                    float* dalpha) {
  GSB_PDL_ENTRY();
  __shared__ double sh[8][3];
  __shared__ double tot[3];
  double a[3] = {0.0, 0.0, 0.0};
  for (int r = threadIdx.x; r < rows; r += blockDim.x) {
    a[0] += dots[(int64_t)r * 3 + 0]; a[1] += dots[(int64_t)r * 3 + 1]; a[2] += dots[(int64_t)r * 3 + 2];
  }
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    double v = a[j];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5][j] = v;
  }
  __syncthreads();
  if (threadIdx.x < 3) {
    double t = 0.0;
    for (int w_ = 0; w_ < 8; ++w_) t += sh[w_][threadIdx.x];
    tot[threadIdx.x] = t;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float w[3];
    softmax3(alpha, w);
    double mean = w[0] * tot[0] + w[1] * tot[1] + w[2] * tot[2];
    for (int j = 0; j < 3; ++j) dalpha[j] += (float)(w[j] * (tot[j] - mean));
  }
}

static int ew_grid(int64_t total) {
  int64_t b = (total + 255) / 256;
  int64_t cap = (int64_t)num_sms() * 16;
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

}  // namespace gsb

using namespace gsb;

// fp32 <-> bf16 casts of flat buffers (gradient buckets of the data-parallel all-reduce travel as bf16), 8 elements per thread
__global__ void __launch_bounds__(256) cast_f32_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int64_t n) {
  GSB_PDL_ENTRY();
  const int64_t n8 = n >> 3, stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += stride) {
    float v[8];
    Vec8<float>::load(src + i * 8, v);
    Vec8<__nv_bfloat16>::store(dst + i * 8, v);
  }
  for (int64_t i = n8 * 8 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) dst[i] = __float2bfloat16_rn(src[i]);
}
__global__ void __launch_bounds__(256) cast_bf16_f32_kernel(const __nv_bfloat16* __restrict__ src, float* __restrict__ dst, int64_t n) {
  GSB_PDL_ENTRY();
  const int64_t n8 = n >> 3, stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += stride) {
    float v[8];
    Vec8<__nv_bfloat16>::load(src + i * 8, v);
    Vec8<float>::store(dst + i * 8, v);
  }
  for (int64_t i = n8 * 8 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) dst[i] = __bfloat162float(src[i]);
}

extern "C" int gsb_cast_flat(const void* src, int src_dtype, void* dst, int dst_dtype, int64_t n, gsb_stream s) {
  GSB_REQUIRE(src && dst && n >= 0, "cast_flat: bad argument");
  GSB_REQUIRE(((uintptr_t)src % 16 == 0) && ((uintptr_t)dst % 16 == 0), "cast_flat: buffers must be 16-byte aligned");
  if (n == 0) return 0;
  const int grid = ew_grid((n + 7) / 8);
  if (src_dtype == GSB_F32 && dst_dtype == GSB_BF16)
    gsb::launch(cast_f32_bf16_kernel, gsb::LaunchCfg(grid, 256, 0, to_stream(s)), (const float*)src, (__nv_bfloat16*)dst, n);
  else if (src_dtype == GSB_BF16 && dst_dtype == GSB_F32)
    gsb::launch(cast_bf16_f32_kernel, gsb::LaunchCfg(grid, 256, 0, to_stream(s)), (const __nv_bfloat16*)src, (float*)dst, n);
  else
    GSB_REQUIRE(false, "cast_flat: unsupported dtype pair %d -> %d", src_dtype, dst_dtype);
  return check_launch("cast_flat");
}

extern "C" int gsb_copy_view(const gsb_view* src, const gsb_view* dst, gsb_stream s) {
  GSB_REQUIRE(src && dst && src->ptr && dst->ptr, "copy_view: null view");
  GSB_REQUIRE(src->n == dst->n && src->h == dst->h && src->w == dst->w && src->c == dst->c, "copy_view: shape mismatch");
  int64_t total = (int64_t)dst->n * dst->h * dst->w * dst->c;
  if (total == 0) return 0;
  gsb::launch(copy_view_kernel, gsb::LaunchCfg(ew_grid(total), 256, 0, to_stream(s)), cv(src), cv(dst), dst->sc == 1 ? 1 : 0);
  return check_launch("copy_view");
}

extern "C" int gsb_pack_weight(const float* src, int d0, int d1, int kh, int kw, int rows_from_d0, void* dst, int dst_dtype,
                               gsb_stream s) {
  GSB_REQUIRE(src && dst, "pack_weight: null pointer");
  const void* srcs[1] = {src};
  void* dsts[1] = {dst};
  const int taps = kh * kw;
  return gsb::run_layout_multi(true, 1, srcs, dsts, &d0, &d1, &taps, &rows_from_d0, nullptr, dst_dtype, 0, to_stream(s));
}

extern "C" int gsb_unpack_wgrad(const float* src, int nslices, int d0, int d1, int kh, int kw, int rows_from_d0, float* dst,
                                int accumulate, gsb_stream s) {
  GSB_REQUIRE(src && dst && nslices >= 1, "unpack_wgrad: bad argument");
  const void* srcs[1] = {src};
  void* dsts[1] = {dst};
  const int taps = kh * kw;
  return gsb::run_layout_multi(false, 1, srcs, dsts, &d0, &d1, &taps, &rows_from_d0, &nslices, GSB_F32, accumulate, to_stream(s));
}

extern "C" int gsb_merge_upconv(const float* w4, const float* w6, const float* w8, const float* b4, const float* b6,
                                const float* b8, const float* alpha, int cin, int cout, float* w_merged, float* b_merged,
                                gsb_stream s) {
  GSB_REQUIRE(w4 && w6 && w8 && alpha && w_merged, "merge_upconv: null pointer");
  GSB_REQUIRE(b_merged == nullptr || (b4 && b6 && b8), "merge_upconv: biases missing");
  int64_t pairs = (int64_t)cin * cout;
  gsb::launch(merge_upconv_kernel, gsb::LaunchCfg(ew_grid(pairs * 8), 256, 0, to_stream(s)), w4, w6, w8, b4, b6, b8, alpha, pairs, cout, w_merged,
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
  double* dots = reinterpret_cast<double*>(workspace);   // one row of 3 partial dot products per CTA
  int64_t pairs = (int64_t)cin * cout;
  int grid = ew_grid(pairs * 8);
  if ((size_t)grid * 3 * sizeof(double) > workspace_bytes) grid = (int)(workspace_bytes / (3 * sizeof(double)));
  gsb::launch(unmerge_upconv_kernel, gsb::LaunchCfg(grid, 256, 0, to_stream(s)), dw_merged, db_merged, w4, w6, w8, b4, b6, b8, alpha,
                                                                      pairs, cout, dw4, dw6, dw8, db4, db6, db8, dots);
  gsb::launch(dalpha_kernel, gsb::LaunchCfg(1, 256, 0, to_stream(s)), (const double*)dots, grid, alpha, dalpha);
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
  int nslices[kMaxLayoutJobs];   // un-pack: the packed source holds this many partial-sum slices, added in slice order
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

// Tiled transposes between the torch layout  T[i0][i1][tap]  (OIHW / IOHW, fp32) and the packed layout
// P[tap][row][col]  (row = i0, col = i1 when rows_from_d0, else swapped).  The element-wise kernels read 4 bytes per
// 32-byte sector on the strided side (8x read amplification: 1.9 ms per Pix2Pix step for 272 M parameters, ncu
// profiles/r01o); here a block stages a 16 (i0) x 32 (i1) x TT (taps) tile in shared memory, touching both sides in
// contiguous runs.  Row pitches 17 / 545 floats keep both transposed read orders bank-conflict free.
constexpr int kTI0 = 16, kTI1 = 32, kTTmax = 16;
constexpr int kTPitch1 = kTTmax + 1, kTPitch0 = kTI1 * kTPitch1 + 1;

static int layout_tiles_host(int d0, int d1, int taps) {
  const int tt = taps < kTTmax ? taps : kTTmax;
  return ((d0 + kTI0 - 1) / kTI0) * ((d1 + kTI1 - 1) / kTI1) * ((taps + tt - 1) / tt);
}

// TT > 0: taps per tile known at compile time (1, 4, 9, 16 cover every layer of the two networks), so the index
// splits below are shifts and constant divisions; TT = 0: generic.
template <bool PACK, int TT>
__device__ __forceinline__ void layout_tile(const float* __restrict__ tsrc, float* __restrict__ tdst, const float* __restrict__ psrc,
                                            void* __restrict__ pdst, int pdtype, int d0, int d1, int taps, int rfd, int tile,
                                            int accumulate, int nslices, float* sm) {
  const int tt = TT > 0 ? TT : (taps < kTTmax ? taps : kTTmax);
  const int n1 = (d1 + kTI1 - 1) / kTI1, nt = (taps + tt - 1) / tt;
  const int tc = tile % nt, b1 = (tile / nt) % n1, b0 = tile / (nt * n1);
  const int i0b = b0 * kTI0, i1b = b1 * kTI1, tb = tc * tt;
  const int rows = rfd ? d0 : d1, cols = rfd ? d1 : d0;
  const int nA = kTI0 * kTI1 * tt;
  const int64_t slice_stride = (int64_t)taps * rows * cols;
  auto torch_side = [&](bool load) {
    for (int idx = threadIdx.x; idx < nA; idx += blockDim.x) {
      const int t = idx % tt, i1l = (idx / tt) % kTI1, i0l = idx / (tt * kTI1);
      const int i0 = i0b + i0l, i1 = i1b + i1l, tap = tb + t;
      if (i0 < d0 && i1 < d1 && tap < taps) {
        const int64_t off = ((int64_t)i0 * d1 + i1) * taps + tap;
        float* cell = sm + i0l * kTPitch0 + i1l * kTPitch1 + t;
        if (load) *cell = tsrc[off];
        else tdst[off] = accumulate ? tdst[off] + *cell : *cell;
      }
    }
  };
  auto packed_side = [&](bool load) {
    for (int idx = threadIdx.x; idx < nA; idx += blockDim.x) {
      int t, i0l, i1l;
      if (rfd) { i1l = idx % kTI1; i0l = (idx / kTI1) % kTI0; t = idx / (kTI1 * kTI0); }
      else     { i0l = idx % kTI0; i1l = (idx / kTI0) % kTI1; t = idx / (kTI0 * kTI1); }
      const int i0 = i0b + i0l, i1 = i1b + i1l, tap = tb + t;
      if (i0 < d0 && i1 < d1 && tap < taps) {
        const int r = rfd ? i0 : i1, c = rfd ? i1 : i0;
        const int64_t off = ((int64_t)tap * rows + r) * cols + c;
        float* cell = sm + i0l * kTPitch0 + i1l * kTPitch1 + t;
        if (load) {
          double acc = (double)psrc[off];   // slice order, fp64: partial sums of a weight gradient nearly cancel
          for (int sl = 1; sl < nslices; ++sl) acc += (double)psrc[(int64_t)sl * slice_stride + off];
          *cell = (float)acc;
        } else st_elem(pdst, off, pdtype, *cell);
      }
    }
  };
  // ---- vector paths (full tiles): 16-byte global accesses on both sides.  The scalar loops above issue one 4-byte load and
  // one 2-byte store per element plus ~30 index instructions: instruction-issue bound (pack of the U-Net's 17 M weights into
  // both layouts 143 us against ~25 us of HBM time, profiles/r01r).
  const bool full = (i0b + kTI0 <= d0) && (i1b + kTI1 <= d1) && (tb + tt <= taps);
  const float* tptr = PACK ? tsrc : tdst;
  const void* pptr = PACK ? (const void*)pdst : (const void*)psrc;
  // a group of 4 consecutive torch-layout floats starting at a multiple of 4 inside the tile is contiguous and 16-byte
  // aligned when the tile spans all taps (rows of 32 * taps floats) or when the tap chunk itself is a multiple of 4
  const bool vec_t = full && (tt == taps ? (((int64_t)d1 * taps) % 4 == 0) : (tt % 4 == 0 && taps % 4 == 0)) &&
                     ((reinterpret_cast<uintptr_t>(tptr) & 15) == 0);
  const bool vec_p = full && (cols % 8 == 0) && ((reinterpret_cast<uintptr_t>(pptr) & 15) == 0) && (slice_stride % 4 == 0);
  auto torch_side_vec = [&](bool load) {
    const int per_row = kTI1 * tt;                       // floats per i0 row of the tile
    if (TT > 0 && load) {
      // all global loads of the thread first (up to 8 x 16 bytes in flight), then the shared-memory scatter
      constexpr int NG = TT > 0 ? (kTI0 * kTI1 * TT / 4 + 255) / 256 : 1;
      float4 v[NG];
#pragma unroll
      for (int u = 0; u < NG; ++u) {
        const int g = threadIdx.x + u * 256;
        if (g < nA / 4) {
          const int e = g * 4, i0l = e / per_row, rem = e - i0l * per_row, i1l = rem / tt, t = rem - i1l * tt;
          v[u] = __ldg(reinterpret_cast<const float4*>(tsrc + ((int64_t)(i0b + i0l) * d1 + i1b + i1l) * taps + tb + t));
        }
      }
#pragma unroll
      for (int u = 0; u < NG; ++u) {
        const int g = threadIdx.x + u * 256;
        if (g < nA / 4) {
          const int e = g * 4, i0l = e / per_row, rem = e - i0l * per_row;
          int i1l = rem / tt, t = rem - i1l * tt;
          float* row = sm + i0l * kTPitch0;
          const float vv[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            row[i1l * kTPitch1 + t] = vv[k];
            if (++t == tt) { t = 0; ++i1l; }
          }
        }
      }
      return;
    }
    for (int g = threadIdx.x; g < nA / 4; g += blockDim.x) {
      const int e = g * 4;
      const int i0l = e / per_row, rem = e - i0l * per_row;
      int i1l = rem / tt, t = rem - i1l * tt;
      const int64_t off = ((int64_t)(i0b + i0l) * d1 + i1b + i1l) * taps + tb + t;
      float* row = sm + i0l * kTPitch0;
      if (load) {
        const float4 v = *reinterpret_cast<const float4*>(tsrc + off);
        const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          row[i1l * kTPitch1 + t] = vv[k];
          if (++t == tt) { t = 0; ++i1l; }
        }
      } else {
        float vv[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          vv[k] = row[i1l * kTPitch1 + t];
          if (++t == tt) { t = 0; ++i1l; }
        }
        float4* d = reinterpret_cast<float4*>(tdst + off);
        if (accumulate) {
          const float4 o = *d;
          vv[0] += o.x; vv[1] += o.y; vv[2] += o.z; vv[3] += o.w;
        }
        *d = make_float4(vv[0], vv[1], vv[2], vv[3]);
      }
    }
  };
  auto packed_side_vec = [&](bool load) {
    // one item = 8 consecutive columns of one packed row: rfd -> (tap, i0l) rows of 32 i1 (4 items), else (tap, i1l) rows
    // of 16 i0 (2 items); tt * 64 items per tile either way
#pragma unroll 4
    for (int item = threadIdx.x; item < tt * 64; item += blockDim.x) {
      int t, i0l, i1l, r, c;
      if (rfd) { const int c8 = item & 3; i0l = (item >> 2) & 15; t = item >> 6; i1l = c8 * 8; r = i0b + i0l; c = i1b + i1l; }
      else     { const int c8 = item & 1; i1l = (item >> 1) & 31; t = item >> 6; i0l = c8 * 8; r = i1b + i1l; c = i0b + i0l; }
      const int64_t off = ((int64_t)(tb + t) * rows + r) * cols + c;
      float* cell = sm + i0l * kTPitch0 + i1l * kTPitch1 + t;
      const int step = rfd ? kTPitch1 : kTPitch0;          // next column = next i1 (rfd) or next i0
      if (load) {
        float4 a = *reinterpret_cast<const float4*>(psrc + off), b = *reinterpret_cast<const float4*>(psrc + off + 4);
        if (nslices > 1) {   // slice order, fp64 accumulation
          double d[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll 2
          for (int sl = 1; sl < nslices; ++sl) {
            const float* q = psrc + (int64_t)sl * slice_stride + off;
            const float4 a2 = __ldg(reinterpret_cast<const float4*>(q)), b2 = __ldg(reinterpret_cast<const float4*>(q + 4));
            d[0] += a2.x; d[1] += a2.y; d[2] += a2.z; d[3] += a2.w;
            d[4] += b2.x; d[5] += b2.y; d[6] += b2.z; d[7] += b2.w;
          }
          a = make_float4((float)d[0], (float)d[1], (float)d[2], (float)d[3]);
          b = make_float4((float)d[4], (float)d[5], (float)d[6], (float)d[7]);
        }
        cell[0] = a.x; cell[step] = a.y; cell[2 * step] = a.z; cell[3 * step] = a.w;
        cell[4 * step] = b.x; cell[5 * step] = b.y; cell[6 * step] = b.z; cell[7 * step] = b.w;
      } else {
        float v[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = cell[k * step];
        if (pdtype == GSB_BF16) Vec8<__nv_bfloat16>::store(reinterpret_cast<__nv_bfloat16*>(pdst) + off, v);
        else Vec8<float>::store(reinterpret_cast<float*>(pdst) + off, v);
      }
    }
  };
  if (PACK) {
    if (vec_t) torch_side_vec(true); else torch_side(true);
    __syncthreads();
    if (vec_p) packed_side_vec(false); else packed_side(false);
  } else {
    if (vec_p) packed_side_vec(true); else packed_side(true);
    __syncthreads();
    if (vec_t) torch_side_vec(false); else torch_side(false);
  }
}

template <bool PACK>
__device__ __forceinline__ void layout_tile_dispatch(const LayoutJobs& J, int j, int tile, float* sm) {
  const float* tsrc = PACK ? reinterpret_cast<const float*>(J.src[j]) : nullptr;
  float* tdst = PACK ? nullptr : reinterpret_cast<float*>(J.dst[j]);
  const float* psrc = PACK ? nullptr : reinterpret_cast<const float*>(J.src[j]);
  void* pdst = PACK ? J.dst[j] : nullptr;
  const int taps = J.taps[j], tt = taps < kTTmax ? taps : kTTmax;
#define GSB_LT(N) layout_tile<PACK, N>(tsrc, tdst, psrc, pdst, J.dst_dtype, J.d0[j], J.d1[j], taps, J.rows_from_d0[j], tile, J.accumulate, PACK ? 1 : J.nslices[j], sm)
  switch (tt) {
    case 16: GSB_LT(16); break;
    case 9: GSB_LT(9); break;
    case 4: GSB_LT(4); break;
    case 1: GSB_LT(1); break;
    default: GSB_LT(0); break;
  }
#undef GSB_LT
}

__global__ void __launch_bounds__(256) pack_weight_multi_kernel(const __grid_constant__ LayoutJobs J) {
  GSB_PDL_ENTRY();
  __shared__ float sm[kTI0 * kTPitch0];
  const int j = find_job(J.blk0, J.count, blockIdx.x);
  layout_tile_dispatch<true>(J, j, blockIdx.x - J.blk0[j], sm);
}

__global__ void __launch_bounds__(256) unpack_wgrad_multi_kernel(const __grid_constant__ LayoutJobs J) {
  GSB_PDL_ENTRY();
  __shared__ float sm[kTI0 * kTPitch0];
  const int j = find_job(J.blk0, J.count, blockIdx.x);
  layout_tile_dispatch<false>(J, j, blockIdx.x - J.blk0[j], sm);
}

// Jobs of an un-pack launch whose source holds many partial-sum slices (small layers split 24..74 ways over the pixels):
// the tile loader below would walk them one after the other per element (8 CTAs x 74 dependent 16-byte loads: 120 us).
// One launch first folds the slices of all such jobs into their slice 0, in place (the source workspace is scratch).
constexpr int kFoldCols = 1024;        // columns per block: one float4 per thread
struct SliceJobs {
  float* ws[kMaxLayoutJobs];
  int64_t n[kMaxLayoutJobs];
  int nslices[kMaxLayoutJobs];
  int blk0[kMaxLayoutJobs + 1];
  int count;
};
// One float4 column group per thread; the slices are added IN SLICE ORDER in fp64, eight independent 16-byte loads in
// flight per thread, no shared memory and no barrier (the form with 8 warps over the slices and a shared-memory reduction
// per 128 columns was barrier / latency bound: ncu profiles/r02k, 12-16 long-scoreboard + 6-9 barrier stall cycles per
// issue at 1.6-3.5 TB/s).  Fixed mapping and order -> deterministic.  n is a multiple of 4 and the workspace 16-byte
// aligned (checked by the caller).
__global__ void __launch_bounds__(256) fold_slices_multi_kernel(const __grid_constant__ SliceJobs J) {
  GSB_PDL_ENTRY();
  const int j = find_job(J.blk0, J.count, blockIdx.x);
  const int ns = J.nslices[j];
  const int64_t n = J.n[j];
  float* ws = J.ws[j];
  const int64_t col = ((int64_t)(blockIdx.x - J.blk0[j]) * 256 + threadIdx.x) * 4;
  if (col >= n) return;
  double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
  const float* src = ws + col;
  for (int s0 = 0; s0 < ns; s0 += 8) {
    float4 v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
      v[i] = s0 + i < ns ? __ldg(reinterpret_cast<const float4*>(src + (int64_t)(s0 + i) * n)) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      a0 += (double)v[i].x; a1 += (double)v[i].y; a2 += (double)v[i].z; a3 += (double)v[i].w;
    }
  }
  *reinterpret_cast<float4*>(ws + col) = make_float4((float)a0, (float)a1, (float)a2, (float)a3);
}

static int run_layout_multi(bool pack, int n, const void* const* src, void* const* dst, const int* d0, const int* d1,
                            const int* taps, const int* rows_from_d0, const int* nslices, int dst_dtype, int accumulate,
                            cudaStream_t st) {
  int launches = 0;
  bool folded[kMaxLayoutJobs * 4];
  for (int i = 0; i < n && i < kMaxLayoutJobs * 4; ++i) folded[i] = false;
  if (!pack && nslices != nullptr) {
    for (int begin = 0; begin < n; begin += kMaxLayoutJobs) {
      SliceJobs S;
      memset(&S, 0, sizeof(S));
      int blocks = 0;
      const int end = n - begin < kMaxLayoutJobs ? n : begin + kMaxLayoutJobs;
      for (int i = begin; i < end; ++i) {
        const int64_t ne = (int64_t)taps[i] * d0[i] * d1[i];
        if (nslices[i] <= 4 || i >= kMaxLayoutJobs * 4 || (ne & 3) != 0 || ((uintptr_t)src[i] & 15) != 0) continue;
        const int k = S.count++;
        S.ws[k] = reinterpret_cast<float*>(const_cast<void*>(src[i]));
        S.n[k] = ne;
        S.nslices[k] = nslices[i];
        S.blk0[k] = blocks;
        blocks += (int)((ne + kFoldCols - 1) / kFoldCols);
        folded[i] = true;
      }
      S.blk0[S.count] = blocks;
      if (S.count > 0 && blocks > 0) {
        gsb::launch(fold_slices_multi_kernel, gsb::LaunchCfg(blocks, 256, 0, st), S);
        ++launches;
      }
    }
  }
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
      J.nslices[k] = (nslices != nullptr && nslices[i] > 1 && !(i < kMaxLayoutJobs * 4 && folded[i])) ? nslices[i] : 1;
      J.blk0[k] = blocks;
      blocks += layout_tiles_host(d0[i], d1[i], taps[i]);
    }
    J.blk0[J.count] = blocks;
    if (blocks == 0) continue;
    if (pack) gsb::launch(pack_weight_multi_kernel, gsb::LaunchCfg(blocks, 256, 0, st), J);
    else gsb::launch(unpack_wgrad_multi_kernel, gsb::LaunchCfg(blocks, 256, 0, st), J);
    ++launches;
  }
  return check_launch(pack ? "pack_weight_multi" : "unpack_wgrad_multi", launches > 0 ? launches : 1);
}
}  // namespace gsb

extern "C" int gsb_pack_weight_multi(int n, const void* const* src, void* const* dst, const int* d0, const int* d1,
                                     const int* taps, const int* rows_from_d0, int dst_dtype, gsb_stream s) {
  if (n <= 0) return 0;
  GSB_REQUIRE(src && dst && d0 && d1 && taps && rows_from_d0, "pack_weight_multi: null argument");
  return gsb::run_layout_multi(true, n, src, dst, d0, d1, taps, rows_from_d0, nullptr, dst_dtype, 0, gsb::to_stream(s));
}

extern "C" int gsb_unpack_wgrad_multi(int n, const void* const* src, const int* nslices, void* const* dst, const int* d0,
                                      const int* d1, const int* taps, const int* rows_from_d0, int accumulate, gsb_stream s) {
  if (n <= 0) return 0;
  GSB_REQUIRE(src && dst && d0 && d1 && taps && rows_from_d0 && nslices, "unpack_wgrad_multi: null argument");
  return gsb::run_layout_multi(false, n, src, dst, d0, d1, taps, rows_from_d0, nslices, GSB_F32, accumulate, gsb::to_stream(s));
}

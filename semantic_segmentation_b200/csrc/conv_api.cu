// C-ABI dispatch for the convolution primitives: tcgen05 path when the shape allows, SIMT otherwise.
#include "common.cuh"

namespace gsb {
int conv_gather_direct(const gsb_view* in, const void* wp, int wdtype, const float* bias, const gsb_view* out,
                       const gsb_conv_geom* g, int act, int accumulate, cudaStream_t st);
int thin_tc_gather_try(const gsb_view* in, const void* wp, const float* bias, const gsb_view* out, const gsb_conv_geom* g, int act,
                       void* scratch, size_t scratch_bytes, cudaStream_t st);
int thin_tc_wgrad_try(const gsb_view* in, const gsb_view* dout, float* ws, int max_slices, int* used, const gsb_conv_geom* g,
                      cudaStream_t st);
int conv_wgrad_direct(const gsb_view* in, const gsb_view* dout, float* ws, int max_slices, int* used, const gsb_conv_geom* g,
                      cudaStream_t st);
bool conv_tc_supported(const gsb_view* in, const gsb_view* out, const gsb_conv_geom* g);
int conv_gather_tc(const gsb_view* in, const void* wp, const float* bias, const gsb_view* out, const gsb_conv_geom* g, int act,
                   float* stats, int* stats_rows, float* scratch, size_t scratch_bytes, cudaStream_t st, const float* scale = nullptr);
int conv_wgrad_tc(const gsb_view* in, const gsb_view* dout, float* ws, int max_slices, int* used, const gsb_conv_geom* g,
                  cudaStream_t st);
int conv_tc_m_tiles(const gsb_view* out);

static int out_extent(int mode, int in, int k, int stride, int pad) {
  return mode == GSB_MODE_CONV ? (in + 2 * pad - k) / stride + 1 : (in - 1) * stride - 2 * pad + k;
}
}  // namespace gsb

using namespace gsb;

static int check_geom(const gsb_conv_geom* g) {
  GSB_REQUIRE(g != nullptr, "conv: null geometry");
  GSB_REQUIRE(g->mode == GSB_MODE_CONV || g->mode == GSB_MODE_CONVT, "conv: bad mode %d", g->mode);
  GSB_REQUIRE(g->kh >= 1 && g->kw >= 1 && g->stride >= 1 && g->pad >= 0, "conv: bad geometry");
  return 0;
}

extern "C" size_t gsb_conv_stats_rows(const gsb_view* out, const gsb_conv_geom* g) {
  (void)g;
  return (size_t)conv_tc_m_tiles(out);
}

extern "C" int gsb_conv_gather(const gsb_view* in, const void* wp, int wdtype, const float* bias, const gsb_view* out,
                               const gsb_conv_geom* g, int act, int accumulate, int use_tc, float* stats_partial,
                               int* stats_rows, void* scratch, size_t scratch_bytes, gsb_stream s) {
  if (check_geom(g)) return -1;
  GSB_REQUIRE(in && out && wp && in->ptr && out->ptr, "conv_gather: null argument");
  GSB_REQUIRE(in->n == out->n, "conv_gather: batch mismatch");
  // gather-conv: `in` is the source; the source->destination relation is (mode, stride, pad)
  if (g->mode == GSB_MODE_CONV) {
    GSB_REQUIRE(out->h == out_extent(GSB_MODE_CONV, in->h, g->kh, g->stride, g->pad) &&
                    out->w == out_extent(GSB_MODE_CONV, in->w, g->kw, g->stride, g->pad),
                "conv_gather: output extent %dx%d does not match conv of %dx%d", out->h, out->w, in->h, in->w);
  } else {
    // transposed gather: any out extent whose forward conv gives back `in` is legal (output_padding)
    GSB_REQUIRE(out_extent(GSB_MODE_CONV, out->h, g->kh, g->stride, g->pad) == in->h &&
                    out_extent(GSB_MODE_CONV, out->w, g->kw, g->stride, g->pad) == in->w,
                "conv_gather: output extent %dx%d does not match transposed conv of %dx%d", out->h, out->w, in->h, in->w);
  }
  if (stats_rows) *stats_rows = 0;
  bool tc_ok = wdtype == GSB_BF16 && !accumulate && gsb_tc_available() && conv_tc_supported(in, out, g);
  if (use_tc == 1) GSB_REQUIRE(tc_ok, "conv_gather: tcgen05 path requested but shape/dtype unsupported");
  if (use_tc != 0 && tc_ok)
    return conv_gather_tc(in, wp, bias, out, g, act, stats_partial, stats_rows, reinterpret_cast<float*>(scratch), scratch_bytes,
                          gsb::to_stream(s));
  GSB_REQUIRE(stats_partial == nullptr || use_tc != 1, "conv_gather: stats need the tcgen05 path");
  if (use_tc != 0 && wdtype == GSB_BF16 && !accumulate && gsb_tc_available()) {   // CW -> 1 transposed 8x8 cell (thin_tc.cu)
    const int r = thin_tc_gather_try(in, wp, bias, out, g, act, scratch, scratch_bytes, gsb::to_stream(s));
    if (r != 0) return r < 0 ? -1 : 0;
  }
  return conv_gather_direct(in, wp, wdtype, bias, out, g, act, accumulate, gsb::to_stream(s));
}

// Convolution with a per-channel affine epilogue: out = act(conv(in) * scale[co] + shift[co]) -- inference with the BatchNorm
// of running statistics folded into the producing kernel (unet_parts.py:16-21 in eval mode: Conv -> BN -> ReLU is one pass,
// the normalised activation is never read back).  Returns 0 = launched, 1 = not applicable for this shape (the caller runs
// gsb_conv_gather + gsb_bn_act), -1 = error.  bf16 weights, tcgen05 halo kernel only (mode CONV, tiled space >= 16 x 16).
extern "C" int gsb_conv_gather_affine(const gsb_view* in, const void* wp, int wdtype, const float* scale, const float* shift,
                                      const gsb_view* out, const gsb_conv_geom* g, int act, gsb_stream s) {
  if (check_geom(g)) return -1;
  GSB_REQUIRE(in && out && wp && scale && in->ptr && out->ptr, "conv_gather_affine: null argument");
  GSB_REQUIRE(in->n == out->n, "conv_gather_affine: batch mismatch");
  if (g->mode != GSB_MODE_CONV || wdtype != GSB_BF16 || !gsb_tc_available() || !conv_tc_supported(in, out, g)) return 1;
  GSB_REQUIRE(out->h == out_extent(GSB_MODE_CONV, in->h, g->kh, g->stride, g->pad) &&
                  out->w == out_extent(GSB_MODE_CONV, in->w, g->kw, g->stride, g->pad),
              "conv_gather_affine: output extent %dx%d does not match conv of %dx%d", out->h, out->w, in->h, in->w);
  return conv_gather_tc(in, wp, shift, out, g, act, nullptr, nullptr, nullptr, 0, gsb::to_stream(s), scale);
}

// Weight gradient, deterministic split reduction: the launch writes `*used_slices` complete fp32 slices
// ws[slice][kh*kw][cout][cin] (one per split of the pixel range) with plain stores; the caller adds them in slice order
// (gsb_sum_slices, or gsb_unpack_wgrad(_multi) with the slice count).  gsb_conv_wgrad_plan returns the slice count the
// launch will use for this geometry on the current device, so the caller can size the workspace.
static int wgrad_dispatch(const gsb_view* in, const gsb_view* dout, float* ws, int max_slices, int* used, const gsb_conv_geom* g,
                          int use_tc, gsb_stream s) {
  if (check_geom(g)) return -1;
  GSB_REQUIRE(in && dout && in->ptr && dout->ptr && used, "conv_wgrad: null argument");
  GSB_REQUIRE(in->n == dout->n, "conv_wgrad: batch mismatch");
  *used = 0;
  bool tc_ok = gsb_tc_available() && conv_tc_supported(in, dout, g);
  if (use_tc == 1) GSB_REQUIRE(tc_ok, "conv_wgrad: tcgen05 path requested but shape/dtype unsupported");
  if (use_tc != 0 && tc_ok) return conv_wgrad_tc(in, dout, ws, max_slices, used, g, gsb::to_stream(s));
  if (use_tc != 0 && gsb_tc_available()) {
    const int r = thin_tc_wgrad_try(in, dout, ws, max_slices, used, g, gsb::to_stream(s));
    if (r != 0) return r < 0 ? -1 : 0;
  }
  return conv_wgrad_direct(in, dout, ws, max_slices, used, g, gsb::to_stream(s));
}

extern "C" int gsb_conv_wgrad_plan(const gsb_view* in, const gsb_view* dout, const gsb_conv_geom* g, int use_tc) {
  int used = 0;
  if (wgrad_dispatch(in, dout, nullptr, 0, &used, g, use_tc, nullptr)) return -1;
  return used;
}

extern "C" int gsb_conv_wgrad(const gsb_view* in, const gsb_view* dout, float* ws, int max_slices, int* used_slices,
                              const gsb_conv_geom* g, int use_tc, gsb_stream s) {
  GSB_REQUIRE(ws != nullptr && max_slices >= 1, "conv_wgrad: workspace required");
  return wgrad_dispatch(in, dout, ws, max_slices, used_slices, g, use_tc, s);
}

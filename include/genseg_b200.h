/* genseg_b200 -- C ABI of the B200-native GenSeg hot path (libgenseg_b200.so).
 *
 * The reference (importZL/semantic_segmentation) has no FFI/plugin seam: its hot path is
 * Python nn.Modules calling torch.nn layers.  This header is the
 * boundary a maintainer binds instead of those library calls; every entry point cites the
 * reference line(s) whose arithmetic it replaces.  Plain pointers and sizes only; the caller
 * owns every buffer; all functions enqueue on the caller's CUDA stream and return
 * 0 on success, <0 on error (message via gsb_last_error()).  There is no CPU fallback.
 */
#ifndef GENSEG_B200_H
#define GENSEG_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GSB_F32 0
#define GSB_BF16 1

#define GSB_ACT_NONE 0
#define GSB_ACT_RELU 1
#define GSB_ACT_LRELU 2 /* negative slope 0.2, networks.py:583,640 */
#define GSB_ACT_TANH 3

#define GSB_MODE_CONV 0  /* gather: iy = oy*stride - pad + ky           (nn.Conv2d) */
#define GSB_MODE_CONVT 1 /* gather: iy = (oy + pad - ky)/stride if exact (nn.ConvTranspose2d) */

/* A 4-D activation view with element strides (NHWC buffers, channel slices of concat
 * buffers and NCHW fp32 tensors at the API edge are all expressible). */
typedef struct {
  void* ptr;
  int32_t dtype; /* GSB_F32 | GSB_BF16 */
  int32_t n, h, w, c;
  int64_t sn, sh, sw, sc; /* strides in elements */
} gsb_view;

typedef struct {
  int32_t mode; /* GSB_MODE_* */
  int32_t kh, kw, stride, pad;
} gsb_conv_geom;

typedef void* gsb_stream; /* cudaStream_t */

int gsb_version(void);
const char* gsb_last_error(void);
/* Set aside n SMs (0..64) for a collective running concurrently on another stream: persistent kernels launched afterwards
 * size their grids for the remaining SMs.  Returns the previous value.  Host-side state, read at launch time. */
int gsb_set_sm_reserve(int n);
/* 1 if the running device is sm_100 and the tcgen05 kernels can launch. */
int gsb_tc_available(void);
/* Number of kernels this library has launched in the calling process (bench evidence). */
unsigned long long gsb_launch_count(void);

/* ---- layout / packing -------------------------------------------------------------- */
/* dst <- cast(src) element-wise over (n,h,w,c) with arbitrary strides (NCHW<->NHWC, dtype). */
int gsb_copy_view(const gsb_view* src, const gsb_view* dst, gsb_stream s);
/* Flat fp32 <-> bf16 cast (16-byte aligned buffers): the gradient buckets of the data-parallel all-reduce travel as bf16. */
int gsb_cast_flat(const void* src, int src_dtype, void* dst, int dst_dtype, int64_t n, gsb_stream s);
/* Pack a torch weight into the GEMM operand layout Wp[tap][rows][cols] (cols contiguous).
 * src is fp32 [d0][d1][kh][kw] (Conv2d: OIHW, unet_parts.py:16; ConvTranspose2d: IOHW,
 * operations.py:18).  rows_from_d0 != 0 -> rows = d0, cols = d1; else rows = d1, cols = d0. */
int gsb_pack_weight(const float* src, int d0, int d1, int kh, int kw, int rows_from_d0, void* dst,
                    int dst_dtype, gsb_stream s);
/* Inverse for gradients: dWp fp32 [nslices][tap][rows][cols] (partial-sum slices written by gsb_conv_wgrad, added here in
 * slice order) -> torch layout fp32 [d0][d1][kh][kw]; accumulate != 0 adds into dst. */
int gsb_unpack_wgrad(const float* src, int nslices, int d0, int d1, int kh, int kw, int rows_from_d0, float* dst,
                     int accumulate, gsb_stream s);
/* Batched forms of the two calls above: n (layer) jobs in one launch; taps = kh*kw; host arrays. */
int gsb_pack_weight_multi(int n, const void* const* src, void* const* dst, const int* d0, const int* d1, const int* taps,
                          const int* rows_from_d0, int dst_dtype, gsb_stream s);
int gsb_unpack_wgrad_multi(int n, const void* const* src, const int* nslices, void* const* dst, const int* d0, const int* d1,
                           const int* taps, const int* rows_from_d0, int accumulate, gsb_stream s);
/* dst[i] (+)= sum over s < nslices of ws[s * n + i], slice order, fp64 accumulation: the deterministic second stage of the
 * split reductions (no atomics anywhere in the library: two runs of a step give the same bits). */
int gsb_sum_slices(const float* ws, int nslices, int64_t n, float* dst, int accumulate, gsb_stream s);
/* Mixed transposed-conv cell (networks.py:495-511, operations.py:14-39): softmax(alpha[3])-
 * weighted, zero-padded sum of the IOHW k4/k6/k8 weights -> one IOHW 8x8 fp32 weight, and
 * merged bias (nullable).  SURVEY fact 4. */
int gsb_merge_upconv(const float* w4, const float* w6, const float* w8, const float* b4, const float* b6,
                     const float* b8, const float* alpha, int cin, int cout, float* w_merged,
                     float* b_merged, gsb_stream s);
/* Backward of gsb_merge_upconv: given dW (8x8 IOHW fp32) and db (nullable), produce
 * dW4/dW6/dW8 (+db4/6/8) and ACCUMULATE dalpha[3] (through the softmax Jacobian). */
int gsb_unmerge_upconv(const float* dw_merged, const float* db_merged, const float* w4, const float* w6,
                       const float* w8, const float* b4, const float* b6, const float* b8,
                       const float* alpha, int cin, int cout, float* dw4, float* dw6, float* dw8,
                       float* db4, float* db6, float* db8, float* dalpha, void* workspace,
                       size_t workspace_bytes, gsb_stream s);

/* ---- convolution (replaces cuDNN behind nn.Conv2d / nn.ConvTranspose2d) ------------- */
/* out = act(gather-conv(in, Wp) + bias).  Wp = [kh*kw][cout][cin], dtype wdtype.
 * Forward and data-gradient of both Conv2d and ConvTranspose2d are this one primitive
 * (mode + packing choose which).  use_tc: 0 = SIMT kernel (any shape, fp32 or bf16),
 * 1 = tcgen05/TMEM/TMA implicit GEMM (bf16 NHWC, cin%64==0, cout%64==0), 2 = auto.
 * stats_partial (nullable, TC path only): fp32 [m_tiles][2][cout] per-CTA sum / sum-of-squares
 * of the bf16-rounded outputs for BatchNorm; *stats_rows receives m_tiles.
 * accumulate != 0: out += result (SIMT path only).
 * scratch (nullable, caller-owned, 16-byte aligned): fp32 partial-sum slices for launches whose few tiles would leave most
 * SMs idle (deep generator levels): the K loop is then split over CTAs, part i stores into slice i, and a second kernel
 * adds the slices in slice order.  Without scratch the launch is not split.  The buffer may be reused by the next call on
 * the same stream. */
int gsb_conv_gather(const gsb_view* in, const void* wp, int wdtype, const float* bias, const gsb_view* out,
                    const gsb_conv_geom* g, int act, int accumulate, int use_tc, float* stats_partial,
                    int* stats_rows, void* scratch, size_t scratch_bytes, gsb_stream s);
/* Convolution with a per-channel affine epilogue, out = act(conv(in) * scale[co] + shift[co]): eval-mode Conv2d -> BatchNorm2d
 * (running statistics) -> ReLU of unet_parts.py:16-21 as ONE pass.  0 = launched, 1 = not applicable for this shape (the
 * caller then runs gsb_conv_gather + gsb_bn_act), -1 = error.  bf16 packed weights, forward convolutions on the tcgen05
 * halo kernel (tiled space >= 16 x 16) only; out may be a channel slice of a wider NHWC buffer. */
int gsb_conv_gather_affine(const gsb_view* in, const void* wp, int wdtype, const float* scale, const float* shift,
                           const gsb_view* out, const gsb_conv_geom* g, int act, gsb_stream s);
size_t gsb_conv_stats_rows(const gsb_view* out, const gsb_conv_geom* g);
/* Weight gradient: sum_pixels dout[pix][cout] * in[pix@tap][cin], with the forward op's gather mode.  The pixel range is
 * split over CTAs; split i STORES its partial sums into slice i of ws[max_slices][kh*kw][cout][cin] (fp32, every used slice
 * written completely, nothing to clear); *used_slices receives the count.  The caller adds the slices in slice order
 * (gsb_unpack_wgrad / _multi / gsb_sum_slices).  gsb_conv_wgrad_plan returns the slice count the launch will use. */
int gsb_conv_wgrad_plan(const gsb_view* in, const gsb_view* dout, const gsb_conv_geom* g, int use_tc);
int gsb_conv_wgrad(const gsb_view* in, const gsb_view* dout, float* ws, int max_slices, int* used_slices,
                   const gsb_conv_geom* g, int use_tc, gsb_stream s);
/* Network-edge convolutions as bandwidth kernels (one side has 1..4 / 1..8 channels).
 * stem: Conv2d(cin<=4 -> y.c, k x k, stride, pad) from the fp32 NCHW input x[n][cin][h][w] with the torch OIHW fp32
 *       weight (unet_model.py:14 / unet_parts.py:16; networks.py:582 outermost, :640); wgrad writes dweight (OIHW).
 * head: 1x1 Conv2d(x.c -> c<=8) + bias to fp32 NCHW logits (OutConv, unet_parts.py:71-77); backward writes
 *       dx (nullable), dweight[c][x.c] and dbias[c] (nullable).
 * The two weight-gradient calls take a caller-owned workspace for their per-CTA partial rows (stem: y.c*cin*k*k floats per
 * row, head: c*x.c + 8 floats per row; gsb_edge_rows() rows; fewer rows = fewer CTA columns), added in row order. */
/* stats_partial (optional): [gsb_stem_stats_rows()][2][y.c] fp32 per-block BatchNorm sum / sum-of-squares of the stored
 * values; *stats_rows receives the rows written (0 when the shape ran through the kernel without fused statistics). */
int gsb_stem_stats_rows(void);
int gsb_stem_conv_fwd(const float* x, int n, int cin, int h, int w, const float* weight, const float* bias,
                      const gsb_view* y, int k, int stride, int pad, int act, float* stats_partial, int* stats_rows,
                      gsb_stream s);
/* Inference form: out = act(conv(x) * scale[co] + shift[co]) in one pass (eval-mode Conv -> BatchNorm -> ReLU of
 * unet_parts.py:16-18 on the network input).  0 = launched, 1 = not applicable (run gsb_stem_conv_fwd + gsb_bn_act), -1 = error. */
int gsb_stem_conv_fwd_affine(const float* x, int n, int cin, int h, int w, const float* weight, const float* scale,
                             const float* shift, const gsb_view* y, int k, int stride, int pad, int act, gsb_stream s);
int gsb_edge_rows(void);
int gsb_stem_conv_wgrad(const float* x, int n, int cin, int h, int w, const gsb_view* dy, int k, int stride, int pad,
                        float* dweight, void* workspace, size_t workspace_bytes, gsb_stream s);
int gsb_head_conv_fwd(const gsb_view* x, const float* weight, const float* bias, int c, float* logits, gsb_stream s);
int gsb_head_conv_bwd(const gsb_view* x, const float* weight, int c, const float* dlogits, const gsb_view* dx,
                      float* dweight, float* dbias, void* workspace, size_t workspace_bytes, gsb_stream s);
/* Per-channel sum over pixels (conv bias gradient): db[c] is STORED.  workspace (nullable) >= gsb_bias_grad_workspace_bytes
 * enables the vectorised two-stage form for NHWC views. */
size_t gsb_bias_grad_workspace_bytes(const gsb_view* dout);
int gsb_bias_grad(const gsb_view* dout, float* db, void* workspace, size_t workspace_bytes, gsb_stream s);

/* ---- BatchNorm2d (+ activation, dropout) -- unet_parts.py:17-21, networks.py:584-607 -- */
/* sums[2][c] (fp64) <- sum x, sum x^2 over (n,h,w); workspace >= gsb_bn_workspace_bytes. */
size_t gsb_bn_workspace_bytes(const gsb_view* x);
/* per_image != 0 (here and in gsb_bn_act / gsb_bn_bwd_partial / gsb_bn_bwd_apply): InstanceNorm2d (networks.py:34-35) --
 * statistics per (image, channel) in the SAME launches: sums are [2][n*c], per-channel arrays [n][c], count = h*w. */
int gsb_bn_reduce(const gsb_view* x, int per_image, double* sums, void* workspace, size_t workspace_bytes, gsb_stream s);
/* Reduce TC-conv per-CTA partials [rows][2][c] fp32 -> sums[2][c] fp64. */
int gsb_bn_reduce_partials(const float* partial, int rows, int c, double* sums, gsb_stream s);
/* Training finalize: mean/biased var from sums & count -> scale = g*invstd, shift = b - mean*scale,
 * saves mean/invstd; running stats updated with momentum and UNBIASED var; nbt += 1 (all nullable
 * except sums/scale/shift).  Eval: pass sums = NULL -> scale/shift from running stats. */
int gsb_bn_finalize(const double* sums, double count, const float* gamma, const float* beta,
                    float* running_mean, float* running_var, int64_t* num_batches_tracked, float momentum,
                    float eps, int c, float* scale, float* shift, float* mean, float* invstd, int centered,
                    gsb_stream s);
/* z = (x - center)*scale + shift;  out1 = act1(drop(z)); out2 (nullable) = act2(z).
 * scale/shift/center nullable (identity / 0).  Two forms of the normalisation: folded (centered = 0 at finalize:
 * shift = beta - mean*scale, center = NULL; the bf16 production mode) and centered (centered = 1: shift = beta,
 * center = mean; the fp32 validation mode evaluates (x - mean)*scale + beta like torch, the folded form loses
 * |mean|/std digits to cancellation).  Dropout(0.5) on the out1 path: explicit keep-mask drop_mask
 * (uint8 dense NHWC, nullable) or, if drop_seed != 0, a stateless Philox4x32-10 mask keyed by
 * (drop_seed, drop_offset + local element index) that backward recomputes; drop_offset = first global sample of this
 * rank * h*w*c makes the masks of a batch sharded over N ranks those of the one-rank run. */
int gsb_bn_act(const gsb_view* x, const float* scale, const float* shift, const float* center, int act1,
               const gsb_view* out1, int act2, const gsb_view* out2, uint64_t drop_seed, int64_t drop_offset,
               const uint8_t* drop_mask, int per_image, gsb_stream s);
/* relu(x * scale + shift) written to out1 AND its MaxPool2d(2) written to pooled, one pass (unet_parts.py:21 -> :34, the
 * DoubleConv -> Down seam); bit-identical to gsb_bn_act(ReLU) followed by gsb_maxpool2_fwd.  Even h and w. */
int gsb_bn_act_pool(const gsb_view* x, const float* scale, const float* shift, const gsb_view* out1, const gsb_view* pooled,
                    gsb_stream s);
/* Explicit Dropout(0.5) keep mask (uint8 dense NHWC, the drop_mask argument of gsb_bn_act / gsb_bn_bwd_*) with the bits the
 * fused Philox path draws for seed = (seed_k + *calls) mod 2^62 + 1 + layer_term (networks.py:604 nn.Dropout(0.5); the
 * per-call seed rule of the generator program) -- the call count is read from device memory, so a captured CUDA graph draws a
 * new mask at every replay.  numel, drop_offset: multiples of 8. */
int gsb_dropout_mask(uint8_t* mask, int64_t numel, uint64_t seed_k, const int64_t* calls, uint64_t layer_term,
                     int64_t drop_offset, gsb_stream s);
/* Backward.  dy = act1'(z)*drop*dA1 + act2'(z)*dA2 (dA2 nullable), z as above.
 * reduce: sums[2][c] fp64 <- sum dy, sum dy*xhat.   apply (has_bn): dx = scale*(dy - s0/cnt - xhat*s1/cnt);
 * (no bn: pass mean = NULL -> dx = dy).  dgamma/dbeta accumulate (nullable). */
int gsb_bn_bwd_reduce(const gsb_view* x, const gsb_view* dA1, const gsb_view* dA2, const float* scale,
                      const float* shift, const float* center, const float* mean, const float* invstd, int act1, int act2,
                      uint64_t drop_seed, int64_t drop_offset, const uint8_t* drop_mask, double* sums, void* workspace,
                      size_t workspace_bytes, gsb_stream s);
int gsb_bn_bwd_apply(const gsb_view* x, const gsb_view* dA1, const gsb_view* dA2, const float* scale,
                     const float* shift, const float* center, const float* mean, const float* invstd, int act1, int act2,
                     uint64_t drop_seed, int64_t drop_offset, const uint8_t* drop_mask, const double* sums, double count,
                     const gsb_view* dx, float* dgamma, float* dbeta, int per_image, gsb_stream s);

/* dbeta[c] (+)= sums[c], dgamma[c] (+)= sums[c + C]: parameter gradients from the LOCAL backward sums (under
 * sync-BN the sums are all-reduced afterwards; the gradient all-reduce then adds the other ranks' parts). */
int gsb_bn_param_grad(const double* sums, int c, float* dgamma, float* dbeta, int accumulate, gsb_stream s);
/* Fused statistic stages (one launch each; with world > 1 the cross-rank sum runs inside the kernel over the peer
 * regions of gsb_peer_create / gsb_peer_open -- pass regions = NULL, rank = 0, world = 1 on a single GPU):
 * gsb_bn_stats_finalize = gsb_bn_reduce_partials + all-reduce + gsb_bn_finalize (count = global element count);
 * gsb_bn_bwd_partial    = stage 1 of gsb_bn_bwd_reduce (rows of per-block sums into workspace, *rows = row count);
 * gsb_bn_bwd_stats      = gsb_bn_reduce_partials + gsb_bn_param_grad (local sums) + all-reduce -> sums[2c]. */
int gsb_bn_stats_finalize(const float* partial, int rows, int c, double count, void* const* regions, int rank, int world,
                          const float* gamma, const float* beta, float* running_mean, float* running_var,
                          int64_t* num_batches_tracked, float momentum, float eps, float* scale, float* shift, float* mean,
                          float* invstd, int centered, gsb_stream s);
int gsb_bn_bwd_partial(const gsb_view* x, const gsb_view* dA1, const gsb_view* dA2, const float* scale, const float* shift,
                       const float* center, const float* mean, const float* invstd, int act1, int act2, uint64_t drop_seed,
                       int64_t drop_offset, const uint8_t* drop_mask, void* workspace, size_t workspace_bytes, int* rows,
                       int per_image, gsb_stream s);
/* gsb_bn_bwd_partial for the BatchNorm + ReLU layer in front of a MaxPool2d(2) (unet_parts.py:34), with the pooling
 * backward folded in: dA (in / out, the skip half of the concat gradient) += the pooled gradient dpool routed to the
 * first maximum of each 2x2 window of relu(scale*x+shift), then the rows of [sum dy, sum dy*xhat] as above.  Replaces
 * gsb_maxpool2_bwd(accumulate) + gsb_bn_bwd_partial (same dA, bit for bit).  bf16, even h and w. */
int gsb_bn_bwd_partial_pool(const gsb_view* x, const gsb_view* dA, const gsb_view* dpool, const float* scale, const float* shift,
                            const float* mean, const float* invstd, void* workspace, size_t workspace_bytes, int* rows,
                            gsb_stream s);
int gsb_bn_bwd_stats(const float* partial, int rows, int c, void* const* regions, int rank, int world, float* dgamma,
                     float* dbeta, int accumulate, double* sums, gsb_stream s);

/* ---- pooling / resampling -- unet_parts.py:34,50,57-67 -------------------------------- */
int gsb_maxpool2_fwd(const gsb_view* x, const gsb_view* y, gsb_stream s);
/* dx (+)= route(dy) to the first maximum of each 2x2 window of x (PyTorch tie rule). */
int gsb_maxpool2_bwd(const gsb_view* x, const gsb_view* dy, const gsb_view* dx, int accumulate,
                     gsb_stream s);
/* Bilinear x2 upsample (align_corners=True) of x written at offset (pad_top, pad_left) inside
 * out (rest of out's window zero-filled): nn.Upsample + F.pad + the cat slice write. */
int gsb_upsample2_fwd(const gsb_view* x, const gsb_view* out, int pad_top, int pad_left, gsb_stream s);
int gsb_upsample2_bwd(const gsb_view* dout, const gsb_view* dx, int pad_top, int pad_left, gsb_stream s);

/* ---- segmentation losses / metrics -- util/dice_score.py, train_end2end_jsrt.py:136,182 -- */
/* logits: fp32 NCHW [n][c][h][w]; labels int64 [n][h][w] (values < max(c,2)).
 * sums (fp64[4]): CE (or BCE) sum, sum p*t, sum p, sum t over the local batch.  The caller may
 * all-reduce sums across ranks before finalize (global-batch Dice). */
int gsb_seg_loss_reduce(const float* logits, const int64_t* labels, int n, int c, int h, int w, double* sums,
                        void* workspace, size_t workspace_bytes, gsb_stream s);
size_t gsb_seg_loss_workspace_bytes(int n, int c, int h, int w);
/* out[3] fp32: total, ce, dice_loss.  count = global n*h*w. */
int gsb_seg_loss_finalize(const double* sums, double count, float* out, gsb_stream s);
/* dlogits = gscale * d(total)/dlogits. */
int gsb_seg_loss_bwd(const float* logits, const int64_t* labels, int n, int c, int h, int w,
                     const double* sums, double count, const float* gscale, float* dlogits, gsb_stream s);
/* evaluate(): mask = sigmoid(x)>0.5 (c==1) or argmax (first max wins); integer counts per
 * (image, class>=1): counts[n][max(c-1,1)][3] int64 = inter, pred, true.  unet/evaluate.py:29-40. */
int gsb_eval_counts(const float* logits, const int64_t* labels, int n, int c, int h, int w, uint8_t* mask,
                    int64_t* counts, gsb_stream s);
/* dice_coeff pieces: per-slice sums[slices][3] fp64 = sum a*b, sum a, sum b over `inner` elements.
 * workspace: fp64 partial rows [rows][slices][3], rows <= 1024 (one per CTA column), added in row order. */
int gsb_dice_sums(const float* a, const float* b, int64_t slices, int64_t inner, double* sums, void* workspace,
                  size_t workspace_bytes, gsb_stream s);
/* gsb_dice_sums without a caller workspace and without any allocation (no cudaMalloc / cudaFree / allocating library
 * call, so it may be called where allocation is forbidden).  It has NO counterpart in the reference.
 * The partial rows live in a 96 KiB file-scope `static __device__` array that every
 * GPU loading the library holds until exit (its device address is resolved once per device with cudaGetSymbolAddress).
 * Limit: slices <= gsb_dice_sums_noalloc_max_slices() (= 64); more is rejected with an error (use gsb_dice_sums).
 * Same result as gsb_dice_sums given the same number of partial rows (rows = min(ceil(inner/4096), 1024,
 * 12288 / (3*slices))).  Every call shares the array and it keeps the previous call's contents (nothing needs
 * clearing: each launch stores every row it later reads); calls that OVERLAP on the GPU (different streams) race. */
int gsb_dice_sums_noalloc(const float* a, const float* b, int64_t slices, int64_t inner, double* sums, gsb_stream s);
size_t gsb_dice_sums_noalloc_max_slices(void);

/* ---- GAN / L1 losses -- networks.py:263-281, pix2pix_model.py:73,127 -------------------- */
/* mode 0 vanilla (BCE-with-logits vs constant label), 1 lsgan (MSE), 2 wgangp (+-mean).
 * loss[0] (fp32) = mean; dpred (nullable) = gscale * dloss/dpred. */
int gsb_gan_loss(const float* pred, int64_t numel, int mode, float label, float* loss, const float* gscale,
                 float* dpred, void* workspace /* one double per CTA, needed when loss != NULL */, size_t workspace_bytes,
                 gsb_stream s);
int gsb_l1_loss(const float* a, const float* b, int64_t numel, float* loss, const float* gscale, float* da,
                void* workspace /* one double per CTA, needed when loss != NULL */, size_t workspace_bytes, gsb_stream s);

/* dx = dout * act'(.) from the activation OUTPUT (contiguous fp32; the Tanh of the generator's
 * outermost block, networks.py:592). */
int gsb_act_bwd(const float* out, const float* dout, int act, float* dx, int64_t numel, gsb_stream s);

/* ---- optimisers -- train_end2end_jsrt.py:71-72, pix2pix_model.py:75-78 ------ */
/* torch.optim.RMSprop(momentum, weight_decay, alpha, eps, centered=False) over a flat fp32 range. */
int gsb_rmsprop_step(float* p, const float* g, float* square_avg, float* mom_buf, int64_t numel, float lr,
                     float alpha, float eps, float weight_decay, float momentum, float grad_scale,
                     gsb_stream s);
/* torch.optim.Adam (no amsgrad). step = 1-based step count. */
int gsb_adam_step(float* p, const float* g, float* exp_avg, float* exp_avg_sq, int64_t numel, float lr,
                  float beta1, float beta2, float eps, float weight_decay, int64_t step, float grad_scale,
                  gsb_stream s);
/* Multi-tensor forms: n tensors in one launch (host arrays of device pointers / element counts). */
int gsb_rmsprop_step_multi(int n, float* const* p, const float* const* g, float* const* square_avg, float* const* mom_buf,
                           const int64_t* numel, float lr, float alpha, float eps, float weight_decay, float momentum,
                           float grad_scale, gsb_stream s);
int gsb_adam_step_multi(int n, float* const* p, const float* const* g, float* const* exp_avg, float* const* exp_avg_sq,
                        const int64_t* numel, float lr, float beta1, float beta2, float eps, float weight_decay, int64_t step,
                        float grad_scale, gsb_stream s);
/* Adam with the step count in DEVICE memory, for training steps replayed as a CUDA graph (torch.optim.Adam(capturable=True)
 * semantics, pix2pix_model.py:75-78): *step_dev is the count of THIS step, advanced by gsb_counter_add before the call. */
int gsb_adam_step_multi_dev(int n, float* const* p, const float* const* g, float* const* exp_avg, float* const* exp_avg_sq,
                            const int64_t* numel, float lr, float beta1, float beta2, float eps, float weight_decay,
                            const int64_t* step_dev, float grad_scale, gsb_stream s);
/* *ctr += delta on the stream (device-side step / call counters of captured training steps). */
int gsb_counter_add(int64_t* ctr, int64_t delta, gsb_stream s);

/* ---- Device-side image chain of the GenSeg end-to-end step (running_files/train_end2end_jsrt.py:186-200).
 * gsb_minmax_f32      : ordered keys of (min, max) over the whole batch tensor (:197), keys = 2 x uint32 on the device.
 * gsb_quantize_hist   : q = uint8(clamp(((x-min)/(max-min))*255 + 0.5, 0, 255)) (:197,:199) + 256-bin histogram per plane
 *                       (x: [planes, plane] fp32; keys NULL = x already normalised).
 * gsb_equalize_lut    : torchvision.transforms.functional.equalize per plane (:200) followed by a 256-entry LUT
 *                       (adjust_gamma(.., 0.5), supplied by the host) -> out_f = value / 255.0f and / or out_u8.
 * gsb_mask_affine     : inverse-mapped bilinear warp (zero outside) of n fp32 masks [n,h,w] by per-image 2x3 matrices
 *                       (output -> input pixel coordinates) + (> threshold) binarisation (:187-195; replaces the CPU imgaug
 *                       round trip).  threshold < 0 writes the interpolated value. */
int gsb_minmax_f32(const float* x, int64_t n, uint32_t* keys, gsb_stream s);
int gsb_quantize_hist(const float* x, const uint32_t* keys, uint8_t* q, uint32_t* hist, int planes, int plane, gsb_stream s);
int gsb_equalize_lut(const uint8_t* q, const uint32_t* hist, uint32_t* ws_hist, const uint8_t* post, float* out_f,
                     uint8_t* out_u8, int planes, int plane, int do_equalize, gsb_stream s);
int gsb_mask_affine(const float* in, const float* mats, float* out, int n, int h, int w, float threshold, gsb_stream s);

/* ---- NVLink peer-memory all-reduce of small fp64 vectors (the sync-BN statistic exchanges and the CE / Dice
 * sums; replaces the nn.DataParallel wrap of models_pix2pix/networks.py:117-121 for the batch-sharded path).
 * Every rank (one process per GPU, same node) creates a region, publishes its CUDA IPC handle, opens its peers'
 * handles, and then calls gsb_peer_allreduce_f64 in the same order on every rank: one kernel per call stores the
 * vector into all regions over NVLink, exchanges sequence flags and sums in rank order (bit-identical on all ranks).
 * n <= gsb_peer_max_elems(), world <= 8.  Graph-capturable (the sequence counter lives in device memory).
 * Every rank must issue the same sequence of exchanges.  A wait longer than GSB_PEER_TIMEOUT_S seconds (default 1800,
 * 0 = no limit) does NOT trap: the kernel records the call's sequence number in its region's error word, keeps its local
 * values and returns; gsb_peer_error(own region, &seq) reads that word (0 = healthy; synchronises the device). */
int64_t gsb_peer_region_bytes(void);
int gsb_peer_max_elems(void);
int gsb_peer_handle_bytes(void);
int gsb_peer_create(void** region, void* handle_out);
int gsb_peer_open(const void* handle, void** peer_region);
int gsb_peer_close(void* peer_region);
int gsb_peer_destroy(void* region);
int gsb_peer_error(void* region, unsigned long long* seq_out);
int gsb_peer_allreduce_f64(double* buf, int n, void* const* regions, int rank, int world, gsb_stream s);

#ifdef __cplusplus
}
#endif
#endif /* GENSEG_B200_H */

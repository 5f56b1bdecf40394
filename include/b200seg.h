/* b200seg.h — C ABI of libb200seg.so, the sm_100a (B200) implementation of the pytorch_segmentation hot path.
 *
 * Plain pointers and sizes only; every call enqueues work on the CUDA stream it is given and returns
 * immediately (0 = OK, non-zero = error, text via b200seg_last_error()).  All device buffers are owned by the
 * caller; the library allocates nothing on the device.  There is no CPU and no cuDNN fallback: an
 * unsupported shape is an error.
 *
 * Data layout: activations are NHWC bf16; a `b200seg_act`-style triple (ptr, c_total, c_offset) addresses a
 * channel slice of a wider buffer so producers write straight into concat buffers (replaces torch.cat at
 * reference models/aspp.py:36, models/deeplabv3plus.py:38, models/unet.py:34,40,46).
 *
 * Each entry point cites the reference call it replaces (paths relative to the reference root).
 */
#ifndef B200SEG_H
#define B200SEG_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* b200seg_stream_t; /* a cudaStream_t */

#define B200SEG_MAX_CONV_SRC 4

/* ---- housekeeping -------------------------------------------------------------------------------------- */
const char* b200seg_last_error(void);
int b200seg_version(void);
/* fills SM count and compute capability of the current device; fails if it is not sm_100 */
int b200seg_device_info(int* sm_count, int* cc_major, int* cc_minor);

/* Diagnostics: when buf != NULL, the epilogue of CTA 0 of every following conv launch writes globaltimer stamps
 * (int64 nanoseconds; 8 per staged 64-channel slab, at most `slots` slabs) into the device buffer; NULL turns
 * it off.  Used by scripts/ to attribute epilogue time; never set on the product path. */
int b200seg_debug_timing(int64_t* buf, int slots);

/* ---- layout / packing ---------------------------------------------------------------------------------- */
/* fp32 NCHW -> bf16 NHWC slice, and back (gradients).  Boundary of the drop-in forward(x: N x C x H x W)
 * (models/deeplabv3plus.py:28, models/unet.py:27). */
int b200seg_nchw_to_nhwc(const float* src, void* dst, int n, int c, int h, int w, int dst_c_total, int dst_c_offset,
                         b200seg_stream_t st);
int b200seg_nhwc_to_nchw(const void* src, float* dst, int n, int c, int h, int w, int src_c_total, int src_c_offset,
                         b200seg_stream_t st);
/* fp32 OIHW conv weight (nn.Conv2d.weight inside ConvNormAct / cls_conv) -> two bf16 packings:
 *   w_fprop [r*s][k_rows][c]                (tap-major, K-major over input channels)
 *   w_dgrad [r*s flipped][c_rows][k_cols]   (taps mirrored, in/out channels swapped) — either may be NULL.
 * k_rows >= k, c_rows >= c and k_cols >= k (a multiple of 8: the dgrad reduction runs over the channels of the
 * padded gradient buffer) are zero padded. */
int b200seg_pack_weights(const float* w_oihw, int k, int c, int r, int s, void* w_fprop, int k_rows, void* w_dgrad,
                         int c_rows, int k_cols, b200seg_stream_t st);

/* Patch gather for few-channel inputs — the 3-channel stems: ResNet-50 conv1 7x7 stride 2 behind
 * models/deeplabv3plus.py:17 (external pytorch_modules.backbones.resnet50) and HRNet's ConvNormAct(3, 64, 3, 2)
 * (models/hrnet.py:260).  fp32 NCHW image -> bf16 NHWC matrix with c_pad channels per output pixel:
 *   col[n][oy][ox][(r*S + s)*C + c] = x[n][c][oy*stride + r - pad][ox*stride + s - pad]   (0 outside / beyond R*S*C)
 * c_pad = R*S*C rounded up to a multiple of 8.  The convolution is then the 1x1 tcgen05 GEMM over col with the
 * weight packed by b200seg_pack_weights_im2col ([k_rows][c_pad], same column order); its weight gradient (fp32
 * [k][c_pad] from b200seg_conv_wgrad with r = s = 1) is scattered back to OIHW by b200seg_unpack_wgrad_im2col. */
int b200seg_im2col_nchw(const float* x, int n, int c, int h, int w, int r, int s, int stride, int pad, void* col,
                        int c_pad, b200seg_stream_t st);
int b200seg_pack_weights_im2col(const float* w_oihw, int k, int c, int r, int s, void* w_fprop, int k_rows, int c_pad,
                                b200seg_stream_t st);
int b200seg_unpack_wgrad_im2col(const float* dw_col, int k, int c, int r, int s, int c_pad, float* dw_oihw,
                                int accumulate, b200seg_stream_t st);
/* Space-to-depth form of a stride-2 r x s conv over c <= 4 channels (ResNet-50's 7x7 conv1, torchvision resnet.py via
 * models/deeplabv3plus.py:17; HRNet's ConvNormAct(3, 64, 3, 2), models/hrnet.py:260) — replaces the gathered matrix:
 * z[n][Y][X][(2b + b')*c + ch] = x[n][ch][2Y + b - pad][2X + b' - pad] (bf16, cs = 4c rounded up to 8 channels, 0 outside
 * the image) on a (oh + ceil(r/2) - 1) x (ow + ceil(s/2) - 1) grid; the conv is then ceil(r/2) row taps over windows of
 * ceil(s/2)*cs contiguous elements: an anchored window source of b200seg_conv_fprop / b200seg_conv_wgrad with
 * x_rows, x_cols = that grid, x_c_total = cs, c = ceil(s/2)*cs, taps ceil(r/2) x 1. */
int b200seg_space_to_depth_nchw(const float* x, int n, int c, int h, int w, int r, int s, int pad, void* z, int cs,
                                b200seg_stream_t st);
/* wp[a][kk][a'*cs + (2b + b')*c + ch] = w[kk][ch][2a + b][2a' + b'] (bf16 [ceil(r/2)][k_rows][ceil(s/2)*cs]) */
int b200seg_pack_weights_s2d(const float* w, int k, int c, int r, int s, void* wp, int k_rows, int cs, b200seg_stream_t st);
/* dw[kk][ch][r][s] (+)= dwin[kk][a'*cs + (2b + b')*c + ch][a] — dwin is b200seg_conv_wgrad's [k][ceil(s/2)*cs][ceil(r/2)][1] */
int b200seg_unpack_wgrad_s2d(const float* dwin, int k, int c, int r, int s, int cs, float* dw, int accumulate,
                             b200seg_stream_t st);

/* Narrow-output 3x3 convolutions (cls_conv, models/deeplabv3plus.py:22,39 and models/unet.py:23,51: 384 -> 21 / 88 -> nc
 * classes) restated so the wide input is read once instead of once per tap (see csrc/taps.cu):
 *   fprop  Z = X * Wz (1x1 GEMM, N = r*s*ncp), then b200seg_tap_scatter_add: y[p][k] = bias[k] + sum_tap Z[p+off(tap)][tap*ncp+k]
 *   bwd    b200seg_tap_gather: col[p][tap*c + ch] = x[p+off(tap)][ch]; dX = col * Wcol (1x1 GEMM); dWcol = col^T X (1x1 wgrad),
 *          b200seg_unpack_wgrad_taps: dW[k][c][tap] = dWcol[mirror(tap)*ncp + k][c]
 * b200seg_pack_weights_taps writes Wz [z_rows >= r*s*ncp][c] and Wcol [c_rows >= c][r*s*ncp] (bf16, either may be NULL). */
int b200seg_tap_gather(const void* x, int x_c_total, int x_c_offset, int c, int n, int h, int w, int r, int s, int dil,
                       void* col, int col_c_total, b200seg_stream_t st);
int b200seg_tap_scatter_add(const void* z, int z_c_total, int ncp, int n, int h, int w, int r, int s, int dil,
                            const float* bias, float* y, b200seg_stream_t st);
int b200seg_pack_weights_taps(const float* w_oihw, int k, int c, int r, int s, int ncp, void* wz, int z_rows, void* wcol,
                              int c_rows, b200seg_stream_t st);
int b200seg_unpack_wgrad_taps(const float* dw_col, int k, int c, int r, int s, int ncp, float* dw_oihw, int accumulate,
                              b200seg_stream_t st);

/* All conv weights of a model re-packed in one or two launches (after an optimizer step every weight is stale at
 * once): per tensor the same packings as b200seg_pack_weights, or — im2col != 0 — as b200seg_pack_weights_im2col. */
typedef struct b200seg_pack_desc {
    const float* w;          /* fp32 OIHW */
    int32_t k, c, r, s;
    void* w_fprop;           /* bf16 [r*s][k_rows][c]   (im2col: [k_rows][c_pad]) */
    void* w_dgrad;           /* bf16 [r*s mirrored][c_rows][k_cols] or NULL */
    int32_t k_rows, c_rows, k_cols;
    int32_t im2col, c_pad;
    int32_t reserved;
} b200seg_pack_desc;
int b200seg_pack_weights_multi(const b200seg_pack_desc* descs, int n, b200seg_stream_t st);

/* ---- optimizer step (the update the reference's external Trainer runs after train.py:64's
 * backward): torch.optim.SGD semantics over ALL parameters in one or two launches.
 *   g' = grad_scale*g + weight_decay*p;  m = first_step ? g' : momentum*m + (1-dampening)*g';
 *   p -= lr * (nesterov ? g' + momentum*m : m)        (momentum == 0: p -= lr*g', m may be NULL)
 * grad_scale folds the 1/world_size of the data-parallel gradient average into the update. */
typedef struct b200seg_sgd_desc {
    float* p;
    const float* g;
    float* m;
    int64_t numel;
} b200seg_sgd_desc;
/* hyper: NULL, or a DEVICE vector {lr, momentum, dampening, weight_decay, grad_scale} read by the kernel in place of the
 * by-value arguments — a captured CUDA graph then follows a learning-rate schedule (the host refreshes the vector before
 * each replay).  Whether momentum / nesterov / first_step apply is fixed by the by-value arguments. */
int b200seg_sgd_multi(const b200seg_sgd_desc* descs, int n, float lr, float momentum, float dampening, float weight_decay,
                      int nesterov, int first_step, float grad_scale, const float* hyper, b200seg_stream_t st);

/* Adam / AdamW for all parameters in one launch per 64 tensors (train.py:96 ``--adam``; torch.optim.Adam(W), amsgrad off):
 *   g' = grad_scale*g (+ weight_decay*p if !decoupled);  decoupled: p *= 1 - lr*weight_decay
 *   m += (1-beta1)*(g'-m);  v = beta2*v + (1-beta2)*g'^2;  p -= (lr/bc1) * m / (sqrt(v)/sqrt_bc2 + eps)
 * hyper = DEVICE vector of 11 floats {lr, beta1, beta2, eps, weight_decay, grad_scale, lr/(1-beta1^t), sqrt(1-beta2^t),
 * 1-beta1, 1-beta2, 1-lr*weight_decay} (derived entries formed by the host in double precision).
 * first_step: m and v are not read (treated as zero). */
typedef struct b200seg_adam_desc {
    float* p;
    const float* g;
    float* m;
    float* v;
    int64_t numel;
} b200seg_adam_desc;
int b200seg_adam_multi(const b200seg_adam_desc* descs, int n, const float* hyper, int decoupled, int first_step,
                       b200seg_stream_t st);

/* dst <- src for many fp32 tensors in one launch: the gradients of a bucket move into their slots of the flat
 * all-reduce buffer (parallel.GradBucketer; the DDP exchange of train.py:108-119).  desc.p = destination, desc.g =
 * source, desc.m unused. */
int b200seg_copy_multi(const b200seg_sgd_desc* descs, int n, b200seg_stream_t st);

/* ---- convolution: tcgen05 implicit GEMM ------------------------------------------------------------------ */
/* One input of a "same"-padded convolution (stride 1 or 2): activation slice + packed weights + filter geometry. */
typedef struct b200seg_conv_src {
    const void* x;        /* bf16 NHWC */
    int32_t x_c_total;    /* channels per pixel of the buffer */
    int32_t x_c_offset;   /* first channel of the slice (multiple of 8) */
    int32_t c;            /* channels in the slice (multiple of 8) */
    const void* w;        /* packed bf16 [r*s][w_rows][c] from b200seg_pack_weights */
    int32_t w_rows;       /* rows per tap in the packed weight (>= output channels) */
    int32_t r, s;         /* filter taps: 1x1 or 3x3 */
    int32_t dil;          /* dilation; padding is ((r-1)/2)*dil */
    int32_t stride;       /* 1 or 2 (2 needs dil == 1).  fprop: x is (stride*h) x (stride*w).  dgrad: x (= dY) is
                             (h/stride) x (w/stride) and the call's n,h,w is the dX grid */
    int32_t x_rows, x_cols; /* 0, 0: regular source.  Otherwise an ANCHORED WINDOW source (fprop only, the single source
                             of the call): x is an [n][x_rows][x_cols][x_c_total] buffer that carries its own padding
                             and tap (i, j) of the r x s (r*s <= 9) filter reads, for output pixel (y, x), the c
                             CONTIGUOUS elements that start at channel x_c_offset of buffer pixel (y + i, x + j).  c may
                             exceed x_c_total: the window then runs on into the following pixels, so neighbouring
                             windows overlap — how the space-to-depth stem (b200seg_space_to_depth_nchw) turns a
                             strided few-channel conv into row taps over 64-element windows without a gathered matrix */
} b200seg_conv_src;

typedef struct b200seg_conv_out {
    void* y;              /* bf16 (or fp32 when out_fp32) NHWC */
    int32_t y_c_total;
    int32_t y_c_offset;
    int32_t k;            /* output channels stored (multiple of 8) */
    int32_t out_fp32;
    const float* scale;   /* per-channel multiplier or NULL (eval-mode BN folded: gamma/sqrt(var+eps)) */
    const float* bias;    /* per-channel addend or NULL (BN shift, or nn.Conv2d bias) — length >= k */
    int32_t relu;
    float* stats;         /* NULL, or fp32 [2*k] zeroed by the caller: the epilogue adds the per-channel sum and sum
                             of squares of the conv result as stored (rounded to bf16 when the output is bf16; before
                             scale/bias/relu) — the training-mode nn.BatchNorm2d statistics of ConvNormAct
                             without a second pass over the output */
} b200seg_conv_out;

/* y = act(scale * sum_src conv(x_src, w_src) + bias) on the n x h x w OUTPUT grid.
 * Replaces nn.Conv2d inside ConvNormAct (models/aspp.py:12,27,29,30; models/deeplabv3plus.py:20;
 * models/unet.py:19-21) and cls_conv (models/deeplabv3plus.py:22,39; models/unet.py:23,51).  With w packed by
 * the dgrad packing and x = dy this is the data gradient; several sources accumulate in TMEM, which is how
 * the five ASPP branches' gradients w.r.t. the shared input are summed without extra passes. */
int b200seg_conv_fprop(int n, int h, int w, int n_src, const b200seg_conv_src* src, const b200seg_conv_out* out,
                       b200seg_stream_t st);
/* Same kernel; separate symbol so call sites say which autograd edge they are (conv backward-data). */
int b200seg_conv_dgrad(int n, int h, int w, int n_src, const b200seg_conv_src* src, const b200seg_conv_out* out,
                       b200seg_stream_t st);

typedef struct b200seg_wgrad_desc {
    int32_t n, h, w;
    const void* dy;       /* bf16 NHWC gradient of the conv output */
    int32_t dy_c_total, dy_c_offset, k;   /* k: channels in the dy slice (multiple of 8) */
    const void* x;        /* bf16 NHWC conv input */
    int32_t x_c_total, x_c_offset, c;
    int32_t r, s, dil;
    int32_t stride;       /* 1 or 2 (2 needs dil == 1): n, h, w is the dY grid, x is (stride*h) x (stride*w) */
    float* dw;            /* fp32 OIHW [k_out][c][r][s] */
    int32_t k_out;        /* logical output channels (<= k) */
    int32_t accumulate;   /* dw += result instead of dw = result */
    void* workspace;
    size_t workspace_bytes;
    int32_t x_rows, x_cols; /* 0, 0: regular.  Otherwise x is an anchored window buffer as in b200seg_conv_src (dil = 1,
                             stride = 1, r*s <= 9): dw[k][c][r][s] is the gradient of the window weights */
} b200seg_wgrad_desc;

size_t b200seg_conv_wgrad_workspace(const b200seg_wgrad_desc* d);
/* dW[k][c][r][s] = sum_pixels dy[p][k] * x[p + tap][c]  (conv backward-weights of the same convs). */
int b200seg_conv_wgrad(const b200seg_wgrad_desc* d, b200seg_stream_t st);
/* The same weight gradient for callers that cannot allocate device memory (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * d->workspace / d->workspace_bytes are ignored: the split-K scratch is a static device array inside the
 * library of B200SEG_NOALLOC_WGRAD_MAX_BYTES bytes (held by every GPU that loads the library until the process exits),
 * and the call makes no cudaMalloc / cudaFree or any other allocating call.  A descriptor whose
 * b200seg_conv_wgrad_workspace() exceeds that limit is rejected (returns 1, b200seg_last_error() names both sizes).
 * Results are those of b200seg_conv_wgrad.  Every call on a device shares the one array, which keeps the last call's
 * contents (nothing in it has to start at zero): calls that overlap on the GPU — two streams, parallel graph branches —
 * race on it; serialise them. */
#define B200SEG_NOALLOC_WGRAD_MAX_BYTES (16u << 20)
int b200seg_conv_wgrad_noalloc(const b200seg_wgrad_desc* d, b200seg_stream_t st);

/* ---- BatchNorm (train mode) around the conv -------------------------------------------------------------- */
/* Per-channel sum and sum of squares of a bf16 NHWC slice into stats[2*c] (fp32; zeroed by the call).
 * nn.BatchNorm2d inside ConvNormAct, training statistics. */
int b200seg_bn_stats(const void* x, int x_c_total, int x_c_offset, int c, int64_t pixels, float* stats,
                     b200seg_stream_t st);
/* stats -> scale/shift for the apply pass, saved mean / inv-std for backward, running-stat update
 * (momentum, unbiased variance) exactly as nn.BatchNorm2d. */
int b200seg_bn_finalize(const float* stats, int c, int64_t count, const float* gamma, const float* beta, float eps,
                        float momentum, float* running_mean, float* running_var, float* scale, float* shift,
                        float* save_mean, float* save_invstd, b200seg_stream_t st);
/* y = relu?(x*scale + shift) from one NHWC slice into another (BN apply + nn.ReLU + concat placement). */
int b200seg_bn_apply(const void* x, int x_c_total, int x_c_offset, void* y, int y_c_total, int y_c_offset, int c,
                     int64_t pixels, const float* scale, const float* shift, int relu, b200seg_stream_t st);
/* b200seg_bn_finalize + b200seg_bn_apply_add in one launch (train mode): scale / shift come straight from the conv
 * epilogue's stats[2c] (count = pixels); vec[4c] receives scale, shift, mean, invstd for the backward and the running
 * statistics move exactly as in b200seg_bn_finalize.  res / res_scale / res_shift as in b200seg_bn_apply_add.
 * mask_bits (optional, [pixels][c/8] bytes): bit j of byte (p, g) = (y[p][8g + j] > 0) — the ReLU mask of the output in
 * 1/16 of its bytes, read by b200seg_grad_add_masked_bits in backward instead of the output itself. */
int b200seg_bn_apply_train(const void* x, int x_c_total, int x_c_offset, void* y, int y_c_total, int y_c_offset, int c,
                           int64_t pixels, const float* stats, const float* gamma, const float* beta, float eps,
                           float momentum, float* running_mean, float* running_var, float* vec, int relu, const void* res,
                           int res_c_total, int res_c_offset, const float* res_scale, const float* res_shift,
                           uint8_t* mask_bits, b200seg_stream_t st);
/* Residual form used by the ResNet bottlenecks of the backbone (pytorch_modules.backbones.resnet50, called at
 * models/deeplabv3plus.py:17-19,29): y = relu?(x*scale + shift + r) with r = res (identity) or res*res_scale +
 * res_shift (the downsample conv's raw output and its BN); res == NULL reduces to b200seg_bn_apply. */
int b200seg_bn_apply_add(const void* x, int x_c_total, int x_c_offset, void* y, int y_c_total, int y_c_offset, int c,
                         int64_t pixels, const float* scale, const float* shift, int relu, const void* res,
                         int res_c_total, int res_c_offset, const float* res_scale, const float* res_shift,
                         b200seg_stream_t st);
/* dx += dy * (y > 0): the identity branch's gradient joins the conv branch's at a residual block's input. */
int b200seg_relu_bwd_add(void* dx, int dx_c_total, int dx_c_offset, const void* dy, int dy_c_total, int dy_c_offset,
                         const void* y, int y_c_total, int y_c_offset, int c, int64_t pixels, b200seg_stream_t st);
/* dx = (dx + dy * [y > 0]) * [xin > 0] over bf16 NHWC slices; dy, y and xin are each optional (NULL).  The gradient
 * join of a residual block (`out += residual; relu(out)`, models/hrnet.py:49-52,97-100 and the torchvision bottleneck
 * behind models/deeplabv3plus.py:17) with the ReLU mask of the block INPUT applied on the way out, so the producing
 * block's BatchNorm backward receives a pre-masked gradient. */
int b200seg_grad_add_masked(void* dx, int dx_c_total, int dx_c_offset, const void* dy, int dy_c_total, int dy_c_offset,
                            const void* y, int y_c_total, int y_c_offset, const void* xin, int xin_c_total,
                            int xin_c_offset, int c, int64_t pixels, b200seg_stream_t st);
/* Same with the block-input mask given as bits (b200seg_bn_apply_train's mask_bits of that tensor). */
int b200seg_grad_add_masked_bits(void* dx, int dx_c_total, int dx_c_offset, const void* dy, int dy_c_total, int dy_c_offset,
                                 const void* y, int y_c_total, int y_c_offset, const uint8_t* xin_bits, int c, int64_t pixels,
                                 b200seg_stream_t st);
/* The same join fused with the FIRST pass of the BatchNorm backward that consumes its result: the joined gradient of a
 * block input is the (pre-masked) output gradient of the ResNet block before it, whose bn3 — and downsample BN, if any —
 * backward begins by reducing it against the saved raw conv output.  dx <- (dx + dy) * mask(xin_bits);
 * red3[0..c) += sum dx, red3[c..2c) += invstd3 * (sum dx*x3 - mean3 * sum dx), likewise (xd, meand, invstdd, redd) when
 * xd != NULL.  The caller then runs b200seg_bn_bwd_apply_masked only (no b200seg_bn_bwd_reduce_*) for those layers.
 * red3 / redd must be zeroed (or carry a partial sum).  dy optional. */
int b200seg_grad_join_reduce(void* dx, int dx_c_total, int dx_c_offset, const void* dy, int dy_c_total, int dy_c_offset,
                             const uint8_t* xin_bits, int c, int64_t pixels, const void* x3, int x3_c_total, int x3_c_offset,
                             const float* mean3, const float* invstd3, float* red3, const void* xd, int xd_c_total,
                             int xd_c_offset, const float* meand, const float* invstdd, float* redd, b200seg_stream_t st);
/* y = relu?(x0 + ... + x_{n-1}), 1 <= n <= 4 bf16 NHWC slices: the multi-resolution fuse of an HRModule
 * (models/hrnet.py:240-249) in one pass. */
int b200seg_add_n_relu(int n_in, const void* const* x, const int* x_c_total, const int* x_c_offset, void* y, int y_c_total,
                       int y_c_offset, int c, int64_t pixels, int relu, b200seg_stream_t st);
/* dx = dy * [y > 0]: the gradient through a ReLU whose output y was stored (fuse ReLU, residual-block ReLU). */
int b200seg_relu_bwd(const void* dy, int dy_c_total, int dy_c_offset, const void* y, int y_c_total, int y_c_offset,
                     void* dx, int dx_c_total, int dx_c_offset, int c, int64_t pixels, b200seg_stream_t st);
/* eval mode (test.py:16, inference.py:24): scale/shift from the running statistics; feeds the conv epilogue so
 * Conv+BN+ReLU is one kernel.  save_mean/save_invstd (optional) receive the running mean and 1/sqrt(var+eps). */
int b200seg_bn_eval_affine(int c, const float* gamma, const float* beta, const float* running_mean,
                           const float* running_var, float eps, float* scale, float* shift, float* save_mean,
                           float* save_invstd, b200seg_stream_t st);
/* Backward of y = relu?(gamma*(x-mean)*invstd + beta): pass 1 reduces sum(dz), sum(dz*xhat) into red[2*c]
 * (dz = dy masked by y > 0), pass 2 writes dx and dgamma/dbeta.  training = 0 treats the statistics as
 * constants (eval-mode BN): dx = scale * dz. */
int b200seg_bn_bwd_reduce(const void* dy, int dy_c_total, int dy_c_offset, const void* x, int x_c_total,
                          int x_c_offset, int c, int64_t pixels, const float* scale, const float* shift,
                          const float* save_mean, const float* save_invstd, int relu, float* red,
                          b200seg_stream_t st);
int b200seg_bn_bwd_apply(const void* dy, int dy_c_total, int dy_c_offset, const void* x, int x_c_total,
                         int x_c_offset, void* dx, int dx_c_total, int dx_c_offset, int c, int64_t pixels,
                         const float* scale, const float* shift, const float* save_mean, const float* save_invstd,
                         int relu, int training, const float* red, float* dgamma, float* dbeta,
                         int accumulate_param_grads, b200seg_stream_t st);

/* Same two passes with the ReLU mask taken from a stored tensor (mask > 0) instead of recomputed from x: after a
 * residual add the sign of the block output is not a function of this BN's input alone. */
int b200seg_bn_bwd_reduce_masked(const void* dy, int dy_c_total, int dy_c_offset, const void* x, int x_c_total,
                                 int x_c_offset, int c, int64_t pixels, const float* scale, const float* shift,
                                 const float* save_mean, const float* save_invstd, int relu, float* red, const void* mask,
                                 int mask_c_total, int mask_c_offset, b200seg_stream_t st);
/* Same, ACCUMULATING into a `red` the caller zeroed (one memset for all BatchNorms of a backward pass instead of one
 * per layer: a memset node between two kernels also breaks their programmatic-dependent-launch overlap). */
int b200seg_bn_bwd_reduce_acc(const void* dy, int dy_c_total, int dy_c_offset, const void* x, int x_c_total, int x_c_offset,
                              int c, int64_t pixels, const float* scale, const float* shift, const float* save_mean,
                              const float* save_invstd, int relu, float* red, const void* mask, int mask_c_total,
                              int mask_c_offset, b200seg_stream_t st);
int b200seg_bn_bwd_apply_masked(const void* dy, int dy_c_total, int dy_c_offset, const void* x, int x_c_total,
                                int x_c_offset, void* dx, int dx_c_total, int dx_c_offset, int c, int64_t pixels,
                                const float* scale, const float* shift, const float* save_mean, const float* save_invstd,
                                int relu, int training, const float* red, float* dgamma, float* dbeta,
                                int accumulate_param_grads, const void* mask, int mask_c_total, int mask_c_offset,
                                b200seg_stream_t st);

/* ---- resampling ------------------------------------------------------------------------------------------ */
/* Bilinear resize of an NHWC bf16 slice into another slice (F.interpolate(..., mode='bilinear') +
 * torch.cat placement: models/deeplabv3plus.py:34-38, models/unet.py:30-50). */
int b200seg_upsample_fwd(const void* x, int x_c_total, int x_c_offset, int n, int h, int w, void* y, int y_c_total,
                         int y_c_offset, int oh, int ow, int c, int align_corners, b200seg_stream_t st);
/* Its adjoint: dx[src] = sum over the destination pixels that read it (gather form, no atomics). */
int b200seg_upsample_bwd(const void* dy, int dy_c_total, int dy_c_offset, int n, int oh, int ow, void* dx,
                         int dx_c_total, int dx_c_offset, int h, int w, int c, int align_corners,
                         b200seg_stream_t st);
/* nn.MaxPool2d(3, stride 2, padding 1) of the ResNet stem (pytorch_modules.backbones.resnet50, called at
 * models/deeplabv3plus.py:17-19,29) on NHWC bf16: (h x w) -> ((h+1)/2 x (w+1)/2).  idx [n][oh][ow][c] (one byte per
 * element) receives the window position of each maximum for the gather-form backward. */
int b200seg_maxpool3x3s2_fwd(const void* x, int x_c_total, int x_c_offset, int n, int h, int w, void* y, int y_c_total,
                             int y_c_offset, uint8_t* idx, int c, b200seg_stream_t st);
int b200seg_maxpool3x3s2_bwd(const void* dy, int dy_c_total, int dy_c_offset, const uint8_t* idx, int n, int h, int w,
                             void* dx, int dx_c_total, int dx_c_offset, int c, b200seg_stream_t st);
/* nn.AdaptiveAvgPool2d(1) (models/aspp.py:11): mean over H*W of an NHWC bf16 slice -> fp32 [n][c]
 * (the H*W range is split across CTAs; fp32 atomics into y). */
int b200seg_gap_fwd(const void* x, int x_c_total, int x_c_offset, int n, int hw, int c, float* y, b200seg_stream_t st);
/* Broadcast a per-image fp32 vector [n][c] over H*W into a bf16 slice: the 1x1 -> HxW bilinear
 * (align_corners=False) of models/aspp.py:16-19, and (with scale = 1/HW, accumulate) the backward of the
 * average pool. */
int b200seg_broadcast_hw(const float* v, int n, int c, float scale, void* y, int y_c_total, int y_c_offset, int hw,
                         int accumulate, b200seg_stream_t st);
/* Sum an NHWC bf16 slice over H*W -> fp32 [n][c] (backward of the broadcast above). */
int b200seg_sum_hw(const void* x, int x_c_total, int x_c_offset, int n, int hw, int c, float* y, b200seg_stream_t st);

/* ---- the image-pooling branch after the pool: fp32 CUDA-core kernels on [n][c] rows ----------------------- */
/* 1x1 ConvNormAct on the N x C x 1 x 1 pooled map (models/aspp.py:12): y[n][k] = sum_c x[n][c] * w[k][c] with w the
 * fp32 OIHW master weight itself; its two gradients; BatchNorm(+ReLU) over the n rows and its backward. */
int b200seg_linear_f32_fwd(const float* x, const float* w, float* y, int n, int c, int k, b200seg_stream_t st);
int b200seg_linear_f32_bwd_data(const float* dy, const float* w, float* dx, int n, int c, int k, b200seg_stream_t st);
int b200seg_linear_f32_bwd_weight(const float* dy, const float* x, float* dw, int n, int c, int k, int accumulate,
                                  b200seg_stream_t st);
int b200seg_bn_rows_fwd(const float* x, int n, int k, const float* gamma, const float* beta, float eps, float momentum,
                        float* running_mean, float* running_var, int training, int relu, float* y, float* save_mean,
                        float* save_invstd, b200seg_stream_t st);
int b200seg_bn_rows_bwd(const float* dy, const float* x, int n, int k, const float* gamma, const float* beta,
                        const float* save_mean, const float* save_invstd, int training, int relu, float* dx, float* dgamma,
                        float* dbeta, b200seg_stream_t st);
/* Per-channel sum over all pixels of a bf16 NHWC slice -> fp32 [c] (bias gradient of cls_conv). */
int b200seg_channel_sum(const void* x, int x_c_total, int x_c_offset, int c, int64_t pixels, float* out,
                        int accumulate, b200seg_stream_t st);

/* ---- logits, loss, metrics ------------------------------------------------------------------------------- */
/* Final bilinear of the class map, NHWC fp32 [n][h][w][c_total] -> NCHW fp32 [n][nc][oh][ow]
 * (models/deeplabv3plus.py:40-43, models/unet.py:52-55; align_corners=True there). */
int b200seg_logits_upsample_fwd(const float* x, int c_total, int n, int h, int w, int nc, float* y, int oh, int ow,
                                int align_corners, b200seg_stream_t st);
/* Adjoint: NCHW fp32 gradient -> NHWC bf16 [n][h][w][c_total] (pad channels zero; c_total <= 64). */
int b200seg_logits_upsample_bwd(const float* dy, int n, int nc, int oh, int ow, void* dx, int c_total, int h, int w,
                                int align_corners, b200seg_stream_t st);
/* Bilinear resize of `planes` = N*C fp32 planes and its adjoint: the F.interpolate of utils/utils.py:18-20 when
 * logits and targets differ in size (identity, and skipped by the caller, when they match). */
int b200seg_resize_nchw_fwd(const float* x, int64_t planes, int h, int w, float* y, int oh, int ow, int align_corners,
                            b200seg_stream_t st);
int b200seg_resize_nchw_bwd(const float* dy, int64_t planes, int oh, int ow, float* dx, int h, int w, int align_corners,
                            b200seg_stream_t st);
/* Measurement aids (bench.py roofline.flops_executed): pipeline iterations of the launches made by this thread's last
 * b200seg_conv_fprop / _dgrad call — run, and what the same tiling runs with no all-padding (tap, tile) skipped; the
 * same pair for a weight-gradient descriptor.  nominal FLOPs x run / dense = FLOPs the tensor pipe was asked for. */
int b200seg_conv_last_plan(int64_t* iters_run, int64_t* iters_dense);
int b200seg_conv_wgrad_plan_info(const b200seg_wgrad_desc* d, int64_t* iters_run, int64_t* iters_dense);

/* Target masks: int64 is what the reference's post_fetch_fn hands on (``segs.long()``, utils/datasets.py:213); uint8 is
 * what BasicDataset.__getitem__ yields before that cast (utils/datasets.py:161-194) and int32 a common intermediate —
 * every function below that takes ``targets`` reads any of the three directly (target_kind), so the 8x wider copy need
 * never exist (33.5 MB -> 4.2 MB per 16 x 512 x 512 batch, host -> device transfer included). */
#define B200SEG_TARGET_I64 0
#define B200SEG_TARGET_I32 1
#define B200SEG_TARGET_U8 2

/* nn.CrossEntropyLoss() of utils/utils.py:12,21 on NCHW fp32 logits and integer targets: accumulates
 * sum_pixels -log softmax[target] into *loss_sum (fp64; caller zeroes it and divides by the pixel count) and
 * writes dlogits = (softmax - onehot) * grad_scale when dlogits != NULL.  Targets outside [0, nc) other than
 * ignore_index set *status = 1. */
int b200seg_softmax_ce(const float* logits, const void* targets, int target_kind, int n, int nc, int64_t hw, int64_t ignore_index,
                       float grad_scale, double* loss_sum, int64_t* valid_count, float* dlogits, int* status,
                       b200seg_stream_t st);
/* Cross-entropy straight from the head's LOW-RESOLUTION fp32 NHWC class map [n][h][w][c_total] (before the final bilinear
 * of models/deeplabv3plus.py:40-43, unet.py:52-55, hrnet.py:400-404): the th x tw logits are interpolated on the fly
 * (same arithmetic as b200seg_logits_upsample_fwd).  _fwd accumulates like b200seg_softmax_ce; _bwd writes the gradient
 * of the class map itself (fp32, zeroed here) = upsample^T((softmax - onehot) * *grad_out / *valid_count) without the
 * full-resolution gradient ever existing in HBM.  th >= h, tw >= w, nc <= 32. */
int b200seg_lowres_ce_fwd(const float* lr, int n, int h, int w, int c_total, int nc, const void* targets, int target_kind, int th, int tw,
                          int align_corners, int64_t ignore_index, double* loss_sum, int64_t* valid_count, int* status,
                          b200seg_stream_t st);
int b200seg_lowres_ce_bwd(const float* lr, int n, int h, int w, int c_total, int nc, const void* targets, int target_kind, int th, int tw,
                          int align_corners, int64_t ignore_index, const float* grad_out, const int64_t* valid_count,
                          float* dlr, b200seg_stream_t st);
/* Training: both of the above in ONE pass over the target tiles (three interpolations of the logits per class instead
 * of five).  The mean's 1/count is not known until every tile is done, so dlr receives the UNSCALED gradient
 * upsample^T(softmax - onehot); b200seg_scale_by_ratio then multiplies it by *grad_out / *valid_count in place (both read
 * on the device).  Accumulators as b200seg_lowres_ce_fwd (zeroed by the caller); dlr is zeroed here. */
int b200seg_lowres_ce_fused(const float* lr, int n, int h, int w, int c_total, int nc, const void* targets, int target_kind,
                            int th, int tw, int align_corners, int64_t ignore_index, double* loss_sum, int64_t* valid_count,
                            int* status, float* dlr, b200seg_stream_t st);
int b200seg_scale_by_ratio(float* y, int64_t count, const float* grad_out, const int64_t* valid_count, b200seg_stream_t st);

/* out[i] = bf16(a[i] + b[i]) (b bf16, may be NULL), count % 4 == 0: the fp32 class-map gradient of
 * b200seg_lowres_ce_bwd as the bf16 NHWC slice b200seg_conv_dgrad / _wgrad consume. */
int b200seg_cast_add_bf16(const float* a, const void* b, void* out, int64_t count, b200seg_stream_t st);
/* *loss = (float)(*loss_sum / *valid_count), NaN when *status != 0: the scalar tail of the mean reduction, on the device. */
int b200seg_ce_finalize(const double* loss_sum, const int64_t* valid_count, const int* status, float* loss,
                        b200seg_stream_t st);
/* Backward of the mean cross-entropy (loss.backward() of train.py:64 reaching utils/utils.py:21): recomputes the
 * softmax from the saved logits and writes dlogits = (softmax - onehot) * (*grad_out) / (*valid_count), both scalars
 * read on the device (no host sync, no separate scaling pass); ignored pixels get 0. */
int b200seg_softmax_ce_bwd(const float* logits, const void* targets, int target_kind, int n, int nc, int64_t hw, int64_t ignore_index,
                           const float* grad_out, const int64_t* valid_count, float* dlogits, b200seg_stream_t st);
/* outputs.max(1)[1] (test.py:31): NCHW fp32 -> int64 [n][hw], lowest index wins ties. */
int b200seg_argmax(const float* logits, int n, int nc, int64_t hw, int64_t* pred, b200seg_stream_t st);
/* cm[t*nc + p] += 1 over all pixels — superset of the tp/fn/fp loop at test.py:34-46 (tp = diag,
 * fn = row sum - diag, fp = column sum - diag).  Labels outside [0, nc) are skipped. */
int b200seg_confusion(const int64_t* pred, const void* targets, int target_kind, int64_t count, int nc, int64_t* cm,
                      b200seg_stream_t st);
/* argmax + confusion in one pass over the logits (never materialises the prediction). */
int b200seg_argmax_confusion(const float* logits, const void* targets, int target_kind, int n, int nc, int64_t hw, int64_t* cm,
                             b200seg_stream_t st);
/* Evaluation from the head's LOW-RESOLUTION class map (test.py:28-46 applied to models/deeplabv3plus.py:40-43's output):
 * prediction = argmax_c bilinear(lr)[c] at every one of the th x tw target pixels, cm[t*nc + p] += 1 — without reading
 * (or needing) the full-resolution logits.  The interpolation is the instruction sequence b200seg_logits_upsample_fwd
 * stores, so the result is bit-identical to b200seg_argmax_confusion over those logits.  Ties: lowest class index (the
 * first maximum, like outputs.max(1)[1]); the first NaN beats everything.  pred (int64 [n][th][tw]) and cm/targets
 * are each optional (one of them must be given). */
int b200seg_class_map_confusion(const float* lr, int n, int h, int w, int c_total, int nc, const void* targets,
                                int target_kind, int th, int tw, int align_corners, int64_t* cm, int64_t* pred,
                                b200seg_stream_t st);

/* ---- input side: the reference's post_fetch_fn (utils/datasets.py:199-213) on the device --------------
 * imgs.float(); imgs -= mean; imgs /= std  with uint8 images as BasicDataset yields them (utils/datasets.py:161-194).
 * layout: 0 = NCHW uint8 (the reference's batches), 1 = NHWC uint8 (what cv2 decodes to).  mean / std: c host floats.
 * Arithmetic is (float(x) - mean) / std with IEEE subtraction and division: bit-identical to the reference's fp32 ops. */
/* -> fp32 NCHW, exactly the tensor post_fetch_fn returns (any model / backbone) */
int b200seg_input_u8_normalize(const uint8_t* x, int layout, int n, int c, int h, int w, const float* mean, const float* std,
                               float* y, b200seg_stream_t st);
/* -> the bf16 space-to-depth operand of a stride-2 r x s stem conv (see b200seg_space_to_depth_nchw): normalisation,
 * fp32 -> bf16 rounding, NCHW -> NHWC and the 2x2 space-to-depth in ONE pass over the 1-byte pixels (12.6 MB in, 34 MB
 * out for 16 x 3 x 512 x 512 instead of an 50 MB fp32 intermediate written and read back).  Padding lanes hold 0. */
int b200seg_input_u8_s2d(const uint8_t* x, int layout, int n, int c, int h, int w, int r, int s, int pad, const float* mean,
                         const float* std, void* z, int cs, b200seg_stream_t st);

/* multi_scale of post_fetch_fn (utils/datasets.py:206-212): F.interpolate(imgs, (oh, ow)) — mode 'nearest' — applied to
 * the 1-byte planes BEFORE the normalisation (nearest sampling commutes with a per-pixel map, so the float result is the
 * reference's bit for bit).  x: [planes][h][w] uint8 (planes = n*c for NCHW images, n for masks), y: [planes][oh][ow].
 * Source index = min(int(floorf(dst * float(in)/out)), in - 1), ATen's fp32 arithmetic. */
int b200seg_resize_nearest_u8(const uint8_t* x, int64_t planes, int h, int w, uint8_t* y, int oh, int ow,
                              b200seg_stream_t st);

#ifdef __cplusplus
}
#endif
#endif /* B200SEG_H */

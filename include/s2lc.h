/* s2lc.h — C ABI of libs2lc.so: the B200-native (sm_100a) EfficientNet-UNet training-step kernels.
 *
 * Drop-in boundary for the hot path of MaxWolf-01/sentinel2-landcover-classification.  The
 * reference has NO native interface for this path (it is 100 % Python over PyTorch ops), so each
 * entry point below cites the reference *call site* it replaces (paths relative to the reference
 * repo root).  INTEGRATION.md shows the ctypes binding a maintainer adds on the reference side.
 *
 * Conventions
 *  - every pointer is a DEVICE pointer owned by the caller; nothing is retained across calls;
 *  - activations are NHWC ("rows" = N*H*W pixels, channels innermost) with an explicit row stride
 *    `ld` in elements, so channel slices of a concat buffer are addressable without copies;
 *  - `dtype`: storage type of activations and packed GEMM weights — S2LC_F32 (verification mode) or
 *    S2LC_BF16 (production mode).  Parameters, statistics and gradients of parameters are fp32;
 *    BatchNorm moment accumulators are fp64;
 *  - all launches are asynchronous on `stream` (a cudaStream_t passed as void*); no host sync;
 *  - return value: 0 = OK, <0 = S2LC_ERR_*; s2lc_last_error() gives the thread-local message.
 *    There is no CPU fallback and no alternate backend: an unsupported shape/arch is an error.
 */
#ifndef S2LC_H
#define S2LC_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define S2LC_OK 0
#define S2LC_ERR_SHAPE (-1)
#define S2LC_ERR_ALIGN (-2)
#define S2LC_ERR_ARCH (-3)
#define S2LC_ERR_CUDA (-4)

#define S2LC_F32 0
#define S2LC_BF16 1

#define S2LC_ACT_NONE 0
#define S2LC_ACT_SILU 1
#define S2LC_ACT_RELU 2

/* A "lazy" forward activation: value(row, c) = act(raw[row*ld + c] * scale[c] + shift[c]) * gate[n, c].
 * This is how a conv -> BatchNorm(train) -> activation chain of the reference
 * (src/modules/efficientnet_unet.py:191-197, :312-345, :168-176) is fused: the producing conv writes
 * `raw` plus per-channel moments, the CONSUMER applies the normalisation on load.               */
typedef struct s2lc_lazy {
  const void* raw;    /* [rows, >=C] storage dtype                                   */
  int64_t ld;         /* row stride in elements                                       */
  const float* scale; /* [C] or NULL (=1)                                             */
  const float* shift; /* [C] or NULL (=0)                                             */
  const float* gate;  /* [N, C] squeeze-excite gate (efficientnet_unet.py:381) or NULL */
  int32_t act;        /* S2LC_ACT_*                                                   */
  int32_t _pad;
} s2lc_lazy;

/* A "lazy" gradient w.r.t. a raw conv output that was followed by BatchNorm(train):
 *   g(row, c) = p[c] * dy[row, c] * rowscale[n] + q[c] * z[row, c] + t[c]
 * (dy = gradient w.r.t. the BatchNorm output, z = the saved raw conv output; p, q, t come from
 * s2lc_bn_bwd_finalize).  z == NULL means g = dy * rowscale (no BatchNorm behind this tensor).  */
typedef struct s2lc_lazygrad {
  const void* dy;
  int64_t ld_dy;
  const void* z;
  int64_t ld_z;
  const float* p;
  const float* q;
  const float* t;
  const float* rowscale; /* [N] per-sample factor (drop-connect, efficientnet_unet.py:390-398) or NULL */
} s2lc_lazygrad;

/* Implicit-GEMM convolution family (CUDA-core fp32 path for S2LC_F32, tcgen05/TMEM path for S2LC_BF16).
 * out[m, n] = sum_{tap, c} A(m, tap, c) * Wp[n, tap*Cpad + c]  (+ bias, + residuals), m over N*H*W rows of A. */
#define S2LC_CONV_1X1 0       /* F.conv2d 1x1: efficientnet_unet.py:315-320, :364-369, :230-235, :121           */
#define S2LC_CONV_3X3 1       /* nn.Conv2d(k=3, p=1): efficientnet_unet.py:170,173 (and its dgrad)              */
#define S2LC_CONVT_FWD 2      /* nn.ConvTranspose2d(k=2, s=2): efficientnet_unet.py:115,119 — scatter epilogue  */
#define S2LC_CONVT_DGRAD 3    /* its data gradient — gathers the 2x2 children of every input pixel             */

#define S2LC_STATS_NONE 0
#define S2LC_STATS_MOMENTS 1  /* stats[c] += sum v, stats[C+c] += sum v*v              (BatchNorm forward)  */
#define S2LC_STATS_BNBWD 2    /* stats[c] += sum v*rs, stats[C+c] += sum v*rs*zprev    (BatchNorm backward) */

typedef struct s2lc_conv_params {
  int32_t mode;            /* S2LC_CONV_* */
  int32_t dtype;           /* S2LC_F32 / S2LC_BF16 */
  int32_t N, H, W;         /* spatial dims of the A operand (rows = N*H*W) */
  int32_t C;               /* channels of A per tap */
  int32_t Cpad;            /* per-tap K stride in the packed weight (multiple of 8, >= C) */
  int32_t Kpad;            /* padded total K of the packed weight (multiple of 64) */
  int32_t Nout;            /* GEMM N: Cout (1x1, 3x3, convT dgrad) or 4*Cout (convT fwd) */
  int32_t Npad;            /* rows of the packed weight (multiple of 16, >= Nout) */
  int32_t _pad0;
  int32_t a_is_grad;       /* 0: A = *a_fwd, 1: A = *a_bwd */
  const s2lc_lazy* a_fwd;
  const s2lc_lazygrad* a_bwd;
  const void* w_packed;    /* [Npad, Kpad] K-major, storage dtype (from s2lc_weight_pack) */
  const float* bias;       /* [Cout] or NULL */
  void* out;               /* [rows_out, ld_out] storage dtype */
  int64_t ld_out;
  const void* res1;        /* optional residual addends, same shape as out */
  int64_t ld_res1;
  const void* res2;
  int64_t ld_res2;
  const s2lc_lazy* prev;   /* backward epilogue: out *= act'(prev.raw*scale+shift); prev.raw is zprev for BNBWD stats */
  double* stats;           /* [2*Nout] fp64 accumulators (pre-zeroed) or NULL */
  int32_t stats_kind;      /* S2LC_STATS_* */
  int32_t _pad;
  const float* stats_rowscale; /* [N] per-sample factor inside BNBWD stats, or NULL */
} s2lc_conv_params;

typedef struct s2lc_wgrad_params {
  int32_t mode;            /* S2LC_CONV_1X1 / S2LC_CONV_3X3 / S2LC_CONVT_FWD (weight gradient of that op) */
  int32_t dtype;
  int32_t N, H, W;         /* spatial dims of the forward INPUT (rows = N*H*W) */
  int32_t Cin, Cout;
  int32_t _pad;
  const s2lc_lazy* a;      /* forward input */
  const s2lc_lazygrad* g;  /* gradient w.r.t. the raw forward output (at 2H x 2W for convT) */
  float* dw;               /* fp32, REFERENCE layout (OIHW conv / IOHW convT), pre-zeroed, accumulated atomically */
  float* dbias;            /* [Cout] pre-zeroed or NULL */
} s2lc_wgrad_params;

/* ---- library ---------------------------------------------------------------------------------- */
int32_t s2lc_version(void);
const char* s2lc_last_error(void);
int32_t s2lc_arch_check(int32_t device);   /* S2LC_ERR_ARCH unless the device is sm_100 */
int32_t s2lc_memset_zero(void* ptr, int64_t bytes, void* stream);

/* ---- weights: fp32 reference layout -> packed [Npad, Kpad] K-major in the compute dtype --------
 * dst[(n1*N2 + n2), tap*Cpad + c] = src[n1*s1 + n2*s2 + tap'*st + c*sc], tap' = flip ? T-1-tap : tap; zero padded. */
int32_t s2lc_weight_pack(const float* src, void* dst, int32_t dtype, int32_t N1, int32_t N2, int32_t T, int32_t C,
                         int32_t Npad, int32_t Cpad, int32_t Kpad, int64_t s1, int64_t s2, int64_t st, int64_t sc,
                         int32_t flip, void* stream);

/* ---- stem: Conv2dSamePadding(Cin -> Co, k=3, s=2, no bias) on the fp32 NCHW input tile ---------
 * replaces efficientnet_unet.py:191-194 (+ the TF-"same" pad of :288-297); writes raw NHWC + moments */
int32_t s2lc_stem_conv_fwd(const float* x_nchw, const float* w_oihw, void* out, double* stats, int32_t N, int32_t Cin,
                           int32_t H, int32_t W, int32_t Co, int32_t dtype, void* stream);
/* im2col of the stem for the bf16 production path: cols[(n, oy, ox), ci*9 + ty*3 + tx] (zero up to `ld` columns), so
 * that Conv2dSamePadding(k=3, s=2) (efficientnet_unet.py:191-194, :288-297) and its weight gradient run as K = 56
 * contractions on the tensor-core engines (s2lc_conv_fwd / s2lc_conv_wgrad in 1x1 mode on the column buffer). */
int32_t s2lc_stem_im2col(const float* x_nchw, void* cols, int64_t ld, int32_t N, int32_t Cin, int32_t H, int32_t W,
                         int32_t dtype, void* stream);
int32_t s2lc_stem_conv_wgrad(const float* x_nchw, const s2lc_lazygrad* g, float* dw_oihw, int32_t N, int32_t Cin,
                             int32_t H, int32_t W, int32_t Co, int32_t dtype, void* stream);

/* ---- BatchNorm (training-mode batch statistics; nn.BatchNorm2d at efficientnet_unet.py:195,236,321,338,370,171,174) */
int32_t s2lc_bn_finalize(const double* stats, const float* gamma, const float* beta, double count, float eps, float momentum, float* running_mean, float* running_var,
                         float* scale, float* shift, float* mean, float* invstd, int32_t C, void* stream);
int32_t s2lc_bn_eval_affine(const float* gamma, const float* beta, const float* running_mean, const float* running_var,
                            float eps, float* scale, float* shift, int32_t C, void* stream);
/* out = act(z*scale+shift) * rowscale[n] + residual   (BN3 + drop-connect + identity, efficientnet_unet.py:382-386) */
int32_t s2lc_bn_apply(const s2lc_lazy* in, const void* residual, int64_t ld_res, const float* rowscale, void* out,
                      int64_t ld_out, int64_t rows, int32_t C, int32_t HW, int32_t dtype, void* stream);
int32_t s2lc_bn_bwd_reduce(const void* dy, int64_t ld_dy, const void* z, int64_t ld_z, const float* rowscale,
                           double* stats, int64_t rows, int32_t C, int32_t HW, int32_t dtype, void* stream);
int32_t s2lc_bn_bwd_finalize(const double* stats, const float* mean, const float* invstd, const float* gamma,
                             double count, float* p, float* q, float* t, float* dgamma, float* dbeta, int32_t C,
                             void* stream);

/* ---- depthwise conv k in {3,5}, stride in {1,2}, TF-"same" padding (efficientnet_unet.py:329-337) */
int32_t s2lc_dwconv_fwd(const s2lc_lazy* in, const float* w, void* out, double* stats, int32_t N, int32_t H, int32_t W,
                        int32_t C, int32_t k, int32_t s, int32_t dtype, void* stream);
/* d(in) = dwconv^T(g) * act'(prev) (+ residual); BNBWD stats of the result against prev.raw */
int32_t s2lc_dwconv_bwd_data(const s2lc_lazygrad* g, const float* w, const s2lc_lazy* prev, const void* res,
                             int64_t ld_res, void* out, int64_t ld_out, double* stats, int32_t N, int32_t H, int32_t W,
                             int32_t C, int32_t k, int32_t s, int32_t dtype, void* stream);
int32_t s2lc_dwconv_bwd_weight(const s2lc_lazygrad* g, const s2lc_lazy* in, float* dw, int32_t N, int32_t H, int32_t W,
                               int32_t C, int32_t k, int32_t s, int32_t dtype, void* stream);
/* Fused streaming backward (autograd of efficientnet_unet.py:329-337 in one pass): tiled over INPUT pixels,
 * writes d(in) = dwconv^T(g) (* act'(in) when apply_act_bwd) (+ residual), the BNBWD statistics of d(in) against
 * in.raw, and accumulates dW[c,ky,kx] += act(in) * g — g, in.raw and d(in) each cross HBM once.               */
int32_t s2lc_dwconv_bwd(const s2lc_lazygrad* g, const float* w, const s2lc_lazy* in, int32_t apply_act_bwd,
                        const void* res, int64_t ld_res, void* out, int64_t ld_out, double* stats, float* dw,
                        int32_t N, int32_t H, int32_t W, int32_t C, int32_t k, int32_t s, int32_t dtype, void* stream);
/* Tile-engine forward (one smem tile per CTA, no row ring); kept to cross-check the streaming kernel. */
int32_t s2lc_dwconv_fwd_tile(const s2lc_lazy* in, const float* w, void* out, double* stats, int32_t N, int32_t H,
                             int32_t W, int32_t C, int32_t k, int32_t s, int32_t dtype, void* stream);

/* ---- squeeze-excite (efficientnet_unet.py:346-362, :381) ------------------------------------- */
int32_t s2lc_se_pool(const s2lc_lazy* in, float* pooled, int32_t N, int32_t HW, int32_t C, int32_t dtype, void* stream);
int32_t s2lc_se_fc_fwd(const float* pooled, const float* w1, const float* b1, const float* w2, const float* b2,
                       float* gate, int32_t N, int32_t C, int32_t S, void* stream);
/* Allocation-free variant of s2lc_se_fc_fwd.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA-graph capture, tight serving loops).  Same arguments and
 * results; the [N, S] hidden scratch is a static 256 KB device array (held by every GPU that loads the library until
 * exit), so N * S <= 65536 or the call fails with S2LC_ERR_SHAPE.  It calls no cudaMalloc / cudaFree.  All calls share
 * the array: calls that overlap on the GPU (different streams / host threads) race on it. */
int32_t s2lc_se_fc_fwd_noalloc(const float* pooled, const float* w1, const float* b1, const float* w2, const float* b2,
                               float* gate, int32_t N, int32_t C, int32_t S, void* stream);
int32_t s2lc_se_bwd_reduce(const void* da, int64_t ld_da, const s2lc_lazy* in, float* dgate, int32_t N, int32_t HW,
                           int32_t C, int32_t dtype, void* stream);
int32_t s2lc_se_fc_bwd(const float* pooled, const float* w1, const float* b1, const float* w2, const float* b2,
                       const float* dgate, float* dpool, float* dw1, float* db1, float* dw2, float* db2,
                       float* scratch /* N*(C+2S) floats */, int32_t N, int32_t C, int32_t S, void* stream);
/* da <- (da*gate + dpool/HW) * act'(u)  in place, + BNBWD stats against in.raw */
int32_t s2lc_se_bwd_apply(void* da, int64_t ld_da, const s2lc_lazy* in, const float* dpool, double* stats, int32_t N,
                          int32_t HW, int32_t C, int32_t dtype, void* stream);

/* ---- implicit-GEMM convolutions ---------------------------------------------------------------- */
int32_t s2lc_conv_fwd(const s2lc_conv_params* p, void* stream);
int32_t s2lc_conv_wgrad(const s2lc_wgrad_params* p, void* stream);

/* ---- classifier out_conv1x1 (C -> num_classes in 2..8, bias): efficientnet_unet.py:121,137.  Per-pixel kernels:
 * forward  logits = W . in + b;  backward (one pass)  dW += dlogits (x) in, db += dlogits,
 * dz = (W^T dlogits) * act'(in) with the BatchNorm-backward moments of dz accumulated into stats[2*C].  */
int32_t s2lc_outconv_fwd(const s2lc_lazy* in, const float* w, const float* bias, void* logits, int64_t ld_out,
                         int64_t rows, int32_t C, int32_t n_classes, int32_t dtype, void* stream);
int32_t s2lc_outconv_bwd(const s2lc_lazy* in, const void* dlogits, int64_t ld_dl, const float* w, float* dw,
                         float* dbias, void* dz, int64_t ld_dz, double* stats, int64_t rows, int32_t C,
                         int32_t n_classes, int32_t dtype, void* stream);

/* ---- plumbing ---------------------------------------------------------------------------------- */
int32_t s2lc_copy_slice(const void* src, int64_t ld_src, void* dst, int64_t ld_dst, int64_t rows, int32_t C,
                        int32_t dtype, void* stream);
int32_t s2lc_add_slice(const void* a, int64_t ld_a, const void* b, int64_t ld_b, void* dst, int64_t ld_dst,
                       int64_t rows, int32_t C, int32_t dtype, void* stream);
int32_t s2lc_nchw_to_slice(const float* x_nchw, void* dst, int64_t ld_dst, int32_t N, int32_t C, int32_t H, int32_t W,
                           int32_t Cfill, int32_t dtype, void* stream);
int32_t s2lc_dropconnect_masks(float* rowscale, int32_t n_blocks, int32_t batch, float rate, uint64_t seed,
                               uint64_t step, void* stream);

/* ---- loss: FocalLoss / weighted CrossEntropy forward + d(logits) in one pass (src/losses.py:36-40, :77-89) */
#define S2LC_LOSS_FOCAL 0
#define S2LC_LOSS_CE 1
int32_t s2lc_loss_fwd_bwd(const void* logits, int64_t ld, const int64_t* y, const float* class_w, float gamma,
                          float label_smoothing, int64_t ignore_index, int32_t kind, int32_t reduce_mean,
                          float* loss_out, float* wsum_scratch, void* dlogits, int64_t ld_d, int64_t rows,
                          int32_t n_classes, int32_t dtype, void* stream);
/* Materialise a lazy gradient: out[row, c] = p[c]*dy[row, c]*rowscale[row / HW] + q[c]*z[row, c] + t[c], same fp32
 * formula and rounding as the conv engines' operand loaders (autograd of BatchNorm2d behind the decoder 3x3 convs,
 * efficientnet_unet.py:168-176).  `out` may alias g->dy.                                                           */
int32_t s2lc_lazygrad_apply(const s2lc_lazygrad* g, void* out, int64_t ld_out, int64_t rows, int32_t C, int32_t HW,
                            int32_t dtype, void* stream);
int32_t s2lc_scale_inplace(void* x, int64_t ld, const float* factor_dev, int64_t rows, int32_t C, int32_t dtype,
                           void* stream);

/* ---- per-step metrics (next row f2) and the uint8 class map of a mosaic writer (next row f4):
 * train_segmentation.py:145 `update_metrics(predictions=logits.argmax(dim=1), labels=y)` -> torchmetrics confusion
 * matrix / IoU / accuracy / F1 (:53-67).  One pass: counts[target * C + argmax] += 1 over ALL pixels whose target is in
 * [0, C) (int64 [C, C], accumulated — the caller zeroes it; ignore_index is applied on the host by dropping that row)
 * and/or pred[row] = argmax (uint8).  `y`/`counts` or `pred` may be NULL.  First maximum wins, as torch.argmax.     */
int32_t s2lc_argmax_confusion(const void* logits, int64_t ld, const int64_t* y, int64_t* counts, uint8_t* pred,
                              int64_t rows, int32_t n_classes, int32_t dtype, void* stream);

/* ---- input pipeline on the device (next row f3): s2osm_dataset.py:51-71 + s2osm_datamodule.py:75-87 +
 * cnes_labell_mappings.py:85-95.  raw int16 [N, C, Hs, Ws] tiles and uint8 [N, Hs, Ws] labels -> crop (crop_yx[n] =
 * top, left; bounds are the caller's contract: 0 <= top <= Hs-crop) -> horizontal / vertical flip (flip_hv[n], may be
 * NULL) -> x = (raw - mean*mpv) * (1/(std*mpv)) fp32 [N, C, crop, crop]; y = lut[label] (lut: 256 x int64 or NULL =
 * identity) int64 [N, crop, crop].  x or y may be NULL to skip that half.                                          */
int32_t s2lc_prepare_batch(const int16_t* raw, const uint8_t* labels, const int32_t* crop_yx, const uint8_t* flip_hv,
                           const float* mean, const float* stdv, const int64_t* lut, float* x, int64_t* y, int32_t N,
                           int32_t C, int32_t Hs, int32_t Ws, int32_t crop, float max_pixel_value, void* stream);

/* ---- mosaic inference writer (next row f4): replaces experiments/inference_demo.py:14-41 (Lightning predict +
 * torch.save of whole logit tensors, argmax later on the CPU) and train_segmentation.py:106-107 predict_step.
 * s2lc_gather_tiles: B tiles of tile x tile px cut from ONE int16 scene [C, Hs, Ws] at origin_yx[b] = (top, left)
 * (may overhang: outside pixels replicate the nearest edge) and normalised as in s2lc_prepare_batch -> x fp32
 * [B, C, tile, tile].  s2lc_argmax_mosaic: argmax over the class axis of NHWC logits [B, tile, tile, ld >= n_classes]
 * written as uint8 into mosaic[Hm, ld_mosaic] at origin_yx[b] + (ky, kx) for the tile centre margin <= ky, kx <
 * tile - margin; pixels landing outside the mosaic are dropped.                                                     */
int32_t s2lc_gather_tiles(const int16_t* scene, const int32_t* origin_yx, const float* mean, const float* stdv,
                          float* x, int32_t B, int32_t C, int32_t Hs, int32_t Ws, int32_t tile, float max_pixel_value,
                          void* stream);
int32_t s2lc_argmax_mosaic(const void* logits, int64_t ld, const int32_t* origin_yx, uint8_t* mosaic,
                           int64_t ld_mosaic, int32_t B, int32_t tile, int32_t margin, int32_t Hm, int32_t Wm,
                           int32_t n_classes, int32_t dtype, void* stream);

/* ---- optimizer (next row f1): torch.optim.Adam with L2 weight decay, train_segmentation.py:110-114.
 * Hyper-parameters are doubles, as torch holds them: 1 - beta2 formed in fp32 from 0.999f is off by 1.3e-5. */
int32_t s2lc_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, double lr,
                       double beta1, double beta2, double eps, double weight_decay, int64_t step, float grad_scale,
                       void* stream);

/* Adam + re-pack of the tensor-core weight operands in ONE pass over the flat parameter buffer (row f1 as SURVEY.md
 * §8 states it: "fused multi-tensor Adam + bf16 weight re-pack in one pass").  `entries` is a DEVICE table sorted by
 * `start`; entry e says: the parameter stored at param[start .. start+numel) (reference layout OIHW / IOHW) has a
 * packed copy dst[Npad, Kpad] (layout of s2lc_weight_pack) whose element (n1*N2 + n2, tap*Cpad + c) is the source
 * element n1*s1 + n2*s2 + tap*st + c*sc; stride[]/role[] list those four source strides in DESCENDING order with
 * role 0 = n1, 1 = n2, 2 = tap, 3 = c (stride 0 = extent 1).  Padding of dst is never written (zero from creation). */
typedef struct {
  int64_t start, numel;
  void* dst;
  int64_t stride[4];
  int32_t role[4];
  int32_t N2, Tt, Cpad, Kpad, flip, dtype;
} s2lc_pack_entry;
int32_t s2lc_adam_pack_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, double lr,
                            double beta1, double beta2, double eps, double weight_decay, int64_t step, float grad_scale,
                            const s2lc_pack_entry* entries, int32_t n_entries, void* stream);

/* ---- diagnostics.  s2lc_simt_fallbacks: number of bf16 conv / weight-gradient calls since load that no tcgen05 engine
 * took and that ran on the CUDA-core verification engine instead (include/s2lc.h top: the product path is tcgen05;
 * the count must stay 0 on the reference architectures — tests and bench.py assert / report it).  reset = 1 clears. */
int64_t s2lc_simt_fallbacks(int32_t reset);
/* Number of launches of the TMA-fed tcgen05 engine (csrc/tconv_tc.cu) since load; reset = 1 clears.  Tests use it to
 * prove that a shape went through TMA rather than through one of the loader-warp engines. */
int64_t s2lc_tma_launches(int32_t reset);
/* SM count of the current device (sizes every persistent grid; read once from the driver, not a constant). */
int32_t s2lc_num_sms(void);
/* Test / profiling switches (bit 0: no streaming engines, 1: force the CUDA-core engine, 2: 7 producer warps,
 * 3: no TMA feed, 4: TMA engines take every plain shape, 5: 32-channel depthwise blocks everywhere, 6: no TMA-store
 * epilogue, 7: generic depthwise kernels instead of the fast ones, 16: generic squeeze-excite backward kernel instead of
 * the packed one; bits 8-15: strip-width cap of the streaming 3x3 weight gradient, from S2LC_SWGRAD_LMAX).  Returns the previous value.
 * Initialised once from S2LC_* environment variables; the default 0 is the product path. */
uint32_t s2lc_set_debug_flags(uint32_t flags);

#ifdef __cplusplus
}
#endif
#endif /* S2LC_H */

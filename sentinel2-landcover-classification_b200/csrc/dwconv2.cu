// dwconv2.cu — streaming depthwise k x k convolution (k in {3,5}, stride in {1,2}, TF-"same" padding):
// forward, and ONE fused backward kernel (data gradient + weight gradient) per MBConv block.
// Reference call site: src/modules/efficientnet_unet.py:329-344 (depthwise conv -> BN -> SiLU) and its autograd.
//
// HBM-bound design:
//  * NHWC; a CTA owns a strip of TW destination pixels x 64 channels and MARCHES DOWN the image in steps of RY
//    rows.  Source rows live in a shared-memory ring (fp32, already transformed: BatchNorm scale/shift + SiLU for
//    the forward, p*dy+q*z+t for the gradient), so every source row is fetched once per strip (no y-halo re-reads)
//    and the transform (MUFU) is evaluated once per element;
//  * the raw 16 B global loads of step t+1 are issued into registers BEFORE the stencil of step t runs and are
//    transformed + stored to the ring after it, so HBM latency overlaps the FMA phase inside one CTA;
//  * lane l owns channel pair (2l, 2l+1): ring reads are 8 B/lane (conflict-free), the math is packed fp32
//    (FFMA2, fma.rn.f32x2), each ring row is read once per warp and applied to all RY x R outputs it touches;
//  * the fused backward tiles over INPUT pixels i: dx[i] = sum_k w[k] g[(i+pad-k)/s] and
//    dW[k] += act(in)[i] * g[(i+pad-k)/s] use the SAME staged gradient value, so g (2 tensors), the saved raw
//    input and dx each cross HBM once (2|D|+2|E| bytes instead of 4|D|+3|E| for two separate kernels);
//  * per-channel BatchNorm moments / dW partials stay in registers across all strips a CTA processes
//    (grid-stride loop) and are flushed once per CTA (smem atomics -> fp64 / fp32 global atomics).
#include "dwconv2.cuh"

using namespace s2lc;

// geometry choice per (K, S, width): R = pixels per warp (strip = 8R pixels wide), RY = rows per step, D = cp.async
// passes in flight; sized so that two CTAs fit an SM (shared memory <= 113 KB) wherever registers allow it.
#define DW2_F(T, K_, S_, R_, RY_, D_, MB, CB_) \
  return dw2::launch_fwd<T, K_, S_, R_, RY_, D_, MB, CB_>(in, w, out, stats, N, H, W, C, st)
// `half`: the last 64-channel block would be at most half full (C = 24, 144, ...) and rows are >= 64 px wide: 32-channel
// blocks with two pixel groups per warp
#define DW2_FWD_BF16                                     \
  if (s == 1) {                                          \
    if (wide) {                                          \
      if (k == 3) { if (half) DW2_F(s2lc::bf16, 3, 1, 4, 4, 2, 2, 32); DW2_F(s2lc::bf16, 3, 1, 4, 4, 2, 2, 64); } \
      if (half5) DW2_F(s2lc::bf16, 5, 1, 4, 2, 2, 2, 32); \
      DW2_F(s2lc::bf16, 5, 1, 4, 2, 2, 2, 64);           \
    }                                                    \
    if (k == 3) DW2_F(s2lc::bf16, 3, 1, 2, 4, 3, 2, 64); \
    DW2_F(s2lc::bf16, 5, 1, 2, 4, 3, 2, 64);             \
  }                                                      \
  if (k == 3) { if (half) DW2_F(s2lc::bf16, 3, 2, 2, 2, 2, 2, 32); DW2_F(s2lc::bf16, 3, 2, 2, 2, 2, 2, 64); } \
  DW2_F(s2lc::bf16, 5, 2, 2, 1, 3, 2, 64);
#define DW2_FWD_F32                                      \
  if (s == 1) {                                          \
    if (k == 3) DW2_F(float, 3, 1, 2, 4, 1, 1, 64);      \
    DW2_F(float, 5, 1, 2, 4, 1, 1, 64);                  \
  }                                                      \
  if (k == 3) DW2_F(float, 3, 2, 2, 2, 1, 1, 64);        \
  DW2_F(float, 5, 2, 2, 1, 1, 1, 64);

// fast kernels (dw2_fwdx_kernel): bf16, exact strip geometry, input = BatchNorm + SiLU (fast == 1) or plain (fast == 2)
#define DW2_FX(K_, S_, R_, RY_, D_, MB, CB_)                                                                    \
  {                                                                                                             \
    int32_t r_ = fast == 1 ? dw2::launch_fwdx<K_, S_, R_, RY_, D_, MB, CB_, true>(in, w, out, stats, N, H, W, C, st)  \
                           : dw2::launch_fwdx<K_, S_, R_, RY_, D_, MB, CB_, false>(in, w, out, stats, N, H, W, C, st); \
    if (r_ <= 0) return r_;                                                                                     \
  }
#define DW2_FWD_FAST                                     \
  if (s == 1) {                                          \
    if (k == 3) { if (half) DW2_FX(3, 1, 4, 4, 2, 2, 32) else DW2_FX(3, 1, 4, 4, 2, 2, 64) } \
    else { if (half5) DW2_FX(5, 1, 4, 2, 2, 2, 32) else DW2_FX(5, 1, 4, 2, 2, 2, 64) }       \
  } else {                                               \
    if (k == 3) { if (half) DW2_FX(3, 2, 2, 2, 2, 2, 32) else DW2_FX(3, 2, 2, 2, 2, 2, 64) } \
    else DW2_FX(5, 2, 2, 1, 3, 2, 64)                    \
  }

extern "C" {

int32_t s2lc_dwconv_fwd(const s2lc_lazy* in, const float* w, void* out, double* stats, int32_t N, int32_t H, int32_t W,
                        int32_t C, int32_t k, int32_t s, int32_t dtype, void* stream) {
  S2LC_REQUIRE((k == 3 || k == 5) && (s == 1 || s == 2), S2LC_ERR_SHAPE, "dwconv: k=%d s=%d unsupported", k, s);
  S2LC_REQUIRE(C % 8 == 0 && in->ld % 8 == 0, S2LC_ERR_ALIGN, "dwconv: C and strides must be multiples of 8 (C=%d)", C);
  S2LC_REQUIRE(in->gate == nullptr, S2LC_ERR_SHAPE, "dwconv: gated input is not a reference configuration");
  cudaStream_t st = (cudaStream_t)stream;
  const bool wide = W > 16;
  // 32-channel blocks: last 64-block at most half full, or rows >= 128 px (two pixel groups per warp are 5-10 % faster there)
  const bool half = W >= 64 && (((C + 63) % 64) < 32 || W >= 128 || (s2lc::debug_flags() & s2lc::DBG_DW_CB32));
  const bool half5 = W >= 64;      // 5x5: 32-channel blocks are 5-8 % faster at 64 px rows (two pixel groups per warp)
  int fast = 0;
  if (dtype == S2LC_BF16 && wide && !(s2lc::debug_flags() & s2lc::DBG_DW_GENERIC)) {
    if (in->act == S2LC_ACT_SILU && in->scale && in->shift) fast = 1;
    if (in->act == S2LC_ACT_NONE && !in->scale && !in->shift) fast = 2;
  }
  if (fast) { DW2_FWD_FAST }
  if (dtype == S2LC_BF16) { DW2_FWD_BF16 }
  if (dtype == S2LC_F32) { DW2_FWD_F32 }
  s2lc::set_error("dwconv: bad dtype %d", dtype);
  return S2LC_ERR_SHAPE;
}

}  // extern "C"

// dwconv2_bwd.cu — C ABI of the fused depthwise backward (kernels: dwconv2.cuh; design notes: dwconv2.cu).
#include "dwconv2.cuh"

using namespace s2lc;

extern "C" {

#define DW2_B(T, K_, S_, R_, RY_, D_, MB, CB_)                                                                          \
  return dw2::launch_bwd<T, K_, S_, R_, RY_, D_, MB, CB_>(g, w, in, apply_act_bwd, res, ld_res, out, ld_out, stats, dw, \
                                                          N, H, W, C, st)
#define DW2_BWD_BF16                                     \
  if (s == 1) {                                          \
    if (wide) {                                          \
      if (k == 3) { if (half) DW2_B(s2lc::bf16, 3, 1, 4, 2, 2, 2, 32); DW2_B(s2lc::bf16, 3, 1, 4, 2, 2, 2, 64); } \
      if (half5) DW2_B(s2lc::bf16, 5, 1, 4, 2, 4, 1, 32); \
      DW2_B(s2lc::bf16, 5, 1, 4, 2, 4, 1, 64);           \
    }                                                    \
    if (k == 3) DW2_B(s2lc::bf16, 3, 1, 2, 4, 2, 2, 64); \
    DW2_B(s2lc::bf16, 5, 1, 2, 2, 4, 1, 64);             \
  }                                                      \
  if (wide) {                                            \
    if (k == 3) { if (half) DW2_B(s2lc::bf16, 3, 2, 4, 4, 2, 2, 32); DW2_B(s2lc::bf16, 3, 2, 4, 4, 2, 2, 64); } \
    DW2_B(s2lc::bf16, 5, 2, 4, 2, 4, 1, 64);             \
  }                                                      \
  if (k == 3) DW2_B(s2lc::bf16, 3, 2, 2, 4, 2, 2, 64);   \
  DW2_B(s2lc::bf16, 5, 2, 2, 2, 4, 1, 64);
// fast kernels (dw2_bwdx_kernel): bf16, exact strip geometry, input = BatchNorm + SiLU with act' applied and no residual
// (fast == 1) or a plain stored tensor without act' (fast == 2); returns 1 when the geometry does not divide
#define DW2_X(K_, S_, R_, RY_, D_, MB, CB_, OB, TP)                                                                      \
  {                                                                                                                      \
    int32_t r_ = fast == 1 ? dw2::launch_bwdx<K_, S_, R_, RY_, D_, MB, CB_, true, OB, TP>(                               \
                                 g, w, in, res, ld_res, out, ld_out, stats, dw, N, H, W, C, st)                          \
                           : dw2::launch_bwdx<K_, S_, R_, RY_, D_, MB, CB_, false, OB, TP>(                              \
                                 g, w, in, res, ld_res, out, ld_out, stats, dw, N, H, W, C, st);                         \
    if (r_ <= 0) return r_;                                                                                              \
  }
// measured (profiles/r02p_dw_variants.txt): two sweeps (one tanh per value, lower register pressure) win on the 3x3
// kernels that run two CTAs per SM — 1.7x on stride 2, where the single sweep spills; the one-CTA-per-SM 5x5 kernels
// are faster with the single sweep (244 registers are available to them anyway) and the one-barrier ring
#define DW2_BWD_FAST                                     \
  if (s == 1) {                                          \
    if (k == 3) { if (half) DW2_X(3, 1, 4, 2, 2, 2, 32, false, true) else DW2_X(3, 1, 4, 2, 2, 2, 64, false, true) } \
    else { if (half5) DW2_X(5, 1, 4, 2, 3, 1, 32, true, false) else DW2_X(5, 1, 4, 2, 3, 1, 64, true, false) }       \
  } else {                                               \
    if (k == 3) { if (half) DW2_X(3, 2, 4, 4, 2, 2, 32, false, true) else DW2_X(3, 2, 4, 4, 2, 2, 64, false, true) } \
    else DW2_X(5, 2, 4, 2, 3, 1, 64, true, false)        \
  }
#define DW2_BWD_F32                                      \
  if (s == 1) {                                          \
    if (k == 3) DW2_B(float, 3, 1, 2, 2, 1, 1, 64);      \
    DW2_B(float, 5, 1, 2, 2, 1, 1, 64);                  \
  }                                                      \
  if (k == 3) DW2_B(float, 3, 2, 2, 2, 1, 1, 64);        \
  DW2_B(float, 5, 2, 2, 2, 1, 1, 64);

/* Fused depthwise backward: dx (data gradient w.r.t. the BatchNorm output feeding the conv, times act' when
 * apply_act_bwd) + BNBWD statistics of dx against in.raw + dW, in one pass. */
int32_t s2lc_dwconv_bwd(const s2lc_lazygrad* g, const float* w, const s2lc_lazy* in, int32_t apply_act_bwd,
                        const void* res, int64_t ld_res, void* out, int64_t ld_out, double* stats, float* dw, int32_t N,
                        int32_t H, int32_t W, int32_t C, int32_t k, int32_t s, int32_t dtype, void* stream) {
  S2LC_REQUIRE((k == 3 || k == 5) && (s == 1 || s == 2), S2LC_ERR_SHAPE, "dwconv: k=%d s=%d unsupported", k, s);
  S2LC_REQUIRE(C % 8 == 0 && g->ld_dy % 8 == 0 && ld_out % 8 == 0 && in->ld % 8 == 0 && (!g->z || g->ld_z % 8 == 0),
               S2LC_ERR_ALIGN, "dwconv: C and strides must be multiples of 8 (C=%d)", C);
  S2LC_REQUIRE(in->gate == nullptr, S2LC_ERR_SHAPE, "dwconv: gated input is not a reference configuration");
  if (s == 2 && ((H | W) & 1)) {
    // odd extents under stride 2 change the TF-"same" padding parity: use the tile kernels (two passes)
    int32_t r = s2lc_dwconv_bwd_weight(g, in, dw, N, H, W, C, k, s, dtype, stream);
    if (r) return r;
    return s2lc_dwconv_bwd_data(g, w, apply_act_bwd ? in : nullptr, res, ld_res, out, ld_out, stats, N, H, W, C, k, s,
                                dtype, stream);
  }
  cudaStream_t st = (cudaStream_t)stream;
  const bool wide = W > 16;
  const bool half = W >= 64 && (((C + 63) % 64) < 32 || W >= 128 || (s2lc::debug_flags() & s2lc::DBG_DW_CB32));
  const bool half5 = W >= 64;
  int fast = 0;
  if (dtype == S2LC_BF16 && wide && !(s2lc::debug_flags() & s2lc::DBG_DW_GENERIC)) {
    if (in->act == S2LC_ACT_SILU && in->scale && in->shift && apply_act_bwd && !res) fast = 1;
    if (in->act == S2LC_ACT_NONE && !in->scale && !in->shift && !apply_act_bwd) fast = 2;
  }
  if (fast) { DW2_BWD_FAST }
  if (dtype == S2LC_BF16) { DW2_BWD_BF16 }
  if (dtype == S2LC_F32) { DW2_BWD_F32 }
  s2lc::set_error("dwconv: bad dtype %d", dtype);
  return S2LC_ERR_SHAPE;
}

}  // extern "C"

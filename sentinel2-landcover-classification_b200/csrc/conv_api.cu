// conv_api.cu — C-ABI entry points of the implicit-GEMM convolution family; validates, flattens the
// parameter block and dispatches to the tcgen05 engine (bf16) or the CUDA-core engine (fp32 / leftovers).
#include <cstdlib>

#include "conv_common.cuh"

using namespace s2lc;

static bool force_simt() { return (debug_flags() & DBG_FORCE_SIMT) != 0; }

extern "C" {

int32_t s2lc_conv_fwd(const s2lc_conv_params* p, void* stream) {
  S2LC_REQUIRE(p != nullptr, S2LC_ERR_SHAPE, "conv_fwd: null params");
  S2LC_REQUIRE(p->mode >= 0 && p->mode <= 3, S2LC_ERR_SHAPE, "conv_fwd: bad mode %d", p->mode);
  S2LC_REQUIRE(p->dtype == S2LC_F32 || p->dtype == S2LC_BF16, S2LC_ERR_SHAPE, "conv_fwd: bad dtype %d", p->dtype);
  S2LC_REQUIRE(p->Cpad % 8 == 0 && p->Cpad >= p->C, S2LC_ERR_ALIGN, "conv_fwd: Cpad=%d must be a multiple of 8 >= C=%d",
               p->Cpad, p->C);
  ConvArgs A{};
  A.mode = p->mode; A.dtype = p->dtype; A.N = p->N; A.H = p->H; A.W = p->W; A.C = p->C; A.Cpad = p->Cpad;
  A.Kpad = p->Kpad; A.Nout = p->Nout; A.Npad = p->Npad; A.a_is_grad = p->a_is_grad;
  A.taps = p->mode == S2LC_CONV_3X3 ? 9 : (p->mode == S2LC_CONVT_DGRAD ? 4 : 1);
  S2LC_REQUIRE(A.Kpad >= A.taps * A.Cpad && A.Npad >= A.Nout, S2LC_ERR_SHAPE, "conv_fwd: packed weight too small");
  A.Cout = p->mode == S2LC_CONVT_FWD ? p->Nout / 4 : p->Nout;
  S2LC_REQUIRE(p->mode != S2LC_CONVT_FWD || p->Nout % 4 == 0, S2LC_ERR_SHAPE, "conv_fwd: convT Nout %% 4");
  if (p->a_is_grad) {
    S2LC_REQUIRE(p->a_bwd != nullptr, S2LC_ERR_SHAPE, "conv_fwd: a_bwd null");
    A.a_bwd = *p->a_bwd;
  } else {
    S2LC_REQUIRE(p->a_fwd != nullptr, S2LC_ERR_SHAPE, "conv_fwd: a_fwd null");
    A.a_fwd = *p->a_fwd;
  }
  A.has_prev = p->prev != nullptr;
  if (p->prev) A.prev = *p->prev;
  A.stats_kind = p->stats ? p->stats_kind : S2LC_STATS_NONE;
  S2LC_REQUIRE(A.stats_kind != S2LC_STATS_BNBWD || A.has_prev, S2LC_ERR_SHAPE, "conv_fwd: BNBWD stats need prev");
  A.w_packed = p->w_packed; A.bias = p->bias; A.out = p->out; A.ld_out = p->ld_out;
  A.res1 = p->res1; A.ld_res1 = p->ld_res1; A.res2 = p->res2; A.ld_res2 = p->ld_res2;
  A.stats = p->stats; A.stats_rowscale = p->stats_rowscale;
  cudaStream_t st = (cudaStream_t)stream;
  if (A.dtype == S2LC_BF16 && !force_simt()) {
    int32_t r = launch_tconv_tc(A, st);
    if (r <= 0) return r;  // taken (0) or error (<0)
    r = launch_sconv_tc(A, st);
    if (r <= 0) return r;
    r = launch_conv_tc(A, st);
    if (r <= 0) return r;
    note_simt_fallback();   // counted, never silent: s2lc_simt_fallbacks() must stay 0 on the reference architectures
  }
  return launch_conv_simt(A, st);
}

int32_t s2lc_conv_wgrad(const s2lc_wgrad_params* p, void* stream) {
  S2LC_REQUIRE(p != nullptr && p->a && p->g && p->dw, S2LC_ERR_SHAPE, "conv_wgrad: null params");
  S2LC_REQUIRE(p->mode == S2LC_CONV_1X1 || p->mode == S2LC_CONV_3X3 || p->mode == S2LC_CONVT_FWD, S2LC_ERR_SHAPE,
               "conv_wgrad: bad mode %d", p->mode);
  WgradArgs A{};
  A.mode = p->mode; A.dtype = p->dtype; A.N = p->N; A.H = p->H; A.W = p->W; A.Cin = p->Cin; A.Cout = p->Cout;
  A.taps = p->mode == S2LC_CONV_3X3 ? 9 : 1;
  A.ngroups = p->mode == S2LC_CONVT_FWD ? 4 : 1;
  A.a = *p->a; A.g = *p->g; A.dw = p->dw; A.dbias = p->dbias;
  cudaStream_t st = (cudaStream_t)stream;
  if (A.dtype == S2LC_BF16 && !force_simt()) {
    int32_t r = launch_twgrad_tc(A, st);
    if (r <= 0) return r;
    r = launch_swgrad_tc(A, st);
    if (r <= 0) return r;
    r = launch_wgrad_tc(A, st);
    if (r <= 0) return r;
    note_simt_fallback();
  }
  return launch_wgrad_simt(A, st);
}

}  // extern "C"

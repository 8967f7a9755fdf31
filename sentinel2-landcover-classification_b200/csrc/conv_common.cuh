// conv_common.cuh — operand loaders and the epilogue of the implicit-GEMM convolution family,
// shared by the CUDA-core engine (igemm_simt.cu) and the tcgen05 engine (igemm_tc.cu).
#pragma once
#include <cuda.h>   // CUtensorMap (types only: no link-time dependency on libcuda)

#include "common.cuh"

namespace s2lc {

// Device-side flattened copy of s2lc_conv_params (nested structs by value).
struct ConvArgs {
  int mode, dtype, N, H, W, C, Cpad, Kpad, Nout, Npad, taps, a_is_grad;
  int Cout;          // channels per output pixel (Nout, or Nout/4 for convT fwd)
  int has_prev, stats_kind;
  s2lc_lazy a_fwd;
  s2lc_lazygrad a_bwd;
  s2lc_lazy prev;
  const void* w_packed;
  const float* bias;
  void* out;
  int64_t ld_out;
  const void* res1;
  int64_t ld_res1;
  const void* res2;
  int64_t ld_res2;
  double* stats;
  const float* stats_rowscale;
};

struct RowCoord {
  int n, h, w;
};

__device__ __forceinline__ RowCoord decode_row(const ConvArgs& P, int64_t m) {
  RowCoord r;
  int hw = P.H * P.W;
  r.n = (int)(m / hw);
  int rem = (int)(m - (int64_t)r.n * hw);
  r.h = rem / P.W;
  r.w = rem - r.h * P.W;
  return r;
}

// source row (in the A tensor) feeding output row `rc` at `tap`; false = zero padding
__device__ __forceinline__ bool src_row(const ConvArgs& P, const RowCoord& rc, int tap, int64_t& srow) {
  if (P.mode == S2LC_CONV_3X3) {
    int hh = rc.h + tap / 3 - 1, ww = rc.w + tap % 3 - 1;
    if (hh < 0 || hh >= P.H || ww < 0 || ww >= P.W) return false;
    srow = ((int64_t)rc.n * P.H + hh) * P.W + ww;
    return true;
  }
  if (P.mode == S2LC_CONVT_DGRAD) {  // A lives at 2H x 2W; tap = dy*2+dx
    srow = ((int64_t)rc.n * (2 * P.H) + 2 * rc.h + (tap >> 1)) * (2 * P.W) + 2 * rc.w + (tap & 1);
    return true;
  }
  srow = ((int64_t)rc.n * P.H + rc.h) * P.W + rc.w;
  return true;
}

template <typename T>
__device__ __forceinline__ float load_a(const ConvArgs& P, int64_t srow, int c, int n) {
  if (P.a_is_grad) return lazygrad_at<T>(P.a_bwd, srow, c, n);
  return lazy_at<T>(P.a_fwd, srow, c, n, P.C);
}

// Applies bias / residuals / act' / store / statistics contribution for one output element.
// (m, rc) = GEMM row and its pixel; n = GEMM column.  Returns the two statistics contributions.
template <typename T>
__device__ __forceinline__ void epilogue_store(const ConvArgs& P, const RowCoord& rc, int64_t m, int n, float acc,
                                               float& s0, float& s1) {
  int col = n;
  int64_t orow = m;
  if (P.mode == S2LC_CONVT_FWD) {
    int q = n / P.Cout;
    col = n - q * P.Cout;
    orow = ((int64_t)rc.n * (2 * P.H) + 2 * rc.h + (q >> 1)) * (2 * P.W) + 2 * rc.w + (q & 1);
  }
  float v = acc;
  if (P.bias) v += P.bias[col];
  if (P.res1) v += ldf<T>(reinterpret_cast<const T*>(P.res1) + orow * P.ld_res1 + col);
  if (P.res2) v += ldf<T>(reinterpret_cast<const T*>(P.res2) + orow * P.ld_res2 + col);
  float zprev = 0.f;
  if (P.has_prev) {
    zprev = ldf<T>(reinterpret_cast<const T*>(P.prev.raw) + orow * P.prev.ld + col);
    float u = zprev;
    if (P.prev.scale) u = u * P.prev.scale[col] + (P.prev.shift ? P.prev.shift[col] : 0.f);
    v *= act_bwd(u, P.prev.act);
  }
  v = roundf_as<T>(v);
  stf<T>(reinterpret_cast<T*>(P.out) + orow * P.ld_out + col, v);
  s0 = 0.f;
  s1 = 0.f;
  if (P.stats_kind == S2LC_STATS_MOMENTS) {
    s0 = v;
    s1 = v * v;
  } else if (P.stats_kind == S2LC_STATS_BNBWD) {
    float d = P.stats_rowscale ? v * P.stats_rowscale[rc.n] : v;
    s0 = d;
    s1 = d * zprev;
  }
}

// ------------------------------------------------------------------------------------------------ wgrad
constexpr int kWgradStem = 4;  // internal mode: 3x3 stride-2 TF-"same" conv on the fp32 NCHW input tile
struct WgradArgs {
  int mode, dtype, N, H, W, Cin, Cout, taps, ngroups;  // ngroups = 4 for convT (q = dy*2+dx), else 1
  int Hin, Win, pad_t, pad_l;                          // stem only: input geometry (H, W are the OUTPUT dims)
  s2lc_lazy a;
  s2lc_lazygrad g;
  float* dw;
  float* dbias;
};

struct WRow {
  int n, h, w;
  int64_t row;
};

__device__ __forceinline__ WRow wdecode_row(const WgradArgs& P, int64_t m) {
  WRow r;
  int hw = P.H * P.W;
  r.n = (int)(m / hw);
  int rem = (int)(m - (int64_t)r.n * hw);
  r.h = rem / P.W;
  r.w = rem - r.h * P.W;
  r.row = m;
  return r;
}

// G(m, n): gradient w.r.t. the raw forward output at the pixel(s) produced from input pixel m
template <typename T>
__device__ __forceinline__ float load_g(const WgradArgs& P, const WRow& r, int n) {
  if (P.mode == S2LC_CONVT_FWD) {
    int q = n / P.Cout, co = n - q * P.Cout;
    int64_t urow = ((int64_t)r.n * (2 * P.H) + 2 * r.h + (q >> 1)) * (2 * P.W) + 2 * r.w + (q & 1);
    return lazygrad_at<T>(P.g, urow, co, r.n);
  }
  return lazygrad_at<T>(P.g, r.row, n, r.n);
}

// A(m, k): forward input seen by output pixel m through tap k / Cin
template <typename T>
__device__ __forceinline__ float load_wa(const WgradArgs& P, const WRow& r, int k) {
  if (P.mode == kWgradStem) {  // k = ci*9 + tap (OIHW order), x is fp32 NCHW
    int ci = k / 9, tap = k - ci * 9;
    int iy = 2 * r.h + tap / 3 - P.pad_t, ix = 2 * r.w + tap % 3 - P.pad_l;
    if (iy < 0 || iy >= P.Hin || ix < 0 || ix >= P.Win) return 0.f;
    return reinterpret_cast<const float*>(P.a.raw)[(((int64_t)r.n * P.Cin + ci) * P.Hin + iy) * P.Win + ix];
  }
  int tap = k / P.Cin, c = k - tap * P.Cin;
  if (P.mode == S2LC_CONV_3X3) {
    int hh = r.h + tap / 3 - 1, ww = r.w + tap % 3 - 1;
    if (hh < 0 || hh >= P.H || ww < 0 || ww >= P.W) return 0.f;
    return lazy_at<T>(P.a, ((int64_t)r.n * P.H + hh) * P.W + ww, c, r.n, P.Cin);
  }
  return lazy_at<T>(P.a, r.row, c, r.n, P.Cin);
}

// index of (n, k) in the reference-layout weight gradient
__device__ __forceinline__ int64_t wgrad_index(const WgradArgs& P, int n, int k) {
  if (P.mode == S2LC_CONV_1X1) return (int64_t)n * P.Cin + k;                 // [O, I, 1, 1]
  if (P.mode == kWgradStem) return (int64_t)n * (P.Cin * 9) + k;              // [O, I, 3, 3], k = ci*9 + tap
  if (P.mode == S2LC_CONV_3X3) {                                              // [O, I, 3, 3], k = tap*Cin + c
    int tap = k / P.Cin, c = k - tap * P.Cin;
    return ((int64_t)n * P.Cin + c) * 9 + tap;
  }
  int q = n / P.Cout, co = n - q * P.Cout;                                    // convT [I, O, 2, 2], k = ci
  return ((int64_t)k * P.Cout + co) * 4 + q;
}

// shared TMA helpers (twgrad_tc.cu): bf16 tensor map, SWIZZLE_128B, zero fill outside the tensor (dims / box fastest
// dimension first, strides in ELEMENTS for dimensions 1..rank-1); per-channel sum of a bf16 [rows, C] tensor -> fp32
bool make_tma_map_bf16(CUtensorMap* map, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_el,
                       const uint32_t* box);
void launch_colsum_bf16(const void* g, int64_t ld, int64_t rows, int C, float* dbias, cudaStream_t st);
void note_simt_fallback();
void note_tma_launch();
int32_t launch_conv_simt(const ConvArgs& A, cudaStream_t st);
int32_t launch_wgrad_simt(const WgradArgs& A, cudaStream_t st);
int32_t launch_conv_tc(const ConvArgs& A, cudaStream_t st);      // igemm_tc.cu; returns 1 if shape not taken
int32_t launch_tconv_tc(const ConvArgs& A, cudaStream_t st);     // tconv_tc.cu (TMA-fed engine, plain operands); 1 if not taken
int32_t launch_sconv_tc(const ConvArgs& A, cudaStream_t st);     // sconv_tc.cu (streaming engine); returns 1 if shape not taken
int32_t launch_wgrad_tc(const WgradArgs& A, cudaStream_t st);
int32_t launch_twgrad_tc(const WgradArgs& A, cudaStream_t st);   // twgrad_tc.cu (TMA-fed engine, plain operands, deep layers)
int32_t launch_swgrad_tc(const WgradArgs& A, cudaStream_t st);   // swgrad_tc.cu (streaming engine); 1 if shape not taken

}  // namespace s2lc

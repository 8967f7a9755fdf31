// stem.cu — the encoder stem: Conv2dSamePadding(C_in -> C_o, k=3, stride 2, no bias) applied directly to the
// fp32 NCHW input tile (reference src/modules/efficientnet_unet.py:191-194 with the run-time TF-"same" padding of
// :288-297).  K = 9*C_in = 54 is too small for the tensor pipe to matter: one thread per output
// pixel, filter in shared memory (broadcast reads), NHWC raw output + BatchNorm moments in the epilogue.
#include "conv_common.cuh"

namespace s2lc {

constexpr int STEM_CIN_MAX = 8;

template <typename T>
__global__ void __launch_bounds__(128) stem_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                       T* __restrict__ out, double* stats, int N, int Cin, int H,
                                                       int W, int Co, int Ho, int Wo, int pad_t, int pad_l) {
  extern __shared__ float smem[];
  float* ws = smem;                  // [Cin*9][Co]
  float* sstat = smem + Cin * 9 * Co;  // [2][Co]
  const int K = Cin * 9;
  for (int i = threadIdx.x; i < K * Co; i += blockDim.x) {
    int k = i / Co, co = i - k * Co;
    ws[i] = w[(int64_t)co * K + k];
  }
  for (int i = threadIdx.x; i < 2 * Co; i += blockDim.x) sstat[i] = 0.f;
  __syncthreads();
  const int64_t total = (int64_t)N * Ho * Wo;
  const int lane = threadIdx.x & 31;
  for (int64_t base = (int64_t)blockIdx.x * blockDim.x; base < total; base += (int64_t)gridDim.x * blockDim.x) {
    int64_t pix = base + threadIdx.x;
    bool valid = pix < total;
    int n = 0, oy = 0, ox = 0;
    if (valid) {
      n = (int)(pix / ((int64_t)Ho * Wo));
      int rem = (int)(pix - (int64_t)n * Ho * Wo);
      oy = rem / Wo;
      ox = rem - oy * Wo;
    }
    float patch[STEM_CIN_MAX * 9];
#pragma unroll
    for (int ci = 0; ci < STEM_CIN_MAX; ++ci) {
#pragma unroll
      for (int t = 0; t < 9; ++t) {
        float v = 0.f;
        if (valid && ci < Cin) {
          int iy = 2 * oy + t / 3 - pad_t, ix = 2 * ox + t % 3 - pad_l;
          if (iy >= 0 && iy < H && ix >= 0 && ix < W) v = x[(((int64_t)n * Cin + ci) * H + iy) * W + ix];
        }
        patch[ci * 9 + t] = v;
      }
    }
    for (int c0 = 0; c0 < Co; c0 += 8) {
      float acc[8] = {};
#pragma unroll
      for (int ci = 0; ci < STEM_CIN_MAX; ++ci) {
        if (ci < Cin) {
#pragma unroll
          for (int t = 0; t < 9; ++t) {
            const float4* wp = reinterpret_cast<const float4*>(ws + (ci * 9 + t) * Co + c0);
            float4 w0 = wp[0], w1 = wp[1];
            float p = patch[ci * 9 + t];
            acc[0] = fmaf(p, w0.x, acc[0]); acc[1] = fmaf(p, w0.y, acc[1]);
            acc[2] = fmaf(p, w0.z, acc[2]); acc[3] = fmaf(p, w0.w, acc[3]);
            acc[4] = fmaf(p, w1.x, acc[4]); acc[5] = fmaf(p, w1.y, acc[5]);
            acc[6] = fmaf(p, w1.z, acc[6]); acc[7] = fmaf(p, w1.w, acc[7]);
          }
        }
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = valid ? roundf_as<T>(acc[j]) : 0.f;
      if (valid) Vec8<T>::store(out + pix * Co + c0, acc);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float s0 = warp_sum(acc[j]), s1 = warp_sum(acc[j] * acc[j]);
        if (lane == 0) {
          atomicAdd(&sstat[c0 + j], s0);
          atomicAdd(&sstat[Co + c0 + j], s1);
        }
      }
    }
  }
  __syncthreads();
  if (stats)
    for (int i = threadIdx.x; i < 2 * Co; i += blockDim.x) atomicAdd(&stats[i], (double)sstat[i]);
}

// im2col of the stem (bf16 production path): cols[(n, oy, ox), k] = x[n, ci, 2*oy + ty - pad_t, 2*ox + tx - pad_l]
// with k = ci*9 + ty*3 + tx (the OIHW flattening of the weight), zero beyond Cin*9 up to `ld`.  One block = 128
// output pixels of one output row: the 3 x Cin input row segments they touch are read once, coalesced, into shared
// memory; every (pixel, 8-column vector) is then one 16 B store, 8 lanes per 128 B row.  The stem conv and its weight
// gradient become a K = 56 contraction on the tensor-core engines (the direct CUDA-core kernels ran at 0.43 TB/s).
template <typename T>
__global__ void __launch_bounds__(256) stem_im2col_kernel(const float* __restrict__ x, T* __restrict__ cols, int64_t ld,
                                                          int N, int Cin, int H, int W, int Ho, int Wo, int pad_t,
                                                          int pad_l) {
  extern __shared__ float patch[];                 // [Cin][3][SEG], SEG = 2*128 + 2
  constexpr int PX = 128, SEG = 2 * PX + 2;
  const int xblocks = (Wo + PX - 1) / PX;
  const int bx = blockIdx.x % xblocks;
  const int row = blockIdx.x / xblocks;            // n * Ho + oy
  const int n = row / Ho, oy = row - n * Ho;
  const int ox0 = bx * PX;
  const int ix0 = 2 * ox0 - pad_l;
  for (int i = threadIdx.x; i < Cin * 3 * SEG; i += blockDim.x) {
    const int ci = i / (3 * SEG), r = i - ci * 3 * SEG, ty = r / SEG, j = r - ty * SEG;
    const int iy = 2 * oy + ty - pad_t, ix = ix0 + j;
    float v = 0.f;
    if (iy >= 0 && iy < H && ix >= 0 && ix < W) v = __ldg(x + (((int64_t)n * Cin + ci) * H + iy) * W + ix);
    patch[i] = v;
  }
  __syncthreads();
  const int nv = (int)(ld / 8), K = Cin * 9;
  for (int i = threadIdx.x; i < PX * nv; i += blockDim.x) {
    const int px = i / nv, v8 = i - px * nv;
    if (ox0 + px >= Wo) continue;
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int k = v8 * 8 + j;
      float val = 0.f;
      if (k < K) {
        const int ci = k / 9, t = k - ci * 9, ty = t / 3, tx = t - ty * 3;
        val = patch[(ci * 3 + ty) * SEG + 2 * px + tx];
      }
      v[j] = val;
    }
    Vec8<T>::store(cols + ((int64_t)row * Wo + ox0 + px) * ld + v8 * 8, v);
  }
}

}  // namespace s2lc

using namespace s2lc;

static void stem_geom(int H, int W, int& Ho, int& Wo, int& pt, int& pl) {
  Ho = (H + 1) / 2; Wo = (W + 1) / 2;
  int ph = (Ho - 1) * 2 + 3 - H; if (ph < 0) ph = 0;
  int pw = (Wo - 1) * 2 + 3 - W; if (pw < 0) pw = 0;
  pt = ph / 2; pl = pw / 2;
}

extern "C" {

int32_t s2lc_stem_conv_fwd(const float* x_nchw, const float* w_oihw, void* out, double* stats, int32_t N, int32_t Cin,
                           int32_t H, int32_t W, int32_t Co, int32_t dtype, void* stream) {
  S2LC_REQUIRE(Cin >= 1 && Cin <= STEM_CIN_MAX, S2LC_ERR_SHAPE, "stem: in_channels=%d unsupported (1..%d)", Cin,
               STEM_CIN_MAX);
  S2LC_REQUIRE(Co % 8 == 0, S2LC_ERR_ALIGN, "stem: Co=%d must be a multiple of 8", Co);
  int Ho, Wo, pt, pl;
  stem_geom(H, W, Ho, Wo, pt, pl);
  int64_t total = (int64_t)N * Ho * Wo;
  int64_t blocks = (total + 127) / 128;
  if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
  size_t smem = sizeof(float) * ((size_t)Cin * 9 * Co + 2 * Co);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == S2LC_F32)
    stem_fwd_kernel<float><<<(int)blocks, 128, smem, st>>>(x_nchw, w_oihw, (float*)out, stats, N, Cin, H, W, Co, Ho,
                                                           Wo, pt, pl);
  else if (dtype == S2LC_BF16)
    stem_fwd_kernel<s2lc::bf16><<<(int)blocks, 128, smem, st>>>(x_nchw, w_oihw, (s2lc::bf16*)out, stats, N, Cin, H, W,
                                                                Co, Ho, Wo, pt, pl);
  else {
    set_error("stem: bad dtype %d", dtype);
    return S2LC_ERR_SHAPE;
  }
  return check_launch("stem_conv_fwd");
}

int32_t s2lc_stem_im2col(const float* x_nchw, void* cols, int64_t ld, int32_t N, int32_t Cin, int32_t H, int32_t W,
                         int32_t dtype, void* stream) {
  S2LC_REQUIRE(Cin >= 1 && Cin * 9 <= ld && ld % 8 == 0, S2LC_ERR_SHAPE, "stem_im2col: Cin=%d does not fit ld=%lld", Cin,
               (long long)ld);
  int Ho, Wo, pt, pl;
  stem_geom(H, W, Ho, Wo, pt, pl);
  const int xblocks = (Wo + 127) / 128;
  const size_t smem = sizeof(float) * (size_t)Cin * 3 * 258;
  const int64_t blocks = (int64_t)N * Ho * xblocks;
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == S2LC_BF16)
    stem_im2col_kernel<s2lc::bf16><<<(unsigned)blocks, 256, smem, st>>>(x_nchw, (s2lc::bf16*)cols, ld, N, Cin, H, W, Ho, Wo, pt, pl);
  else if (dtype == S2LC_F32)
    stem_im2col_kernel<float><<<(unsigned)blocks, 256, smem, st>>>(x_nchw, (float*)cols, ld, N, Cin, H, W, Ho, Wo, pt, pl);
  else {
    set_error("stem_im2col: bad dtype %d", dtype);
    return S2LC_ERR_SHAPE;
  }
  return check_launch("stem_im2col");
}

int32_t s2lc_stem_conv_wgrad(const float* x_nchw, const s2lc_lazygrad* g, float* dw_oihw, int32_t N, int32_t Cin,
                             int32_t H, int32_t W, int32_t Co, int32_t dtype, void* stream) {
  S2LC_REQUIRE(g != nullptr && dw_oihw != nullptr, S2LC_ERR_SHAPE, "stem_wgrad: null");
  WgradArgs A{};
  int Ho, Wo, pt, pl;
  stem_geom(H, W, Ho, Wo, pt, pl);
  A.mode = kWgradStem; A.dtype = dtype; A.N = N; A.H = Ho; A.W = Wo; A.Cin = Cin; A.Cout = Co;
  A.taps = 9; A.ngroups = 1; A.Hin = H; A.Win = W; A.pad_t = pt; A.pad_l = pl;
  A.a.raw = x_nchw; A.g = *g; A.dw = dw_oihw; A.dbias = nullptr;
  return launch_wgrad_simt(A, (cudaStream_t)stream);
}

}  // extern "C"

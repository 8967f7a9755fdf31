// outconv.cu — the classifier `out_conv1x1` (C -> num_classes, 1x1, bias; reference
// src/modules/efficientnet_unet.py:121,137) forward and backward as per-pixel HBM-bound kernels.
//
// N = num_classes (2..8) is not a tensor-core shape: at 512x512 the op moves 64 B of activations per pixel and does
// 256 FMAs, so it is bound by HBM and the only thing that matters is touching every byte once:
//   forward : logits = W . act(bn(z)) + b                      reads z (C ch), writes logits (N ch)
//   backward: dW += dlogits (x) act(bn(z)),  db += dlogits,     reads z, dlogits; writes dz;
//             dz = (W^T dlogits) * act'(bn(z)), + BatchNorm-backward moments (sum dz, sum dz*z) of dz
// (one pass instead of the three launches wgrad + dgrad-with-epilogue of the generic engine).
// Thread layout: C/8 consecutive lanes share a pixel, each owns one 16 B chunk (8 channels) => a warp reads
// 32 consecutive 16 B chunks (512 B, fully coalesced); partial dot products are combined with shuffles.
#include "common.cuh"

namespace s2lc {

constexpr int OC_THREADS = 256;

template <typename T, int NCLS>
__global__ void __launch_bounds__(OC_THREADS) outconv_fwd_kernel(s2lc_lazy in, const float* __restrict__ w,
                                                                 const float* __restrict__ bias, T* __restrict__ out,
                                                                 int64_t ld_out, int64_t rows, int C) {
  const int lpp = C >> 3;                       // lanes per pixel (power of two <= 32)
  const int tid = blockIdx.x * OC_THREADS + threadIdx.x;
  const int c8 = (tid & (lpp - 1)) * 8;
  float wr[NCLS][8], sc[8], sh[8];
#pragma unroll
  for (int k = 0; k < NCLS; ++k)
#pragma unroll
    for (int j = 0; j < 8; ++j) wr[k][j] = w[k * C + c8 + j];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    sc[j] = in.scale ? in.scale[c8 + j] : 1.f;
    sh[j] = in.shift ? in.shift[c8 + j] : 0.f;
  }
  const T* raw = reinterpret_cast<const T*>(in.raw);
  const int ppw = 32 / lpp;                                        // pixels per warp
  const int64_t stride = (int64_t)gridDim.x * (OC_THREADS / 32) * ppw;
  for (int64_t base = (int64_t)(tid >> 5) * ppw; base < rows; base += stride) {   // warp-uniform (shuffles below)
    const int64_t r = base + ((tid & 31) / lpp);
    const bool valid = r < rows;
    float v[8];
    if (valid) {
      Vec8<T>::load(raw + r * in.ld + c8, v);
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = 0.f;
    }
    float acc[NCLS];
#pragma unroll
    for (int k = 0; k < NCLS; ++k) acc[k] = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float a = act_fwd_t<T>(fmaf(v[j], sc[j], sh[j]), in.act);
      if (sizeof(T) == 2) a = roundf_as<T>(a);   // the tensor-core engine rounds its A operand to bf16: same values
#pragma unroll
      for (int k = 0; k < NCLS; ++k) acc[k] = fmaf(a, wr[k][j], acc[k]);
    }
    for (int o = lpp >> 1; o > 0; o >>= 1) {
#pragma unroll
      for (int k = 0; k < NCLS; ++k) acc[k] += __shfl_xor_sync(0xffffffffu, acc[k], o);
    }
    if (valid && c8 == 0) {
#pragma unroll
      for (int k = 0; k < NCLS; ++k) stf<T>(out + r * ld_out + k, acc[k] + (bias ? bias[k] : 0.f));
    }
  }
}

template <typename T, int NCLS>
__global__ void __launch_bounds__(OC_THREADS) outconv_bwd_kernel(s2lc_lazy in, const T* __restrict__ dl, int64_t ld_dl,
                                                                 const float* __restrict__ w, float* __restrict__ dw,
                                                                 float* __restrict__ dbias, T* __restrict__ dz,
                                                                 int64_t ld_dz, double* __restrict__ stats,
                                                                 int64_t rows, int C) {
  extern __shared__ float sred[];               // [NCLS*C] dW | [2*C] stats | [NCLS] db
  const int lpp = C >> 3;
  const int tid = blockIdx.x * OC_THREADS + threadIdx.x;
  const int c8 = (tid & (lpp - 1)) * 8;
  const int nred = NCLS * C + 2 * C + NCLS;
  for (int i = threadIdx.x; i < nred; i += OC_THREADS) sred[i] = 0.f;
  float wr[NCLS][8], sc[8], sh[8];
#pragma unroll
  for (int k = 0; k < NCLS; ++k)
#pragma unroll
    for (int j = 0; j < 8; ++j) wr[k][j] = w[k * C + c8 + j];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    sc[j] = in.scale ? in.scale[c8 + j] : 1.f;
    sh[j] = in.shift ? in.shift[c8 + j] : 0.f;
  }
  float gw[NCLS][8], s0[8], s1[8], gb[NCLS];
#pragma unroll
  for (int k = 0; k < NCLS; ++k) {
    gb[k] = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) gw[k][j] = 0.f;
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) { s0[j] = 0.f; s1[j] = 0.f; }
  const T* raw = reinterpret_cast<const T*>(in.raw);
  const int64_t ppi = (int64_t)gridDim.x * OC_THREADS / lpp;
  for (int64_t r = tid / lpp; r < rows; r += ppi) {
    float z[8], g[NCLS], o[8];
    Vec8<T>::load(raw + r * in.ld + c8, z);
#pragma unroll
    for (int k = 0; k < NCLS; ++k) g[k] = ldf<T>(dl + r * ld_dl + k);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float u = fmaf(z[j], sc[j], sh[j]);
      float a = act_fwd_t<T>(u, in.act);
      if (sizeof(T) == 2) a = roundf_as<T>(a);
      float d = 0.f;
#pragma unroll
      for (int k = 0; k < NCLS; ++k) {
        gw[k][j] = fmaf(g[k], a, gw[k][j]);
        d = fmaf(g[k], wr[k][j], d);
      }
      d = roundf_as<T>(d * act_bwd_t<T>(u, in.act));
      o[j] = d;
      s0[j] += d;
      s1[j] = fmaf(d, z[j], s1[j]);
    }
    Vec8<T>::store(dz + r * ld_dz + c8, o);
    if (c8 == 0) {
#pragma unroll
      for (int k = 0; k < NCLS; ++k) gb[k] += g[k];
    }
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < NCLS; ++k)
#pragma unroll
    for (int j = 0; j < 8; ++j) atomicAdd(&sred[k * C + c8 + j], gw[k][j]);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    atomicAdd(&sred[NCLS * C + c8 + j], s0[j]);
    atomicAdd(&sred[NCLS * C + C + c8 + j], s1[j]);
  }
  if (c8 == 0) {
#pragma unroll
    for (int k = 0; k < NCLS; ++k) atomicAdd(&sred[NCLS * C + 2 * C + k], gb[k]);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < nred; i += OC_THREADS) {
    const float v = sred[i];
    if (i < NCLS * C) atomicAdd(dw + i, v);
    else if (i < NCLS * C + 2 * C) { if (stats) atomicAdd(stats + (i - NCLS * C), (double)v); }
    else if (dbias) atomicAdd(dbias + (i - NCLS * C - 2 * C), v);
  }
}

static int oc_grid(int64_t rows, int lpp) {
  int64_t threads = rows * lpp;
  int64_t blocks = (threads + OC_THREADS - 1) / OC_THREADS;
  const int64_t cap = (int64_t)kNumSMs * 8;        // 8 resident CTAs of 256 threads per SM, grid-stride beyond
  return (int)(blocks < cap ? (blocks < 1 ? 1 : blocks) : cap);
}

}  // namespace s2lc

using namespace s2lc;

#define OC_CHECK(C, N)                                                                                             \
  S2LC_REQUIRE((C) % 8 == 0 && (C) >= 8 && (C) <= 256 && (((C) / 8) & ((C) / 8 - 1)) == 0, S2LC_ERR_SHAPE,          \
               "outconv: C=%d must be 8 * a power of two (8..256)", (C));                                          \
  S2LC_REQUIRE((N) >= 2 && (N) <= 8, S2LC_ERR_SHAPE, "outconv: num_classes=%d unsupported (2..8)", (N));

#define OC_SWITCH(N, BODY)                                                        \
  switch (N) {                                                                    \
    case 2: { constexpr int NC = 2; BODY; } break;                                \
    case 3: { constexpr int NC = 3; BODY; } break;                                \
    case 4: { constexpr int NC = 4; BODY; } break;                                \
    case 5: { constexpr int NC = 5; BODY; } break;                                \
    case 6: { constexpr int NC = 6; BODY; } break;                                \
    case 7: { constexpr int NC = 7; BODY; } break;                                \
    default: { constexpr int NC = 8; BODY; } break;                               \
  }

extern "C" {

int32_t s2lc_outconv_fwd(const s2lc_lazy* in, const float* w, const float* bias, void* logits, int64_t ld_out,
                         int64_t rows, int32_t C, int32_t n_classes, int32_t dtype, void* stream) {
  S2LC_REQUIRE(in && w && logits, S2LC_ERR_SHAPE, "outconv_fwd: null argument");
  OC_CHECK(C, n_classes)
  S2LC_REQUIRE(in->ld % 8 == 0, S2LC_ERR_ALIGN, "outconv_fwd: row stride must be a multiple of 8");
  cudaStream_t st = (cudaStream_t)stream;
  const int grid = oc_grid(rows, C / 8);
  if (dtype == S2LC_F32) {
    OC_SWITCH(n_classes, (outconv_fwd_kernel<float, NC><<<grid, OC_THREADS, 0, st>>>(*in, w, bias, (float*)logits,
                                                                                      ld_out, rows, C)))
  } else if (dtype == S2LC_BF16) {
    OC_SWITCH(n_classes, (outconv_fwd_kernel<bf16, NC><<<grid, OC_THREADS, 0, st>>>(*in, w, bias, (bf16*)logits,
                                                                                     ld_out, rows, C)))
  } else {
    set_error("outconv_fwd: bad dtype %d", dtype);
    return S2LC_ERR_SHAPE;
  }
  return check_launch("outconv_fwd");
}

int32_t s2lc_outconv_bwd(const s2lc_lazy* in, const void* dlogits, int64_t ld_dl, const float* w, float* dw,
                         float* dbias, void* dz, int64_t ld_dz, double* stats, int64_t rows, int32_t C,
                         int32_t n_classes, int32_t dtype, void* stream) {
  S2LC_REQUIRE(in && dlogits && w && dw && dz, S2LC_ERR_SHAPE, "outconv_bwd: null argument");
  OC_CHECK(C, n_classes)
  S2LC_REQUIRE(in->ld % 8 == 0 && ld_dz % 8 == 0, S2LC_ERR_ALIGN, "outconv_bwd: row strides must be multiples of 8");
  cudaStream_t st = (cudaStream_t)stream;
  const int grid = oc_grid(rows, C / 8) > kNumSMs * 4 ? kNumSMs * 4 : oc_grid(rows, C / 8);
  const size_t smem = sizeof(float) * ((size_t)n_classes * C + 2 * C + n_classes);
  if (dtype == S2LC_F32) {
    OC_SWITCH(n_classes, (outconv_bwd_kernel<float, NC><<<grid, OC_THREADS, smem, st>>>(
                             *in, (const float*)dlogits, ld_dl, w, dw, dbias, (float*)dz, ld_dz, stats, rows, C)))
  } else if (dtype == S2LC_BF16) {
    OC_SWITCH(n_classes, (outconv_bwd_kernel<bf16, NC><<<grid, OC_THREADS, smem, st>>>(
                             *in, (const bf16*)dlogits, ld_dl, w, dw, dbias, (bf16*)dz, ld_dz, stats, rows, C)))
  } else {
    set_error("outconv_bwd: bad dtype %d", dtype);
    return S2LC_ERR_SHAPE;
  }
  return check_launch("outconv_bwd");
}

}  // extern "C"

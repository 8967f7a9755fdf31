// common.cuh — device helpers shared by all s2lc kernels (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/s2lc.h"

namespace s2lc {

typedef __nv_bfloat16 bf16;

void set_error(const char* fmt, ...);
int32_t check_launch(const char* what);

#define S2LC_REQUIRE(cond, code, ...)   \
  do {                                  \
    if (!(cond)) {                      \
      s2lc::set_error(__VA_ARGS__);     \
      return (code);                    \
    }                                   \
  } while (0)

// SM count of the CURRENT device, read once per device from the driver (sizes every persistent grid; a partitioned or
// different-SKU part must not silently get B200's 148).  Host side only.
int num_sms();
// Test / profiling switches (s2lc_set_debug_flags; initialised ONCE from the environment at first use, never read
// per launch): the defaults (0) are the product path.
enum : uint32_t { DBG_NO_STREAM = 1u, DBG_FORCE_SIMT = 2u, DBG_SCONV_PROD7 = 4u, DBG_NO_TMA = 8u, DBG_TMA_ALL = 16u, DBG_DW_CB32 = 32u, DBG_NO_TMA_STORE = 64u, DBG_DW_GENERIC = 128u, DBG_STREAM_GENERIC = 1u << 16 /* bits 8-15 are the strip-width cap */ };
uint32_t debug_flags();          // bits above; bits 8-15 = strip-width cap of the streaming 3x3 weight gradient (0 = none)
#define kNumSMs (::s2lc::num_sms())

// ------------------------------------------------------------------ scalar load/store by storage type
template <typename T> __device__ __forceinline__ float ldf(const T* p);
template <> __device__ __forceinline__ float ldf<float>(const float* p) { return *p; }
template <> __device__ __forceinline__ float ldf<bf16>(const bf16* p) { return __bfloat162float(*p); }
template <typename T> __device__ __forceinline__ void stf(T* p, float v);
template <> __device__ __forceinline__ void stf<float>(float* p, float v) { *p = v; }
template <> __device__ __forceinline__ void stf<bf16>(bf16* p, float v) { *p = __float2bfloat16_rn(v); }
// value as it will be read back from storage (used so that statistics see the stored, rounded value)
template <typename T> __device__ __forceinline__ float roundf_as(float v);
template <> __device__ __forceinline__ float roundf_as<float>(float v) { return v; }
template <> __device__ __forceinline__ float roundf_as<bf16>(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }

// 8-channel vector access (16 B for bf16, 32 B for fp32); pointers must be 16 B aligned
template <typename T> struct Vec8;
template <> struct Vec8<bf16> {
  static __device__ __forceinline__ void load(const bf16* p, float (&v)[8]) {
    uint4 u = *reinterpret_cast<const uint4*>(p);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float2 f = __bfloat1622float2(h[i]);
      v[2 * i] = f.x;
      v[2 * i + 1] = f.y;
    }
  }
  static __device__ __forceinline__ void store(bf16* p, const float (&v)[8]) {
    uint4 u;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
    *reinterpret_cast<uint4*>(p) = u;
  }
};
template <> struct Vec8<float> {
  static __device__ __forceinline__ void load(const float* p, float (&v)[8]) {
    float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
  static __device__ __forceinline__ void store(float* p, const float (&v)[8]) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
  }
};

// ------------------------------------------------------------------ programmatic dependent launch
// A kernel launched with this configuration is staged while its predecessor in the stream drains (the launch latency
// and the stream-serialisation gap, ~2.4 us per launch measured, are hidden); it MUST call
// cudaGridDependencySynchronize() before it touches global memory.  No kernel here triggers early, so the dependent
// grid starts only after every block of the predecessor has exited.  (An early cudaTriggerProgrammaticLaunchCompletion()
// at kernel entry was measured and is SLOWER, 54.5 -> 56.5 ms per step: the staged blocks wait on the SMs and take
// the slots of the grid they wait for.)
static cudaLaunchAttribute g_pdl_attr;
static inline cudaLaunchConfig_t pdl_config(dim3 grid, int block, cudaStream_t st, size_t smem = 0) {
  g_pdl_attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
  g_pdl_attr.val.programmaticStreamSerializationAllowed = 1;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = dim3(block);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cfg.attrs = &g_pdl_attr;
  cfg.numAttrs = 1;
  return cfg;
}


#define PDL_LAUNCH(kern, grid, block, smem, st, ...)                      \
  do {                                                                    \
    cudaLaunchConfig_t cfg_ = s2lc::pdl_config(grid, block, st, smem);    \
    cudaLaunchKernelEx(&cfg_, kern, __VA_ARGS__);                         \
  } while (0)

// ------------------------------------------------------------------ activations
// exact form (expf, IEEE division): the fp32 verification engines and the tiny squeeze-excite FCs; the bf16-storage
// streaming kernels use the one-MUFU forms below through act_fwd_t / act_bwd_t
__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + expf(-x)); }
__device__ __forceinline__ float act_fwd(float u, int act) {
  if (act == S2LC_ACT_SILU) return u * sigmoidf_(u);
  if (act == S2LC_ACT_RELU) return fmaxf(u, 0.f);
  return u;
}
__device__ __forceinline__ float act_bwd(float u, int act) {  // d act(u) / du
  if (act == S2LC_ACT_SILU) {
    float s = sigmoidf_(u);
    return s * (1.f + u * (1.f - s));
  }
  if (act == S2LC_ACT_RELU) return u > 0.f ? 1.f : 0.f;
  return 1.f;
}

// bf16-storage engines: sigmoid through ONE MUFU (tanh.approx, relative error ~2^-11 — far below the bf16 rounding of
// the tensors these values are stored to) instead of ex2 + rcp; the fp32 verification engines keep the exact form.
__device__ __forceinline__ float tanh_approx(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float silu_fast(float u) {
  const float h = 0.5f * u;
  return fmaf(h, tanh_approx(h), h);
}
__device__ __forceinline__ float act_fwd_fast(float u, int act) {
  if (act == S2LC_ACT_SILU) return silu_fast(u);
  if (act == S2LC_ACT_RELU) return fmaxf(u, 0.f);
  return u;
}
__device__ __forceinline__ float act_bwd_fast(float u, int act) {
  if (act == S2LC_ACT_SILU) {
    const float s = fmaf(0.5f, tanh_approx(0.5f * u), 0.5f);
    return s * fmaf(u, 1.f - s, 1.f);
  }
  if (act == S2LC_ACT_RELU) return u > 0.f ? 1.f : 0.f;
  return 1.f;
}

// packed fp32 pairs (FFMA2 / FMUL2 / FADD2: one issue slot for two lanes of math — the streaming kernels are issue-bound)
__device__ __forceinline__ float2 f2fma(const float2& a, const float2& b, const float2& c) {  // a*b + c
  float2 d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;"
      : "=l"(reinterpret_cast<unsigned long long&>(d))
      : "l"(reinterpret_cast<const unsigned long long&>(a)), "l"(reinterpret_cast<const unsigned long long&>(b)),
        "l"(reinterpret_cast<const unsigned long long&>(c)));
  return d;
}
__device__ __forceinline__ float2 f2mul(const float2& a, const float2& b) {
  float2 d;
  asm("mul.rn.f32x2 %0, %1, %2;"
      : "=l"(reinterpret_cast<unsigned long long&>(d))
      : "l"(reinterpret_cast<const unsigned long long&>(a)), "l"(reinterpret_cast<const unsigned long long&>(b)));
  return d;
}
__device__ __forceinline__ float2 f2add(const float2& a, const float2& b) {
  float2 d;
  asm("add.rn.f32x2 %0, %1, %2;"
      : "=l"(reinterpret_cast<unsigned long long&>(d))
      : "l"(reinterpret_cast<const unsigned long long&>(a)), "l"(reinterpret_cast<const unsigned long long&>(b)));
  return d;
}
__device__ __forceinline__ float2 bf2_unpack(uint32_t w) {
  return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u));
}
__device__ __forceinline__ uint32_t bf2_pack(const float2& v) {
  __nv_bfloat162 h = __floats2bfloat162_rn(v.x, v.y);
  return *reinterpret_cast<uint32_t*>(&h);
}

// by storage type: bf16 -> one-MUFU forms, fp32 -> exact forms
template <typename T> __device__ __forceinline__ float act_fwd_t(float u, int act) {
  return sizeof(T) == 2 ? act_fwd_fast(u, act) : act_fwd(u, act);
}
template <typename T> __device__ __forceinline__ float act_bwd_t(float u, int act) {
  return sizeof(T) == 2 ? act_bwd_fast(u, act) : act_bwd(u, act);
}

// ------------------------------------------------------------------ lazy sources, element-wise (generic path)
// value(row, c) of a forward lazy activation; n = sample index of the row, C = channel count (gate row stride)
template <typename T>
__device__ __forceinline__ float lazy_at(const s2lc_lazy& L, int64_t row, int c, int n, int C) {
  float v = ldf<T>(reinterpret_cast<const T*>(L.raw) + row * L.ld + c);
  if (L.scale) v = v * L.scale[c] + (L.shift ? L.shift[c] : 0.f);
  v = act_fwd(v, L.act);
  if (L.gate) v *= L.gate[(int64_t)n * C + c];
  return v;
}
// pre-activation u(row, c) = raw*scale+shift of a lazy activation (for act')
template <typename T>
__device__ __forceinline__ float lazy_pre(const s2lc_lazy& L, int64_t row, int c) {
  float v = ldf<T>(reinterpret_cast<const T*>(L.raw) + row * L.ld + c);
  if (L.scale) v = v * L.scale[c] + (L.shift ? L.shift[c] : 0.f);
  return v;
}
template <typename T>
__device__ __forceinline__ float lazygrad_at(const s2lc_lazygrad& G, int64_t row, int c, int n) {
  float v = ldf<T>(reinterpret_cast<const T*>(G.dy) + row * G.ld_dy + c);
  if (G.rowscale) v *= G.rowscale[n];
  if (G.z) v = G.p[c] * v + G.q[c] * ldf<T>(reinterpret_cast<const T*>(G.z) + row * G.ld_z + c) + G.t[c];
  return v;
}

// ------------------------------------------------------------------ reductions
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

}  // namespace s2lc

// dwconv2.cuh — kernels and launchers of the streaming depthwise convolution (see dwconv2.cu for the design notes).
#pragma once
#include "common.cuh"

namespace s2lc {
namespace dw2 {

// Channels per CTA block are a template parameter CB in {64, 32}.  CB = 64: lane = channel pair, a warp row is one
// pixel.  CB = 32: 16 lanes per pixel and the two half-warps own different pixels of the strip (strip = 16R wide), so
// layers with 24 / 144 channels fill 75 % / 90 % of the lanes instead of 37 % / 75 %.
constexpr int NT = 256;  // threads per CTA

__device__ __forceinline__ void ffma2(float2& d, const float2& a, const float2& b) {
  asm("fma.rn.f32x2 %0, %1, %2, %0;"
      : "+l"(reinterpret_cast<unsigned long long&>(d))
      : "l"(reinterpret_cast<const unsigned long long&>(a)), "l"(reinterpret_cast<const unsigned long long&>(b)));
}
__device__ __forceinline__ float tanh_fast(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// FAST (bf16 storage): sigmoid via one MUFU.TANH, error ~2^-11, far below the bf16 rounding of the stored tensors
template <bool FAST>
__device__ __forceinline__ float act_f(float u, int act) {
  if (act == S2LC_ACT_SILU) {
    if (FAST) {
      float h = 0.5f * u;
      return fmaf(h, tanh_fast(h), h);
    }
    return u * sigmoidf_(u);
  }
  if (act == S2LC_ACT_RELU) return fmaxf(u, 0.f);
  return u;
}
template <bool FAST>
__device__ __forceinline__ float act_d(float u, int act) {
  if (act == S2LC_ACT_SILU) {
    float s = FAST ? fmaf(0.5f, tanh_fast(0.5f * u), 0.5f) : sigmoidf_(u);
    return s * (1.f + u * (1.f - s));
  }
  if (act == S2LC_ACT_RELU) return u > 0.f ? 1.f : 0.f;
  return 1.f;
}

constexpr int floordiv(int a, int b) { return (a >= 0) ? a / b : -((-a + b - 1) / b); }

// channel pair of a destination-side tensor
template <typename T> struct Pair;
template <> struct Pair<bf16> {
  uint32_t u;
  __device__ __forceinline__ void load(const bf16* p) { u = __ldg(reinterpret_cast<const uint32_t*>(p)); }
  __device__ __forceinline__ float2 get() const {
    return make_float2(__uint_as_float(u << 16), __uint_as_float(u & 0xffff0000u));
  }
  __device__ __forceinline__ void zero() { u = 0; }
};
template <> struct Pair<float> {
  float2 f;
  __device__ __forceinline__ void load(const float* p) { f = __ldg(reinterpret_cast<const float2*>(p)); }
  __device__ __forceinline__ float2 get() const { return f; }
  __device__ __forceinline__ void zero() { f = make_float2(0.f, 0.f); }
};
template <typename T> __device__ __forceinline__ void st_pair(T* p, float a, float b);
template <> __device__ __forceinline__ void st_pair<bf16>(bf16* p, float a, float b) {
  *reinterpret_cast<__nv_bfloat162*>(p) = __floats2bfloat162_rn(a, b);
}
template <> __device__ __forceinline__ void st_pair<float>(float* p, float a, float b) {
  *reinterpret_cast<float2*>(p) = make_float2(a, b);
}

// store a channel pair and return it as it will be read back from storage
template <typename T> __device__ __forceinline__ float2 st_pair_rt(T* p, float2 v);
template <> __device__ __forceinline__ float2 st_pair_rt<bf16>(bf16* p, float2 v) {
  __nv_bfloat162 h = __floats2bfloat162_rn(v.x, v.y);
  *reinterpret_cast<__nv_bfloat162*>(p) = h;
  uint32_t u = *reinterpret_cast<uint32_t*>(&h);
  return make_float2(__uint_as_float(u << 16), __uint_as_float(u & 0xffff0000u));
}
template <> __device__ __forceinline__ float2 st_pair_rt<float>(float* p, float2 v) {
  *reinterpret_cast<float2*>(p) = v;
  return v;
}

// ---------------------------------------------------------------------------------------------------------
// Ring loader.  issue(): rows [rel0, rel0+nrows) of a WW-pixel x 64-channel window of the source tensor(s) are
// fetched with cp.async (16 B, L1-bypassing) into a raw STAGE in shared memory where every thread owns private
// slots (so no barrier is needed between arrival and use, only cp.async.wait_group); several stages are in
// flight at once.  commit(): the owning thread reads its slots back, applies the transform and writes fp32 into
// ring slot (rel % NRING).  NSRC = 1: forward activation (aux = scale, shift); NSRC = 2: lazy gradient
// (aux = p, q, t; second source = z).
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)),
               "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

template <typename T> struct Raw8;
template <> struct Raw8<bf16> {
  static __device__ __forceinline__ void unpack(const unsigned char* p, float (&v)[8]) {
    uint4 u = *reinterpret_cast<const uint4*>(p);
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      v[2 * i] = __uint_as_float(w[i] << 16);
      v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
  }
};
template <> struct Raw8<float> {
  static __device__ __forceinline__ void unpack(const unsigned char* p, float (&v)[8]) {
    float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 16);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
};

__device__ __forceinline__ float2 fmul2(const float2& a, const float2& b) {
  float2 d;
  asm("mul.rn.f32x2 %0, %1, %2;"
      : "=l"(reinterpret_cast<unsigned long long&>(d))
      : "l"(reinterpret_cast<const unsigned long long&>(a)), "l"(reinterpret_cast<const unsigned long long&>(b)));
  return d;
}
__device__ __forceinline__ float2 ffma2r(const float2& a, const float2& b, const float2& c) {  // a*b + c
  float2 d = c;
  ffma2(d, a, b);
  return d;
}

// Thread <-> vector mapping of a pass (NEW rows x WW pixels x 8 vectors of 8 channels):
//  * WIDE (WW >= 32): thread (px = tid/8, v8 = tid%8) owns the same pixel column in every row (slot j = row j, no
//    divisions, one bounds check per item) and the WW-32 halo columns go to ONE extra slot of the low threads;
//  * narrow windows: generic q = tid + j*256 -> (row, px) by constant division.
template <typename T, int NSRC, int WW, int NEW, int NRING, bool FAST, int CB>
struct Loader {
  static constexpr int CBV = CB / 8;                         // 8-channel vectors per pixel
  static constexpr int MC = NT / CBV;                        // main pixel columns: one per CBV threads
  static constexpr bool WIDE = WW >= MC;
  static constexpr int XC = WIDE ? WW - MC : 0;              // halo columns beyond the main ones
  static constexpr int NX = NEW * XC * CBV;                  // their vectors per pass
  static constexpr int NXT = NX > 0 ? 1 : 0;
  static_assert(NX <= NT, "halo columns of a pass must fit one slot");
  static constexpr int NVT = WIDE ? NEW + NXT : (NEW * WW * CBV + NT - 1) / NT;
  static constexpr int VB = 8 * sizeof(T);                   // bytes per 8-channel vector
  // main slots: one 16 B (bf16) vector per thread, row and source; the halo slot is compact (NX threads use it)
  static constexpr int MAIN_SLOTS = WIDE ? NEW : NVT;
  static constexpr int STAGE_BYTES = MAIN_SLOTS * NSRC * NT * VB + (WIDE ? NX * NSRC * VB : 0);

  // per-item state
  const T* pm[2];       // main column: pointer to (n, row 0, gx, c8) of each source
  const T* pe[2];       // halo column of this thread
  int64_t rstride[2];   // elements per source row
  int gy0, Hs;
  bool okm, oke;
  int erow, epx;        // halo slot: row inside the pass, pixel column
  // generic path
  int gx0, Ws, n, cb0, C;
  int64_t ld[2];
  const T* base[2];

  __device__ __forceinline__ void begin_item(const T* s0, int64_t ld0, const T* s1, int64_t ld1, int n_, int Hs_,
                                             int Ws_, int gy0_, int gx0_, int cb0_, int C_) {
    gy0 = gy0_; Hs = Hs_;
    const int c8 = cb0_ + (threadIdx.x % CBV) * 8;
    if (WIDE) {
      const int px = threadIdx.x / CBV;
      const int gx = gx0_ + px;
      okm = gx >= 0 && gx < Ws_ && c8 < C_;
      const int64_t img = (int64_t)n_ * Hs_ * Ws_;
      pm[0] = s0 + (img + gx) * ld0 + c8;
      pm[1] = s1 + (img + gx) * ld1 + c8;
      rstride[0] = (int64_t)Ws_ * ld0;
      rstride[1] = (int64_t)Ws_ * ld1;
      if (NXT) {
        erow = threadIdx.x / (XC * CBV > 0 ? XC * CBV : 1);
        epx = MC + ((threadIdx.x - erow * (XC * CBV)) / CBV);
        const int gxe = gx0_ + epx;
        oke = (int)threadIdx.x < NX && gxe >= 0 && gxe < Ws_ && c8 < C_;
        pe[0] = s0 + (img + gxe) * ld0 + c8;
        pe[1] = s1 + (img + gxe) * ld1 + c8;
      }
    } else {
      gx0 = gx0_; Ws = Ws_; n = n_; cb0 = cb0_; C = C_;
      ld[0] = ld0; ld[1] = ld1; base[0] = s0; base[1] = s1;
    }
  }

  static __device__ __forceinline__ unsigned char* slot(unsigned char* stage, int j, int s) {
    if (WIDE && j == NEW) return stage + ((size_t)NEW * NSRC * NT + s * NX + threadIdx.x) * VB;
    return stage + ((size_t)(j * NSRC + s) * NT + threadIdx.x) * VB;
  }
  static __device__ __forceinline__ void fetch(unsigned char* stage, int j, const T* a, const T* b) {
    cp_async16(slot(stage, j, 0), a);
    if (sizeof(T) == 4) cp_async16(slot(stage, j, 0) + 16, a + 4);
    if (NSRC == 2) {
      cp_async16(slot(stage, j, NSRC - 1), b);
      if (sizeof(T) == 4) cp_async16(slot(stage, j, NSRC - 1) + 16, b + 4);
    }
  }

  __device__ __forceinline__ void issue(unsigned char* stage, int rel0, int nrows, bool do_commit = true) const {
    if (WIDE) {
#pragma unroll
      for (int j = 0; j < NEW; ++j) {
        const int gy = gy0 + rel0 + j;
        if (j < nrows && okm && gy >= 0 && gy < Hs) fetch(stage, j, pm[0] + gy * rstride[0], pm[1] + gy * rstride[1]);
      }
      if (NXT) {
        const int gy = gy0 + rel0 + erow;
        if (oke && erow < nrows && gy >= 0 && gy < Hs) fetch(stage, NEW, pe[0] + gy * rstride[0], pe[1] + gy * rstride[1]);
      }
    } else {
      const int c8 = cb0 + (threadIdx.x % CBV) * 8;
#pragma unroll
      for (int j = 0; j < NVT; ++j) {
        int q = threadIdx.x + j * NT;
        int row = q / (WW * CBV);
        int px = (q - row * (WW * CBV)) / CBV;
        int gy = gy0 + rel0 + row, gx = gx0 + px;
        if (row < nrows && gy >= 0 && gy < Hs && gx >= 0 && gx < Ws && c8 < C) {
          int64_t r = ((int64_t)n * Hs + gy) * Ws + gx;
          fetch(stage, j, base[0] + r * ld[0] + c8, base[1] + r * ld[1] + c8);
        }
      }
    }
    if (do_commit) cp_async_commit();
  }

  // transform one staged vector (or zeros) and store it to the fp32 ring
  // mode 0: v = act(v*aux0 + aux1)  [FAST SiLU: aux pre-scaled by 1/2, v = h*tanh(h)+h];
  // mode 1: v = (aux0*rs) * v + aux1 * z + aux2;  mode 2: v = v * rs
  static __device__ __forceinline__ void xform_store(const unsigned char* stage, int j, bool ok, float* dst,
                                                     const float* aux, int mode, int act, float rs) {
    const int v8 = threadIdx.x % CBV;
    float v[8];
    if (ok) {
      Raw8<T>::unpack(slot(const_cast<unsigned char*>(stage), j, 0), v);
      const float4* a0 = reinterpret_cast<const float4*>(aux + v8 * 8);
      const float4* a1 = reinterpret_cast<const float4*>(aux + CB + v8 * 8);
      float4 s0 = a0[0], s1 = a0[1], h0 = a1[0], h1 = a1[1];
      const float2 A[4] = {{s0.x, s0.y}, {s0.z, s0.w}, {s1.x, s1.y}, {s1.z, s1.w}};
      const float2 B[4] = {{h0.x, h0.y}, {h0.z, h0.w}, {h1.x, h1.y}, {h1.z, h1.w}};
      if (mode == 0) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          float2 u = ffma2r(make_float2(v[2 * i], v[2 * i + 1]), A[i], B[i]);
          if (FAST && act == S2LC_ACT_SILU) {
            float2 t = make_float2(tanh_fast(u.x), tanh_fast(u.y));
            u = ffma2r(u, t, u);
          } else {
            u.x = act_f<false>(u.x, act);
            u.y = act_f<false>(u.y, act);
          }
          v[2 * i] = u.x; v[2 * i + 1] = u.y;
        }
      } else if (mode == 1) {
        float z[8];
        Raw8<T>::unpack(slot(const_cast<unsigned char*>(stage), j, NSRC - 1), z);
        const float4* a2 = reinterpret_cast<const float4*>(aux + 2 * CB + v8 * 8);
        float4 t0 = a2[0], t1 = a2[1];
        const float2 Tt[4] = {{t0.x, t0.y}, {t0.z, t0.w}, {t1.x, t1.y}, {t1.z, t1.w}};
        const float2 rs2 = make_float2(rs, rs);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          float2 u = ffma2r(B[i], make_float2(z[2 * i], z[2 * i + 1]), Tt[i]);
          u = ffma2r(fmul2(A[i], rs2), make_float2(v[2 * i], v[2 * i + 1]), u);
          v[2 * i] = u.x; v[2 * i + 1] = u.y;
        }
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] *= rs;
      }
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = 0.f;
    }
    float4* d4 = reinterpret_cast<float4*>(dst + v8 * 8);
    d4[0] = make_float4(v[0], v[1], v[2], v[3]);
    d4[1] = make_float4(v[4], v[5], v[6], v[7]);
  }

  // rbase = ring slot of row rel0
  __device__ __forceinline__ void commit(const unsigned char* stage, float* tile, const float* aux, int rel0, int rbase,
                                         int nrows, int mode, int act, float rs) const {
    if (WIDE) {
      const int px = threadIdx.x / CBV;
#pragma unroll
      for (int j = 0; j < NEW; ++j) {
        if (j >= nrows) break;
        const int gy = gy0 + rel0 + j;
        int sr = rbase + j;
        if (sr >= NRING) sr -= NRING;
        xform_store(stage, j, okm && gy >= 0 && gy < Hs, tile + ((size_t)sr * WW + px) * CB, aux, mode, act, rs);
      }
      if (NXT) {
        if ((int)threadIdx.x < NX && erow < nrows) {
          const int gy = gy0 + rel0 + erow;
          int sr = rbase + erow;
          if (sr >= NRING) sr -= NRING;
          xform_store(stage, NEW, oke && gy >= 0 && gy < Hs, tile + ((size_t)sr * WW + epx) * CB, aux, mode, act, rs);
        }
      }
    } else {
      const int c8 = cb0 + (threadIdx.x % CBV) * 8;
#pragma unroll
      for (int j = 0; j < NVT; ++j) {
        int q = threadIdx.x + j * NT;
        int row = q / (WW * CBV);
        int px = (q - row * (WW * CBV)) / CBV;
        if (row >= nrows) continue;
        int gy = gy0 + rel0 + row, gx = gx0 + px;
        bool ok = gy >= 0 && gy < Hs && gx >= 0 && gx < Ws && c8 < C;
        int sr = rbase + row;
        if (sr >= NRING) sr -= NRING;
        xform_store(stage, j, ok, tile + ((size_t)sr * WW + px) * CB, aux, mode, act, rs);
      }
    }
  }

  // ---- compile-time variants (fast bf16 kernels): MODE 0 = SiLU(v*aux0+aux1) with aux pre-scaled by 1/2,
  // MODE 1 = aux0*v + aux1*z + aux2 (aux0 already holds p*rowscale), MODE 2 = aux0*v (aux0 = rowscale)
  template <int MODE>
  static __device__ __forceinline__ void xform_store_ct(const unsigned char* stage, int j, bool ok, float* dst,
                                                        const float* aux) {
    const int v8 = threadIdx.x % CBV;
    float v[8];
    if (ok) {
      Raw8<T>::unpack(slot(const_cast<unsigned char*>(stage), j, 0), v);
      const float4* a0 = reinterpret_cast<const float4*>(aux + v8 * 8);
      float4 s0 = a0[0], s1 = a0[1];
      const float2 A[4] = {{s0.x, s0.y}, {s0.z, s0.w}, {s1.x, s1.y}, {s1.z, s1.w}};
      if (MODE == 2) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          float2 u = fmul2(make_float2(v[2 * i], v[2 * i + 1]), A[i]);
          v[2 * i] = u.x; v[2 * i + 1] = u.y;
        }
      } else {
        const float4* a1 = reinterpret_cast<const float4*>(aux + CB + v8 * 8);
        float4 h0 = a1[0], h1 = a1[1];
        const float2 B[4] = {{h0.x, h0.y}, {h0.z, h0.w}, {h1.x, h1.y}, {h1.z, h1.w}};
        if (MODE == 0) {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            float2 u = ffma2r(make_float2(v[2 * i], v[2 * i + 1]), A[i], B[i]);
            float2 t = make_float2(tanh_fast(u.x), tanh_fast(u.y));
            u = ffma2r(u, t, u);
            v[2 * i] = u.x; v[2 * i + 1] = u.y;
          }
        } else {
          float z[8];
          Raw8<T>::unpack(slot(const_cast<unsigned char*>(stage), j, NSRC - 1), z);
          const float4* a2 = reinterpret_cast<const float4*>(aux + 2 * CB + v8 * 8);
          float4 t0 = a2[0], t1 = a2[1];
          const float2 Tt[4] = {{t0.x, t0.y}, {t0.z, t0.w}, {t1.x, t1.y}, {t1.z, t1.w}};
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            float2 u = ffma2r(B[i], make_float2(z[2 * i], z[2 * i + 1]), Tt[i]);
            u = ffma2r(A[i], make_float2(v[2 * i], v[2 * i + 1]), u);
            v[2 * i] = u.x; v[2 * i + 1] = u.y;
          }
        }
      }
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = 0.f;
    }
    float4* d4 = reinterpret_cast<float4*>(dst + v8 * 8);
    d4[0] = make_float4(v[0], v[1], v[2], v[3]);
    d4[1] = make_float4(v[4], v[5], v[6], v[7]);
  }

  // nrows == NEW unless PARTIAL
  template <int MODE, bool PARTIAL>
  __device__ __forceinline__ void commit_ct(const unsigned char* stage, float* tile, const float* aux, int rel0,
                                            int rbase, int nrows) const {
    if (WIDE) {
      const int px = threadIdx.x / CBV;
#pragma unroll
      for (int j = 0; j < NEW; ++j) {
        if (PARTIAL && j >= nrows) break;
        const int gy = gy0 + rel0 + j;
        int sr = rbase + j;
        if (sr >= NRING) sr -= NRING;
        xform_store_ct<MODE>(stage, j, okm && gy >= 0 && gy < Hs, tile + ((size_t)sr * WW + px) * CB, aux);
      }
      if (NXT) {
        if ((int)threadIdx.x < NX && (!PARTIAL || erow < nrows)) {
          const int gy = gy0 + rel0 + erow;
          int sr = rbase + erow;
          if (sr >= NRING) sr -= NRING;
          xform_store_ct<MODE>(stage, NEW, oke && gy >= 0 && gy < Hs, tile + ((size_t)sr * WW + epx) * CB, aux);
        }
      }
    } else {
      const int c8 = cb0 + (threadIdx.x % CBV) * 8;
#pragma unroll
      for (int j = 0; j < NVT; ++j) {
        int q = threadIdx.x + j * NT;
        int row = q / (WW * CBV);
        int px = (q - row * (WW * CBV)) / CBV;
        if (row >= nrows) continue;
        int gy = gy0 + rel0 + row, gx = gx0 + px;
        bool ok = gy >= 0 && gy < Hs && gx >= 0 && gx < Ws && c8 < C;
        int sr = rbase + row;
        if (sr >= NRING) sr -= NRING;
        xform_store_ct<MODE>(stage, j, ok, tile + ((size_t)sr * WW + px) * CB, aux);
      }
    }
  }
};

struct Geom {
  int N, C;
  int Hs, Ws;        // source tensor (staged in the ring)
  int Hd, Wd;        // destination tensor (written)
  int pad_t, pad_l;  // forward TF-"same" padding
  int xs, ychunks, YC;
};

template <int CB>
__device__ __forceinline__ void flush_stats(float2 s0, float2 s1, float* sstat, int cb0, int C, double* stats) {
  // sstat: [2][CB] floats, zeroed at kernel start
  const int lane = (threadIdx.x & 31) % (CB / 2);
  atomicAdd(&sstat[2 * lane], s0.x);
  atomicAdd(&sstat[2 * lane + 1], s0.y);
  atomicAdd(&sstat[CB + 2 * lane], s1.x);
  atomicAdd(&sstat[CB + 2 * lane + 1], s1.y);
  __syncthreads();
  if (threadIdx.x < 2 * CB) {
    int s = threadIdx.x / CB, c = cb0 + threadIdx.x % CB;
    if (c < C) atomicAdd(&stats[(int64_t)s * C + c], (double)sstat[threadIdx.x]);
  }
}

// ---------------------------------------------------------------------------------------------------------
// Forward: out[o] = sum_k w[k] * act(bn(in))[o*S + k - pad]; epilogue: raw output + BatchNorm moments.
// R = pixels per warp (strip = 8R wide), RY = rows per step, D = cp.async passes in flight.
template <typename T, int K, int S, int R, int RY, int D, int CB>
struct FwdCfg {
  static constexpr int NH = 64 / CB;               // pixel groups per warp (half-warps own different pixels)
  static constexpr int TW = 8 * R * NH;            // destination pixels per strip (8 warps x NH x R)
  static constexpr int WW = (TW - 1) * S + K;      // staged window width
  static constexpr int WIN = (RY - 1) * S + K;     // staged window height per step
  static constexpr int NEW = RY * S;               // new source rows per step
  static constexpr int NRING = WIN;
  static constexpr int VW = (R - 1) * S + K;       // window width per warp
  static constexpr int PRO = WIN - NEW;            // rows staged before the first step
  static constexpr int PRO_PASSES = (PRO + NEW - 1) / NEW;
  static constexpr int NSTAGE = D + PRO_PASSES;
  using Ld = Loader<T, 1, WW, NEW, NRING, sizeof(T) == 2, CB>;
  static constexpr size_t RING_BYTES = sizeof(float) * (size_t)NRING * WW * CB;
  static constexpr size_t SMEM = RING_BYTES + sizeof(float) * 4 * CB + (size_t)NSTAGE * Ld::STAGE_BYTES;
};

template <typename T, int K, int S, int R, int RY, int D, int MINB, int CB>
__global__ void __launch_bounds__(NT, MINB) dw2_fwd_kernel(s2lc_lazy in, const float* __restrict__ w,
                                                           T* __restrict__ out, double* stats, Geom g) {
  cudaGridDependencySynchronize();   // programmatic dependent launch (common.cuh)
  using Cfg = FwdCfg<T, K, S, R, RY, D, CB>;
  using Ld = typename Cfg::Ld;
  constexpr int WW = Cfg::WW, WIN = Cfg::WIN, NEW = Cfg::NEW, NRING = Cfg::NRING, VW = Cfg::VW, PRO = Cfg::PRO,
                PRO_PASSES = Cfg::PRO_PASSES, NSTAGE = Cfg::NSTAGE;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* tile = reinterpret_cast<float*>(smem_raw);
  float* aux = tile + (size_t)NRING * WW * CB;  // [2][CB] scale, shift
  float* sstat = aux + 2 * CB;                   // [2][CB]
  unsigned char* stages = reinterpret_cast<unsigned char*>(sstat + 2 * CB);
  constexpr int HL = CB / 2, NH = Cfg::NH;              // lanes per pixel group, pixel groups per warp
  const int lane = (threadIdx.x & 31) % HL, pg = (threadIdx.x >> 5) * NH + (threadIdx.x & 31) / HL;
  const int cb0 = blockIdx.y * CB;
  const int c = cb0 + 2 * lane;
  const bool cvalid = c < g.C;
  if (threadIdx.x < CB) {
    int cc = cb0 + threadIdx.x;
    const float f = (sizeof(T) == 2 && in.act == S2LC_ACT_SILU) ? 0.5f : 1.f;   // h = u/2 for the tanh form of SiLU
    aux[threadIdx.x] = f * ((cc < g.C && in.scale) ? in.scale[cc] : 1.f);
    aux[CB + threadIdx.x] = f * ((cc < g.C && in.shift) ? in.shift[cc] : 0.f);
  }
  if (threadIdx.x < 2 * CB) sstat[threadIdx.x] = 0.f;
  float2 wr[K * K];
#pragma unroll
  for (int t = 0; t < K * K; ++t)
    wr[t] = cvalid ? make_float2(w[(int64_t)c * K * K + t], w[(int64_t)(c + 1) * K * K + t]) : make_float2(0.f, 0.f);
  float2 s0 = make_float2(0.f, 0.f), s1 = make_float2(0.f, 0.f);
  const T* src = reinterpret_cast<const T*>(in.raw);
  Ld ld;
  const int items = g.N * g.ychunks * g.xs;
  int iss = 0, com = 0;  // stage cursors (issue / commit), both advance once per pass
  auto next = [](int& x) { int r = x; x = (x + 1 == NSTAGE) ? 0 : x + 1; return r; };

  for (int item = blockIdx.x; item < items; item += gridDim.x) {
    int n = item / (g.ychunks * g.xs);
    int rem = item - n * g.ychunks * g.xs;
    int y0 = (rem / g.xs) * g.YC, x0 = (rem % g.xs) * Cfg::TW;
    int yend = min(y0 + g.YC, g.Hd);
    int nsteps = (yend - y0 + RY - 1) / RY;
    const int gy0 = y0 * S - g.pad_t, gx0 = x0 * S - g.pad_l;
    ld.begin_item(src, in.ld, src, in.ld, n, g.Hs, g.Ws, gy0, gx0, cb0, g.C);
    auto issue = [&](int rel0, int nr) { ld.issue(stages + (size_t)next(iss) * Ld::STAGE_BYTES, rel0, nr); };
    auto commit = [&](int rel0, int rb, int nr) {
      ld.commit(stages + (size_t)next(com) * Ld::STAGE_BYTES, tile, aux, rel0, rb, nr, 0, in.act, 1.f);
    };
    // prologue: the PRO leading rows, then steps 0 .. D-1 in flight
#pragma unroll
    for (int pp = 0; pp < PRO_PASSES; ++pp) issue(pp * NEW, min(NEW, PRO - pp * NEW));
#pragma unroll
    for (int d = 0; d < D; ++d) {
      if (d < nsteps) issue(PRO + d * NEW, NEW);
      else cp_async_commit();
    }
    cp_async_wait<D - 1>();
    __syncthreads();  // previous item's readers are done with the ring; aux/sstat initialised
#pragma unroll
    for (int pp = 0; pp < PRO_PASSES; ++pp) commit(pp * NEW, (pp * NEW) % NRING, min(NEW, PRO - pp * NEW));
    commit(PRO, PRO % NRING, NEW);
    int cslot = (PRO + NEW) % NRING;  // ring slot the next pass is committed to
    __syncthreads();
    int base = 0;  // ring slot of the first window row of this step
    for (int t = 0; t < nsteps; ++t) {
      if (t + D < nsteps) issue(PRO + (t + D) * NEW, NEW);
      else cp_async_commit();
      float2 acc[RY][R];
#pragma unroll
      for (int dy = 0; dy < RY; ++dy)
#pragma unroll
        for (int r = 0; r < R; ++r) acc[dy][r] = make_float2(0.f, 0.f);
      const float* tl = tile + (pg * R * S) * CB + 2 * lane;
#pragma unroll
      for (int j = 0; j < WIN; ++j) {
        int slot = base + j;
        if (slot >= NRING) slot -= NRING;
        const float* rp = tl + (size_t)slot * WW * CB;
        float2 v[VW];
#pragma unroll
        for (int i = 0; i < VW; ++i) v[i] = *reinterpret_cast<const float2*>(rp + i * CB);
#pragma unroll
        for (int dy = 0; dy < RY; ++dy) {
          const int ky = j - dy * S;
          if (ky < 0 || ky >= K) continue;
#pragma unroll
          for (int r = 0; r < R; ++r)
#pragma unroll
            for (int kx = 0; kx < K; ++kx) ffma2(acc[dy][r], wr[ky * K + kx], v[r * S + kx]);
        }
      }
      if (cvalid) {
        const int ox0 = x0 + pg * R;
        const float2 one2 = make_float2(1.f, 1.f);
#pragma unroll
        for (int dy = 0; dy < RY; ++dy) {
          int oy = y0 + t * RY + dy;
          if (oy >= yend) break;
          T* op = out + (((int64_t)n * g.Hd + oy) * g.Wd + ox0) * g.C + c;
#pragma unroll
          for (int r = 0; r < R; ++r) {
            if (ox0 + r >= g.Wd) break;
            float2 ab = st_pair_rt<T>(op + (int64_t)r * g.C, acc[dy][r]);   // stored (rounded) values
            ffma2(s0, ab, one2);
            ffma2(s1, ab, ab);
          }
        }
      }
      base += NEW;
      if (base >= NRING) base -= NRING;
      if (t + 1 < nsteps) {
        cp_async_wait<D - 1>();   // pass of step t+1 has landed in this thread's slots
        __syncthreads();          // all warps finished reading the rows that are about to be replaced
        commit(PRO + (t + 1) * NEW, cslot, NEW);
        cslot += NEW;
        if (cslot >= NRING) cslot -= NRING;
        __syncthreads();
      }
    }
  }
  if (stats) {
    __syncthreads();
    flush_stats<CB>(s0, s1, sstat, cb0, g.C, stats);
  }
}

// ---------------------------------------------------------------------------------------------------------
// Fused backward, tiled over INPUT pixels i (destination = input-resolution tensor, source ring = lazy gradient g
// at output resolution):  dx[i] = act'(u[i]) * sum_k w[k] g[(i+PAD-k)/S]  (+ residual),
//                         dW[k] += act(u[i]) * g[(i+PAD-k)/S],   taps with (i+PAD-k) % S == 0 only.
// PAD is the forward top/left padding: (K-1)/2 for S=1, (K-2)/2 for S=2 with even H, W (host-checked).
template <typename T, int K, int S, int R, int RY, int D, int CB>
struct BwdCfg {
  static constexpr int PAD = S == 1 ? (K - 1) / 2 : (K - 2) / 2;
  static constexpr int FL = floordiv(PAD - K + 1, S);          // first source offset (in source pixels)
  static constexpr int NH = 64 / CB;
  static constexpr int TW = 8 * R * NH;
  static constexpr int WW = (TW - 1 + PAD) / S - FL + 1;
  static constexpr int WIN = (RY - 1 + PAD) / S - FL + 1;
  static constexpr int NEW = RY / S;
  static constexpr int NRING = WIN;
  static constexpr int VW = (R - 1 + PAD) / S - FL + 1;
  static constexpr int PRO = WIN - NEW;
  static constexpr int PRO_PASSES = (PRO + NEW - 1) / NEW;
  static constexpr int NSTAGE = D + PRO_PASSES;
  using Ld = Loader<T, 2, WW, NEW, NRING, sizeof(T) == 2, CB>;
  static constexpr size_t RING_BYTES = sizeof(float) * (size_t)NRING * WW * CB;
  static constexpr size_t SMEM =
      RING_BYTES + sizeof(float) * (7 * CB + K * K * CB) + (size_t)NSTAGE * Ld::STAGE_BYTES;
};

template <typename T, int K, int S, int R, int RY, int D, int MINB, int CB>
__global__ void __launch_bounds__(NT, MINB) dw2_bwd_kernel(s2lc_lazygrad gin, const float* __restrict__ w,
                                                           s2lc_lazy in, int apply_act_bwd,
                                                           const T* __restrict__ res, int64_t ld_res,
                                                           T* __restrict__ out, int64_t ld_out, double* stats,
                                                           float* __restrict__ dw, Geom g) {
  cudaGridDependencySynchronize();   // programmatic dependent launch (common.cuh)
  using Cfg = BwdCfg<T, K, S, R, RY, D, CB>;
  using Ld = typename Cfg::Ld;
  constexpr bool FAST = sizeof(T) == 2;
  constexpr int WW = Cfg::WW, WIN = Cfg::WIN, NEW = Cfg::NEW, NRING = Cfg::NRING, VW = Cfg::VW, PAD = Cfg::PAD,
                FL = Cfg::FL, PRO = Cfg::PRO, PRO_PASSES = Cfg::PRO_PASSES, NSTAGE = Cfg::NSTAGE;
  static_assert(RY % S == 0 && R % S == 0, "strip geometry must be stride-aligned");
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* tile = reinterpret_cast<float*>(smem_raw);
  float* aux = tile + (size_t)NRING * WW * CB;  // [3][CB] p, q, t
  float* iaux = aux + 3 * CB;                    // [2][CB] input scale, shift
  float* sstat = iaux + 2 * CB;                  // [2][CB]
  float* sdw = sstat + 2 * CB;                   // [K*K][CB]
  unsigned char* stages = reinterpret_cast<unsigned char*>(sdw + K * K * CB);
  constexpr int HL = CB / 2, NH = Cfg::NH;
  const int lane = (threadIdx.x & 31) % HL, pg = (threadIdx.x >> 5) * NH + (threadIdx.x & 31) / HL;
  const int cb0 = blockIdx.y * CB;
  const int c = cb0 + 2 * lane;
  const bool cvalid = c < g.C;
  const int gmode = gin.z ? 1 : 2;
  if (threadIdx.x < CB) {
    int cc = cb0 + threadIdx.x;
    bool v = cc < g.C;
    aux[threadIdx.x] = (v && gin.z) ? gin.p[cc] : 1.f;
    aux[CB + threadIdx.x] = (v && gin.z) ? gin.q[cc] : 0.f;
    aux[2 * CB + threadIdx.x] = (v && gin.z) ? gin.t[cc] : 0.f;
    iaux[threadIdx.x] = (v && in.scale) ? in.scale[cc] : 1.f;
    iaux[CB + threadIdx.x] = (v && in.shift) ? in.shift[cc] : 0.f;
  }
  if (threadIdx.x < 2 * CB) sstat[threadIdx.x] = 0.f;
  for (int i = threadIdx.x; i < K * K * CB; i += NT) sdw[i] = 0.f;
  float2 wr[K * K], wacc[K * K];
#pragma unroll
  for (int t = 0; t < K * K; ++t) {
    wr[t] = cvalid ? make_float2(w[(int64_t)c * K * K + t], w[(int64_t)(c + 1) * K * K + t]) : make_float2(0.f, 0.f);
    wacc[t] = make_float2(0.f, 0.f);
  }
  __syncthreads();
  const float2 isc = *reinterpret_cast<const float2*>(iaux + 2 * lane);
  const float2 ish = *reinterpret_cast<const float2*>(iaux + CB + 2 * lane);
  float2 s0 = make_float2(0.f, 0.f), s1 = make_float2(0.f, 0.f);
  const T* sdy = reinterpret_cast<const T*>(gin.dy);
  const T* sz = reinterpret_cast<const T*>(gin.z ? gin.z : gin.dy);
  const T* zin = reinterpret_cast<const T*>(in.raw);
  Ld ld;
  const int items = g.N * g.ychunks * g.xs;
  int iss = 0, com = 0;  // stage cursors (issue / commit), both advance once per pass
  auto next = [](int& x) { int r = x; x = (x + 1 == NSTAGE) ? 0 : x + 1; return r; };

  for (int item = blockIdx.x; item < items; item += gridDim.x) {
    int n = item / (g.ychunks * g.xs);
    int rem = item - n * g.ychunks * g.xs;
    int y0 = (rem / g.xs) * g.YC, x0 = (rem % g.xs) * Cfg::TW;
    int yend = min(y0 + g.YC, g.Hd);
    int nsteps = (yend - y0 + RY - 1) / RY;
    const int gy0 = y0 / S + FL, gx0 = x0 / S + FL;
    const float rs = gin.rowscale ? gin.rowscale[n] : 1.f;
    ld.begin_item(sdy, gin.ld_dy, sz, gin.ld_z, n, g.Hs, g.Ws, gy0, gx0, cb0, g.C);
    auto issue = [&](int rel0, int nr) { ld.issue(stages + (size_t)next(iss) * Ld::STAGE_BYTES, rel0, nr); };
    auto commit = [&](int rel0, int rb, int nr) {
      ld.commit(stages + (size_t)next(com) * Ld::STAGE_BYTES, tile, aux, rel0, rb, nr, gmode, 0, rs);
    };
    // destination-side operand (saved raw input of the depthwise conv at this warp's RY x R pixels), prefetched
    // one step ahead in registers
    Pair<T> zr[RY][R], zn[RY][R];
    auto load_z = [&](Pair<T> (&z)[RY][R], int t) {
#pragma unroll
      for (int dy = 0; dy < RY; ++dy) {
        int iy = y0 + t * RY + dy;
#pragma unroll
        for (int r = 0; r < R; ++r) {
          int ix = x0 + pg * R + r;
          z[dy][r].zero();
          if (cvalid && iy < yend && ix < g.Wd) z[dy][r].load(zin + (((int64_t)n * g.Hd + iy) * g.Wd + ix) * in.ld + c);
        }
      }
    };
#pragma unroll
    for (int pp = 0; pp < PRO_PASSES; ++pp) issue(pp * NEW, min(NEW, PRO - pp * NEW));
#pragma unroll
    for (int d = 0; d < D; ++d) {
      if (d < nsteps) issue(PRO + d * NEW, NEW);
      else cp_async_commit();
    }
    load_z(zn, 0);
    cp_async_wait<D - 1>();
    __syncthreads();
#pragma unroll
    for (int pp = 0; pp < PRO_PASSES; ++pp) commit(pp * NEW, (pp * NEW) % NRING, min(NEW, PRO - pp * NEW));
    commit(PRO, PRO % NRING, NEW);
    int cslot = (PRO + NEW) % NRING;
    __syncthreads();
    int base = 0;
    for (int t = 0; t < nsteps; ++t) {
      if (t + D < nsteps) issue(PRO + (t + D) * NEW, NEW);
      else cp_async_commit();
#pragma unroll
      for (int dy = 0; dy < RY; ++dy)
#pragma unroll
        for (int r = 0; r < R; ++r) zr[dy][r] = zn[dy][r];
      if (t + 1 < nsteps) load_z(zn, t + 1);
      float2 a[RY][R], acc[RY][R];
#pragma unroll
      for (int dy = 0; dy < RY; ++dy) {
        int iy = y0 + t * RY + dy;
#pragma unroll
        for (int r = 0; r < R; ++r) {
          int ix = x0 + pg * R + r;
          float2 z = zr[dy][r].get();
          bool v = cvalid && iy < yend && ix < g.Wd;
          float2 u = ffma2r(z, isc, ish);
          a[dy][r].x = v ? act_f<FAST>(u.x, in.act) : 0.f;
          a[dy][r].y = v ? act_f<FAST>(u.y, in.act) : 0.f;
          acc[dy][r] = make_float2(0.f, 0.f);
        }
      }
      const float* tl = tile + ((pg * R) / S) * CB + 2 * lane;
#pragma unroll
      for (int j = 0; j < WIN; ++j) {
        int slot = base + j;
        if (slot >= NRING) slot -= NRING;
        const float* rp = tl + (size_t)slot * WW * CB;
        float2 v[VW];
#pragma unroll
        for (int i = 0; i < VW; ++i) v[i] = *reinterpret_cast<const float2*>(rp + i * CB);
#pragma unroll
        for (int dy = 0; dy < RY; ++dy) {
          const int ky = dy + PAD - S * (FL + j);
          if (ky < 0 || ky >= K) continue;
#pragma unroll
          for (int r = 0; r < R; ++r)
#pragma unroll
            for (int i = 0; i < VW; ++i) {
              const int kx = r + PAD - S * (FL + i);
              if (kx < 0 || kx >= K) continue;
              ffma2(acc[dy][r], wr[ky * K + kx], v[i]);
              ffma2(wacc[ky * K + kx], a[dy][r], v[i]);
            }
        }
      }
      if (cvalid) {
        const int ix0 = x0 + pg * R;
        const float2 one2 = make_float2(1.f, 1.f);
        const bool silu_fast = FAST && in.act == S2LC_ACT_SILU;
#pragma unroll
        for (int dy = 0; dy < RY; ++dy) {
          int iy = y0 + t * RY + dy;
          if (iy >= yend) break;
          const int64_t orow0 = ((int64_t)n * g.Hd + iy) * g.Wd + ix0;
          T* op = out + orow0 * ld_out + c;
          const T* rp = res ? res + orow0 * ld_res + c : nullptr;
#pragma unroll
          for (int r = 0; r < R; ++r) {
            if (ix0 + r >= g.Wd) break;
            float2 z = zr[dy][r].get();
            float2 dx = acc[dy][r];
            if (apply_act_bwd) {
              float2 u = ffma2r(z, isc, ish);
              float2 d;
              if (silu_fast) {  // s = (1 + tanh(u/2))/2;  act' = s * (1 + u*(1-s))
                const float2 half2 = make_float2(0.5f, 0.5f);
                float2 hu = fmul2(u, half2);
                float2 sg = ffma2r(make_float2(tanh_fast(hu.x), tanh_fast(hu.y)), half2, half2);
                float2 oms = ffma2r(sg, make_float2(-1.f, -1.f), one2);
                d = fmul2(sg, ffma2r(u, oms, one2));
              } else {
                d = make_float2(act_d<false>(u.x, in.act), act_d<false>(u.y, in.act));
              }
              dx = fmul2(dx, d);
            }
            if (rp) {
              Pair<T> e;
              e.load(rp + (int64_t)r * ld_res);
              float2 ef = e.get();
              dx.x += ef.x; dx.y += ef.y;
            }
            float2 st = st_pair_rt<T>(op + (int64_t)r * ld_out, dx);
            ffma2(s0, st, one2);
            ffma2(s1, st, z);
          }
        }
      }
      base += NEW;
      if (base >= NRING) base -= NRING;
      if (t + 1 < nsteps) {
        cp_async_wait<D - 1>();
        __syncthreads();
        commit(PRO + (t + 1) * NEW, cslot, NEW);
        cslot += NEW;
        if (cslot >= NRING) cslot -= NRING;
        __syncthreads();
      }
    }
  }
  __syncthreads();
  if (dw) {
#pragma unroll
    for (int t = 0; t < K * K; ++t) {
      atomicAdd(&sdw[t * CB + 2 * lane], wacc[t].x);
      atomicAdd(&sdw[t * CB + 2 * lane + 1], wacc[t].y);
    }
  }
  if (stats) flush_stats<CB>(s0, s1, sstat, cb0, g.C, stats);  // contains a __syncthreads
  else __syncthreads();
  if (dw) {
    for (int i = threadIdx.x; i < K * K * CB; i += NT) {
      int t = i / CB, cc = cb0 + i % CB;
      if (cc < g.C) atomicAdd(&dw[(int64_t)cc * K * K + t], sdw[i]);
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// Fast forward (bf16 storage, exact strip geometry): dw2_fwd_kernel with the launch-uniform choices made at compile
// time — input = BatchNorm + SiLU (SILU: packed transform with pre-halved scale / shift) or a plain stored tensor, no
// per-pixel bounds tests in the epilogue, output pointers advanced by strides.
template <int K, int S, int R, int RY, int D, int MINB, int CB, bool SILU>
__global__ void __launch_bounds__(NT, MINB) dw2_fwdx_kernel(s2lc_lazy in, const float* __restrict__ w,
                                                            bf16* __restrict__ out, double* stats, Geom g) {
  cudaGridDependencySynchronize();   // programmatic dependent launch (common.cuh)
  using T = bf16;
  using Cfg = FwdCfg<T, K, S, R, RY, D, CB>;
  using Ld = typename Cfg::Ld;
  constexpr int WW = Cfg::WW, WIN = Cfg::WIN, NEW = Cfg::NEW, NRING = Cfg::NRING, VW = Cfg::VW, PRO = Cfg::PRO,
                PRO_PASSES = Cfg::PRO_PASSES, NSTAGE = Cfg::NSTAGE;
  constexpr int MODE = SILU ? 0 : 2;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* tile = reinterpret_cast<float*>(smem_raw);
  float* aux = tile + (size_t)NRING * WW * CB;  // [2][CB] scale / 2, shift / 2 (SILU) or 1, 0
  float* sstat = aux + 2 * CB;                   // [2][CB]
  unsigned char* stages = reinterpret_cast<unsigned char*>(sstat + 2 * CB);
  constexpr int HL = CB / 2, NH = Cfg::NH;
  const int lane = (threadIdx.x & 31) % HL, pg = (threadIdx.x >> 5) * NH + (threadIdx.x & 31) / HL;
  const int cb0 = blockIdx.y * CB;
  const int c = cb0 + 2 * lane;
  const bool cvalid = c < g.C;
  if (threadIdx.x < CB) {
    const int cc = cb0 + threadIdx.x;
    aux[threadIdx.x] = SILU ? 0.5f * (cc < g.C ? in.scale[cc] : 1.f) : 1.f;
    aux[CB + threadIdx.x] = SILU ? 0.5f * (cc < g.C ? in.shift[cc] : 0.f) : 0.f;
  }
  if (threadIdx.x < 2 * CB) sstat[threadIdx.x] = 0.f;
  float2 wr[K * K];
#pragma unroll
  for (int t = 0; t < K * K; ++t)
    wr[t] = cvalid ? make_float2(w[(int64_t)c * K * K + t], w[(int64_t)(c + 1) * K * K + t]) : make_float2(0.f, 0.f);
  float2 s0 = make_float2(0.f, 0.f), s1 = make_float2(0.f, 0.f);
  const T* src = reinterpret_cast<const T*>(in.raw);
  Ld ld;
  const int items = g.N * g.ychunks * g.xs;
  const int nsteps = g.YC / RY;                       // exact geometry
  const int64_t orow = (int64_t)g.Wd * g.C;
  int iss = 0, com = 0;
  auto next = [](int& x) { int r = x; x = (x + 1 == NSTAGE) ? 0 : x + 1; return r; };
  const float2 one2 = make_float2(1.f, 1.f);

  for (int item = blockIdx.x; item < items; item += gridDim.x) {
    const int n = item / (g.ychunks * g.xs);
    const int rem = item - n * g.ychunks * g.xs;
    const int y0 = (rem / g.xs) * g.YC, x0 = (rem % g.xs) * Cfg::TW;
    const int gy0 = y0 * S - g.pad_t, gx0 = x0 * S - g.pad_l;
    ld.begin_item(src, in.ld, src, in.ld, n, g.Hs, g.Ws, gy0, gx0, cb0, g.C);
    auto issue = [&](int rel0, int nr) { ld.issue(stages + (size_t)next(iss) * Ld::STAGE_BYTES, rel0, nr); };
    T* op = out + (((int64_t)n * g.Hd + y0) * g.Wd + x0 + pg * R) * g.C + c;   // advanced by RY rows per step
#pragma unroll
    for (int pp = 0; pp < PRO_PASSES; ++pp) issue(pp * NEW, min(NEW, PRO - pp * NEW));
#pragma unroll
    for (int d = 0; d < D; ++d) {
      if (d < nsteps) issue(PRO + d * NEW, NEW);
      else cp_async_commit();
    }
    cp_async_wait<D - 1>();
    __syncthreads();  // previous item's readers are done with the ring; aux/sstat initialised
#pragma unroll
    for (int pp = 0; pp < PRO_PASSES; ++pp)
      ld.template commit_ct<MODE, true>(stages + (size_t)next(com) * Ld::STAGE_BYTES, tile, aux, pp * NEW,
                                        (pp * NEW) % NRING, min(NEW, PRO - pp * NEW));
    ld.template commit_ct<MODE, false>(stages + (size_t)next(com) * Ld::STAGE_BYTES, tile, aux, PRO, PRO % NRING, NEW);
    int cslot = (PRO + NEW) % NRING;
    __syncthreads();
    int base = 0;
    for (int t = 0; t < nsteps; ++t) {
      if (t + D < nsteps) issue(PRO + (t + D) * NEW, NEW);
      else cp_async_commit();
      float2 acc[RY][R];
#pragma unroll
      for (int dy = 0; dy < RY; ++dy)
#pragma unroll
        for (int r = 0; r < R; ++r) acc[dy][r] = make_float2(0.f, 0.f);
      const float* tl = tile + (pg * R * S) * CB + 2 * lane;
#pragma unroll
      for (int j = 0; j < WIN; ++j) {
        int slot = base + j;
        if (slot >= NRING) slot -= NRING;
        const float* rp = tl + (size_t)slot * WW * CB;
        float2 v[VW];
#pragma unroll
        for (int i = 0; i < VW; ++i) v[i] = *reinterpret_cast<const float2*>(rp + i * CB);
#pragma unroll
        for (int dy = 0; dy < RY; ++dy) {
          const int ky = j - dy * S;
          if (ky < 0 || ky >= K) continue;
#pragma unroll
          for (int r = 0; r < R; ++r)
#pragma unroll
            for (int kx = 0; kx < K; ++kx) ffma2(acc[dy][r], wr[ky * K + kx], v[r * S + kx]);
        }
      }
      if (cvalid) {
#pragma unroll
        for (int dy = 0; dy < RY; ++dy) {
          T* opr = op + dy * orow;
#pragma unroll
          for (int r = 0; r < R; ++r) {
            const float2 ab = st_pair_rt<T>(opr + r * g.C, acc[dy][r]);   // stored (rounded) values
            ffma2(s0, ab, one2);
            ffma2(s1, ab, ab);
          }
        }
      }
      op += RY * orow;
      base += NEW;
      if (base >= NRING) base -= NRING;
      if (t + 1 < nsteps) {
        cp_async_wait<D - 1>();
        __syncthreads();
        ld.template commit_ct<MODE, false>(stages + (size_t)next(com) * Ld::STAGE_BYTES, tile, aux,
                                           PRO + (t + 1) * NEW, cslot, NEW);
        cslot += NEW;
        if (cslot >= NRING) cslot -= NRING;
        __syncthreads();
      }
    }
  }
  if (stats) {
    __syncthreads();
    flush_stats<CB>(s0, s1, sstat, cb0, g.C, stats);
  }
}

// ---------------------------------------------------------------------------------------------------------
// Fast fused backward (bf16 storage, the shapes of the b5 / b7 encoders): same algorithm and ring as dw2_bwd_kernel
// with everything that is uniform over a launch resolved at compile time — the strip geometry divides the image (no
// per-pixel bounds tests), the input transform is either BatchNorm + SiLU with act' applied (SILU) or a plain stored
// tensor (no act', optional residual), p * rowscale is folded once per strip, and the raw input tile of a step
// arrives with cp.async next to the gradient rows (16 B per thread and row instead of RY x R 4-byte LDGs with
// rebuilt addresses).  The step loop drops from ~1480 to ~870 SASS instructions for 8 pixel pairs
// (tools/sass_lines.py) — these kernels are issue-bound.
//
// Shared-memory plan: per stage the loader's gradient slots followed by the RY x TW x CB raw input tile of the same
// step (bf16, pixel-major).  CB = 32: a pixel record is 64 B and the two half-warps of a warp read pixels 4 apart,
// i.e. the same 16 banks — odd groups of 4 pixels therefore swap the two records of every 128 B line (record of
// pixel p lives at slot p ^ ((p >> 2) & 1)).
// ONEBAR (kernels that run one CTA per SM): the ring holds NEW rows more than the window, so the pass of step t+1 is
// transformed into rows that step t does not read — one barrier per step instead of two, and the transform of one
// warp overlaps the stencil of another.
// TWOPH: two sweeps over the window (data gradient, epilogue, weight gradient) — one tanh per value and lower
// register pressure; wins where two CTAs share an SM (3x3), loses on the 5x5 kernels (profiles/r02p_dw_variants.txt).
template <int K, int S, int R, int RY, int D, int CB, bool ONEBAR>
struct BwdxCfg {
  using B = BwdCfg<bf16, K, S, R, RY, D, CB>;
  static constexpr int WW = B::WW, WIN = B::WIN, NEW = B::NEW, VW = B::VW, PAD = B::PAD, FL = B::FL, PRO = B::PRO,
                       PRO_PASSES = B::PRO_PASSES, TW = B::TW, NH = B::NH;
  static constexpr int NRING = WIN + (ONEBAR ? NEW : 0);
  static constexpr int NSTAGE = D + PRO_PASSES;
  using Ld = Loader<bf16, 2, WW, NEW, NRING, true, CB>;
  static constexpr int PXB = CB * 2;                                     // bytes per pixel record
  static constexpr int ZROWB = TW * PXB;
  static constexpr int ZBYTES = RY * ZROWB;
  static constexpr int STRIDE = Ld::STAGE_BYTES + ZBYTES;
  static constexpr size_t RING_BYTES = sizeof(float) * (size_t)NRING * WW * CB;
  static constexpr size_t SMEM = RING_BYTES + sizeof(float) * (7 * CB + K * K * CB) + (size_t)NSTAGE * STRIDE;
  static_assert(TW == NT / (CB / 8), "one input-tile vector per thread and row");
  static_assert(Ld::STAGE_BYTES % 16 == 0, "input tile must stay 16 B aligned");
  static_assert(R == 4, "the record swap assumes groups of 4 pixels per lane group");
  static_assert(SMEM <= 232448, "shared memory of one CTA");
};

template <int K, int S, int R, int RY, int D, int MINB, int CB, bool SILU, bool ONEBAR, bool TWOPH>
__global__ void __launch_bounds__(NT, MINB) dw2_bwdx_kernel(s2lc_lazygrad gin, const float* __restrict__ w,
                                                            s2lc_lazy in, const bf16* __restrict__ res,
                                                            int64_t ld_res, bf16* __restrict__ out, int64_t ld_out,
                                                            double* stats, float* __restrict__ dw, Geom g) {
  cudaGridDependencySynchronize();   // programmatic dependent launch (common.cuh)
  using T = bf16;
  using X = BwdxCfg<K, S, R, RY, D, CB, ONEBAR>;
  using Ld = typename X::Ld;
  constexpr int WW = X::WW, WIN = X::WIN, NEW = X::NEW, NRING = X::NRING, VW = X::VW, PAD = X::PAD, FL = X::FL,
                PRO = X::PRO, PRO_PASSES = X::PRO_PASSES, NSTAGE = X::NSTAGE;
  constexpr int CBV = CB / 8;
  static_assert(RY % S == 0 && R % S == 0, "strip geometry must be stride-aligned");
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* tile = reinterpret_cast<float*>(smem_raw);
  float* aux = tile + (size_t)NRING * WW * CB;  // [3][CB] p * rowscale (per strip), q, t
  float* iaux = aux + 3 * CB;                    // [CB] p
  float* sstat = iaux + 2 * CB;                  // [2][CB]
  float* sdw = sstat + 2 * CB;                   // [K*K][CB]
  unsigned char* stages = reinterpret_cast<unsigned char*>(sdw + K * K * CB);
  constexpr int HL = CB / 2, NH = X::NH;
  const int lane = (threadIdx.x & 31) % HL, pg = (threadIdx.x >> 5) * NH + (threadIdx.x & 31) / HL;
  const int cb0 = blockIdx.y * CB;
  const int c = cb0 + 2 * lane;
  const bool cvalid = c < g.C;
  const bool lazyg = gin.z != nullptr;
  const bool has_res = !SILU && res != nullptr && cvalid;
  // input-tile roles: writer = (pixel zpx, 8-channel vector zv8) once per row; reader = this lane's channel pair
  const int zpx = threadIdx.x / CBV, zv8 = threadIdx.x % CBV;
  const bool zok = cb0 + zv8 * 8 < g.C;
  const int zsw = CB == 32 ? (pg & 1) : 0;            // this lane group's records are swapped pairwise
  const int zwr = Ld::STAGE_BYTES + (CB == 32 ? (zpx ^ ((zpx >> 2) & 1)) : zpx) * X::PXB + zv8 * 16;
  const int zrd = Ld::STAGE_BYTES + pg * R * X::PXB + lane * 4;
  const int zrd_e = zrd + zsw * X::PXB, zrd_o = zrd - zsw * X::PXB;   // bases of the even / odd pixels of the group
  if (threadIdx.x < CB) {
    int cc = cb0 + threadIdx.x;
    bool v = cc < g.C;
    iaux[threadIdx.x] = (v && lazyg) ? gin.p[cc] : 1.f;
    aux[CB + threadIdx.x] = (v && lazyg) ? gin.q[cc] : 0.f;
    aux[2 * CB + threadIdx.x] = (v && lazyg) ? gin.t[cc] : 0.f;
  }
  if (threadIdx.x < 2 * CB) sstat[threadIdx.x] = 0.f;
  for (int i = threadIdx.x; i < K * K * CB; i += NT) sdw[i] = 0.f;
  if (!zok) {  // channel vectors beyond C are never fetched: keep them zero
#pragma unroll
    for (int sg = 0; sg < NSTAGE; ++sg)
#pragma unroll
      for (int dy = 0; dy < RY; ++dy)
        *reinterpret_cast<uint4*>(stages + (size_t)sg * X::STRIDE + zwr + dy * X::ZROWB) = make_uint4(0, 0, 0, 0);
  }
  float2 wr[K * K], wacc[K * K];
#pragma unroll
  for (int t = 0; t < K * K; ++t) {
    wr[t] = cvalid ? make_float2(w[(int64_t)c * K * K + t], w[(int64_t)(c + 1) * K * K + t]) : make_float2(0.f, 0.f);
    wacc[t] = make_float2(0.f, 0.f);
  }
  // h = u / 2 = z * (scale / 2) + shift / 2 feeds tanh directly
  float2 isc = make_float2(0.f, 0.f), ish = make_float2(0.f, 0.f);
  if (SILU && cvalid) {
    isc = make_float2(0.5f * in.scale[c], 0.5f * in.scale[c + 1]);
    ish = make_float2(0.5f * in.shift[c], 0.5f * in.shift[c + 1]);
  }
  float2 s0 = make_float2(0.f, 0.f), s1 = make_float2(0.f, 0.f);
  const T* sdy = reinterpret_cast<const T*>(gin.dy);
  const T* sz = reinterpret_cast<const T*>(lazyg ? gin.z : gin.dy);
  const T* zin = reinterpret_cast<const T*>(in.raw);
  Ld ld;
  const int items = g.N * g.ychunks * g.xs;
  const int nsteps = g.YC / RY;                      // exact geometry: every strip has YC rows
  const int64_t zrow = (int64_t)g.Wd * in.ld, orow = (int64_t)g.Wd * ld_out, rrow = (int64_t)g.Wd * ld_res;
  const int ldo = (int)ld_out, ldr = (int)ld_res;
  int iss = 0, com = 0;
  auto next = [](int& x) { int r = x; x = (x + 1 == NSTAGE) ? 0 : x + 1; return r; };
  const float2 one2 = make_float2(1.f, 1.f), half2 = make_float2(0.5f, 0.5f), mone2 = make_float2(-1.f, -1.f);
  auto zpair = [](const unsigned char* p) {
    const uint32_t u = *reinterpret_cast<const uint32_t*>(p);
    return make_float2(__uint_as_float(u << 16), __uint_as_float(u & 0xffff0000u));
  };

  for (int item = blockIdx.x; item < items; item += gridDim.x) {
    const int n = item / (g.ychunks * g.xs);
    const int rem = item - n * g.ychunks * g.xs;
    const int y0 = (rem / g.xs) * g.YC, x0 = (rem % g.xs) * X::TW;
    const int gy0 = y0 / S + FL, gx0 = x0 / S + FL;
    ld.begin_item(sdy, gin.ld_dy, sz, gin.ld_z, n, g.Hs, g.Ws, gy0, gx0, cb0, g.C);
    const int64_t pixi = ((int64_t)n * g.Hd + y0) * g.Wd + x0;
    const T* zsrc = zin + (pixi + zpx) * in.ld + cb0 + zv8 * 8;   // this thread's input-tile vector, row 0 of the strip
    T* op = out + (pixi + pg * R) * ld_out + c;                    // advanced by RY rows per step
    const T* rp = res ? res + (pixi + pg * R) * ld_res + c : nullptr;
    // the input tiles of a stage (and, with ONEBAR, the ring rows) are shared between threads: the previous strip's
    // last step must be over before this strip's prologue overwrites them
    __syncthreads();
    if (threadIdx.x < CB) aux[threadIdx.x] = iaux[threadIdx.x] * (gin.rowscale ? gin.rowscale[n] : 1.f);
    // pass of step ts = its gradient rows + its input tile, one cp.async group
    auto issue_step = [&](int ts) {
      unsigned char* stg = stages + (size_t)next(iss) * X::STRIDE;
      ld.issue(stg, PRO + ts * NEW, NEW, false);
      if (zok) {
        const T* zs = zsrc + (int64_t)ts * RY * zrow;
#pragma unroll
        for (int dy = 0; dy < RY; ++dy) cp_async16(stg + zwr + dy * X::ZROWB, zs + dy * zrow);
      }
      cp_async_commit();
    };
    auto commit_step = [&](const unsigned char* stg, int ts, int slot) {
      if (lazyg) ld.template commit_ct<1, false>(stg, tile, aux, PRO + ts * NEW, slot, NEW);
      else ld.template commit_ct<2, false>(stg, tile, aux, PRO + ts * NEW, slot, NEW);
    };
#pragma unroll
    for (int pp = 0; pp < PRO_PASSES; ++pp)
      ld.issue(stages + (size_t)next(iss) * X::STRIDE, pp * NEW, min(NEW, PRO - pp * NEW));
#pragma unroll
    for (int d = 0; d < D; ++d) {
      if (d < nsteps) issue_step(d);
      else cp_async_commit();
    }
    cp_async_wait<D - 1>();
    __syncthreads();   // aux of this strip is visible; step 0's input tile has landed for every thread
#pragma unroll
    for (int pp = 0; pp < PRO_PASSES; ++pp) {
      const unsigned char* stg = stages + (size_t)next(com) * X::STRIDE;
      if (lazyg) ld.template commit_ct<1, true>(stg, tile, aux, pp * NEW, (pp * NEW) % NRING, min(NEW, PRO - pp * NEW));
      else ld.template commit_ct<2, true>(stg, tile, aux, pp * NEW, (pp * NEW) % NRING, min(NEW, PRO - pp * NEW));
    }
    const unsigned char* zcur = stages + (size_t)next(com) * X::STRIDE;   // stage of the current step
    commit_step(zcur, 0, PRO % NRING);
    int cslot = (PRO + NEW) % NRING;
    __syncthreads();
    int base = 0;
    for (int t = 0; t < nsteps; ++t) {
      if (t + D < nsteps) issue_step(t + D);
      else cp_async_commit();
      const unsigned char* znext = zcur;
      if (ONEBAR && t + 1 < nsteps) {
        // rows of step t+1 go to ring slots outside this step's window: no barrier before the transform
        cp_async_wait<D - 1>();
        znext = stages + (size_t)next(com) * X::STRIDE;
        commit_step(znext, t + 1, cslot);
        cslot += NEW;
        if (cslot >= NRING) cslot -= NRING;
      }
      const unsigned char* zs_e = zcur + zrd_e;
      const unsigned char* zs_o = zcur + zrd_o;
      const float* tl = tile + ((pg * R) / S) * CB + 2 * lane;
      float2 a[RY][R], acc[RY][R];
#pragma unroll
      for (int dy = 0; dy < RY; ++dy)
#pragma unroll
        for (int r = 0; r < R; ++r) {
          acc[dy][r] = make_float2(0.f, 0.f);
          if (!TWOPH) {  // single sweep: the activation is needed before the stencil (tanh evaluated twice per value)
            const float2 z = zpair(((r & 1) ? zs_o : zs_e) + dy * X::ZROWB + r * X::PXB);
            if (SILU) {
              const float2 h = ffma2r(z, isc, ish);
              a[dy][r] = ffma2r(h, make_float2(tanh_fast(h.x), tanh_fast(h.y)), h);
            } else {
              a[dy][r] = z;
            }
          }
        }
      // sweep 1: data gradient (and, single-sweep variant, the weight gradient)
#pragma unroll
      for (int j = 0; j < WIN; ++j) {
        int slot = base + j;
        if (slot >= NRING) slot -= NRING;
        const float* rpz = tl + (size_t)slot * WW * CB;
        float2 v[VW];
#pragma unroll
        for (int i = 0; i < VW; ++i) v[i] = *reinterpret_cast<const float2*>(rpz + i * CB);
#pragma unroll
        for (int dy = 0; dy < RY; ++dy) {
          const int ky = dy + PAD - S * (FL + j);
          if (ky < 0 || ky >= K) continue;
#pragma unroll
          for (int r = 0; r < R; ++r)
#pragma unroll
            for (int i = 0; i < VW; ++i) {
              const int kx = r + PAD - S * (FL + i);
              if (kx < 0 || kx >= K) continue;
              ffma2(acc[dy][r], wr[ky * K + kx], v[i]);
              if (!TWOPH) ffma2(wacc[ky * K + kx], a[dy][r], v[i]);
            }
        }
      }
      // epilogue; two-sweep variant: it also yields the activation for sweep 2 (one tanh per value)
#pragma unroll
      for (int dy = 0; dy < RY; ++dy) {
        T* opr = op + dy * orow;
        const T* rpr = rp + dy * rrow;
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const float2 z = zpair(((r & 1) ? zs_o : zs_e) + dy * X::ZROWB + r * X::PXB);
          float2 dx = acc[dy][r];
          if (SILU) {  // s = (1 + tanh(h))/2;  act = h*tanh(h) + h;  act' = s * (1 + h*(1 - tanh(h)))
            const float2 h = ffma2r(z, isc, ish);
            const float2 th = make_float2(tanh_fast(h.x), tanh_fast(h.y));
            if (TWOPH) a[dy][r] = ffma2r(h, th, h);
            const float2 sg = ffma2r(th, half2, half2);
            const float2 e = ffma2r(h, ffma2r(th, mone2, one2), one2);
            dx = fmul2(dx, fmul2(sg, e));
          } else {
            if (TWOPH) a[dy][r] = z;
            if (has_res) {   // (lanes beyond C must not touch memory: the last pixel's record ends the buffer)
              Pair<T> ev;
              ev.load(rpr + r * ldr);
              const float2 ef = ev.get();
              dx.x += ef.x; dx.y += ef.y;
            }
          }
          if (cvalid) {
            const float2 st = st_pair_rt<T>(opr + r * ldo, dx);
            ffma2(s0, st, one2);
            ffma2(s1, st, z);
          }
        }
      }
      if (TWOPH) {  // sweep 2: weight gradient
#pragma unroll
        for (int j = 0; j < WIN; ++j) {
          int slot = base + j;
          if (slot >= NRING) slot -= NRING;
          const float* rpz = tl + (size_t)slot * WW * CB;
          float2 v[VW];
#pragma unroll
          for (int i = 0; i < VW; ++i) v[i] = *reinterpret_cast<const float2*>(rpz + i * CB);
#pragma unroll
          for (int dy = 0; dy < RY; ++dy) {
            const int ky = dy + PAD - S * (FL + j);
            if (ky < 0 || ky >= K) continue;
#pragma unroll
            for (int r = 0; r < R; ++r)
#pragma unroll
              for (int i = 0; i < VW; ++i) {
                const int kx = r + PAD - S * (FL + i);
                if (kx < 0 || kx >= K) continue;
                ffma2(wacc[ky * K + kx], a[dy][r], v[i]);
              }
          }
        }
      }
      op += RY * orow;
      if (has_res) rp += RY * rrow;
      base += NEW;
      if (base >= NRING) base -= NRING;
      if (t + 1 < nsteps) {
        if (ONEBAR) {
          zcur = znext;
          __syncthreads();
        } else {
          cp_async_wait<D - 1>();
          __syncthreads();
          zcur = stages + (size_t)next(com) * X::STRIDE;
          commit_step(zcur, t + 1, cslot);
          cslot += NEW;
          if (cslot >= NRING) cslot -= NRING;
          __syncthreads();
        }
      }
    }
  }
  __syncthreads();
  if (dw) {
#pragma unroll
    for (int t = 0; t < K * K; ++t) {
      atomicAdd(&sdw[t * CB + 2 * lane], wacc[t].x);
      atomicAdd(&sdw[t * CB + 2 * lane + 1], wacc[t].y);
    }
  }
  if (stats) flush_stats<CB>(s0, s1, sstat, cb0, g.C, stats);  // contains a __syncthreads
  else __syncthreads();
  if (dw) {
    for (int i = threadIdx.x; i < K * K * CB; i += NT) {
      int t = i / CB, cc = cb0 + i % CB;
      if (cc < g.C) atomicAdd(&dw[(int64_t)cc * K * K + t], sdw[i]);
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// host side
template <typename Kern>
static int32_t set_smem(Kern kernel, size_t bytes) {
  if (bytes > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e != cudaSuccess) {
      set_error("cudaFuncSetAttribute(smem=%zu): %s", bytes, cudaGetErrorString(e));
      return S2LC_ERR_CUDA;
    }
  }
  return S2LC_OK;
}

// strip height YC (multiple of RY): as tall as possible while the launch still has >= ~8 items per resident CTA
static void plan_strips(Geom& g, int TW, int RY, int cblocks, int resident) {
  g.xs = (g.Wd + TW - 1) / TW;
  int yc = (g.Hd + RY - 1) / RY * RY;
  const int min_yc = 8 * RY;
  const long want = 8L * resident;
  while (yc / 2 >= min_yc && (yc / 2) % RY == 0 &&
         (long)g.N * g.xs * ((g.Hd + yc - 1) / yc) * cblocks < want)
    yc /= 2;
  g.YC = yc;
  g.ychunks = (g.Hd + yc - 1) / yc;
}

template <typename Kern>
static int resident_ctas(Kern kern, size_t smem) {
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, NT, smem) != cudaSuccess || per_sm < 1) per_sm = 1;
  return per_sm * kNumSMs;
}

template <typename T, int K, int S, int R, int RY, int D, int MINB, int CB>
static int32_t launch_fwd(const s2lc_lazy* in, const float* w, void* out, double* stats, int N, int H, int W, int C,
                          cudaStream_t st) {
  using Cfg = FwdCfg<T, K, S, R, RY, D, CB>;
  auto kern = dw2_fwd_kernel<T, K, S, R, RY, D, MINB, CB>;
  int32_t r = set_smem(kern, Cfg::SMEM);
  if (r) return r;
  Geom g{};
  g.N = N; g.C = C; g.Hs = H; g.Ws = W;
  g.Hd = (H + S - 1) / S; g.Wd = (W + S - 1) / S;
  int ph = (g.Hd - 1) * S + K - H; if (ph < 0) ph = 0;
  int pw = (g.Wd - 1) * S + K - W; if (pw < 0) pw = 0;
  g.pad_t = ph / 2; g.pad_l = pw / 2;
  const int cblocks = (C + CB - 1) / CB;
  static int resident = resident_ctas(kern, Cfg::SMEM);
  plan_strips(g, Cfg::TW, RY, cblocks, resident);
  long items = (long)g.N * g.xs * g.ychunks;
  int gx = resident / cblocks;   // never more CTAs than are co-resident: a partial second wave doubles the time
  if (gx < 1) gx = 1;
  if (gx > items) gx = (int)items;
  PDL_LAUNCH(kern, dim3(gx, cblocks), NT, Cfg::SMEM, st, *in, w, (T*)out, stats, g);
  return check_launch("dwconv2_fwd");
}

template <typename T, int K, int S, int R, int RY, int D, int MINB, int CB>
static int32_t launch_bwd(const s2lc_lazygrad* gin, const float* w, const s2lc_lazy* in, int apply_act_bwd,
                          const void* res, int64_t ld_res, void* out, int64_t ld_out, double* stats, float* dw, int N,
                          int H, int W, int C, cudaStream_t st) {
  using Cfg = BwdCfg<T, K, S, R, RY, D, CB>;
  auto kern = dw2_bwd_kernel<T, K, S, R, RY, D, MINB, CB>;
  int32_t r = set_smem(kern, Cfg::SMEM);
  if (r) return r;
  Geom g{};
  g.N = N; g.C = C; g.Hd = H; g.Wd = W;
  g.Hs = (H + S - 1) / S; g.Ws = (W + S - 1) / S;
  g.pad_t = g.pad_l = Cfg::PAD;
  const int cblocks = (C + CB - 1) / CB;
  static int resident = resident_ctas(kern, Cfg::SMEM);
  plan_strips(g, Cfg::TW, RY, cblocks, resident);
  long items = (long)g.N * g.xs * g.ychunks;
  int gx = resident / cblocks;   // never more CTAs than are co-resident: a partial second wave doubles the time
  if (gx < 1) gx = 1;
  if (gx > items) gx = (int)items;
  PDL_LAUNCH(kern, dim3(gx, cblocks), NT, Cfg::SMEM, st, *gin, w, *in, (int)apply_act_bwd, (const T*)res, ld_res, (T*)out,
             ld_out, stats, dw, g);
  return check_launch("dwconv2_bwd");
}

// fast forward: returns 1 when the geometry is not exact (the caller then uses the generic kernel)
template <int K, int S, int R, int RY, int D, int MINB, int CB, bool SILU>
static int32_t launch_fwdx(const s2lc_lazy* in, const float* w, void* out, double* stats, int N, int H, int W, int C,
                           cudaStream_t st) {
  using Cfg = FwdCfg<bf16, K, S, R, RY, D, CB>;
  auto kern = dw2_fwdx_kernel<K, S, R, RY, D, MINB, CB, SILU>;
  Geom g{};
  g.N = N; g.C = C; g.Hs = H; g.Ws = W;
  g.Hd = (H + S - 1) / S; g.Wd = (W + S - 1) / S;
  int ph = (g.Hd - 1) * S + K - H; if (ph < 0) ph = 0;
  int pw = (g.Wd - 1) * S + K - W; if (pw < 0) pw = 0;
  g.pad_t = ph / 2; g.pad_l = pw / 2;
  const int cblocks = (C + CB - 1) / CB;
  int32_t r = set_smem(kern, Cfg::SMEM);
  if (r) return r;
  static int resident = resident_ctas(kern, Cfg::SMEM);
  plan_strips(g, Cfg::TW, RY, cblocks, resident);
  if (g.Wd % Cfg::TW != 0 || g.Hd % g.YC != 0 || g.YC % RY != 0) return 1;
  long items = (long)g.N * g.xs * g.ychunks;
  int gx = resident / cblocks;
  if (gx < 1) gx = 1;
  if (gx > items) gx = (int)items;
  PDL_LAUNCH(kern, dim3(gx, cblocks), NT, Cfg::SMEM, st, *in, w, (bf16*)out, stats, g);
  return check_launch("dwconv2_fwdx");
}

// fast variant: returns 1 when the geometry is not exact (the caller then uses the generic kernel)
template <int K, int S, int R, int RY, int D, int MINB, int CB, bool SILU, bool ONEBAR, bool TWOPH>
static int32_t launch_bwdx(const s2lc_lazygrad* gin, const float* w, const s2lc_lazy* in, const void* res,
                           int64_t ld_res, void* out, int64_t ld_out, double* stats, float* dw, int N, int H, int W,
                           int C, cudaStream_t st) {
  using X = BwdxCfg<K, S, R, RY, D, CB, ONEBAR>;
  static_assert(MINB == 1 || X::SMEM <= 115712, "two CTAs per SM must fit");
  auto kern = dw2_bwdx_kernel<K, S, R, RY, D, MINB, CB, SILU, ONEBAR, TWOPH>;
  Geom g{};
  g.N = N; g.C = C; g.Hd = H; g.Wd = W;
  g.Hs = (H + S - 1) / S; g.Ws = (W + S - 1) / S;
  g.pad_t = g.pad_l = X::PAD;
  const int cblocks = (C + CB - 1) / CB;
  int32_t r = set_smem(kern, X::SMEM);
  if (r) return r;
  static int resident = resident_ctas(kern, X::SMEM);
  plan_strips(g, X::TW, RY, cblocks, resident);
  if (W % X::TW != 0 || H % g.YC != 0 || g.YC % RY != 0) return 1;
  long items = (long)g.N * g.xs * g.ychunks;
  int gx = resident / cblocks;
  if (gx < 1) gx = 1;
  if (gx > items) gx = (int)items;
  PDL_LAUNCH(kern, dim3(gx, cblocks), NT, X::SMEM, st, *gin, w, *in, (const bf16*)res, ld_res, (bf16*)out, ld_out, stats,
             dw, g);
  return check_launch("dwconv2_bwdx");
}

}  // namespace dw2
}  // namespace s2lc

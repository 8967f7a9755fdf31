// dwconv.cu — depthwise k x k convolution (k in {3,5}, stride in {1,2}, TF-"same" asymmetric padding),
// forward, data-gradient and weight-gradient, fused with the BatchNorm/SiLU of the neighbouring layers.
//
// HBM-bound design (reference call site: src/modules/efficientnet_unet.py:329-344):
//  * NHWC; a CTA stages one (input tile + halo) x 64-channel block in shared memory ONCE, applying the
//    producer's BatchNorm scale/shift + SiLU on the way in (16 B vector global loads, 16 B smem stores),
//    so the activated tensor is never materialised in HBM and SiLU is evaluated once per element;
//  * a warp then owns a strip of 8 output pixels; lane l owns channel pair (2l, 2l+1): smem reads are
//    4 B/lane, conflict-free, each value loaded once per strip row and reused across the k taps from
//    registers (sliding window); the k*k*2 filter taps live in registers;
//  * the epilogue writes the raw conv output (coalesced 128 B per pixel per warp) and accumulates the
//    per-channel BatchNorm moments in registers across all tiles a CTA processes (persistent grid-stride
//    loop), flushed once per CTA: smem atomics -> 128 fp64 global atomics.
#include "common.cuh"

namespace s2lc {

constexpr int DW_CB = 64;  // channels per CTA block
constexpr int DW_TW = 16;  // output tile width
constexpr int DW_R = 8;    // output pixels per strip

template <typename T> __device__ __forceinline__ float2 ld2(const T* p);
template <> __device__ __forceinline__ float2 ld2<bf16>(const bf16* p) {
  return __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(p));
}
template <> __device__ __forceinline__ float2 ld2<float>(const float* p) { return *reinterpret_cast<const float2*>(p); }
template <typename T> __device__ __forceinline__ void st2(T* p, float a, float b);
template <> __device__ __forceinline__ void st2<bf16>(bf16* p, float a, float b) {
  *reinterpret_cast<__nv_bfloat162*>(p) = __floats2bfloat162_rn(a, b);
}
template <> __device__ __forceinline__ void st2<float>(float* p, float a, float b) {
  *reinterpret_cast<float2*>(p) = make_float2(a, b);
}

struct DwGeom {
  int N, H, W, C;      // input-resolution tensor
  int Ho, Wo;          // output-resolution tensor
  int pad_t, pad_l;
  int tiles_x, tiles_y;
};

// Stage an (IH x IW) pixel x 64-channel tile into smem.  `fetch(row, c8, n, v)` yields 8 transformed
// channel values of global pixel row `row` (NHWC row index) starting at channel c8.
template <typename T, typename Fetch>
__device__ __forceinline__ void stage_tile(T* tile, int IH, int IW, int y0, int x0, int n, int Hs, int Ws, int cb0,
                                           int C, Fetch fetch) {
  const int total = IH * IW * (DW_CB / 8);
  for (int idx = threadIdx.x; idx < total; idx += blockDim.x) {
    int v8 = idx & 7, pix = idx >> 3;
    int ty = pix / IW, tx = pix - ty * IW;
    int gy = y0 + ty, gx = x0 + tx;
    int c8 = cb0 + v8 * 8;
    float v[8];
    if (gy >= 0 && gy < Hs && gx >= 0 && gx < Ws && c8 < C) {
      fetch(((int64_t)n * Hs + gy) * Ws + gx, c8, n, v);
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = 0.f;
    }
    Vec8<T>::store(tile + (int64_t)pix * DW_CB + v8 * 8, v);
  }
}

template <typename T>
struct LazyFetch {
  s2lc_lazy L;
  __device__ __forceinline__ void operator()(int64_t row, int c8, int n, float (&v)[8]) const {
    Vec8<T>::load(reinterpret_cast<const T*>(L.raw) + row * L.ld + c8, v);
    if (L.scale) {
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = v[j] * L.scale[c8 + j] + (L.shift ? L.shift[c8 + j] : 0.f);
    }
    if (L.act != S2LC_ACT_NONE) {
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = act_fwd_t<T>(v[j], L.act);
    }
  }
};
template <typename T>
struct GradFetch {
  s2lc_lazygrad G;
  __device__ __forceinline__ void operator()(int64_t row, int c8, int n, float (&v)[8]) const {
    Vec8<T>::load(reinterpret_cast<const T*>(G.dy) + row * G.ld_dy + c8, v);
    if (G.rowscale) {
      float rs = G.rowscale[n];
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] *= rs;
    }
    if (G.z) {
      float z[8];
      Vec8<T>::load(reinterpret_cast<const T*>(G.z) + row * G.ld_z + c8, z);
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = G.p[c8 + j] * v[j] + G.q[c8 + j] * z[j] + G.t[c8 + j];
    }
  }
};

// sliding-window stencil for one strip of R outputs and one channel pair
template <typename T, int K, int S>
__device__ __forceinline__ void stencil_strip(const T* base, int IW, const float (&w)[K * K][2],
                                              float (&acc)[DW_R][2]) {
#pragma unroll
  for (int ky = 0; ky < K; ++ky) {
#pragma unroll
    for (int j = 0; j < (DW_R - 1) * S + K; ++j) {
      float2 v = ld2<T>(base + (int64_t)(ky * IW + j) * DW_CB);
#pragma unroll
      for (int kx = 0; kx < K; ++kx) {
        if (j - kx >= 0 && (j - kx) % S == 0 && (j - kx) / S < DW_R) {
          acc[(j - kx) / S][0] = fmaf(w[ky * K + kx][0], v.x, acc[(j - kx) / S][0]);
          acc[(j - kx) / S][1] = fmaf(w[ky * K + kx][1], v.y, acc[(j - kx) / S][1]);
        }
      }
    }
  }
}

__device__ __forceinline__ void flush_pair_stats(float s0a, float s0b, float s1a, float s1b, float* sstat, int cb0,
                                                 int C, double* stats) {
  // sstat: [2][64] floats, zeroed by the caller before the tile loop
  const int lane = threadIdx.x & 31;
  atomicAdd(&sstat[2 * lane], s0a);
  atomicAdd(&sstat[2 * lane + 1], s0b);
  atomicAdd(&sstat[DW_CB + 2 * lane], s1a);
  atomicAdd(&sstat[DW_CB + 2 * lane + 1], s1b);
  __syncthreads();
  if (threadIdx.x < 2 * DW_CB) {
    int s = threadIdx.x / DW_CB, c = cb0 + threadIdx.x % DW_CB;
    if (c < C) atomicAdd(&stats[(int64_t)s * C + c], (double)sstat[threadIdx.x]);
  }
}

// MODE 0: forward (in = lazy activation, out = raw conv output + MOMENTS stats)
// MODE 1: data gradient for stride 1 (in = lazy gradient, flipped taps; epilogue act'(prev) + residual + BNBWD stats)
template <typename T, int K, int S, int TH, int MODE>
__global__ void __launch_bounds__(256) dwconv_kernel(s2lc_lazy in, s2lc_lazygrad gin, const float* __restrict__ w,
                                                     s2lc_lazy prev, int has_prev, const T* __restrict__ res,
                                                     int64_t ld_res, T* __restrict__ out, int64_t ld_out,
                                                     double* stats, DwGeom g) {
  constexpr int IH = (TH - 1) * S + K, IW = (DW_TW - 1) * S + K;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* tile = reinterpret_cast<T*>(smem_raw);
  float* sstat = reinterpret_cast<float*>(smem_raw + sizeof(T) * IH * IW * DW_CB);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int cb0 = blockIdx.y * DW_CB;
  const int c = cb0 + 2 * lane;
  const bool cvalid = c < g.C;
  if (threadIdx.x < 2 * DW_CB) sstat[threadIdx.x] = 0.f;

  float wr[K * K][2];
#pragma unroll
  for (int t = 0; t < K * K; ++t) {
    int ts = MODE == 1 ? (K * K - 1 - t) : t;  // flipped taps for the data gradient
    wr[t][0] = cvalid ? w[(int64_t)c * K * K + ts] : 0.f;
    wr[t][1] = cvalid ? w[(int64_t)(c + 1) * K * K + ts] : 0.f;
  }
  float psc0 = 1.f, psc1 = 1.f, psh0 = 0.f, psh1 = 0.f;
  if (MODE == 1 && has_prev && cvalid && prev.scale) {
    psc0 = prev.scale[c]; psc1 = prev.scale[c + 1];
    psh0 = prev.shift ? prev.shift[c] : 0.f; psh1 = prev.shift ? prev.shift[c + 1] : 0.f;
  }
  float s0a = 0.f, s0b = 0.f, s1a = 0.f, s1b = 0.f;

  // geometry of the tensor being read (source) and written (destination)
  const int Hs = MODE == 0 ? g.H : g.Ho, Ws = MODE == 0 ? g.W : g.Wo;
  const int Hd = MODE == 0 ? g.Ho : g.H, Wd = MODE == 0 ? g.Wo : g.W;
  const int off_y = MODE == 0 ? g.pad_t : (K - 1 - g.pad_t);
  const int off_x = MODE == 0 ? g.pad_l : (K - 1 - g.pad_l);
  const int n_tiles = g.N * g.tiles_y * g.tiles_x;
  LazyFetch<T> lf{in};
  GradFetch<T> gf{gin};

  for (int tile_id = blockIdx.x; tile_id < n_tiles; tile_id += gridDim.x) {
    int n = tile_id / (g.tiles_y * g.tiles_x);
    int rem = tile_id - n * g.tiles_y * g.tiles_x;
    int ty0 = (rem / g.tiles_x) * TH, tx0 = (rem % g.tiles_x) * DW_TW;
    __syncthreads();  // previous tile's readers are done
    if (MODE == 0)
      stage_tile<T>(tile, IH, IW, ty0 * S - off_y, tx0 * S - off_x, n, Hs, Ws, cb0, g.C, lf);
    else
      stage_tile<T>(tile, IH, IW, ty0 * S - off_y, tx0 * S - off_x, n, Hs, Ws, cb0, g.C, gf);
    __syncthreads();
    constexpr int STRIPS = TH * (DW_TW / DW_R);
    for (int sidx = warp; sidx < STRIPS; sidx += 8) {
      int sy = sidx / (DW_TW / DW_R), sx = (sidx % (DW_TW / DW_R)) * DW_R;
      int oy = ty0 + sy;
      if (oy >= Hd) continue;  // warp-uniform
      float acc[DW_R][2] = {};
      stencil_strip<T, K, S>(tile + ((int64_t)(sy * S) * IW + sx * S) * DW_CB + 2 * lane, IW, wr, acc);
      if (!cvalid) continue;
#pragma unroll
      for (int r = 0; r < DW_R; ++r) {
        int ox = tx0 + sx + r;
        if (ox >= Wd) break;
        int64_t orow = ((int64_t)n * Hd + oy) * Wd + ox;
        float a = acc[r][0], b = acc[r][1];
        if (MODE == 0) {
          a = roundf_as<T>(a); b = roundf_as<T>(b);
          st2<T>(out + orow * ld_out + c, a, b);
          s0a += a; s0b += b; s1a += a * a; s1b += b * b;
        } else {
          float za = 0.f, zb = 0.f;
          if (has_prev) {
            float2 z = ld2<T>(reinterpret_cast<const T*>(prev.raw) + orow * prev.ld + c);
            za = z.x; zb = z.y;
            a *= act_bwd_t<T>(za * psc0 + psh0, prev.act);
            b *= act_bwd_t<T>(zb * psc1 + psh1, prev.act);
          }
          if (res) {
            float2 e = ld2<T>(res + orow * ld_res + c);
            a += e.x; b += e.y;
          }
          a = roundf_as<T>(a); b = roundf_as<T>(b);
          st2<T>(out + orow * ld_out + c, a, b);
          s0a += a; s0b += b; s1a += a * za; s1b += b * zb;
        }
      }
    }
  }
  if (stats) {
    __syncthreads();
    flush_pair_stats(s0a, s0b, s1a, s1b, sstat, cb0, g.C, stats);
  }
}

// Data gradient for stride 2: tile over INPUT pixels, gather the parity-matching taps from the staged
// gradient tile.  All predicates are warp-uniform (a warp works on one pixel strip, lanes = channels).
template <typename T, int K, int TH>
__global__ void __launch_bounds__(256) dwconv_bwd_data_s2_kernel(s2lc_lazygrad gin, const float* __restrict__ w,
                                                                 s2lc_lazy prev, int has_prev,
                                                                 const T* __restrict__ res, int64_t ld_res,
                                                                 T* __restrict__ out, int64_t ld_out, double* stats,
                                                                 DwGeom g) {
  constexpr int S = 2;
  constexpr int GH = (TH + K - 2) / S + 2, GW = (DW_TW + K - 2) / S + 2;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* tile = reinterpret_cast<T*>(smem_raw);
  float* sstat = reinterpret_cast<float*>(smem_raw + sizeof(T) * GH * GW * DW_CB);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int cb0 = blockIdx.y * DW_CB;
  const int c = cb0 + 2 * lane;
  const bool cvalid = c < g.C;
  if (threadIdx.x < 2 * DW_CB) sstat[threadIdx.x] = 0.f;
  float wr[K * K][2];
#pragma unroll
  for (int t = 0; t < K * K; ++t) {
    wr[t][0] = cvalid ? w[(int64_t)c * K * K + t] : 0.f;
    wr[t][1] = cvalid ? w[(int64_t)(c + 1) * K * K + t] : 0.f;
  }
  float psc0 = 1.f, psc1 = 1.f, psh0 = 0.f, psh1 = 0.f;
  if (has_prev && cvalid && prev.scale) {
    psc0 = prev.scale[c]; psc1 = prev.scale[c + 1];
    psh0 = prev.shift ? prev.shift[c] : 0.f; psh1 = prev.shift ? prev.shift[c + 1] : 0.f;
  }
  float s0a = 0.f, s0b = 0.f, s1a = 0.f, s1b = 0.f;
  const int n_tiles = g.N * g.tiles_y * g.tiles_x;
  GradFetch<T> gf{gin};
  for (int tile_id = blockIdx.x; tile_id < n_tiles; tile_id += gridDim.x) {
    int n = tile_id / (g.tiles_y * g.tiles_x);
    int rem = tile_id - n * g.tiles_y * g.tiles_x;
    int iy0 = (rem / g.tiles_x) * TH, ix0 = (rem % g.tiles_x) * DW_TW;
    // first output row/col that any pixel of this tile can touch (floor division, may be negative)
    int ay = iy0 + g.pad_t - (K - 1), ax = ix0 + g.pad_l - (K - 1);
    int goy0 = ay >= 0 ? ay / S : -((-ay + S - 1) / S);
    int gox0 = ax >= 0 ? ax / S : -((-ax + S - 1) / S);
    __syncthreads();
    stage_tile<T>(tile, GH, GW, goy0, gox0, n, g.Ho, g.Wo, cb0, g.C, gf);
    __syncthreads();
    constexpr int STRIPS = TH * (DW_TW / DW_R);
    for (int sidx = warp; sidx < STRIPS; sidx += 8) {
      int sy = sidx / (DW_TW / DW_R), sx = (sidx % (DW_TW / DW_R)) * DW_R;
      int iy = iy0 + sy;
      if (iy >= g.H) continue;
      float acc[DW_R][2] = {};
#pragma unroll
      for (int ky = 0; ky < K; ++ky) {
        int tyy = iy + g.pad_t - ky;
        if (tyy < 0 || (tyy & 1)) continue;  // warp-uniform
        int gy = tyy / S - goy0;             // row inside the staged tile (zero-filled outside the tensor)
#pragma unroll
        for (int r = 0; r < DW_R; ++r) {
#pragma unroll
          for (int kx = 0; kx < K; ++kx) {
            int txx = ix0 + sx + r + g.pad_l - kx;
            if (txx < 0 || (txx & 1)) continue;
            int gx = txx / S - gox0;
            float2 v = ld2<T>(tile + ((int64_t)gy * GW + gx) * DW_CB + 2 * lane);
            acc[r][0] = fmaf(wr[ky * K + kx][0], v.x, acc[r][0]);
            acc[r][1] = fmaf(wr[ky * K + kx][1], v.y, acc[r][1]);
          }
        }
      }
      if (!cvalid) continue;
#pragma unroll
      for (int r = 0; r < DW_R; ++r) {
        int ix = ix0 + sx + r;
        if (ix >= g.W) break;
        int64_t orow = ((int64_t)n * g.H + iy) * g.W + ix;
        float a = acc[r][0], b = acc[r][1];
        float za = 0.f, zb = 0.f;
        if (has_prev) {
          float2 z = ld2<T>(reinterpret_cast<const T*>(prev.raw) + orow * prev.ld + c);
          za = z.x; zb = z.y;
          a *= act_bwd_t<T>(za * psc0 + psh0, prev.act);
          b *= act_bwd_t<T>(zb * psc1 + psh1, prev.act);
        }
        if (res) {
          float2 e = ld2<T>(res + orow * ld_res + c);
          a += e.x; b += e.y;
        }
        a = roundf_as<T>(a); b = roundf_as<T>(b);
        st2<T>(out + orow * ld_out + c, a, b);
        s0a += a; s0b += b; s1a += a * za; s1b += b * zb;
      }
    }
  }
  if (stats) {
    __syncthreads();
    flush_pair_stats(s0a, s0b, s1a, s1b, sstat, cb0, g.C, stats);
  }
}

// Weight gradient: dW[c, ky, kx] = sum_{n, oy, ox} g[n, oy, ox, c] * in_act[n, oy*S+ky-pt, ox*S+kx-pl, c]
template <typename T, int K, int S, int TH>
__global__ void __launch_bounds__(256) dwconv_bwd_weight_kernel(s2lc_lazygrad gin, s2lc_lazy in,
                                                                float* __restrict__ dw, DwGeom g) {
  constexpr int IH = (TH - 1) * S + K, IW = (DW_TW - 1) * S + K;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* tile = reinterpret_cast<T*>(smem_raw);
  float* sred = reinterpret_cast<float*>(smem_raw + sizeof(T) * IH * IW * DW_CB);  // [K*K][64]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int cb0 = blockIdx.y * DW_CB;
  const int c = cb0 + 2 * lane;
  const bool cvalid = c < g.C;
  for (int i = threadIdx.x; i < K * K * DW_CB; i += blockDim.x) sred[i] = 0.f;
  float gp0 = 1.f, gp1 = 1.f, gq0 = 0.f, gq1 = 0.f, gt0 = 0.f, gt1 = 0.f;
  if (gin.z && cvalid) {
    gp0 = gin.p[c]; gp1 = gin.p[c + 1]; gq0 = gin.q[c]; gq1 = gin.q[c + 1]; gt0 = gin.t[c]; gt1 = gin.t[c + 1];
  }
  float wacc[K * K][2] = {};
  const int n_tiles = g.N * g.tiles_y * g.tiles_x;
  LazyFetch<T> lf{in};
  for (int tile_id = blockIdx.x; tile_id < n_tiles; tile_id += gridDim.x) {
    int n = tile_id / (g.tiles_y * g.tiles_x);
    int rem = tile_id - n * g.tiles_y * g.tiles_x;
    int ty0 = (rem / g.tiles_x) * TH, tx0 = (rem % g.tiles_x) * DW_TW;
    __syncthreads();
    stage_tile<T>(tile, IH, IW, ty0 * S - g.pad_t, tx0 * S - g.pad_l, n, g.H, g.W, cb0, g.C, lf);
    __syncthreads();
    constexpr int STRIPS = TH * (DW_TW / DW_R);
    const float rs = gin.rowscale ? gin.rowscale[n] : 1.f;
    for (int sidx = warp; sidx < STRIPS; sidx += 8) {
      int sy = sidx / (DW_TW / DW_R), sx = (sidx % (DW_TW / DW_R)) * DW_R;
      int oy = ty0 + sy;
      if (oy >= g.Ho || !cvalid) continue;
      float gv[DW_R][2];
#pragma unroll
      for (int r = 0; r < DW_R; ++r) {
        int ox = tx0 + sx + r;
        gv[r][0] = 0.f; gv[r][1] = 0.f;
        if (ox < g.Wo) {
          int64_t row = ((int64_t)n * g.Ho + oy) * g.Wo + ox;
          float2 d = ld2<T>(reinterpret_cast<const T*>(gin.dy) + row * gin.ld_dy + c);
          d.x *= rs; d.y *= rs;
          if (gin.z) {
            float2 z = ld2<T>(reinterpret_cast<const T*>(gin.z) + row * gin.ld_z + c);
            d.x = gp0 * d.x + gq0 * z.x + gt0;
            d.y = gp1 * d.y + gq1 * z.y + gt1;
          }
          gv[r][0] = d.x; gv[r][1] = d.y;
        }
      }
      const T* base = tile + ((int64_t)(sy * S) * IW + sx * S) * DW_CB + 2 * lane;
#pragma unroll
      for (int ky = 0; ky < K; ++ky) {
#pragma unroll
        for (int j = 0; j < (DW_R - 1) * S + K; ++j) {
          float2 v = ld2<T>(base + (int64_t)(ky * IW + j) * DW_CB);
#pragma unroll
          for (int kx = 0; kx < K; ++kx) {
            if (j - kx >= 0 && (j - kx) % S == 0 && (j - kx) / S < DW_R) {
              wacc[ky * K + kx][0] = fmaf(gv[(j - kx) / S][0], v.x, wacc[ky * K + kx][0]);
              wacc[ky * K + kx][1] = fmaf(gv[(j - kx) / S][1], v.y, wacc[ky * K + kx][1]);
            }
          }
        }
      }
    }
  }
  __syncthreads();
#pragma unroll
  for (int t = 0; t < K * K; ++t) {
    atomicAdd(&sred[t * DW_CB + 2 * lane], wacc[t][0]);
    atomicAdd(&sred[t * DW_CB + 2 * lane + 1], wacc[t][1]);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < K * K * DW_CB; i += blockDim.x) {
    int t = i / DW_CB, cc = cb0 + i % DW_CB;
    if (cc < g.C) atomicAdd(&dw[(int64_t)cc * K * K + t], sred[i]);
  }
}

static DwGeom make_geom(int N, int H, int W, int C, int k, int s, int TH, bool tile_over_input) {
  DwGeom g;
  g.N = N; g.H = H; g.W = W; g.C = C;
  g.Ho = (H + s - 1) / s; g.Wo = (W + s - 1) / s;
  int ph = (g.Ho - 1) * s + k - H; if (ph < 0) ph = 0;
  int pw = (g.Wo - 1) * s + k - W; if (pw < 0) pw = 0;
  g.pad_t = ph / 2; g.pad_l = pw / 2;
  int th = tile_over_input ? H : g.Ho, tw = tile_over_input ? W : g.Wo;
  g.tiles_y = (th + TH - 1) / TH; g.tiles_x = (tw + DW_TW - 1) / DW_TW;
  return g;
}

static int dw_grid_x(const DwGeom& g, int cblocks, int ctas_per_sm) {
  int n_tiles = g.N * g.tiles_y * g.tiles_x;
  int want = (kNumSMs * ctas_per_sm + cblocks - 1) / cblocks;
  if (want < 1) want = 1;
  return n_tiles < want ? n_tiles : want;
}

template <typename K>
static int32_t set_smem(K kernel, size_t bytes) {
  if (bytes > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e != cudaSuccess) {
      set_error("cudaFuncSetAttribute(smem=%zu): %s", bytes, cudaGetErrorString(e));
      return S2LC_ERR_CUDA;
    }
  }
  return S2LC_OK;
}

template <typename T, int K, int S, int TH, int MODE>
static int32_t launch_dw(const s2lc_lazy* in, const s2lc_lazygrad* gin, const float* w, const s2lc_lazy* prev,
                         const void* res, int64_t ld_res, void* out, int64_t ld_out, double* stats, int N, int H, int W,
                         int C, cudaStream_t st) {
  constexpr int IH = (TH - 1) * S + K, IW = (DW_TW - 1) * S + K;
  size_t smem = sizeof(T) * IH * IW * DW_CB + 2 * DW_CB * sizeof(float);
  DwGeom g = make_geom(N, H, W, C, K, S, TH, MODE == 1);
  auto kern = dwconv_kernel<T, K, S, TH, MODE>;
  int32_t r = set_smem(kern, smem);
  if (r) return r;
  int cblocks = (C + DW_CB - 1) / DW_CB;
  int per_sm = (int)(200 * 1024 / smem); if (per_sm > 4) per_sm = 4; if (per_sm < 1) per_sm = 1;
  dim3 grid(dw_grid_x(g, cblocks, per_sm), cblocks);
  s2lc_lazy zin{}; s2lc_lazygrad zg{}; s2lc_lazy zp{};
  kern<<<grid, 256, smem, st>>>(in ? *in : zin, gin ? *gin : zg, w, prev ? *prev : zp, prev != nullptr,
                                (const T*)res, ld_res, (T*)out, ld_out, stats, g);
  return check_launch("dwconv");
}

template <typename T, int K, int TH>
static int32_t launch_dw_bwd_s2(const s2lc_lazygrad* gin, const float* w, const s2lc_lazy* prev, const void* res,
                                int64_t ld_res, void* out, int64_t ld_out, double* stats, int N, int H, int W, int C,
                                cudaStream_t st) {
  constexpr int GH = (TH + K - 2) / 2 + 2, GW = (DW_TW + K - 2) / 2 + 2;
  size_t smem = sizeof(T) * GH * GW * DW_CB + 2 * DW_CB * sizeof(float);
  DwGeom g = make_geom(N, H, W, C, K, 2, TH, true);
  auto kern = dwconv_bwd_data_s2_kernel<T, K, TH>;
  int32_t r = set_smem(kern, smem);
  if (r) return r;
  int cblocks = (C + DW_CB - 1) / DW_CB;
  dim3 grid(dw_grid_x(g, cblocks, 4), cblocks);
  s2lc_lazy zp{};
  kern<<<grid, 256, smem, st>>>(*gin, w, prev ? *prev : zp, prev != nullptr, (const T*)res, ld_res, (T*)out, ld_out,
                                stats, g);
  return check_launch("dwconv_bwd_data_s2");
}

template <typename T, int K, int S, int TH>
static int32_t launch_dw_wgrad(const s2lc_lazygrad* gin, const s2lc_lazy* in, float* dw, int N, int H, int W, int C,
                               cudaStream_t st) {
  constexpr int IH = (TH - 1) * S + K, IW = (DW_TW - 1) * S + K;
  size_t smem = sizeof(T) * IH * IW * DW_CB + K * K * DW_CB * sizeof(float);
  DwGeom g = make_geom(N, H, W, C, K, S, TH, false);
  auto kern = dwconv_bwd_weight_kernel<T, K, S, TH>;
  int32_t r = set_smem(kern, smem);
  if (r) return r;
  int cblocks = (C + DW_CB - 1) / DW_CB;
  int per_sm = (int)(200 * 1024 / smem); if (per_sm > 3) per_sm = 3; if (per_sm < 1) per_sm = 1;
  dim3 grid(dw_grid_x(g, cblocks, per_sm), cblocks);
  kern<<<grid, 256, smem, st>>>(*gin, *in, dw, g);
  return check_launch("dwconv_bwd_weight");
}

}  // namespace s2lc

using namespace s2lc;

#define DW_REQ(C, lds)                                                                                           \
  S2LC_REQUIRE((k == 3 || k == 5) && (s == 1 || s == 2), S2LC_ERR_SHAPE, "dwconv: k=%d s=%d unsupported", k, s); \
  S2LC_REQUIRE((C) % 8 == 0 && (lds), S2LC_ERR_ALIGN, "dwconv: C and strides must be multiples of 8 (C=%d)", (C));

// stride-2 tiles are 4 rows tall for fp32 (smem) and 8 for bf16
#define DW_DISPATCH(FN, ...)                                                                   \
  if (dtype == S2LC_F32) {                                                                     \
    if (k == 3 && s == 1) return FN(float, 3, 1, 8, __VA_ARGS__);                              \
    if (k == 5 && s == 1) return FN(float, 5, 1, 8, __VA_ARGS__);                              \
    if (k == 3 && s == 2) return FN(float, 3, 2, 4, __VA_ARGS__);                              \
    return FN(float, 5, 2, 4, __VA_ARGS__);                                                    \
  } else if (dtype == S2LC_BF16) {                                                             \
    if (k == 3 && s == 1) return FN(s2lc::bf16, 3, 1, 8, __VA_ARGS__);                         \
    if (k == 5 && s == 1) return FN(s2lc::bf16, 5, 1, 8, __VA_ARGS__);                         \
    if (k == 3 && s == 2) return FN(s2lc::bf16, 3, 2, 8, __VA_ARGS__);                         \
    return FN(s2lc::bf16, 5, 2, 8, __VA_ARGS__);                                               \
  }                                                                                            \
  s2lc::set_error("dwconv: bad dtype %d", dtype);                                              \
  return S2LC_ERR_SHAPE;

extern "C" {

int32_t s2lc_dwconv_fwd_tile(const s2lc_lazy* in, const float* w, void* out, double* stats, int32_t N, int32_t H, int32_t W,
                        int32_t C, int32_t k, int32_t s, int32_t dtype, void* stream) {
  DW_REQ(C, in->ld % 8 == 0)
  cudaStream_t st = (cudaStream_t)stream;
#define FWD(T, K, S, TH, ...) \
  launch_dw<T, K, S, TH, 0>(in, nullptr, w, nullptr, nullptr, 0, out, C, stats, N, H, W, C, st)
  DW_DISPATCH(FWD, 0)
#undef FWD
}

int32_t s2lc_dwconv_bwd_data(const s2lc_lazygrad* g, const float* w, const s2lc_lazy* prev, const void* res,
                             int64_t ld_res, void* out, int64_t ld_out, double* stats, int32_t N, int32_t H, int32_t W,
                             int32_t C, int32_t k, int32_t s, int32_t dtype, void* stream) {
  DW_REQ(C, g->ld_dy % 8 == 0 && ld_out % 8 == 0)
  cudaStream_t st = (cudaStream_t)stream;
  if (s == 1) {
#define BWD1(T, K, S, TH, ...) \
  launch_dw<T, K, 1, 8, 1>(nullptr, g, w, prev, res, ld_res, out, ld_out, stats, N, H, W, C, st)
    if (dtype == S2LC_F32) { if (k == 3) return BWD1(float, 3, 1, 8, 0); return BWD1(float, 5, 1, 8, 0); }
    if (dtype == S2LC_BF16) { if (k == 3) return BWD1(s2lc::bf16, 3, 1, 8, 0); return BWD1(s2lc::bf16, 5, 1, 8, 0); }
#undef BWD1
  } else {
#define BWD2(T, K) launch_dw_bwd_s2<T, K, 8>(g, w, prev, res, ld_res, out, ld_out, stats, N, H, W, C, st)
    if (dtype == S2LC_F32) { if (k == 3) return BWD2(float, 3); return BWD2(float, 5); }
    if (dtype == S2LC_BF16) { if (k == 3) return BWD2(s2lc::bf16, 3); return BWD2(s2lc::bf16, 5); }
#undef BWD2
  }
  s2lc::set_error("dwconv_bwd_data: bad dtype %d", dtype);
  return S2LC_ERR_SHAPE;
}

int32_t s2lc_dwconv_bwd_weight(const s2lc_lazygrad* g, const s2lc_lazy* in, float* dw, int32_t N, int32_t H, int32_t W,
                               int32_t C, int32_t k, int32_t s, int32_t dtype, void* stream) {
  DW_REQ(C, g->ld_dy % 8 == 0 && in->ld % 8 == 0)
  cudaStream_t st = (cudaStream_t)stream;
#define WG(T, K, S, TH, ...) launch_dw_wgrad<T, K, S, TH>(g, in, dw, N, H, W, C, st)
  DW_DISPATCH(WG, 0)
#undef WG
}

}  // extern "C"

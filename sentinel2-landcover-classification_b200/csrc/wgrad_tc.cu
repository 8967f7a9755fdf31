// wgrad_tc.cu — tcgen05 / TMEM weight-gradient engine for the 1x1, 3x3 and 2x2-transposed convolutions.
//
//   dW[n, (tap, c)] = sum_{pixels m} G(m, n) * A(m shifted by tap, c)
//
// (reference: autograd of F.conv2d / nn.ConvTranspose2d at src/modules/efficientnet_unet.py:297, :115-119,
// :168-176).  The reduction runs over PIXELS, i.e. over the row index of both NHWC tensors, so both GEMM operands
// are "MN-major": their non-reduced dimension (channels) is the contiguous one in HBM.  tcgen05.mma reads MN-major
// operands natively (instruction-descriptor bits 15/16), so a 16 B global load of 8 channels of one pixel goes —
// after the lazy transform — into shared memory as ONE 16 B store with no transpose:
//     G = p*dy*rowscale + q*z + t            (BatchNorm backward + drop-connect, include/s2lc.h s2lc_lazygrad)
//     A = act(raw*scale + shift) * gate      (BatchNorm + SiLU/ReLU + squeeze-excite, s2lc_lazy)
// Layout in smem: SWIZZLE_128B MN-major atoms of 8 pixels x 64 channels (1 KB); a 64-pixel pipeline block of a
// 64-channel group is 8 KB (SBO = 1 KB between 8-pixel groups), channel groups are LBO = 8 KB apart.
//
// Warp specialisation: 16 loader warps (gather + transform + st.shared, arrive on `full` mbarriers), 1 MMA warp
// (one elected lane issues tcgen05.mma, tcgen05.commit releases the slots), accumulators live in TMEM for the
// whole pixel slab; the loader warps become the epilogue (tcgen05.ld -> fp32 atomicAdd into the reference-layout
// gradient).  For the 3x3 convs ONE CTA accumulates all nine taps (9 accumulators of [128 x 48] fp32 = 432 TMEM
// columns) so the gradient tile G is fetched once per pixel block and the nine shifted A tiles come from L1/L2.
// Split-K over pixel slabs fills the 148 SMs; there is no CPU or library fallback.
#include "conv_common.cuh"
#include "tc_common.cuh"
#include "stream_common.cuh"

namespace s2lc {

using namespace tc;

constexpr int WG_BK = 64;                 // pixels per pipeline block (GEMM K per stage)
constexpr int WG_LOAD_WARPS = 16;
constexpr int WG_LOAD_THREADS = WG_LOAD_WARPS * 32;
constexpr int WG_THREADS = WG_LOAD_THREADS + 32;
constexpr int WG_GROUP_BYTES = 8192;      // 64 pixels x 64 channels bf16
constexpr int WG_MAX_STAGES = 8;

struct WgPlan {
  int g_on_m;       // 1: G channels on the M (TMEM lane) side and A on the N side; 0: swapped
  int nt;           // N-side tile width in channels (multiple of 8, <= 256)
  int ntpad;        // MMA N (multiple of 16)
  int n_ntiles;     // tiles along the N side
  int taps;         // 9 (3x3) or 1
  int stages;       // ring depth; one stage = the G tile + `taps` A tiles of one 64-pixel block
  int gsz, asz;     // tile bytes
  int tmem_cols;
  int slabs;
  int64_t px_per_slab;   // multiple of WG_BK
};

__device__ __forceinline__ uint32_t wg_off(int k, int cc) {   // pixel k (0..63), 16 B chunk cc of the tile
  return (uint32_t)((cc >> 3) * WG_GROUP_BYTES + (k >> 3) * 1024 + (k & 7) * 128 + (((cc & 7) ^ (k & 7)) << 4));
}
__device__ __forceinline__ uint4 wg_ldg(const bf16* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }
__device__ __forceinline__ void wg_unpack(const uint4& u, float (&v)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 f = __bfloat1622float2(h[i]);
    v[2 * i] = f.x;
    v[2 * i + 1] = f.y;
  }
}
__device__ __forceinline__ uint4 wg_pack(const float (&v)[8]) {
  uint4 u;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
  return u;
}
__device__ __forceinline__ void wg_ld8(const float* p, float (&v)[8]) {
  float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p + 4));
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void loaders_sync() { asm volatile("bar.sync 1, %0;" ::"n"(WG_LOAD_THREADS) : "memory"); }

// TAPS: 9 (3x3, all taps accumulated by one CTA) or 1.  MAXPG / MAXPA: compile-time bound on the number of loader
// passes over the G / A tile (a pass = 512 threads x one 16 B chunk); every global load of a pixel block is issued
// before the first transform so that ~13 independent 16 B loads per thread are in flight (HBM latency hiding).
template <int TAPS, int MAXPG, int MAXPA>
__global__ void __launch_bounds__(WG_THREADS, 1) wgrad_tc_kernel(WgradArgs P, WgPlan K) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* smem =
      reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int stage_bytes = K.gsz + TAPS * K.asz;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + K.stages * stage_bytes);
  uint64_t* empty = full + WG_MAX_STAGES;
  uint64_t* done = empty + WG_MAX_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done + 1);
  float* sbias = reinterpret_cast<float*>(tmem_slot + 2);   // [256]

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t M = (int64_t)P.N * P.H * P.W;
  const int64_t mb = (int64_t)blockIdx.x * K.px_per_slab;
  const int64_t me = (mb + K.px_per_slab < M) ? mb + K.px_per_slab : M;
  const int nblk = me > mb ? (int)((me - mb + WG_BK - 1) / WG_BK) : 0;
  const int mt = blockIdx.y / K.n_ntiles, nti = blockIdx.y % K.n_ntiles;
  const int Ng = P.ngroups * P.Cout;
  const int Ca = (P.Cin + 7) & ~7;                       // A channels incl. the zero pad of the buffer
  // channel windows of this CTA
  const int g_c0 = K.g_on_m ? mt * 128 : nti * K.nt;
  const int a_c0 = K.g_on_m ? nti * K.nt : mt * 128;
  const int g_tw = K.g_on_m ? 128 : K.nt, a_tw = K.g_on_m ? K.nt : 128;
  const int g_cw = (Ng - g_c0) < g_tw ? (Ng - g_c0) : g_tw;
  const int a_cw = (Ca - a_c0) < a_tw ? (Ca - a_c0) : a_tw;

  if (tid == 0) {
    for (int s = 0; s < K.stages; ++s) { mbar_init(&full[s], WG_LOAD_WARPS); mbar_init(&empty[s], 1); }
    mbar_init(done, 1);
    fence_barrier_init();
  }
  for (int i = tid; i < 256; i += WG_THREADS) sbias[i] = 0.f;
  if (warp == WG_LOAD_WARPS) tmem_alloc(tmem_slot, (uint32_t)K.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == WG_LOAD_WARPS) {
    // =============================================================== MMA issuer (one elected lane)
    if (nblk > 0) {    // converged warp: descriptors stay in uniform registers, one elected lane issues
      const bool leader = elect_one();
      const uint32_t idesc = idesc_bf16(128, (uint32_t)K.ntpad, 1, 1);   // both operands MN-major
      const uint64_t dbase = smem_desc_sw128(0, 1024, WG_GROUP_BYTES);
      for (int blk = 0; blk < nblk; ++blk) {
        const int s = blk % K.stages;
        mbar_wait(&full[s], (uint32_t)((blk / K.stages) & 1));
        tc_fence_after();
        const uint32_t g_addr = smem_u32(smem + s * stage_bytes);
#pragma unroll
        for (int t = 0; t < TAPS; ++t) {
          const uint32_t a_addr = g_addr + K.gsz + t * K.asz;
          const uint32_t m_addr = K.g_on_m ? g_addr : a_addr, n_addr = K.g_on_m ? a_addr : g_addr;
#pragma unroll
          for (int k4 = 0; k4 < WG_BK / 16; ++k4) {
            const uint64_t dm = dbase + (uint64_t)((m_addr + k4 * 2048) >> 4);
            const uint64_t dn = dbase + (uint64_t)((n_addr + k4 * 2048) >> 4);
            if (leader) umma_bf16(tmem_base + (uint32_t)(t * K.ntpad), dm, dn, idesc, (blk > 0 || k4 > 0) ? 1u : 0u);
          }
        }
        if (leader) umma_commit(&empty[s]);
        __syncwarp();
      }
      if (leader) umma_commit(done);
      __syncwarp();
    }
  } else {
    // =============================================================== loader warps
    // All index arithmetic is 32-bit and INCREMENTAL: each thread computes the (sample, row, column) of its first
    // pixel once (the only divisions of the kernel) and then walks it forward by one pass / one 64-pixel block with
    // compare-and-carry steps — per-chunk 64-bit divisions made the first version of this loop ALU-bound.
    const bf16* dyb = reinterpret_cast<const bf16*>(P.g.dy);
    const bf16* zb = reinterpret_cast<const bf16*>(P.g.z);
    const bf16* ab = reinterpret_cast<const bf16*>(P.a.raw);
    const int H = P.H, W = P.W, HW = P.H * P.W;
    const int mbi = (int)mb, mei = (int)me;
    // fixed (channel chunk, first pixel row) of this thread in the G tile and in the A tile
    const int CWg = g_cw >> 3, CWa = a_cw >> 3;
    const int rppG = WG_LOAD_THREADS / CWg, rppA = WG_LOAD_THREADS / CWa;
    const int cG = tid % CWg, rG = tid / CWg, cA = tid % CWa, rA = tid / CWa;
    const bool actG = rG < rppG, actA = rA < rppA;
    const int gch = g_c0 + cG * 8, ach = a_c0 + cA * 8;
    const bool is_ct = P.mode == S2LC_CONVT_FWD;
    int g_q = 0, g_co = gch;                              // convT: G channel = (2x2 child q, co)
    if (is_ct) { g_q = gch / P.Cout; g_co = gch - g_q * P.Cout; }
    const int g_qy = g_q >> 1, g_qx = g_q & 1;
    const bool g_aff = zb != nullptr, a_aff = P.a.scale != nullptr, a_gate = P.a.gate != nullptr;
    const bool g_rs = P.g.rowscale != nullptr;
    const int a_act = P.a.act;
    const bool a_xf = a_aff || a_gate || a_act != S2LC_ACT_NONE;
    const bool do_bias = P.dbias != nullptr && (K.g_on_m ? nti == 0 : mt == 0);
    const bool g_xf = g_aff || g_rs || do_bias;
    const bool g_pos = is_ct || g_rs, a_pos = TAPS == 9 || a_gate;   // who needs (n, h, w)
    float bsum[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    // steps of the coordinate walk
    const int blkH = WG_BK / W, blkW = WG_BK - blkH * W;
    const int pgH = rppG / W, pgW = rppG - pgH * W, paH = rppA / W, paW = rppA - paH * W;
    auto walk = [&](int& n, int& h, int& w, int dh, int dw) {
      w += dw;
      if (w >= W) { w -= W; ++h; }
      h += dh;
      while (h >= H) { h -= H; ++n; }
    };
    int gn0 = 0, gh0 = 0, gw0 = 0, an0 = 0, ah0 = 0, aw0 = 0;   // coordinates of this thread's pass-0 pixel
    {
      int m = mbi + rG;
      gn0 = m / HW; int r = m - gn0 * HW; gh0 = r / W; gw0 = r - gh0 * W;
      m = mbi + rA;
      an0 = m / HW; r = m - an0 * HW; ah0 = r / W; aw0 = r - ah0 * W;
    }

    for (int blk = 0; blk < nblk; ++blk) {
      const int px0 = mbi + blk * WG_BK;
      const int s = blk % K.stages;
      unsigned char* dstG = smem + s * stage_bytes;
      unsigned char* dstA = dstG + K.gsz;
      // ---------------------------------------------------------- issue every load of this pixel block
      uint4 rdy[MAXPG], rz[MAXPG], ra[TAPS][MAXPA];
      uint32_t okG = 0, okA = 0;
      {
        int n = gn0, h = gh0, w = gw0;
#pragma unroll
        for (int p = 0; p < MAXPG; ++p) {
          rdy[p] = make_uint4(0, 0, 0, 0);
          rz[p] = make_uint4(0, 0, 0, 0);
          const int k = rG + p * rppG;
          const int m = px0 + k;
          if (actG && k < WG_BK && m < mei) {
            int64_t row = m;
            if (is_ct) row = ((int64_t)(n * 2 * H + 2 * h + g_qy)) * (2 * W) + 2 * w + g_qx;
            okG |= 1u << p;
            rdy[p] = wg_ldg(dyb + row * P.g.ld_dy + g_co);
            if (g_aff) rz[p] = wg_ldg(zb + row * P.g.ld_z + g_co);
          }
          if (g_pos) walk(n, h, w, pgH, pgW);
        }
      }
      {
        int n = an0, h = ah0, w = aw0;
#pragma unroll
        for (int p = 0; p < MAXPA; ++p) {
          const int k = rA + p * rppA;
          const int m = px0 + k;
          const bool in = actA && k < WG_BK && m < mei;
          const bf16* src = ab + (int64_t)m * P.a.ld + ach;
#pragma unroll
          for (int t = 0; t < TAPS; ++t) {
            ra[t][p] = make_uint4(0, 0, 0, 0);
            if (TAPS == 9) {
              const int dy = t / 3 - 1, dx = t % 3 - 1;
              const int hh = h + dy, ww = w + dx;
              if (in && hh >= 0 && hh < H && ww >= 0 && ww < W) {
                okA |= 1u << (t * MAXPA + p);
                ra[t][p] = wg_ldg(src + (int64_t)(dy * W + dx) * P.a.ld);
              }
            } else if (in) {
              okA |= 1u << p;
              ra[t][p] = wg_ldg(src);
            }
          }
          if (a_pos) walk(n, h, w, paH, paW);
        }
      }
      // ---------------------------------------------------------- stage free?  transform + store everything
      if (blk >= K.stages) mbar_wait(&empty[s], (uint32_t)((blk / K.stages - 1) & 1));
      {
        int n = gn0, h = gh0, w = gw0;
#pragma unroll
        for (int p = 0; p < MAXPG; ++p) {
          const int k = rG + p * rppG;
          if (actG && k < WG_BK) {
            if (((okG >> p) & 1u) && g_xf) {
              float v[8];
              wg_unpack(rdy[p], v);
              if (g_rs) {
                float rs = __ldg(P.g.rowscale + n);
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] *= rs;
              }
              if (g_aff) {
                float z[8], c[8];
                wg_unpack(rz[p], z);
                wg_ld8(P.g.q + g_co, c);
#pragma unroll
                for (int j = 0; j < 8; ++j) z[j] *= c[j];
                wg_ld8(P.g.t + g_co, c);
#pragma unroll
                for (int j = 0; j < 8; ++j) z[j] += c[j];
                wg_ld8(P.g.p + g_co, c);
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] = fmaf(c[j], v[j], z[j]);
              }
              if (do_bias) {
#pragma unroll
                for (int j = 0; j < 8; ++j) bsum[j] += v[j];
              }
              rdy[p] = wg_pack(v);
            }
            *reinterpret_cast<uint4*>(dstG + wg_off(k, cG)) = rdy[p];
          }
          if (g_rs) walk(n, h, w, pgH, pgW);
        }
      }
      {
        int n = an0, h = ah0, w = aw0;
#pragma unroll
        for (int p = 0; p < MAXPA; ++p) {
          const int k = rA + p * rppA;
          if (actA && k < WG_BK) {
            float sc[8], sh[8], gt8[8];
            if (a_aff) { wg_ld8(P.a.scale + ach, sc); wg_ld8(P.a.shift + ach, sh); }
            if (a_gate && px0 + k < mei) wg_ld8(P.a.gate + (int64_t)n * P.Cin + ach, gt8);   // tail pixels: n >= N
#pragma unroll
            for (int t = 0; t < TAPS; ++t) {
              if (((okA >> (t * MAXPA + p)) & 1u) && a_xf) {
                float v[8];
                wg_unpack(ra[t][p], v);
                if (a_aff) {
#pragma unroll
                  for (int j = 0; j < 8; ++j) v[j] = fmaf(v[j], sc[j], sh[j]);
                }
                if (a_act != S2LC_ACT_NONE) {
#pragma unroll
                  for (int j = 0; j < 8; ++j) v[j] = act_fwd_fast(v[j], a_act);
                }
                if (a_gate) {
#pragma unroll
                  for (int j = 0; j < 8; ++j) v[j] *= gt8[j];
                }
                ra[t][p] = wg_pack(v);
              }
              *reinterpret_cast<uint4*>(dstA + t * K.asz + wg_off(k, cA)) = ra[t][p];
            }
          }
          if (a_gate) walk(n, h, w, paH, paW);
        }
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(&full[s]);
      if (g_pos) walk(gn0, gh0, gw0, blkH, blkW);
      if (a_pos) walk(an0, ah0, aw0, blkH, blkW);
    }

    // =============================================================== epilogue (same warps)
    if (nblk > 0) {
      mbar_wait(done, 0);
      tc_fence_after();
      const int q = warp & 3, half = warp >> 2;
      const int mrow = q * 32 + lane;
      const int nchunks = K.ntpad / 16;
      for (int t = 0; t < TAPS; ++t) {
        for (int cj = half; cj < nchunks; cj += WG_LOAD_WARPS / 4) {   // 4 warps share a TMEM lane quarter
          float acc[16];
          tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(t * K.ntpad + cj * 16), acc);
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int ncol = cj * 16 + j;
            const int gl = K.g_on_m ? mrow : ncol, al = K.g_on_m ? ncol : mrow;
            if (gl < g_cw && al < a_cw && a_c0 + al < P.Cin) {
              int64_t idx = wgrad_index(P, g_c0 + gl, t * P.Cin + a_c0 + al);
              atomicAdd(P.dw + idx, acc[j]);
            }
          }
        }
      }
      if (do_bias) {
        if (actG) {
#pragma unroll
          for (int j = 0; j < 8; ++j) atomicAdd(&sbias[cG * 8 + j], bsum[j]);
        }
        loaders_sync();
        for (int i = tid; i < g_cw; i += WG_LOAD_THREADS) atomicAdd(P.dbias + ((g_c0 + i) % P.Cout), sbias[i]);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == WG_LOAD_WARPS) tmem_dealloc(tmem_base, (uint32_t)K.tmem_cols);
}

static bool wg_aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
static int up(int x, int m) { return (x + m - 1) / m * m; }

int32_t launch_wgrad_tc(const WgradArgs& A, cudaStream_t st) {
  // shapes this engine takes; the rest (fp32 verification mode, the 4-class classifier, the stem) stay on CUDA cores
  if (A.dtype != S2LC_BF16) return 1;
  if (A.mode != S2LC_CONV_1X1 && A.mode != S2LC_CONV_3X3 && A.mode != S2LC_CONVT_FWD) return 1;
  const int Ng = A.ngroups * A.Cout, Ca = up(A.Cin, 8);
  if (A.Cout % 8 != 0 || A.a.ld < Ca || A.a.ld % 8 != 0 || A.g.ld_dy % 8 != 0) return 1;
  if (A.g.z && A.g.ld_z % 8 != 0) return 1;
  if (!wg_aligned16(A.a.raw) || !wg_aligned16(A.g.dy) || (A.g.z && !wg_aligned16(A.g.z))) return 1;
  if (A.a.scale && !A.a.shift) return 1;
  if (A.a.gate && A.Cin % 8 != 0) return 1;
  const int64_t M = (int64_t)A.N * A.H * A.W;
  if (M < WG_BK) return 1;

  WgPlan K{};
  if (A.mode == S2LC_CONV_3X3) {
    K.g_on_m = 1; K.taps = 9;
    K.n_ntiles = (Ca + 47) / 48;
    K.nt = up((Ca + K.n_ntiles - 1) / K.n_ntiles, 8);
  } else {
    K.taps = 1;
    // orientation with the cheaper tensor-pipe cost (M tiles are always 128 lanes wide)
    auto cost = [](int cm, int cn, int& ntiles, int& nt) {
      ntiles = (cn + 255) / 256;
      nt = up((cn + ntiles - 1) / ntiles, 8);
      return (int64_t)((cm + 127) / 128) * ntiles * up(nt, 16);
    };
    int nta, nt_a, ntb, nt_b;
    int64_t c1 = cost(Ng, Ca, nta, nt_a), c0 = cost(Ca, Ng, ntb, nt_b);
    if (c1 <= c0) { K.g_on_m = 1; K.n_ntiles = nta; K.nt = nt_a; }
    else { K.g_on_m = 0; K.n_ntiles = ntb; K.nt = nt_b; }
  }
  K.ntpad = up(K.nt, 16);
  const int n_side_bytes = (K.nt + 63) / 64 * WG_GROUP_BYTES;
  K.gsz = K.g_on_m ? 2 * WG_GROUP_BYTES : n_side_bytes;
  K.asz = K.g_on_m ? n_side_bytes : 2 * WG_GROUP_BYTES;
  const int budget = 208 * 1024;
  const int stage_bytes = K.gsz + K.taps * K.asz;
  K.stages = budget / stage_bytes;
  if (K.stages > 4) K.stages = 4;
  if (K.stages < 2) return 1;
  K.tmem_cols = 32;
  while (K.tmem_cols < K.taps * K.ntpad) K.tmem_cols *= 2;
  if (K.tmem_cols > 512) return 1;

  const int n_mtiles = ((K.g_on_m ? Ng : Ca) + 127) / 128;
  const int tiles = n_mtiles * K.n_ntiles;
  // split the pixels so that tiles * slabs fills a whole number of 148-SM waves as evenly as possible
  const int64_t nblocks = (M + WG_BK - 1) / WG_BK;
  int best_slabs = 1;
  double best_eff = -1.0;
  for (int w = 1; w <= 3; ++w) {
    int64_t s = (int64_t)kNumSMs * w / tiles;
    if (s < 1) s = 1;
    if (s > nblocks) s = nblocks;
    double eff = (double)(tiles * s) / (double)(((tiles * s + kNumSMs - 1) / kNumSMs) * kNumSMs);
    if (eff > best_eff + 0.05) { best_eff = eff; best_slabs = (int)s; }
  }
  K.px_per_slab = ((nblocks + best_slabs - 1) / best_slabs) * WG_BK;
  K.slabs = (int)((M + K.px_per_slab - 1) / K.px_per_slab);

  const size_t smem = 1024 + (size_t)K.stages * stage_bytes + (2 * WG_MAX_STAGES + 1) * sizeof(uint64_t) +
                      2 * sizeof(uint32_t) + 256 * sizeof(float) + 64;
  auto k9 = wgrad_tc_kernel<9, 2, 1>;   // 512 loader threads: 128-ch G tile = 2 passes, <=64-ch A tile = 1 pass
  auto k1 = wgrad_tc_kernel<1, 4, 4>;   // up to 256-ch tiles = 4 passes
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(k9, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k1, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) {
      set_error("wgrad_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
      return S2LC_ERR_CUDA;
    }
    attr_set = true;
  }
  dim3 grid((unsigned)K.slabs, (unsigned)tiles);
  if (K.taps == 9)
    k9<<<grid, WG_THREADS, smem, st>>>(A, K);
  else
    k1<<<grid, WG_THREADS, smem, st>>>(A, K);
  return check_launch("wgrad_tc");
}

}  // namespace s2lc

// swgrad_tc.cu — STREAMING tcgen05 weight-gradient engine (persistent, warp-specialised) for the shallow layers:
//
//   dW[co, (tap, ci)] = sum over pixels  G(px, co) * A(px shifted by tap, ci)
//
// (reference: autograd of nn.Conv2d(k=3, p=1) at src/modules/efficientnet_unet.py:168-176 and of the 1x1 expand /
//  project / head convs at :312-327, :363-375, :229-238.)
//
// The reduction runs over PIXELS, so both tensor-core operands are MN-major: a pixel is one 128-byte line of a
// SWIZZLE_128B tile and 64 channels are contiguous inside it — exactly the layout sconv_tc.cu uses for its K-major
// A operand, so the producers are the same: cp.async straight to the swizzled line, in-place lazy transform
//   A = act(raw*scale + shift) * gate          G = p*dy*rowscale + q*z + t   (z staged beside the ring)
// by the thread that issued the copy, fence.proxy.async, mbarrier.  One CTA per SM walks its share of the pixels and
// keeps ALL its accumulators in TMEM until the end (one fp32 atomicAdd pass per CTA into the reference-layout
// gradient):
//   3x3 : units are strips of 128 px x RU rows; the A rows roll through a ring (each row is read from HBM once for
//         its 3 output rows x 9 taps; a tap is a shifted start address: +dx lines, slot of row r+dy); G rows have
//         their own shorter ring.  M = co (<= 64, upper half of the 128 lanes reads a shared zero chunk),
//         N = the CTA's slice of <= 48 input channels, 9 accumulators of N columns.
//   1x1 : units are tiles of 128 / 64 / 32 pixels; the wider operand sits on M in up to 4 tiles of 128 channels,
//         the narrower one on N; one accumulator per M tile.
// Roles: warps 0-14 producers (the in-place transforms are issue-latency bound: 7 producer warps left the SM's four
// schedulers half empty and a ring slot took ~6k cycles whatever its length), warp 15 MMA issuer (converged warp,
// elected lane); when the stream ends producer warps 0-7 become the epilogue.
#include <cstdlib>

#include "conv_common.cuh"
#include "stream_common.cuh"

namespace s2lc {

using namespace tc;

constexpr int SW_PROD_WARPS = 15;
constexpr int SW_PROD_THREADS = SW_PROD_WARPS * 32;
constexpr int SW_MMA_WARP = SW_PROD_WARPS;
constexpr int SW_EPI_WARPS = 8;          // producer warps 0-7 drain TMEM once their stream has ended
constexpr int SW_THREADS = (SW_PROD_WARPS + 1) * 32;
constexpr int SW_MAX_RING = 8;

struct WPlan {
  int taps;            // 9 or 1
  int lines;           // pixels per G slot / per 1x1 tile (K of one slot)
  int a_lines;         // lines of an A slot that are loaded: 130 (3x3) or `lines`
  int CWa, CWg;        // 16 B chunks per pixel that are loaded for A (the CTA's slice) and for G
  int a_stride;        // bytes between the 64-channel chunks of an A slot
  int g_stride;        // same for G
  int a_slot, g_slot;  // bytes per slot
  int zslot;           // bytes per z staging slot (0: no BatchNorm behind G)
  int ringA, ringG, pf;
  int nsplit, nt;      // CTAs per unit (= asplit * gsplit) and the A slice width in channels
  int asplit, g_w;     // A slices; G slice width in channels (G slices = nsplit / asplit)
  int m_on_g;          // 1: G channels on M (TMEM lanes), A on N;  0: swapped
  int n_mtiles;
  int Npad;            // MMA N
  int nacc;            // accumulators: taps (3x3) or n_mtiles (1x1)
  int tmem_cols;
  int g_lbo;           // 3x3: 0 = G slot has 2 chunks; else G slot has 1 chunk and the upper lanes read the zero chunk
  int RU, HS, WS, units, ctas_per_slice;
  int children, lw;    // transposed conv: 4 children (2x2 output positions), each its own CTA set; lw = log2(W)
};

struct WUnit {
  int n, h0, rows, w0;
  int64_t m0;
  int nload, nout;
};
__device__ __forceinline__ WUnit wunit_geom(const WgradArgs& P, const WPlan& K, int u) {
  WUnit g;
  if (K.taps == 9) {
    const int ws = u % K.WS;
    const int t = u / K.WS;
    const int hs = t % K.HS;
    g.n = t / K.HS;
    g.h0 = hs * K.RU;
    g.rows = (P.H - g.h0) < K.RU ? (P.H - g.h0) : K.RU;
    g.w0 = ws * K.lines;
    g.m0 = 0;
    g.nload = g.rows + 2;
    g.nout = g.rows;
  } else {
    g.m0 = (int64_t)u * K.lines;
    g.n = (int)((uint32_t)g.m0 / (uint32_t)(P.H * P.W));      // M < 2^31 (checked by the host)
    g.h0 = 0; g.rows = 1; g.w0 = 0;
    g.nload = 1;
    g.nout = 1;
  }
  return g;
}

struct WCursor {
  int u, j;
  bool ok;
  WUnit ug;
};

template <int TAPS, bool CT>   // CT: 2x2 transposed conv (1x1 path with strided gradient rows, four children)
__global__ void __launch_bounds__(SW_THREADS, 1) swgrad_kernel(WgradArgs P, WPlan K) {
  cudaGridDependencySynchronize();   // programmatic dependent launch (common.cuh)
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* smem =
      reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  unsigned char* ringA = smem;                                     // [ringA][a_slot]
  unsigned char* ringG = ringA + K.ringA * K.a_slot;               // [ringG][g_slot]
  unsigned char* zchunk = ringG + K.ringG * K.g_slot;              // 16 KB of zeros (3x3 with narrow G), else empty
  unsigned char* zring = zchunk + (K.g_lbo ? 16384 : 0);           // [pf+1][zslot]
  float* sbias = reinterpret_cast<float*>(zring + (K.pf + 1) * K.zslot);   // [256]
  uint64_t* full = reinterpret_cast<uint64_t*>(sbias + 256);
  uint64_t* emptyA = full + SW_MAX_RING;
  uint64_t* emptyG = emptyA + SW_MAX_RING;
  uint64_t* done = emptyG + SW_MAX_RING;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done + 1);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int nsl = K.nsplit * K.children;
  const int sl = blockIdx.x % nsl, cta = blockIdx.x / nsl;
  const int child = sl / K.nsplit, slice = sl - child * K.nsplit;      // child = 2*qy + qx of a transposed conv
  const int qy = child >> 1, qx = child & 1;
  const int a_c0 = (slice % K.asplit) * K.nt;                      // first input channel of this CTA's A slice
  const int g_c0 = (slice / K.asplit) * K.g_w;                     // first output channel of its G slice
  constexpr int R = TAPS == 9 ? 3 : 1;
  const uint32_t ringA_u32 = smem_u32(ringA), ringG_u32 = smem_u32(ringG), z_u32 = smem_u32(zring);
  const int Ca = (P.Cin + 7) & ~7;
  int CWa = K.CWa;                                                  // chunks of this slice that exist in the buffer
  {
    const int left = (Ca - a_c0) >> 3;
    CWa = left < CWa ? left : CWa;
  }
  int CWg_ = K.CWg;
  {
    const int left = (P.Cout - g_c0) >> 3;
    CWg_ = left < CWg_ ? left : CWg_;
  }

  if (tid == 0) {
    for (int s = 0; s < K.ringA; ++s) { mbar_init(&full[s], SW_PROD_WARPS); mbar_init(&emptyA[s], 1); }
    for (int s = 0; s < K.ringG; ++s) mbar_init(&emptyG[s], 1);
    mbar_init(done, 1);
    fence_barrier_init();
  }
  if (warp == SW_MMA_WARP) tmem_alloc(tmem_slot, (uint32_t)K.tmem_cols);
  {
    const uint4 z4 = make_uint4(0, 0, 0, 0);
    const int nz = (K.ringA * K.a_slot + K.ringG * K.g_slot + (K.g_lbo ? 16384 : 0)) >> 4;
    for (int i = tid; i < nz; i += SW_THREADS) sts16(ringA_u32 + (uint32_t)i * 16u, z4);
    for (int i = tid; i < 256; i += SW_THREADS) sbias[i] = 0.f;
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  auto cur_init = [&](WCursor& c) {
    c.u = cta; c.j = 0;
    c.ok = c.u < K.units;
    if (c.ok) c.ug = wunit_geom(P, K, c.u);
  };
  auto cur_next = [&](WCursor& c) {
    if (++c.j == c.ug.nload) {
      c.u += K.ctas_per_slice; c.j = 0;
      c.ok = c.u < K.units;
      if (c.ok) c.ug = wunit_geom(P, K, c.u);
    }
  };

  if (warp < SW_PROD_WARPS) {
    // ================================================================== producers
    const bf16* a_ptr = reinterpret_cast<const bf16*>(P.a.raw);
    const int a_ld = (int)P.a.ld;
    const bf16* dy_ptr = reinterpret_cast<const bf16*>(P.g.dy);
    const bf16* z_ptr = reinterpret_cast<const bf16*>(P.g.z);     // same row stride as dy (checked by the host)
    const int g_ld = (int)P.g.ld_dy;
    const float* gate = P.a.gate;
    const float* rowscale = P.g.rowscale;
    const int act = P.a.act;
    const bool a_aff = P.a.scale != nullptr;
    const bool a_xf = a_aff || gate != nullptr || act != S2LC_ACT_NONE;
    const bool do_bias = P.dbias != nullptr && (slice % K.asplit) == 0;
    constexpr bool is_ct = CT;
    const int HWi = P.H * P.W;
    const bool g_xf = z_ptr != nullptr || rowscale != nullptr || do_bias;
    // fixed chunk per thread, separately for the A part and the G part of a load
    const int dla = SW_PROD_THREADS / CWa, nact_a = dla * CWa;
    const int la0 = tid / CWa, cha = tid - la0 * CWa;
    const bool act_a = tid < nact_a;
    const int nit_a = act_a ? (K.a_lines - la0 + dla - 1) / dla : 0;
    const int CWg = CWg_, dlg = SW_PROD_THREADS / CWg, nact_g = dlg * CWg;
    const int lg0 = tid / CWg, chg = tid - lg0 * CWg;
    const bool act_g = tid < nact_g;
    const int nit_g = act_g ? (K.lines - lg0 + dlg - 1) / dlg : 0;
    const uint32_t dst_ca = (uint32_t)((cha >> 3) * K.a_stride), cha7 = (uint32_t)(cha & 7);
    const uint32_t dst_cg = (uint32_t)((chg >> 3) * K.g_stride), chg7 = (uint32_t)(chg & 7);
    float sa[8], sb[8], gp[8], gq[8], gt[8], bsum[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) { sa[k] = 1.f; sb[k] = 0.f; gp[k] = 1.f; gq[k] = 0.f; gt[k] = 0.f; bsum[k] = 0.f; }
    if (act_a && a_aff) { sc_ldf8(P.a.scale + a_c0 + cha * 8, sa); sc_ldf8(P.a.shift + a_c0 + cha * 8, sb); }
    if (act_g && z_ptr) {
      sc_ldf8(P.g.p + g_c0 + chg * 8, gp); sc_ldf8(P.g.q + g_c0 + chg * 8, gq); sc_ldf8(P.g.t + g_c0 + chg * 8, gt);
    }
    const int zslots = K.pf + 1;
    int ia = 0, ig = 0, iz = 0, fa = 0, fg = 0, fz = 0;     // ring positions: issue / finish cursors
    uint32_t pha = 0, phg = 0;
    bool wrapA = false, wrapG = false;

    // geometry of one load: A row0 + valid A line range, G row0 + validity
    auto geom = [&](const WCursor& c, int64_t& arow0, int& lo, int& hi, int64_t& grow0, bool& gok) {
      if (TAPS == 9) {
        const int hr = c.ug.h0 - 1 + c.j;
        arow0 = ((int64_t)c.ug.n * P.H + hr) * P.W + c.ug.w0 - 1;
        lo = c.ug.w0 == 0 ? 1 : 0;
        hi = c.ug.w0 + K.lines == P.W ? K.lines : K.lines + 1;
        if (hr < 0 || hr >= P.H) { lo = 1; hi = 0; }
        grow0 = arow0 + 1;                                    // G row = the output row h0 + j - 1
        gok = c.j >= 1 && c.j <= c.ug.rows;
      } else {
        arow0 = c.ug.m0; lo = 0; hi = K.lines - 1;
        grow0 = c.ug.m0; gok = true;
      }
    };
    auto issue = [&](const WCursor& c) {
      int64_t arow0, grow0; int lo, hi; bool gok;
      geom(c, arow0, lo, hi, grow0, gok);
      const uint32_t sA = ringA_u32 + (uint32_t)(ia * K.a_slot) + dst_ca;
      const bf16* abase = a_ptr + arow0 * a_ld + a_c0 + cha * 8;
      int line = la0;
#pragma unroll 2
      for (int i = 0; i < nit_a; ++i, line += dla) {
        const uint32_t dst = sA + (uint32_t)(line * 128) + ((cha7 ^ (uint32_t)(line & 7)) << 4);
        if (line >= lo && line <= hi) cp_async16(dst, abase + line * a_ld);
        else sts16(dst, make_uint4(0, 0, 0, 0));
      }
      if (gok) {
        const uint32_t sG = ringG_u32 + (uint32_t)(ig * K.g_slot) + dst_cg;
        const uint32_t zG = z_u32 + (uint32_t)(iz * K.zslot) + (uint32_t)tid * 16u;
        // transposed conv: input pixel (n, h, w) pairs with the gradient row of its child (n, 2h+qy, 2w+qx)
        int pi0 = 0;
        if (is_ct) {
          pi0 = (int)(c.ug.m0 - (int64_t)c.ug.n * HWi);
          grow0 = (int64_t)c.ug.n * 4 * HWi;
        }
        const bf16* dbase = dy_ptr + grow0 * g_ld + g_c0 + chg * 8;
        const bf16* zbase = z_ptr ? z_ptr + grow0 * g_ld + g_c0 + chg * 8 : nullptr;
        line = lg0;
#pragma unroll 2
        for (int i = 0; i < nit_g; ++i, line += dlg) {
          const uint32_t dst = sG + (uint32_t)(line * 128) + ((chg7 ^ (uint32_t)(line & 7)) << 4);
          int roff = line;
          if (is_ct) {
            const int pi = pi0 + line;
            roff = (2 * (pi >> K.lw) + qy) * (2 * P.W) + 2 * (pi & (P.W - 1)) + qx;
          }
          cp_async16(dst, dbase + roff * g_ld);
          if (zbase) cp_async16(zG + (uint32_t)(i * nact_g) * 16u, zbase + roff * g_ld);
        }
      }
      cp_async_commit();
    };
    auto finish = [&](const WCursor& c) {
      int64_t arow0, grow0; int lo, hi; bool gok;
      geom(c, arow0, lo, hi, grow0, gok);
      if (a_xf && act_a) {
        const uint32_t sA = ringA_u32 + (uint32_t)(fa * K.a_slot) + dst_ca;
        float g8[8];
        if (gate) sc_ldf8(gate + (int64_t)c.ug.n * P.Cin + a_c0 + cha * 8, g8);
        int line = la0;
#pragma unroll 2
        for (int i = 0; i < nit_a; ++i, line += dla) {
          if (line >= lo && line <= hi) {
            const uint32_t dst = sA + (uint32_t)(line * 128) + ((cha7 ^ (uint32_t)(line & 7)) << 4);
            float v[8];
            sc_unpack8(lds16(dst), v);
            if (a_aff) {
#pragma unroll
              for (int k = 0; k < 8; ++k) v[k] = fmaf(v[k], sa[k], sb[k]);
            }
            if (act == S2LC_ACT_RELU) {
#pragma unroll
              for (int k = 0; k < 8; ++k) v[k] = fmaxf(v[k], 0.f);
            } else if (act == S2LC_ACT_SILU) {
#pragma unroll
              for (int k = 0; k < 8; ++k) v[k] = silu_fast(v[k]);
            }
            if (gate) {
#pragma unroll
              for (int k = 0; k < 8; ++k) v[k] *= g8[k];
            }
            sts16(dst, sc_pack8(v));
          }
        }
      }
      if (g_xf && gok && act_g) {
        const uint32_t sG = ringG_u32 + (uint32_t)(fg * K.g_slot) + dst_cg;
        const uint32_t zG = z_u32 + (uint32_t)(fz * K.zslot) + (uint32_t)tid * 16u;
        const float rs = rowscale ? __ldg(rowscale + c.ug.n) : 1.f;
        int line = lg0;
#pragma unroll 2
        for (int i = 0; i < nit_g; ++i, line += dlg) {
          const uint32_t dst = sG + (uint32_t)(line * 128) + ((chg7 ^ (uint32_t)(line & 7)) << 4);
          float v[8];
          sc_unpack8(lds16(dst), v);
          if (rowscale) {
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] *= rs;
          }
          if (z_ptr) {
            float z[8];
            sc_unpack8(lds16(zG + (uint32_t)(i * nact_g) * 16u), z);
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = fmaf(gp[k], v[k], fmaf(gq[k], z[k], gt[k]));
          }
          if (do_bias) {
#pragma unroll
            for (int k = 0; k < 8; ++k) bsum[k] += v[k];
          }
          sts16(dst, sc_pack8(v));
        }
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) sc_mbar_arrive(&full[fa]);
    };

    WCursor ci, cf;
    cur_init(ci);
    cur_init(cf);
    int inflight = 0;
    while (cf.ok) {
      while (ci.ok && inflight <= K.pf) {
        const bool gok_i = TAPS == 9 ? (ci.j >= 1 && ci.j <= ci.ug.rows) : true;   // this load carries a G row
        if (wrapA) mbar_wait(&emptyA[ia], pha);
        if (gok_i && wrapG) mbar_wait(&emptyG[ig], phg);
        issue(ci);
        ++inflight;
        cur_next(ci);
        if (++ia == K.ringA) { ia = 0; if (wrapA) pha ^= 1u; wrapA = true; }
        if (gok_i && ++ig == K.ringG) { ig = 0; if (wrapG) phg ^= 1u; wrapG = true; }
        if (++iz == zslots) iz = 0;
      }
      switch (inflight - 1) {
        case 0: cp_async_wait<0>(); break;
        case 1: cp_async_wait<1>(); break;
        case 2: cp_async_wait<2>(); break;
        default: cp_async_wait<3>(); break;
      }
      const bool gok_f = TAPS == 9 ? (cf.j >= 1 && cf.j <= cf.ug.rows) : true;
      finish(cf);
      --inflight;
      cur_next(cf);
      if (++fa == K.ringA) fa = 0;
      if (gok_f && ++fg == K.ringG) fg = 0;
      if (++fz == zslots) fz = 0;
    }
    if (do_bias) {   // block-level reduction of the bias gradient: smem atomics, then one global atomic per channel
      if (act_g) {
#pragma unroll
        for (int k = 0; k < 8; ++k) atomicAdd(&sbias[chg * 8 + k], bsum[k]);
      }
      asm volatile("bar.sync 1, %0;" ::"n"(SW_PROD_THREADS) : "memory");
      for (int i = tid; i < CWg * 8; i += SW_PROD_THREADS) atomicAdd(P.dbias + g_c0 + i, sbias[i]);
    }
  } else if (warp == SW_MMA_WARP) {
    // ================================================================== MMA issuer (converged warp, elected lane)
    {
      const bool leader = elect_one();
      const uint32_t idesc = idesc_bf16(128, (uint32_t)K.Npad, 1, 1);     // both operands MN-major
      const uint64_t dA = smem_desc_sw128(0, 1024, (uint32_t)K.a_stride);
      const uint64_t dG = smem_desc_sw128(0, 1024, (uint32_t)K.g_stride);
      const uint32_t zc16 = smem_u32(zchunk) >> 4;
      const uint32_t a_lo = ringA_u32 >> 4, g_lo = ringG_u32 >> 4;
      const uint32_t aslot16 = (uint32_t)K.a_slot >> 4, gslot16 = (uint32_t)K.g_slot >> 4;
      const int nks = K.lines >> 4;                                       // k-steps (16 pixels) per slot
      int fs = 0;  uint32_t fph = 0;
      int sA_first = 0;                     // ring slot of load 0 of the current unit
      int g0 = 0;                           // G ring slot of the current output (one G row per output)
      uint32_t any = 0;                     // 0 until the first MMA of this CTA has been issued into every accumulator
      for (int u = cta; u < K.units; u += K.ctas_per_slice) {
        const WUnit ug = wunit_geom(P, K, u);
        int waited = 0;
        int r0 = sA_first;
        for (int o = 0; o < ug.nout; ++o) {
          while (waited < o + R) {
            mbar_wait(&full[fs], fph);
            if (++fs == K.ringA) { fs = 0; fph ^= 1u; }
            ++waited;
          }
          tc_fence_after();
          const uint32_t g_addr = g_lo + (uint32_t)g0 * gslot16;
          if (TAPS == 9) {
            // M operand = G row (co on the TMEM lanes); upper 64 lanes read chunk 1 of the slot or the shared zero chunk
            uint64_t dM = dG + (uint64_t)g_addr;
            if (K.g_lbo) dM = smem_desc_sw128(0, 1024, (zc16 - g_addr) << 4) + (uint64_t)g_addr;
            int rs = r0;
#pragma unroll
            for (int dy1 = 0; dy1 < 3; ++dy1) {
              const uint32_t a_row = a_lo + (uint32_t)rs * aslot16;
#pragma unroll
              for (int dx1 = 0; dx1 < 3; ++dx1) {
                const uint64_t dN = dA + (uint64_t)(a_row + (uint32_t)(dx1 * 8));
                const uint32_t d_tmem = tmem_base + (uint32_t)((dy1 * 3 + dx1) * K.Npad);
#pragma unroll 2
                for (int ks = 0; ks < nks; ++ks) {
                  if (leader) umma_bf16(d_tmem, dM + (uint64_t)(ks * 128), dN + (uint64_t)(ks * 128), idesc, (any | (uint32_t)ks) ? 1u : 0u);
                }
              }
              if (++rs == K.ringA) rs = 0;
            }
          } else {
            const uint32_t a_addr = a_lo + (uint32_t)r0 * aslot16;
            for (int mt = 0; mt < K.n_mtiles; ++mt) {
              // 128 channels of the M-side operand = 2 chunks of its slot
              const uint64_t dM = K.m_on_g ? dG + (uint64_t)(g_addr + (uint32_t)(2 * mt) * ((uint32_t)K.g_stride >> 4))
                                           : dA + (uint64_t)(a_addr + (uint32_t)(2 * mt) * ((uint32_t)K.a_stride >> 4));
              const uint64_t dN = K.m_on_g ? dA + (uint64_t)a_addr : dG + (uint64_t)g_addr;
              const uint32_t d_tmem = tmem_base + (uint32_t)(mt * K.Npad);
#pragma unroll 2
              for (int ks = 0; ks < nks; ++ks) {
                if (leader) umma_bf16(d_tmem, dM + (uint64_t)(ks * 128), dN + (uint64_t)(ks * 128), idesc, (any | (uint32_t)ks) ? 1u : 0u);
              }
            }
          }
          any = 1;
          if (leader) {
            umma_commit(&emptyA[r0]);
            umma_commit(&emptyG[g0]);
            if (TAPS == 9 && o == ug.nout - 1) {
              int r1 = r0 + 1 == K.ringA ? 0 : r0 + 1;
              int r2 = r1 + 1 == K.ringA ? 0 : r1 + 1;
              umma_commit(&emptyA[r1]);
              umma_commit(&emptyA[r2]);
            }
          }
          __syncwarp();
          if (++r0 == K.ringA) r0 = 0;
          if (++g0 == K.ringG) g0 = 0;
        }
        sA_first = (sA_first + ug.nload) % K.ringA;
      }
      if (leader) umma_commit(done);
      __syncwarp();
    }
  }
  if (warp < SW_EPI_WARPS) {
    // ================================================================== epilogue: once, after the last MMA
    const int half = warp >> 2, q = warp & 3;
    const int ct_base = CT ? child * P.Cout : 0;   // convT: gradient row = (child, co)
    mbar_wait(done, 0);
    tc_fence_after();
    const int mrow = q * 32 + lane;
    const int nchunks = K.Npad / 16;
    for (int acc = 0; acc < K.nacc; ++acc) {
      const int tap = TAPS == 9 ? acc : 0, mt = TAPS == 9 ? 0 : acc;
      const int mch = mt * 128 + mrow;                       // channel index on the M side
      for (int cj = half; cj < nchunks; cj += 2) {
        float v[16];
        tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * K.Npad + cj * 16), v);
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const int nch = cj * 16 + j;
          const int gl = K.m_on_g ? mch : nch, al = K.m_on_g ? nch : mch;   // channel inside the G / A slice
          const int co = gl + g_c0, ci = al + a_c0;
          // rows / columns beyond the slice widths belong to another CTA (and are zero here)
          if (gl < K.g_w && al < K.nt && co < P.Cout && ci < P.Cin)
            atomicAdd(P.dw + wgrad_index(P, ct_base + co, tap * P.Cin + ci), v[j]);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == SW_MMA_WARP) tmem_dealloc(tmem_base, (uint32_t)K.tmem_cols);
}

static bool sw_aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
static int sw_up(int x, int m) { return (x + m - 1) / m * m; }

// Returns 0 if launched, <0 on error, 1 if this shape is left to the slab engine (wgrad_tc.cu).
int32_t launch_swgrad_tc(const WgradArgs& A, cudaStream_t st) {
  if (debug_flags() & DBG_NO_STREAM) return 1;   // test hook: compare against the slab engine
  if (A.dtype != S2LC_BF16 || (A.mode != S2LC_CONV_3X3 && A.mode != S2LC_CONV_1X1 && A.mode != S2LC_CONVT_FWD)) return 1;
  const int children = A.mode == S2LC_CONVT_FWD ? 4 : 1;
  int lw = 0;
  while ((1 << lw) < A.W) ++lw;
  if (children == 4 && ((1 << lw) != A.W || A.a.gate || A.g.rowscale)) return 1;   // row arithmetic by shifts
  if ((int64_t)A.N * A.H * A.W * children >= (1ll << 31)) return 1;
  const int Ca = sw_up(A.Cin, 8);
  if (A.Cout % 8 != 0 || A.a.ld < Ca || A.a.ld % 8 != 0 || A.g.ld_dy % 8 != 0) return 1;
  if (!sw_aligned16(A.a.raw) || !sw_aligned16(A.g.dy)) return 1;
  if (A.g.z && (A.g.ld_z != A.g.ld_dy || !sw_aligned16(A.g.z))) return 1;
  if (A.a.scale && (!A.a.shift || A.Cin % 8 != 0)) return 1;
  if (A.a.gate && A.Cin % 8 != 0) return 1;
  const int64_t HW = (int64_t)A.H * A.W, M = HW * A.N;
  const int budget = 226 * 1024;
  const int fixed = 1024 + 256 * 4 + (3 * SW_MAX_RING + 1) * 8 + 64;
  WPlan K{};
  K.taps = A.mode == S2LC_CONV_3X3 ? 9 : 1;
  K.CWg = A.Cout / 8;
  int ctas = kNumSMs;
  if (K.taps == 9) {
    if (A.W % 32 != 0 || (A.Cout > 128 && A.Cout % 128 != 0)) return 1;
    // G (output channels) sits on the 128 TMEM lanes: layers wider than 128 are cut into G slices of 128 channels,
    // A (input channels) into slices of <= 48 (nine accumulators share the 512 TMEM columns).  One CTA = one
    // (A slice, G slice) pair over its share of the pixel strips.
    const int gsplit = (A.Cout + 127) / 128;
    K.asplit = (Ca + 47) / 48;
    K.nsplit = K.asplit * gsplit;
    if (K.nsplit > kNumSMs) return 1;
    K.g_w = A.Cout < 128 ? A.Cout : 128;
    K.CWg = K.g_w / 8;
    K.nt = sw_up((Ca + K.asplit - 1) / K.asplit, 8);
    K.Npad = sw_up(K.nt, 16);
    if (9 * K.Npad > 512) return 1;
    K.CWa = K.nt / 8;
    K.g_lbo = A.Cout <= 64 ? 1 : 0;                                 // narrow G: upper 64 lanes read the zero chunk
    K.m_on_g = 1; K.n_mtiles = 1; K.nacc = 9;
    // strip width = pixels per ring slot: the widest of 128 / 64 / 32 that divides W and whose rings fit shared memory
    bool ok = false;
    int Lmax = 128;
    if ((debug_flags() >> 8) & 0xffu) Lmax = (int)((debug_flags() >> 8) & 0xffu);   // profiling hook: strip-width cap
    for (int L = Lmax; L >= 32 && !ok; L >>= 1) {
      if (A.W % L != 0) continue;
      K.lines = L; K.a_lines = L + 2;
      K.a_slot = sw_up((L + 2) * 128, 1024);                        // one 64-channel chunk (nt <= 48)
      K.a_stride = K.a_slot;
      K.g_stride = L * 128;
      K.g_slot = K.g_lbo ? L * 128 : 2 * L * 128;
      K.zslot = A.g.z ? L * K.CWg * 16 : 0;
      for (int slack = 1; slack >= 0 && !ok; --slack) {   // a slot of slack matters more than a deeper prefetch
        for (int pf = (L < 128 ? 3 : 2); pf >= 1 && !ok; --pf) {
          const int ra = pf + 3 + slack, rg = pf + 2 + slack;
          if (ra > SW_MAX_RING || rg > SW_MAX_RING) continue;
          if (fixed + ra * K.a_slot + rg * K.g_slot + (K.g_lbo ? 16384 : 0) + (pf + 1) * K.zslot <= budget) {
            K.pf = pf; K.ringA = ra; K.ringG = rg; ok = true;
          }
        }
      }
    }
    if (!ok) return 1;
    ctas = kNumSMs / K.nsplit;
    K.WS = A.W / K.lines;
    int64_t best_cost = -1;
    for (int ru = 8; ru <= 128; ru *= 2) {
      const int r = ru < A.H ? ru : A.H;
      const int hs = (A.H + r - 1) / r;
      const int64_t units = (int64_t)A.N * K.WS * hs;
      const int c = units < ctas ? (int)units : ctas;
      const int64_t cost = ((units + c - 1) / c) * (r + 2);
      if (best_cost < 0 || cost < best_cost) { best_cost = cost; K.RU = r; K.HS = hs; K.units = (int)units; }
      if (ru >= A.H) break;
    }
  } else {
    // the wider operand goes on M (TMEM lanes) in tiles of 128 channels, the narrower one on N.  Deep layers do not
    // fit one CTA's TMEM (n_mtiles * Npad <= 512 columns): the [Cm x Cn] gradient is cut into msplit x nsplit_n
    // blocks, each block x pixel-range is one CTA.  Cost model: bytes one CTA streams.
    K.m_on_g = A.Cout >= Ca ? 1 : 0;
    const int cm = K.m_on_g ? A.Cout : Ca, cn = K.m_on_g ? Ca : A.Cout;
    const int zm = (K.m_on_g && A.g.z) ? 2 : 1, zn = (!K.m_on_g && A.g.z) ? 2 : 1;   // G is read twice (dy and z)
    bool ok = false;
    double best = 0.0;
    for (int ms = 1; ms <= 16; ++ms) {
      const int mcta = sw_up((cm + ms - 1) / ms, 128);
      if ((cm + mcta - 1) / mcta != ms) continue;
      const int n_mt = mcta / 128;
      if (n_mt > 4) continue;
      for (int ns = 1; ns <= 8; ++ns) {
        const int ncta = sw_up((cn + ns - 1) / ns, 8);
        if ((cn + ncta - 1) / ncta != ns) continue;
        const int npad = sw_up(ncta, 16);
        if (npad > 256 || n_mt * npad > 512 || ms * ns * children > kNumSMs) continue;
        if (A.dbias && (K.m_on_g ? mcta : ncta) > 256) continue;
        const int cw_m = (mcta < cm ? mcta : sw_up(cm, 8)) / 8, cw_n = ncta / 8;
        if (cw_m > SW_PROD_THREADS || cw_n > SW_PROD_THREADS) continue;
        const int chunks_m = 2 * n_mt, chunks_n = (npad + 63) / 64;
        const int cwg = K.m_on_g ? cw_m : cw_n;
        const int kslabs = kNumSMs / (ms * ns * children);
        for (int lines = 128; lines >= 32; lines /= 2) {
          if (HW % lines != 0 || M / lines < 2 * kslabs) continue;
          const int slot = (chunks_m + chunks_n) * lines * 128;
          const int zslot = A.g.z ? lines * cwg * 16 : 0;
          int pf = 0;
          for (int q = 3; q >= 1 && !pf; --q)
            if (fixed + (q + 2) * slot + (q + 1) * zslot <= budget) pf = q;
          if (!pf) continue;
          // bytes per CTA (relative) + a penalty for short slots (per-slot overhead)
          const int mw = mcta < cm ? mcta : cm, nw = ncta < cn ? ncta : cn;   // widest slice actually streamed
          const double cost = (double)(mw * zm + nw * zn) / kslabs * (1.0 + 16.0 / lines);
          if (!ok || cost < best) {
            ok = true; best = cost;
            K.nsplit = ms * ns; K.n_mtiles = n_mt; K.Npad = npad; K.nacc = n_mt;
            K.asplit = K.m_on_g ? ns : ms;
            K.nt = K.m_on_g ? ncta : mcta; K.g_w = K.m_on_g ? mcta : ncta;
            K.CWa = K.m_on_g ? cw_n : cw_m; K.CWg = K.m_on_g ? cw_m : cw_n;
            K.lines = lines; K.a_lines = lines; K.a_stride = lines * 128; K.g_stride = lines * 128;
            K.a_slot = (K.m_on_g ? chunks_n : chunks_m) * lines * 128;
            K.g_slot = (K.m_on_g ? chunks_m : chunks_n) * lines * 128;
            K.zslot = zslot; K.pf = pf; K.ringA = pf + 2; K.ringG = pf + 2;
          }
          break;   // the largest feasible slot for this (ms, ns)
        }
      }
    }
    if (!ok) return 1;
    K.g_lbo = 0;
    K.RU = 1; K.HS = 1; K.WS = 1;
    K.units = (int)(M / K.lines);
    ctas = kNumSMs / (K.nsplit * children);
  }
  K.children = children; K.lw = lw;
  K.tmem_cols = 32;
  while (K.tmem_cols < K.nacc * K.Npad) K.tmem_cols *= 2;
  if (K.units < ctas) ctas = K.units;
  K.ctas_per_slice = ctas;
  size_t smem = (size_t)fixed + (size_t)K.ringA * K.a_slot + (size_t)K.ringG * K.g_slot +
                (K.g_lbo ? 16384 : 0) + (size_t)(K.pf + 1) * K.zslot;
  if (smem < 120 * 1024) smem = 120 * 1024;   // one CTA per SM always: each CTA may hold all 512 TMEM columns
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(swgrad_kernel<9, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(swgrad_kernel<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(swgrad_kernel<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) {
      set_error("swgrad_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
      return S2LC_ERR_CUDA;
    }
    attr_set = true;
  }
  if (K.taps == 9) PDL_LAUNCH((swgrad_kernel<9, false>), dim3(ctas * K.nsplit), SW_THREADS, smem, st, A, K);
  else if (children == 4) PDL_LAUNCH((swgrad_kernel<1, true>), dim3(ctas * K.nsplit * K.children), SW_THREADS, smem, st, A, K);
  else PDL_LAUNCH((swgrad_kernel<1, false>), dim3(ctas * K.nsplit), SW_THREADS, smem, st, A, K);
  return check_launch("swgrad_tc");
}

}  // namespace s2lc

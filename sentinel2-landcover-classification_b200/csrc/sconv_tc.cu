// sconv_tc.cu — STREAMING tcgen05 convolution: persistent, warp-specialised, for the shallow layers whose cost is
// HBM traffic, not flops (decoder 3x3 convs at >= 128 px wide rows, MBConv 1x1 expand/project convs with K <= 512).
//
// (reference call sites: src/modules/efficientnet_unet.py:168-176 `_double_conv`, :312-327 expand, :363-375 project,
//  and the autograd data gradients of both.)
//
// One CTA per SM walks a static list of work units.  A unit is
//   3x3 : a strip of 128 pixels x RU image rows of one sample.  Input rows ROLL through a shared-memory ring: every
//         input row segment (128 + 2 halo pixels, all channels) is fetched from HBM once and serves the three output
//         rows and nine taps that touch it.  A tap is just a shifted START ADDRESS of the same ring slot: the slot keeps
//         pixel p on 128-byte line p of a SWIZZLE_128B K-major tile, so (dy, dx) = (slot of row r+dy, line offset dx+1)
//         (the hardware swizzle is a function of absolute smem address bits, so no descriptor base offset is needed).
//   1x1 : one tile of 128 consecutive pixels per ring slot.
// Roles (544 threads):
//   warps 0-7   producers: cp.async (16 B, straight to the swizzled destination — no staging copy, no registers),
//               prefetching `pf` slots ahead; when a slot has landed each thread transforms ITS OWN chunks in place
//               (BatchNorm scale/shift + SiLU/ReLU + SE gate, or the BatchNorm-backward affine p*dy + q*z + t with z
//               staged beside it), fence.proxy.async, arrive on the slot's `full` mbarrier;
//   warp 8      MMA issuer: one lane issues tcgen05.mma for all taps / K chunks of an output tile into one of `nacc`
//               TMEM accumulators; tcgen05.commit frees the oldest ring slot and publishes the accumulator;
//   warps 9-16  two epilogue groups (alternate tiles): tcgen05.ld -> bias, residuals, act'(prev), bf16 store, BatchNorm
//               moments (butterfly column sums -> smem -> fp64 global atomics once per unit).
// Packed weights of the CTA's N slice stay resident in shared memory for the whole kernel.
#include <cstdlib>

#include "conv_common.cuh"
#include "tc_common.cuh"

namespace s2lc {

using namespace tc;

constexpr int SC_PROD_WARPS = 8;
constexpr int SC_PROD_THREADS = SC_PROD_WARPS * 32;
constexpr int SC_MMA_WARP = SC_PROD_WARPS;
constexpr int SC_EPI_WARP0 = SC_PROD_WARPS + 1;
constexpr int SC_EPI_WARPS = 8;
constexpr int SC_THREADS = (SC_PROD_WARPS + 1 + SC_EPI_WARPS) * 32;
constexpr int SC_MAX_RING = 8;
constexpr int SC_MAX_ACC = 4;
constexpr int SC_MAX_ITEMS = 10;   // 16 B chunks per producer thread per slot: ceil(130 * 16 / 256) = 9

struct SPlan {
  int taps;          // 9 or 1
  int Kc;            // 64-channel K chunks of the A operand
  int CW;            // 16 B chunks per pixel (C / 8)
  int lines;         // 130 (3x3) or 128 (1x1) lines per slot chunk
  int chunk_stride;  // bytes between K chunks of a slot (1024-aligned)
  int slot_bytes;    // Kc * chunk_stride
  int zslot_bytes;   // linear staging of z (BatchNorm-backward operand), 0 if unused
  int ring, pf;      // ring slots, prefetch distance (pf < ring - (3x3 ? 2 : 0))
  int nsplit;        // CTAs sharing a unit along N
  int Ncta;          // output channels per CTA (multiple of 16)
  int nacc;          // TMEM accumulators
  int tmem_cols;
  int w_bytes;       // resident weight bytes
  int RU;            // 3x3: output rows per unit
  int HS;            // 3x3: units per strip (ceil(H / RU))
  int WS;            // 3x3: strips per image row (W / 128)
  int units;         // total units (per N slice)
  int ctas_per_slice;
};

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void sc_mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void sts16(uint32_t addr, const uint4& v) {
  asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 lds16(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void sc_unpack8(const uint4& u, float (&v)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 f = __bfloat1622float2(h[i]);
    v[2 * i] = f.x;
    v[2 * i + 1] = f.y;
  }
}
__device__ __forceinline__ uint4 sc_pack8(const float (&v)[8]) {
  uint4 u;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
  return u;
}
__device__ __forceinline__ void sc_ldf8(const float* p, float (&v)[8]) {
  float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p + 4));
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
// SWIZZLE_128B K-major descriptor whose start may sit on any 128 B line of a 1024 B-aligned tile
// MEASURED on B200: the 128B swizzle is a function of the absolute shared-memory address bits, so a start that is
// shifted by whole 128 B lines needs NO matrix-base-offset field (setting it to the start line's phase gives garbage;
// tools/diag_sconv.py is the experiment).
__device__ __forceinline__ uint64_t sc_desc(uint32_t saddr) { return smem_desc_sw128(saddr, 1024, 16); }
// 16 per-lane column values -> column sums over the warp (see igemm_tc.cu)
__device__ __forceinline__ float sc_butterfly16(const float (&a)[16], int lane, int& col) {
  float b8[8], b4[4], b2[2];
  const bool h16 = lane & 16, h8 = lane & 8, h4 = lane & 4, h2 = lane & 2;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    float send = h16 ? a[i] : a[i + 8], keep = h16 ? a[i + 8] : a[i];
    b8[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float send = h8 ? b8[i] : b8[i + 4], keep = h8 ? b8[i + 4] : b8[i];
    b4[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
  }
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    float send = h4 ? b4[i] : b4[i + 2], keep = h4 ? b4[i + 2] : b4[i];
    b2[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
  }
  float send = h2 ? b2[0] : b2[1], keep = h2 ? b2[1] : b2[0];
  float r = keep + __shfl_xor_sync(0xffffffffu, send, 2);
  r += __shfl_xor_sync(0xffffffffu, r, 1);
  col = (h16 ? 8 : 0) + (h8 ? 4 : 0) + (h4 ? 2 : 0) + (h2 ? 1 : 0);
  return r;
}

// ---- the static work sequence every role walks identically -------------------------------------------------------
// A unit contributes `nload` ring loads and `nout` output tiles; output o of a unit consumes loads o .. o+R-1.
struct UnitGeom {
  int n, h0, rows, w0;   // 3x3: sample, first output row, output rows, first pixel column
  int64_t m0;            // 1x1: first pixel of the tile
  int nload, nout;
};
__device__ __forceinline__ UnitGeom unit_geom(const ConvArgs& P, const SPlan& K, int u) {
  UnitGeom g;
  if (K.taps == 9) {
    const int ws = u % K.WS;
    const int t = u / K.WS;
    const int hs = t % K.HS;
    g.n = t / K.HS;
    g.h0 = hs * K.RU;
    g.rows = (P.H - g.h0) < K.RU ? (P.H - g.h0) : K.RU;
    g.w0 = ws * 128;
    g.m0 = 0;
    g.nload = g.rows + 2;
    g.nout = g.rows;
  } else {
    g.m0 = (int64_t)u * 128;
    g.n = (int)(g.m0 / ((int64_t)P.H * P.W));
    g.h0 = 0; g.rows = 1; g.w0 = 0;
    g.nload = 1;
    g.nout = 1;
  }
  return g;
}

struct SCursor {
  int u, j, g;
  bool ok;
  UnitGeom ug;
};

__global__ void __launch_bounds__(SC_THREADS, 1) sconv_kernel(ConvArgs P, SPlan K) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* smem =
      reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  unsigned char* wsm = smem;                                   // resident weights
  unsigned char* ring = wsm + K.w_bytes;                       // [ring][slot_bytes]
  unsigned char* zring = ring + K.ring * K.slot_bytes;         // [pf+1][zslot_bytes]
  float* svec = reinterpret_cast<float*>(zring + (K.pf + 1) * K.zslot_bytes);   // [3][Kc*64] operand affine vectors
  float* sstat = svec + 3 * K.Kc * 64;                         // [2 groups][2][256]
  uint64_t* full = reinterpret_cast<uint64_t*>(sstat + 2 * 2 * 256);
  uint64_t* empty = full + SC_MAX_RING;
  uint64_t* acc_full = empty + SC_MAX_RING;
  uint64_t* acc_empty = acc_full + SC_MAX_ACC;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + SC_MAX_ACC);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int slice = blockIdx.x % K.nsplit, cta = blockIdx.x / K.nsplit;
  const int n0 = slice * K.Ncta;
  const int R = K.taps == 9 ? 3 : 1;
  const uint32_t ring_u32 = smem_u32(ring), w_u32 = smem_u32(wsm);

  // ------------------------------------------------------------------ one-time setup (all threads)
  if (tid == 0) {
    for (int s = 0; s < K.ring; ++s) { mbar_init(&full[s], SC_PROD_WARPS); mbar_init(&empty[s], 1); }
    for (int b = 0; b < K.nacc; ++b) { mbar_init(&acc_full[b], 1); mbar_init(&acc_empty[b], 4); }
    fence_barrier_init();
  }
  if (warp == SC_MMA_WARP) tmem_alloc(tmem_slot, (uint32_t)K.tmem_cols);
  {
    const uint4 z4 = make_uint4(0, 0, 0, 0);
    const int nz = (K.ring * K.slot_bytes) >> 4;
    for (int i = tid; i < nz; i += SC_THREADS) sts16(ring_u32 + (uint32_t)i * 16u, z4);
    // weights: block (tap, kc) = [Ncta rows x 128 B] SWIZZLE_128B K-major
    const bf16* wp = reinterpret_cast<const bf16*>(P.w_packed);
    const int nw = K.taps * K.Kc * K.Ncta * 8;
    for (int i = tid; i < nw; i += SC_THREADS) {
      const int c8 = i & 7, r = (i >> 3) % K.Ncta, blk = (i >> 3) / K.Ncta;
      const int tap = blk / K.Kc, kc = blk - tap * K.Kc;
      const int kk = tap * P.Cpad + kc * 64 + c8 * 8;
      const bool kin = K.taps == 9 ? (kc * 64 + c8 * 8 < P.Cpad) : (kk < P.Kpad);
      uint4 w = z4;
      if (n0 + r < P.Npad && kin) w = __ldg(reinterpret_cast<const uint4*>(wp + (int64_t)(n0 + r) * P.Kpad + kk));
      sts16(w_u32 + (uint32_t)(blk * K.Ncta * 128) + sw128_off((uint32_t)r, (uint32_t)c8), w);
    }
    const int nv = K.Kc * 64;
    for (int i = tid; i < nv; i += SC_THREADS) {
      float a = 1.f, b = 0.f, c = 0.f;
      if (i < P.C) {
        if (P.a_is_grad) {
          if (P.a_bwd.z) { a = P.a_bwd.p[i]; b = P.a_bwd.q[i]; c = P.a_bwd.t[i]; }
        } else if (P.a_fwd.scale) {
          a = P.a_fwd.scale[i]; b = P.a_fwd.shift[i];
        }
      }
      svec[i] = a; svec[nv + i] = b; svec[2 * nv + i] = c;
    }
    for (int i = tid; i < 2 * 2 * 256; i += SC_THREADS) sstat[i] = 0.f;
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  auto cur_init = [&](SCursor& c) {
    c.u = cta; c.j = 0; c.g = 0;
    c.ok = c.u < K.units;
    if (c.ok) c.ug = unit_geom(P, K, c.u);
  };
  auto cur_next = [&](SCursor& c) {
    ++c.j; ++c.g;
    if (c.j == c.ug.nload) {
      c.u += K.ctas_per_slice; c.j = 0;
      c.ok = c.u < K.units;
      if (c.ok) c.ug = unit_geom(P, K, c.u);
    }
  };

  if (warp < SC_PROD_WARPS) {
    // ================================================================== producers
    const bool is_grad = P.a_is_grad != 0;
    const bf16* a_ptr = reinterpret_cast<const bf16*>(is_grad ? P.a_bwd.dy : P.a_fwd.raw);
    const int64_t a_ld = is_grad ? P.a_bwd.ld_dy : P.a_fwd.ld;
    const bf16* z_ptr = is_grad ? reinterpret_cast<const bf16*>(P.a_bwd.z) : nullptr;
    const int64_t z_ld = P.a_bwd.ld_z;
    const float* gate = is_grad ? nullptr : P.a_fwd.gate;
    const float* rowscale = is_grad ? P.a_bwd.rowscale : nullptr;
    const int act = is_grad ? S2LC_ACT_NONE : P.a_fwd.act;
    const bool aff = is_grad ? (z_ptr != nullptr) : (P.a_fwd.scale != nullptr);
    const bool xf = aff || gate != nullptr || rowscale != nullptr || act != S2LC_ACT_NONE;
    const int CW = K.CW, total = K.lines * CW, nv = K.Kc * 64;
    const int line0 = tid / CW, c0 = tid - line0 * CW, dl = SC_PROD_THREADS / CW, dc = SC_PROD_THREADS - dl * CW;
    const int zslots = K.pf + 1;
    const uint32_t z_u32 = smem_u32(zring);

    auto slot_addr = [&](int g) { return ring_u32 + (uint32_t)((g % K.ring) * K.slot_bytes); };
    // validity + global row of (cursor, line)
    auto src_of = [&](const SCursor& c, int line, int64_t& row) -> bool {
      if (K.taps == 9) {
        const int hr = c.ug.h0 - 1 + c.j, w = c.ug.w0 - 1 + line;
        if (hr < 0 || hr >= P.H || w < 0 || w >= P.W) return false;
        row = ((int64_t)c.ug.n * P.H + hr) * P.W + w;
        return true;
      }
      row = c.ug.m0 + line;
      return true;
    };
    auto issue = [&](const SCursor& c) {
      const uint32_t sa = slot_addr(c.g);
      const uint32_t za = z_u32 + (uint32_t)((c.g % zslots) * K.zslot_bytes);
      int line = line0, ch = c0, it = tid;
#pragma unroll 1
      for (int i = 0; i < SC_MAX_ITEMS && it < total; ++i) {
        const uint32_t dst = sa + (uint32_t)((ch >> 3) * K.chunk_stride + line * 128 + (((ch & 7) ^ (line & 7)) << 4));
        int64_t row;
        if (src_of(c, line, row)) {
          cp_async16(dst, a_ptr + row * a_ld + ch * 8);
          if (z_ptr) cp_async16(za + (uint32_t)it * 16u, z_ptr + row * z_ld + ch * 8);
        } else {
          sts16(dst, make_uint4(0, 0, 0, 0));
        }
        it += SC_PROD_THREADS; line += dl; ch += dc;
        if (ch >= CW) { ch -= CW; ++line; }
      }
      cp_async_commit();
    };
    auto finish = [&](const SCursor& c) {
      if (xf) {
        const uint32_t sa = slot_addr(c.g);
        const uint32_t za = z_u32 + (uint32_t)((c.g % zslots) * K.zslot_bytes);
        const float rs = rowscale ? __ldg(rowscale + c.ug.n) : 1.f;
        int line = line0, ch = c0, it = tid;
#pragma unroll 1
        for (int i = 0; i < SC_MAX_ITEMS && it < total; ++i) {
          int64_t row;
          if (src_of(c, line, row)) {
            const uint32_t dst = sa + (uint32_t)((ch >> 3) * K.chunk_stride + line * 128 + (((ch & 7) ^ (line & 7)) << 4));
            float v[8];
            sc_unpack8(lds16(dst), v);
            const float* s0 = svec + ch * 8;
            if (is_grad) {
              if (rowscale) {
#pragma unroll
                for (int k = 0; k < 8; ++k) v[k] *= rs;
              }
              if (z_ptr) {
                float z[8];
                sc_unpack8(lds16(za + (uint32_t)it * 16u), z);
#pragma unroll
                for (int k = 0; k < 8; ++k) v[k] = fmaf(s0[k], v[k], fmaf(s0[nv + k], z[k], s0[2 * nv + k]));
              }
            } else {
              if (aff) {
#pragma unroll
                for (int k = 0; k < 8; ++k) v[k] = fmaf(v[k], s0[k], s0[nv + k]);
              }
              if (act != S2LC_ACT_NONE) {
#pragma unroll
                for (int k = 0; k < 8; ++k) v[k] = act_fwd(v[k], act);
              }
              if (gate) {
                float g8[8];
                sc_ldf8(gate + (int64_t)c.ug.n * P.C + ch * 8, g8);
#pragma unroll
                for (int k = 0; k < 8; ++k) v[k] *= g8[k];
              }
            }
            sts16(dst, sc_pack8(v));
          }
          it += SC_PROD_THREADS; line += dl; ch += dc;
          if (ch >= CW) { ch -= CW; ++line; }
        }
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) sc_mbar_arrive(&full[c.g % K.ring]);
    };

    SCursor ci, cf;
    cur_init(ci);
    cur_init(cf);
    int inflight = 0;
    while (cf.ok) {
      while (ci.ok && inflight <= K.pf) {
        if (ci.g >= K.ring) mbar_wait(&empty[ci.g % K.ring], (uint32_t)((ci.g / K.ring - 1) & 1));
        issue(ci);
        ++inflight;
        cur_next(ci);
      }
      // oldest group must have landed: allow (inflight - 1) younger groups to stay pending
      switch (inflight - 1) {
        case 0: cp_async_wait<0>(); break;
        case 1: cp_async_wait<1>(); break;
        case 2: cp_async_wait<2>(); break;
        default: cp_async_wait<3>(); break;
      }
      finish(cf);
      --inflight;
      cur_next(cf);
    }
  } else if (warp == SC_MMA_WARP) {
    // ================================================================== MMA issuer
    if (lane == 0) {
      const uint32_t idesc = idesc_bf16(128, (uint32_t)K.Ncta, 0, 0);
      int t = 0, gbase = 0, gw = 0;
      for (int u = cta; u < K.units; u += K.ctas_per_slice) {
        const UnitGeom ug = unit_geom(P, K, u);
        for (int o = 0; o < ug.nout; ++o, ++t) {
          const int need = gbase + o + R - 1;
          while (gw <= need) { mbar_wait(&full[gw % K.ring], (uint32_t)((gw / K.ring) & 1)); ++gw; }
          const int b = t % K.nacc;
          if (t >= K.nacc) mbar_wait(&acc_empty[b], (uint32_t)((t / K.nacc - 1) & 1));
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + (uint32_t)(b * K.Ncta);
          uint32_t accum = 0;
          for (int tap = 0; tap < K.taps; ++tap) {
            int dy1 = 0, dx1 = 0;                       // dy + 1, dx + 1
            if (K.taps == 9) { dy1 = tap / 3; dx1 = tap - dy1 * 3; }
            const uint32_t a_base = ring_u32 + (uint32_t)(((gbase + o + dy1) % K.ring) * K.slot_bytes + dx1 * 128);
            for (int kc = 0; kc < K.Kc; ++kc) {
              int rem = P.C - kc * 64;
              rem = rem > 64 ? 64 : rem;
              const int ks = (rem + 15) >> 4;
              const uint32_t b_base = w_u32 + (uint32_t)((tap * K.Kc + kc) * K.Ncta * 128);
              for (int k4 = 0; k4 < ks; ++k4) {
                umma_bf16(d_tmem, sc_desc(a_base + (uint32_t)(kc * K.chunk_stride + k4 * 32)),
                          smem_desc_sw128(b_base + (uint32_t)(k4 * 32), 1024, 16), idesc, accum);
                accum = 1;
              }
            }
          }
          umma_commit(&acc_full[b]);
          umma_commit(&empty[(gbase + o) % K.ring]);
          if (R == 3 && o == ug.nout - 1) {
            umma_commit(&empty[(gbase + o + 1) % K.ring]);
            umma_commit(&empty[(gbase + o + 2) % K.ring]);
          }
        }
        gbase += ug.nload;
      }
    }
  } else {
    // ================================================================== epilogue groups
    const int ew = warp - SC_EPI_WARP0, grp = ew >> 2, q = warp & 3;   // q = TMEM lane quarter this warp may read
    float* gstat = sstat + grp * 512;                                   // [2][256] of this group
    const int gtid = (ew & 3) * 32 + lane;
    const int nchunks = K.Ncta / 16;
    bf16* outp = reinterpret_cast<bf16*>(P.out);
    int t = 0, since_flush = 0;
    auto flush = [&]() {
      asm volatile("bar.sync %0, 128;" ::"r"(2 + grp) : "memory");
      for (int i = gtid; i < K.Ncta; i += 128) {
        if (n0 + i < P.Nout) {
          atomicAdd(&P.stats[n0 + i], (double)gstat[i]);
          atomicAdd(&P.stats[P.Nout + n0 + i], (double)gstat[256 + i]);
        }
        gstat[i] = 0.f;
        gstat[256 + i] = 0.f;
      }
      asm volatile("bar.sync %0, 128;" ::"r"(2 + grp) : "memory");
    };
    for (int u = cta; u < K.units; u += K.ctas_per_slice) {
      const UnitGeom ug = unit_geom(P, K, u);
      for (int o = 0; o < ug.nout; ++o, ++t) {
        if ((t & 1) != grp) continue;
        const int b = t % K.nacc;
        mbar_wait(&acc_full[b], (uint32_t)((t / K.nacc) & 1));
        tc_fence_after();
        const int r = q * 32 + lane;
        int64_t orow;
        int n_s;
        if (K.taps == 9) {
          n_s = ug.n;
          orow = ((int64_t)ug.n * P.H + ug.h0 + o) * P.W + ug.w0 + r;
        } else {
          n_s = ug.n;
          orow = ug.m0 + r;
        }
        for (int cj = 0; cj < nchunks; ++cj) {
          const int col0 = n0 + cj * 16;
          int nvalid = P.Nout - col0;
          nvalid = nvalid > 16 ? 16 : nvalid;
          if (nvalid <= 0) break;   // warp-uniform
          float acc[16];
          tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(b * K.Ncta + cj * 16), acc);
          const bool full16 = nvalid == 16;
          if (P.bias) {
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (j < nvalid) acc[j] += __ldg(P.bias + col0 + j);
          }
          auto add_rows = [&](const void* base, int64_t ld) {
            const bf16* p = reinterpret_cast<const bf16*>(base) + orow * ld + col0;
            if (full16 && ((reinterpret_cast<uintptr_t>(p) & 15) == 0)) {
              float e[8];
              sc_unpack8(__ldg(reinterpret_cast<const uint4*>(p)), e);
#pragma unroll
              for (int j = 0; j < 8; ++j) acc[j] += e[j];
              sc_unpack8(__ldg(reinterpret_cast<const uint4*>(p + 8)), e);
#pragma unroll
              for (int j = 0; j < 8; ++j) acc[8 + j] += e[j];
            } else {
#pragma unroll
              for (int j = 0; j < 16; ++j)
                if (j < nvalid) acc[j] += __bfloat162float(p[j]);
            }
          };
          if (P.res1) add_rows(P.res1, P.ld_res1);
          if (P.res2) add_rows(P.res2, P.ld_res2);
          float zp[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) zp[j] = 0.f;
          if (P.has_prev) {
            const bf16* p = reinterpret_cast<const bf16*>(P.prev.raw) + orow * P.prev.ld + col0;
            if (full16 && ((reinterpret_cast<uintptr_t>(p) & 15) == 0)) {
              float e[8];
              sc_unpack8(__ldg(reinterpret_cast<const uint4*>(p)), e);
#pragma unroll
              for (int j = 0; j < 8; ++j) zp[j] = e[j];
              sc_unpack8(__ldg(reinterpret_cast<const uint4*>(p + 8)), e);
#pragma unroll
              for (int j = 0; j < 8; ++j) zp[8 + j] = e[j];
            } else {
#pragma unroll
              for (int j = 0; j < 16; ++j)
                if (j < nvalid) zp[j] = __bfloat162float(p[j]);
            }
            if (P.prev.act != S2LC_ACT_NONE) {
#pragma unroll
              for (int j = 0; j < 16; ++j)
                if (j < nvalid) {
                  float uu = zp[j];
                  if (P.prev.scale) uu = fmaf(uu, __ldg(P.prev.scale + col0 + j), __ldg(P.prev.shift + col0 + j));
                  acc[j] *= act_bwd(uu, P.prev.act);
                }
            }
          }
#pragma unroll
          for (int j = 0; j < 16; ++j) acc[j] = roundf_as<bf16>(acc[j]);
          bf16* op = outp + orow * P.ld_out + col0;
          if (full16 && ((reinterpret_cast<uintptr_t>(op) & 15) == 0)) {
            float lo[8], hi[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) { lo[j] = acc[j]; hi[j] = acc[8 + j]; }
            *reinterpret_cast<uint4*>(op) = sc_pack8(lo);
            *reinterpret_cast<uint4*>(op + 8) = sc_pack8(hi);
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (j < nvalid) op[j] = __float2bfloat16_rn(acc[j]);
          }
          if (P.stats_kind != S2LC_STATS_NONE) {
            float s0[16], s1[16];
            if (P.stats_kind == S2LC_STATS_MOMENTS) {
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                const float v = j < nvalid ? acc[j] : 0.f;
                s0[j] = v; s1[j] = v * v;
              }
            } else {
              const float rs = P.stats_rowscale ? __ldg(P.stats_rowscale + n_s) : 1.f;
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                const float v = j < nvalid ? acc[j] * rs : 0.f;
                s0[j] = v; s1[j] = v * zp[j];
              }
            }
            int col;
            const float t0 = sc_butterfly16(s0, lane, col);
            const float t1 = sc_butterfly16(s1, lane, col);
            if ((lane & 1) == 0) {
              atomicAdd(&gstat[cj * 16 + col], t0);
              atomicAdd(&gstat[256 + cj * 16 + col], t1);
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) sc_mbar_arrive(&acc_empty[b]);
        ++since_flush;
      }
      if (P.stats_kind != S2LC_STATS_NONE && since_flush >= 16) { flush(); since_flush = 0; }
    }
    if (P.stats_kind != S2LC_STATS_NONE && since_flush > 0) flush();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == SC_MMA_WARP) tmem_dealloc(tmem_base, (uint32_t)K.tmem_cols);
}

static bool sc_aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
static int sc_up(int x, int m) { return (x + m - 1) / m * m; }

// Returns 0 if launched, <0 on error, 1 if this shape is left to the tile-at-a-time engine (igemm_tc.cu).
int32_t launch_sconv_tc(const ConvArgs& A, cudaStream_t st) {
  {
    const char* e = getenv("S2LC_NO_STREAM");   // test hook: compare against the tile-at-a-time engine
    if (e && e[0] == '1') return 1;
  }
  if (A.dtype != S2LC_BF16 || (A.mode != S2LC_CONV_3X3 && A.mode != S2LC_CONV_1X1)) return 1;
  if (A.C % 8 != 0 || A.Kpad % 64 != 0 || !sc_aligned16(A.w_packed)) return 1;
  if (A.a_is_grad) {
    if (A.a_bwd.ld_dy % 8 != 0 || !sc_aligned16(A.a_bwd.dy)) return 1;
    if (A.a_bwd.z && (A.a_bwd.ld_z % 8 != 0 || !sc_aligned16(A.a_bwd.z))) return 1;
  } else {
    if (A.a_fwd.ld % 8 != 0 || !sc_aligned16(A.a_fwd.raw)) return 1;
    if (A.a_fwd.scale && !A.a_fwd.shift) return 1;
  }
  if (A.has_prev && A.prev.scale && !A.prev.shift) return 1;
  const int64_t HW = (int64_t)A.H * A.W, M = HW * A.N;
  SPlan K{};
  K.taps = A.mode == S2LC_CONV_3X3 ? 9 : 1;
  K.Kc = (A.C + 63) / 64;
  K.CW = A.C / 8;
  if (K.taps == 9) {
    if (A.W % 128 != 0 || A.Cpad != K.Kc * 64 || K.Kc > 2) return 1;
    K.lines = 130;
    K.chunk_stride = 17 * 1024;
  } else {
    if (HW % 128 != 0 || M < 128 * 2 * kNumSMs || K.Kc > 8 || K.Kc * 64 > A.Kpad) return 1;
    K.lines = 128;
    K.chunk_stride = 16 * 1024;
  }
  if ((K.lines * K.CW + SC_PROD_THREADS - 1) / SC_PROD_THREADS > SC_MAX_ITEMS) return 1;
  K.slot_bytes = K.Kc * K.chunk_stride;
  const bool has_z = A.a_is_grad && A.a_bwd.z != nullptr;
  K.zslot_bytes = has_z ? K.lines * K.CW * 16 : 0;
  const int R = K.taps == 9 ? 3 : 1;
  const int n16 = sc_up(A.Nout, 16);
  const int budget = 226 * 1024;
  const int fixed = 1024 + 3 * K.Kc * 64 * 4 + 2 * 2 * 256 * 4 + (2 * SC_MAX_RING + 2 * SC_MAX_ACC) * 8 + 64;
  bool found = false;
  for (int ns = 1; ns <= 4 && !found; ++ns) {
    const int ncta = sc_up((n16 + ns - 1) / ns, 16);
    if (ncta > 256) continue;
    const int wb = K.taps * K.Kc * ncta * 128;
    for (int ring = 6; ring >= R + 1; --ring) {
      int pf = ring - R;
      if (pf > 3) pf = 3;
      if (fixed + wb + ring * K.slot_bytes + (pf + 1) * K.zslot_bytes <= budget) {
        K.nsplit = ns; K.Ncta = ncta; K.w_bytes = wb; K.ring = ring; K.pf = pf;
        found = true;
        break;
      }
    }
  }
  if (!found) return 1;
  K.nacc = 512 / K.Ncta;
  if (K.nacc > SC_MAX_ACC) K.nacc = SC_MAX_ACC;
  if (K.nacc == 3) K.nacc = 2;
  if (K.nacc < 2) return 1;
  K.tmem_cols = 32;
  while (K.tmem_cols < K.nacc * K.Ncta) K.tmem_cols *= 2;
  int ctas = kNumSMs / K.nsplit;
  if (K.taps == 9) {
    K.WS = A.W / 128;
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
    K.RU = 1; K.HS = 1; K.WS = 1;
    K.units = (int)(M / 128);
  }
  if (K.units < ctas) ctas = K.units;
  K.ctas_per_slice = ctas;
  const size_t smem = (size_t)fixed + K.w_bytes + (size_t)K.ring * K.slot_bytes + (size_t)(K.pf + 1) * K.zslot_bytes;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(sconv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) {
      set_error("sconv_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
      return S2LC_ERR_CUDA;
    }
    attr_set = true;
  }
  sconv_kernel<<<ctas * K.nsplit, SC_THREADS, smem, st>>>(A, K);
  return check_launch("sconv_tc");
}

}  // namespace s2lc

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
// Roles (512 threads):
//   warps 0-6   producers: cp.async (16 B, straight to the swizzled destination — no staging copy, no registers),
//               prefetching `pf` slots ahead; when a slot has landed each thread transforms ITS OWN chunks in place
//               (BatchNorm scale/shift + SiLU/ReLU + SE gate, or the BatchNorm-backward affine p*dy + q*z + t with z
//               staged beside it), fence.proxy.async, arrive on the slot's `full` mbarrier;
//   warp 7      MMA issuer: one lane issues tcgen05.mma for all taps / K chunks of an output tile into one of `nacc`
//               TMEM accumulators; tcgen05.commit frees the oldest ring slot and publishes the accumulator;
//   warps 8-15  two epilogue groups (alternate 16-column chunks): tcgen05.ld -> bias, residuals, act'(prev), bf16 store, BatchNorm
//               moments (butterfly column sums -> smem -> fp64 global atomics once per unit).
// Packed weights of the CTA's N slice stay resident in shared memory for the whole kernel.
#include <cstdlib>
#include <type_traits>

#include "conv_common.cuh"
#include "stream_common.cuh"

namespace s2lc {

using namespace tc;

// 16 warps -> 128 registers per thread.  Warp 15 issues the MMAs; warps [0, prod_warps) are producers and warps
// [prod_warps, 15) form epilogue groups of four (one warp per TMEM lane quarter).  The split is a plan parameter:
// 7 + 8 (two epilogue groups) by default, 3 + 12 (three groups) for thin-A / wide-N layers (1x1 expand convs), where
// the producers finish a slot in a fraction of the time the epilogue needs for its 128 x N tile.
constexpr int SC_WARPS = 16;
constexpr int SC_MMA_WARP = SC_WARPS - 1;
constexpr int SC_THREADS = SC_WARPS * 32;
constexpr int SC_MAX_EGROUPS = 3;
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
  int prod_warps;    // producer warps: 7 (two epilogue groups), 3 (three groups, thin A / wide N) or 11 (one group)
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
  int KS;                  // k-steps (K = 16 each) per tap
  uint32_t koff_a[32];     // descriptor-address offsets (>> 4) of k-step s inside a slot / a weight tap
  uint32_t koff_b[32];
};

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

// TAPS: 9 (3x3, rolling rows) or 1 (1x1, one tile per slot).
// REGSTAT: every epilogue group owns ONE 16-column chunk (Ncta <= 32), so BatchNorm moments accumulate in registers
// across tiles and are reduced over the warp only when they are flushed.
template <int TAPS, bool REGSTAT>
__global__ void __launch_bounds__(SC_THREADS, 1) sconv_kernel(ConvArgs P, SPlan K) {
  cudaGridDependencySynchronize();   // programmatic dependent launch (common.cuh)
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* smem =
      reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  unsigned char* wsm = smem;                                   // resident weights
  unsigned char* ring = wsm + K.w_bytes;                       // [ring][slot_bytes]
  unsigned char* zring = ring + K.ring * K.slot_bytes;         // [pf+1][zslot_bytes]
  float* svec = reinterpret_cast<float*>(zring + (K.pf + 1) * K.zslot_bytes);   // [3][Kc*64] operand affine vectors
  float* evec = svec + 3 * K.Kc * 64;                          // [3][256] epilogue vectors: bias, prev scale, prev shift
  const int sgroups = K.prod_warps == 3 ? SC_MAX_EGROUPS : 2;   // epilogue groups that own a statistics area
  float* sstat = evec + 3 * 256;                               // [sgroups][2][256]
  uint64_t* full = reinterpret_cast<uint64_t*>(sstat + sgroups * 2 * 256);
  uint64_t* empty = full + SC_MAX_RING;
  uint64_t* acc_full = empty + SC_MAX_RING;
  uint64_t* acc_empty = acc_full + SC_MAX_ACC;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + SC_MAX_ACC);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int slice = blockIdx.x % K.nsplit, cta = blockIdx.x / K.nsplit;
  const int n0 = slice * K.Ncta;
  constexpr int R = TAPS == 9 ? 3 : 1;
  const uint32_t ring_u32 = smem_u32(ring), w_u32 = smem_u32(wsm);
  const int nchunks = K.Ncta / 16;
  const int epi_groups = (SC_MMA_WARP - K.prod_warps) >> 2;
  const int egroups = nchunks < epi_groups ? nchunks : epi_groups;

  // ------------------------------------------------------------------ one-time setup (all threads)
  if (tid == 0) {
    for (int s = 0; s < K.ring; ++s) { mbar_init(&full[s], (uint32_t)K.prod_warps); mbar_init(&empty[s], 1); }
    for (int b = 0; b < K.nacc; ++b) { mbar_init(&acc_full[b], 1); mbar_init(&acc_empty[b], 4 * egroups); }
    fence_barrier_init();
  }
  if (warp == SC_MMA_WARP) tmem_alloc(tmem_slot, (uint32_t)K.tmem_cols);
  {
    const uint4 z4 = make_uint4(0, 0, 0, 0);
    const int nz = (K.ring * K.slot_bytes) >> 4;
    for (int i = tid; i < nz; i += SC_THREADS) sts16(ring_u32 + (uint32_t)i * 16u, z4);
    // weights: block (tap, kc) = [Ncta rows x 128 B] SWIZZLE_128B K-major
    const bf16* wp = reinterpret_cast<const bf16*>(P.w_packed);
    const int nw = TAPS * K.Kc * K.Ncta * 8;
    for (int i = tid; i < nw; i += SC_THREADS) {
      const int c8 = i & 7, r = (i >> 3) % K.Ncta, blk = (i >> 3) / K.Ncta;
      const int tap = blk / K.Kc, kc = blk - tap * K.Kc;
      const int kk = tap * P.Cpad + kc * 64 + c8 * 8;
      const bool kin = TAPS == 9 ? (kc * 64 + c8 * 8 < P.Cpad) : (kk < P.Kpad);
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
    for (int i = tid; i < 256; i += SC_THREADS) {
      const int col = n0 + i;
      const bool in = i < K.Ncta && col < P.Nout;
      evec[i] = (in && P.bias) ? P.bias[col] : 0.f;
      evec[256 + i] = (in && P.has_prev && P.prev.scale) ? P.prev.scale[col] : 1.f;
      evec[512 + i] = (in && P.has_prev && P.prev.scale) ? P.prev.shift[col] : 0.f;
    }
    for (int i = tid; i < sgroups * 2 * 256; i += SC_THREADS) sstat[i] = 0.f;
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

  if (warp < K.prod_warps) {
    // ================================================================== producers
    const bool is_grad = P.a_is_grad != 0;
    const bf16* a_ptr = reinterpret_cast<const bf16*>(is_grad ? P.a_bwd.dy : P.a_fwd.raw);
    const int a_ld = (int)(is_grad ? P.a_bwd.ld_dy : P.a_fwd.ld);
    const bf16* z_ptr = is_grad ? reinterpret_cast<const bf16*>(P.a_bwd.z) : nullptr;   // same row stride as dy
    const float* gate = is_grad ? nullptr : P.a_fwd.gate;
    const float* rowscale = is_grad ? P.a_bwd.rowscale : nullptr;
    const int act = is_grad ? S2LC_ACT_NONE : P.a_fwd.act;
    const bool aff = is_grad ? (z_ptr != nullptr) : (P.a_fwd.scale != nullptr);
    const bool xf = aff || gate != nullptr || rowscale != nullptr || act != S2LC_ACT_NONE;
    // every ACTIVE producer thread owns one fixed 8-channel chunk `ch` (so its affine vectors live in registers) and
    // walks the lines of a slot with stride dl
    const int CW = K.CW, nv = K.Kc * 64;
    const int dl = (K.prod_warps * 32) / CW, nact = dl * CW;
    const bool active = tid < nact;
    const int line0 = tid / CW, ch = tid - line0 * CW;
    const int nit = active ? (K.lines - line0 + dl - 1) / dl : 0;   // lines this thread owns in every slot
    const int zslots = K.pf + 1;
    const uint32_t z_u32 = smem_u32(zring);
    const uint32_t dst_c = (uint32_t)((ch >> 3) * K.chunk_stride), ch7 = (uint32_t)(ch & 7);
    float va[8], vb[8], vc[8];
    {
      const float* s0 = svec + (active ? ch * 8 : 0);
#pragma unroll
      for (int k = 0; k < 8; ++k) { va[k] = s0[k]; vb[k] = s0[nv + k]; vc[k] = s0[2 * nv + k]; }
    }
    int islot = 0, fslot = 0, izs = 0, fzs = 0;      // ring / z-ring positions of the issue and finish cursors
    uint32_t iphase = 0;                              // parity to wait for on `empty` (valid once the ring wrapped)
    bool wrapped = false;

    // geometry of one load: base row (pixel index of line 0), valid line range [lo, hi]
    auto load_geom = [&](const SCursor& c, int64_t& row0, int& lo, int& hi) {
      if (TAPS == 9) {
        const int hr = c.ug.h0 - 1 + c.j;
        row0 = ((int64_t)c.ug.n * P.H + hr) * P.W + c.ug.w0 - 1;
        lo = c.ug.w0 == 0 ? 1 : 0;
        hi = c.ug.w0 + 128 == P.W ? 128 : 129;
        if (hr < 0 || hr >= P.H) { lo = 1; hi = 0; }   // whole row is zero padding
      } else {
        row0 = c.ug.m0; lo = 0; hi = 127;
      }
    };
    auto issue = [&](const SCursor& c) {
      const uint32_t sa = ring_u32 + (uint32_t)(islot * K.slot_bytes) + dst_c;
      const uint32_t za = z_u32 + (uint32_t)(izs * K.zslot_bytes) + (uint32_t)tid * 16u;
      int64_t row0; int lo, hi;
      load_geom(c, row0, lo, hi);
      const bf16* abase = a_ptr + row0 * a_ld + ch * 8;
      const bf16* zbase = z_ptr ? z_ptr + row0 * a_ld + ch * 8 : nullptr;
      {
        int line = line0;
#pragma unroll 2
        for (int i = 0; i < nit; ++i, line += dl) {
          const uint32_t dst = sa + (uint32_t)(line * 128) + ((ch7 ^ (uint32_t)(line & 7)) << 4);
          if (line >= lo && line <= hi) {
            const int off = line * a_ld;
            cp_async16(dst, abase + off);
            if (zbase) cp_async16(za + (uint32_t)(i * nact) * 16u, zbase + off);
          } else {
            sts16(dst, make_uint4(0, 0, 0, 0));
          }
        }
      }
      cp_async_commit();
    };
    auto finish = [&](const SCursor& c) {
      if (xf && active) {
        const uint32_t sa = ring_u32 + (uint32_t)(fslot * K.slot_bytes) + dst_c;
        const uint32_t za = z_u32 + (uint32_t)(fzs * K.zslot_bytes) + (uint32_t)tid * 16u;
        int64_t row0; int lo, hi;
        load_geom(c, row0, lo, hi);
        const float rs = rowscale ? __ldg(rowscale + c.ug.n) : 1.f;
        float g8[8];
        if (gate) sc_ldf8(gate + (int64_t)c.ug.n * P.C + ch * 8, g8);
        int line = line0;
#pragma unroll 2
        for (int i = 0; i < nit; ++i, line += dl) {
          if (line >= lo && line <= hi) {
            const uint32_t dst = sa + (uint32_t)(line * 128) + ((ch7 ^ (uint32_t)(line & 7)) << 4);
            float v[8];
            sc_unpack8(lds16(dst), v);
            if (is_grad) {
              if (rowscale) {
#pragma unroll
                for (int k = 0; k < 8; ++k) v[k] *= rs;
              }
              if (z_ptr) {
                float z[8];
                sc_unpack8(lds16(za + (uint32_t)(i * nact) * 16u), z);
#pragma unroll
                for (int k = 0; k < 8; ++k) v[k] = fmaf(va[k], v[k], fmaf(vb[k], z[k], vc[k]));
              }
            } else {
              if (aff) {
#pragma unroll
                for (int k = 0; k < 8; ++k) v[k] = fmaf(v[k], va[k], vb[k]);
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
            }
            sts16(dst, sc_pack8(v));
          }
        }
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) sc_mbar_arrive(&full[fslot]);
    };

    SCursor ci, cf;
    cur_init(ci);
    cur_init(cf);
    int inflight = 0;
    while (cf.ok) {
      while (ci.ok && inflight <= K.pf) {
        if (wrapped) mbar_wait(&empty[islot], iphase);
        issue(ci);
        ++inflight;
        cur_next(ci);
        if (++islot == K.ring) { islot = 0; if (wrapped) iphase ^= 1u; wrapped = true; }
        if (++izs == zslots) izs = 0;
      }
      switch (inflight - 1) {   // the oldest group must have landed; younger ones may stay in flight
        case 0: cp_async_wait<0>(); break;
        case 1: cp_async_wait<1>(); break;
        case 2: cp_async_wait<2>(); break;
        default: cp_async_wait<3>(); break;
      }
      finish(cf);
      --inflight;
      cur_next(cf);
      if (++fslot == K.ring) fslot = 0;
      if (++fzs == zslots) fzs = 0;
    }
  } else if (warp == SC_MMA_WARP) {
    // ================================================================== MMA issuer.  The WHOLE warp walks the loop
    // converged (so every address is warp-uniform and lives in uniform registers); one elected lane issues.
    {
      const bool leader = elect_one();
      const uint32_t idesc = idesc_bf16(128, (uint32_t)K.Ncta, 0, 0);
      const uint64_t dbase = smem_desc_sw128(0, 1024, 16);          // address bits are added below (addr >> 4 < 2^14)
      const uint32_t a_lo = ring_u32 >> 4, w_lo = w_u32 >> 4;
      const uint32_t slot16 = (uint32_t)K.slot_bytes >> 4;
      const uint32_t wblk16 = (uint32_t)(K.Ncta * 128) >> 4;        // one (tap, kc) weight block
      int fs = 0;  uint32_t fph = 0;        // next `full` barrier to observe
      int s_first = 0;                      // ring slot of load 0 of the current unit
      int b = 0;   uint32_t bph = 0;  bool bwrapped = false;
      for (int u = cta; u < K.units; u += K.ctas_per_slice) {
        const UnitGeom ug = unit_geom(P, K, u);
        int waited = 0;                     // loads of this unit already observed
        int r0 = s_first;
        for (int o = 0; o < ug.nout; ++o) {
          while (waited < o + R) {
            mbar_wait(&full[fs], fph);
            if (++fs == K.ring) { fs = 0; fph ^= 1u; }
            ++waited;
          }
          if (bwrapped) mbar_wait(&acc_empty[b], bph);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + (uint32_t)(b * K.Ncta);
          uint32_t accum = 0;
          int rs = r0;
#pragma unroll
          for (int dy1 = 0; dy1 < R; ++dy1) {
            const uint32_t a_row = a_lo + (uint32_t)rs * slot16;
#pragma unroll
            for (int dx1 = 0; dx1 < R; ++dx1) {
              const uint64_t a_tap = dbase + (uint64_t)(a_row + (uint32_t)(dx1 * 8));          // + dx1 lines of 128 B
              const uint64_t b_tap = dbase + (uint64_t)(w_lo + (uint32_t)((dy1 * R + dx1) * K.Kc) * wblk16);
#pragma unroll 2
              for (int ks = 0; ks < K.KS; ++ks) {
                if (leader) umma_bf16(d_tmem, a_tap + K.koff_a[ks], b_tap + K.koff_b[ks], idesc, accum);
                accum = 1;
              }
            }
            if (++rs == K.ring) rs = 0;
          }
          if (leader) {
            umma_commit(&acc_full[b]);
            umma_commit(&empty[r0]);
            if (R == 3 && o == ug.nout - 1) {
              int r1 = r0 + 1 == K.ring ? 0 : r0 + 1;
              int r2 = r1 + 1 == K.ring ? 0 : r1 + 1;
              umma_commit(&empty[r1]);
              umma_commit(&empty[r2]);
            }
          }
          __syncwarp();
          if (++r0 == K.ring) r0 = 0;
          if (++b == K.nacc) { b = 0; if (bwrapped) bph ^= 1u; bwrapped = true; }
        }
        s_first += ug.nload;
        while (s_first >= K.ring) s_first -= K.ring;
      }
    }
  } else {
    // ================================================================== epilogue: groups of 4 lane-quarter warps;
    // group g owns the 16-column chunks cj = g, g+egroups, ... of EVERY tile.
    const int ew = warp - K.prod_warps, grp = ew >> 2, q = warp & 3;   // q = TMEM lane quarter this warp may read
    if (grp < egroups) {
      float* gstat = sstat + grp * 512;                                 // [2][256] of this group
      const int gtid = (ew & 3) * 32 + lane;
      bf16* outp = reinterpret_cast<bf16*>(P.out);
      const bf16* prevp = P.has_prev ? reinterpret_cast<const bf16*>(P.prev.raw) : nullptr;
      const bf16* res1p = reinterpret_cast<const bf16*>(P.res1);
      const bf16* res2p = reinterpret_cast<const bf16*>(P.res2);
      const int prev_act = P.has_prev ? P.prev.act : S2LC_ACT_NONE;
      const int r = q * 32 + lane;
      int b = 0;  uint32_t bph = 0;
      int since_flush = 0;
      float rs0[16], rs1[16];                     // REGSTAT accumulators (chunk `grp`)
#pragma unroll
      for (int j = 0; j < 16; ++j) { rs0[j] = 0.f; rs1[j] = 0.f; }
      auto flush = [&]() {
        if (REGSTAT) {
          int col;
          const float t0 = sc_butterfly16(rs0, lane, col);
          const float t1 = sc_butterfly16(rs1, lane, col);
          if ((lane & 1) == 0) {
            atomicAdd(&gstat[grp * 16 + col], t0);
            atomicAdd(&gstat[256 + grp * 16 + col], t1);
          }
#pragma unroll
          for (int j = 0; j < 16; ++j) { rs0[j] = 0.f; rs1[j] = 0.f; }
        }
        asm volatile("bar.sync %0, 128;" ::"r"(2 + grp) : "memory");
        for (int i = gtid; i < K.Ncta; i += 128) {
          const float v0 = gstat[i], v1 = gstat[256 + i];
          if (n0 + i < P.Nout && (v0 != 0.f || v1 != 0.f)) {
            atomicAdd(&P.stats[n0 + i], (double)v0);
            atomicAdd(&P.stats[P.Nout + n0 + i], (double)v1);
          }
          gstat[i] = 0.f;
          gstat[256 + i] = 0.f;
        }
        asm volatile("bar.sync %0, 128;" ::"r"(2 + grp) : "memory");
      };
      const bool has_bias = P.bias != nullptr;
      const int stats_kind = P.stats_kind;
      // one 16-column chunk of one tile row.  FULL: all 16 columns exist (16 B vector path, no per-column predicates)
      auto chunk = [&](auto full_tag, int lc0, int col0, int nvalid, int64_t orow, int b, float srs) {
        constexpr bool FULL = decltype(full_tag)::value;
        // issue the global loads of this chunk first, then read the accumulator
        uint4 pz0 = make_uint4(0, 0, 0, 0), pz1 = pz0, ra0 = pz0, ra1 = pz0, rb0 = pz0, rb1 = pz0;
        if (FULL) {
          if (prevp) {
            const uint4* p = reinterpret_cast<const uint4*>(prevp + orow * P.prev.ld + col0);
            pz0 = __ldg(p); pz1 = __ldg(p + 1);
          }
          if (res1p) {
            const uint4* p = reinterpret_cast<const uint4*>(res1p + orow * P.ld_res1 + col0);
            ra0 = __ldg(p); ra1 = __ldg(p + 1);
          }
          if (res2p) {
            const uint4* p = reinterpret_cast<const uint4*>(res2p + orow * P.ld_res2 + col0);
            rb0 = __ldg(p); rb1 = __ldg(p + 1);
          }
        }
        float acc[16];
        tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(b * K.Ncta + lc0), acc);
        if (has_bias) {
          const float4* bv = reinterpret_cast<const float4*>(evec + lc0);
#pragma unroll
          for (int j4 = 0; j4 < 4; ++j4) {
            const float4 t = bv[j4];
            acc[4 * j4] += t.x; acc[4 * j4 + 1] += t.y; acc[4 * j4 + 2] += t.z; acc[4 * j4 + 3] += t.w;
          }
        }
        float zp[16];
        if (FULL) {
          float e[8];
          if (res1p) {
            sc_unpack8(ra0, e);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] += e[j];
            sc_unpack8(ra1, e);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[8 + j] += e[j];
          }
          if (res2p) {
            sc_unpack8(rb0, e);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] += e[j];
            sc_unpack8(rb1, e);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[8 + j] += e[j];
          }
          sc_unpack8(pz0, *reinterpret_cast<float(*)[8]>(&zp[0]));
          sc_unpack8(pz1, *reinterpret_cast<float(*)[8]>(&zp[8]));
        } else {   // ragged last chunk: scalar path
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            zp[j] = 0.f;
            if (j < nvalid) {
              if (res1p) acc[j] += __bfloat162float(res1p[orow * P.ld_res1 + col0 + j]);
              if (res2p) acc[j] += __bfloat162float(res2p[orow * P.ld_res2 + col0 + j]);
              if (prevp) zp[j] = __bfloat162float(prevp[orow * P.prev.ld + col0 + j]);
            }
          }
        }
        if (prev_act == S2LC_ACT_RELU) {
#pragma unroll
          for (int j = 0; j < 16; ++j)
            acc[j] = fmaf(zp[j], evec[256 + lc0 + j], evec[512 + lc0 + j]) > 0.f ? acc[j] : 0.f;
        } else if (prev_act == S2LC_ACT_SILU) {
#pragma unroll
          for (int j = 0; j < 16; ++j)
            acc[j] *= act_bwd_fast(fmaf(zp[j], evec[256 + lc0 + j], evec[512 + lc0 + j]), S2LC_ACT_SILU);
        }
        // round to the stored precision (statistics must see what the consumer will read back)
        uint4 o0, o1;
        {
          __nv_bfloat162* h0 = reinterpret_cast<__nv_bfloat162*>(&o0);
          __nv_bfloat162* h1 = reinterpret_cast<__nv_bfloat162*>(&o1);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            h0[j] = __floats2bfloat162_rn(acc[2 * j], acc[2 * j + 1]);
            h1[j] = __floats2bfloat162_rn(acc[8 + 2 * j], acc[8 + 2 * j + 1]);
          }
        }
        bf16* op = outp + orow * P.ld_out + col0;
        if (FULL) {
          *reinterpret_cast<uint4*>(op) = o0;
          *reinterpret_cast<uint4*>(op + 8) = o1;
        } else {
          const bf16* hv0 = reinterpret_cast<const bf16*>(&o0);
          const bf16* hv1 = reinterpret_cast<const bf16*>(&o1);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            if (j < nvalid) op[j] = hv0[j];
            if (8 + j < nvalid) op[8 + j] = hv1[j];
          }
        }
        if (stats_kind != S2LC_STATS_NONE) {
          float v[16];
          sc_unpack8(o0, *reinterpret_cast<float(*)[8]>(&v[0]));
          sc_unpack8(o1, *reinterpret_cast<float(*)[8]>(&v[8]));
          if (!FULL) {
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (j >= nvalid) v[j] = 0.f;
          }
          if (REGSTAT) {
            if (stats_kind == S2LC_STATS_MOMENTS) {
#pragma unroll
              for (int j = 0; j < 16; ++j) { rs0[j] += v[j]; rs1[j] = fmaf(v[j], v[j], rs1[j]); }
            } else {
#pragma unroll
              for (int j = 0; j < 16; ++j) { const float x = v[j] * srs; rs0[j] += x; rs1[j] = fmaf(x, zp[j], rs1[j]); }
            }
          } else {
            float s0[16], s1[16];
            if (stats_kind == S2LC_STATS_MOMENTS) {
#pragma unroll
              for (int j = 0; j < 16; ++j) { s0[j] = v[j]; s1[j] = v[j] * v[j]; }
            } else {
#pragma unroll
              for (int j = 0; j < 16; ++j) { const float x = v[j] * srs; s0[j] = x; s1[j] = x * zp[j]; }
            }
            int col;
            const float t0 = sc_butterfly16(s0, lane, col);
            const float t1 = sc_butterfly16(s1, lane, col);
            if ((lane & 1) == 0) {
              atomicAdd(&gstat[lc0 + col], t0);
              atomicAdd(&gstat[256 + lc0 + col], t1);
            }
          }
        }
      };
      for (int u = cta; u < K.units; u += K.ctas_per_slice) {
        const UnitGeom ug = unit_geom(P, K, u);
        int64_t orow = TAPS == 9 ? ((int64_t)ug.n * P.H + ug.h0) * P.W + ug.w0 + r : ug.m0 + r;
        const float srs = (P.stats_kind == S2LC_STATS_BNBWD && P.stats_rowscale) ? __ldg(P.stats_rowscale + ug.n) : 1.f;
        for (int o = 0; o < ug.nout; ++o, orow += P.W) {
          mbar_wait(&acc_full[b], bph);
          tc_fence_after();
          for (int cj = grp; cj < nchunks; cj += egroups) {
            const int lc0 = cj * 16, col0 = n0 + lc0;
            const int nvalid = P.Nout - col0;
            if (nvalid <= 0) break;   // warp-uniform
            if (nvalid >= 16) chunk(std::true_type{}, lc0, col0, 16, orow, b, srs);
            else chunk(std::false_type{}, lc0, col0, nvalid, orow, b, srs);
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) sc_mbar_arrive(&acc_empty[b]);
          if (++b == K.nacc) { b = 0; bph ^= 1u; }
          ++since_flush;
        }
        if (P.stats_kind != S2LC_STATS_NONE && since_flush >= 32) { flush(); since_flush = 0; }
      }
      if (P.stats_kind != S2LC_STATS_NONE && since_flush > 0) flush();
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == SC_MMA_WARP) tmem_dealloc(tmem_base, (uint32_t)K.tmem_cols);
}

static bool sc_aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
static int sc_up(int x, int m) { return (x + m - 1) / m * m; }

// Returns 0 if launched, <0 on error, 1 if this shape is left to the tile-at-a-time engine (igemm_tc.cu).
int32_t launch_sconv_tc(const ConvArgs& A, cudaStream_t st) {
  if (debug_flags() & DBG_NO_STREAM) return 1;   // test hook: compare against the tile-at-a-time engine
  if (A.dtype != S2LC_BF16 || (A.mode != S2LC_CONV_3X3 && A.mode != S2LC_CONV_1X1)) return 1;
  if (A.C % 8 != 0 || A.Kpad % 64 != 0 || !sc_aligned16(A.w_packed)) return 1;
  if (A.a_is_grad) {
    if (A.a_bwd.ld_dy % 8 != 0 || !sc_aligned16(A.a_bwd.dy)) return 1;
    if (A.a_bwd.z && (A.a_bwd.ld_z != A.a_bwd.ld_dy || !sc_aligned16(A.a_bwd.z))) return 1;
  } else {
    if (A.a_fwd.ld % 8 != 0 || !sc_aligned16(A.a_fwd.raw)) return 1;
    if (A.a_fwd.scale && !A.a_fwd.shift) return 1;
  }
  if (A.has_prev && A.prev.scale && !A.prev.shift) return 1;
  if (A.ld_out % 8 != 0 || !sc_aligned16(A.out)) return 1;          // the epilogue moves 16 B vectors
  if (A.has_prev && (A.prev.ld % 8 != 0 || !sc_aligned16(A.prev.raw))) return 1;
  if (A.res1 && (A.ld_res1 % 8 != 0 || !sc_aligned16(A.res1))) return 1;
  if (A.res2 && (A.ld_res2 % 8 != 0 || !sc_aligned16(A.res2))) return 1;
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
  K.slot_bytes = K.Kc * K.chunk_stride;
  const bool has_z = A.a_is_grad && A.a_bwd.z != nullptr;
  K.zslot_bytes = has_z ? K.lines * K.CW * 16 : 0;
  const int R = K.taps == 9 ? 3 : 1;
  const int n16 = sc_up(A.Nout, 16);
  const int budget = 226 * 1024;
  const int fixed = 1024 + 3 * K.Kc * 64 * 4 + 3 * 256 * 4 + 2 * 2 * 256 * 4 + (2 * SC_MAX_RING + 2 * SC_MAX_ACC) * 8 + 64;
  bool found = false;
  for (int ns = 1; ns <= 4 && !found; ++ns) {
    const int ncta = sc_up((n16 + ns - 1) / ns, 16);
    if (ncta > 256) continue;
    const int wb = K.taps * K.Kc * ncta * 128;
    // 1x1 with K > 256: a ring of two 80-96 KB whole-K slots leaves no load in flight behind the tile being multiplied
    // (384->64 at 64x64 ran 2.3x slower than on the tile engine) -> the tile engine takes those; up to K = 256 two
    // slots still beat it (240->40 at 128x128: 0.19 vs 0.25 ms)
    for (int ring = 6; ring >= (K.taps == 9 ? R + 1 : (K.Kc >= 5 ? 3 : 2)); --ring) {
      int pf = ring - R - 1;          // one slot of slack so producers never wait on the tile they just fed
      if (pf > 3) pf = 3;
      if (pf < 1) pf = 1;
      if (fixed + wb + ring * K.slot_bytes + (pf + 1) * K.zslot_bytes <= budget) {
        K.nsplit = ns; K.Ncta = ncta; K.w_bytes = wb; K.ring = ring; K.pf = pf;
        found = true;
        break;
      }
    }
  }
  if (!found) return 1;
  // thin A, wide N (1x1 expand convs and their mirror images in the backward pass): three producer warps keep the
  // ring full (<= 8 chunks of 16 B per thread and slot) and a third epilogue group takes a third of the columns
  K.prod_warps = 7;
  if (K.taps == 1 && K.lines * K.CW * (has_z ? 2 : 1) <= 96 * 11 && K.Ncta >= 48) K.prod_warps = 3;
  // wide A, narrow N (1x1 project convs behind SE gate + SiLU, expand-conv data gradients with the BatchNorm-backward
  // operand): the transform is the long pole and one epilogue group drains <= 4 column chunks per tile
  else if (K.taps == 1 && K.Ncta > 32 && K.Ncta <= 64 && K.CW * (has_z ? 2 : 1) >= 24) K.prod_warps = 11;
  if (debug_flags() & DBG_SCONV_PROD7) K.prod_warps = 7;   // A/B hook for the profiles
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
  K.KS = 0;
  for (int kc = 0; kc < K.Kc; ++kc) {
    int rem = A.C - kc * 64;
    rem = rem > 64 ? 64 : rem;
    for (int k4 = 0; k4 < (rem + 15) / 16; ++k4, ++K.KS) {
      K.koff_a[K.KS] = (uint32_t)((kc * K.chunk_stride + k4 * 32) >> 4);
      K.koff_b[K.KS] = (uint32_t)((kc * K.Ncta * 128 + k4 * 32) >> 4);
    }
  }
  size_t smem = (size_t)fixed + K.w_bytes + (size_t)K.ring * K.slot_bytes + (size_t)(K.pf + 1) * K.zslot_bytes;
  if (K.prod_warps == 3) {                       // third statistics area
    if (smem + 2 * 256 * 4 > (size_t)budget) K.prod_warps = 7;
    else smem += 2 * 256 * 4;
  }
  const bool regstat = K.Ncta <= 32;
  void (*kern)(ConvArgs, SPlan) =
      K.taps == 9 ? (regstat ? sconv_kernel<9, true> : sconv_kernel<9, false>)
                  : (regstat ? sconv_kernel<1, true> : sconv_kernel<1, false>);
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaSuccess;
    void (*all[4])(ConvArgs, SPlan) = {sconv_kernel<9, true>, sconv_kernel<9, false>, sconv_kernel<1, true>,
                                       sconv_kernel<1, false>};
    for (int i = 0; i < 4 && e == cudaSuccess; ++i)
      e = cudaFuncSetAttribute(all[i], cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) {
      set_error("sconv_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
      return S2LC_ERR_CUDA;
    }
    attr_set = true;
  }
  PDL_LAUNCH(kern, dim3(ctas * K.nsplit), SC_THREADS, smem, st, A, K);
  return check_launch("sconv_tc");
}

}  // namespace s2lc

// tconv_tc.cu — TMA-fed, warp-specialised, persistent tcgen05 implicit-GEMM convolution for MATERIALISED operands:
// the compute-heavy contractions (decoder 3x3 convs and their data gradients at <= 128 px wide rows, deep 1x1
// expand / project / head convs and their data gradients, 2x2 transposed convs).
// Reference call sites: src/modules/efficientnet_unet.py:168-176 (_double_conv), :115-119 (ConvTranspose2d),
// :312-327 / :363-375 (expand / project), :229-238 (conv_head), and the autograd data gradients of all of them.
//
//   out[m, n] = sum_{tap, c} A[pixel(m) + tap, c] * Wp[n, tap*Cpad + c]
//
// The A operand is a stored bf16 NHWC tensor (an activation written once by s2lc_bn_apply, a concat buffer, or a
// gradient written once by s2lc_lazygrad_apply), so no transform sits between HBM/L2 and the tensor core and the
// whole operand path is TMA:
//   * A: one 4-D tensor map (C, W, H, N); the tile of tap (dy, dx), K chunk kc is the BOX (64 ch, bw px, bh rows) at
//        (kc*64, x0+dx-1, y0+dy-1, n) — im2col by shifted boxes, zero padding = TMA out-of-bounds fill; 1x1 and
//        transposed convs use a 2-D map (C, M).  SWIZZLE_128B, so the box lands as the K-major tile UMMA reads.
//   * B: one 2-D map over the packed weights [Npad, Kpad], box (64, BN).
// Roles (14 warps, one CTA per SM, static tile schedule): warp 0 = TMA producer (one lane), warp 1 = MMA issuer (one
// lane; owns TMEM), warps 2-13 = epilogue (three warps per TMEM lane quarter, interleaved 16-column chunks): tcgen05.ld ->
// bias, residuals, act'(prev), bf16 16 B stores (incl. the 2x2 scatter of the transposed conv), BatchNorm moments.
// Two TMEM accumulators: the epilogue of tile i overlaps the main loop of tile i+1.
#include <cuda.h>

#include <cstring>
#include <type_traits>

#include "conv_common.cuh"
#include "stream_common.cuh"

namespace s2lc {

using namespace tc;

constexpr int TV_EPI_WARPS = 12;
constexpr int TV_WARPS = 2 + TV_EPI_WARPS;
constexpr int TV_THREADS = TV_WARPS * 32;
constexpr int TV_MAX_STAGES = 8;
constexpr int TV_A_BYTES = 128 * 128;   // 128 pixels x 64 channels bf16

struct TPlan {
  int taps;          // 1, 9 (3x3) or 4 (data gradient of a 2x2 stride-2 transposed conv)
  int boxed;         // 1: a tile is a pixel BOX (bw x bh) of one sample (3x3, ConvT data gradient); 0: 128 consecutive pixels
  int KC;            // 64-channel chunks per tap
  int ksteps_last;   // K = 16 steps in the last chunk of a tap (1..4)
  int BN;            // output columns per tile (multiple of 16, <= 256)
  int n_tiles, m_tiles, total_tiles;
  int bw, bh, bw_shift;       // 3x3: pixel box of a tile (bw * bh = 128, bw a power of two)
  int tiles_x, tiles_y;
  int stages, stage_bytes;
  int tmem_cols;
  int epi_tma;       // 1: epilogue through a shared-memory staging tile, TMA bulk stores and a column pass for the moments
  int stage_out;     // bytes of the staging tile (ceil(BN / 64) boxes of 128 rows x 128 B)
};

// ---- TMA / mbarrier transaction wrappers ---------------------------------------------------------------------
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2,
                                            int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], "
      "[%2];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tma_load_5d(uint32_t dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2,
                                            int c3, int c4) {
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, "
      "%7}], [%2];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
      : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}

struct TTile {
  int mt, nt;        // tile coordinates
  int n, y0, x0;     // 3x3: sample and first pixel of the box
  int64_t m0;        // 1x1: first pixel
};
__device__ __forceinline__ TTile tile_of(const ConvArgs& P, const TPlan& K, int t) {
  TTile T;
  T.mt = t / K.n_tiles;
  T.nt = t - T.mt * K.n_tiles;
  T.n = 0; T.y0 = 0; T.x0 = 0;
  T.m0 = (int64_t)T.mt * 128;
  if (K.boxed) {
    const int per = K.tiles_x * K.tiles_y;
    T.n = T.mt / per;
    const int r = T.mt - T.n * per;
    const int ty = r / K.tiles_x;
    T.y0 = ty * K.bh;
    T.x0 = (r - ty * K.tiles_x) * K.bw;
  }
  return T;
}

__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, uint32_t src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.tile.bulk_group [%0, {%2, %3}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(map)), "r"(src), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_store_4d(const CUtensorMap* map, uint32_t src, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(map)), "r"(src), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

__global__ void __launch_bounds__(TV_THREADS, 1) tconv_kernel(ConvArgs P, TPlan K, const __grid_constant__ CUtensorMap tmA,
                                                              const __grid_constant__ CUtensorMap tmB,
                                                              const __grid_constant__ CUtensorMap tmO) {
  cudaGridDependencySynchronize();   // programmatic dependent launch (common.cuh)
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* smem =
      reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  // per-channel statistics of ALL column tiles this CTA visits, [2][n_tiles * BN] floats, flushed once at the end
  // (a flush per tile is two fp64 atomics per column from every CTA on the same addresses: on the thin layers —
  // 16 k tiles of 128 x 144 — the L2 atomic unit, not the tensor core, set the pace)
  unsigned char* ostage = smem + K.stages * K.stage_bytes;     // staging tile of the TMA-store epilogue (1024 B aligned)
  float* sstat = reinterpret_cast<float*>(ostage + K.stage_out);
  const int NS = K.n_tiles * K.BN;
  uint64_t* full = reinterpret_cast<uint64_t*>(sstat + 2 * NS);
  uint64_t* empty = full + TV_MAX_STAGES;
  uint64_t* acc_full = empty + TV_MAX_STAGES;
  uint64_t* acc_empty = acc_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t M = (int64_t)P.N * P.H * P.W;

  if (tid == 0) {
    for (int s = 0; s < K.stages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(&acc_full[b], 1); mbar_init(&acc_empty[b], TV_EPI_WARPS); }
    fence_barrier_init();
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1) tmem_alloc(tmem_slot, (uint32_t)K.tmem_cols);
  for (int i = tid; i < 2 * NS; i += TV_THREADS) sstat[i] = 0.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t smem_base = smem_u32(smem);
  const int chunks = K.taps * K.KC;                 // K chunks (pipeline steps) per tile

  if (warp == 0) {
    // ================================================================== TMA producer
    if (elect_one()) {
      int s = 0;
      uint32_t ph = 0;
      const uint32_t bytes = (uint32_t)K.stage_bytes;
      for (int t = blockIdx.x; t < K.total_tiles; t += gridDim.x) {
        const TTile T = tile_of(P, K, t);
        const int n0 = T.nt * K.BN;
        for (int tap = 0; tap < K.taps; ++tap) {
          const int dy = K.taps == 9 ? tap / 3 - 1 : 0, dx = K.taps == 9 ? tap - (tap / 3) * 3 - 1 : 0;
          for (int kc = 0; kc < K.KC; ++kc) {
            mbar_wait(&empty[s], ph ^ 1u);
            mbar_expect_tx(&full[s], bytes);
            const uint32_t sa = smem_base + (uint32_t)(s * K.stage_bytes);
            if (K.taps == 9) tma_load_4d(sa, &tmA, &full[s], kc * 64, T.x0 + dx, T.y0 + dy, T.n);
            else if (K.taps == 4)   // child (qy, qx) = tap of the 2H x 2W gradient, viewed as (c, qx, w, qy, n*H + h)
              tma_load_5d(sa, &tmA, &full[s], kc * 64, tap & 1, T.x0, tap >> 1, T.n * P.H + T.y0);
            else tma_load_2d(sa, &tmA, &full[s], kc * 64, (int)T.m0);
            tma_load_2d(sa + TV_A_BYTES, &tmB, &full[s], tap * P.Cpad + kc * 64, n0);
            if (++s == K.stages) { s = 0; ph ^= 1u; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================================================================== MMA issuer (converged warp, one lane issues)
    const bool leader = elect_one();
    const uint32_t idesc = idesc_bf16(128, (uint32_t)K.BN, 0, 0);
    const uint64_t dbase = smem_desc_sw128(0, 1024, 16);
    int s = 0;
    uint32_t ph = 0;
    int b = 0;
    uint32_t bph = 0;
    for (int t = blockIdx.x; t < K.total_tiles; t += gridDim.x) {
      mbar_wait(&acc_empty[b], bph ^ 1u);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)(b * K.BN);
      uint32_t accum = 0;
      for (int c = 0; c < chunks; ++c) {
        const int kc = c % K.KC;
        const int ks = (kc == K.KC - 1) ? K.ksteps_last : 4;
        mbar_wait(&full[s], ph);
        tc_fence_after();
        const uint32_t sa = smem_base + (uint32_t)(s * K.stage_bytes);
        const uint64_t da = dbase + (uint64_t)(sa >> 4), db = dbase + (uint64_t)((sa + TV_A_BYTES) >> 4);
        if (leader) {
#pragma unroll 4
          for (int k4 = 0; k4 < ks; ++k4) {
            umma_bf16(d_tmem, da + (uint64_t)(k4 * 2), db + (uint64_t)(k4 * 2), idesc, accum);
            accum = 1;
          }
          umma_commit(&empty[s]);
          if (c == chunks - 1) umma_commit(&acc_full[b]);
        }
        __syncwarp();
        if (++s == K.stages) { s = 0; ph ^= 1u; }
      }
      if (++b == 2) { b = 0; bph ^= 1u; }
    }
  } else {
    // ================================================================== epilogue
    constexpr int NSUB = TV_EPI_WARPS / 4;                      // warps per TMEM lane quarter
    const int ew = warp - 2, q = warp & 3, sub = ew >> 2;       // q = TMEM lane quarter this warp may read
    const int r = q * 32 + lane;                                // accumulator row = pixel of the tile
    bf16* outp = reinterpret_cast<bf16*>(P.out);
    const bf16* prevp = P.has_prev ? reinterpret_cast<const bf16*>(P.prev.raw) : nullptr;
    const bf16* res1p = reinterpret_cast<const bf16*>(P.res1);
    const bf16* res2p = reinterpret_cast<const bf16*>(P.res2);
    const int prev_act = P.has_prev ? P.prev.act : S2LC_ACT_NONE;
    const bool prev_aff = P.has_prev && P.prev.scale != nullptr;
    const int stats_kind = P.stats_kind;
    const int nchunks = K.BN / 16;
    const int hw = P.H * P.W;
    // global operands of one 16-column chunk of this thread's row, fetched ONE CHUNK AHEAD of their use (and, for the
    // first chunk of a tile, before the wait for the accumulator): a 16 B load per row and tensor is all latency
    struct Ops {
      uint4 pz0, pz1, ra0, ra1, rb0, rb1;
      int64_t orow;
      int col0, nvalid;
    };
    int b = 0;
    uint32_t bph = 0;
    if (K.epi_tma) {
      // ---------------------------------------------------------------- TMA-store epilogue (plain forward convs and
      // data gradients without residual / act' operands).  Per 16-column chunk a thread only adds the bias, rounds and
      // writes 2 x 16 B into the swizzled staging tile; one lane then stores the tile with bulk tensor copies (whole
      // 128 B lines, rows / columns outside the tensor are clipped by the copy) and the BatchNorm moments come from a
      // column pass over the staged bf16 values: ~5 instructions per element pair instead of two 16-lane butterflies
      // per chunk.  The thin-K layers (expand convs, project-conv data gradients) are bound by exactly this code.
      const int etid = ew * 32 + lane;
      constexpr int NE = TV_EPI_WARPS * 32;
      const uint32_t ost = smem_u32(ostage);
      const int boxes = (K.BN + 63) >> 6;
      const int CP = K.BN >> 1;                                   // column pairs of the tile
      const int G = NE / CP > 0 ? NE / CP : 1;                    // row groups of the column pass
      const uint32_t rsw = (uint32_t)(r & 7);
      for (int t = blockIdx.x; t < K.total_tiles; t += gridDim.x) {
        const TTile T = tile_of(P, K, t);
        const int n0 = T.nt * K.BN;
        int rows_valid = 128;
        if (!K.boxed) {
          const int64_t left = M - T.m0;
          rows_valid = left < 128 ? (int)left : 128;
        }
        mbar_wait(&acc_full[b], bph);
        tc_fence_after();
        // the previous tile's bulk stores must have finished READING the staging tile
        if (etid == 0) tma_store_wait_read();
        asm volatile("bar.sync 1, %0;" ::"n"(NE) : "memory");
        for (int cj = sub; cj < nchunks; cj += NSUB) {
          const int col0 = n0 + cj * 16;
          const int nvalid = P.Nout - col0;
          if (nvalid <= 0) break;
          float acc[16];
          tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(b * K.BN + cj * 16), acc);
          if (P.bias) {
            if (nvalid >= 16) {
              const float4* bv = reinterpret_cast<const float4*>(P.bias + col0);
#pragma unroll
              for (int j4 = 0; j4 < 4; ++j4) {
                const float4 tb = __ldg(bv + j4);
                acc[4 * j4] += tb.x; acc[4 * j4 + 1] += tb.y; acc[4 * j4 + 2] += tb.z; acc[4 * j4 + 3] += tb.w;
              }
            } else {
#pragma unroll
              for (int j = 0; j < 16; ++j)
                if (j < nvalid) acc[j] += __ldg(P.bias + col0 + j);
            }
          }
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
          // tile column cj*16 .. +15 = box cj/4, 16 B chunks 2*(cj%4), +1 of row r's 128 B line (SWIZZLE_128B)
          const uint32_t line = ost + (uint32_t)((cj >> 2) * 16384) + (uint32_t)(r >> 3) * 1024u + rsw * 128u;
          const uint32_t ch = (uint32_t)((cj & 3) * 2);
          sts16(line + (((ch) ^ rsw) << 4), o0);
          sts16(line + (((ch + 1) ^ rsw) << 4), o1);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) sc_mbar_arrive(&acc_empty[b]);
        fence_proxy_async();
        asm volatile("bar.sync 1, %0;" ::"n"(NE) : "memory");
        if (etid == 0) {
          for (int j = 0; j < boxes; ++j) {
            if (n0 + j * 64 >= P.Nout) break;
            if (K.boxed) tma_store_4d(&tmO, ost + (uint32_t)(j * 16384), n0 + j * 64, T.x0, T.y0, T.n);
            else tma_store_2d(&tmO, ost + (uint32_t)(j * 16384), n0 + j * 64, (int)T.m0);
          }
          tma_store_commit();
        }
        if (stats_kind == S2LC_STATS_MOMENTS) {
          const int cp = etid % CP, g = etid / CP;
          if (g < G && n0 + 2 * cp < P.Nout) {
            // word (row rr, column pair cp): box cp/32, 16 B chunk (cp%32)/4, word cp%4
            const uint32_t base = ost + (uint32_t)((cp >> 5) * 16384) + (uint32_t)((cp & 3) * 4);
            const uint32_t c16 = (uint32_t)((cp & 31) >> 2);
            float2 a0 = make_float2(0.f, 0.f), a1 = make_float2(0.f, 0.f);
            for (int rr = g; rr < rows_valid; rr += G) {
              uint32_t w;
              asm volatile("ld.shared.b32 %0, [%1];" : "=r"(w)
                           : "r"(base + (uint32_t)(rr >> 3) * 1024u + (uint32_t)(rr & 7) * 128u + ((c16 ^ (uint32_t)(rr & 7)) << 4)));
              const float x0 = __uint_as_float(w << 16), x1 = __uint_as_float(w & 0xffff0000u);
              a0.x += x0; a0.y += x1;
              a1.x = fmaf(x0, x0, a1.x); a1.y = fmaf(x1, x1, a1.y);
            }
            float* gs = sstat + n0 + 2 * cp;
            atomicAdd(gs, a0.x); atomicAdd(gs + 1, a0.y);
            atomicAdd(gs + NS, a1.x); atomicAdd(gs + NS + 1, a1.y);
          }
        }
        if (++b == 2) { b = 0; bph ^= 1u; }
      }
      if (etid == 0) tma_store_wait_all();
    } else
    for (int t = blockIdx.x; t < K.total_tiles; t += gridDim.x) {
      const TTile T = tile_of(P, K, t);
      const int n0 = T.nt * K.BN;
      // this thread's output pixel
      int64_t m;
      int sn, ph_, pw_;
      bool valid = true;
      if (K.boxed) {
        ph_ = T.y0 + (r >> K.bw_shift);
        pw_ = T.x0 + (r & (K.bw - 1));
        sn = T.n;
        m = ((int64_t)sn * P.H + ph_) * P.W + pw_;
      } else {
        m = T.m0 + r;
        valid = m < M;
        const int64_t mm = valid ? m : 0;
        sn = (int)(mm / hw);
        const int rem = (int)(mm - (int64_t)sn * hw);
        ph_ = rem / P.W;
        pw_ = rem - ph_ * P.W;
      }
      auto fetch = [&](int cj, Ops& o) {
        const int ncol0 = n0 + cj * 16;
        int nv = P.Nout - ncol0;
        o.nvalid = nv > 16 ? 16 : nv;
        o.col0 = ncol0;
        o.orow = m;
        if (P.mode == S2LC_CONVT_FWD) {
          const int qq = ncol0 / P.Cout;
          o.col0 = ncol0 - qq * P.Cout;
          o.orow = ((int64_t)sn * (2 * P.H) + 2 * ph_ + (qq >> 1)) * (2 * P.W) + 2 * pw_ + (qq & 1);
        }
        o.pz0 = make_uint4(0, 0, 0, 0);
        o.pz1 = o.pz0; o.ra0 = o.pz0; o.ra1 = o.pz0; o.rb0 = o.pz0; o.rb1 = o.pz0;
        if (valid && o.nvalid == 16) {
          if (prevp) {
            const uint4* p = reinterpret_cast<const uint4*>(prevp + o.orow * P.prev.ld + o.col0);
            o.pz0 = __ldg(p); o.pz1 = __ldg(p + 1);
          }
          if (res1p) {
            const uint4* p = reinterpret_cast<const uint4*>(res1p + o.orow * P.ld_res1 + o.col0);
            o.ra0 = __ldg(p); o.ra1 = __ldg(p + 1);
          }
          if (res2p) {
            const uint4* p = reinterpret_cast<const uint4*>(res2p + o.orow * P.ld_res2 + o.col0);
            o.rb0 = __ldg(p); o.rb1 = __ldg(p + 1);
          }
        }
      };
      const float srs = (stats_kind == S2LC_STATS_BNBWD && P.stats_rowscale) ? __ldg(P.stats_rowscale + sn) : 1.f;
      float* gstat = sstat + n0;
      Ops cur, nxt;
      if (sub < nchunks) fetch(sub, cur);
      mbar_wait(&acc_full[b], bph);
      tc_fence_after();
      for (int cj = sub; cj < nchunks; cj += NSUB) {
        if (cur.nvalid <= 0) break;                             // warp-uniform: columns beyond Nout
        if (cj + NSUB < nchunks) fetch(cj + NSUB, nxt);
        const int nvalid = cur.nvalid, col0 = cur.col0;
        const int64_t orow = cur.orow;
        const bool fullc = nvalid == 16;
        float acc[16];
        tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(b * K.BN + cj * 16), acc);
        float zp[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) zp[j] = 0.f;
        if (P.bias) {
          if (fullc) {
            const float4* bv = reinterpret_cast<const float4*>(P.bias + col0);
#pragma unroll
            for (int j4 = 0; j4 < 4; ++j4) {
              const float4 tb = __ldg(bv + j4);
              acc[4 * j4] += tb.x; acc[4 * j4 + 1] += tb.y; acc[4 * j4 + 2] += tb.z; acc[4 * j4 + 3] += tb.w;
            }
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (j < nvalid) acc[j] += __ldg(P.bias + col0 + j);
          }
        }
        if (valid) {
          if (fullc) {
            float e[8];
            if (res1p) {
              sc_unpack8(cur.ra0, e);
#pragma unroll
              for (int j = 0; j < 8; ++j) acc[j] += e[j];
              sc_unpack8(cur.ra1, e);
#pragma unroll
              for (int j = 0; j < 8; ++j) acc[8 + j] += e[j];
            }
            if (res2p) {
              sc_unpack8(cur.rb0, e);
#pragma unroll
              for (int j = 0; j < 8; ++j) acc[j] += e[j];
              sc_unpack8(cur.rb1, e);
#pragma unroll
              for (int j = 0; j < 8; ++j) acc[8 + j] += e[j];
            }
            if (prevp) {
              sc_unpack8(cur.pz0, *reinterpret_cast<float(*)[8]>(&zp[0]));
              sc_unpack8(cur.pz1, *reinterpret_cast<float(*)[8]>(&zp[8]));
            }
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              if (j < nvalid) {
                if (res1p) acc[j] += __bfloat162float(res1p[orow * P.ld_res1 + col0 + j]);
                if (res2p) acc[j] += __bfloat162float(res2p[orow * P.ld_res2 + col0 + j]);
                if (prevp) zp[j] = __bfloat162float(prevp[orow * P.prev.ld + col0 + j]);
              }
            }
          }
          if (prev_act != S2LC_ACT_NONE) {
            float u[16];
            if (prev_aff && fullc) {
              const float4* sv = reinterpret_cast<const float4*>(P.prev.scale + col0);
              const float4* hv = reinterpret_cast<const float4*>(P.prev.shift + col0);
#pragma unroll
              for (int j4 = 0; j4 < 4; ++j4) {
                const float4 a4 = __ldg(sv + j4), b4 = __ldg(hv + j4);
                u[4 * j4] = fmaf(zp[4 * j4], a4.x, b4.x);
                u[4 * j4 + 1] = fmaf(zp[4 * j4 + 1], a4.y, b4.y);
                u[4 * j4 + 2] = fmaf(zp[4 * j4 + 2], a4.z, b4.z);
                u[4 * j4 + 3] = fmaf(zp[4 * j4 + 3], a4.w, b4.w);
              }
            } else {
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                u[j] = zp[j];
                if (prev_aff && j < nvalid) u[j] = fmaf(zp[j], __ldg(P.prev.scale + col0 + j), __ldg(P.prev.shift + col0 + j));
              }
            }
            if (prev_act == S2LC_ACT_RELU) {
#pragma unroll
              for (int j = 0; j < 16; ++j) acc[j] = u[j] > 0.f ? acc[j] : 0.f;
            } else {
#pragma unroll
              for (int j = 0; j < 16; ++j) acc[j] *= act_bwd_fast(u[j], S2LC_ACT_SILU);
            }
          }
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
        if (valid) {
          bf16* op = outp + orow * P.ld_out + col0;
          if (fullc) {
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
        }
        if (stats_kind != S2LC_STATS_NONE) {                    // warp-uniform
          float v[16], s0[16], s1[16];
          sc_unpack8(o0, *reinterpret_cast<float(*)[8]>(&v[0]));
          sc_unpack8(o1, *reinterpret_cast<float(*)[8]>(&v[8]));
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const float x = (valid && j < nvalid) ? v[j] * srs : 0.f;
            s0[j] = x;
            s1[j] = stats_kind == S2LC_STATS_MOMENTS ? x * x : x * zp[j];
          }
          int col;
          const float t0 = sc_butterfly16(s0, lane, col);
          const float t1 = sc_butterfly16(s1, lane, col);
          if ((lane & 1) == 0) {
            atomicAdd(&gstat[cj * 16 + col], t0);
            atomicAdd(&gstat[NS + cj * 16 + col], t1);
          }
        }
        cur = nxt;
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) sc_mbar_arrive(&acc_empty[b]);
      if (++b == 2) { b = 0; bph ^= 1u; }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, (uint32_t)K.tmem_cols);
  if (P.stats_kind != S2LC_STATS_NONE) {
    for (int i = tid; i < NS; i += TV_THREADS) {
      const float v0 = sstat[i], v1 = sstat[NS + i];
      if (i < P.Nout && (v0 != 0.f || v1 != 0.f)) {
        atomicAdd(&P.stats[i], (double)v0);
        atomicAdd(&P.stats[P.Nout + i], (double)v1);
      }
    }
  }
}

// ---- host side --------------------------------------------------------------------------------------------------
typedef CUresult (*TmapEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                 const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
// cuTensorMapEncodeTiled through the runtime's driver entry point query: libs2lc.so keeps no link-time dependency
// on libcuda (it must load on a box without a driver, where only the symbol table is checked).
static TmapEncodeFn tmap_encode_fn() {
  static TmapEncodeFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<TmapEncodeFn>(p);
  }
  return fn;
}

static bool tv_aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// bf16 tensor map, SWIZZLE_128B, zero fill outside the tensor.  dims / box: fastest dimension first; strides in
// ELEMENTS for dimensions 1..rank-1.
static bool tv_make_map(CUtensorMap* map, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_el,
                        const uint32_t* box) {
  TmapEncodeFn enc = tmap_encode_fn();
  if (!enc) return false;
  cuuint64_t gdim[5], gstr[4];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) { gdim[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
  for (int i = 0; i + 1 < rank; ++i) gstr[i] = strides_el[i] * 2;
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), gdim, gstr, bx, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

// Returns 0 if launched, <0 on error, 1 if this shape / operand is left to the other engines.
int32_t launch_tconv_tc(const ConvArgs& A, cudaStream_t st) {
  const uint32_t dbg = debug_flags();
  if (dbg & (DBG_NO_TMA | DBG_NO_STREAM)) return 1;
  if (A.dtype != S2LC_BF16) return 1;
  if (A.mode != S2LC_CONV_1X1 && A.mode != S2LC_CONV_3X3 && A.mode != S2LC_CONVT_FWD && A.mode != S2LC_CONVT_DGRAD) return 1;
  // PLAIN operand only: a stored tensor, nothing to apply on the way to the tensor core
  const void* a_ptr;
  int64_t a_ld;
  if (A.a_is_grad) {
    if (A.a_bwd.z || A.a_bwd.rowscale) return 1;
    a_ptr = A.a_bwd.dy; a_ld = A.a_bwd.ld_dy;
  } else {
    if (A.a_fwd.scale || A.a_fwd.gate || A.a_fwd.act != S2LC_ACT_NONE) return 1;
    a_ptr = A.a_fwd.raw; a_ld = A.a_fwd.ld;
  }
  if (A.C % 8 != 0 || a_ld % 8 != 0 || !tv_aligned16(a_ptr) || A.Kpad % 8 != 0 || !tv_aligned16(A.w_packed)) return 1;
  if (A.ld_out % 8 != 0 || !tv_aligned16(A.out)) return 1;
  if (A.has_prev && (A.prev.ld % 8 != 0 || !tv_aligned16(A.prev.raw) || (A.prev.scale && !A.prev.shift))) return 1;
  if (A.res1 && (A.ld_res1 % 8 != 0 || !tv_aligned16(A.res1))) return 1;
  if (A.res2 && (A.ld_res2 % 8 != 0 || !tv_aligned16(A.res2))) return 1;
  if (A.mode == S2LC_CONVT_FWD && A.Cout % 16 != 0) return 1;
  const int64_t M = (int64_t)A.N * A.H * A.W;
  if (M >= (1ll << 31)) return 1;
  TPlan K{};
  K.taps = A.mode == S2LC_CONV_3X3 ? 9 : (A.mode == S2LC_CONVT_DGRAD ? 4 : 1);
  K.boxed = K.taps != 1;
  K.KC = (A.C + 63) / 64;
  K.ksteps_last = (A.C - (K.KC - 1) * 64 + 15) / 16;
  // Which plain shapes this engine takes (tools/tma_sweep.sh, profiles/r02b_tma_sweep.txt): 3x3 layers up to 128 px
  // wide rows, and 256 px rows with more than one 64-channel chunk (wider / thinner ones are HBM-bound and belong to
  // the rolling-row engine, sconv_tc.cu); 1x1 and transposed convs from 176 / 128 input channels (below that the
  // tile is all epilogue and the 12-epilogue-warp streaming engine is as fast or faster).
  const bool take_all = (dbg & DBG_TMA_ALL) != 0;
  // epilogue without per-row global operands -> shared-memory staging + TMA bulk stores (see the kernel)
  const bool epi_ok = A.mode != S2LC_CONVT_FWD && !A.has_prev && !A.res1 && !A.res2 && A.Nout % 8 == 0 &&
                      (A.stats_kind == S2LC_STATS_NONE || A.stats_kind == S2LC_STATS_MOMENTS) && !(dbg & DBG_NO_TMA_STORE);
  if (!take_all) {
    // thin-K 1x1 layers (expand convs) are all epilogue: with the TMA-store epilogue this engine writes 2.2-2.5 TB/s
    // where the streaming engine reaches 1.1-1.5 (profiles/r02g_tma_sweep.txt); without it, or with N < 64, it loses
    if (A.mode == S2LC_CONV_1X1 && A.C < 176 && !(epi_ok && A.Nout >= 64)) return 1;
    if (A.mode == S2LC_CONVT_FWD && A.C < 128) return 1;
  }
  if (K.boxed) {
    if (K.taps == 9 && !take_all && !(A.W <= 128 || (A.W == 256 && A.C > 64))) return 1;
    K.bw = A.W < 128 ? A.W : 128;
    if ((K.bw & (K.bw - 1)) != 0 || K.bw < 8 || A.W % K.bw != 0) return 1;
    K.bh = 128 / K.bw;
    if (A.H % K.bh != 0) return 1;
    K.bw_shift = 0;
    while ((1 << K.bw_shift) < K.bw) ++K.bw_shift;
    K.tiles_x = A.W / K.bw;
    K.tiles_y = A.H / K.bh;
    K.m_tiles = A.N * K.tiles_x * K.tiles_y;
  } else {
    K.bw = 128; K.bh = 1; K.bw_shift = 7; K.tiles_x = K.tiles_y = 1;
    K.m_tiles = (int)((M + 127) / 128);
  }
  const int n16 = (A.Nout + 15) / 16 * 16;
  int nt = (n16 + 255) / 256;
  // fewer, wider column tiles reuse the A tile better; but a launch that cannot fill the SMs once is cut finer
  while ((int64_t)K.m_tiles * nt < kNumSMs && (n16 + nt) / (nt + 1) >= 64) ++nt;
  K.BN = ((n16 + nt - 1) / nt + 15) / 16 * 16;
  // TMA-store epilogue: forward convs / data gradients whose epilogue has no per-row global operand (no residual, no
  // act'(prev)), bf16 rows of whole 16 B vectors.  With several column tiles the tile width must be a multiple of the
  // 64-column store box (a partial last box would spill into the next tile's columns); the copy clips at Nout.
  const int chunks_k = K.taps * K.KC;
  if (epi_ok) {
    // the staging tile takes shared memory from the operand pipeline: keep >= min(4, k chunks + 1) stages
    const int need = chunks_k + 1 < 4 ? chunks_k + 1 : 4;
    const int cand[4] = {nt > 1 ? (K.BN + 63) / 64 * 64 : K.BN, 192, 128, 64};
    // narrower tiles only for thin-K layers (all epilogue); compute-bound layers keep their 256-wide tiles
    for (int i = 0; i < ((nt > 1 && chunks_k <= 6) ? 4 : 1); ++i) {
      const int bn = cand[i];
      if (bn > 256 || (i > 0 && bn >= cand[0])) continue;
      const int so = (bn + 63) / 64 * 16384;
      const int nt_e = (A.Nout + bn - 1) / bn;
      const int fixed_e = 1024 + 2 * nt_e * bn * 4 + (2 * TV_MAX_STAGES + 4) * 8 + 64;
      const int st_e = (226 * 1024 - fixed_e - so) / (TV_A_BYTES + bn * 128);
      if (st_e >= need && st_e >= 2) { K.epi_tma = 1; K.stage_out = so; K.BN = bn; break; }
    }
  }
  K.n_tiles = (A.Nout + K.BN - 1) / K.BN;
  K.total_tiles = K.m_tiles * K.n_tiles;
  K.stage_bytes = TV_A_BYTES + K.BN * 128;
  const int fixed = 1024 + 2 * K.n_tiles * K.BN * 4 + (2 * TV_MAX_STAGES + 4) * 8 + 64;
  K.stages = (226 * 1024 - fixed - K.stage_out) / K.stage_bytes;
  if (K.stages > TV_MAX_STAGES) K.stages = TV_MAX_STAGES;
  if (K.stages < 2) return 1;
  K.tmem_cols = 32;
  while (K.tmem_cols < 2 * K.BN) K.tmem_cols *= 2;

  CUtensorMap tmA, tmB, tmO;
  memset(&tmO, 0, sizeof(tmO));
  bool ok;
  if (K.taps == 9) {
    const uint64_t dims[4] = {(uint64_t)A.C, (uint64_t)A.W, (uint64_t)A.H, (uint64_t)A.N};
    const uint64_t str[3] = {(uint64_t)a_ld, (uint64_t)A.W * a_ld, (uint64_t)A.H * A.W * a_ld};
    const uint32_t box[4] = {64, (uint32_t)K.bw, (uint32_t)K.bh, 1};
    ok = tv_make_map(&tmA, a_ptr, 4, dims, str, box);
  } else if (K.taps == 4) {
    // the gradient lives at 2H x 2W: view (c, qx, w, qy, n*H + h), a tap is the phase (qy, qx)
    const uint64_t W = (uint64_t)A.W, ld = (uint64_t)a_ld;
    const uint64_t dims[5] = {(uint64_t)A.C, 2, W, 2, (uint64_t)A.N * A.H};
    const uint64_t str[4] = {ld, 2 * ld, 2 * W * ld, 4 * W * ld};
    const uint32_t box[5] = {64, 1, (uint32_t)K.bw, 1, (uint32_t)K.bh};
    ok = tv_make_map(&tmA, a_ptr, 5, dims, str, box);
  } else {
    const uint64_t dims[2] = {(uint64_t)A.C, (uint64_t)M};
    const uint64_t str[1] = {(uint64_t)a_ld};
    const uint32_t box[2] = {64, 128};
    ok = tv_make_map(&tmA, a_ptr, 2, dims, str, box);
  }
  {
    const uint64_t dims[2] = {(uint64_t)A.Kpad, (uint64_t)A.Npad};
    const uint64_t str[1] = {(uint64_t)A.Kpad};
    const uint32_t box[2] = {64, (uint32_t)K.BN};
    ok = ok && tv_make_map(&tmB, A.w_packed, 2, dims, str, box);
  }
  if (ok && K.epi_tma) {
    if (K.boxed) {
      const uint64_t dims[4] = {(uint64_t)A.Nout, (uint64_t)A.W, (uint64_t)A.H, (uint64_t)A.N};
      const uint64_t str[3] = {(uint64_t)A.ld_out, (uint64_t)A.W * A.ld_out, (uint64_t)A.H * A.W * A.ld_out};
      const uint32_t box[4] = {64, (uint32_t)K.bw, (uint32_t)K.bh, 1};
      ok = tv_make_map(&tmO, A.out, 4, dims, str, box);
    } else {
      const uint64_t dims[2] = {(uint64_t)A.Nout, (uint64_t)M};
      const uint64_t str[1] = {(uint64_t)A.ld_out};
      const uint32_t box[2] = {64, 128};
      ok = tv_make_map(&tmO, A.out, 2, dims, str, box);
    }
  }
  if (!ok) return 1;   // no driver entry point / encoding refused: the other engines take the shape
  const size_t smem = (size_t)fixed + (size_t)K.stage_out + (size_t)K.stages * K.stage_bytes;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(tconv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) {
      set_error("tconv_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
      return S2LC_ERR_CUDA;
    }
    attr_set = true;
  }
  int grid = kNumSMs < K.total_tiles ? kNumSMs : K.total_tiles;
  PDL_LAUNCH(tconv_kernel, dim3(grid), TV_THREADS, smem, st, A, K, tmA, tmB, tmO);
  note_tma_launch();
  return check_launch("tconv_tc");
}

}  // namespace s2lc

// twgrad_tc.cu — TMA-fed, warp-specialised, persistent tcgen05 WEIGHT-GRADIENT engine for materialised operands on the
// deep (compute-bound, L2-resident) layers:
//
//   dW[co, tap, ci] = sum over pixels p  G[p, co] * A[p + tap, ci]
//
// (reference: autograd of nn.Conv2d(k=3, p=1) at src/modules/efficientnet_unet.py:168-176, of the 1x1 expand / project /
//  head convs at :312-327, :363-375, :229-238 and of nn.ConvTranspose2d(k=2, s=2) at :115-119.)
//
// The reduction runs over PIXELS, so both tensor-core operands are MN-major: a pixel is one 128-byte line of a
// SWIZZLE_128B tile holding 64 channels — exactly what a TMA box (64 channels x bw x bh pixels) writes.  A work item
// is (tap, 128 channels of the M side, <= 256 channels of the N side, a share of the pixels); its K loop walks 64-pixel
// chunks, each one pipeline stage of 2 + ceil(NB/64) boxes.  A tap of a 3x3 layer is the A box SHIFTED by (dy, dx)
// with out-of-bounds zero fill (no im2col copy, no halo logic); a child of a transposed conv is the G box at the
// (qy, qx) phase of a 5-D view (c, qx, w, qy, n*h) of the gradient.  Roles as in tconv_tc.cu: warp 0 = TMA producer,
// warp 1 = MMA issuer, warps 2-9 = epilogue (tcgen05.ld -> fp32 atomicAdd into the reference-layout gradient), two
// TMEM accumulators so the epilogue of one item overlaps the main loop of the next.
#include <cuda.h>

#include "conv_common.cuh"
#include "stream_common.cuh"

namespace s2lc {

using namespace tc;

constexpr int TG_EPI_WARPS = 8;
constexpr int TG_THREADS = (2 + TG_EPI_WARPS) * 32;
constexpr int TG_MAX_STAGES = 8;
constexpr int TG_CHUNK = 64 * 128;     // one box: 64 pixels x 64 channels bf16

struct GPlan {
  int taps;             // 9 (3x3), 4 (children of a transposed conv), 1
  int kind;             // 0 = 1x1 (2-D maps), 1 = 3x3, 2 = transposed conv
  int m_on_g;           // 1: G channels on M (TMEM lanes), A channels on N; 0: swapped
  int CM, CN;           // channel counts of the M / N side
  int NB, NBc;          // N tile (multiple of 16, <= 256) and its 64-channel chunks
  int m_tiles, n_tiles, ksplit, items;
  int kchunks, cps;     // 64-pixel chunks in total / per split
  int bw, bh, tiles_x, tiles_y;
  int stages, stage_bytes, tmem_cols;
  int npack;            // > 1: the N operand is `npack` boxes of <= 64 channels each, one per x-shift of a 3x3 tap row
                        // (3) or per child of a transposed conv (4): ONE wide MMA per k-step instead of 3 / 4 narrow
                        // ones (a tcgen05.mma costs ~85 cycles however narrow it is: thin layers are bound by the count)
  int mboxes;           // boxes of the M operand that exist (1 when it has <= 64 channels)
};

__device__ __forceinline__ void tg_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tg_load_2d(uint32_t dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tg_load_4d(uint32_t dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2,
                                           int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], "
      "[%2];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tg_load_5d(uint32_t dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2,
                                           int c3, int c4) {
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, "
      "%7}], [%2];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
      : "memory");
}

struct GItem {
  int tap, mt, nt, ks;
};
__device__ __forceinline__ GItem item_of(const GPlan& K, int it) {
  GItem I;
  I.nt = it % K.n_tiles;
  int r = it / K.n_tiles;
  I.mt = r % K.m_tiles;
  r /= K.m_tiles;
  I.tap = r % K.taps;
  I.ks = r / K.taps;
  return I;
}

__global__ void __launch_bounds__(TG_THREADS, 1) twgrad_kernel(WgradArgs P, GPlan K,
                                                               const __grid_constant__ CUtensorMap tmA,
                                                               const __grid_constant__ CUtensorMap tmG) {
  cudaGridDependencySynchronize();   // programmatic dependent launch (common.cuh)
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* smem =
      reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + K.stages * K.stage_bytes);
  uint64_t* empty = full + TG_MAX_STAGES;
  uint64_t* acc_full = empty + TG_MAX_STAGES;
  uint64_t* acc_empty = acc_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  if (tid == 0) {
    for (int s = 0; s < K.stages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(&acc_full[b], 1); mbar_init(&acc_empty[b], TG_EPI_WARPS); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, (uint32_t)K.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t smem_base = smem_u32(smem);

  if (warp == 0) {
    // ================================================================== TMA producer
    if (elect_one()) {
      const CUtensorMap* mapM = K.m_on_g ? &tmG : &tmA;
      const CUtensorMap* mapN = K.m_on_g ? &tmA : &tmG;
      const int per = K.tiles_x * K.tiles_y;
      int s = 0;
      uint32_t ph = 0;
      for (int it = blockIdx.x; it < K.items; it += gridDim.x) {
        const GItem I = item_of(K, it);
        const int c_begin = I.ks * K.cps;
        const int c_end = min(c_begin + K.cps, K.kchunks);
        // tap geometry: 3x3 shifts the A box, a transposed conv selects the (qy, qx) phase of the G box
        // (packed: a 3x3 item is a whole tap ROW dy, a transposed-conv item covers all four children)
        const int dy = K.kind == 1 ? (K.npack > 1 ? I.tap - 1 : I.tap / 3 - 1) : 0;
        const int dx = (K.kind == 1 && K.npack == 1) ? I.tap - (I.tap / 3) * 3 - 1 : 0;
        const int qy = I.tap >> 1, qx = I.tap & 1;
        for (int c = c_begin; c < c_end; ++c) {
          int n = 0, y0 = 0, x0 = 0;
          if (K.kind != 0) {
            n = c / per;
            const int r = c - n * per;
            const int ty = r / K.tiles_x;
            y0 = ty * K.bh;
            x0 = (r - ty * K.tiles_x) * K.bw;
          }
          mbar_wait(&empty[s], ph ^ 1u);
          tg_expect_tx(&full[s], (uint32_t)((K.mboxes + K.NBc) * TG_CHUNK));
          const uint32_t sa = smem_base + (uint32_t)(s * K.stage_bytes);
          // one box of operand `g` (1 = gradient, 0 = activation) starting at channel c0
          auto box = [&](uint32_t dst, const CUtensorMap* map, bool g, int c0) {
            if (K.kind == 0) tg_load_2d(dst, map, &full[s], c0, c * 64);
            else if (K.kind == 1) {
              if (g) tg_load_4d(dst, map, &full[s], c0, x0, y0, n);
              else tg_load_4d(dst, map, &full[s], c0, x0 + dx, y0 + dy, n);
            } else {
              if (g) tg_load_5d(dst, map, &full[s], c0, qx, x0, qy, n * P.H + y0);
              else tg_load_4d(dst, map, &full[s], c0, x0, y0, n);
            }
          };
          box(sa, mapM, K.m_on_g != 0, I.mt * 128);
          if (K.mboxes > 1) box(sa + TG_CHUNK, mapM, K.m_on_g != 0, I.mt * 128 + 64);
          if (K.npack == 1) {
            for (int j = 0; j < K.NBc; ++j)
              box(sa + 2 * TG_CHUNK + (uint32_t)(j * TG_CHUNK), mapN, K.m_on_g == 0, I.nt * K.NB + j * 64);
          } else if (K.kind == 1) {     // N = activation at x-shifts -1, 0, +1 (out-of-bounds columns are zero-filled)
            for (int j = 0; j < 3; ++j)
              tg_load_4d(sa + 2 * TG_CHUNK + (uint32_t)(j * TG_CHUNK), mapN, &full[s], 0, x0 + j - 1, y0 + dy, n);
          } else {                      // N = gradient of the four children (qy, qx) of a transposed conv
            for (int j = 0; j < 4; ++j)
              tg_load_5d(sa + 2 * TG_CHUNK + (uint32_t)(j * TG_CHUNK), mapN, &full[s], 0, j & 1, x0, j >> 1, n * P.H + y0);
          }
          if (++s == K.stages) { s = 0; ph ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ================================================================== MMA issuer
    const bool leader = elect_one();
    const uint32_t idesc = idesc_bf16(128, (uint32_t)K.NB, 1, 1);        // both operands MN-major
    const uint64_t dbase = smem_desc_sw128(0, 1024, TG_CHUNK);           // LBO = distance between 64-channel chunks
    int s = 0;
    uint32_t ph = 0;
    int b = 0;
    uint32_t bph = 0;
    for (int it = blockIdx.x; it < K.items; it += gridDim.x) {
      const GItem I = item_of(K, it);
      const int c_begin = I.ks * K.cps;
      const int c_end = min(c_begin + K.cps, K.kchunks);
      mbar_wait(&acc_empty[b], bph ^ 1u);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)(b * K.NB);
      uint32_t accum = 0;
      for (int c = c_begin; c < c_end; ++c) {
        mbar_wait(&full[s], ph);
        tc_fence_after();
        const uint32_t sa = smem_base + (uint32_t)(s * K.stage_bytes);
        const uint64_t dM = dbase + (uint64_t)(sa >> 4), dN = dbase + (uint64_t)((sa + 2 * TG_CHUNK) >> 4);
        if (leader) {
#pragma unroll
          for (int k4 = 0; k4 < 4; ++k4) {                               // 16 pixels = 16 lines of 128 B per k-step
            umma_bf16(d_tmem, dM + (uint64_t)(k4 * 128), dN + (uint64_t)(k4 * 128), idesc, accum);
            accum = 1;
          }
          umma_commit(&empty[s]);
          if (c == c_end - 1) umma_commit(&acc_full[b]);
        }
        __syncwarp();
        if (++s == K.stages) { s = 0; ph ^= 1u; }
      }
      if (++b == 2) { b = 0; bph ^= 1u; }
    }
  } else {
    // ================================================================== epilogue
    const int ew = warp - 2, q = warp & 3, half = ew >> 2;
    const int r = q * 32 + lane;
    const int nchunks = K.NB / 16;
    int b = 0;
    uint32_t bph = 0;
    for (int it = blockIdx.x; it < K.items; it += gridDim.x) {
      const GItem I = item_of(K, it);
      const int mch = I.mt * 128 + r;
      const int tap3 = K.kind == 1 ? I.tap : 0;
      const int ct_base = K.kind == 2 ? I.tap * P.Cout : 0;
      const int packed = K.npack > 1;
      mbar_wait(&acc_full[b], bph);
      tc_fence_after();
      for (int cj = half; cj < nchunks; cj += 2) {
        float v[16];
        tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(b * K.NB + cj * 16), v);
        if (mch < K.CM) {
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            int nch = I.nt * K.NB + cj * 16 + j;
            if (packed) {
              // column = (box jb, channel nch of the N side): jb is the x-shift of tap (dy, jb) or the child index
              const int jb = nch >> 6;
              nch &= 63;
              if (nch < K.CN) {
                if (K.kind == 1) atomicAdd(P.dw + wgrad_index(P, mch, (I.tap * 3 + jb) * P.Cin + nch), v[j]);
                else atomicAdd(P.dw + wgrad_index(P, jb * P.Cout + nch, mch), v[j]);
              }
            } else if (nch < K.CN) {
              const int co = K.m_on_g ? mch : nch, ci = K.m_on_g ? nch : mch;
              atomicAdd(P.dw + wgrad_index(P, ct_base + co, tap3 * P.Cin + ci), v[j]);
            }
          }
        }
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
}

// bias gradient = per-channel sum of the (materialised) gradient: [rows, C] -> dbias[C] (fp32 atomics, one per block)
__global__ void __launch_bounds__(256) tg_colsum_kernel(const bf16* __restrict__ g, int64_t ld, int64_t rows, int C,
                                                        float* __restrict__ dbias) {
  extern __shared__ float sm[];
  const int CV = C / 8;
  const int cvb = CV < 256 ? CV : 256, rows_par = 256 / cvb;
  const int c_local = threadIdx.x % cvb, r_local = threadIdx.x / cvb;
  for (int cv0 = 0; cv0 < CV; cv0 += cvb) {
    const int cv = cv0 + c_local;
    float acc[8] = {};
    if (r_local < rows_par && cv < CV) {
      for (int64_t r = (int64_t)blockIdx.x * rows_par + r_local; r < rows; r += (int64_t)gridDim.x * rows_par) {
        float v[8];
        sc_unpack8(__ldg(reinterpret_cast<const uint4*>(g + r * ld + cv * 8)), v);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += v[j];
      }
    }
    for (int i = threadIdx.x; i < cvb * 8; i += 256) sm[i] = 0.f;
    __syncthreads();
    if (r_local < rows_par && cv < CV) {
#pragma unroll
      for (int j = 0; j < 8; ++j) atomicAdd(&sm[c_local * 8 + j], acc[j]);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < cvb * 8; i += 256)
      if (cv0 * 8 + i < C) atomicAdd(&dbias[cv0 * 8 + i], sm[i]);
    __syncthreads();
  }
}

// ---- host side --------------------------------------------------------------------------------------------------
typedef CUresult (*TgEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                               const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                               CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static TgEncodeFn tg_encode_fn() {
  static TgEncodeFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<TgEncodeFn>(p);
  }
  return fn;
}
static bool tg_make_map(CUtensorMap* map, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_el,
                        const uint32_t* box) {
  TgEncodeFn enc = tg_encode_fn();
  if (!enc) return false;
  cuuint64_t gdim[5], gstr[4];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) { gdim[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
  for (int i = 0; i + 1 < rank; ++i) gstr[i] = strides_el[i] * 2;
  return enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), gdim, gstr, bx, es,
             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}
bool make_tma_map_bf16(CUtensorMap* map, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_el,
                       const uint32_t* box) {
  return tg_make_map(map, base, rank, dims, strides_el, box);
}
void launch_colsum_bf16(const void* g, int64_t ld, int64_t rows, int C, float* dbias, cudaStream_t st) {
  const int cv = C / 8;
  const int cvb = cv < 256 ? cv : 256, rows_par = 256 / cvb;
  int64_t blocks = (rows + (int64_t)rows_par * 16 - 1) / ((int64_t)rows_par * 16);
  if (blocks > 4 * kNumSMs) blocks = 4 * kNumSMs;
  if (blocks < 1) blocks = 1;
  tg_colsum_kernel<<<(int)blocks, 256, cvb * 8 * sizeof(float), st>>>(reinterpret_cast<const bf16*>(g), ld, rows, C, dbias);
}
static bool tg_aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// Returns 0 if launched, <0 on error, 1 if this shape / operand is left to the loader-warp engines.
int32_t launch_twgrad_tc(const WgradArgs& A, cudaStream_t st) {
  const uint32_t dbg = debug_flags();
  if (dbg & (DBG_NO_TMA | DBG_NO_STREAM)) return 1;
  if (A.dtype != S2LC_BF16) return 1;
  if (A.mode != S2LC_CONV_1X1 && A.mode != S2LC_CONV_3X3 && A.mode != S2LC_CONVT_FWD) return 1;
  // PLAIN operands only
  if (A.a.scale || A.a.gate || A.a.act != S2LC_ACT_NONE || A.g.z || A.g.rowscale) return 1;
  // (the channel COUNTS may be anything: the tensor maps carry them as extents and zero-fill beyond; rows must start on
  //  16 B boundaries)
  if (A.a.ld % 8 != 0 || A.g.ld_dy % 8 != 0 || A.a.ld < A.Cin || A.g.ld_dy < A.Cout || !tg_aligned16(A.a.raw) ||
      !tg_aligned16(A.g.dy))
    return 1;
  if (A.dbias && A.Cout % 8 != 0) return 1;       // the bias-gradient pass reads 8-channel vectors
  const bool take_all = (dbg & DBG_TMA_ALL) != 0;
  const int HW = A.H * A.W;
  const int cmin = A.Cin < A.Cout ? A.Cin : A.Cout;
  GPlan K{};
  K.kind = A.mode == S2LC_CONV_1X1 ? 0 : (A.mode == S2LC_CONV_3X3 ? 1 : 2);
  K.taps = K.kind == 1 ? 9 : (K.kind == 2 ? 4 : 1);
  // deep layers only: both operands stay in L2 while every tap / channel tile re-reads them, and the work is
  // compute-bound; the shallow, HBM-bound layers belong to the rolling-row engine (swgrad_tc.cu)
  if (!take_all) {
    // (tools/tma_wgrad_sweep.sh, profiles/r02c_tma_wgrad_sweep.txt: 3x3 at 128 px is a tie, at 256 px a loss)
    if (K.kind == 1 && HW > 64 * 64 && !(A.Cin <= 64 && A.Cin % 8 == 0 && A.Cout <= 128 && A.W <= 256)) return 1;   // thin layers: packed taps
    if (K.kind == 0 && (HW > 128 * 128 || cmin < 32 || (HW > 64 * 64 && cmin < 32))) return 1;
    if (K.kind == 2 && (HW > 128 * 128 || cmin < 64) && !(A.Cout <= 64 && A.Cin <= 128)) return 1;
  }
  const int64_t M = (int64_t)A.N * HW;
  if (M >= (1ll << 31) / 4) return 1;
  K.bw = 64; K.bh = 1; K.tiles_x = K.tiles_y = 1;
  if (K.kind != 0) {
    K.bw = A.W < 64 ? A.W : 64;
    if ((K.bw & (K.bw - 1)) != 0 || K.bw < 8 || A.W % K.bw != 0) return 1;
    K.bh = 64 / K.bw;
    if (A.H % K.bh != 0) return 1;
    K.tiles_x = A.W / K.bw;
    K.tiles_y = A.H / K.bh;
    K.kchunks = A.N * K.tiles_x * K.tiles_y;
  } else {
    K.kchunks = (int)((M + 63) / 64);
  }
  K.m_on_g = A.Cout >= A.Cin ? 1 : 0;
  K.npack = 1;
  // thin layers: pack the x-shifts of a tap row (3x3, <= 64 input channels) or the four children (transposed conv,
  // <= 64 output channels) along N
  // (Cin a multiple of 8: a 38-channel slice of a 40-wide concat buffer measured slower than the rolling-row engine)
  if (K.kind == 1 && A.Cin <= 64 && A.Cin % 8 == 0 && A.Cout <= 128) { K.npack = 3; K.m_on_g = 1; K.taps = 3; }
  if (K.kind == 2 && A.Cout <= 64 && A.Cin <= 128) { K.npack = 4; K.m_on_g = 0; K.taps = 1; }
  K.CM = K.m_on_g ? A.Cout : A.Cin;
  K.CN = K.m_on_g ? A.Cin : A.Cout;
  K.m_tiles = (K.CM + 127) / 128;
  K.mboxes = K.CM > 64 ? 2 : 1;
  if (K.npack > 1) {
    K.NB = K.npack * 64; K.n_tiles = 1; K.NBc = K.npack;
  } else {
    const int n16 = (K.CN + 15) / 16 * 16;
    const int nt = (n16 + 255) / 256;
    K.NB = ((n16 + nt - 1) / nt + 15) / 16 * 16;
    K.n_tiles = (K.CN + K.NB - 1) / K.NB;
    K.NBc = (K.NB + 63) / 64;
  }
  K.stage_bytes = (2 + K.NBc) * TG_CHUNK;
  const int fixed = 1024 + (2 * TG_MAX_STAGES + 4) * 8 + 64;
  K.stages = (226 * 1024 - fixed) / K.stage_bytes;
  if (K.stages > TG_MAX_STAGES) K.stages = TG_MAX_STAGES;
  if (K.stages < 2) return 1;
  K.tmem_cols = 32;
  while (K.tmem_cols < 2 * K.NB) K.tmem_cols *= 2;
  // split the pixels so that the items fill whole waves: minimise waves * (chunks per item + epilogue of ~8 chunks)
  const int items0 = K.taps * K.m_tiles * K.n_tiles;
  const int sms = kNumSMs;
  int best_ks = 1;
  long best_cost = -1;
  for (int ks = 1; ks <= 160 && ks <= K.kchunks; ++ks) {
    const int cps = (K.kchunks + ks - 1) / ks;
    if (ks > 1 && cps < 4) break;
    const long waves = ((long)items0 * ks + sms - 1) / sms;
    const long cost = waves * (cps + 8);
    if (best_cost < 0 || cost < best_cost) { best_cost = cost; best_ks = ks; }
  }
  K.ksplit = best_ks;
  K.cps = (K.kchunks + K.ksplit - 1) / K.ksplit;
  K.ksplit = (K.kchunks + K.cps - 1) / K.cps;       // no empty splits
  K.items = items0 * K.ksplit;

  CUtensorMap tmA, tmG;
  bool ok;
  const uint64_t lda = (uint64_t)A.a.ld, ldg = (uint64_t)A.g.ld_dy;
  if (K.kind == 0) {
    const uint64_t da[2] = {(uint64_t)A.Cin, (uint64_t)M}, sa[1] = {lda};
    const uint64_t dg[2] = {(uint64_t)A.Cout, (uint64_t)M}, sg[1] = {ldg};
    const uint32_t box[2] = {64, 64};
    ok = tg_make_map(&tmA, A.a.raw, 2, da, sa, box) && tg_make_map(&tmG, A.g.dy, 2, dg, sg, box);
  } else {
    const uint64_t W = (uint64_t)A.W, H = (uint64_t)A.H, N = (uint64_t)A.N;
    const uint64_t da[4] = {(uint64_t)A.Cin, W, H, N}, sa[3] = {lda, W * lda, H * W * lda};
    const uint32_t box4[4] = {64, (uint32_t)K.bw, (uint32_t)K.bh, 1};
    ok = tg_make_map(&tmA, A.a.raw, 4, da, sa, box4);
    if (K.kind == 1) {
      const uint64_t dg[4] = {(uint64_t)A.Cout, W, H, N}, sg[3] = {ldg, W * ldg, H * W * ldg};
      ok = ok && tg_make_map(&tmG, A.g.dy, 4, dg, sg, box4);
    } else {
      // gradient of a 2x2 stride-2 transposed conv lives at 2H x 2W: view (c, qx, w, qy, n*H + h)
      const uint64_t dg[5] = {(uint64_t)A.Cout, 2, W, 2, N * H}, sg[4] = {ldg, 2 * ldg, 2 * W * ldg, 4 * W * ldg};
      const uint32_t box5[5] = {64, 1, (uint32_t)K.bw, 1, (uint32_t)K.bh};
      ok = ok && tg_make_map(&tmG, A.g.dy, 5, dg, sg, box5);
    }
  }
  if (!ok) return 1;
  const size_t smem = (size_t)fixed + (size_t)K.stages * K.stage_bytes;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(twgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) {
      set_error("twgrad_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
      return S2LC_ERR_CUDA;
    }
    attr_set = true;
  }
  if (A.dbias) launch_colsum_bf16(A.g.dy, A.g.ld_dy, M * (K.kind == 2 ? 4 : 1), A.Cout, A.dbias, st);
  const int grid = sms < K.items ? sms : K.items;
  PDL_LAUNCH(twgrad_kernel, dim3(grid), TG_THREADS, smem, st, A, K, tmA, tmG);
  note_tma_launch();
  return check_launch("twgrad_tc");
}

}  // namespace s2lc

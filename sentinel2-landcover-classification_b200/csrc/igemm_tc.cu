// igemm_tc.cu — tcgen05 / TMEM implicit-GEMM convolution engine (bf16 operands, fp32 accumulation in TMEM).
//
//   out[m, n] = sum_{tap, c} A(m, tap, c) * Wp[n, tap*Cpad + c]        m over N*H*W pixels, 128 per CTA
//
// One kernel serves the 1x1 expand/project/head convs, the decoder 3x3 convs, the 2x2 transposed convs and all
// of their data gradients (reference call sites: src/modules/efficientnet_unet.py:312-327, :363-375, :229-238,
// :168-176, :115-119).  What makes it more than a GEMM is the A-operand PROLOGUE: the operand is never read as a
// stored activation, it is produced on the fly from the producer's raw conv output by the loader warps
//   forward : A = act(raw*scale+shift) * gate[n, c]      (BatchNorm(train) + SiLU/ReLU + squeeze-excite gate)
//   backward: A = p*dy*rowscale[n] + q*z + t              (BatchNorm backward + drop-connect)
// gathered through the conv geometry (3x3 halo with zero padding, 2x2 transposed-conv children), rounded to bf16
// and written straight into the 128B-swizzled K-major shared-memory layout tcgen05.mma consumes.  TMA cannot do
// that transform, so the operand path is: 16 B coalesced global loads -> registers -> transform -> st.shared,
// published to the tensor core with fence.proxy.async + mbarrier; one elected thread issues the MMAs, and
// tcgen05.commit releases each shared-memory stage.  The EPILOGUE reads the accumulator with tcgen05.ld and fuses
// bias, residual adds, act'(previous layer), the bf16 store (incl. the 2x2 scatter of the transposed conv into a
// channel slice of the concat buffer) and the per-channel BatchNorm moments (butterfly warp reduction).
#include "conv_common.cuh"
#include "tc_common.cuh"
#include "stream_common.cuh"

namespace s2lc {

using namespace tc;

constexpr int TC_BM = 128;
constexpr int TC_BK = 64;
constexpr int TC_THREADS = 256;
constexpr int TC_A_BYTES = TC_BM * TC_BK * 2;  // 16 KB

__device__ __forceinline__ void unpack8(const uint4& u, float (&v)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 f = __bfloat1622float2(h[i]);
    v[2 * i] = f.x;
    v[2 * i + 1] = f.y;
  }
}
__device__ __forceinline__ uint4 pack8(const float (&v)[8]) {
  uint4 u;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
  return u;
}
__device__ __forceinline__ uint4 ldg16(const void* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }
__device__ __forceinline__ void ldf8(const float* p, float (&v)[8]) {
  float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p + 4));
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}

// butterfly: 16 per-lane column values -> column sums over the 32 lanes; result for column `col` lands in the two
// lanes whose bits [4:1] spell `col`.  16 shuffles instead of 80.
__device__ __forceinline__ float butterfly16(const float (&a)[16], int lane, int& col) {
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

template <int STAGES>
__global__ void __launch_bounds__(TC_THREADS, 2) conv_tc_kernel(ConvArgs P, int BN, int tmem_cols) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  // carve: [stage: A 16 KB | B BN*128] x STAGES, then stats, barriers
  const uint32_t stage_bytes = TC_A_BYTES + (uint32_t)BN * 128u;
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  float* sstat = reinterpret_cast<float*>(smem + STAGES * stage_bytes);   // [2][256]
  uint64_t* bar_empty = reinterpret_cast<uint64_t*>(sstat + 512);         // [STAGES]
  uint64_t* bar_done = bar_empty + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_done + 1);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t M = (int64_t)P.N * P.H * P.W;
  const int64_t m0 = (int64_t)blockIdx.x * TC_BM;
  const int n0 = blockIdx.y * BN;

  for (int i = tid; i < 512; i += TC_THREADS) sstat[i] = 0.f;
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s) mbar_init(&bar_empty[s], 1);
    mbar_init(bar_done, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(tmem_slot, (uint32_t)tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // ---- loader roles: 16 B chunk c of rows rb + 32*i
  const int c8 = tid & 7, rb = tid >> 3;
  RowCoord rc[4];
  bool rvalid[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int64_t m = m0 + rb + 32 * i;
    rvalid[i] = m < M;
    rc[i] = decode_row(P, rvalid[i] ? m : 0);
  }
  const bf16* wp = reinterpret_cast<const bf16*>(P.w_packed);
  const int nk = P.Kpad / TC_BK;
  const uint32_t idesc = idesc_bf16(TC_BM, (uint32_t)BN, 0, 0);

  for (int kc = 0; kc < nk; ++kc) {
    const int s = kc % STAGES;
    if (kc >= STAGES) mbar_wait(&bar_empty[s], (uint32_t)((kc / STAGES - 1) & 1));
    unsigned char* sa = smem + s * stage_bytes;
    unsigned char* sb = sa + TC_A_BYTES;
    // ---------------- A operand: gather + transform + swizzled store
    {
      const int k = kc * TC_BK + c8 * 8;
      const int tap = k / P.Cpad, ch = k - tap * P.Cpad;
      const bool kvalid = tap < P.taps && ch < P.C;
      uint4 raw[4], zraw[4];
      int64_t srow[4];
      bool ok[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        ok[i] = kvalid && rvalid[i] && src_row(P, rc[i], tap, srow[i]);
        raw[i] = make_uint4(0, 0, 0, 0);
        zraw[i] = make_uint4(0, 0, 0, 0);
      }
      if (!P.a_is_grad) {
        const bf16* base = reinterpret_cast<const bf16*>(P.a_fwd.raw);
#pragma unroll
        for (int i = 0; i < 4; ++i)
          if (ok[i]) raw[i] = ldg16(base + srow[i] * P.a_fwd.ld + ch);
        const bool has_aff = P.a_fwd.scale != nullptr, has_gate = P.a_fwd.gate != nullptr;
        const int act = P.a_fwd.act;
        if (has_aff || has_gate || act != S2LC_ACT_NONE) {
          float sc[8], sh[8];
          if (has_aff && kvalid) {
            ldf8(P.a_fwd.scale + ch, sc);
            ldf8(P.a_fwd.shift + ch, sh);
          }
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            if (!ok[i]) continue;
            float v[8];
            unpack8(raw[i], v);
            if (has_aff) {
#pragma unroll
              for (int j = 0; j < 8; ++j) v[j] = fmaf(v[j], sc[j], sh[j]);
            }
            if (act != S2LC_ACT_NONE) {
#pragma unroll
              for (int j = 0; j < 8; ++j) v[j] = act_fwd_fast(v[j], act);
            }
            if (has_gate) {
              float g[8];
              ldf8(P.a_fwd.gate + (int64_t)rc[i].n * P.C + ch, g);
#pragma unroll
              for (int j = 0; j < 8; ++j) v[j] *= g[j];
            }
            raw[i] = pack8(v);
          }
        }
      } else {
        const bf16* dyb = reinterpret_cast<const bf16*>(P.a_bwd.dy);
        const bf16* zb = reinterpret_cast<const bf16*>(P.a_bwd.z);
#pragma unroll
        for (int i = 0; i < 4; ++i)
          if (ok[i]) {
            raw[i] = ldg16(dyb + srow[i] * P.a_bwd.ld_dy + ch);
            if (zb) zraw[i] = ldg16(zb + srow[i] * P.a_bwd.ld_z + ch);
          }
        if (zb || P.a_bwd.rowscale) {
          float pp[8], qq[8], tt[8];
          if (zb && kvalid) {
            ldf8(P.a_bwd.p + ch, pp);
            ldf8(P.a_bwd.q + ch, qq);
            ldf8(P.a_bwd.t + ch, tt);
          }
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            if (!ok[i]) continue;
            float v[8];
            unpack8(raw[i], v);
            if (P.a_bwd.rowscale) {
              float rs = __ldg(P.a_bwd.rowscale + rc[i].n);
#pragma unroll
              for (int j = 0; j < 8; ++j) v[j] *= rs;
            }
            if (zb) {
              float z[8];
              unpack8(zraw[i], z);
#pragma unroll
              for (int j = 0; j < 8; ++j) v[j] = fmaf(pp[j], v[j], fmaf(qq[j], z[j], tt[j]));
            }
            raw[i] = pack8(v);
          }
        }
      }
#pragma unroll
      for (int i = 0; i < 4; ++i)
        *reinterpret_cast<uint4*>(sa + sw128_off((uint32_t)(rb + 32 * i), (uint32_t)c8)) = raw[i];
    }
    // ---------------- B operand: packed weights [Npad, Kpad] K-major
    for (int r = rb; r < BN; r += 32) {
      int gn = n0 + r;
      uint4 w = make_uint4(0, 0, 0, 0);
      if (gn < P.Npad) w = ldg16(wp + (int64_t)gn * P.Kpad + kc * TC_BK + c8 * 8);
      *reinterpret_cast<uint4*>(sb + sw128_off((uint32_t)r, (uint32_t)c8)) = w;
    }
    fence_proxy_async();
    __syncthreads();
    if (warp == 0) {   // converged warp: descriptors stay in uniform registers, one elected lane issues
      tc_fence_after();
      const bool leader = elect_one();
      const uint32_t a_addr = smem_u32(sa), b_addr = smem_u32(sb);
#pragma unroll
      for (int k4 = 0; k4 < TC_BK / 16; ++k4) {
        uint64_t da = smem_desc_sw128(a_addr + k4 * 32, 1024, 16);
        uint64_t db = smem_desc_sw128(b_addr + k4 * 32, 1024, 16);
        if (leader) umma_bf16(tmem_base, da, db, idesc, (kc > 0 || k4 > 0) ? 1u : 0u);
      }
      if (leader) {
        umma_commit(&bar_empty[s]);
        if (kc == nk - 1) umma_commit(bar_done);
      }
      __syncwarp();
    }
  }

  // ---------------- epilogue
  mbar_wait(bar_done, 0);
  tc_fence_after();
  const int q = warp & 3, half = warp >> 2;
  const int64_t m = m0 + q * 32 + lane;
  const bool mvalid = m < M;
  const RowCoord mr = decode_row(P, mvalid ? m : 0);
  const int nchunks = BN / 16;
  bf16* outp = reinterpret_cast<bf16*>(P.out);
  for (int cj = half; cj < nchunks; cj += 2) {
    float acc[16];
    tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(cj * 16), acc);
    const int ncol0 = n0 + cj * 16;
    int nvalid = P.Nout - ncol0;
    nvalid = nvalid > 16 ? 16 : nvalid;
    float s0[16], s1[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) { s0[j] = 0.f; s1[j] = 0.f; }
    if (mvalid && nvalid > 0) {
      int col0 = ncol0;
      int64_t orow = m;
      if (P.mode == S2LC_CONVT_FWD) {
        int qq = ncol0 / P.Cout;
        col0 = ncol0 - qq * P.Cout;
        orow = ((int64_t)mr.n * (2 * P.H) + 2 * mr.h + (qq >> 1)) * (2 * P.W) + 2 * mr.w + (qq & 1);
      }
      const bool full = nvalid == 16;
      if (P.bias) {
#pragma unroll
        for (int j = 0; j < 16; ++j)
          if (j < nvalid) acc[j] += __ldg(P.bias + col0 + j);
      }
      auto add_rows = [&](const void* base, int64_t ld) {
        const bf16* p = reinterpret_cast<const bf16*>(base) + orow * ld + col0;
        if (full && ((reinterpret_cast<uintptr_t>(p) & 15) == 0)) {
          float e[8];
          unpack8(ldg16(p), e);
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[j] += e[j];
          unpack8(ldg16(p + 8), e);
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
        if (full && ((reinterpret_cast<uintptr_t>(p) & 15) == 0)) {
          float e[8];
          unpack8(ldg16(p), e);
#pragma unroll
          for (int j = 0; j < 8; ++j) zp[j] = e[j];
          unpack8(ldg16(p + 8), e);
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
              float u = zp[j];
              if (P.prev.scale) u = fmaf(u, __ldg(P.prev.scale + col0 + j), __ldg(P.prev.shift + col0 + j));
              acc[j] *= act_bwd_fast(u, P.prev.act);
            }
        }
      }
#pragma unroll
      for (int j = 0; j < 16; ++j) acc[j] = roundf_as<bf16>(acc[j]);
      bf16* op = outp + orow * P.ld_out + col0;
      if (full && ((reinterpret_cast<uintptr_t>(op) & 15) == 0)) {
        float lo[8], hi[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) { lo[j] = acc[j]; hi[j] = acc[8 + j]; }
        *reinterpret_cast<uint4*>(op) = pack8(lo);
        *reinterpret_cast<uint4*>(op + 8) = pack8(hi);
      } else {
#pragma unroll
        for (int j = 0; j < 16; ++j)
          if (j < nvalid) op[j] = __float2bfloat16_rn(acc[j]);
      }
      if (P.stats_kind == S2LC_STATS_MOMENTS) {
#pragma unroll
        for (int j = 0; j < 16; ++j)
          if (j < nvalid) { s0[j] = acc[j]; s1[j] = acc[j] * acc[j]; }
      } else if (P.stats_kind == S2LC_STATS_BNBWD) {
        const float rs = P.stats_rowscale ? __ldg(P.stats_rowscale + mr.n) : 1.f;
#pragma unroll
        for (int j = 0; j < 16; ++j)
          if (j < nvalid) { s0[j] = acc[j] * rs; s1[j] = acc[j] * rs * zp[j]; }
      }
    }
    if (P.stats_kind != S2LC_STATS_NONE) {   // warp-uniform
      int col;
      float t0 = butterfly16(s0, lane, col);
      float t1 = butterfly16(s1, lane, col);
      if ((lane & 1) == 0) {
        atomicAdd(&sstat[cj * 16 + col], t0);
        atomicAdd(&sstat[256 + cj * 16 + col], t1);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (P.stats_kind != S2LC_STATS_NONE) {
    for (int i = tid; i < BN; i += TC_THREADS) {
      if (n0 + i < P.Nout) {
        atomicAdd(&P.stats[n0 + i], (double)sstat[i]);
        atomicAdd(&P.stats[P.Nout + n0 + i], (double)sstat[256 + i]);
      }
    }
  }
  if (warp == 0) tmem_dealloc(tmem_base, (uint32_t)tmem_cols);
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

int32_t launch_conv_tc(const ConvArgs& A, cudaStream_t st) {
  // shapes this engine takes; anything else falls to the CUDA-core engine
  if (A.dtype != S2LC_BF16 || A.C % 8 != 0 || A.Kpad % TC_BK != 0 || !aligned16(A.w_packed)) return 1;
  if (A.a_is_grad) {
    if (A.a_bwd.ld_dy % 8 != 0 || !aligned16(A.a_bwd.dy)) return 1;
    if (A.a_bwd.z && (A.a_bwd.ld_z % 8 != 0 || !aligned16(A.a_bwd.z))) return 1;
  } else {
    if (A.a_fwd.ld % 8 != 0 || !aligned16(A.a_fwd.raw)) return 1;
    if (A.a_fwd.scale && !A.a_fwd.shift) return 1;
  }
  if (A.has_prev && A.prev.scale && !A.prev.shift) return 1;
  if (A.mode == S2LC_CONVT_FWD && A.Cout % 16 != 0) return 1;
  const int n16 = (A.Nout + 15) / 16 * 16;
  const int tiles = (n16 + 255) / 256;
  int BN = ((n16 + tiles - 1) / tiles + 15) / 16 * 16;
  int tmem_cols = 32;
  while (tmem_cols < BN) tmem_cols *= 2;
  constexpr int STAGES = 2;
  size_t smem = 1024 + (size_t)STAGES * (TC_A_BYTES + (size_t)BN * 128) + 512 * sizeof(float) + 64;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(conv_tc_kernel<STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (e != cudaSuccess) {
      set_error("conv_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
      return S2LC_ERR_CUDA;
    }
    attr_set = true;
  }
  const int64_t M = (int64_t)A.N * A.H * A.W;
  dim3 grid((unsigned)((M + TC_BM - 1) / TC_BM), (unsigned)((A.Nout + BN - 1) / BN));
  conv_tc_kernel<STAGES><<<grid, TC_THREADS, smem, st>>>(A, BN, tmem_cols);
  return check_launch("conv_tc");
}

}  // namespace s2lc

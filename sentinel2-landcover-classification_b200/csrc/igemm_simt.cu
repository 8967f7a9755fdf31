// igemm_simt.cu — CUDA-core fp32-accumulate implicit-GEMM convolutions (forward/dgrad and wgrad).
// This is the VERIFICATION-mode engine (S2LC_F32, true fp32 math, no TF32) and the reference
// implementation the tcgen05 kernels in igemm_tc.cu are checked against on the GPU; it also serves
// S2LC_BF16 for shapes the tensor-core path does not take.  Operand loaders and the epilogue are
// shared with the tensor-core path through conv_common.cuh so both engines compute the same function.
#include "conv_common.cuh"

namespace s2lc {

constexpr int BM = 64, BN = 64, BK = 16;

// out[m, n] = sum_k A(m, k) * Wp[n, k]
template <typename T>
__global__ void __launch_bounds__(256) conv_simt_kernel(ConvArgs P) {
  __shared__ float As[BK][BM + 4];
  __shared__ float Bs[BK][BN + 4];
  __shared__ float sstat[2][BN];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int64_t m0 = (int64_t)blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  const int64_t M = (int64_t)P.N * P.H * P.W;
  if (tid < 2 * BN) (&sstat[0][0])[tid] = 0.f;

  // A-load role: one row, 4 consecutive k
  const int a_row = tid >> 2, a_kq = (tid & 3) * 4;
  const int64_t am = m0 + a_row;
  RowCoord rc = decode_row(P, am < M ? am : 0);
  const bool a_valid = am < M;
  // B-load role
  const int b_n = tid >> 2, b_kq = (tid & 3) * 4;
  const T* wp = reinterpret_cast<const T*>(P.w_packed);

  float acc[4][4] = {};
  const int Ktot = P.taps * P.Cpad;
  for (int k0 = 0; k0 < Ktot; k0 += BK) {
    {
      int k = k0 + a_kq;
      float v[4] = {0.f, 0.f, 0.f, 0.f};
      if (a_valid && k < Ktot) {
        int tap = k / P.Cpad, c0 = k % P.Cpad;
        int64_t srow;
        if (src_row(P, rc, tap, srow)) {
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (c0 + j < P.C) v[j] = load_a<T>(P, srow, c0 + j, rc.n);
        }
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) As[a_kq + j][a_row] = v[j];
    }
    {
      int n = n0 + b_n, k = k0 + b_kq;
#pragma unroll
      for (int j = 0; j < 4; ++j)
        Bs[b_kq + j][b_n] = (n < P.Npad && k + j < P.Kpad) ? ldf<T>(wp + (int64_t)n * P.Kpad + k + j) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = As[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = Bs[kk][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }

  // epilogue
  float cs0[4] = {}, cs1[4] = {};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int64_t m = m0 + ty * 4 + i;
    if (m >= M) continue;
    RowCoord r = decode_row(P, m);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int n = n0 + tx * 4 + j;
      if (n >= P.Nout) continue;
      float s0, s1;
      epilogue_store<T>(P, r, m, n, acc[i][j], s0, s1);
      cs0[j] += s0;
      cs1[j] += s1;
    }
  }
  if (P.stats_kind != S2LC_STATS_NONE) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      atomicAdd(&sstat[0][tx * 4 + j], cs0[j]);
      atomicAdd(&sstat[1][tx * 4 + j], cs1[j]);
    }
    __syncthreads();
    if (tid < BN && n0 + tid < P.Nout) {
      atomicAdd(&P.stats[n0 + tid], (double)sstat[0][tid]);
      atomicAdd(&P.stats[P.Nout + n0 + tid], (double)sstat[1][tid]);
    }
  }
}

// dW[n, k] = sum_m G(m, n) * A(m, k);  n over Cout (or (q, Cout) for convT), k over (tap, Cin)
template <typename T>
__global__ void __launch_bounds__(256) wgrad_simt_kernel(WgradArgs P, int64_t m_per_slab) {
  constexpr int MK = 16;
  __shared__ float Gs[MK][BN + 4];
  __shared__ float As[MK][BM + 4];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int n0 = blockIdx.x * BN;   // over Ng
  const int k0 = blockIdx.y * BM;   // over Ktot = taps * Cin
  const int64_t M = (int64_t)P.N * P.H * P.W;
  const int64_t mb = (int64_t)blockIdx.z * m_per_slab;
  const int64_t me = (mb + m_per_slab < M) ? mb + m_per_slab : M;
  const int Ng = P.ngroups * P.Cout, Ktot = P.taps * P.Cin;  // stem: taps = 9, k = ci*9+tap
  const int l_m = tid >> 4, l_q = (tid & 15) * 4;
  float acc[4][4] = {};   // [n][k]
  float bsum[4] = {};
  const bool do_bias = P.dbias != nullptr && blockIdx.y == 0;
  for (int64_t mc = mb; mc < me; mc += MK) {
    int64_t m = mc + l_m;
    bool mv = m < me;
    WRow r = wdecode_row(P, mv ? m : 0);
    {  // G
      float v[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        int n = n0 + l_q + j;
        if (mv && n < Ng) v[j] = load_g<T>(P, r, n);
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) Gs[l_m][l_q + j] = v[j];
      if (do_bias) {
#pragma unroll
        for (int j = 0; j < 4; ++j) bsum[j] += v[j];
      }
    }
    {  // A
      float v[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        int k = k0 + l_q + j;
        if (mv && k < Ktot) v[j] = load_wa<T>(P, r, k);
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) As[l_m][l_q + j] = v[j];
    }
    __syncthreads();
#pragma unroll
    for (int mm = 0; mm < MK; ++mm) {
      float g[4], a[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) g[i] = Gs[mm][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) a[j] = As[mm][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(g[i], a[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int n = n0 + ty * 4 + i;
    if (n >= Ng) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int k = k0 + tx * 4 + j;
      if (k >= Ktot) continue;
      atomicAdd(P.dw + wgrad_index(P, n, k), acc[i][j]);
    }
  }
  if (do_bias) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int n = n0 + l_q + j;
      if (n < Ng && bsum[j] != 0.f) atomicAdd(P.dbias + (n % P.Cout), bsum[j]);
    }
  }
}

int32_t launch_conv_simt(const ConvArgs& A, cudaStream_t st) {
  int64_t M = (int64_t)A.N * A.H * A.W;
  dim3 grid((unsigned)((M + BM - 1) / BM), (unsigned)((A.Nout + BN - 1) / BN));
  if (A.dtype == S2LC_F32)
    conv_simt_kernel<float><<<grid, 256, 0, st>>>(A);
  else
    conv_simt_kernel<bf16><<<grid, 256, 0, st>>>(A);
  return check_launch("conv_simt");
}

int32_t launch_wgrad_simt(const WgradArgs& A, cudaStream_t st) {
  int64_t M = (int64_t)A.N * A.H * A.W;
  int Ng = A.ngroups * A.Cout, Ktot = A.taps * A.Cin;
  int gx = (Ng + BN - 1) / BN, gy = (Ktot + BM - 1) / BM;
  int64_t want = (int64_t)kNumSMs * 8 / ((int64_t)gx * gy);
  if (want < 1) want = 1;
  int64_t max_slabs = (M + 255) / 256;
  int64_t slabs = want < max_slabs ? want : max_slabs;
  int64_t per = ((M + slabs - 1) / slabs + 15) / 16 * 16;
  slabs = (M + per - 1) / per;
  dim3 grid(gx, gy, (unsigned)slabs);
  if (A.dtype == S2LC_F32)
    wgrad_simt_kernel<float><<<grid, 256, 0, st>>>(A, per);
  else
    wgrad_simt_kernel<bf16><<<grid, 256, 0, st>>>(A, per);
  return check_launch("wgrad_simt");
}

}  // namespace s2lc

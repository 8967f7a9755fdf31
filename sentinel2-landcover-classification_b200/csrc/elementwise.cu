// elementwise.cu — library plumbing + the HBM-bound per-pixel / per-channel kernels:
// BatchNorm finalize/apply/backward-reduce, squeeze-excite, copies, loss, Adam.
// All activation kernels use 8-channel vectors (16 B bf16 / 32 B fp32) with threads laid out so that
// consecutive lanes touch consecutive 16 B chunks of an NHWC row (coalesced), and a FIXED channel
// vector per thread so per-channel reductions stay in registers until one flush per CTA.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>

#include "common.cuh"

namespace s2lc {

static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
int32_t check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: %s", what, cudaGetErrorString(e));
    return S2LC_ERR_CUDA;
  }
  return S2LC_OK;
}

int num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

static uint32_t g_dbg = 0;
static bool g_dbg_init = false;
uint32_t debug_flags() {
  if (!g_dbg_init) {
    uint32_t f = 0;
    const char* e;
    if ((e = getenv("S2LC_NO_STREAM")) && e[0] == '1') f |= DBG_NO_STREAM;
    if ((e = getenv("S2LC_FORCE_SIMT")) && e[0] == '1') f |= DBG_FORCE_SIMT;
    if ((e = getenv("S2LC_SCONV_PROD")) && e[0] == '7') f |= DBG_SCONV_PROD7;
    if ((e = getenv("S2LC_NO_TMA")) && e[0] == '1') f |= DBG_NO_TMA;
    if ((e = getenv("S2LC_TMA_ALL")) && e[0] == '1') f |= DBG_TMA_ALL;
    if ((e = getenv("S2LC_DW_CB32")) && e[0] == '1') f |= DBG_DW_CB32;
    if ((e = getenv("S2LC_NO_TMA_STORE")) && e[0] == '1') f |= DBG_NO_TMA_STORE;
    if ((e = getenv("S2LC_DW_GENERIC")) && e[0] == '1') f |= DBG_DW_GENERIC;
    if ((e = getenv("S2LC_STREAM_GENERIC")) && e[0] == '1') f |= DBG_STREAM_GENERIC;
    if ((e = getenv("S2LC_SWGRAD_LMAX")) && (atoi(e) == 64 || atoi(e) == 32)) f |= (uint32_t)atoi(e) << 8;
    g_dbg = f;
    g_dbg_init = true;
  }
  return g_dbg;
}
static long long g_simt_fallbacks = 0;
void note_simt_fallback() { __atomic_fetch_add(&g_simt_fallbacks, 1, __ATOMIC_RELAXED); }
static long long g_tma_launches = 0;
void note_tma_launch() { __atomic_fetch_add(&g_tma_launches, 1, __ATOMIC_RELAXED); }

// Thread layout for [rows, C] tensors processed in 8-channel vectors.
struct ChanIter {
  int cvb;       // channel vectors handled concurrently by a block (<= blockDim)
  int rows_par;  // rows processed concurrently by a block
  int c_local, r_local;
  bool active;
  __device__ ChanIter(int CV) {
    cvb = CV < (int)blockDim.x ? CV : (int)blockDim.x;
    rows_par = blockDim.x / cvb;
    c_local = threadIdx.x % cvb;
    r_local = threadIdx.x / cvb;
    active = r_local < rows_par;
  }
};

// Flush per-thread per-channel partials (8 channels x NS stats) to fp64 global accumulators.
// smem must hold NS * cvb * 8 floats.  stats layout: [NS][C].
template <int NS>
__device__ __forceinline__ void flush_stats(float (&acc)[NS][8], const ChanIter& it, int cv0, int C, double* stats,
                                            float* smem) {
  const int n = NS * it.cvb * 8;
  for (int i = threadIdx.x; i < n; i += blockDim.x) smem[i] = 0.f;
  __syncthreads();
  if (it.active && cv0 + it.c_local < C / 8) {
#pragma unroll
    for (int s = 0; s < NS; ++s)
#pragma unroll
      for (int j = 0; j < 8; ++j) atomicAdd(&smem[(s * it.cvb + it.c_local) * 8 + j], acc[s][j]);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    int s = i / (it.cvb * 8), r = i % (it.cvb * 8);
    int c = cv0 * 8 + r;
    if (c < C) atomicAdd(&stats[(int64_t)s * C + c], (double)smem[i]);
  }
  __syncthreads();
}

// ------------------------------------------------------------------------------------------ weight pack
template <typename T>
__global__ void weight_pack_kernel(const float* __restrict__ src, T* __restrict__ dst, int N1, int N2, int Tt, int C,
                                   int Npad, int Cpad, int Kpad, int64_t s1, int64_t s2, int64_t st, int64_t sc,
                                   int flip) {
  int64_t total = (int64_t)Npad * Kpad;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int n = (int)(i / Kpad), k = (int)(i % Kpad);
    int tap = k / Cpad, c = k % Cpad;
    float v = 0.f;
    if (n < N1 * N2 && tap < Tt && c < C) {
      int n1 = n / N2, n2 = n % N2;
      int tp = flip ? (Tt - 1 - tap) : tap;
      v = src[n1 * s1 + n2 * s2 + tp * st + c * sc];
    }
    stf<T>(dst + i, v);
  }
}

// ------------------------------------------------------------------------------------------ BatchNorm
__global__ void bn_finalize_kernel(const double* __restrict__ stats, const float* __restrict__ gamma,
                                   const float* __restrict__ beta, double count, float eps, float momentum,
                                   float* running_mean, float* running_var, float* scale, float* shift, float* mean_o,
                                   float* invstd_o, int C) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  cudaGridDependencySynchronize();   // launched programmatically dependent: the producers' statistics are complete
  if (c >= C) return;
  double mean = stats[c] / count;
  double var = stats[C + c] / count - mean * mean;
  if (var < 0) var = 0;
  double invstd = rsqrt(var + (double)eps);
  float sc = (float)(gamma[c] * invstd);
  scale[c] = sc;
  shift[c] = (float)(beta[c] - mean * gamma[c] * invstd);
  if (mean_o) mean_o[c] = (float)mean;
  if (invstd_o) invstd_o[c] = (float)invstd;
  if (running_mean) {  // torch semantics: running_var uses the unbiased estimator
    double unbiased = count > 1 ? var * count / (count - 1) : var;
    running_mean[c] = (float)((1.0 - momentum) * running_mean[c] + momentum * mean);
    running_var[c] = (float)((1.0 - momentum) * running_var[c] + momentum * unbiased);
  }
}

__global__ void bn_eval_affine_kernel(const float* gamma, const float* beta, const float* rm, const float* rv,
                                      float eps, float* scale, float* shift, int C) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  float inv = rsqrtf(rv[c] + eps);
  scale[c] = gamma[c] * inv;
  shift[c] = beta[c] - rm[c] * gamma[c] * inv;
}

__global__ void bn_bwd_finalize_kernel(const double* __restrict__ stats, const float* __restrict__ mean,
                                       const float* __restrict__ invstd, const float* __restrict__ gamma, double count,
                                       float* p, float* q, float* t, float* dgamma, float* dbeta, int C) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  cudaGridDependencySynchronize();
  if (c >= C) return;
  // stats[c] = sum dy, stats[C+c] = sum dy*z  ->  sum dy*zhat = r*(sum dy z - mu*sum dy)
  double sdy = stats[c], sdyz = stats[C + c];
  double mu = mean[c], r = invstd[c], g = gamma[c];
  double sdyzh = r * (sdyz - mu * sdy);
  double m1 = sdy / count, m2 = sdyzh / count;
  // dz = g r (dy - m1 - zhat m2),  zhat = (z - mu) r
  p[c] = (float)(g * r);
  q[c] = (float)(-g * r * r * m2);
  t[c] = (float)(g * r * (mu * r * m2 - m1));
  if (dgamma) dgamma[c] = (float)sdyzh;
  if (dbeta) dbeta[c] = (float)sdy;
}

template <typename T>
__global__ void bn_apply_kernel(s2lc_lazy in, const T* __restrict__ res, int64_t ld_res,
                                const float* __restrict__ rowscale, T* __restrict__ out, int64_t ld_out, int64_t rows,
                                int C, int HW) {
  const int CV = C / 8;
  ChanIter it(CV);
  const T* raw = reinterpret_cast<const T*>(in.raw);
  for (int cv0 = 0; cv0 < CV; cv0 += it.cvb) {
    int cv = cv0 + it.c_local;
    if (!it.active || cv >= CV) continue;
    float sc[8], sh[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      sc[j] = in.scale ? in.scale[cv * 8 + j] : 1.f;
      sh[j] = in.shift ? in.shift[cv * 8 + j] : 0.f;
    }
    for (int64_t r = (int64_t)blockIdx.x * it.rows_par + it.r_local; r < rows; r += (int64_t)gridDim.x * it.rows_par) {
      float v[8];
      Vec8<T>::load(raw + r * in.ld + cv * 8, v);
      float rs = rowscale ? rowscale[r / HW] : 1.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = act_fwd_t<T>(v[j] * sc[j] + sh[j], in.act) * rs;
      if (in.gate) {   // squeeze-excite gate [N, C]: the materialised operand of a project conv
        float g[8];
        Vec8<float>::load(in.gate + (r / HW) * C + cv * 8, g);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] *= g[j];
      }
      if (res) {
        float e[8];
        Vec8<T>::load(res + r * ld_res + cv * 8, e);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] += e[j];
      }
      Vec8<T>::store(out + r * ld_out + cv * 8, v);
    }
  }
}

template <typename T>
__global__ void bn_bwd_reduce_kernel(const T* __restrict__ dy, int64_t ld_dy, const T* __restrict__ z, int64_t ld_z,
                                     const float* __restrict__ rowscale, double* stats, int64_t rows, int C, int HW) {
  extern __shared__ float smem[];
  const int CV = C / 8;
  ChanIter it(CV);
  for (int cv0 = 0; cv0 < CV; cv0 += it.cvb) {
    int cv = cv0 + it.c_local;
    float acc[2][8] = {};
    if (it.active && cv < CV) {
      for (int64_t r = (int64_t)blockIdx.x * it.rows_par + it.r_local; r < rows;
           r += (int64_t)gridDim.x * it.rows_par) {
        float a[8], b[8];
        Vec8<T>::load(dy + r * ld_dy + cv * 8, a);
        Vec8<T>::load(z + r * ld_z + cv * 8, b);
        float rs = rowscale ? rowscale[r / HW] : 1.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float d = a[j] * rs;
          acc[0][j] += d;
          acc[1][j] += d * b[j];
        }
      }
    }
    flush_stats<2>(acc, it, cv0, C, stats, smem);
  }
}

// ------------------------------------------------------------------------------------------ squeeze-excite
// The streaming squeeze-excite kernels are latency-bound unless several rows are in flight per thread AND enough
// blocks are resident: rows are fetched as RAW 16 B vectors (4 registers each, converted only when consumed) SE_U at a
// time, registers are capped at 64 (4 blocks of 256 threads per SM) and the grid never exceeds the resident blocks.
template <typename T> struct RawV8;
template <> struct RawV8<bf16> {
  uint4 u;
  __device__ __forceinline__ void load(const bf16* p) { u = __ldg(reinterpret_cast<const uint4*>(p)); }
  __device__ __forceinline__ void unpack(float (&v)[8]) const {
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      v[2 * i] = __uint_as_float(w[i] << 16);
      v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
  }
};
template <> struct RawV8<float> {
  float4 a, b;
  __device__ __forceinline__ void load(const float* p) {
    a = __ldg(reinterpret_cast<const float4*>(p));
    b = __ldg(reinterpret_cast<const float4*>(p + 4));
  }
  __device__ __forceinline__ void unpack(float (&v)[8]) const {
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
};
template <typename T> struct SeU { static constexpr int value = sizeof(T) == 2 ? 4 : 2; };   // rows in flight
constexpr int SE_BLOCKS_PER_SM = 4;
constexpr int SE_APPLY_BLOCKS_PER_SM = 3;   // se_bwd_apply holds 4 per-channel vectors + 2 accumulators: 85 registers
// pooled[n, c] = (1/HW) * sum_{rows of sample n} act(z*scale+shift), split over gridDim.x blocks per sample
template <typename T, bool WITH_DA>
__global__ void __launch_bounds__(256, SE_BLOCKS_PER_SM) se_pool_kernel(s2lc_lazy in, const T* __restrict__ da, int64_t ld_da, float* __restrict__ out, int HW,
                               int C, float mul, float* __restrict__ part, unsigned* __restrict__ counter, int VS) {
  cudaGridDependencySynchronize();
  // grid: (blocks per sample, N).  WITH_DA: out[n,c] = sum da*act(u)  (d gate), else mean of act(u).
  // The rows of a sample are dealt to VS VIRTUAL splits, VS a function of (HW, C) only; a block walks the virtual
  // splits blockIdx.x, blockIdx.x + gridDim.x, ...  Every partial therefore sums the same rows in the same order
  // whatever the batch size or the grid: the result is bit-identical run to run AND across batch sizes.
  const int CV = C / 8;
  ChanIter it(CV);
  const int n = blockIdx.y;
  const T* raw = reinterpret_cast<const T*>(in.raw) + (int64_t)n * HW * in.ld;
  const T* dap = WITH_DA ? da + (int64_t)n * HW * ld_da : nullptr;
  extern __shared__ float smem[];
  for (int vs = blockIdx.x; vs < VS; vs += gridDim.x) {
    for (int cv0 = 0; cv0 < CV; cv0 += it.cvb) {
      int cv = cv0 + it.c_local;
      float acc[8] = {};
      if (it.active && cv < CV) {
        float sc[8], sh[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          sc[j] = in.scale ? in.scale[cv * 8 + j] : 1.f;
          sh[j] = in.shift ? in.shift[cv * 8 + j] : 0.f;
        }
        constexpr int U = SeU<T>::value;
        const int stride = VS * it.rows_par;
        for (int r = vs * it.rows_par + it.r_local; r < HW; r += U * stride) {
          RawV8<T> rv[U], rd[U];
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const int rr = r + u * stride;
            if (rr < HW) {
              rv[u].load(raw + (int64_t)rr * in.ld + cv * 8);
              if (WITH_DA) rd[u].load(dap + (int64_t)rr * ld_da + cv * 8);
            }
          }
#pragma unroll
          for (int u = 0; u < U; ++u) {
            if (r + u * stride < HW) {
              float v[8], d[8];
              rv[u].unpack(v);
              if (WITH_DA) rd[u].unpack(d);
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const float a = act_fwd_t<T>(fmaf(v[j], sc[j], sh[j]), in.act);
                acc[j] = WITH_DA ? fmaf(d[j], a, acc[j]) : acc[j] + a;
              }
            }
          }
        }
      }
      // DETERMINISTIC reduction: the rows_par threads that share a channel vector are summed in index order through
      // shared memory, ...
      const int nsm = it.cvb * 8;
      if (it.active) {
#pragma unroll
        for (int j = 0; j < 8; ++j) smem[(it.r_local * it.cvb + it.c_local) * 8 + j] = (cv < CV) ? acc[j] : 0.f;
      }
      __syncthreads();
      for (int i = threadIdx.x; i < nsm; i += blockDim.x) {
        float t = 0.f;
        for (int r = 0; r < it.rows_par; ++r) t += smem[r * nsm + i];
        int c = cv0 * 8 + i;
        if (c < C) part[((int64_t)n * VS + vs) * C + c] = t;
      }
      __syncthreads();
    }
  }
  // ... and the VS partials of a sample by the LAST block to finish, again in index order.
  __shared__ unsigned ticket;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) ticket = atomicAdd(&counter[n], 1u);
  __syncthreads();
  if (ticket == gridDim.x - 1) {
    __threadfence();
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      float t = 0.f;
      for (int sidx = 0; sidx < VS; ++sidx) t += __ldcg(&part[((int64_t)n * VS + sidx) * C + c]);
      out[(int64_t)n * C + c] = t * mul;
    }
    if (threadIdx.x == 0) counter[n] = 0u;    // ready for the next launch on this stream
  }
}

// The squeeze-excite FCs are tiny ([N, C] x [C, S], S = C/24): what matters is spreading them over the SMs.
// Grid = (channel blocks of 256, N); every block recomputes the S hidden units of its sample (cheap) and owns
// 256 of the C channels.
constexpr int SE_CB = 256;

// Hidden pre-activations hp[n, s] = W1[s, :] . pooled[n, :] + b1[s] for ALL samples, one warp per (s, n): grid
// (ceil(S / 8), N).  Its own launch because every (channel block, sample) CTA of the gate kernels used to recompute all
// S hidden units of its sample — C/256 times the W1 traffic (12 x 1.5 MB per sample at C = 3072), which set the time of
// the deep blocks (0.15 ms per forward call, 0.22 ms per backward call).  `zero` (optional, [N, S]) is cleared for the
// backward accumulation that follows.
__global__ void __launch_bounds__(256) se_hidden_kernel(const float* __restrict__ pooled, const float* __restrict__ w1,
                                                        const float* __restrict__ b1, float* __restrict__ hp,
                                                        float* __restrict__ zero, int C, int S) {
  cudaGridDependencySynchronize();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n = blockIdx.y, s = blockIdx.x * 8 + warp;
  if (s >= S) return;
  const float4* wr = reinterpret_cast<const float4*>(w1 + (int64_t)s * C);
  const float4* xr = reinterpret_cast<const float4*>(pooled + (int64_t)n * C);
  float a = 0.f;
  for (int c4 = lane; c4 < C / 4; c4 += 32) {
    const float4 w = __ldg(wr + c4), x = __ldg(xr + c4);
    a += w.x * x.x + w.y * x.y + w.z * x.z + w.w * x.w;
  }
  a = warp_sum(a);
  if (lane == 0) {
    hp[(int64_t)n * S + s] = a + b1[s];
    if (zero) zero[(int64_t)n * S + s] = 0.f;
  }
}

// gate = sigmoid(W2 silu(hp) + b2), grid (channel blocks of 256, N)
__global__ void __launch_bounds__(256) se_fc_fwd_kernel(const float* __restrict__ hp, const float* __restrict__ w2,
                                                        const float* __restrict__ b2, float* __restrict__ gate, int C,
                                                        int S) {
  cudaGridDependencySynchronize();
  extern __shared__ float smem[];
  float* hs = smem;  // [S]
  const int n = blockIdx.y, c = blockIdx.x * SE_CB + threadIdx.x;
  for (int s = threadIdx.x; s < S; s += blockDim.x) hs[s] = act_fwd(hp[(int64_t)n * S + s], S2LC_ACT_SILU);
  __syncthreads();
  if (c < C) {
    float a = b2[c];
    const float* wr = w2 + (int64_t)c * S;
    for (int s = 0; s < S; ++s) a = fmaf(__ldg(wr + s), hs[s], a);
    gate[(int64_t)n * C + c] = sigmoidf_(a);
  }
}

// backward, step 1: dgp[n, c] = dgate * g(1-g);
// dacc[n, s] += sum over this block's channels of dgp[c] * W2[c, s]   (dacc zeroed by se_hidden_kernel)
__global__ void __launch_bounds__(256) se_fc_bwd1_kernel(const float* __restrict__ w2, const float* __restrict__ b2,
                                                         const float* __restrict__ dgate, float* __restrict__ sc_dgp,
                                                         const float* __restrict__ sc_hp, float* __restrict__ sc_dacc,
                                                         int C, int S) {
  cudaGridDependencySynchronize();
  extern __shared__ float smem[];
  float* hs = smem;               // [S] silu(hp)
  float* dg = hs + S;             // [SE_CB]
  const int n = blockIdx.y, c0 = blockIdx.x * SE_CB, c = c0 + threadIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  for (int s = threadIdx.x; s < S; s += blockDim.x) hs[s] = act_fwd(sc_hp[(int64_t)n * S + s], S2LC_ACT_SILU);
  __syncthreads();
  float d = 0.f;
  if (c < C) {
    float a = b2[c];
    const float* wr = w2 + (int64_t)c * S;
    for (int s = 0; s < S; ++s) a = fmaf(__ldg(wr + s), hs[s], a);
    const float g = sigmoidf_(a);
    d = dgate[(int64_t)n * C + c] * g * (1.f - g);
    sc_dgp[(int64_t)n * C + c] = d;
  }
  dg[threadIdx.x] = d;
  __syncthreads();
  const int cn = (C - c0) < SE_CB ? (C - c0) : SE_CB;
  for (int s = warp; s < S; s += nw) {
    float a = 0.f;
    for (int i = lane; i < cn; i += 32) a = fmaf(dg[i], __ldg(w2 + (int64_t)(c0 + i) * S + s), a);
    a = warp_sum(a);
    if (lane == 0) atomicAdd(&sc_dacc[(int64_t)n * S + s], a);
  }
}

// backward, step 2: dhp[s] = dacc[s] * silu'(hp[s]);  dpool[n, c] = sum_s dhp[s] * W1[s, c]
__global__ void __launch_bounds__(256) se_fc_bwd2_kernel(const float* __restrict__ w1, const float* __restrict__ sc_hp,
                                                         const float* __restrict__ sc_dacc, float* __restrict__ dpool,
                                                         int C, int S) {
  cudaGridDependencySynchronize();
  extern __shared__ float smem[];
  float* dh = smem;  // [S]
  const int n = blockIdx.y, c = blockIdx.x * SE_CB + threadIdx.x;
  for (int s = threadIdx.x; s < S; s += blockDim.x)
    dh[s] = sc_dacc[(int64_t)n * S + s] * act_bwd(sc_hp[(int64_t)n * S + s], S2LC_ACT_SILU);
  __syncthreads();
  if (c < C) {
    float a = 0.f;
    for (int s = 0; s < S; ++s) a = fmaf(dh[s], __ldg(w1 + (int64_t)s * C + c), a);
    dpool[(int64_t)n * C + c] = a;
  }
}

// weight part: dw2[c,s] = sum_n dgp[n,c] silu(hp[n,s]); dw1[s,c] = sum_n dhp[n,s] pooled[n,c]; biases.
// grid (channel blocks of 256, S): the per-sample factors silu(hp[n,s]) and dhp[n,s] of the block's hidden unit are
// formed ONCE in shared memory (the first version evaluated two expf + two divisions per (n, s, c): 56 us at
// C = 3072, S = 128), the loop over samples is two coalesced loads and two FMAs; same summation order as before.
__global__ void __launch_bounds__(256) se_fc_bwd_weight_kernel(const float* __restrict__ pooled,
                                                               const float* __restrict__ sc_dgp,
                                                               const float* __restrict__ sc_hp,
                                                               const float* __restrict__ sc_dacc,
                                                               float* __restrict__ dw1, float* __restrict__ db1,
                                                               float* __restrict__ dw2, float* __restrict__ db2, int N,
                                                               int C, int S) {
  cudaGridDependencySynchronize();
  extern __shared__ float smem[];
  float* sh = smem;       // [N] silu(hp[n, s])
  float* dh = smem + N;   // [N] dacc[n, s] * silu'(hp[n, s])
  const int s = blockIdx.y, c = blockIdx.x * SE_CB + threadIdx.x;
  for (int n = threadIdx.x; n < N; n += blockDim.x) {
    const float hpv = sc_hp[(int64_t)n * S + s];
    sh[n] = act_fwd(hpv, S2LC_ACT_SILU);
    dh[n] = sc_dacc[(int64_t)n * S + s] * act_bwd(hpv, S2LC_ACT_SILU);
  }
  __syncthreads();
  if (c >= C) return;
  float a2 = 0.f, a1 = 0.f, bb2 = 0.f, bb1 = 0.f;
  for (int n = 0; n < N; ++n) {
    const float g = sc_dgp[(int64_t)n * C + c], d = dh[n];
    a2 = fmaf(g, sh[n], a2);
    a1 = fmaf(d, pooled[(int64_t)n * C + c], a1);
    bb2 += g;
    bb1 += d;
  }
  dw2[(int64_t)c * S + s] = a2;
  dw1[(int64_t)s * C + c] = a1;
  if (s == 0) db2[c] = bb2;
  if (c == 0) db1[s] = bb1;
}

// da <- (da*gate + dpool/HW) * act'(u), BNBWD stats against z
template <typename T>
__global__ void __launch_bounds__(256, SE_APPLY_BLOCKS_PER_SM) se_bwd_apply_kernel(T* __restrict__ da, int64_t ld_da, s2lc_lazy in, const float* __restrict__ dpool,
                                    double* stats, int HW, int C, float inv_hw) {
  cudaGridDependencySynchronize();
  extern __shared__ float smem[];
  const int CV = C / 8;
  ChanIter it(CV);
  const int n = blockIdx.y;
  const T* raw = reinterpret_cast<const T*>(in.raw) + (int64_t)n * HW * in.ld;
  T* dap = da + (int64_t)n * HW * ld_da;
  for (int cv0 = 0; cv0 < CV; cv0 += it.cvb) {
    int cv = cv0 + it.c_local;
    float acc[2][8] = {};
    if (it.active && cv < CV) {
      float sc[8], sh[8], g[8], dp[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        int c = cv * 8 + j;
        sc[j] = in.scale ? in.scale[c] : 1.f;
        sh[j] = in.shift ? in.shift[c] : 0.f;
        g[j] = in.gate ? in.gate[(int64_t)n * C + c] : 1.f;
        dp[j] = dpool ? dpool[(int64_t)n * C + c] * inv_hw : 0.f;
      }
      constexpr int U = 2;
      const int stride = gridDim.x * it.rows_par;
      for (int r = blockIdx.x * it.rows_par + it.r_local; r < HW; r += U * stride) {
        RawV8<T> rz[U], rd[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int rr = r + u * stride;
          if (rr < HW) {
            rz[u].load(raw + (int64_t)rr * in.ld + cv * 8);
            rd[u].load(dap + (int64_t)rr * ld_da + cv * 8);
          }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int rr = r + u * stride;
          if (rr < HW) {
            float z[8], d[8];
            rz[u].unpack(z);
            rd[u].unpack(d);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              float pre = z[j] * sc[j] + sh[j];
              float v = roundf_as<T>((d[j] * g[j] + dp[j]) * act_bwd_t<T>(pre, in.act));
              d[j] = v;
              acc[0][j] += v;
              acc[1][j] += v * z[j];
            }
            Vec8<T>::store(dap + (int64_t)rr * ld_da + cv * 8, d);
          }
        }
      }
    }
    flush_stats<2>(acc, it, cv0, C, stats, smem);
  }
}

// ------------------------------------------------------------------------------------------ packed + pipelined form
// se_bwd_apply for bf16 storage + SiLU (every squeeze-excite block of a training step): the loads of sweep i+1 are
// issued BEFORE sweep i is consumed (register double buffer) and the math runs on packed fp32 pairs (FFMA2 / FMUL2:
// the same operations per channel in the same order as the generic kernel, so results are bit-identical): +18 % on
// every stage shape (profiles/r02q_stream_sweep.txt).  The same treatment was measured on se_pool / se_bwd_reduce /
// bn_apply and did NOT pay (equal or slower: they lose more from the registers than they gain), so those stay generic.
// act'(u) of SiLU, the packed form of act_bwd_fast
__device__ __forceinline__ float2 dsilu2(const float2& z, const float2& sc, const float2& sh) {
  const float2 half2 = make_float2(0.5f, 0.5f), one2 = make_float2(1.f, 1.f), mone2 = make_float2(-1.f, -1.f);
  const float2 u = f2fma(z, sc, sh);
  const float2 h = f2mul(u, half2);
  const float2 s = f2fma(half2, make_float2(tanh_approx(h.x), tanh_approx(h.y)), half2);
  return f2mul(s, f2fma(u, f2fma(s, mone2, one2), one2));
}
__device__ __forceinline__ uint4 ldg16(const bf16* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }

// da <- (da*gate + dpool/HW) * act'(u), BNBWD stats against z: packed + pipelined form of se_bwd_apply_kernel
template <int U, int MINB>
__global__ void __launch_bounds__(256, MINB) se_bwd_apply_p_kernel(bf16* __restrict__ da, int64_t ld_da, s2lc_lazy in,
                                                                   const float* __restrict__ dpool, double* stats,
                                                                   int HW, int C, float inv_hw) {
  cudaGridDependencySynchronize();
  extern __shared__ float smem[];
  const int CV = C / 8;
  ChanIter it(CV);
  const int n = blockIdx.y;
  const bf16* raw = reinterpret_cast<const bf16*>(in.raw) + (int64_t)n * HW * in.ld;
  bf16* dap = da + (int64_t)n * HW * ld_da;
  for (int cv0 = 0; cv0 < CV; cv0 += it.cvb) {
    const int cv = cv0 + it.c_local;
    float acc[2][8] = {};
    if (it.active && cv < CV) {
      float2 sc[4], sh[4], g[4], dp[4], a0[4], a1[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int c = cv * 8 + 2 * i;
        sc[i] = in.scale ? make_float2(in.scale[c], in.scale[c + 1]) : make_float2(1.f, 1.f);
        sh[i] = in.shift ? make_float2(in.shift[c], in.shift[c + 1]) : make_float2(0.f, 0.f);
        g[i] = in.gate ? make_float2(in.gate[(int64_t)n * C + c], in.gate[(int64_t)n * C + c + 1]) : make_float2(1.f, 1.f);
        dp[i] = dpool ? make_float2(dpool[(int64_t)n * C + c] * inv_hw, dpool[(int64_t)n * C + c + 1] * inv_hw)
                      : make_float2(0.f, 0.f);
        a0[i] = a1[i] = make_float2(0.f, 0.f);
      }
      const int stride = gridDim.x * it.rows_par;
      const bf16* pz = raw + cv * 8;
      bf16* pd = dap + cv * 8;
      uint4 cz[U], cd[U], nz[U], nd[U];
      int r = blockIdx.x * it.rows_par + it.r_local;
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int rr = r + u * stride;
        if (rr < HW) {
          cz[u] = ldg16(pz + (int64_t)rr * in.ld);
          cd[u] = *reinterpret_cast<const uint4*>(pd + (int64_t)rr * ld_da);
        }
      }
      for (; r < HW; r += U * stride) {
#pragma unroll
        for (int u = 0; u < U; ++u) {   // rows of the next sweep: never written by this sweep (distinct rows)
          const int rr = r + (U + u) * stride;
          if (rr < HW) {
            nz[u] = ldg16(pz + (int64_t)rr * in.ld);
            nd[u] = *reinterpret_cast<const uint4*>(pd + (int64_t)rr * ld_da);
          }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int rr = r + u * stride;
          if (rr < HW) {
            const uint32_t zw[4] = {cz[u].x, cz[u].y, cz[u].z, cz[u].w};
            const uint32_t dw_[4] = {cd[u].x, cd[u].y, cd[u].z, cd[u].w};
            uint32_t o[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const float2 zz = bf2_unpack(zw[i]);
              const float2 val = f2fma(bf2_unpack(dw_[i]), g[i], dp[i]);
              o[i] = bf2_pack(f2mul(val, dsilu2(zz, sc[i], sh[i])));
              const float2 vr = bf2_unpack(o[i]);
              a0[i] = f2add(a0[i], vr);
              a1[i] = f2fma(vr, zz, a1[i]);
            }
            *reinterpret_cast<uint4*>(pd + (int64_t)rr * ld_da) = make_uint4(o[0], o[1], o[2], o[3]);
          }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) { cz[u] = nz[u]; cd[u] = nd[u]; }
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        acc[0][2 * i] = a0[i].x; acc[0][2 * i + 1] = a0[i].y;
        acc[1][2 * i] = a1[i].x; acc[1][2 * i + 1] = a1[i].y;
      }
    }
    flush_stats<2>(acc, it, cv0, C, stats, smem);
  }
}

// ------------------------------------------------------------------------------------------ copies
template <typename T>
__global__ void copy_slice_kernel(const T* __restrict__ a, int64_t ld_a, const T* __restrict__ b, int64_t ld_b,
                                  T* __restrict__ dst, int64_t ld_dst, int64_t rows, int C) {
  const int CV = C / 8;
  int64_t total = rows * CV;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = i / CV;
    int cv = (int)(i % CV);
    float v[8];
    Vec8<T>::load(a + r * ld_a + cv * 8, v);
    if (b) {
      float w[8];
      Vec8<T>::load(b + r * ld_b + cv * 8, w);
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] += w[j];
    }
    Vec8<T>::store(dst + r * ld_dst + cv * 8, v);
  }
}

template <typename T>
__global__ void nchw_to_slice_kernel(const float* __restrict__ x, T* __restrict__ dst, int64_t ld_dst, int N, int C,
                                     int H, int W, int Cfill) {
  int64_t HW = (int64_t)H * W, total = (int64_t)N * HW;
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < total; r += (int64_t)gridDim.x * blockDim.x) {
    int64_t n = r / HW, hw = r % HW;
    if (sizeof(T) == 2 && Cfill == 8 && ((reinterpret_cast<uintptr_t>(dst + r * ld_dst) & 15) == 0)) {
      // the usual case (6 bands into an 8-channel slice of the concat buffer): one 16 B store per pixel
      float v[8];
#pragma unroll
      for (int c = 0; c < 8; ++c) v[c] = c < C ? __ldg(x + (n * C + c) * HW + hw) : 0.f;
      Vec8<T>::store(dst + r * ld_dst, v);
      continue;
    }
    for (int c = 0; c < Cfill; ++c) {
      float v = c < C ? x[(n * C + c) * HW + hw] : 0.f;
      stf<T>(dst + r * ld_dst + c, v);
    }
  }
}

__device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
__global__ void dropconnect_kernel(float* rowscale, int n_blocks, int batch, float rate, uint64_t seed,
                                   uint64_t step) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_blocks * batch) return;
  int b = i / batch;
  float keep = 1.f - rate * ((float)b / (float)n_blocks);
  uint64_t h = splitmix64(splitmix64(seed ^ (step * 0xD1B54A32D192ED03ull)) + (uint64_t)i);
  float u = (float)(h >> 40) * (1.0f / 16777216.0f);  // [0,1)
  rowscale[i] = floorf(keep + u) / keep;
}

// ------------------------------------------------------------------------------------------ loss
// FocalLoss.__call__ (reference src/losses.py:77-89) and weighted nn.CrossEntropyLoss (:36-40), forward value and
// d(loss)/d(logits) in one pass over the logits.  One thread per pixel; logits row = CLS contiguous values.
template <typename T, int CLS>
__global__ void loss_kernel(const T* __restrict__ logits, int64_t ld, const int64_t* __restrict__ y,
                            const float* __restrict__ cw, float gamma, float ls, int64_t ignore_index, int kind,
                            float out_scale_const, const float* __restrict__ wsum, float* __restrict__ loss_out,
                            T* __restrict__ dlogits, int64_t ld_d, int64_t rows) {
  float w[CLS];
  float wtot = 0.f;
#pragma unroll
  for (int c = 0; c < CLS; ++c) {
    w[c] = cw ? cw[c] : 1.f;
    wtot += w[c];
  }
  // mean-reduction scale: focal -> 1/rows (ALL pixels, losses.py:131-134); CE -> 1/sum of target weights
  float scale = out_scale_const;
  if (wsum) scale = 1.f / fmaxf(*wsum, 1e-30f);
  float local = 0.f;
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < rows; r += (int64_t)gridDim.x * blockDim.x) {
    float z[CLS];
#pragma unroll
    for (int c = 0; c < CLS; ++c) z[c] = ldf<T>(logits + r * ld + c);
    int64_t t = y[r];
    float d[CLS];
    bool ign = (t == ignore_index) || t < 0 || t >= CLS;
    if (ign) {
#pragma unroll
      for (int c = 0; c < CLS; ++c) d[c] = 0.f;
    } else {
      float m = z[0];
#pragma unroll
      for (int c = 1; c < CLS; ++c) m = fmaxf(m, z[c]);
      float se = 0.f;
#pragma unroll
      for (int c = 0; c < CLS; ++c) se += expf(z[c] - m);
      float lse = m + logf(se);
      float p[CLS], lp_t = 0.f, p_sum_neg_lp = 0.f, w_t = 0.f, wlp = 0.f;
#pragma unroll
      for (int c = 0; c < CLS; ++c) {
        float lp = z[c] - lse;
        p[c] = expf(lp);
        p_sum_neg_lp -= lp;
        wlp -= w[c] * lp;
        if (c == (int)t) {
          lp_t = lp;
          w_t = w[c];
        }
      }
      if (kind == S2LC_LOSS_FOCAL) {
        // ce (unweighted, label-smoothed) -> pt = exp(-ce) -> alpha_t (1-pt)^gamma ce
        float ce = (1.f - ls) * (-lp_t) + (ls / CLS) * p_sum_neg_lp;
        float pt = expf(-ce);
        float om = fmaxf(1.f - pt, 0.f);
        float f = powf(om, gamma);
        float fl = w_t * f * ce;
        // d fl / d ce = alpha [ (1-pt)^g + g (1-pt)^(g-1) pt ce ]
        float fm1 = (gamma == 0.f) ? 0.f : (om > 0.f ? gamma * powf(om, gamma - 1.f) : (gamma == 1.f ? 1.f : 0.f));
        float dfl = w_t * (f + fm1 * pt * ce);
        local += fl;
#pragma unroll
        for (int c = 0; c < CLS; ++c) d[c] = dfl * (p[c] - (c == (int)t ? (1.f - ls) : 0.f) - ls / CLS) * scale;
      } else {
        float li = (1.f - ls) * w_t * (-lp_t) + (ls / CLS) * wlp;
        local += li;
#pragma unroll
        for (int c = 0; c < CLS; ++c)
          d[c] = ((1.f - ls) * w_t * (p[c] - (c == (int)t ? 1.f : 0.f)) + (ls / CLS) * (p[c] * wtot - w[c])) * scale;
      }
    }
#pragma unroll
    for (int c = 0; c < CLS; ++c) stf<T>(dlogits + r * ld_d + c, d[c]);
  }
  local = warp_sum(local);
  __shared__ float part[32];
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = local;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < (blockDim.x >> 5) ? part[threadIdx.x] : 0.f;
    v = warp_sum(v);
    if (threadIdx.x == 0) atomicAdd(loss_out, v * scale);
  }
}

// sum of class weights of the non-ignored targets (denominator of the weighted CE mean)
__global__ void loss_wsum_kernel(const int64_t* __restrict__ y, const float* __restrict__ cw, int64_t ignore_index,
                                 int cls, float* wsum, int64_t rows) {
  float local = 0.f;
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < rows; r += (int64_t)gridDim.x * blockDim.x) {
    int64_t t = y[r];
    if (t != ignore_index && t >= 0 && t < cls) local += cw ? cw[t] : 1.f;
  }
  local = warp_sum(local);
  __shared__ float part[32];
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = local;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < (blockDim.x >> 5) ? part[threadIdx.x] : 0.f;
    v = warp_sum(v);
    if (threadIdx.x == 0) atomicAdd(wsum, v);
  }
}

template <typename T>
__global__ void scale_inplace_kernel(T* x, int64_t ld, const float* factor, int64_t rows, int C) {
  float f = *factor;
  int64_t total = rows * C;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    T* p = x + (i / C) * ld + (i % C);
    stf<T>(p, ldf<T>(p) * f);
  }
}

// ------------------------------------------------------------------------------------------ Adam (L2 decay)
__global__ void adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                            float* __restrict__ v, int64_t n, float step_size, float b1, float b2, float omb1, float omb2, float eps,
                            float wd, float bc2_sqrt, float gscale) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float pi = p[i];
    float gi = g[i] * gscale + wd * pi;
    float mi = b1 * m[i] + omb1 * gi;
    float vi = b2 * v[i] + omb2 * gi * gi;
    m[i] = mi;
    v[i] = vi;
    float denom = sqrtf(vi) / bc2_sqrt + eps;
    p[i] = pi - step_size * (mi / denom);
  }
}

// Adam + bf16/fp32 GEMM-operand re-pack in ONE pass over the flat parameter buffer.  `ent` is a table
// of pack jobs sorted by `start` (flat offset of the source parameter); a parameter with a forward and a data-gradient
// operand has two consecutive entries with the same start.  Each thread updates its elements and scatters the new value
// to every packed copy: source offset -> (n1, n2, tap, c) by peeling the source strides in descending order.
__device__ __forceinline__ void pack_scatter(const s2lc_pack_entry& e, int64_t r, float val) {
  int idx[4] = {0, 0, 0, 0};
#pragma unroll
  for (int d = 0; d < 4; ++d) {
    const int64_t s = e.stride[d];
    if (s > 0) {
      const int64_t q = r / s;
      idx[e.role[d]] = (int)q;
      r -= q * s;
    }
  }
  // a pack may take a SLICE of the source along n2 (the data-gradient operand of the first decoder conv keeps the
  // 32 up-sampled input channels of a 38-channel weight): elements outside the slice have no packed copy
  if (idx[1] >= e.N2 || idx[3] >= e.Cpad) return;
  const int n = idx[0] * e.N2 + idx[1];
  const int tap = e.flip ? (e.Tt - 1 - idx[2]) : idx[2];
  const int64_t di = (int64_t)n * e.Kpad + (int64_t)tap * e.Cpad + idx[3];
  if (e.dtype == S2LC_BF16) reinterpret_cast<bf16*>(e.dst)[di] = __float2bfloat16_rn(val);
  else reinterpret_cast<float*>(e.dst)[di] = val;
}

__global__ void __launch_bounds__(256) adam_pack_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                        float* __restrict__ m, float* __restrict__ v, int64_t n,
                                                        float step_size, float b1, float b2, float omb1, float omb2,
                                                        float eps, float wd, float bc2_sqrt, float gscale,
                                                        const s2lc_pack_entry* __restrict__ ent, int n_ent) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float pi = p[i];
    float gi = g[i] * gscale + wd * pi;
    float mi = b1 * m[i] + omb1 * gi;
    float vi = b2 * v[i] + omb2 * gi * gi;
    m[i] = mi;
    v[i] = vi;
    float denom = sqrtf(vi) / bc2_sqrt + eps;
    pi = pi - step_size * (mi / denom);
    p[i] = pi;
    // last entry with start <= i
    int lo = 0, hi = n_ent;
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (ent[mid].start <= i) lo = mid + 1; else hi = mid;
    }
    for (int j = lo - 1; j >= 0; --j) {
      const s2lc_pack_entry& e = ent[j];
      if (i >= e.start + e.numel) break;          // entries with one start are contiguous; sources never overlap
      pack_scatter(e, i - e.start, pi);
      if (j > 0 && ent[j - 1].start != e.start) break;
    }
  }
}

// ------------------------------------------------------------------------------------------ lazy gradient -> tensor
// g = p*dy*rowscale + q*z + t written out once (in place over dy is allowed).  Used in front of the compute-heavy
// decoder 3x3 layers: their weight- and data-gradient engines split the channels over many CTAs, each of which would
// otherwise redo this transform on its copy of the operand (3-14x per element), and those producers are
// instruction-bound, not HBM-bound.  Same fp32 formula and the same single rounding as the engines' loaders.
template <typename T>
__global__ void __launch_bounds__(256) lazygrad_apply_kernel(s2lc_lazygrad g, T* __restrict__ out, int64_t ld_out,
                                                             int64_t rows, int C, int HW) {
  const int CV = C / 8;
  const int64_t total = rows * CV;
  const T* dy = reinterpret_cast<const T*>(g.dy);
  const T* z = reinterpret_cast<const T*>(g.z);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / CV;
    const int c = (int)(i - r * CV) * 8;
    float v[8];
    Vec8<T>::load(dy + r * g.ld_dy + c, v);
    if (g.rowscale) {
      const float rs = __ldg(g.rowscale + r / HW);
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] *= rs;
    }
    if (z) {
      float zz[8], pp[8], qq[8], tt[8];
      Vec8<T>::load(z + r * g.ld_z + c, zz);
      Vec8<float>::load(g.p + c, pp);
      Vec8<float>::load(g.q + c, qq);
      Vec8<float>::load(g.t + c, tt);
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = fmaf(pp[j], v[j], fmaf(qq[j], zz[j], tt[j]));
    }
    Vec8<T>::store(out + r * ld_out + c, v);
  }
}

// Streaming forms of the two kernels above for rows < 2^30 (every configuration of the reference): a thread keeps ONE
// channel vector (per-channel constants in registers, no 64-bit division per element — the flat-index forms spend
// ~200 instructions per 48 bytes on two emulated divisions and were issue-bound at 3.4-4.0 TB/s), rows strided over
// the grid with U raw 16 B loads per source in flight before the first is consumed.  Same fp32 formulas, same
// single rounding.
template <typename T, int U>
__global__ void __launch_bounds__(256, 3) lazygrad_apply2_kernel(s2lc_lazygrad g, T* __restrict__ out, int64_t ld_out,
                                                              int rows, int C, int HW) {
  cudaGridDependencySynchronize();
  const int CV = C / 8;
  ChanIter it(CV);
  const T* dy = reinterpret_cast<const T*>(g.dy);
  const T* z = reinterpret_cast<const T*>(g.z);
  const int stride = gridDim.x * it.rows_par;
  for (int cv0 = 0; cv0 < CV; cv0 += it.cvb) {
    const int cv = cv0 + it.c_local;
    if (!it.active || cv >= CV) continue;
    float pp[8], qq[8], tt[8];
    if (z) {
      Vec8<float>::load(g.p + cv * 8, pp);
      Vec8<float>::load(g.q + cv * 8, qq);
      Vec8<float>::load(g.t + cv * 8, tt);
    }
    for (int r = blockIdx.x * it.rows_par + it.r_local; r < rows; r += U * stride) {
      RawV8<T> rd[U], rz[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int rr = r + u * stride;
        if (rr < rows) {
          rd[u].load(dy + (int64_t)rr * g.ld_dy + cv * 8);
          if (z) rz[u].load(z + (int64_t)rr * g.ld_z + cv * 8);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int rr = r + u * stride;
        if (rr < rows) {
          float v[8];
          rd[u].unpack(v);
          if (g.rowscale) {
            const float rs = __ldg(g.rowscale + rr / HW);
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] *= rs;
          }
          if (z) {
            float zz[8];
            rz[u].unpack(zz);
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = fmaf(pp[j], v[j], fmaf(qq[j], zz[j], tt[j]));
          }
          Vec8<T>::store(out + (int64_t)rr * ld_out + cv * 8, v);
        }
      }
    }
  }
}

template <typename T, int U>
__global__ void __launch_bounds__(256, 3) bn_apply2_kernel(s2lc_lazy in, const T* __restrict__ res, int64_t ld_res,
                                                        const float* __restrict__ rowscale, T* __restrict__ out,
                                                        int64_t ld_out, int rows, int C, int HW) {
  cudaGridDependencySynchronize();
  const int CV = C / 8;
  ChanIter it(CV);
  const T* raw = reinterpret_cast<const T*>(in.raw);
  const int stride = gridDim.x * it.rows_par;
  const bool per_sample = rowscale != nullptr || in.gate != nullptr;
  for (int cv0 = 0; cv0 < CV; cv0 += it.cvb) {
    const int cv = cv0 + it.c_local;
    if (!it.active || cv >= CV) continue;
    float sc[8], sh[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      sc[j] = in.scale ? in.scale[cv * 8 + j] : 1.f;
      sh[j] = in.shift ? in.shift[cv * 8 + j] : 0.f;
    }
    for (int r = blockIdx.x * it.rows_par + it.r_local; r < rows; r += U * stride) {
      RawV8<T> rv[U], re[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int rr = r + u * stride;
        if (rr < rows) {
          rv[u].load(raw + (int64_t)rr * in.ld + cv * 8);
          if (res) re[u].load(res + (int64_t)rr * ld_res + cv * 8);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int rr = r + u * stride;
        if (rr < rows) {
          const int n = per_sample ? rr / HW : 0;
          const float rs = rowscale ? rowscale[n] : 1.f;
          float v[8];
          rv[u].unpack(v);
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] = act_fwd_t<T>(v[j] * sc[j] + sh[j], in.act) * rs;
          if (in.gate) {
            float gt[8];
            Vec8<float>::load(in.gate + (int64_t)n * C + cv * 8, gt);
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] *= gt[j];
          }
          if (res) {
            float e[8];
            re[u].unpack(e);
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] += e[j];
          }
          Vec8<T>::store(out + (int64_t)rr * ld_out + cv * 8, v);
        }
      }
    }
  }
}

// Threads per block of the ChanIter layout such that no thread idles: a multiple of the channel vectors CV = C/8
// (<= 256), or an even split of CV when it exceeds 256.  With a fixed 256 threads C = 1056 (CV = 132) leaves 124
// threads idle and C = 768 a quarter of them.
static inline int chan_block(int C) {
  const int CV = C / 8;
  if (CV <= 256) return CV * (256 / CV);
  const int k = (CV + 255) / 256;
  return (CV + k - 1) / k;
}
// blocks of the row-strided streaming kernels: every block gets >= 2 sweeps of U rows, at most 8 blocks per SM
static inline int stream_grid(int64_t rows, int C, int U) {
  const int t = chan_block(C);
  const int cvb = C / 8 < t ? C / 8 : t;
  const int rows_par = t / cvb;
  int64_t b = (rows + (int64_t)rows_par * U * 2 - 1) / ((int64_t)rows_par * U * 2);
  if (b < 1) b = 1;
  if (b > kNumSMs * 8) b = kNumSMs * 8;
  return (int)b;
}

static inline int grid_for(int64_t work_items, int per_block, int max_blocks = kNumSMs * 8) {
  int64_t b = (work_items + per_block - 1) / per_block;
  if (b < 1) b = 1;
  if (b > max_blocks) b = max_blocks;
  return (int)b;
}

}  // namespace s2lc

using namespace s2lc;

#define DISPATCH_T(dtype, ...)                                   \
  if ((dtype) == S2LC_F32) {                                     \
    typedef float T;                                             \
    __VA_ARGS__;                                                 \
  } else if ((dtype) == S2LC_BF16) {                             \
    typedef s2lc::bf16 T;                                        \
    __VA_ARGS__;                                                 \
  } else {                                                       \
    s2lc::set_error("bad dtype %d", (int)(dtype));               \
    return S2LC_ERR_SHAPE;                                       \
  }

template <typename T>
static int32_t launch_loss(const void* logits, int64_t ld, const int64_t* y, const float* cw, float gamma, float ls,
                           int64_t ii, int kind, float scale, const float* wsum, float* loss_out, void* dl,
                           int64_t ld_d, int64_t rows, int cls, cudaStream_t st) {
  int grid = grid_for(rows, 256, kNumSMs * 8);
#define L(CLS)                                                                                                       \
  case CLS:                                                                                                          \
    loss_kernel<T, CLS><<<grid, 256, 0, st>>>((const T*)logits, ld, y, cw, gamma, ls, ii, kind, scale, wsum, loss_out, \
                                              (T*)dl, ld_d, rows);                                                    \
    break;
  switch (cls) {
    L(2) L(3) L(4) L(5) L(6) L(7) L(8)
    default:
      set_error("loss: n_classes=%d unsupported (2..8)", cls);
      return S2LC_ERR_SHAPE;
  }
#undef L
  return check_launch("loss_fwd_bwd");
}

extern "C" {

int32_t s2lc_version(void) { return 100; }
const char* s2lc_last_error(void) { return s2lc::g_err; }

int32_t s2lc_arch_check(int32_t device) {
  cudaDeviceProp prop;
  cudaError_t e = cudaGetDeviceProperties(&prop, device);
  if (e != cudaSuccess) {
    set_error("cudaGetDeviceProperties(%d): %s", device, cudaGetErrorString(e));
    return S2LC_ERR_CUDA;
  }
  S2LC_REQUIRE(prop.major == 10, S2LC_ERR_ARCH, "libs2lc is built for sm_100a only; device %d is sm_%d%d", device,
               prop.major, prop.minor);
  return S2LC_OK;
}

int32_t s2lc_memset_zero(void* ptr, int64_t bytes, void* stream) {
  cudaError_t e = cudaMemsetAsync(ptr, 0, (size_t)bytes, (cudaStream_t)stream);
  if (e != cudaSuccess) {
    set_error("cudaMemsetAsync: %s", cudaGetErrorString(e));
    return S2LC_ERR_CUDA;
  }
  return S2LC_OK;
}

int32_t s2lc_weight_pack(const float* src, void* dst, int32_t dtype, int32_t N1, int32_t N2, int32_t T_, int32_t C,
                         int32_t Npad, int32_t Cpad, int32_t Kpad, int64_t s1, int64_t s2, int64_t st, int64_t sc,
                         int32_t flip, void* stream) {
  S2LC_REQUIRE(Npad >= N1 * N2 && Cpad >= C && Kpad >= T_ * Cpad, S2LC_ERR_SHAPE, "weight_pack: bad padding");
  int64_t total = (int64_t)Npad * Kpad;
  DISPATCH_T(dtype, (weight_pack_kernel<T><<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(
                        src, (T*)dst, N1, N2, T_, C, Npad, Cpad, Kpad, s1, s2, st, sc, flip)));
  return check_launch("weight_pack");
}

int32_t s2lc_bn_finalize(const double* stats, const float* gamma, const float* beta, double count, float eps,
                         float momentum, float* running_mean, float* running_var, float* scale, float* shift,
                         float* mean, float* invstd, int32_t C, void* stream) {
  // programmatic dependent launch: these [C]-sized kernels are pure launch latency (252 of them per training step);
  // their grid is scheduled while the producer of the statistics drains and waits in cudaGridDependencySynchronize()
  cudaLaunchConfig_t cfg = pdl_config(dim3((C + 127) / 128), 128, (cudaStream_t)stream);
  cudaLaunchKernelEx(&cfg, bn_finalize_kernel, stats, gamma, beta, count, eps, momentum, running_mean, running_var, scale,
                     shift, mean, invstd, (int)C);
  return check_launch("bn_finalize");
}

int32_t s2lc_bn_eval_affine(const float* gamma, const float* beta, const float* running_mean,
                            const float* running_var, float eps, float* scale, float* shift, int32_t C,
                            void* stream) {
  bn_eval_affine_kernel<<<(C + 127) / 128, 128, 0, (cudaStream_t)stream>>>(gamma, beta, running_mean, running_var,
                                                                            eps, scale, shift, C);
  return check_launch("bn_eval_affine");
}

int32_t s2lc_bn_bwd_finalize(const double* stats, const float* mean, const float* invstd, const float* gamma,
                             double count, float* p, float* q, float* t, float* dgamma, float* dbeta, int32_t C,
                             void* stream) {
  cudaLaunchConfig_t cfg = pdl_config(dim3((C + 127) / 128), 128, (cudaStream_t)stream);
  cudaLaunchKernelEx(&cfg, bn_bwd_finalize_kernel, stats, mean, invstd, gamma, count, p, q, t, dgamma, dbeta, (int)C);
  return check_launch("bn_bwd_finalize");
}

int32_t s2lc_bn_apply(const s2lc_lazy* in, const void* residual, int64_t ld_res, const float* rowscale, void* out,
                      int64_t ld_out, int64_t rows, int32_t C, int32_t HW, int32_t dtype, void* stream) {
  S2LC_REQUIRE(C % 8 == 0 && in->ld % 8 == 0 && ld_out % 8 == 0 && (!residual || ld_res % 8 == 0), S2LC_ERR_ALIGN,
               "bn_apply: C and strides must be multiples of 8 (C=%d)", C);
  if (rows > 0 && rows < (1 << 30) && HW > 0) {
    DISPATCH_T(dtype, PDL_LAUNCH((bn_apply2_kernel<T, 4>), dim3(stream_grid(rows, C, 4)), chan_block(C), 0,
                                 (cudaStream_t)stream, *in, (const T*)residual, ld_res, rowscale, (T*)out, ld_out,
                                 (int)rows, (int)C, (int)HW));
    return check_launch("bn_apply");
  }
  DISPATCH_T(dtype, (bn_apply_kernel<T><<<grid_for(rows * (C / 8), 256 * 4), 256, 0, (cudaStream_t)stream>>>(
                        *in, (const T*)residual, ld_res, rowscale, (T*)out, ld_out, rows, C, HW)));
  return check_launch("bn_apply");
}

static inline size_t stats_smem(int C, int ns) {
  int cvb = C / 8 < 256 ? C / 8 : 256;
  return (size_t)ns * cvb * 8 * sizeof(float);
}

int32_t s2lc_bn_bwd_reduce(const void* dy, int64_t ld_dy, const void* z, int64_t ld_z, const float* rowscale,
                           double* stats, int64_t rows, int32_t C, int32_t HW, int32_t dtype, void* stream) {
  S2LC_REQUIRE(C % 8 == 0 && ld_dy % 8 == 0 && ld_z % 8 == 0, S2LC_ERR_ALIGN, "bn_bwd_reduce: alignment (C=%d)", C);
  DISPATCH_T(dtype, (bn_bwd_reduce_kernel<T><<<grid_for(rows * (C / 8), 256 * 8, kNumSMs * 4), 256, stats_smem(C, 2),
                                               (cudaStream_t)stream>>>((const T*)dy, ld_dy, (const T*)z, ld_z,
                                                                       rowscale, stats, rows, C, HW)));
  return check_launch("bn_bwd_reduce");
}

int32_t s2lc_lazygrad_apply(const s2lc_lazygrad* g, void* out, int64_t ld_out, int64_t rows, int32_t C, int32_t HW,
                            int32_t dtype, void* stream) {
  S2LC_REQUIRE(g != nullptr && g->dy != nullptr && out != nullptr, S2LC_ERR_SHAPE, "lazygrad_apply: null pointer");
  S2LC_REQUIRE(C % 8 == 0 && g->ld_dy % 8 == 0 && ld_out % 8 == 0 && (!g->z || g->ld_z % 8 == 0), S2LC_ERR_ALIGN,
               "lazygrad_apply: alignment (C=%d)", C);
  S2LC_REQUIRE(!g->z || (g->p && g->q && g->t), S2LC_ERR_SHAPE, "lazygrad_apply: z needs p, q, t");
  S2LC_REQUIRE(HW > 0, S2LC_ERR_SHAPE, "lazygrad_apply: HW");
  if (rows == 0) return S2LC_OK;
  if (rows < (1 << 30)) {
    DISPATCH_T(dtype, PDL_LAUNCH((lazygrad_apply2_kernel<T, 4>), dim3(stream_grid(rows, C, 4)), chan_block(C), 0,
                                 (cudaStream_t)stream, *g, (T*)out, ld_out, (int)rows, (int)C, (int)HW));
    return check_launch("lazygrad_apply");
  }
  DISPATCH_T(dtype, (lazygrad_apply_kernel<T><<<grid_for(rows * (C / 8), 256, kNumSMs * 8), 256, 0,
                                                (cudaStream_t)stream>>>(*g, (T*)out, ld_out, rows, C, HW)));
  return check_launch("lazygrad_apply");
}

static inline int se_splits(int N, int HW, int C, int blocks_per_sm = SE_BLOCKS_PER_SM) {
  int cvb = C / 8 < 256 ? C / 8 : 256;
  int rows_par = 256 / cvb;
  int want = (kNumSMs * blocks_per_sm) / N;   // floor: one wave of resident blocks, never a partial second one
  int maxs = (HW + rows_par * 4 - 1) / (rows_par * 4);
  int s = want < maxs ? want : maxs;
  return s < 1 ? 1 : s;
}

// Virtual splits of the deterministic pooling: a power of two that depends on (HW, C) ONLY (<= 32, >= 4 row groups of
// work each), and the grid along x = the largest power of two <= min(VS, resident blocks / N).
static inline int se_vsplits(int HW, int C) {
  int cvb = C / 8 < 256 ? C / 8 : 256;
  int rows_par = 256 / cvb;
  int maxs = (HW + rows_par * 4 - 1) / (rows_par * 4);
  int vs = 1;
  while (vs * 2 <= maxs && vs < 32) vs *= 2;
  return vs;
}
static inline int se_pool_blocks(int N, int VS, int blocks_per_sm = SE_BLOCKS_PER_SM) {
  int want = (kNumSMs * blocks_per_sm) / N;
  int b = 1;
  while (b * 2 <= want && b * 2 <= VS) b *= 2;
  return b;
}

// Library-owned scratch of the deterministic squeeze-excite reductions: per-block partials [N][splits][C] and one
// arrival counter per sample (zero between launches).  Cached per device, grown on demand (include/s2lc.h: "the
// library never allocates persistent device memory except a cached, size-keyed workspace pool").  One stream per
// process launches these kernels, so consecutive launches reuse the buffer in stream order.
struct SeScratch { float* part = nullptr; size_t floats = 0; unsigned* counter = nullptr; int n = 0; };
static SeScratch g_se[64];
static int32_t se_scratch(int N, size_t floats, float** part, unsigned** counter) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { set_error("se scratch: bad device"); return S2LC_ERR_CUDA; }
  SeScratch& sc = g_se[dev];
  if (floats > sc.floats) {
    if (sc.part) cudaFree(sc.part);
    sc.floats = floats + floats / 2;
    if (cudaMalloc(&sc.part, sc.floats * sizeof(float)) != cudaSuccess) { sc.part = nullptr; sc.floats = 0; set_error("se scratch: cudaMalloc"); return S2LC_ERR_CUDA; }
  }
  if (N > sc.n) {
    if (sc.counter) cudaFree(sc.counter);
    sc.n = N < 1024 ? 1024 : 2 * N;
    if (cudaMalloc(&sc.counter, sc.n * sizeof(unsigned)) != cudaSuccess || cudaMemset(sc.counter, 0, sc.n * sizeof(unsigned)) != cudaSuccess) {
      sc.counter = nullptr; sc.n = 0; set_error("se scratch: cudaMalloc"); return S2LC_ERR_CUDA;
    }
  }
  *part = sc.part; *counter = sc.counter;
  return S2LC_OK;
}

int32_t s2lc_se_pool(const s2lc_lazy* in, float* pooled, int32_t N, int32_t HW, int32_t C, int32_t dtype,
                     void* stream) {
  S2LC_REQUIRE(C % 8 == 0 && in->ld % 8 == 0, S2LC_ERR_ALIGN, "se_pool: alignment (C=%d)", C);
  const int VS = se_vsplits(HW, C);
  dim3 grid(se_pool_blocks(N, VS), N);
  float* part; unsigned* counter;
  int32_t rc = se_scratch(N, (size_t)N * VS * C, &part, &counter);
  if (rc) return rc;
  DISPATCH_T(dtype, PDL_LAUNCH((se_pool_kernel<T, false>), grid, chan_block(C), 256 * 8 * sizeof(float),
                               (cudaStream_t)stream, *in, (const T*)nullptr, (int64_t)0, pooled, (int)HW, (int)C,
                               1.f / (float)HW, part, counter, VS));
  return check_launch("se_pool");
}

int32_t s2lc_se_bwd_reduce(const void* da, int64_t ld_da, const s2lc_lazy* in, float* dgate, int32_t N, int32_t HW,
                           int32_t C, int32_t dtype, void* stream) {
  S2LC_REQUIRE(C % 8 == 0 && in->ld % 8 == 0 && ld_da % 8 == 0, S2LC_ERR_ALIGN, "se_bwd_reduce: alignment (C=%d)", C);
  const int VS = se_vsplits(HW, C);
  dim3 grid(se_pool_blocks(N, VS), N);
  float* part; unsigned* counter;
  int32_t rc = se_scratch(N, (size_t)N * VS * C, &part, &counter);
  if (rc) return rc;
  DISPATCH_T(dtype, PDL_LAUNCH((se_pool_kernel<T, true>), grid, chan_block(C), 256 * 8 * sizeof(float),
                               (cudaStream_t)stream, *in, (const T*)da, ld_da, dgate, (int)HW, (int)C, 1.f, part, counter,
                               VS));
  return check_launch("se_bwd_reduce");
}

// library-owned [N, S] scratch of the forward hidden pre-activations (one stream per process launches these kernels)
static float* g_se_hp[64];
static size_t g_se_hp_floats[64];
static int32_t se_hp_scratch(size_t floats, float** out) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { set_error("se hidden scratch: bad device"); return S2LC_ERR_CUDA; }
  if (floats > g_se_hp_floats[dev]) {
    if (g_se_hp[dev]) cudaFree(g_se_hp[dev]);
    g_se_hp_floats[dev] = floats < 16384 ? 16384 : 2 * floats;
    if (cudaMalloc(&g_se_hp[dev], g_se_hp_floats[dev] * sizeof(float)) != cudaSuccess) {
      g_se_hp[dev] = nullptr; g_se_hp_floats[dev] = 0; set_error("se hidden scratch: cudaMalloc"); return S2LC_ERR_CUDA;
    }
  }
  *out = g_se_hp[dev];
  return S2LC_OK;
}

int32_t s2lc_se_fc_fwd(const float* pooled, const float* w1, const float* b1, const float* w2, const float* b2,
                       float* gate, int32_t N, int32_t C, int32_t S, void* stream) {
  S2LC_REQUIRE(C % 4 == 0, S2LC_ERR_ALIGN, "se_fc_fwd: C=%d must be a multiple of 4", C);
  float* hp;
  int32_t rc = se_hp_scratch((size_t)N * S, &hp);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  float* nozero = nullptr;
  cudaLaunchConfig_t c1 = pdl_config(dim3((S + 7) / 8, N), 256, st);
  cudaLaunchKernelEx(&c1, se_hidden_kernel, pooled, w1, b1, hp, nozero, (int)C, (int)S);
  cudaLaunchConfig_t c2 = pdl_config(dim3((C + SE_CB - 1) / SE_CB, N), 256, st, sizeof(float) * (size_t)S);
  cudaLaunchKernelEx(&c2, se_fc_fwd_kernel, (const float*)hp, w2, b2, gate, (int)C, (int)S);
  return check_launch("se_fc_fwd");
}

// ---- allocation-free variant of s2lc_se_fc_fwd (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA-graph capture, tight serving loops).  The [N, S] hidden
// pre-activations live in a file-scope device array instead of the grown-on-demand pool above: every GPU that loads
// the library holds these 256 KB until exit.  ALL calls share the array and it keeps the last call's contents; the
// first kernel overwrites every element the second one reads, so nothing has to start at zero — but calls that
// overlap on the GPU (two streams, two host threads) race on it.
#define S2LC_SE_NOALLOC_MAX_FLOATS 65536          /* N * S <= 65536 (b7 at bs 256: 256 x 160 = 40960) */
static __device__ __align__(256) float g_se_hp_static[S2LC_SE_NOALLOC_MAX_FLOATS];
static float* g_se_hp_static_ptr[64];             // device address of the array, resolved once per device

int32_t s2lc_se_fc_fwd_noalloc(const float* pooled, const float* w1, const float* b1, const float* w2, const float* b2,
                               float* gate, int32_t N, int32_t C, int32_t S, void* stream) {
  S2LC_REQUIRE(C % 4 == 0, S2LC_ERR_ALIGN, "se_fc_fwd_noalloc: C=%d must be a multiple of 4", C);
  S2LC_REQUIRE(N > 0 && S > 0 && (int64_t)N * S <= S2LC_SE_NOALLOC_MAX_FLOATS, S2LC_ERR_SHAPE,
               "se_fc_fwd_noalloc: N*S = %lld exceeds the static scratch of %d floats (use s2lc_se_fc_fwd)",
               (long long)N * S, S2LC_SE_NOALLOC_MAX_FLOATS);
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { set_error("se_fc_fwd_noalloc: bad device"); return S2LC_ERR_CUDA; }
  if (!g_se_hp_static_ptr[dev]) {
    // the array's name is a host-side stand-in here: the kernels need its address on the CURRENT device
    float* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_se_hp_static);
    if (e != cudaSuccess || !ptr) { set_error("se_fc_fwd_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e)); return S2LC_ERR_CUDA; }
    g_se_hp_static_ptr[dev] = ptr;
  }
  float* hp = g_se_hp_static_ptr[dev];
  cudaStream_t st = (cudaStream_t)stream;
  se_hidden_kernel<<<dim3((S + 7) / 8, N), 256, 0, st>>>(pooled, w1, b1, hp, nullptr, C, S);
  dim3 grid((C + SE_CB - 1) / SE_CB, N);
  se_fc_fwd_kernel<<<grid, 256, sizeof(float) * (size_t)S, st>>>(hp, w2, b2, gate, C, S);
  return check_launch("se_fc_fwd_noalloc");
}

int32_t s2lc_se_fc_bwd(const float* pooled, const float* w1, const float* b1, const float* w2, const float* b2,
                       const float* dgate, float* dpool, float* dw1, float* db1, float* dw2, float* db2,
                       float* scratch, int32_t N, int32_t C, int32_t S, void* stream) {
  S2LC_REQUIRE(C % 4 == 0, S2LC_ERR_ALIGN, "se_fc_bwd: C=%d must be a multiple of 4", C);
  S2LC_REQUIRE(N >= 1 && N <= 4096 && S >= 1 && S <= 65535, S2LC_ERR_SHAPE, "se_fc_bwd: N=%d S=%d out of range", N, S);
  float* sc_dgp = scratch;                       // [N, C]
  float* sc_hp = scratch + (size_t)N * C;        // [N, S] hidden pre-activation
  float* sc_dacc = sc_hp + (size_t)N * S;        // [N, S] sum_c dgp * W2 (accumulated atomically)
  cudaStream_t st = (cudaStream_t)stream;
  const dim3 grid((C + SE_CB - 1) / SE_CB, N);
  cudaLaunchConfig_t c0 = pdl_config(dim3((S + 7) / 8, N), 256, st);
  cudaLaunchKernelEx(&c0, se_hidden_kernel, pooled, w1, b1, sc_hp, sc_dacc, (int)C, (int)S);
  cudaLaunchConfig_t c1 = pdl_config(grid, 256, st, sizeof(float) * (size_t)(S + SE_CB));
  cudaLaunchKernelEx(&c1, se_fc_bwd1_kernel, w2, b2, dgate, sc_dgp, (const float*)sc_hp, sc_dacc, (int)C, (int)S);
  cudaLaunchConfig_t c2 = pdl_config(grid, 256, st, sizeof(float) * (size_t)S);
  cudaLaunchKernelEx(&c2, se_fc_bwd2_kernel, w1, (const float*)sc_hp, (const float*)sc_dacc, dpool, (int)C, (int)S);
  cudaLaunchConfig_t c3 = pdl_config(dim3((C + SE_CB - 1) / SE_CB, S), 256, st, sizeof(float) * 2 * (size_t)N);
  cudaLaunchKernelEx(&c3, se_fc_bwd_weight_kernel, pooled, (const float*)sc_dgp, (const float*)sc_hp,
                     (const float*)sc_dacc, dw1, db1, dw2, db2, (int)N, (int)C, (int)S);
  return check_launch("se_fc_bwd");
}

int32_t s2lc_se_bwd_apply(void* da, int64_t ld_da, const s2lc_lazy* in, const float* dpool, double* stats, int32_t N,
                          int32_t HW, int32_t C, int32_t dtype, void* stream) {
  S2LC_REQUIRE(C % 8 == 0 && in->ld % 8 == 0 && ld_da % 8 == 0, S2LC_ERR_ALIGN, "se_bwd_apply: alignment (C=%d)", C);
  dim3 grid(se_splits(N, HW, C, SE_APPLY_BLOCKS_PER_SM), N);
  if (dtype == S2LC_BF16 && in->act == S2LC_ACT_SILU && !(debug_flags() & DBG_STREAM_GENERIC)) {
    constexpr int MINB = 2;
    grid.x = se_splits(N, HW, C, MINB);
    PDL_LAUNCH((se_bwd_apply_p_kernel<2, MINB>), grid, chan_block(C), stats_smem(C, 2), (cudaStream_t)stream, (bf16*)da,
               ld_da, *in, dpool, stats, (int)HW, (int)C, 1.f / (float)HW);
    return check_launch("se_bwd_apply");
  }
  DISPATCH_T(dtype, PDL_LAUNCH((se_bwd_apply_kernel<T>), grid, chan_block(C), stats_smem(C, 2), (cudaStream_t)stream,
                               (T*)da, ld_da, *in, dpool, stats, (int)HW, (int)C, 1.f / (float)HW));
  return check_launch("se_bwd_apply");
}

int32_t s2lc_copy_slice(const void* src, int64_t ld_src, void* dst, int64_t ld_dst, int64_t rows, int32_t C,
                        int32_t dtype, void* stream) {
  S2LC_REQUIRE(C % 8 == 0 && ld_src % 8 == 0 && ld_dst % 8 == 0, S2LC_ERR_ALIGN, "copy_slice: alignment (C=%d)", C);
  DISPATCH_T(dtype, (copy_slice_kernel<T><<<grid_for(rows * (C / 8), 256 * 2), 256, 0, (cudaStream_t)stream>>>(
                        (const T*)src, ld_src, nullptr, 0, (T*)dst, ld_dst, rows, C)));
  return check_launch("copy_slice");
}

int32_t s2lc_add_slice(const void* a, int64_t ld_a, const void* b, int64_t ld_b, void* dst, int64_t ld_dst,
                       int64_t rows, int32_t C, int32_t dtype, void* stream) {
  S2LC_REQUIRE(C % 8 == 0 && ld_a % 8 == 0 && ld_b % 8 == 0 && ld_dst % 8 == 0, S2LC_ERR_ALIGN,
               "add_slice: alignment (C=%d)", C);
  DISPATCH_T(dtype, (copy_slice_kernel<T><<<grid_for(rows * (C / 8), 256 * 2), 256, 0, (cudaStream_t)stream>>>(
                        (const T*)a, ld_a, (const T*)b, ld_b, (T*)dst, ld_dst, rows, C)));
  return check_launch("add_slice");
}

int32_t s2lc_nchw_to_slice(const float* x_nchw, void* dst, int64_t ld_dst, int32_t N, int32_t C, int32_t H, int32_t W,
                           int32_t Cfill, int32_t dtype, void* stream) {
  int64_t rows = (int64_t)N * H * W;
  DISPATCH_T(dtype, (nchw_to_slice_kernel<T><<<grid_for(rows, 256), 256, 0, (cudaStream_t)stream>>>(
                        x_nchw, (T*)dst, ld_dst, N, C, H, W, Cfill)));
  return check_launch("nchw_to_slice");
}

int32_t s2lc_dropconnect_masks(float* rowscale, int32_t n_blocks, int32_t batch, float rate, uint64_t seed,
                               uint64_t step, void* stream) {
  int n = n_blocks * batch;
  dropconnect_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(rowscale, n_blocks, batch, rate, seed, step);
  return check_launch("dropconnect_masks");
}

int32_t s2lc_loss_fwd_bwd(const void* logits, int64_t ld, const int64_t* y, const float* class_w, float gamma,
                          float label_smoothing, int64_t ignore_index, int32_t kind, int32_t reduce_mean,
                          float* loss_out, float* wsum_scratch, void* dlogits, int64_t ld_d, int64_t rows,
                          int32_t n_classes, int32_t dtype, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  cudaMemsetAsync(loss_out, 0, sizeof(float), st);
  float scale = 1.f;
  const float* wsum = nullptr;
  if (reduce_mean) {
    if (kind == S2LC_LOSS_FOCAL) {
      scale = 1.f / (float)rows;
    } else {
      S2LC_REQUIRE(wsum_scratch != nullptr, S2LC_ERR_SHAPE, "loss: CE mean needs wsum_scratch");
      cudaMemsetAsync(wsum_scratch, 0, sizeof(float), st);
      loss_wsum_kernel<<<grid_for(rows, 256, kNumSMs * 8), 256, 0, st>>>(y, class_w, ignore_index, n_classes,
                                                                         wsum_scratch, rows);
      wsum = wsum_scratch;
    }
  }
  if (dtype == S2LC_F32)
    return launch_loss<float>(logits, ld, y, class_w, gamma, label_smoothing, ignore_index, kind, scale, wsum,
                              loss_out, dlogits, ld_d, rows, n_classes, st);
  if (dtype == S2LC_BF16)
    return launch_loss<s2lc::bf16>(logits, ld, y, class_w, gamma, label_smoothing, ignore_index, kind, scale, wsum,
                                   loss_out, dlogits, ld_d, rows, n_classes, st);
  set_error("bad dtype %d", dtype);
  return S2LC_ERR_SHAPE;
}

int32_t s2lc_scale_inplace(void* x, int64_t ld, const float* factor_dev, int64_t rows, int32_t C, int32_t dtype,
                           void* stream) {
  DISPATCH_T(dtype, (scale_inplace_kernel<T><<<grid_for(rows * C, 256 * 4), 256, 0, (cudaStream_t)stream>>>(
                        (T*)x, ld, factor_dev, rows, C)));
  return check_launch("scale_inplace");
}

int32_t s2lc_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, double lr,
                       double beta1, double beta2, double eps, double weight_decay, int64_t step, float grad_scale,
                       void* stream) {
  double bc1 = 1.0 - pow(beta1, (double)step);
  double bc2 = 1.0 - pow(beta2, (double)step);
  adam_kernel<<<grid_for(n, 256 * 4), 256, 0, (cudaStream_t)stream>>>(
      param, grad, exp_avg, exp_avg_sq, n, (float)(lr / bc1), (float)beta1, (float)beta2, (float)(1.0 - beta1),
      (float)(1.0 - beta2), (float)eps, (float)weight_decay, (float)sqrt(bc2), grad_scale);
  return check_launch("adam_step");
}

int64_t s2lc_simt_fallbacks(int32_t reset) {
  long long v = __atomic_load_n(&s2lc::g_simt_fallbacks, __ATOMIC_RELAXED);
  if (reset) __atomic_store_n(&s2lc::g_simt_fallbacks, 0, __ATOMIC_RELAXED);
  return (int64_t)v;
}
int64_t s2lc_tma_launches(int32_t reset) {
  long long v = __atomic_load_n(&s2lc::g_tma_launches, __ATOMIC_RELAXED);
  if (reset) __atomic_store_n(&s2lc::g_tma_launches, 0, __ATOMIC_RELAXED);
  return (int64_t)v;
}
int32_t s2lc_num_sms(void) { return s2lc::num_sms(); }
uint32_t s2lc_set_debug_flags(uint32_t flags) {
  uint32_t old = s2lc::debug_flags();
  s2lc::g_dbg = flags;
  return old;
}

int32_t s2lc_adam_pack_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, double lr,
                            double beta1, double beta2, double eps, double weight_decay, int64_t step, float grad_scale,
                            const s2lc_pack_entry* entries, int32_t n_entries, void* stream) {
  S2LC_REQUIRE(n_entries == 0 || entries != nullptr, S2LC_ERR_SHAPE, "adam_pack_step: null entry table");
  double bc1 = 1.0 - pow(beta1, (double)step);
  double bc2 = 1.0 - pow(beta2, (double)step);
  adam_pack_kernel<<<grid_for(n, 256 * 4), 256, 0, (cudaStream_t)stream>>>(
      param, grad, exp_avg, exp_avg_sq, n, (float)(lr / bc1), (float)beta1, (float)beta2, (float)(1.0 - beta1),
      (float)(1.0 - beta2), (float)eps, (float)weight_decay, (float)sqrt(bc2), grad_scale, entries, n_entries);
  return check_launch("adam_pack_step");
}

}  // extern "C"

// stream_common.cuh — helpers shared by the streaming tcgen05 engines (sconv_tc.cu, swgrad_tc.cu):
// cp.async, 16 B shared-memory moves, bf16x8 pack/unpack, warp election, butterfly column sums.
#pragma once
#include "common.cuh"
#include "tc_common.cuh"

namespace s2lc {

using namespace tc;

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void sc_mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool elect_one() {   // one lane of the (converged) warp
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
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

}  // namespace s2lc

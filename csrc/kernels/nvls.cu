// C3 / C4 / K17 on NVLS: in-switch reduction and multicast over NVSwitch.
//
// A multicast address (`mc`) aliases the SAME offset of a symmetric buffer on every GPU of the
// group: a store to it is replicated by the switch into all copies (`multimem.st`), a
// `multimem.ld_reduce` returns the element-wise sum of all copies computed inside the switch.
// The buffers and the multicast mapping come from torch's symmetric-memory rendezvous (control
// plane, parallel/nvls.py); the kernels below are the data plane.
//
// Reference roles: the "CP" copy vertex / Tee broadcast tree (DryadVertex/VertexHost/vertex/
// vertexHost/vertexHost.cpp:331-571, GraphManager/stagemanager/DrDynamicBroadcast.cpp:125-130) and
// the merge-to-one aggregation stage consumed everywhere (LinqToDryad/DryadLinqQueryNode.cs:
// 2605-2610): here one store per 16 bytes reaches all GPUs, and one load returns the all-GPU sum.
#include "common.cuh"
#include "../../include/dryad_b200.h"

namespace dryad { namespace nvls {

__device__ __forceinline__ float4 mc_ld_reduce_f32x4(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void mc_st_f32x4(float* mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
               ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ unsigned long long mc_ld_reduce_u64(const unsigned long long* mc) {
  unsigned long long v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.u64 %0, [%1];" : "=l"(v) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void mc_st_u64(unsigned long long* mc, unsigned long long v) {
  asm volatile("multimem.st.relaxed.sys.global.u64 [%0], %1;" ::"l"(mc), "l"(v) : "memory");
}

// Two-shot all-reduce in place: this rank owns slice `rank` of the buffer, pulls the all-GPU sum of
// the slice out of the switch and multicasts it back, so every copy ends up with IDENTICAL bits
// (one reduction per element, performed once).  Callers barrier before (all partials written) and
// after (all slices broadcast).
__global__ void __launch_bounds__(256) allreduce_f32(float* mc, int64_t n4, int rank, int world) {
  const int64_t per = (n4 + world - 1) / world;
  const int64_t lo = per * rank, hi = min(n4, lo + per);
  int64_t i = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < hi; i += stride) mc_st_f32x4(mc + i * 4, mc_ld_reduce_f32x4(mc + i * 4));
}
__global__ void __launch_bounds__(256) allreduce_u64(unsigned long long* mc, int64_t n, int rank, int world) {
  const int64_t per = (n + world - 1) / world;
  const int64_t lo = per * rank, hi = min(n, lo + per);
  int64_t i = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < hi; i += stride) mc_st_u64(mc + i, mc_ld_reduce_u64(mc + i));
}

// Multicast copy: src (local, 16-byte aligned) -> the same range of every GPU's copy.
__global__ void __launch_bounds__(256) broadcast16(float* mc_dst, const float4* __restrict__ src, int64_t n16) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n16; i += stride) mc_st_f32x4(mc_dst + i * 4, __ldcs(src + i));
}

}}  // namespace dryad::nvls

extern "C" {
using namespace dryad::nvls;

int dryad_nvls_allreduce_f32(float* mc, int64_t n, int rank, int world, void* stream) {
  if (!mc || n < 0 || (n & 3) || ((uintptr_t)mc & 15) || world < 1 || rank < 0 || rank >= world) return DRYAD_ERR_BAD_ARG;
  if (n == 0) return DRYAD_OK;
  const int64_t per = (n / 4 + world - 1) / world;
  int64_t g = (per + 255) / 256; if (g < 1) g = 1; if (g > 148 * 4) g = 148 * 4;
  allreduce_f32<<<(int)g, 256, 0, (cudaStream_t)stream>>>(mc, n / 4, rank, world);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_nvls_allreduce_u64(uint64_t* mc, int64_t n, int rank, int world, void* stream) {
  if (!mc || n < 0 || ((uintptr_t)mc & 7) || world < 1 || rank < 0 || rank >= world) return DRYAD_ERR_BAD_ARG;
  if (n == 0) return DRYAD_OK;
  const int64_t per = (n + world - 1) / world;
  int64_t g = (per + 255) / 256; if (g < 1) g = 1; if (g > 148 * 4) g = 148 * 4;
  allreduce_u64<<<(int)g, 256, 0, (cudaStream_t)stream>>>((unsigned long long*)mc, n, rank, world);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_nvls_broadcast(void* mc_dst, const void* src, int64_t nbytes, void* stream) {
  if (!mc_dst || nbytes < 0 || (nbytes & 15) || ((uintptr_t)mc_dst & 15) || ((uintptr_t)src & 15)) return DRYAD_ERR_BAD_ARG;
  if (nbytes == 0) return DRYAD_OK;
  const int64_t n16 = nbytes / 16;
  int64_t g = (n16 + 1023) / 1024; if (g < 1) g = 1; if (g > 148 * 8) g = 148 * 8;
  broadcast16<<<(int)g, 256, 0, (cudaStream_t)stream>>>((float*)mc_dst, (const float4*)src, n16);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

}  // extern "C"

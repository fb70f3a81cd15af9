// Shared device/host helpers for the dryad_b200 sm_100a kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/dryad_b200.h"

#define DRYAD_MAX_PEERS 8

#define DRYAD_CUDA_TRY(expr)                      \
  do {                                            \
    cudaError_t _e = (expr);                      \
    if (_e != cudaSuccess) return (int)_e;        \
  } while (0)

#define DRYAD_LAUNCH_CHECK()                      \
  do {                                            \
    cudaError_t _e = cudaGetLastError();          \
    if (_e != cudaSuccess) return (int)_e;        \
  } while (0)

namespace dryad {

__host__ __device__ __forceinline__ uint64_t mix64(uint64_t x) {
  // splitmix64 finalizer: the engine-wide 64-bit key hash (K14).
  x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ULL;
  x ^= x >> 27; x *= 0x94d049bb133111ebULL;
  x ^= x >> 31;
  return x;
}

// Bucket rule of the reference's HashPartition: (hash & 0x7FFFFFFF) % n
// (LinqToDryad/DryadLinqVertex.cs:4861-4862).  We fold the 64-bit hash to 32 bits first.
__host__ __device__ __forceinline__ uint32_t hash_bucket(uint64_t key, uint32_t nparts) {
  uint64_t h = mix64(key);
  uint32_t h32 = (uint32_t)(h ^ (h >> 32));
  return (h32 & 0x7FFFFFFFu) % nparts;
}

__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31; }
__device__ __forceinline__ uint32_t lanemask_lt() {
  uint32_t m; asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m)); return m;
}

__device__ __forceinline__ void st_release_sys(uint64_t* p, uint64_t v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_acquire_sys(const uint64_t* p) {
  uint64_t v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_acquire_gpu_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_relaxed_gpu_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_gpu_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Streaming 16-byte global load that does not allocate in L1.
__device__ __forceinline__ uint4 ld_stream_v4(const uint4* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}

template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Bounded spin: returns false if `cycles` elapse (protocol failure rather than a hung box).
__device__ __forceinline__ bool spin_until_ge(const uint64_t* flag, uint64_t want, long long cycles) {
  long long t0 = clock64();
  while (ld_acquire_sys(flag) < want) {
    if (clock64() - t0 > cycles) return false;
    __nanosleep(64);
  }
  return true;
}

static inline int div_up(long long a, long long b) { return (int)((a + b - 1) / b); }

}  // namespace dryad

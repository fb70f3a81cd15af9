// Micro-benchmark 2: L2 residency of the aggregation table.  Arbitrary (non power-of-two) table
// capacity via multiply-shift range reduction, streaming (evict-first) input loads, relaxed
// gpu-scope probe loads, lane-paired REDs into 32-byte slots vs SoA.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)
#define EMPTY (~0ull)
__device__ __forceinline__ uint64_t mix64(uint64_t x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdull; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ull; x ^= x >> 33; return x;
}
__global__ void gen(uint64_t* keys, int64_t* vals, int64_t n, uint64_t nkeys) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, st = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += st) {
    uint64_t h = mix64((uint64_t)i * 0x9E3779B97F4A7C15ull + 12345);
    keys[i] = (h % nkeys) * 0x9E3779B97F4A7C15ull;   // sparse 64-bit keys
    vals[i] = (int64_t)((h >> 40) % 1000);
  }
}
struct Slot { uint64_t key; unsigned long long sum; unsigned long long cnt; uint64_t pad; };
__global__ void init_soa(uint64_t* k, unsigned long long* s, unsigned long long* c, uint64_t cap) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i < cap; i += (uint64_t)gridDim.x * blockDim.x) { k[i] = EMPTY; s[i] = 0; c[i] = 0; }
}
__global__ void init_aos(Slot* t, uint64_t cap) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i < cap; i += (uint64_t)gridDim.x * blockDim.x) { t[i].key = EMPTY; t[i].sum = 0; t[i].cnt = 0; t[i].pad = 0; }
}
__device__ __forceinline__ uint64_t ld_relaxed(const uint64_t* p) {
  uint64_t v; asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory"); return v;
}
__device__ __forceinline__ uint64_t ld_stream(const uint64_t* p) {
  uint64_t v; v = __ldcs((const unsigned long long*)p); return v;
}
template <int STRIDE>
__device__ __forceinline__ uint64_t probe(uint64_t* tk, uint64_t cap, uint64_t key) {
  uint64_t s = __umul64hi(mix64(key), cap);
  for (;;) {
    uint64_t k = ld_relaxed(tk + s * STRIDE);
    if (k == key) return s;
    if (k == EMPTY) {
      unsigned long long old = atomicCAS((unsigned long long*)(tk + s * STRIDE), EMPTY, (unsigned long long)key);
      if (old == EMPTY || old == key) return s;
    }
    if (++s == cap) s = 0;
  }
}
// SoA
template <bool STREAM>
__global__ void __launch_bounds__(256) soa(uint64_t* tk, unsigned long long* ts, unsigned long long* tc, uint64_t cap,
                                           const uint64_t* __restrict__ keys, const int64_t* __restrict__ vals, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, st = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += st) {
    uint64_t k = STREAM ? ld_stream(keys + i) : keys[i];
    int64_t v = STREAM ? (int64_t)ld_stream((const uint64_t*)vals + i) : vals[i];
    uint64_t s = probe<1>(tk, cap, k);
    atomicAdd(ts + s, (unsigned long long)v);
    atomicAdd(tc + s, 1ull);
  }
}
// AoS paired
template <bool STREAM>
__global__ void __launch_bounds__(256) aosp(Slot* t, uint64_t cap,
                                            const uint64_t* __restrict__ keys, const int64_t* __restrict__ vals, int64_t n) {
  int64_t base = (int64_t)blockIdx.x * blockDim.x, st = (int64_t)gridDim.x * blockDim.x;
  const int lane = threadIdx.x & 31;
  for (; base < n; base += st) {
    int64_t i = base + threadIdx.x;
    uint64_t s = EMPTY; int64_t v = 0;
    if (i < n) {
      uint64_t k = STREAM ? ld_stream(keys + i) : keys[i];
      v = STREAM ? (int64_t)ld_stream((const uint64_t*)vals + i) : vals[i];
      s = probe<4>((uint64_t*)t, cap, k);
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      int src = h * 16 + (lane >> 1);
      uint64_t ss = __shfl_sync(0xffffffffu, s, src);
      int64_t vv = __shfl_sync(0xffffffffu, v, src);
      if (ss != EMPTY) {
        unsigned long long* a = (lane & 1) ? &t[ss].cnt : &t[ss].sum;
        atomicAdd(a, (lane & 1) ? 1ull : (unsigned long long)vv);
      }
    }
  }
}
// AoS unpaired
template <bool STREAM>
__global__ void __launch_bounds__(256) aosu(Slot* t, uint64_t cap,
                                            const uint64_t* __restrict__ keys, const int64_t* __restrict__ vals, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, st = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += st) {
    uint64_t k = STREAM ? ld_stream(keys + i) : keys[i];
    int64_t v = STREAM ? (int64_t)ld_stream((const uint64_t*)vals + i) : vals[i];
    uint64_t s = probe<4>((uint64_t*)t, cap, k);
    atomicAdd(&t[s].sum, (unsigned long long)v);
    atomicAdd(&t[s].cnt, 1ull);
  }
}
// dense: 2 REDs to direct-indexed tables (keys taken modulo cap)
template <bool STREAM>
__global__ void __launch_bounds__(256) dense(unsigned long long* ts, unsigned long long* tc, uint64_t cap,
                                             const uint64_t* __restrict__ keys, const int64_t* __restrict__ vals, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, st = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += st) {
    uint64_t k = STREAM ? ld_stream(keys + i) : keys[i];
    int64_t v = STREAM ? (int64_t)ld_stream((const uint64_t*)vals + i) : vals[i];
    uint64_t s = __umul64hi(mix64(k), cap);
    atomicAdd(ts + s, (unsigned long long)v);
    atomicAdd(tc + s, 1ull);
  }
}
__global__ void check_soa(const uint64_t* k, const unsigned long long* s, const unsigned long long* c, uint64_t cap, unsigned long long* out) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; unsigned long long a = 0, b = 0, g = 0;
  for (; i < cap; i += (uint64_t)gridDim.x * blockDim.x) if (k[i] != EMPTY) { a += s[i]; b += c[i]; g++; }
  atomicAdd(out, a); atomicAdd(out + 1, b); atomicAdd(out + 2, g);
}
__global__ void check_aos(const Slot* t, uint64_t cap, unsigned long long* out) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; unsigned long long a = 0, b = 0, g = 0;
  for (; i < cap; i += (uint64_t)gridDim.x * blockDim.x) if (t[i].key != EMPTY) { a += t[i].sum; b += t[i].cnt; g++; }
  atomicAdd(out, a); atomicAdd(out + 1, b); atomicAdd(out + 2, g);
}
int main(int argc, char** argv) {
  int64_t n = argc > 1 ? atoll(argv[1]) : 125000000;
  uint64_t nkeys = argc > 2 ? strtoull(argv[2], 0, 10) : (1ull << 20);
  double lf = argc > 3 ? atof(argv[3]) : 1.3;
  uint64_t cap = (uint64_t)(nkeys * lf) + 64;
  uint64_t* keys; int64_t* vals; CK(cudaMalloc(&keys, n * 8)); CK(cudaMalloc(&vals, n * 8));
  gen<<<2048, 256>>>(keys, vals, n, nkeys);
  uint64_t* tk; unsigned long long *ts, *tc, *out; Slot* t;
  CK(cudaMalloc(&tk, cap * 8)); CK(cudaMalloc(&ts, cap * 8)); CK(cudaMalloc(&tc, cap * 8)); CK(cudaMalloc(&t, cap * 32)); CK(cudaMalloc(&out, 64));
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  printf("n=%lld nkeys=%llu cap=%llu (x%.2f)  SoA %.1f MB  AoS %.1f MB\n", (long long)n, (unsigned long long)nkeys, (unsigned long long)cap, lf, cap * 24 / 1e6, cap * 32 / 1e6);
  auto run = [&](const char* name, int kind, auto launch) {
    float best = 1e30f; unsigned long long h[3];
    for (int rep = 0; rep < 4; ++rep) {
      if (kind == 0) init_soa<<<1024, 256>>>(tk, ts, tc, cap); else init_aos<<<1024, 256>>>(t, cap);
      CK(cudaDeviceSynchronize());
      cudaEventRecord(e0); launch(); cudaEventRecord(e1); CK(cudaDeviceSynchronize());
      float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    CK(cudaMemset(out, 0, 64));
    if (kind == 0) check_soa<<<1024, 256>>>(tk, ts, tc, cap, out); else check_aos<<<1024, 256>>>(t, cap, out);
    CK(cudaMemcpy(h, out, 24, cudaMemcpyDeviceToHost));
    printf("  %-26s %8.3f ms  %7.2f Grows/s   cnt=%llu groups=%llu\n", name, best, n / (best * 1e6), h[1], h[2]);
  };
  for (int gm = 8; gm <= 16; gm *= 2) {
    int g = 148 * gm; printf(" grid=148x%d\n", gm);
    run("dense", 0, [&] { dense<false><<<g, 256>>>(ts, tc, cap, keys, vals, n); });
    run("dense stream", 0, [&] { dense<true><<<g, 256>>>(ts, tc, cap, keys, vals, n); });
    run("soa", 0, [&] { soa<false><<<g, 256>>>(tk, ts, tc, cap, keys, vals, n); });
    run("soa stream", 0, [&] { soa<true><<<g, 256>>>(tk, ts, tc, cap, keys, vals, n); });
    run("aos unpaired stream", 1, [&] { aosu<true><<<g, 256>>>(t, cap, keys, vals, n); });
    run("aos paired", 1, [&] { aosp<false><<<g, 256>>>(t, cap, keys, vals, n); });
    run("aos paired stream", 1, [&] { aosp<true><<<g, 256>>>(t, cap, keys, vals, n); });
  }
  return 0;
}

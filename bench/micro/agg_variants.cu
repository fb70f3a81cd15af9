// Micro-benchmark: how fast can one B200 fold (key, value) rows into a 1 M-group Sum/Count table?
// Variants differ in table layout (SoA / 32-byte AoS slots), rows in flight per thread, and in how
// the two reductions of a row are issued (two RED instructions, or one RED instruction whose lane
// pairs hit the same 32-byte sector).  Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a
//   -o bench/micro/agg_variants.bin bench/micro/agg_variants.cu ; run on the GPU box.
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
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t st = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += st) {
    uint64_t h = mix64((uint64_t)i * 0x9E3779B97F4A7C15ull + 12345);
    keys[i] = h % nkeys;
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

__device__ __forceinline__ uint64_t foi(uint64_t* keys, uint64_t stride_words, uint64_t mask, uint64_t key) {
  uint64_t s = mix64(key) & mask;
  for (;;) {
    uint64_t k = *reinterpret_cast<volatile uint64_t*>(keys + s * stride_words);
    if (k == key) return s;
    if (k == EMPTY) {
      unsigned long long old = atomicCAS((unsigned long long*)(keys + s * stride_words), (unsigned long long)EMPTY, (unsigned long long)key);
      if (old == EMPTY || old == key) return s;
    }
    s = (s + 1) & mask;
  }
}

// V0: SoA, one row per thread per iteration
__global__ void __launch_bounds__(256) v0(uint64_t* tk, unsigned long long* ts, unsigned long long* tc, uint64_t mask,
                                          const uint64_t* __restrict__ keys, const int64_t* __restrict__ vals, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, st = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += st) {
    uint64_t s = foi(tk, 1, mask, keys[i]);
    atomicAdd(ts + s, (unsigned long long)vals[i]);
    atomicAdd(tc + s, 1ull);
  }
}

// V1: SoA, R rows in flight per thread
template <int R>
__global__ void __launch_bounds__(256) v1(uint64_t* tk, unsigned long long* ts, unsigned long long* tc, uint64_t mask,
                                          const uint64_t* __restrict__ keys, const int64_t* __restrict__ vals, int64_t n) {
  int64_t base = ((int64_t)blockIdx.x * blockDim.x) * R + threadIdx.x, st = (int64_t)gridDim.x * blockDim.x * R;
  for (; base < n; base += st) {
    uint64_t k[R]; int64_t v[R]; uint64_t s[R]; uint64_t cur[R];
#pragma unroll
    for (int j = 0; j < R; ++j) { int64_t i = base + (int64_t)j * blockDim.x; bool ok = i < n; k[j] = ok ? keys[i] : EMPTY; v[j] = ok ? vals[i] : 0; }
#pragma unroll
    for (int j = 0; j < R; ++j) { s[j] = mix64(k[j]) & mask; }
#pragma unroll
    for (int j = 0; j < R; ++j) { cur[j] = (k[j] != EMPTY) ? *reinterpret_cast<volatile uint64_t*>(tk + s[j]) : EMPTY; }
#pragma unroll
    for (int j = 0; j < R; ++j) {
      if (k[j] == EMPTY) continue;
      if (cur[j] != k[j]) s[j] = foi(tk, 1, mask, k[j]);
      atomicAdd(ts + s[j], (unsigned long long)v[j]);
      atomicAdd(tc + s[j], 1ull);
    }
  }
}

// V2: AoS 32-byte slots, R rows in flight, two REDs per row
template <int R>
__global__ void __launch_bounds__(256) v2(Slot* t, uint64_t mask,
                                          const uint64_t* __restrict__ keys, const int64_t* __restrict__ vals, int64_t n) {
  int64_t base = ((int64_t)blockIdx.x * blockDim.x) * R + threadIdx.x, st = (int64_t)gridDim.x * blockDim.x * R;
  uint64_t* tk = (uint64_t*)t;
  for (; base < n; base += st) {
    uint64_t k[R]; int64_t v[R]; uint64_t s[R]; uint64_t cur[R];
#pragma unroll
    for (int j = 0; j < R; ++j) { int64_t i = base + (int64_t)j * blockDim.x; bool ok = i < n; k[j] = ok ? keys[i] : EMPTY; v[j] = ok ? vals[i] : 0; }
#pragma unroll
    for (int j = 0; j < R; ++j) { s[j] = mix64(k[j]) & mask; }
#pragma unroll
    for (int j = 0; j < R; ++j) { cur[j] = (k[j] != EMPTY) ? *reinterpret_cast<volatile uint64_t*>(tk + s[j] * 4) : EMPTY; }
#pragma unroll
    for (int j = 0; j < R; ++j) {
      if (k[j] == EMPTY) continue;
      if (cur[j] != k[j]) s[j] = foi(tk, 4, mask, k[j]);
      atomicAdd(&t[s[j]].sum, (unsigned long long)v[j]);
      atomicAdd(&t[s[j]].cnt, 1ull);
    }
  }
}

// V3: AoS, lane-paired: one RED instruction carries (sum, count) of 16 rows; the two lanes of a
// pair address the same 32-byte sector.
template <int R>
__global__ void __launch_bounds__(256) v3(Slot* t, uint64_t mask,
                                          const uint64_t* __restrict__ keys, const int64_t* __restrict__ vals, int64_t n) {
  int64_t base = ((int64_t)blockIdx.x * blockDim.x) * R + threadIdx.x, st = (int64_t)gridDim.x * blockDim.x * R;
  uint64_t* tk = (uint64_t*)t;
  const int lane = threadIdx.x & 31;
  for (; base - threadIdx.x < n; base += st) {      // warp-uniform trip count
    uint64_t k[R]; int64_t v[R]; uint64_t s[R]; uint64_t cur[R];
#pragma unroll
    for (int j = 0; j < R; ++j) { int64_t i = base + (int64_t)j * blockDim.x; bool ok = i < n; k[j] = ok ? keys[i] : EMPTY; v[j] = ok ? vals[i] : 0; }
#pragma unroll
    for (int j = 0; j < R; ++j) { s[j] = mix64(k[j]) & mask; }
#pragma unroll
    for (int j = 0; j < R; ++j) { cur[j] = (k[j] != EMPTY) ? *reinterpret_cast<volatile uint64_t*>(tk + s[j] * 4) : EMPTY; }
#pragma unroll
    for (int j = 0; j < R; ++j) {
      if (k[j] != EMPTY && cur[j] != k[j]) s[j] = foi(tk, 4, mask, k[j]);
      if (k[j] == EMPTY) s[j] = EMPTY;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        int src = h * 16 + (lane >> 1);
        uint64_t ss = __shfl_sync(0xffffffffu, s[j], src);
        int64_t vv = __shfl_sync(0xffffffffu, v[j], src);
        if (ss != EMPTY) {
          unsigned long long* a = (lane & 1) ? &t[ss].cnt : &t[ss].sum;
          atomicAdd(a, (lane & 1) ? 1ull : (unsigned long long)vv);
        }
      }
    }
  }
}

// V6: dense direct-index tables (what index_add_ does): 2 REDs per row, no probe
__global__ void __launch_bounds__(256) v6(unsigned long long* ts, unsigned long long* tc,
                                          const uint64_t* __restrict__ keys, const int64_t* __restrict__ vals, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, st = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += st) {
    uint64_t s = keys[i];
    atomicAdd(ts + s, (unsigned long long)vals[i]);
    atomicAdd(tc + s, 1ull);
  }
}
// V7: one RED per row only (count) into the dense table: the raw RED ceiling
__global__ void __launch_bounds__(256) v7(unsigned long long* tc, const uint64_t* __restrict__ keys, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, st = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += st) atomicAdd(tc + keys[i], 1ull);
}
// V8: probe only (AoS), no REDs: cost of the key lookups
template <int R>
__global__ void __launch_bounds__(256) v8(Slot* t, uint64_t mask, const uint64_t* __restrict__ keys, int64_t n, unsigned long long* sink) {
  int64_t base = ((int64_t)blockIdx.x * blockDim.x) * R + threadIdx.x, st = (int64_t)gridDim.x * blockDim.x * R;
  uint64_t* tk = (uint64_t*)t; unsigned long long acc = 0;
  for (; base < n; base += st) {
#pragma unroll
    for (int j = 0; j < R; ++j) { int64_t i = base + (int64_t)j * blockDim.x; if (i < n) { uint64_t k = keys[i]; acc += foi(tk, 4, mask, k); } }
  }
  if (acc == 0x1234567) *sink = acc;
}

__global__ void check_soa(const uint64_t* k, const unsigned long long* s, const unsigned long long* c, uint64_t cap, unsigned long long* out) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  unsigned long long a = 0, b = 0, g = 0;
  for (; i < cap; i += (uint64_t)gridDim.x * blockDim.x) if (k[i] != EMPTY) { a += s[i]; b += c[i]; g++; }
  atomicAdd(out, a); atomicAdd(out + 1, b); atomicAdd(out + 2, g);
}
__global__ void check_aos(const Slot* t, uint64_t cap, unsigned long long* out) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  unsigned long long a = 0, b = 0, g = 0;
  for (; i < cap; i += (uint64_t)gridDim.x * blockDim.x) if (t[i].key != EMPTY) { a += t[i].sum; b += t[i].cnt; g++; }
  atomicAdd(out, a); atomicAdd(out + 1, b); atomicAdd(out + 2, g);
}

int main(int argc, char** argv) {
  int64_t n = argc > 1 ? atoll(argv[1]) : 125000000;
  uint64_t nkeys = argc > 2 ? strtoull(argv[2], 0, 10) : (1ull << 20);
  uint64_t cap = 1; while (cap < 2 * nkeys) cap <<= 1;
  uint64_t mask = cap - 1;
  uint64_t* keys; int64_t* vals; CK(cudaMalloc(&keys, n * 8)); CK(cudaMalloc(&vals, n * 8));
  gen<<<2048, 256>>>(keys, vals, n, nkeys);
  uint64_t* tk; unsigned long long *ts, *tc, *out; Slot* t;
  CK(cudaMalloc(&tk, cap * 8)); CK(cudaMalloc(&ts, cap * 8)); CK(cudaMalloc(&tc, cap * 8)); CK(cudaMalloc(&t, cap * 32)); CK(cudaMalloc(&out, 64));
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  int sms = 148;
  auto run = [&](const char* name, int kind, auto launch) {
    float best = 1e30f; unsigned long long h[3];
    for (int rep = 0; rep < 4; ++rep) {
      if (kind == 0) init_soa<<<1024, 256>>>(tk, ts, tc, cap); else init_aos<<<1024, 256>>>(t, cap);
      if (kind == 2) { CK(cudaMemset(ts, 0, cap * 8)); CK(cudaMemset(tc, 0, cap * 8)); CK(cudaMemset(tk, 0, cap * 8)); }
      CK(cudaDeviceSynchronize());
      cudaEventRecord(e0); launch(); cudaEventRecord(e1); CK(cudaDeviceSynchronize());
      float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    CK(cudaMemset(out, 0, 64));
    if (kind == 0 || kind == 2) check_soa<<<1024, 256>>>(tk, ts, tc, cap, out); else check_aos<<<1024, 256>>>(t, cap, out);
    CK(cudaMemcpy(h, out, 24, cudaMemcpyDeviceToHost));
    printf("%-28s %8.3f ms  %7.2f Grows/s   sum=%llu cnt=%llu groups=%llu\n", name, best, n / (best * 1e6), h[0], h[1], h[2]);
  };
  int g1 = sms * 8;
  run("v0 soa r1", 0, [&] { v0<<<g1, 256>>>(tk, ts, tc, mask, keys, vals, n); });
  run("v1 soa r4", 0, [&] { v1<4><<<g1, 256>>>(tk, ts, tc, mask, keys, vals, n); });
  run("v1 soa r8", 0, [&] { v1<8><<<g1, 256>>>(tk, ts, tc, mask, keys, vals, n); });
  run("v2 aos r1", 1, [&] { v2<1><<<g1, 256>>>(t, mask, keys, vals, n); });
  run("v2 aos r4", 1, [&] { v2<4><<<g1, 256>>>(t, mask, keys, vals, n); });
  run("v2 aos r8", 1, [&] { v2<8><<<g1, 256>>>(t, mask, keys, vals, n); });
  run("v3 aos paired r1", 1, [&] { v3<1><<<g1, 256>>>(t, mask, keys, vals, n); });
  run("v3 aos paired r4", 1, [&] { v3<4><<<g1, 256>>>(t, mask, keys, vals, n); });
  run("v3 aos paired r8", 1, [&] { v3<8><<<g1, 256>>>(t, mask, keys, vals, n); });
  run("v6 dense 2 RED", 2, [&] { v6<<<g1, 256>>>(ts, tc, keys, vals, n); });
  run("v7 dense 1 RED", 2, [&] { v7<<<g1, 256>>>(tc, keys, n); });
  run("v8 probe only r4", 1, [&] { v8<4><<<g1, 256>>>(t, mask, keys, n, out); });
  // bigger grids
  run("v2 aos r4 grid x2", 1, [&] { v2<4><<<g1 * 2, 256>>>(t, mask, keys, vals, n); });
  run("v3 aos paired r4 grid x2", 1, [&] { v3<4><<<g1 * 2, 256>>>(t, mask, keys, vals, n); });
  return 0;
}

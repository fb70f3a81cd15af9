// Micro-benchmark 3: bucketised key array (4 keys = one 32-byte sector per probe, 256-bit loads)
// + 16-byte {sum,count} value pairs updated by lane-paired REDs.
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
    keys[i] = (h % nkeys) * 0x9E3779B97F4A7C15ull;
    vals[i] = (int64_t)((h >> 40) % 1000);
  }
}
__global__ void init_tbl(uint64_t* k, unsigned long long* v, uint64_t cap) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i < cap; i += (uint64_t)gridDim.x * blockDim.x) { k[i] = EMPTY; v[2 * i] = 0; v[2 * i + 1] = 0; }
}
__device__ __forceinline__ void ld_bucket(const uint64_t* p, uint64_t (&k)[4]) {
  asm volatile("ld.global.cg.v4.u64 {%0, %1, %2, %3}, [%4];" : "=l"(k[0]), "=l"(k[1]), "=l"(k[2]), "=l"(k[3]) : "l"(p) : "memory");
}
// resolve `key` given the already-loaded first bucket; returns slot index
__device__ __forceinline__ uint64_t resolve(uint64_t* tk, uint64_t nbuckets, uint64_t b, uint64_t (&k)[4], uint64_t key) {
  for (;;) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (k[j] == key) return b * 4 + j;
      if (k[j] == EMPTY) {
        unsigned long long old = atomicCAS((unsigned long long*)(tk + b * 4 + j), EMPTY, (unsigned long long)key);
        if (old == EMPTY || old == key) return b * 4 + j;
      }
    }
    if (++b == nbuckets) b = 0;
    ld_bucket(tk + b * 4, k);
  }
}
template <int R, bool PAIRED>
__global__ void __launch_bounds__(256) bucket_agg(uint64_t* tk, unsigned long long* tv, uint64_t nbuckets,
                                                  const uint64_t* __restrict__ keys, const int64_t* __restrict__ vals, int64_t n) {
  int64_t base = (int64_t)blockIdx.x * blockDim.x * R, st = (int64_t)gridDim.x * blockDim.x * R;
  const int lane = threadIdx.x & 31;
  for (; base < n; base += st) {
    uint64_t key[R], b[R], s[R]; int64_t v[R]; uint64_t kk[R][4];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      int64_t i = base + (int64_t)r * blockDim.x + threadIdx.x;
      bool ok = i < n;
      key[r] = ok ? __ldcs((const unsigned long long*)keys + i) : EMPTY;
      v[r] = ok ? (int64_t)__ldcs((const unsigned long long*)vals + i) : 0;
    }
#pragma unroll
    for (int r = 0; r < R; ++r) { b[r] = __umul64hi(mix64(key[r]), nbuckets); if (key[r] != EMPTY) ld_bucket(tk + b[r] * 4, kk[r]); }
#pragma unroll
    for (int r = 0; r < R; ++r) {
      s[r] = (key[r] != EMPTY) ? resolve(tk, nbuckets, b[r], kk[r], key[r]) : EMPTY;
      if (PAIRED) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          int src = h * 16 + (lane >> 1);
          uint64_t ss = __shfl_sync(0xffffffffu, s[r], src);
          int64_t vv = __shfl_sync(0xffffffffu, v[r], src);
          if (ss != EMPTY) atomicAdd(tv + ss * 2 + (lane & 1), (lane & 1) ? 1ull : (unsigned long long)vv);
        }
      } else if (s[r] != EMPTY) {
        atomicAdd(tv + s[r] * 2, (unsigned long long)v[r]);
        atomicAdd(tv + s[r] * 2 + 1, 1ull);
      }
    }
  }
}
// ceiling: no probe, paired REDs straight to the hashed slot
__global__ void __launch_bounds__(256) dense_paired(unsigned long long* tv, uint64_t cap,
                                                    const uint64_t* __restrict__ keys, const int64_t* __restrict__ vals, int64_t n) {
  int64_t base = (int64_t)blockIdx.x * blockDim.x, st = (int64_t)gridDim.x * blockDim.x;
  const int lane = threadIdx.x & 31;
  for (; base < n; base += st) {
    int64_t i = base + threadIdx.x;
    uint64_t s = EMPTY; int64_t v = 0;
    if (i < n) { s = __umul64hi(mix64(__ldcs((const unsigned long long*)keys + i)), cap); v = (int64_t)__ldcs((const unsigned long long*)vals + i); }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      int src = h * 16 + (lane >> 1);
      uint64_t ss = __shfl_sync(0xffffffffu, s, src);
      int64_t vv = __shfl_sync(0xffffffffu, v, src);
      if (ss != EMPTY) atomicAdd(tv + ss * 2 + (lane & 1), (lane & 1) ? 1ull : (unsigned long long)vv);
    }
  }
}
// ceiling 2: one row per lane, a single 16-byte-aligned pair, but sum only (1 RED/row)
__global__ void __launch_bounds__(256) dense_one(unsigned long long* tv, uint64_t cap,
                                                 const uint64_t* __restrict__ keys, const int64_t* __restrict__ vals, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, st = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += st) {
    uint64_t s = __umul64hi(mix64(__ldcs((const unsigned long long*)keys + i)), cap);
    atomicAdd(tv + s * 2, (unsigned long long)__ldcs((const unsigned long long*)vals + i));
  }
}
__global__ void check(const uint64_t* k, const unsigned long long* v, uint64_t cap, unsigned long long* out) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; unsigned long long a = 0, b = 0, g = 0;
  for (; i < cap; i += (uint64_t)gridDim.x * blockDim.x) { if (k[i] != EMPTY) g++; a += v[2 * i]; b += v[2 * i + 1]; }
  atomicAdd(out, a); atomicAdd(out + 1, b); atomicAdd(out + 2, g);
}
int main(int argc, char** argv) {
  int64_t n = argc > 1 ? atoll(argv[1]) : 125000000;
  uint64_t nkeys = argc > 2 ? strtoull(argv[2], 0, 10) : (1ull << 20);
  double lf = argc > 3 ? atof(argv[3]) : 2.0;
  uint64_t nb = ((uint64_t)(nkeys * lf) + 3) / 4 + 16, cap = nb * 4;
  uint64_t* keys; int64_t* vals; CK(cudaMalloc(&keys, n * 8)); CK(cudaMalloc(&vals, n * 8));
  gen<<<2048, 256>>>(keys, vals, n, nkeys);
  uint64_t* tk; unsigned long long *tv, *out;
  CK(cudaMalloc(&tk, cap * 8)); CK(cudaMalloc(&tv, cap * 16)); CK(cudaMalloc(&out, 64));
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  printf("n=%lld nkeys=%llu cap=%llu (x%.2f) table %.1f MB\n", (long long)n, (unsigned long long)nkeys, (unsigned long long)cap, lf, cap * 24 / 1e6);
  auto run = [&](const char* name, auto launch) {
    float best = 1e30f; unsigned long long h[3];
    for (int rep = 0; rep < 4; ++rep) {
      init_tbl<<<1024, 256>>>(tk, tv, cap); CK(cudaDeviceSynchronize());
      cudaEventRecord(e0); launch(); cudaEventRecord(e1); CK(cudaDeviceSynchronize());
      float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    CK(cudaMemset(out, 0, 64)); check<<<1024, 256>>>(tk, tv, cap, out);
    CK(cudaMemcpy(h, out, 24, cudaMemcpyDeviceToHost));
    printf("  %-26s %8.3f ms  %7.2f Grows/s   sum=%llu cnt=%llu groups=%llu\n", name, best, n / (best * 1e6), h[0], h[1], h[2]);
  };
  int g = 148 * 8;
  run("dense paired", [&] { dense_paired<<<g, 256>>>(tv, cap, keys, vals, n); });
  run("dense one RED", [&] { dense_one<<<g, 256>>>(tv, cap, keys, vals, n); });
  run("bucket paired r1", [&] { bucket_agg<1, true><<<g, 256>>>(tk, tv, nb, keys, vals, n); });
  run("bucket paired r2", [&] { bucket_agg<2, true><<<g, 256>>>(tk, tv, nb, keys, vals, n); });
  run("bucket paired r4", [&] { bucket_agg<4, true><<<g, 256>>>(tk, tv, nb, keys, vals, n); });
  run("bucket unpaired r1", [&] { bucket_agg<1, false><<<g, 256>>>(tk, tv, nb, keys, vals, n); });
  run("bucket paired r1 g16", [&] { bucket_agg<1, true><<<g * 2, 256>>>(tk, tv, nb, keys, vals, n); });
  run("bucket paired r2 g4", [&] { bucket_agg<2, true><<<g / 2, 256>>>(tk, tv, nb, keys, vals, n); });
  return 0;
}

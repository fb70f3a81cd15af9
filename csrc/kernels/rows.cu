// Row/record kernels around the sort path: key hashing (K14), range sampler + separator pick
// (K3), big-endian key extraction, tie repair for keys wider than 64 bits, row gather by
// permutation, order verification.
//
// Reference counterparts: DryadLinqSampler.Phase1Sampling / RangeSamplerCore
// (LinqToDryad/DryadLinqSampler.cs:42-106, 139-246), key hashing through IEqualityComparer
// (LinqToDryad/DryadLinqVertex.cs:4801) and Hash64 (LinqToDryad/Hash64.cs), the record movement
// that Array.Sort does implicitly on object references (DryadLinqVertex.cs:9321-9818).
#include "common.cuh"
#include "../../include/dryad_b200.h"

namespace dryad { namespace rows {

// ---- K14 ------------------------------------------------------------------------------------
__global__ void hash_u64_kernel(const uint64_t* __restrict__ keys, uint64_t* __restrict__ out,
                                int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) out[i] = mix64(keys[i]);
}

// Variable-length byte keys (offsets has n+1 entries).  FNV-1a over bytes, then mix64, so that
// equal strings hash equally on host (oracle) and device.
__global__ void hash_bytes_kernel(const unsigned char* __restrict__ bytes,
                                  const int64_t* __restrict__ offsets, uint64_t* __restrict__ out,
                                  int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    int64_t b = offsets[i], e = offsets[i + 1];
    uint64_t h = 0xcbf29ce484222325ULL;
    for (int64_t j = b; j < e; ++j) { h ^= bytes[j]; h *= 0x100000001b3ULL; }
    out[i] = mix64(h);
  }
}

// ---- big-endian key extraction from fixed-width rows ----------------------------------------
__global__ void extract_be64_kernel(const unsigned char* __restrict__ rowsp, int64_t n, int width,
                                    int key_off, uint64_t* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const bool aligned = ((reinterpret_cast<uintptr_t>(rowsp) | (uintptr_t)width | (uintptr_t)key_off) & 3) == 0;
  for (; i < n; i += stride) {
    const unsigned char* k = rowsp + (size_t)i * width + key_off;
    uint64_t v = 0;
    if (aligned) {          // two word loads + byte swaps instead of eight byte loads
      const uint32_t* w = reinterpret_cast<const uint32_t*>(k);
      v = ((uint64_t)__byte_perm(__ldg(w), 0, 0x0123) << 32) | (uint64_t)__byte_perm(__ldg(w + 1), 0, 0x0123);
    } else {
#pragma unroll
      for (int b = 0; b < 8; ++b) v = (v << 8) | (uint64_t)k[b];
    }
    out[i] = v;
  }
}

// ---- K3 sampler -----------------------------------------------------------------------------
// Bernoulli(rate) over row indices with a counter-based generator keyed by (seed, index): a
// retried stage emits the same sample (the reference seeds from the vertex id for the same
// reason, DryadLinqSampler.cs:46-53).
__global__ void sample_keys_kernel(const uint64_t* __restrict__ keys, int64_t n, uint64_t seed,
                                   uint64_t threshold, uint64_t* __restrict__ out,
                                   unsigned long long* __restrict__ out_count, int64_t cap) {
  int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t base = i0 - threadIdx.x % 32; base < n; base += stride) {
    int64_t i = base + (threadIdx.x & 31);
    bool take = false;
    if (i < n) take = mix64(seed * 0x9E3779B97F4A7C15ULL + (uint64_t)i + 1) < threshold;
    uint32_t m = __ballot_sync(0xffffffffu, take);
    if (m) {
      unsigned long long pos = 0;
      if (lane_id() == 0) pos = atomicAdd(out_count, (unsigned long long)__popc(m));
      pos = __shfl_sync(0xffffffffu, pos, 0);
      if (take) {
        int64_t o = (int64_t)pos + __popc(m & lanemask_lt());
        if (o < cap) out[o] = keys[i];
      }
    }
  }
}

// Same, keys read big-endian out of rows (no materialised key column needed).
__global__ void sample_rows_kernel(const unsigned char* __restrict__ rowsp, int width, int key_off,
                                   int64_t n, uint64_t seed, uint64_t threshold,
                                   uint64_t* __restrict__ out,
                                   unsigned long long* __restrict__ out_count, int64_t cap) {
  int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t base = i0 - threadIdx.x % 32; base < n; base += stride) {
    int64_t i = base + (threadIdx.x & 31);
    bool take = false;
    if (i < n) take = mix64(seed * 0x9E3779B97F4A7C15ULL + (uint64_t)i + 1) < threshold;
    uint32_t m = __ballot_sync(0xffffffffu, take);
    if (m) {
      unsigned long long pos = 0;
      if (lane_id() == 0) pos = atomicAdd(out_count, (unsigned long long)__popc(m));
      pos = __shfl_sync(0xffffffffu, pos, 0);
      if (take) {
        int64_t o = (int64_t)pos + __popc(m & lanemask_lt());
        if (o < cap) {
          const unsigned char* k = rowsp + (size_t)i * width + key_off;
          uint64_t v = 0;
#pragma unroll
          for (int b = 0; b < 8; ++b) v = (v << 8) | (uint64_t)k[b];
          out[o] = v;
        }
      }
    }
  }
}

// Separator pick from ascending-sorted samples (RangeSamplerCore, DryadLinqSampler.cs:185-245).
__global__ void select_separators_kernel(const uint64_t* __restrict__ sorted, int64_t cnt,
                                         int pcount, int descending, uint64_t* __restrict__ seps) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= pcount - 1) return;
  if (cnt == 0) { seps[i] = descending ? 0ull : ~0ull; return; }
  if (cnt < pcount) {
    if (descending) seps[i] = (i < cnt) ? sorted[cnt - 1 - i] : sorted[0];
    else seps[i] = (i < cnt) ? sorted[i] : sorted[cnt - 1];
    return;
  }
  int64_t intv = cnt / pcount;
  seps[i] = descending ? sorted[cnt - intv * (i + 1)] : sorted[intv * (i + 1)];
}

// Same pick with the sample count still on the device: cnt = sum over r of min(counts[r], cap_each).
// `sorted` holds every rank's fixed-size sample slab (unused slots pre-filled with ~0, so they sort
// behind the real samples) — no host round trip between sampling and the shuffle.
__global__ void select_separators_dev_kernel(const uint64_t* __restrict__ sorted, const int64_t* __restrict__ counts,
                                             int ncounts, int64_t cap_each, int pcount, int descending,
                                             uint64_t* __restrict__ seps) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= pcount - 1) return;
  int64_t cnt = 0;
  for (int r = 0; r < ncounts; ++r) { int64_t c = counts[r]; cnt += c < cap_each ? c : cap_each; }
  if (cnt == 0) { seps[i] = descending ? 0ull : ~0ull; return; }
  if (cnt < pcount) {
    if (descending) seps[i] = (i < cnt) ? sorted[cnt - 1 - i] : sorted[0];
    else seps[i] = (i < cnt) ? sorted[i] : sorted[cnt - 1];
    return;
  }
  int64_t intv = cnt / pcount;
  seps[i] = descending ? sorted[cnt - intv * (i + 1)] : sorted[intv * (i + 1)];
}

// ---- tie repair -----------------------------------------------------------------------------
// After sorting by the 64-bit key prefix, rows whose prefixes collide are ordered by the
// remaining `extra` key bytes (then by original index, keeping the sort stable).  One thread per
// run of equal prefixes; runs are short for any realistic key distribution, and a run longer than
// `max_run` is reported instead of being processed quadratically.
__global__ void fix_ties_kernel(const uint64_t* __restrict__ keys, uint32_t* __restrict__ idx,
                                int64_t n, const unsigned char* __restrict__ rowsp, int width,
                                int extra_off, int extra, int max_run, int descending,
                                uint64_t prefix_mask, int32_t* __restrict__ long_runs) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    uint64_t k = keys[i] & prefix_mask;
    if (i + 1 >= n || (keys[i + 1] & prefix_mask) != k) continue;
    if (i > 0 && (keys[i - 1] & prefix_mask) == k) continue;
    int64_t e = i + 1;
    while (e < n && (keys[e] & prefix_mask) == k) ++e;
    int64_t len = e - i;
    if (len > max_run) { atomicAdd(long_runs, 1); continue; }
    // insertion sort on (extra bytes, idx)
    for (int64_t a = i + 1; a < e; ++a) {
      uint32_t ia = idx[a];
      const unsigned char* ra = rowsp + (size_t)ia * width + extra_off;
      int64_t b = a - 1;
      while (b >= i) {
        uint32_t ib = idx[b];
        const unsigned char* rb = rowsp + (size_t)ib * width + extra_off;
        int cmp = 0;
        for (int t = 0; t < extra && cmp == 0; ++t) cmp = (int)rb[t] - (int)ra[t];
        if (descending) cmp = -cmp;
        if (cmp == 0) cmp = (ib > ia) ? 1 : -1;
        if (cmp <= 0) break;
        idx[b + 1] = ib;
        --b;
      }
      idx[b + 1] = ia;
    }
  }
}

// ---- row gather -----------------------------------------------------------------------------
// out[i] = in[idx[i]] for fixed-width rows (width % 4 == 0).  A warp moves 32 rows at a time;
// lane j carries word (it*32 + j) of the 32-row group so the store side is fully coalesced and
// the load side reads each row as one contiguous burst.
__global__ void __launch_bounds__(256) gather_rows_kernel(const uint32_t* __restrict__ in,
                                                          const uint32_t* __restrict__ idx,
                                                          uint32_t* __restrict__ out, int64_t n,
                                                          int words) {
  const int lane = threadIdx.x & 31;
  int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  int64_t groups = (n + 31) / 32;
  for (int64_t g = warp; g < groups; g += nwarps) {
    int64_t r = g * 32 + lane;
    uint32_t my = (r < n) ? idx[r] : 0u;
    int rows_here = (int)min((int64_t)32, n - g * 32);
    int total = rows_here * words;
    uint32_t* o = out + (size_t)g * 32 * words;
    for (int j0 = 0; j0 < total; j0 += 32) {   // warp-uniform trip count: shfl needs all lanes
      int j = j0 + lane;
      int jc = j < total ? j : total - 1;
      int row = jc / words, w = jc - row * words;
      uint32_t src = __shfl_sync(0xffffffffu, my, row);
      if (j < total) o[j] = in[(size_t)src * words + w];
    }
  }
}

// Compile-time row width: all WORDS loads of a 32-row group are issued before the first store, so
// each warp keeps WORDS independent 128-byte requests in flight (memory-level parallelism).
template <int WORDS>
__global__ void __launch_bounds__(256) gather_rows_fixed_kernel(const uint32_t* __restrict__ in,
                                                                const uint32_t* __restrict__ idx,
                                                                uint32_t* __restrict__ out, int64_t n) {
  const int lane = threadIdx.x & 31;
  int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  int64_t groups = n / 32;                      // full groups only; the tail uses the generic kernel
  for (int64_t g = warp; g < groups; g += nwarps) {
    const uint32_t my = idx[g * 32 + lane];
    uint32_t v[WORDS];
#pragma unroll
    for (int it = 0; it < WORDS; ++it) {
      const int j = it * 32 + lane;
      const int row = j / WORDS, w = j - row * WORDS;
      const uint32_t src = __shfl_sync(0xffffffffu, my, row);
      v[it] = __ldg(in + (size_t)src * WORDS + w);
    }
    uint32_t* o = out + (size_t)g * 32 * WORDS;
#pragma unroll
    for (int it = 0; it < WORDS; ++it) o[it * 32 + lane] = v[it];
  }
}

// Window of the big-endian key: out32[i] = (be64(row key) >> shift) as uint32.
__global__ void extract_be_window_kernel(const unsigned char* __restrict__ rowsp, int64_t n, int width, int key_off,
                                         int shift, uint32_t* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    const unsigned char* k = rowsp + (size_t)i * width + key_off;
    uint64_t v = 0;
#pragma unroll
    for (int b = 0; b < 8; ++b) v = (v << 8) | (uint64_t)k[b];
    out[i] = (uint32_t)(v >> shift);
  }
}
__global__ void shift_u64_to_u32_kernel(const uint64_t* __restrict__ in, int64_t n, int shift, uint32_t* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) out[i] = (uint32_t)(in[i] >> shift);
}

// Tie repair over 32-bit sorted window keys (see fix_ties_kernel).
__global__ void fix_ties32_kernel(const uint32_t* __restrict__ keys, uint32_t* __restrict__ idx, int64_t n,
                                  const unsigned char* __restrict__ rowsp, int width, int extra_off, int extra,
                                  int max_run, int descending, int32_t* __restrict__ long_runs) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    const uint32_t k = keys[i];
    if (i + 1 >= n || keys[i + 1] != k) continue;
    if (i > 0 && keys[i - 1] == k) continue;
    int64_t e = i + 1;
    while (e < n && keys[e] == k) ++e;
    if (e - i > max_run) { atomicAdd(long_runs, 1); continue; }
    for (int64_t a = i + 1; a < e; ++a) {
      uint32_t ia = idx[a];
      const unsigned char* ra = rowsp + (size_t)ia * width + extra_off;
      int64_t b = a - 1;
      while (b >= i) {
        uint32_t ib = idx[b];
        const unsigned char* rb = rowsp + (size_t)ib * width + extra_off;
        int cmp = 0;
        for (int t = 0; t < extra && cmp == 0; ++t) cmp = (int)rb[t] - (int)ra[t];
        if (descending) cmp = -cmp;
        if (cmp == 0) cmp = (ib > ia) ? 1 : -1;
        if (cmp <= 0) break;
        idx[b + 1] = ib;
        --b;
      }
      idx[b + 1] = ia;
    }
  }
}

// Tie repair that needs no rows: runs of equal 32-bit windows are ordered by the full 64-bit key
// prefix of each slot (keys64[idx]), then by slot (stable).  Pairs whose 64-bit prefixes are equal
// cannot be decided without the remaining key bytes: they are counted in `unresolved` when the key
// is wider than the prefix (`wide_key`), so the caller can fall back to the row-based repair.
__global__ void fix_ties32_keys_kernel(const uint32_t* __restrict__ keys, uint32_t* __restrict__ idx, int64_t n,
                                       const uint64_t* __restrict__ keys64, int max_run, int wide_key,
                                       int32_t* __restrict__ long_runs, int32_t* __restrict__ unresolved) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    const uint32_t k = keys[i];
    if (i + 1 >= n || keys[i + 1] != k) continue;
    if (i > 0 && keys[i - 1] == k) continue;
    int64_t e = i + 1;
    while (e < n && keys[e] == k) ++e;
    if (e - i > max_run) { atomicAdd(long_runs, 1); continue; }
    int eq = 0;
    for (int64_t a = i + 1; a < e; ++a) {
      const uint32_t ia = idx[a];
      const uint64_t ka = keys64[ia];
      int64_t b = a - 1;
      while (b >= i) {
        const uint32_t ib = idx[b];
        const uint64_t kb = keys64[ib];
        if (kb == ka) eq = 1;
        if (kb < ka || (kb == ka && ib < ia)) break;
        idx[b + 1] = ib;
        --b;
      }
      idx[b + 1] = ia;
    }
    if (eq && wide_key) atomicAdd(unresolved, 1);
  }
}

// Element gather for 4- and 8-byte columns.
template <typename T>
__global__ void gather_elems_kernel(const T* __restrict__ in, const uint32_t* __restrict__ idx,
                                    T* __restrict__ out, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) out[i] = in[idx[i]];
}

// ---- verification ---------------------------------------------------------------------------
// result[0] = number of adjacent pairs out of order on the first key_len bytes (bytewise),
// result[1] = wrapping sum of the 64-bit key prefixes (a permutation-invariant checksum).
__global__ void check_sorted_rows_kernel(const unsigned char* __restrict__ rowsp, int64_t n,
                                         int width, int key_off, int key_len, int descending,
                                         unsigned long long* __restrict__ result) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  unsigned long long bad = 0, sum = 0;
  for (; i < n; i += stride) {
    const unsigned char* a = rowsp + (size_t)i * width + key_off;
    uint64_t v = 0;
    for (int b = 0; b < 8 && b < key_len; ++b) v = (v << 8) | (uint64_t)a[b];
    sum += v;
    if (i + 1 < n) {
      const unsigned char* c = a + width;
      int cmp = 0;
      for (int b = 0; b < key_len && cmp == 0; ++b) cmp = (int)a[b] - (int)c[b];
      if (descending) cmp = -cmp;
      if (cmp > 0) ++bad;
    }
  }
  bad = warp_sum(bad); sum = warp_sum(sum);
  if (lane_id() == 0) { if (bad) atomicAdd(&result[0], bad); atomicAdd(&result[1], sum); }
}

static inline int grid_for(int64_t n, int threads, int per_thread = 4) {
  int64_t g = (n + (int64_t)threads * per_thread - 1) / ((int64_t)threads * per_thread);
  if (g < 1) g = 1;
  if (g > 148 * 16) g = 148 * 16;
  return (int)g;
}

}}  // namespace dryad::rows

extern "C" {
using namespace dryad::rows;

int dryad_hash_u64(const uint64_t* keys, uint64_t* out, int64_t n, void* stream) {
  if (n <= 0) return DRYAD_OK;
  hash_u64_kernel<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(keys, out, n);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_hash_bytes(const uint8_t* bytes, const int64_t* offsets, uint64_t* out, int64_t n,
                     void* stream) {
  if (n <= 0) return DRYAD_OK;
  hash_bytes_kernel<<<grid_for(n, 256, 1), 256, 0, (cudaStream_t)stream>>>(bytes, offsets, out, n);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_extract_be64(const uint8_t* rowsp, int64_t n, int width, int key_off, uint64_t* out,
                       void* stream) {
  if (n <= 0) return DRYAD_OK;
  if (width <= 0 || key_off < 0 || key_off + 8 > width) return DRYAD_ERR_BAD_ARG;
  extract_be64_kernel<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(rowsp, n, width, key_off, out);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_sample_keys(const uint64_t* keys, int64_t n, uint64_t seed, double rate, uint64_t* out,
                      unsigned long long* out_count, int64_t cap, void* stream) {
  if (n <= 0) return DRYAD_OK;
  if (!(rate >= 0.0 && rate <= 1.0)) return DRYAD_ERR_BAD_ARG;
  uint64_t thr = rate >= 1.0 ? ~0ull : (uint64_t)(rate * 18446744073709551616.0);
  sample_keys_kernel<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(keys, n, seed, thr, out, out_count, cap);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_sample_rows(const uint8_t* rowsp, int width, int key_off, int64_t n, uint64_t seed,
                      double rate, uint64_t* out, unsigned long long* out_count, int64_t cap,
                      void* stream) {
  if (n <= 0) return DRYAD_OK;
  if (!(rate >= 0.0 && rate <= 1.0) || key_off < 0 || key_off + 8 > width) return DRYAD_ERR_BAD_ARG;
  uint64_t thr = rate >= 1.0 ? ~0ull : (uint64_t)(rate * 18446744073709551616.0);
  sample_rows_kernel<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(rowsp, width, key_off, n, seed, thr, out, out_count, cap);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_select_separators(const uint64_t* sorted, int64_t cnt, int pcount, int descending,
                            uint64_t* seps, void* stream) {
  if (pcount < 1) return DRYAD_ERR_BAD_ARG;
  if (pcount == 1) return DRYAD_OK;
  select_separators_kernel<<<(pcount + 254) / 256, 256, 0, (cudaStream_t)stream>>>(sorted, cnt, pcount, descending, seps);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_select_separators_dev(const uint64_t* sorted, const int64_t* counts, int ncounts, int64_t cap_each,
                                int pcount, int descending, uint64_t* seps, void* stream) {
  if (pcount < 1 || ncounts < 1 || cap_each < 0) return DRYAD_ERR_BAD_ARG;
  if (pcount == 1) return DRYAD_OK;
  select_separators_dev_kernel<<<(pcount + 254) / 256, 256, 0, (cudaStream_t)stream>>>(sorted, counts, ncounts, cap_each, pcount, descending, seps);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_fix_ties(const uint64_t* keys, uint32_t* idx, int64_t n, const uint8_t* rowsp, int width,
                   int extra_off, int extra, int max_run, int descending, uint64_t prefix_mask,
                   int32_t* long_runs, void* stream) {
  if (n <= 1 || extra <= 0) return DRYAD_OK;
  if (extra_off < 0 || extra_off + extra > width) return DRYAD_ERR_BAD_ARG;
  fix_ties_kernel<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(keys, idx, n, rowsp, width, extra_off, extra, max_run, descending, prefix_mask, long_runs);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_gather_rows(const void* in, const uint32_t* idx, void* out, int64_t n, int width,
                      void* stream) {
  if (n <= 0) return DRYAD_OK;
  if (width <= 0 || (width & 3)) return DRYAD_ERR_BAD_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  if (width == 4) gather_elems_kernel<uint32_t><<<grid_for(n, 256), 256, 0, st>>>((const uint32_t*)in, idx, (uint32_t*)out, n);
  else if (width == 8) gather_elems_kernel<uint64_t><<<grid_for(n, 256), 256, 0, st>>>((const uint64_t*)in, idx, (uint64_t*)out, n);
  else {
    int64_t groups = (n + 31) / 32;
    int64_t blocks = (groups + 7) / 8;
    if (blocks > 148 * 32) blocks = 148 * 32;
    const int words = width / 4;
    int64_t full = (n / 32) * 32;
    if (words == 25 && full > 0) {                 // 100-byte records: unrolled fast path + generic tail
      gather_rows_fixed_kernel<25><<<(int)blocks, 256, 0, st>>>((const uint32_t*)in, idx, (uint32_t*)out, n);
      DRYAD_LAUNCH_CHECK();
      if (n > full)
        gather_rows_kernel<<<1, 256, 0, st>>>((const uint32_t*)in, idx + full, (uint32_t*)out + (size_t)full * words, n - full, words);
    } else {
      gather_rows_kernel<<<(int)blocks, 256, 0, st>>>((const uint32_t*)in, idx, (uint32_t*)out, n, words);
    }
  }
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_extract_be_window(const uint8_t* rowsp, int64_t n, int width, int key_off, int shift, uint32_t* out, void* stream) {
  if (n <= 0) return DRYAD_OK;
  if (width <= 0 || key_off < 0 || key_off + 8 > width || shift < 0 || shift > 32) return DRYAD_ERR_BAD_ARG;
  extract_be_window_kernel<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(rowsp, n, width, key_off, shift, out);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_shift_u64_to_u32(const uint64_t* in, int64_t n, int shift, uint32_t* out, void* stream) {
  if (n <= 0) return DRYAD_OK;
  if (shift < 0 || shift > 32) return DRYAD_ERR_BAD_ARG;
  shift_u64_to_u32_kernel<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(in, n, shift, out);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_fix_ties32(const uint32_t* keys, uint32_t* idx, int64_t n, const uint8_t* rowsp, int width, int extra_off,
                     int extra, int max_run, int descending, int32_t* long_runs, void* stream) {
  if (n <= 1 || extra <= 0) return DRYAD_OK;
  if (extra_off < 0 || extra_off + extra > width) return DRYAD_ERR_BAD_ARG;
  fix_ties32_kernel<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(keys, idx, n, rowsp, width, extra_off, extra, max_run, descending, long_runs);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_fix_ties32_keys(const uint32_t* keys, uint32_t* idx, int64_t n, const uint64_t* keys64, int max_run,
                          int wide_key, int32_t* long_runs, int32_t* unresolved, void* stream) {
  if (n <= 1) return DRYAD_OK;
  fix_ties32_keys_kernel<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(keys, idx, n, keys64, max_run, wide_key, long_runs, unresolved);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_check_sorted_rows(const uint8_t* rowsp, int64_t n, int width, int key_off, int key_len,
                            int descending, unsigned long long* result2, void* stream) {
  if (n <= 0) return DRYAD_OK;
  if (key_off < 0 || key_len <= 0 || key_off + key_len > width) return DRYAD_ERR_BAD_ARG;
  check_sorted_rows_kernel<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(rowsp, n, width, key_off, key_len, descending, result2);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

}  // extern "C"

// K4: LSD radix sort of (key, u32 value) pairs — one-sweep with decoupled look-back.
//
// Replaces the reference's per-partition Sort / ParallelSort (2 Mi-element chunks of Array.Sort
// followed by 16-way merges; LinqToDryad/DryadLinqVertex.cs:293-316, 9321-9818) with a GPU sort
// that reads and writes every pair exactly once per 8-bit digit:
//   1. rs_histogram   : one pass over the keys builds the histograms of ALL digits.
//   2. rs_scan        : exclusive scan of each 256-bin histogram -> global digit bases.
//   3. rs_onesweep    : per digit, a single kernel that ranks a tile in shared memory, publishes
//                       its per-digit counts, resolves its global offsets by looking back at the
//                       preceding tiles' published counts, and scatters through shared memory so
//                       global writes are contiguous runs per digit.
// No CUB / thrust: the north star forbids a library sort on this path.
#include "common.cuh"
#include "../../include/dryad_b200.h"
#include <mutex>

namespace dryad { namespace rs {

constexpr int RADIX_BITS = 8;
constexpr int RADIX = 1 << RADIX_BITS;
constexpr int THREADS = 512;
constexpr int WARPS = THREADS / 32;
constexpr int ITEMS = 12;
constexpr int TILE = THREADS * ITEMS;

constexpr uint32_t FLAG_AGG = 1u << 30;
constexpr uint32_t FLAG_INC = 2u << 30;
constexpr uint32_t VAL_MASK = (1u << 30) - 1;

template <typename K>
__device__ __forceinline__ uint32_t digit_of(K key, int shift, uint32_t mask) {
  return (uint32_t)(key >> shift) & mask;
}

// ---- 1. all-digit histogram ------------------------------------------------------------------
template <typename K>
__global__ void __launch_bounds__(512) rs_histogram(const K* __restrict__ keys, int64_t n,
                                                    uint32_t* __restrict__ hist, int begin_bit,
                                                    int npasses, int end_bit) {
  extern __shared__ uint32_t sh[];  // [npasses][RADIX]
  for (int i = threadIdx.x; i < npasses * RADIX; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    K k = keys[i];
#pragma unroll 1
    for (int p = 0; p < npasses; ++p) {
      int shift = begin_bit + p * RADIX_BITS;
      int bits = min(RADIX_BITS, end_bit - shift);
      atomicAdd(&sh[p * RADIX + digit_of(k, shift, (1u << bits) - 1)], 1u);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < npasses * RADIX; i += blockDim.x)
    if (sh[i]) atomicAdd(&hist[i], sh[i]);
}

// ---- 2. exclusive scan of each digit histogram -----------------------------------------------
__global__ void __launch_bounds__(RADIX) rs_scan(const uint32_t* __restrict__ hist,
                                                 uint32_t* __restrict__ base) {
  __shared__ uint32_t s[RADIX];
  int p = blockIdx.x, t = threadIdx.x;
  uint32_t v = hist[p * RADIX + t];
  s[t] = v;
  __syncthreads();
  for (int o = 1; o < RADIX; o <<= 1) {
    uint32_t a = (t >= o) ? s[t - o] : 0;
    __syncthreads();
    s[t] += a;
    __syncthreads();
  }
  base[p * RADIX + t] = s[t] - v;
}

// ---- 3. one-sweep pass -----------------------------------------------------------------------
template <typename K, bool HAS_VALS>
__global__ void __launch_bounds__(THREADS, 2)
rs_onesweep(const K* __restrict__ keys_in, K* __restrict__ keys_out,
            const uint32_t* __restrict__ vals_in, uint32_t* __restrict__ vals_out, int64_t n,
            int shift, uint32_t mask, int iota_vals, const uint32_t* __restrict__ digit_base,
            uint32_t* __restrict__ status, uint32_t* __restrict__ tile_counter) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  K* s_keys = reinterpret_cast<K*>(smem_raw);                               // [TILE]
  uint32_t* s_vals = reinterpret_cast<uint32_t*>(s_keys + TILE);            // [TILE] (if vals)
  uint32_t* s_warp_hist = s_vals + (HAS_VALS ? TILE : 0);                   // [WARPS][RADIX]
  uint32_t* s_digit_start = s_warp_hist + WARPS * RADIX;                    // [RADIX]
  int64_t* s_global_off = reinterpret_cast<int64_t*>(s_digit_start + RADIX);// [RADIX]
  __shared__ uint32_t s_tile;
  __shared__ uint32_t s_warp_tot[RADIX / 32];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_tile = atomicAdd(tile_counter, 1u);
  for (int i = tid; i < WARPS * RADIX; i += THREADS) s_warp_hist[i] = 0;
  __syncthreads();
  const uint32_t tile = s_tile;
  const int64_t tile_base = (int64_t)tile * TILE;
  const int valid = (int)min((int64_t)TILE, n - tile_base);

  // Warp-striped load: item i of lane l lives at warp*ITEMS*32 + i*32 + l.
  K key[ITEMS];
  uint32_t val[ITEMS];
  uint32_t rank[ITEMS];
  const int wbase = warp * ITEMS * 32 + lane;
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) {
    int idx = wbase + i * 32;
    key[i] = (idx < valid) ? keys_in[tile_base + idx] : (K)~(K)0;
    if (HAS_VALS) {
      if (iota_vals) val[i] = (uint32_t)(tile_base + idx);
      else val[i] = (idx < valid) ? vals_in[tile_base + idx] : 0u;
    }
  }

  // Warp-local stable ranking per digit (match_any).
  uint32_t* my_hist = s_warp_hist + warp * RADIX;
  const uint32_t lt = lanemask_lt();
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) {
    uint32_t d = digit_of(key[i], shift, mask);
    uint32_t peers = __match_any_sync(0xffffffffu, d);
    uint32_t prev = my_hist[d];
    rank[i] = prev + __popc(peers & lt);
    __syncwarp();
    if ((peers >> lane) <= 1u) my_hist[d] = prev + __popc(peers);  // highest lane of the group
    __syncwarp();
  }
  __syncthreads();

  // Per digit: exclusive scan over warps, tile count, then block scan over digits.
  uint32_t count = 0;
  if (tid < RADIX) {
    uint32_t sum = 0;
#pragma unroll
    for (int w = 0; w < WARPS; ++w) {
      uint32_t c = s_warp_hist[w * RADIX + tid];
      s_warp_hist[w * RADIX + tid] = sum;
      sum += c;
    }
    count = sum;
    // publish the tile aggregate as early as possible
    st_relaxed_gpu_u32(&status[(size_t)tile * RADIX + tid],
                       (tile == 0 ? FLAG_INC : FLAG_AGG) | count);
    uint32_t incl = count;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t a = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += a;
    }
    if (lane == 31) s_warp_tot[warp] = incl;
    s_digit_start[tid] = incl - count;  // warp-local exclusive for now
  }
  __syncthreads();
  if (tid < RADIX) {
    uint32_t add = 0;
#pragma unroll
    for (int w = 0; w < RADIX / 32; ++w) add += (w < warp) ? s_warp_tot[w] : 0;
    uint32_t dstart = s_digit_start[tid] + add;
    s_digit_start[tid] = dstart;

    // Decoupled look-back over preceding tiles for this digit.
    uint32_t excl = 0;
    if (tile > 0) {
      int64_t look = (int64_t)tile - 1;
      while (true) {
        uint32_t v = ld_relaxed_gpu_u32(&status[(size_t)look * RADIX + tid]);
        uint32_t f = v & ~VAL_MASK;
        if (f == 0) continue;
        excl += v & VAL_MASK;
        if (f == FLAG_INC) break;
        --look;
      }
      st_relaxed_gpu_u32(&status[(size_t)tile * RADIX + tid], FLAG_INC | (excl + count));
    }
    s_global_off[tid] = (int64_t)digit_base[tid] + (int64_t)excl - (int64_t)dstart;
  }
  __syncthreads();

  // Place into tile-sorted order in shared memory.
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) {
    uint32_t d = digit_of(key[i], shift, mask);
    uint32_t p = s_digit_start[d] + my_hist[d] + rank[i];
    s_keys[p] = key[i];
    if (HAS_VALS) s_vals[p] = val[i];
  }
  __syncthreads();

  // Contiguous runs per digit -> coalesced global writes.
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) {
    int j = tid + i * THREADS;
    if (j < valid) {
      K k = s_keys[j];
      int64_t o = s_global_off[digit_of(k, shift, mask)] + j;
      keys_out[o] = k;
      if (HAS_VALS) vals_out[o] = s_vals[j];
    }
  }
}

template <typename K>
static size_t smem_bytes(bool has_vals) {
  return sizeof(K) * TILE + (has_vals ? 4 * TILE : 0) + 4 * WARPS * RADIX + 4 * RADIX + 8 * RADIX;
}

struct Layout {
  int npasses;
  int64_t tiles;
  size_t off_hist, off_base, off_counter, off_status, total;
};

static Layout layout(int64_t n, int begin_bit, int end_bit) {
  Layout L;
  L.npasses = (end_bit - begin_bit + RADIX_BITS - 1) / RADIX_BITS;
  L.tiles = (n + TILE - 1) / TILE;
  if (L.tiles < 1) L.tiles = 1;
  size_t o = 0;
  L.off_hist = o;    o += (size_t)L.npasses * RADIX * 4;
  L.off_base = o;    o += (size_t)L.npasses * RADIX * 4;
  L.off_counter = o; o += 256;  // one u32 per pass, padded
  L.off_status = o;  o += (size_t)L.tiles * RADIX * 4;
  L.total = (o + 255) & ~(size_t)255;
  return L;
}

template <typename K>
static int sort_impl(K* keys, K* keys_alt, uint32_t* vals, uint32_t* vals_alt, int64_t n,
                     int begin_bit, int end_bit, int iota_vals, void* scratch,
                     size_t scratch_bytes, int* result_in_alt, cudaStream_t st) {
  if (n < 0 || begin_bit < 0 || end_bit > (int)sizeof(K) * 8 || begin_bit >= end_bit)
    return DRYAD_ERR_BAD_ARG;
  if (n >= (1ll << 30)) return DRYAD_ERR_TOO_LARGE;
  if (result_in_alt) *result_in_alt = 0;
  if (n == 0) return DRYAD_OK;
  Layout L = layout(n, begin_bit, end_bit);
  if (L.npasses > 16) return DRYAD_ERR_BAD_ARG;
  if (scratch_bytes < L.total || scratch == nullptr) return DRYAD_ERR_SCRATCH;
  const bool has_vals = (vals != nullptr);
  unsigned char* sc = (unsigned char*)scratch;
  uint32_t* hist = (uint32_t*)(sc + L.off_hist);
  uint32_t* base = (uint32_t*)(sc + L.off_base);
  uint32_t* counters = (uint32_t*)(sc + L.off_counter);
  uint32_t* status = (uint32_t*)(sc + L.off_status);

  DRYAD_CUDA_TRY(cudaMemsetAsync(sc, 0, L.off_status, st));
  int64_t hg = (n + 512 * 8 - 1) / (512 * 8);
  int hgrid = (int)(hg < 148 * 4 ? hg : 148 * 4);
  rs_histogram<K><<<hgrid, 512, (size_t)L.npasses * RADIX * 4, st>>>(keys, n, hist, begin_bit,
                                                                    L.npasses, end_bit);
  DRYAD_LAUNCH_CHECK();
  rs_scan<<<L.npasses, RADIX, 0, st>>>(hist, base);
  DRYAD_LAUNCH_CHECK();

  size_t smem = smem_bytes<K>(has_vals);
  auto kern_v = rs_onesweep<K, true>;
  auto kern_k = rs_onesweep<K, false>;
  if (has_vals)
    DRYAD_CUDA_TRY(cudaFuncSetAttribute(kern_v, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  else
    DRYAD_CUDA_TRY(cudaFuncSetAttribute(kern_k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));

  K* kin = keys; K* kout = keys_alt;
  uint32_t* vin = vals; uint32_t* vout = vals_alt;
  for (int p = 0; p < L.npasses; ++p) {
    int shift = begin_bit + p * RADIX_BITS;
    int bits = (end_bit - shift) < RADIX_BITS ? (end_bit - shift) : RADIX_BITS;
    uint32_t mask = (1u << bits) - 1;
    DRYAD_CUDA_TRY(cudaMemsetAsync(status, 0, (size_t)L.tiles * RADIX * 4, st));
    if (has_vals)
      kern_v<<<(unsigned)L.tiles, THREADS, smem, st>>>(kin, kout, vin, vout, n, shift, mask,
                                                       (p == 0 && iota_vals) ? 1 : 0,
                                                       base + p * RADIX, status, counters + p);
    else
      kern_k<<<(unsigned)L.tiles, THREADS, smem, st>>>(kin, kout, nullptr, nullptr, n, shift, mask,
                                                       0, base + p * RADIX, status, counters + p);
    DRYAD_LAUNCH_CHECK();
    K* tk = kin; kin = kout; kout = tk;
    uint32_t* tv = vin; vin = vout; vout = tv;
  }
  if (result_in_alt) *result_in_alt = (L.npasses & 1);
  return DRYAD_OK;
}

}}  // namespace dryad::rs

extern "C" {

size_t dryad_radix_sort_scratch_bytes(int64_t n, int begin_bit, int end_bit) {
  if (n <= 0 || end_bit <= begin_bit) return 256;
  return dryad::rs::layout(n, begin_bit, end_bit).total;
}

int dryad_radix_sort_pairs_u64(uint64_t* keys, uint64_t* keys_alt, uint32_t* vals,
                               uint32_t* vals_alt, int64_t n, int begin_bit, int end_bit,
                               int iota_vals, void* scratch, size_t scratch_bytes,
                               int* result_in_alt, void* stream) {
  return dryad::rs::sort_impl<uint64_t>(keys, keys_alt, vals, vals_alt, n, begin_bit, end_bit,
                                        iota_vals, scratch, scratch_bytes, result_in_alt,
                                        (cudaStream_t)stream);
}

int dryad_radix_sort_pairs_u32(uint32_t* keys, uint32_t* keys_alt, uint32_t* vals,
                               uint32_t* vals_alt, int64_t n, int begin_bit, int end_bit,
                               int iota_vals, void* scratch, size_t scratch_bytes,
                               int* result_in_alt, void* stream) {
  return dryad::rs::sort_impl<uint32_t>(keys, keys_alt, vals, vals_alt, n, begin_bit, end_bit,
                                        iota_vals, scratch, scratch_bytes, result_in_alt,
                                        (cudaStream_t)stream);
}

// ---- allocation-free variant ---------------------------------------------------------------------
// Some callers may not allocate device memory at all: a process whose own pool already owns all
// of HBM, a latency-critical thread that must never wait on the driver's allocator (cudaMalloc and
// cudaFree take a device-wide lock and can synchronise the device), or a plain C program that
// links this library without any allocator of its own.  For them the sort's working storage
// (digit histograms, tile counters, look-back status words) lives in a fixed array that is part of
// the loaded module.  Every GPU that loads the library carries these DRYAD_SORT_NOALLOC_SCRATCH
// bytes until the process exits.
static __device__ __align__(256) unsigned char g_sort_noalloc_scratch[DRYAD_SORT_NOALLOC_SCRATCH];

namespace dryad { namespace rs {
static std::mutex g_noalloc_mu;
static void* g_noalloc_ptr[64] = {nullptr};   // device address of the array, per device ordinal
}}

int dryad_radix_sort_pairs_u64_noalloc(uint64_t* keys, uint64_t* keys_alt, uint32_t* vals,
                                       uint32_t* vals_alt, int64_t n, int begin_bit, int end_bit,
                                       int iota_vals, int* result_in_alt, void* stream) {
  using namespace dryad::rs;
  if (n > DRYAD_SORT_NOALLOC_MAX_N) return DRYAD_ERR_TOO_LARGE;   // beyond the built-in scratch
  if (n > 0 && end_bit > begin_bit && layout(n, begin_bit, end_bit).total > DRYAD_SORT_NOALLOC_SCRATCH)
    return DRYAD_ERR_TOO_LARGE;
  int dev = -1;
  DRYAD_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return DRYAD_ERR_BAD_ARG;
  void* scratch = nullptr;
  {
    // One-time set-up per device, with that device current: in host code the array's name is only
    // a handle, so ask the runtime for the address the kernels must use.
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (g_noalloc_ptr[dev] == nullptr) {
      void* p = nullptr;
      DRYAD_CUDA_TRY(cudaGetSymbolAddress(&p, g_sort_noalloc_scratch));
      g_noalloc_ptr[dev] = p;
    }
    scratch = g_noalloc_ptr[dev];
  }
  // sort_impl clears the parts of the scratch that must start at zero (histograms, tile counters,
  // status words) on `stream` before each use, so stale contents of the previous call are harmless.
  return sort_impl<uint64_t>(keys, keys_alt, vals, vals_alt, n, begin_bit, end_bit, iota_vals,
                             scratch, (size_t)DRYAD_SORT_NOALLOC_SCRATCH, result_in_alt,
                             (cudaStream_t)stream);
}

}  // extern "C"

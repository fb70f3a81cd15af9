// K5 (k-way merge of sorted runs, merge-path partitioned) and K9 (merge join on key-sorted inputs).
//
// Reference counterparts: MergeSort / SequentialMergeSort / ParallelMergeSort
// (LinqToDryad/DryadLinqVertex.cs:319-423, 9819-10140, SortHelper.HeapMergeSort :10588) merge the
// sorted outputs of k producer ports; MergeJoin / MergeGroupJoin (:898-1230) co-iterate two
// key-sorted inputs.  On the GPU a 2-way merge is cut into independent tiles by binary searches on
// the merge-path diagonals (every tile knows which slice of A and of B it consumes); k runs are
// merged by a tree of 2-way merges.  Merge join = per outer key the [lower, upper) bound in the
// sorted inner keys, then an exclusive scan and an emit pass.
#include "common.cuh"
#include "../../include/dryad_b200.h"

namespace dryad { namespace mg {

constexpr int THREADS = 256, VT = 8, TILE = THREADS * VT;

// Largest a in [max(0,d-nb), min(d,na)] such that A[a-1] <= B[d-a] (ties take A first: stable).
__device__ __forceinline__ int64_t merge_path(const uint64_t* A, int64_t na, const uint64_t* B, int64_t nb, int64_t d) {
  int64_t lo = d > nb ? d - nb : 0, hi = d < na ? d : na;
  while (lo < hi) {
    int64_t a = (lo + hi) >> 1;
    if (A[a] <= B[d - 1 - a]) lo = a + 1; else hi = a;
  }
  return lo;
}

__global__ void __launch_bounds__(THREADS) merge_pairs_kernel(const uint64_t* __restrict__ ka, const uint32_t* __restrict__ va, int64_t na,
                                                              const uint64_t* __restrict__ kb, const uint32_t* __restrict__ vb, int64_t nb,
                                                              uint64_t* __restrict__ ko, uint32_t* __restrict__ vo) {
  __shared__ uint64_t s_k[TILE + 2];
  __shared__ uint32_t s_v[TILE + 2];
  __shared__ int64_t s_a0, s_a1;
  const int64_t total = na + nb;
  const int64_t d0 = (int64_t)blockIdx.x * TILE;
  const int64_t d1 = min(total, d0 + TILE);
  if (threadIdx.x == 0) s_a0 = merge_path(ka, na, kb, nb, d0);
  if (threadIdx.x == 32) s_a1 = merge_path(ka, na, kb, nb, d1);
  __syncthreads();
  const int64_t a0 = s_a0, a1 = s_a1, b0 = d0 - a0, b1 = d1 - a1;
  const int la = (int)(a1 - a0), lb = (int)(b1 - b0);
  for (int i = threadIdx.x; i < la; i += THREADS) { s_k[i] = ka[a0 + i]; s_v[i] = va ? va[a0 + i] : 0u; }
  for (int i = threadIdx.x; i < lb; i += THREADS) { s_k[la + i] = kb[b0 + i]; s_v[la + i] = vb ? vb[b0 + i] : 0u; }
  __syncthreads();
  const uint64_t* A = s_k; const uint64_t* B = s_k + la;
  const int d = min(threadIdx.x * VT, la + lb);
  int ai = (int)merge_path(A, la, B, lb, d), bi = d - ai;
  uint64_t rk[VT]; uint32_t rv[VT];
#pragma unroll
  for (int i = 0; i < VT; ++i) {
    bool takeA = ai < la && (bi >= lb || A[ai] <= B[bi]);
    int src = takeA ? ai : la + bi;
    rk[i] = (ai < la || bi < lb) ? s_k[src] : 0; rv[i] = (ai < la || bi < lb) ? s_v[src] : 0;
    if (takeA) ++ai; else ++bi;
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < VT; ++i) { s_k[threadIdx.x * VT + i] = rk[i]; s_v[threadIdx.x * VT + i] = rv[i]; }
  __syncthreads();
  for (int i = threadIdx.x; i < (int)(d1 - d0); i += THREADS) { ko[d0 + i] = s_k[i]; if (vo) vo[d0 + i] = s_v[i]; }
}

// For every probe key: [lower, upper) bounds in ascending `sorted`.
__global__ void bounds_kernel(const uint64_t* __restrict__ sorted, int64_t m, const uint64_t* __restrict__ probe, int64_t n,
                              int64_t* __restrict__ lower, int64_t* __restrict__ upper) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    const uint64_t k = probe[i];
    int64_t lo = 0, hi = m;
    while (lo < hi) { int64_t mid = (lo + hi) >> 1; if (sorted[mid] < k) lo = mid + 1; else hi = mid; }
    lower[i] = lo;
    hi = m;
    int64_t l2 = lo;
    while (l2 < hi) { int64_t mid = (l2 + hi) >> 1; if (sorted[mid] <= k) l2 = mid + 1; else hi = mid; }
    upper[i] = l2;
  }
}

// Emit (outer i, inner j) for j in [lower[i], upper[i]) at offsets[i].
__global__ void emit_ranges_kernel(const int64_t* __restrict__ lower, const int64_t* __restrict__ upper,
                                   const int64_t* __restrict__ offsets, int64_t n, int32_t* __restrict__ out_outer,
                                   int32_t* __restrict__ out_inner) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    int64_t o = offsets[i];
    for (int64_t j = lower[i]; j < upper[i]; ++j, ++o) { out_outer[o] = (int32_t)i; out_inner[o] = (int32_t)j; }
  }
}

}}  // namespace dryad::mg

extern "C" {
using namespace dryad::mg;

// Stable 2-way merge of ascending (key, u32 value) runs; values may be NULL.
int dryad_merge_pairs_u64(const uint64_t* ka, const uint32_t* va, int64_t na, const uint64_t* kb, const uint32_t* vb,
                          int64_t nb, uint64_t* ko, uint32_t* vo, void* stream) {
  if (na < 0 || nb < 0) return DRYAD_ERR_BAD_ARG;
  int64_t total = na + nb;
  if (total == 0) return DRYAD_OK;
  int64_t blocks = (total + TILE - 1) / TILE;
  merge_pairs_kernel<<<(unsigned)blocks, THREADS, 0, (cudaStream_t)stream>>>(ka, va, na, kb, vb, nb, ko, vo);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_sorted_bounds(const uint64_t* sorted, int64_t m, const uint64_t* probe, int64_t n, int64_t* lower,
                        int64_t* upper, void* stream) {
  if (n <= 0) return DRYAD_OK;
  int64_t g = (n + 1023) / 1024; if (g > 148 * 16) g = 148 * 16;
  bounds_kernel<<<(int)g, 256, 0, (cudaStream_t)stream>>>(sorted, m, probe, n, lower, upper);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_emit_ranges(const int64_t* lower, const int64_t* upper, const int64_t* offsets, int64_t n,
                      int32_t* out_outer, int32_t* out_inner, void* stream) {
  if (n <= 0) return DRYAD_OK;
  int64_t g = (n + 1023) / 1024; if (g > 148 * 16) g = 148 * 16;
  emit_ranges_kernel<<<(int)g, 256, 0, (cudaStream_t)stream>>>(lower, upper, offsets, n, out_outer, out_inner);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

}  // extern "C"

// K1 + K2 + C1: fused partition-and-scatter ("shuffle") over NVLink peer memory.
//
// Reference behaviour being replaced: every HashPartition / RangePartition vertex computes a
// port per record — (hash & 0x7FFFFFFF) % nPorts (LinqToDryad/DryadLinqVertex.cs:4798-4907) or
// a left-most binary search in the separator keys (DryadLinqVertex.cs:4911-5151,
// DryadLinqUtil.cs:112-139) — and writes the record to one file per consumer; consumers fetch
// P×Q files over HTTP after the producer finished (GraphBuilder.cs:481-504,
// HttpScheduler.cs:65-91).  Here the bucket computation and the all-to-all are one kernel chain
// with no NCCL call and no host round trip:
//
//   sh_count   : per-block histogram of destination partitions.
//   sh_publish : scan over blocks; writes this rank's per-partition totals into EVERY rank's
//                count matrix with P2P stores and raises a release.sys flag per peer.
//   sh_scatter : waits (acquire.sys) for all ranks' totals, derives the exact, stable write
//                offset of every record, groups each tile by destination in shared memory and
//                stores contiguous, 16-byte aligned runs straight into the destination
//                partition buffers — which live in peer GPUs' HBM.
//   sh_barrier : epoch-numbered signal/wait across ranks so consumers may read.
//
// Offsets are exact (count → scan → write), so receive buffers are used densely, the result is
// deterministic and stable (source rank, then source order), and skew can only fail loudly
// (status = overflow) instead of corrupting memory.
#include "common.cuh"
#include "../../include/dryad_b200.h"
#include <stdlib.h>

namespace dryad { namespace sh {

constexpr int THREADS = 256;
constexpr int TILE = 256;
constexpr int WARPS = THREADS / 32;
constexpr int MAX_PARTS = 256;
constexpr int MAX_COLS = 8;
constexpr long long SPIN_CYCLES = 20000000000ll;  // ~10 s at 2 GHz: protocol failure, not a hang

struct Params {
  int ncols, nparts, mode, key_col, key_kind, key_off, emit_key, rank, world;
  int nblocks, tiles_per_block;
  int64_t n;
  const unsigned char* src[MAX_COLS];
  uint32_t width[MAX_COLS];
  void* const* dst_table;
  uint64_t* const* dst_key_table;
  const uint64_t* separators;
  const int64_t* capacity;
  int64_t* counts_local;
  int64_t* const* counts_peers;
  uint64_t* flags_local;
  uint64_t* const* flags_peers;
  uint64_t epoch;
  uint32_t* block_counts;  // [nparts][nblocks]
  int32_t* status;
  int max_width;
  uint64_t* key_out;       // optional: the count pass also writes every row's derived key here
  int filter_mod, filter_rem;   // filter_mod > 0: this scatter pass moves only partitions p with p % filter_mod == filter_rem
};

__device__ __forceinline__ uint64_t load_key(const Params& P, int64_t row) {
  const unsigned char* base = P.src[P.key_col] + (size_t)row * P.width[P.key_col];
  if (P.key_kind == 0) return *reinterpret_cast<const uint64_t*>(base);
  const unsigned char* k = base + P.key_off;
  if ((reinterpret_cast<uintptr_t>(k) & 3) == 0) {
    const uint32_t* w = reinterpret_cast<const uint32_t*>(k);
    return ((uint64_t)__byte_perm(__ldg(w), 0, 0x0123) << 32) | (uint64_t)__byte_perm(__ldg(w + 1), 0, 0x0123);
  }
  uint64_t v = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) v = (v << 8) | (uint64_t)k[i];
  return v;
}

__device__ __forceinline__ uint64_t key_from_smem(const Params& P, const unsigned char* row) {
  if (P.key_kind == 0) return *reinterpret_cast<const uint64_t*>(row);
  const unsigned char* k = row + P.key_off;
  if ((reinterpret_cast<uintptr_t>(k) & 3) == 0) {   // two word loads + byte swaps
    const uint32_t* w = reinterpret_cast<const uint32_t*>(k);
    return ((uint64_t)__byte_perm(w[0], 0, 0x0123) << 32) | (uint64_t)__byte_perm(w[1], 0, 0x0123);
  }
  uint64_t v = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) v = (v << 8) | (uint64_t)k[i];
  return v;
}

// Destination partition of a key.  `seps` points at shared memory (nparts-1 separators).
__device__ __forceinline__ uint32_t part_of(uint64_t key, int mode, int nparts,
                                            const uint64_t* seps) {
  if (mode == DRYAD_PART_HASH) return hash_bucket(key, (uint32_t)nparts);
  if (mode == DRYAD_PART_DIRECT) return (uint32_t)(key % (uint64_t)nparts);
  // Left-most rule: ascending -> #separators strictly below key; descending -> strictly above.
  int lo = 0, hi = nparts - 1;
  const bool desc = (mode == DRYAD_PART_RANGE_DESC);
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    uint64_t s = seps[mid];
    bool go_right = desc ? (s > key) : (s < key);
    if (go_right) lo = mid + 1; else hi = mid;
  }
  return (uint32_t)lo;
}

__global__ void __launch_bounds__(THREADS) sh_count(Params P) {
  __shared__ uint32_t s_cnt[MAX_PARTS];
  __shared__ uint64_t s_seps[MAX_PARTS];
  const int tid = threadIdx.x;
  for (int i = tid; i < P.nparts; i += THREADS) s_cnt[i] = 0;
  if (P.mode == DRYAD_PART_RANGE_ASC || P.mode == DRYAD_PART_RANGE_DESC)
    for (int i = tid; i < P.nparts - 1; i += THREADS) s_seps[i] = P.separators[i];
  __syncthreads();
  int64_t row0 = (int64_t)blockIdx.x * P.tiles_per_block * TILE;
  int64_t row1 = min(P.n, row0 + (int64_t)P.tiles_per_block * TILE);
  for (int64_t r0 = row0; r0 < row1; r0 += THREADS) {
    const int64_t r = r0 + tid;
    const bool in = r < row1;
    const uint32_t amask = __ballot_sync(0xffffffffu, in);
    if (in) {
      const uint64_t key = load_key(P, r);
      if (P.key_out) P.key_out[r] = key;
      uint32_t p = part_of(key, P.mode, P.nparts, s_seps);
      uint32_t peers = __match_any_sync(amask, p);
      if ((peers >> lane_id()) <= 1u) atomicAdd(&s_cnt[p], (uint32_t)__popc(peers));
    }
  }
  __syncthreads();
  for (int i = tid; i < P.nparts; i += THREADS)
    P.block_counts[(size_t)i * P.nblocks + blockIdx.x] = s_cnt[i];
}

// One block per partition: block p turns partition p's per-block counts into exclusive offsets
// (chunked block scan), writes the total into EVERY rank's count matrix with P2P stores; the last
// block to finish raises the release.sys flag on every peer.
__global__ void __launch_bounds__(256) sh_publish(Params P, uint32_t* done_counter) {
  __shared__ uint32_t s_warp[8];
  __shared__ uint64_t s_carry;
  __shared__ bool s_last;
  const int p = blockIdx.x, t = threadIdx.x, lane = t & 31, warp = t >> 5;
  uint32_t* row = P.block_counts + (size_t)p * P.nblocks;
  if (t == 0) s_carry = 0;
  __syncthreads();
  for (int base = 0; base < P.nblocks; base += 256) {
    const int i = base + t;
    const uint32_t v = (i < P.nblocks) ? row[i] : 0u;
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t a = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += a;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t before = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) before += (w < warp) ? s_warp[w] : 0u;
    const uint64_t carry = s_carry;
    if (i < P.nblocks) row[i] = (uint32_t)(carry + before + incl - v);
    __syncthreads();
    if (t == 255) s_carry = carry + before + incl;
    __syncthreads();
  }
  if (t < P.world) P.counts_peers[t][(size_t)P.rank * P.nparts + p] = (int64_t)s_carry;
  __threadfence_system();
  __syncthreads();
  if (t == 0) s_last = atomicAdd(done_counter, 1u) == gridDim.x - 1;
  __syncthreads();
  if (s_last) {
    __threadfence_system();
    if (t < P.world) st_release_sys(&P.flags_peers[t][P.rank], P.epoch);
  }
}

__global__ void __launch_bounds__(THREADS) sh_scatter(Params P) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int vb = (int)blockIdx.x;
  // layout: in[TILE*maxw] | out[TILE*maxw + 32*nparts + 16] | small arrays
  const int maxw = P.max_width;
  unsigned char* s_in = smem;
  unsigned char* s_out = s_in + (size_t)TILE * maxw;
  size_t out_bytes = ((size_t)TILE * maxw + 32 * (size_t)P.nparts + 16 + 15) & ~(size_t)15;
  const int npa = (P.nparts + 3) & ~3;                                      // arrays sized by nparts
  uint64_t* s_seps = reinterpret_cast<uint64_t*>(s_out + out_bytes);        // [npa]
  int64_t* s_gbase = reinterpret_cast<int64_t*>(s_seps + npa);              // [npa] rows
  uint32_t* s_wcnt = reinterpret_cast<uint32_t*>(s_gbase + npa);            // [WARPS][np]
  uint32_t* s_tcnt = s_wcnt + WARPS * npa;                                  // [npa]
  uint32_t* s_tstart = s_tcnt + npa;                                        // [npa+4] rows
  uint32_t* s_ustart = s_tstart + npa + 4;                                  // [npa+4] 16B units
  uint32_t* s_uoff = s_ustart + npa + 4;                                    // [npa] byte offset of run in s_out
  __shared__ int s_abort;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int np = P.nparts;

  // Wait for every rank's totals, then derive this (rank, block)'s base row in each partition.
  if (tid == 0) s_abort = 0;
  __syncthreads();
  if (tid < P.world) {
    if (!spin_until_ge(&P.flags_local[tid], P.epoch, SPIN_CYCLES)) { s_abort = 2; }
  }
  __syncthreads();
  if (s_abort) { if (tid == 0) atomicMax(P.status, 2); return; }
  for (int p = tid; p < np; p += THREADS) {
    int64_t before = 0, total = 0;
    for (int r = 0; r < P.world; ++r) {
      int64_t c = reinterpret_cast<const volatile int64_t*>(P.counts_local)[(size_t)r * np + p];
      if (r < P.rank) before += c;
      total += c;
    }
    if (total > P.capacity[p]) s_abort = 1;
    s_gbase[p] = before + (int64_t)P.block_counts[(size_t)p * P.nblocks + vb];
  }
  if (P.mode == DRYAD_PART_RANGE_ASC || P.mode == DRYAD_PART_RANGE_DESC)
    for (int i = tid; i < np - 1; i += THREADS) s_seps[i] = P.separators[i];
  __syncthreads();
  if (s_abort) { if (tid == 0) atomicMax(P.status, 1); return; }

  const int kc = P.key_col;
  const uint32_t kw = P.width[kc];
  const int total_cols = P.ncols + (P.emit_key ? 1 : 0);

  for (int t = 0; t < P.tiles_per_block; ++t) {
    int64_t row0 = ((int64_t)vb * P.tiles_per_block + t) * TILE;
    if (row0 >= P.n) break;
    const int valid = (int)min((int64_t)TILE, P.n - row0);

    // --- stage the key column, compute destination + stable rank ------------------------------
    {
      const unsigned char* g = P.src[kc] + (size_t)row0 * kw;
      const int words = valid * (int)(kw / 4);
      const uint32_t* g4 = reinterpret_cast<const uint32_t*>(g);
      uint32_t* s4 = reinterpret_cast<uint32_t*>(s_in);
      if ((reinterpret_cast<uintptr_t>(g) & 15) == 0) {
        const int v16 = words / 4;
        const uint4* g16 = reinterpret_cast<const uint4*>(g);
        uint4* s16 = reinterpret_cast<uint4*>(s_in);
        for (int i = tid; i < v16; i += THREADS) s16[i] = ld_stream_v4(g16 + i);
        for (int i = v16 * 4 + tid; i < words; i += THREADS) s4[i] = g4[i];
      } else {
        for (int i = tid; i < words; i += THREADS) s4[i] = g4[i];
      }
    }
    for (int i = tid; i < WARPS * np; i += THREADS) s_wcnt[i] = 0;
    __syncthreads();

    uint64_t key = 0; uint32_t part = 0, rank = 0;
    bool active = tid < valid;
    if (active) {
      key = key_from_smem(P, s_in + (size_t)tid * kw);
      part = part_of(key, P.mode, np, s_seps);
      if (P.filter_mod > 0 && (int)(part % (uint32_t)P.filter_mod) != P.filter_rem) active = false;   // another pass moves it
    }
    {
      uint32_t amask = __ballot_sync(0xffffffffu, active);
      if (active) {
        uint32_t peers = __match_any_sync(amask, part);
        rank = __popc(peers & lanemask_lt());
        if ((peers >> lane) <= 1u) s_wcnt[warp * np + part] = __popc(peers);
      }
    }
    __syncthreads();
    // per partition: exclusive scan over warps + tile count
    for (int p = tid; p < np; p += THREADS) {
      uint32_t sum = 0;
#pragma unroll
      for (int w = 0; w < WARPS; ++w) { uint32_t c = s_wcnt[w * np + p]; s_wcnt[w * np + p] = sum; sum += c; }
      s_tcnt[p] = sum;
    }
    __syncthreads();
    if (tid == 0) {  // np <= 256: a serial scan is cheap next to the tile's 25 KB of traffic
      uint32_t acc = 0;
      for (int p = 0; p < np; ++p) { s_tstart[p] = acc; acc += s_tcnt[p]; }
      s_tstart[np] = acc;
    }
    __syncthreads();
    if (active) rank += s_tstart[part] + s_wcnt[warp * np + part];  // tile-sorted row index
    const uint32_t rank_in_part = active ? rank - s_tstart[part] : 0;

    // --- per column: group by destination in shared memory, then aligned run stores ---------
    for (int ci = 0; ci < total_cols; ++ci) {
      // key column first (it is already staged), then the others, then the derived-key output
      const bool is_key_out = (ci == P.ncols);
      const int c = is_key_out ? P.ncols : (ci == 0 ? kc : (ci <= kc ? ci - 1 : ci));
      const uint32_t w = is_key_out ? 8u : P.width[c];
      if (ci > 0) __syncthreads();  // previous column's run stores are done with s_out/s_u*
      // unit layout for this column: run p starts at a shared offset congruent to its global
      // destination address mod 16, in its own 16-byte units.
      if (tid == 0) {
        uint32_t unit = 0;
        for (int p = 0; p < np; ++p) {
          s_ustart[p] = unit;
          uint32_t cnt = s_tcnt[p];
          if (cnt == 0) { s_uoff[p] = unit * 16; continue; }
          unsigned char* dbase = is_key_out ? reinterpret_cast<unsigned char*>(P.dst_key_table[p])
                                            : reinterpret_cast<unsigned char*>(P.dst_table[(size_t)c * np + p]);
          uintptr_t a = reinterpret_cast<uintptr_t>(dbase) + (uintptr_t)s_gbase[p] * w;
          uint32_t mis = (uint32_t)(a & 15);
          s_uoff[p] = unit * 16 + mis;
          unit += (mis + cnt * w + 15) / 16;
        }
        s_ustart[np] = unit;
      }
      if (!is_key_out && ci > 0) {
        const unsigned char* g = P.src[c] + (size_t)row0 * w;
        const int words = valid * (int)(w / 4);
        const uint32_t* g4 = reinterpret_cast<const uint32_t*>(g);
        uint32_t* s4 = reinterpret_cast<uint32_t*>(s_in);
        if ((reinterpret_cast<uintptr_t>(g) & 15) == 0) {
          const int v16 = words / 4;
          const uint4* g16 = reinterpret_cast<const uint4*>(g);
          uint4* s16 = reinterpret_cast<uint4*>(s_in);
          for (int i = tid; i < v16; i += THREADS) s16[i] = ld_stream_v4(g16 + i);
          for (int i = v16 * 4 + tid; i < words; i += THREADS) s4[i] = g4[i];
        } else {
          for (int i = tid; i < words; i += THREADS) s4[i] = g4[i];
        }
      }
      __syncthreads();
      // permute rows into destination-grouped order
      if (active) {
        uint32_t* d = reinterpret_cast<uint32_t*>(s_out + s_uoff[part] + (size_t)rank_in_part * w);
        if (is_key_out) {
          d[0] = (uint32_t)key; d[1] = (uint32_t)(key >> 32);
        } else {
          const uint32_t* s = reinterpret_cast<const uint32_t*>(s_in + (size_t)tid * w);
          const int ww = (int)(w / 4);
          for (int k = 0; k < ww; ++k) d[k] = s[k];
        }
      }
      __syncthreads();
      // write out: warp w streams the runs of partitions w, w+8, ...  (16-byte aligned units)
      for (int p = warp; p < np; p += WARPS) {
        const uint32_t cnt = s_tcnt[p];
        if (cnt == 0) continue;
        unsigned char* dbase = is_key_out ? reinterpret_cast<unsigned char*>(P.dst_key_table[p])
                                          : reinterpret_cast<unsigned char*>(P.dst_table[(size_t)c * np + p]);
        const uint32_t run_s = s_uoff[p];                    // shared byte offset of run start
        const uint32_t run_e = run_s + cnt * w;              // exclusive
        const uintptr_t ga = reinterpret_cast<uintptr_t>(dbase) + (uintptr_t)s_gbase[p] * w;
        const uint32_t u0 = run_s >> 4, u1 = (run_e + 15) >> 4;
        for (uint32_t u = u0 + lane; u < u1; u += 32) {
          const uint32_t ub = u * 16;
          const uintptr_t gu = ga + (intptr_t)ub - (intptr_t)run_s;
          if (ub >= run_s && ub + 16 <= run_e) {
            *reinterpret_cast<uint4*>(gu) = *reinterpret_cast<const uint4*>(s_out + ub);
          } else {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              uint32_t bb = ub + 4 * k;
              if (bb >= run_s && bb < run_e)
                *reinterpret_cast<uint32_t*>(gu + 4 * k) = *reinterpret_cast<const uint32_t*>(s_out + bb);
            }
          }
        }
      }
    }
    __syncthreads();
    for (int p = tid; p < np; p += THREADS) s_gbase[p] += s_tcnt[p];
    __syncthreads();
  }
  __threadfence_system();
}

// ---- TMA-pipelined scatter (single payload column, the wide-record path) ----------------------
// Same semantics as sh_scatter, restructured so neither the HBM reads nor the NVLink writes stall
// the SM: tile t+1 is fetched by a bulk async copy (cp.async.bulk global->shared, mbarrier
// complete_tx) while tile t is ranked and permuted, and the destination runs of tile t leave
// through bulk async stores (cp.async.bulk shared->global, to local or PEER HBM) that drain in the
// background while the CTA already works on the next tile (double-buffered in/out tiles).
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
  uint32_t done;
  do {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(addr), "r"(parity) : "memory");
  } while (!done);
}
__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_store(void* gdst, const void* smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
               ::"l"(gdst), "r"(smem_u32(smem_src)), "r"(bytes) : "memory");
}

__global__ void __launch_bounds__(THREADS) sh_scatter_tma(Params P) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int np = P.nparts;
  const int npa = (np + 3) & ~3;
  const uint32_t w = P.width[0];
  const size_t in_bytes = ((size_t)TILE * w + 127) & ~(size_t)127;
  const size_t out_bytes = ((size_t)TILE * w + 32 * (size_t)np + 16 + 127) & ~(size_t)127;
  const size_t kout_bytes = P.emit_key ? (((size_t)TILE * 8 + 32 * (size_t)np + 16 + 127) & ~(size_t)127) : 0;
  unsigned char* s_in = smem;                                   // [2][in_bytes]
  unsigned char* s_out = s_in + 2 * in_bytes;                   // [2][out_bytes]
  unsigned char* s_kout = s_out + 2 * out_bytes;                // [2][kout_bytes]
  uint64_t* s_seps = reinterpret_cast<uint64_t*>(s_kout + 2 * kout_bytes);
  int64_t* s_gbase = reinterpret_cast<int64_t*>(s_seps + npa);
  uint32_t* s_wcnt = reinterpret_cast<uint32_t*>(s_gbase + npa);
  uint32_t* s_tcnt = s_wcnt + WARPS * npa;
  uint32_t* s_tstart = s_tcnt + npa;
  uint32_t* s_uoff = s_tstart + npa + 4;                         // payload run offsets in s_out
  uint32_t* s_koff = s_uoff + npa;                               // key run offsets in s_kout
  __shared__ uint64_t s_full[2];
  __shared__ int s_abort;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) { s_abort = 0; mbar_init(&s_full[0], 1); mbar_init(&s_full[1], 1);
                  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  __syncthreads();

  // Persistent form: the grid may be smaller than the number of row blocks the count pass used
  // (a capped CTA count leaves SMs to kernels on other streams while this link-bound kernel
  // runs); every CTA walks the virtual blocks vb = blockIdx.x, blockIdx.x + gridDim.x, ...
  const int64_t ntiles_total = (P.n + TILE - 1) / TILE;
  uint32_t g = 0;                       // tiles processed by this CTA so far: buffer = g & 1
  auto issue_load = [&](int64_t tile, uint32_t gi) {
    const int64_t row0 = tile * TILE;
    const int valid = (int)min((int64_t)TILE, P.n - row0);
    const uint32_t bytes = ((uint32_t)valid * w) & ~15u;
    uint64_t* bar = &s_full[gi & 1];
    mbar_expect_tx(bar, bytes);
    if (bytes) bulk_load(s_in + (gi & 1) * in_bytes, P.src[0] + (size_t)row0 * w, bytes, bar);
  };
  auto tiles_of = [&](int vb) {
    int mt = (int)min((int64_t)P.tiles_per_block, ntiles_total - (int64_t)vb * P.tiles_per_block);
    return mt < 0 ? 0 : mt;
  };
  // prefetch the first tile while waiting for the peers' counts
  const int vb0 = (int)blockIdx.x;
  const bool have_first = vb0 < P.nblocks && tiles_of(vb0) > 0;
  if (tid == 0 && have_first) issue_load((int64_t)vb0 * P.tiles_per_block, 0);

  if (tid < P.world) {
    if (!spin_until_ge(&P.flags_local[tid], P.epoch, SPIN_CYCLES)) { s_abort = 2; }
  }
  __syncthreads();
  if (s_abort) { if (tid == 0) atomicMax(P.status, 2); if (tid == 0 && have_first) mbar_wait(&s_full[0], 0); return; }
  for (int p = tid; p < np; p += THREADS) {
    int64_t total = 0;
    for (int r = 0; r < P.world; ++r) total += reinterpret_cast<const volatile int64_t*>(P.counts_local)[(size_t)r * np + p];
    if (total > P.capacity[p]) s_abort = 1;
  }
  if (P.mode == DRYAD_PART_RANGE_ASC || P.mode == DRYAD_PART_RANGE_DESC)
    for (int i = tid; i < np - 1; i += THREADS) s_seps[i] = P.separators[i];
  __syncthreads();
  if (s_abort) { if (tid == 0) atomicMax(P.status, 1); if (tid == 0 && have_first) mbar_wait(&s_full[0], 0); return; }

  for (int vb = vb0; vb < P.nblocks; vb += gridDim.x) {
  const int64_t tile0 = (int64_t)vb * P.tiles_per_block;
  const int my_tiles = tiles_of(vb);
  if (vb != vb0 && tid == 0 && my_tiles > 0) issue_load(tile0, g);   // in[g&1] was released at the end of the previous tile
  for (int p = tid; p < np; p += THREADS) {
    int64_t before = 0;
    for (int r = 0; r < P.rank; ++r) before += reinterpret_cast<const volatile int64_t*>(P.counts_local)[(size_t)r * np + p];
    s_gbase[p] = before + (int64_t)P.block_counts[(size_t)p * P.nblocks + vb];
  }
  __syncthreads();

  for (int t = 0; t < my_tiles; ++t, ++g) {
    const int buf = g & 1;
    const int64_t row0 = (tile0 + t) * TILE;
    const int valid = (int)min((int64_t)TILE, P.n - row0);
    unsigned char* in = s_in + buf * in_bytes;
    unsigned char* out = s_out + buf * out_bytes;
    unsigned char* kout = s_kout + buf * kout_bytes;
    if (tid == 0 && t + 1 < my_tiles) issue_load(tile0 + t + 1, g + 1);       // in[buf^1] was released at the end of the previous tile
    mbar_wait(&s_full[buf], (g >> 1) & 1);
    {   // tail bytes that are not a multiple of 16 (last partial tile only)
      const uint32_t total_b = (uint32_t)valid * w, bulk_b = total_b & ~15u;
      for (uint32_t b = bulk_b + 4 * tid; b < total_b; b += 4 * THREADS)
        *reinterpret_cast<uint32_t*>(in + b) = *reinterpret_cast<const uint32_t*>(P.src[0] + (size_t)row0 * w + b);
    }
    for (int i = tid; i < WARPS * np; i += THREADS) s_wcnt[i] = 0;
    __syncthreads();

    uint64_t key = 0; uint32_t part = 0, rank = 0;
    bool active = tid < valid;
    if (active) {
      key = key_from_smem(P, in + (size_t)tid * w); part = part_of(key, P.mode, np, s_seps);
      if (P.filter_mod > 0 && (int)(part % (uint32_t)P.filter_mod) != P.filter_rem) active = false;   // another pass moves it
    }
    {
      uint32_t amask = __ballot_sync(0xffffffffu, active);
      if (active) {
        uint32_t peers = __match_any_sync(amask, part);
        rank = __popc(peers & lanemask_lt());
        if ((peers >> lane) <= 1u) s_wcnt[warp * np + part] = __popc(peers);
      }
    }
    __syncthreads();
    for (int p = tid; p < np; p += THREADS) {
      uint32_t sum = 0;
#pragma unroll
      for (int wp = 0; wp < WARPS; ++wp) { uint32_t c = s_wcnt[wp * np + p]; s_wcnt[wp * np + p] = sum; sum += c; }
      s_tcnt[p] = sum;
    }
    __syncthreads();
    if (tid == 0) {
      // the bulk stores that read out[buf] two tiles ago must have finished reading shared memory
      asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
      uint32_t acc = 0, off = 0, koff = 0;
      for (int p = 0; p < np; ++p) {
        s_tstart[p] = acc; const uint32_t cnt = s_tcnt[p]; acc += cnt;
        unsigned char* dbase = reinterpret_cast<unsigned char*>(P.dst_table[p]);
        uint32_t mis = (uint32_t)((reinterpret_cast<uintptr_t>(dbase) + (uintptr_t)s_gbase[p] * w) & 15);
        s_uoff[p] = off + mis;
        off += cnt ? ((mis + cnt * w + 15) & ~15u) : 0;
        if (P.emit_key) {
          uint32_t kmis = (uint32_t)((reinterpret_cast<uintptr_t>(P.dst_key_table[p]) + (uintptr_t)s_gbase[p] * 8) & 15);
          s_koff[p] = koff + kmis;
          koff += cnt ? ((kmis + cnt * 8 + 15) & ~15u) : 0;
        }
      }
      s_tstart[np] = acc;
    }
    __syncthreads();
    if (active) {
      const uint32_t rip = rank + s_wcnt[warp * np + part];        // rank inside the destination run
      uint32_t* d = reinterpret_cast<uint32_t*>(out + s_uoff[part] + (size_t)rip * w);
      const uint32_t* sp = reinterpret_cast<const uint32_t*>(in + (size_t)tid * w);
      const int ww = (int)(w / 4);
      for (int k = 0; k < ww; ++k) d[k] = sp[k];
      if (P.emit_key) {
        uint32_t* kd = reinterpret_cast<uint32_t*>(kout + s_koff[part] + (size_t)rip * 8);
        kd[0] = (uint32_t)key; kd[1] = (uint32_t)(key >> 32);
      }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic smem writes -> async proxy
    __syncthreads();
    // stores: heads/tails by lanes, 16-byte aligned bodies by bulk async copies (one issuing thread)
    for (int p = warp; p < np; p += WARPS) {
      const uint32_t cnt = s_tcnt[p];
      if (cnt == 0) continue;
      for (int col = 0; col < (P.emit_key ? 2 : 1); ++col) {
        const uint32_t cw = col ? 8u : w;
        unsigned char* sbase = col ? kout : out;
        const uint32_t run_s = col ? s_koff[p] : s_uoff[p], run_e = run_s + cnt * cw;
        unsigned char* dbase = col ? reinterpret_cast<unsigned char*>(P.dst_key_table[p]) : reinterpret_cast<unsigned char*>(P.dst_table[p]);
        const uintptr_t ga = reinterpret_cast<uintptr_t>(dbase) + (uintptr_t)s_gbase[p] * cw;
        const uint32_t body_s = (run_s + 15) & ~15u, body_e = run_e & ~15u;
        // head words [run_s, body_s) and tail words [body_e, run_e): at most 3 + 3 words
        const uint32_t b = (lane < 4) ? run_s + 4 * lane : body_e + 4 * (lane - 4);
        const bool in_head = lane < 4 && b < body_s && b < run_e;
        const bool in_tail = lane >= 4 && lane < 8 && b < run_e && b >= body_s;
        if (body_e <= body_s) {                                   // run shorter than one aligned unit
          for (uint32_t bb = run_s + 4 * lane; bb < run_e; bb += 128)
            *reinterpret_cast<uint32_t*>(ga + (bb - run_s)) = *reinterpret_cast<const uint32_t*>(sbase + bb);
        } else {
          if (in_head || in_tail) *reinterpret_cast<uint32_t*>(ga + (b - run_s)) = *reinterpret_cast<const uint32_t*>(sbase + b);
          if (lane == 0) bulk_store(reinterpret_cast<void*>(ga + (body_s - run_s)), sbase + body_s, body_e - body_s);
        }
      }
    }
    __syncwarp();
    if (lane == 0) asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    __syncthreads();
    for (int p = tid; p < np; p += THREADS) s_gbase[p] += s_tcnt[p];
    // every warp leader owns its own bulk groups: before ITS out[buf] slot is rewritten (tile t+2)
    // it must have drained the reads of tile t; warp leaders wait here, one tile late
    if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
    __syncthreads();
  }
  }
  if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");   // writes performed
  __threadfence_system();
}

// Epoch barrier across ranks: every rank's prior stream work is complete and visible.
__global__ void sh_barrier(uint64_t* const* flags_peers, uint64_t* flags_local, int rank,
                           int world, int slot, uint64_t epoch, int32_t* status) {
  int t = threadIdx.x;
  __threadfence_system();
  if (t < world) {
    st_release_sys(&flags_peers[t][(size_t)slot * DRYAD_MAX_PEERS + rank], epoch);
    if (!spin_until_ge(&flags_local[(size_t)slot * DRYAD_MAX_PEERS + t], epoch, SPIN_CYCLES))
      atomicMax(status, 2);
  }
}

static size_t scatter_smem(int maxw, int nparts) {
  size_t b = (size_t)TILE * maxw;                       // in
  b += ((size_t)TILE * maxw + 32 * (size_t)nparts + 16 + 15) & ~(size_t)15;  // out
  const size_t npa = ((size_t)nparts + 3) & ~(size_t)3;
  b = (b + 15) & ~(size_t)15;
  b += 8 * npa + 8 * npa;                               // seps, gbase
  b += 4 * (WARPS * npa + npa + (npa + 4) * 2 + npa);
  return (b + 15) & ~(size_t)15;
}

static size_t scatter_tma_smem(int w, int nparts, int emit_key) {
  size_t in_b = ((size_t)TILE * w + 127) & ~(size_t)127;
  size_t out_b = ((size_t)TILE * w + 32 * (size_t)nparts + 16 + 127) & ~(size_t)127;
  size_t k_b = emit_key ? (((size_t)TILE * 8 + 32 * (size_t)nparts + 16 + 127) & ~(size_t)127) : 0;
  size_t npa = ((size_t)nparts + 3) & ~(size_t)3;
  size_t b = 2 * in_b + 2 * out_b + 2 * k_b;
  b += 8 * npa + 8 * npa;
  b += 4 * (WARPS * npa + npa + (npa + 4) + npa + npa);
  return (b + 127) & ~(size_t)127;
}

static int pick_blocks(int64_t n, int* tiles_per_block) {
  int64_t tiles = (n + TILE - 1) / TILE;
  if (tiles < 1) tiles = 1;
  int64_t nb = tiles < 148 * 8 ? tiles : 148 * 8;
  int64_t tpb = (tiles + nb - 1) / nb;
  nb = (tiles + tpb - 1) / tpb;
  *tiles_per_block = (int)tpb;
  return (int)nb;
}

}}  // namespace dryad::sh

extern "C" {

size_t dryad_shuffle_scratch_bytes(int64_t n, int nparts) {
  int tpb;
  int nb = dryad::sh::pick_blocks(n, &tpb);
  return ((((size_t)nparts * nb * 4) + 255) & ~(size_t)255) + 256;   // + done counter
}

int dryad_shuffle_run(const DryadShuffleDesc* d, void* stream) {
  using namespace dryad::sh;
  cudaStream_t st = (cudaStream_t)stream;
  if (!d || d->ncols < 1 || d->ncols > MAX_COLS || d->nparts < 1 || d->nparts > MAX_PARTS ||
      d->world < 1 || d->world > DRYAD_MAX_PEERS || d->rank < 0 || d->rank >= d->world || d->n < 0)
    return DRYAD_ERR_BAD_ARG;
  Params P;
  P.ncols = d->ncols; P.nparts = d->nparts; P.mode = d->mode; P.key_col = d->key_col;
  P.key_kind = d->key_kind; P.key_off = d->key_off; P.emit_key = d->emit_key;
  P.rank = d->rank; P.world = d->world; P.n = d->n;
  int maxw = d->emit_key ? 8 : 0;
  for (int c = 0; c < d->ncols; ++c) {
    if (d->width[c] == 0 || (d->width[c] & 3) || d->width[c] > 512) return DRYAD_ERR_BAD_ARG;
    P.src[c] = (const unsigned char*)d->src[c];
    P.width[c] = d->width[c];
    if ((int)d->width[c] > maxw) maxw = (int)d->width[c];
  }
  if (d->key_col < 0 || d->key_col >= d->ncols) return DRYAD_ERR_BAD_ARG;
  // key_kind 0: the key is the little-endian u64 at the start of the key column's row (rows of
  // several 8-byte words are allowed: packed {key, payload...} records)
  if (d->key_kind == 0 && (d->width[d->key_col] < 8 || (d->width[d->key_col] & 7))) return DRYAD_ERR_BAD_ARG;
  if (d->key_kind == 1 && (d->key_off < 0 || d->key_off + 8 > (int)d->width[d->key_col]))
    return DRYAD_ERR_BAD_ARG;
  P.max_width = maxw;
  P.dst_table = d->dst_table; P.dst_key_table = d->dst_key_table;
  P.separators = d->separators; P.capacity = d->capacity;
  P.counts_local = d->counts_local; P.counts_peers = d->counts_peers;
  P.flags_local = d->flags_local; P.flags_peers = d->flags_peers; P.epoch = d->epoch;
  P.status = d->status;
  P.nblocks = pick_blocks(d->n, &P.tiles_per_block);
  P.key_out = d->key_out;
  P.filter_mod = d->filter_mod; P.filter_rem = d->filter_rem;
  if (d->filter_mod < 0 || (d->filter_mod > 0 && (d->filter_rem < 0 || d->filter_rem >= d->filter_mod))) return DRYAD_ERR_BAD_ARG;
  const int vblocks = P.nblocks;
  size_t counts_bytes = (((size_t)P.nparts * P.nblocks * 4) + 255) & ~(size_t)255;
  if (!d->scratch || d->scratch_bytes < counts_bytes + 256) return DRYAD_ERR_SCRATCH;
  P.block_counts = (uint32_t*)d->scratch;
  uint32_t* done_counter = (uint32_t*)((unsigned char*)d->scratch + counts_bytes);
  const int phases = d->phases ? d->phases : (DRYAD_SHUFFLE_PLAN | DRYAD_SHUFFLE_SCATTER | DRYAD_SHUFFLE_BARRIER);
  const int slot = d->barrier_slot ? d->barrier_slot : 1;
  if (slot < 0 || slot >= DRYAD_FLAG_SLOTS) return DRYAD_ERR_BAD_ARG;

  if (phases & DRYAD_SHUFFLE_PLAN) {
    DRYAD_CUDA_TRY(cudaMemsetAsync(done_counter, 0, 4, st));
    sh_count<<<P.nblocks, THREADS, 0, st>>>(P);
    DRYAD_LAUNCH_CHECK();
    sh_publish<<<P.nparts, 256, 0, st>>>(P, done_counter);
    DRYAD_LAUNCH_CHECK();
  }
  if (phases & DRYAD_SHUFFLE_SCATTER) {
    const char* no_tma = getenv("DRYAD_SHUFFLE_NO_TMA");
    const bool tma_ok = d->ncols == 1 && d->width[0] >= 16 && (((uintptr_t)d->src[0]) & 15) == 0 &&
                        scatter_tma_smem((int)d->width[0], P.nparts, d->emit_key) <= 200 * 1024 && !(no_tma && no_tma[0] == '1');
    if (tma_ok) {
      size_t smem = scatter_tma_smem((int)d->width[0], P.nparts, d->emit_key);
      DRYAD_CUDA_TRY(cudaFuncSetAttribute(sh_scatter_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      int grid = vblocks;
      if (d->max_ctas > 0 && grid > d->max_ctas) grid = d->max_ctas;   // persistent CTAs walk the row blocks
      if (grid > 0) sh_scatter_tma<<<grid, THREADS, smem, st>>>(P);
    } else {
      size_t smem = scatter_smem(maxw, P.nparts);
      DRYAD_CUDA_TRY(cudaFuncSetAttribute(sh_scatter, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      if (vblocks > 0) sh_scatter<<<vblocks, THREADS, smem, st>>>(P);
    }
    DRYAD_LAUNCH_CHECK();
  }
  if (phases & DRYAD_SHUFFLE_BARRIER) {
    sh_barrier<<<1, 32, 0, st>>>(d->flags_peers, d->flags_local, d->rank, d->world, slot,
                                 d->barrier_epoch ? d->barrier_epoch : d->epoch, d->status);
    DRYAD_LAUNCH_CHECK();
  }
  return DRYAD_OK;
}

int dryad_signal_barrier(uint64_t* const* flags_peers, uint64_t* flags_local, int rank, int world,
                         int slot, uint64_t epoch, int32_t* status, void* stream) {
  if (world < 1 || world > DRYAD_MAX_PEERS || slot < 0 || slot >= DRYAD_FLAG_SLOTS)
    return DRYAD_ERR_BAD_ARG;
  dryad::sh::sh_barrier<<<1, 32, 0, (cudaStream_t)stream>>>(flags_peers, flags_local, rank, world,
                                                            slot, epoch, status);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

}  // extern "C"

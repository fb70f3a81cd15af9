// K6 + C2 (hash aggregate, fused reduce-scatter of partials), K8 (hash join), K10 (hash sets).
//
// Reference counterparts: AccumulateDictionary — a chained table with seed/accumulate in place
// (LinqToDryad/DryadLinqCollection.cs:519-722) driven by ParallelHashGroupByPartialAccumulate /
// FullAccumulate (LinqToDryad/DryadLinqVertex.cs:5718-6416); the partial → HashPartition → Merge →
// final GroupBy plan (LinqToDryad/DryadLinqQueryGen.cs:2179-2296); ParallelHashJoin
// (DryadLinqVertex.cs:6703-7315); hash-set Distinct/Union/Intersect/Except (DryadLinqVertex.cs:
// 1239-1660).
//
// Design: one open-addressing table per GPU in HBM.  Keys live in 4-key buckets (one 32-byte L2
// sector, fetched with a single 256-bit load), bucket = mulhi(mix64(key), cap/4), linear probing
// over buckets; `cap` is any multiple of 4.  Aggregate values are slot-major: slot s owns the L
// consecutive int64 words vals[s*L .. s*L+L) (L = lane stride = next power of two >= nlanes), so
// all aggregates of a group share one sector.  A row's L reductions are issued by L adjacent
// threads of ONE warp instruction (the probing thread hands slot and values over by shuffle):
// the L2 sees one reduction request per row instead of one per aggregate.  Measured on B200
// (bench/micro/agg_variants*.cu, 125 M rows into 1 M groups, Sum+Count): linear probing + SoA
// lanes 20 Grows/s, bucketised + paired 67-85 Grows/s, no-probe ceiling 141 Grows/s.
// The decomposable combine across GPUs is NOT a separate shuffle: `agg_merge_peers` walks the
// local partial table and applies each partial directly to the key-owner's final table in the
// owner's HBM with system-scope atomics over NVLink — the reduce-scatter and the final accumulate
// are the same instruction.
#include "common.cuh"
#include "../../include/dryad_b200.h"

namespace dryad { namespace ht {

constexpr uint64_t EMPTY = ~0ull;  // reserved key; the engine remaps a genuine ~0 key on ingest
constexpr uint64_t NOSLOT = ~0ull;

struct AggDesc {
  int nlanes;                          // lanes the kernels update (user lanes + a hidden row counter in direct mode)
  int nlanes_out;                      // user lanes (what extract / merge hand on)
  int in_stride;                       // elements between consecutive rows of the inputs (1 = columns)
  int direct;                          // 1: slot = key - key_lo, no key array, no probing
  int occ_lane;                        // direct mode: the COUNT lane whose value > 0 marks an occupied slot
  int src_stride;                      // merge: lane stride of the LOCAL table (may differ from the owners')
  uint64_t key_lo;
  int op[DRYAD_MAX_LANES];
  const int64_t* in[DRYAD_MAX_LANES];  // input columns (may be null for COUNT)
};

struct PeerTables {
  int world;
  uint64_t* keys[DRYAD_MAX_PEERS];
  int64_t* vals[DRYAD_MAX_PEERS];
  uint64_t cap;
};

__device__ __forceinline__ int64_t identity_of(int op) {
  switch (op) {
    case DRYAD_AGG_MIN_I64: return INT64_MAX;
    case DRYAD_AGG_MAX_I64: return INT64_MIN;
    case DRYAD_AGG_MIN_F64: return __double_as_longlong(__longlong_as_double(0x7ff0000000000000ll));
    case DRYAD_AGG_MAX_F64: return __double_as_longlong(__longlong_as_double(0xfff0000000000000ll));
    default: return 0;  // sums and counts (0.0 is all-zero bits)
  }
}

__device__ __forceinline__ void atomic_min_f64(double* addr, double v) {
  unsigned long long* a = (unsigned long long*)addr;
  unsigned long long old = *a, assumed;
  do {
    assumed = old;
    if (__longlong_as_double(assumed) <= v) break;
    old = atomicCAS(a, assumed, (unsigned long long)__double_as_longlong(v));
  } while (assumed != old);
}
__device__ __forceinline__ void atomic_max_f64(double* addr, double v) {
  unsigned long long* a = (unsigned long long*)addr;
  unsigned long long old = *a, assumed;
  do {
    assumed = old;
    if (__longlong_as_double(assumed) >= v) break;
    old = atomicCAS(a, assumed, (unsigned long long)__double_as_longlong(v));
  } while (assumed != old);
}

// `combine`: true when folding partial aggregates (COUNT partials add instead of counting rows).
// COUNT and SUM_I64 share ONE reduction instruction so that a warp whose lanes alternate between
// the two (the Sum+Count pair layout) issues a single RED.
template <bool SYSTEM>
__device__ __forceinline__ void apply(int op, int64_t* slot, int64_t v, bool combine,
                                      int32_t* status) {
  if (op <= DRYAD_AGG_SUM_I64) {
    const unsigned long long add = (op == DRYAD_AGG_COUNT && !combine) ? 1ull : (unsigned long long)v;
    if (SYSTEM) atomicAdd_system((unsigned long long*)slot, add);
    else atomicAdd((unsigned long long*)slot, add);
    return;
  }
  switch (op) {
    case DRYAD_AGG_SUM_I64_CHECKED: {
      long long old = (long long)(SYSTEM ? atomicAdd_system((unsigned long long*)slot, (unsigned long long)v)
                                         : atomicAdd((unsigned long long*)slot, (unsigned long long)v));
      long long nw = (long long)((unsigned long long)old + (unsigned long long)v);
      if (((old ^ nw) & (v ^ nw)) < 0) atomicMax(status, 3);  // AddChecked overflow
      break;
    }
    case DRYAD_AGG_SUM_F64:
      if (SYSTEM) atomicAdd_system((double*)slot, __longlong_as_double(v));
      else atomicAdd((double*)slot, __longlong_as_double(v));
      break;
    case DRYAD_AGG_MIN_I64:
      if (SYSTEM) atomicMin_system((long long*)slot, (long long)v); else atomicMin((long long*)slot, (long long)v);
      break;
    case DRYAD_AGG_MAX_I64:
      if (SYSTEM) atomicMax_system((long long*)slot, (long long)v); else atomicMax((long long*)slot, (long long)v);
      break;
    case DRYAD_AGG_MIN_F64: atomic_min_f64((double*)slot, __longlong_as_double(v)); break;
    case DRYAD_AGG_MAX_F64: atomic_max_f64((double*)slot, __longlong_as_double(v)); break;
  }
}

// One bucket = 4 keys = one sector.  Local tables: a single 256-bit L2 load (a stale EMPTY is
// harmless: the CAS that follows returns the truth, and a slot only ever changes EMPTY -> key).
// Peer tables: four relaxed system-scope loads.
template <bool SYSTEM>
__device__ __forceinline__ void load_bucket(const uint64_t* p, uint64_t (&k)[4]) {
  if (SYSTEM) {
    // peer HBM: one 32-byte NVLink read (served by the home GPU's L2, the point of coherence for the
    // system-scope CAS/RED that follow; never cached on this side: .cg bypasses L1 and peer lines do
    // not live in the local L2).  Four scalar relaxed.sys loads cost four round-trip packets.
    asm volatile("ld.global.cg.v4.u64 {%0, %1, %2, %3}, [%4];"
                 : "=l"(k[0]), "=l"(k[1]), "=l"(k[2]), "=l"(k[3]) : "l"(p) : "memory");
  } else {
    asm volatile("ld.global.cg.v4.u64 {%0, %1, %2, %3}, [%4];"
                 : "=l"(k[0]), "=l"(k[1]), "=l"(k[2]), "=l"(k[3]) : "l"(p) : "memory");
  }
}

__device__ __forceinline__ uint64_t home_bucket(uint64_t key, uint64_t nbuckets) {
  return __umul64hi(mix64(key), nbuckets);
}

// Find-or-claim the slot of `key`.  Returns cap on a full table.  Slots of a bucket are claimed in
// order, so a key is always stored before the first EMPTY slot of its probe sequence.
template <bool SYSTEM>
__device__ __forceinline__ uint64_t find_or_insert(uint64_t* keys, uint64_t cap, uint64_t key) {
  const uint64_t nb = cap >> 2;
  uint64_t b = home_bucket(key, nb);
  for (uint64_t probes = 0; probes < nb; ++probes) {
    uint64_t k[4];
    load_bucket<SYSTEM>(keys + b * 4, k);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (k[j] == key) return b * 4 + j;
      if (k[j] == EMPTY) {
        unsigned long long old = SYSTEM
            ? atomicCAS_system((unsigned long long*)(keys + b * 4 + j), (unsigned long long)EMPTY, (unsigned long long)key)
            : atomicCAS((unsigned long long*)(keys + b * 4 + j), (unsigned long long)EMPTY, (unsigned long long)key);
        if (old == EMPTY || old == key) return b * 4 + j;
      }
    }
    if (++b == nb) b = 0;
  }
  return cap;
}

// Read-only lookup in a finished table.  Returns cap when absent.
__device__ __forceinline__ uint64_t find_slot(const uint64_t* keys, uint64_t cap, uint64_t key) {
  const uint64_t nb = cap >> 2;
  uint64_t b = home_bucket(key, nb);
  for (uint64_t probes = 0; probes < nb; ++probes) {
    uint64_t k[4];
    asm volatile("ld.global.nc.v4.u64 {%0, %1, %2, %3}, [%4];"
                 : "=l"(k[0]), "=l"(k[1]), "=l"(k[2]), "=l"(k[3]) : "l"(keys + b * 4));
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (k[j] == key) return b * 4 + j;
      if (k[j] == EMPTY) return cap;
    }
    if (++b == nb) b = 0;
  }
  return cap;
}

__device__ __forceinline__ int64_t ld_stream(const int64_t* p) {
  return (int64_t)__ldcs(reinterpret_cast<const unsigned long long*>(p));
}

// The four slot keys of bucket b and a 4-bit occupancy mask: the key array of a hashed table
// (EMPTY marks a free slot), or (direct tables) key_lo + slot for the slots whose row counter is
// non-zero — there every 64-bit value, ~0 included, is a legitimate key.
__device__ __forceinline__ uint32_t bucket_keys(const uint64_t* __restrict__ keys, const int64_t* __restrict__ vals,
                                                const AggDesc& A, int stride, uint64_t b, uint64_t (&k)[4]) {
  uint32_t mask = 0;
  if (!A.direct) {
    asm volatile("ld.global.nc.v4.u64 {%0, %1, %2, %3}, [%4];"
                 : "=l"(k[0]), "=l"(k[1]), "=l"(k[2]), "=l"(k[3]) : "l"(keys + b * 4));
#pragma unroll
    for (int j = 0; j < 4; ++j) mask |= (uint32_t)(k[j] != EMPTY) << j;
  } else {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      k[j] = A.key_lo + b * 4 + j;
      mask |= (uint32_t)(vals[(b * 4 + j) * stride + A.occ_lane] != 0) << j;
    }
  }
  return mask;
}

template <int L>
__global__ void agg_init_kernel(uint64_t* keys, uint64_t cap, AggDesc A, int64_t* vals) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < cap * L; i += stride) {
    const int l = (int)(i & (L - 1));
    if (l == 0 && keys) keys[i / L] = EMPTY;
    vals[i] = (l < A.nlanes) ? identity_of(A.op[l]) : 0;
  }
}

// Hand each probed row to a group of L adjacent lanes; lane (g*L + l) applies aggregate l.
template <int L, bool SYSTEM>
__device__ __forceinline__ void apply_rows(const AggDesc& A, int64_t* vals, uint64_t s, const int64_t (&v)[L],
                                           bool combine, int32_t* status) {
  const int lane = (int)lane_id();
  const int sub = lane & (L - 1);
  int op = -1;                               // unrolled select: no dynamic index into the params
#pragma unroll
  for (int l = 0; l < L; ++l) if (l == sub && l < A.nlanes) op = A.op[l];
#pragma unroll
  for (int h = 0; h < L; ++h) {
    const int src = h * (32 / L) + lane / L;
    const uint64_t ss = __shfl_sync(0xffffffffu, s, src);
    int64_t vv = 0;
#pragma unroll
    for (int l = 0; l < L; ++l) {
      const int64_t t = __shfl_sync(0xffffffffu, v[l], src);
      if (l == sub) vv = t;
    }
    if (ss != NOSLOT && op >= 0) apply<SYSTEM>(op, vals + ss * L + sub, vv, combine, status);
  }
}

template <int L>
__global__ void __launch_bounds__(256) agg_update_kernel(uint64_t* keys, uint64_t cap, AggDesc A, int64_t* vals,
                                                         const uint64_t* __restrict__ in_keys,
                                                         int64_t n, int combine, int32_t* status,
                                                         const int32_t* skip_if_set = nullptr) {
  if (skip_if_set && *reinterpret_cast<const volatile int32_t*>(skip_if_set) != 0) return;   // the front-table kernel owns these rows
  int64_t base = (int64_t)blockIdx.x * blockDim.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; base < n; base += stride) {          // block-uniform trip count (shuffles below)
    const int64_t i = base + threadIdx.x;
    uint64_t s = NOSLOT;
    int64_t v[L];
#pragma unroll
    for (int l = 0; l < L; ++l) v[l] = 1;
    if (i < n) {
      const uint64_t key = (uint64_t)ld_stream(reinterpret_cast<const int64_t*>(in_keys) + i * A.in_stride);
#pragma unroll
      for (int l = 0; l < L; ++l)
        if (l < A.nlanes && A.in[l]) v[l] = ld_stream(A.in[l] + i * A.in_stride);
      if (A.direct) {
        s = key - A.key_lo;                      // dense key range: the key IS the slot
        if (s >= cap) { atomicMax(status, 4); s = NOSLOT; }
      } else {
        s = find_or_insert<false>(keys, cap, key);
        if (s == cap) { atomicMax(status, 1); s = NOSLOT; }
      }
    }
    apply_rows<L, false>(A, vals, s, v, combine != 0, status);
  }
}

// ---- shared-memory front table: low-cardinality / skewed keys ----------------------------------
// With few distinct keys every global reduction lands on a handful of addresses and the L2
// serialises them (24 M tokens into 4 K words: 20 ms).  Each CTA therefore keeps a small table in
// shared memory (1024 slots): rows whose key is resident there are folded with shared-memory
// atomics; rows that miss go to the global table as usual (same lane-paired reductions).  A CTA
// whose hit rate stays under a quarter for two consecutive tiles (high-cardinality data) stops
// probing the front table.  At the end the front table is flushed into the global one as partial
// aggregates.  Reference role: the per-thread AccumulateDictionary instances of
// ParallelHashGroupByPartialAccumulate (LinqToDryad/DryadLinqVertex.cs:5718-6032).
constexpr int FRONT_SLOTS = 1024;
constexpr int FRONT_PROBES = 4;

__device__ __forceinline__ void apply_shared(int op, int64_t* slot, int64_t v, bool combine) {
  switch (op) {
    case DRYAD_AGG_COUNT: atomicAdd((unsigned long long*)slot, combine ? (unsigned long long)v : 1ull); break;
    case DRYAD_AGG_SUM_I64: atomicAdd((unsigned long long*)slot, (unsigned long long)v); break;
    case DRYAD_AGG_SUM_F64: atomicAdd((double*)slot, __longlong_as_double(v)); break;
    case DRYAD_AGG_MIN_I64: atomicMin((long long*)slot, (long long)v); break;
    case DRYAD_AGG_MAX_I64: atomicMax((long long*)slot, (long long)v); break;
    case DRYAD_AGG_MIN_F64: atomic_min_f64((double*)slot, __longlong_as_double(v)); break;
    case DRYAD_AGG_MAX_F64: atomic_max_f64((double*)slot, __longlong_as_double(v)); break;
    default: break;       // checked sums never use the front table (host side)
  }
}

template <int L>
__global__ void __launch_bounds__(256) agg_update_front_kernel(uint64_t* keys, uint64_t cap, AggDesc A, int64_t* vals,
                                                               const uint64_t* __restrict__ in_keys,
                                                               int64_t n, int combine, int32_t* status,
                                                               const int32_t* run_if_set, int32_t* probe_stats) {
  // run_if_set: skip unless the probe launch found the keys to be low-cardinality.
  // probe_stats (probe launch only): [0] hits, [1] rows, [2] finished CTAs, [3] verdict.
  if (run_if_set && *reinterpret_cast<const volatile int32_t*>(run_if_set) == 0) return;
  extern __shared__ __align__(16) unsigned char front_raw[];
  uint64_t* fkeys = reinterpret_cast<uint64_t*>(front_raw);                    // [FRONT_SLOTS]
  int64_t* fvals = reinterpret_cast<int64_t*>(fkeys + FRONT_SLOTS);            // [FRONT_SLOTS][L]
  __shared__ int s_hits, s_mode, s_low, s_fill;
  for (int i = threadIdx.x; i < FRONT_SLOTS; i += blockDim.x) fkeys[i] = EMPTY;
  for (int i = threadIdx.x; i < FRONT_SLOTS * L; i += blockDim.x) {
    const int l = i & (L - 1);
    fvals[i] = (l < A.nlanes) ? identity_of(A.op[l]) : 0;
  }
  if (threadIdx.x == 0) { s_hits = 0; s_mode = 1; s_low = 0; s_fill = 0; }
  __syncthreads();
  int tot_hits = 0, tot_rows = 0;                       // thread 0 only

  int64_t base = (int64_t)blockIdx.x * blockDim.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  // ---- phase 1: front table in use (left for good once the hit rate stays low) ----------------
  for (; base < n; base += stride) {
    if (s_mode == 0) break;                             // block-uniform: written before the last barrier
    const int64_t i = base + threadIdx.x;
    uint64_t s = NOSLOT;
    int64_t v[L];
#pragma unroll
    for (int l = 0; l < L; ++l) v[l] = 1;
    if (i < n) {
      const uint64_t key = (uint64_t)ld_stream(reinterpret_cast<const int64_t*>(in_keys) + i * A.in_stride);
#pragma unroll
      for (int l = 0; l < L; ++l)
        if (l < A.nlanes && A.in[l]) v[l] = ld_stream(A.in[l] + i * A.in_stride);
      bool done = false, hit = false;
      if (key != EMPTY) {
        uint32_t f = (uint32_t)(mix64(key) >> 40) & (FRONT_SLOTS - 1);
        for (int pr = 0; pr < FRONT_PROBES && !done; ++pr) {
          uint64_t cur = *reinterpret_cast<volatile uint64_t*>(&fkeys[f]);
          if (cur == EMPTY) {
            // new keys are admitted only while the table is at most 3/4 full (short probe chains)
            if (*reinterpret_cast<volatile int*>(&s_fill) >= FRONT_SLOTS * 3 / 4) break;
            const unsigned long long old = atomicCAS((unsigned long long*)&fkeys[f], (unsigned long long)EMPTY, (unsigned long long)key);
            if (old == (unsigned long long)EMPTY) { atomicAdd(&s_fill, 1); cur = key; } else { cur = (uint64_t)old; hit = cur == key; }
          } else {
            hit = cur == key;
          }
          if (cur == key) {
#pragma unroll
            for (int l = 0; l < L; ++l)
              if (l < A.nlanes) apply_shared(A.op[l], &fvals[f * L + l], v[l], combine != 0);
            done = true;
          }
          f = (f + 1) & (FRONT_SLOTS - 1);
        }
        if (hit) atomicAdd(&s_hits, 1);                // a first insertion is not a hit
      }
      if (!done) {
        if (A.direct) {
          s = key - A.key_lo;
          if (s >= cap) { atomicMax(status, 4); s = NOSLOT; }
        } else {
          s = find_or_insert<false>(keys, cap, key);
          if (s == cap) { atomicMax(status, 1); s = NOSLOT; }
        }
      }
    }
    apply_rows<L, false>(A, vals, s, v, combine != 0, status);
    __syncthreads();
    if (threadIdx.x == 0) {
      tot_hits += s_hits; tot_rows += (int)min((int64_t)blockDim.x, n - base);
      s_low = (s_hits * 4 < (int)blockDim.x) ? s_low + 1 : 0;
      if (s_low >= 2 && !probe_stats) s_mode = 0;      // the probe launch keeps counting
      s_hits = 0;
    }
    __syncthreads();
  }
  // ---- phase 2: plain global path (same as agg_update_kernel) ------------------------------------
  for (; base < n; base += stride) {
    const int64_t i = base + threadIdx.x;
    uint64_t s = NOSLOT;
    int64_t v[L];
#pragma unroll
    for (int l = 0; l < L; ++l) v[l] = 1;
    if (i < n) {
      const uint64_t key = (uint64_t)ld_stream(reinterpret_cast<const int64_t*>(in_keys) + i * A.in_stride);
#pragma unroll
      for (int l = 0; l < L; ++l)
        if (l < A.nlanes && A.in[l]) v[l] = ld_stream(A.in[l] + i * A.in_stride);
      if (A.direct) {
        s = key - A.key_lo;
        if (s >= cap) { atomicMax(status, 4); s = NOSLOT; }
      } else {
        s = find_or_insert<false>(keys, cap, key);
        if (s == cap) { atomicMax(status, 1); s = NOSLOT; }
      }
    }
    apply_rows<L, false>(A, vals, s, v, combine != 0, status);
  }
  __syncthreads();
  // flush the front table: partial aggregates into the global table
  for (int f = threadIdx.x; f < FRONT_SLOTS; f += blockDim.x) {
    const uint64_t key = fkeys[f];
    if (key == EMPTY) continue;
    uint64_t s;
    if (A.direct) { s = key - A.key_lo; if (s >= cap) { atomicMax(status, 4); continue; } }
    else { s = find_or_insert<false>(keys, cap, key); if (s == cap) { atomicMax(status, 1); continue; } }
    for (int l = 0; l < A.nlanes; ++l) apply<false>(A.op[l], vals + s * L + l, fvals[f * L + l], true, status);
  }
  if (probe_stats && threadIdx.x == 0) {
    atomicAdd(&probe_stats[0], tot_hits);
    atomicAdd(&probe_stats[1], tot_rows);
    __threadfence();
    if (atomicAdd(&probe_stats[2], 1) == (int)gridDim.x - 1) {
      __threadfence();
      const int h = *reinterpret_cast<volatile int32_t*>(&probe_stats[0]), r = *reinterpret_cast<volatile int32_t*>(&probe_stats[1]);
      probe_stats[3] = (h * 5 >= r * 2) ? 1 : 0;        // >= 40 % of the probed rows hit a resident key
    }
  }
}

// K6+C2 fused: fold every local partial into the owner rank's table over NVLink.
template <int L>
__global__ void __launch_bounds__(256) agg_merge_peers_kernel(const uint64_t* __restrict__ keys,
                                                              uint64_t cap, AggDesc A, const int64_t* __restrict__ vals,
                                                              PeerTables T, int32_t* status) {
  uint64_t base = (uint64_t)blockIdx.x * blockDim.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  const int lane = (int)lane_id();
  const int sub = lane & (L - 1);
  int op = -1;                               // unrolled select: no dynamic index into the params
#pragma unroll
  for (int l = 0; l < L; ++l) if (l == sub && l < A.nlanes_out) op = A.op[l];
  for (; base < cap; base += stride) {
    const uint64_t i = base + threadIdx.x;
    uint64_t key = EMPTY;
    bool occupied = false;
    if (i < cap) {
      if (A.direct) { key = A.key_lo + i; occupied = vals[i * A.src_stride + A.occ_lane] != 0; }
      else { key = keys[i]; occupied = key != EMPTY; }
    }
    uint64_t s = NOSLOT;
    int owner = 0;
    if (occupied) {
      owner = (int)hash_bucket(key, (uint32_t)T.world);
      s = find_or_insert<true>(T.keys[owner], T.cap, key);
      if (s == T.cap) { atomicMax(status, 1); s = NOSLOT; }
    }
#pragma unroll
    for (int h = 0; h < L; ++h) {
      const int src = h * (32 / L) + lane / L;
      const uint64_t ss = __shfl_sync(0xffffffffu, s, src);
      const int ow = __shfl_sync(0xffffffffu, owner, src);
      const uint64_t row = base + (threadIdx.x & ~31u) + src;   // local slot probed by lane `src`
      if (ss != NOSLOT && op >= 0)
        apply<true>(op, T.vals[ow] + ss * L + sub, vals[row * A.src_stride + sub], true, status);
    }
  }
}

// Compaction of the occupied slots.  One thread = one bucket (a single 256-bit load); occupancy
// counts are scanned across the block so that a tile of 1024 slots costs ONE atomic on the output
// cursor (per-warp atomics on one address serialise in the L2: 0.2 ms for a 3 M-slot table).
template <int L>
__global__ void __launch_bounds__(256) agg_extract_kernel(const uint64_t* __restrict__ keys,
                                                          uint64_t cap, AggDesc A, const int64_t* __restrict__ vals,
                                                          uint64_t* __restrict__ out_keys,
                                                          int64_t* const* __restrict__ out_lanes,
                                                          unsigned long long* __restrict__ out_count,
                                                          int64_t out_cap) {
  __shared__ uint32_t warp_tot[8];
  __shared__ unsigned long long tile_base;
  const uint64_t nb = cap >> 2;
  const int lane = (int)lane_id(), warp = threadIdx.x >> 5;
  for (uint64_t b0 = (uint64_t)blockIdx.x * 256; b0 < nb; b0 += (uint64_t)gridDim.x * 256) {
    const uint64_t b = b0 + threadIdx.x;
    uint64_t k[4] = {EMPTY, EMPTY, EMPTY, EMPTY};
    const uint32_t occ = (b < nb) ? bucket_keys(keys, vals, A, L, b, k) : 0u;
    const uint32_t mine = __popc(occ);
    uint32_t incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += t;
    }
    if (lane == 31) warp_tot[warp] = incl;
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t tot = 0;
#pragma unroll
      for (int w = 0; w < 8; ++w) { const uint32_t t = warp_tot[w]; warp_tot[w] = tot; tot += t; }
      tile_base = tot ? atomicAdd(out_count, (unsigned long long)tot) : 0ull;
    }
    __syncthreads();
    int64_t o = (int64_t)tile_base + warp_tot[warp] + (incl - mine);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (!((occ >> j) & 1)) continue;
      if (o < out_cap) {
        out_keys[o] = k[j];
        for (int l = 0; l < A.nlanes_out; ++l) out_lanes[l][o] = vals[(b * 4 + j) * L + l];
      }
      ++o;
    }
    __syncthreads();                 // warp_tot / tile_base are reused by the next tile
  }
}

// Same compaction, but every occupied slot becomes ONE packed row {key, vals[0..nlanes)} of
// (1 + nlanes) words: the unit the fused shuffle moves when partial aggregates travel to their
// key owners as a push (stores) instead of remote probes.
template <int L>
__global__ void __launch_bounds__(256) agg_extract_rows_kernel(const uint64_t* __restrict__ keys,
                                                               uint64_t cap, AggDesc A, const int64_t* __restrict__ vals,
                                                               int64_t* __restrict__ out_rows,
                                                               unsigned long long* __restrict__ out_count,
                                                               int64_t out_cap) {
  __shared__ uint32_t warp_tot[8];
  __shared__ unsigned long long tile_base;
  const uint64_t nb = cap >> 2;
  const int lane = (int)lane_id(), warp = threadIdx.x >> 5;
  const int rw = 1 + A.nlanes_out;
  for (uint64_t b0 = (uint64_t)blockIdx.x * 256; b0 < nb; b0 += (uint64_t)gridDim.x * 256) {
    const uint64_t b = b0 + threadIdx.x;
    uint64_t k[4] = {EMPTY, EMPTY, EMPTY, EMPTY};
    const uint32_t occ = (b < nb) ? bucket_keys(keys, vals, A, L, b, k) : 0u;
    const uint32_t mine = __popc(occ);
    uint32_t incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += t;
    }
    if (lane == 31) warp_tot[warp] = incl;
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t tot = 0;
#pragma unroll
      for (int w = 0; w < 8; ++w) { const uint32_t t = warp_tot[w]; warp_tot[w] = tot; tot += t; }
      tile_base = tot ? atomicAdd(out_count, (unsigned long long)tot) : 0ull;
    }
    __syncthreads();
    int64_t o = (int64_t)tile_base + warp_tot[warp] + (incl - mine);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (!((occ >> j) & 1)) continue;
      if (o < out_cap) {
        out_rows[o * rw] = (int64_t)k[j];
        for (int l = 0; l < A.nlanes_out; ++l) out_rows[o * rw + 1 + l] = vals[(b * 4 + j) * L + l];
      }
      ++o;
    }
    __syncthreads();
  }
}

// ---- K8 hash join (multimap: head[slot] -> row chain) ---------------------------------------
__global__ void join_init_kernel(uint64_t* keys, int32_t* head, uint64_t cap) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < cap; i += stride) { keys[i] = EMPTY; head[i] = -1; }
}

__global__ void join_build_kernel(uint64_t* keys, int32_t* head, uint64_t cap,
                                  const uint64_t* __restrict__ build_keys, int32_t* next, int64_t n,
                                  int32_t* status) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    uint64_t s = find_or_insert<false>(keys, cap, build_keys[i]);
    if (s == cap) { atomicMax(status, 1); continue; }
    const int32_t prev = atomicExch(&head[s], (int32_t)i);
    next[i] = prev;
    if (prev >= 0 && *reinterpret_cast<volatile int32_t*>(status) < 2) atomicMax(status, 2);   // duplicate build key
  }
}

// mode 0: count matches per probe row; mode 1: emit (probe, build) index pairs at offsets[i].
__global__ void join_probe_kernel(const uint64_t* __restrict__ keys, const int32_t* __restrict__ head,
                                  uint64_t cap, const int32_t* __restrict__ next,
                                  const uint64_t* __restrict__ probe_keys, int64_t n, int mode,
                                  int64_t* __restrict__ counts_or_offsets,
                                  int32_t* __restrict__ out_probe, int32_t* __restrict__ out_build) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    uint64_t s = find_slot(keys, cap, probe_keys[i]);
    if (mode == 0) {
      int64_t c = 0;
      if (s != cap) for (int32_t j = head[s]; j >= 0; j = next[j]) ++c;
      counts_or_offsets[i] = c;
    } else if (s != cap) {
      int64_t o = counts_or_offsets[i];
      for (int32_t j = head[s]; j >= 0; j = next[j]) { out_probe[o] = (int32_t)i; out_build[o] = j; ++o; }
    }
  }
}

// K8 + C4 fused: broadcast join build.  Every rank owns a slice of the small (dimension) table;
// this kernel replicates the slice's payload columns into EVERY rank's replica buffers and inserts
// each key into EVERY rank's hash table, all with stores / system-scope atomics on peer HBM over
// NVLink — the broadcast and the hash build are one pass over the slice (reference: Tee +
// CrossProduct + Merge(parCount) then HashJoin, LinqToDryad/DryadLinqQueryGen.cs:2957-3001,
// LinqToDryad/DryadLinqVertex.cs:6703-7315).
struct BcastTables {
  int world, ncols;
  uint64_t* keys[DRYAD_MAX_PEERS];
  int32_t* head[DRYAD_MAX_PEERS];
  int32_t* next[DRYAD_MAX_PEERS];
  int64_t* cols[DRYAD_MAX_PEERS][DRYAD_MAX_LANES];
  const int64_t* src[DRYAD_MAX_LANES];
  uint64_t cap;
};

__global__ void __launch_bounds__(256) join_build_bcast_kernel(BcastTables T, const uint64_t* __restrict__ keys,
                                                               int64_t n, int64_t row_base, int32_t* status) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    const uint64_t key = keys[i];
    const int32_t g = (int32_t)(row_base + i);
    for (int r = 0; r < T.world; ++r) {
      for (int c = 0; c < T.ncols; ++c) T.cols[r][c][g] = T.src[c][i];
      uint64_t s = find_or_insert<true>(T.keys[r], T.cap, key);
      if (s == T.cap) { atomicMax(status, 1); continue; }
      T.next[r][g] = atomicExch_system(&T.head[r][s], g);
    }
  }
}

// K17 + C4: replicate this rank's slice of (key, payload...) columns into every rank's replica
// buffers with 16-byte P2P stores (the "CP" copy vertex of the reference's broadcast tree,
// DryadVertex/VertexHost/vertex/vertexHost/vertexHost.cpp:331-571, done by the producer itself).
struct ReplTables {
  int world, ncols;
  int64_t* dst[DRYAD_MAX_PEERS][DRYAD_MAX_LANES + 1];
  const int64_t* src[DRYAD_MAX_LANES + 1];
};
__global__ void __launch_bounds__(256) replicate_slice_kernel(ReplTables T, int64_t n, int64_t row_base) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride)
    for (int c = 0; c < T.ncols; ++c) {
      const int64_t v = T.src[c][i];
      for (int r = 0; r < T.world; ++r) T.dst[r][c][row_base + i] = v;
    }
}

// Probe for a build side with UNIQUE keys (dimension tables): one pass, no count/scan/emit —
// out_row[i] = matching build row (0 when absent), matched[i] = 1/0.
__global__ void join_probe_unique_kernel(const uint64_t* __restrict__ keys, const int32_t* __restrict__ head,
                                         uint64_t cap, const uint64_t* __restrict__ probe_keys, int64_t n,
                                         int32_t* __restrict__ out_row, uint8_t* __restrict__ matched) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    uint64_t s = find_slot(keys, cap, probe_keys[i]);
    int32_t r = (s != cap) ? head[s] : -1;
    out_row[i] = r < 0 ? 0 : r;
    matched[i] = r >= 0;
  }
}

// Set membership: out[i] = 1 if probe key is present.
__global__ void set_contains_kernel(const uint64_t* __restrict__ keys, uint64_t cap,
                                    const uint64_t* __restrict__ probe_keys, int64_t n,
                                    uint8_t* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) out[i] = find_slot(keys, cap, probe_keys[i]) != cap;
}

// Distinct that keeps the FIRST occurrence index per key (deterministic): lane 0 = MIN index.
static inline int grid_for(int64_t n) {
  int64_t g = (n + 256 * 4 - 1) / (256 * 4);
  if (g < 1) g = 1;
  if (g > 148 * 16) g = 148 * 16;
  return (int)g;
}

static inline int lane_stride(int nlanes) { return nlanes <= 1 ? 1 : nlanes <= 2 ? 2 : nlanes <= 4 ? 4 : 8; }

static int fill_desc(AggDesc& A, int nlanes, const int* ops, const int64_t* const* in) {
  if (nlanes < 0 || nlanes > DRYAD_MAX_LANES) return DRYAD_ERR_BAD_ARG;
  A.nlanes = nlanes; A.nlanes_out = nlanes;
  A.in_stride = 1; A.direct = 0; A.occ_lane = 0; A.key_lo = 0;
  A.src_stride = lane_stride(nlanes);
  for (int l = 0; l < DRYAD_MAX_LANES; ++l) { A.op[l] = 0; A.in[l] = nullptr; }
  for (int l = 0; l < nlanes; ++l) {
    if (ops[l] < 0 || ops[l] > DRYAD_AGG_MAX_F64) return DRYAD_ERR_BAD_ARG;
    A.op[l] = ops[l];
    A.in[l] = in ? in[l] : nullptr;
  }
  return DRYAD_OK;
}

// Adaptive update.  With a 4-word work area: (1) a small PROBE launch folds the first rows through
// the shared-memory front table and records how many rows hit a resident key; (2) the rest of the
// rows are covered by two launches of which exactly one does work, chosen on the device by the
// probe's verdict — the plain lane-paired kernel for high-cardinality keys, the front-table kernel
// for few / skewed keys.  No host synchronisation; each kernel keeps its own register and shared
// memory budget.  Without a work area: the plain kernel.
constexpr int64_t PROBE_ROWS = 8 * 1024;      // 8 CTAs x 4 tiles: a few microseconds
template <int L>
static int launch_update(uint64_t* keys, uint64_t cap, AggDesc A, int64_t* vals, const uint64_t* in_keys,
                         int64_t n, int combine, int32_t* status, cudaStream_t st, int32_t* work = nullptr) {
  bool checked = false;
  for (int l = 0; l < A.nlanes; ++l) checked |= A.op[l] == DRYAD_AGG_SUM_I64_CHECKED;
  static const bool front_off = [] { const char* e = getenv("DRYAD_AGG_FRONT"); return e && e[0] == '0'; }();
  auto grid = [](int64_t rows) { int64_t g = (rows + 255) / 256; if (g > 148 * 16) g = 148 * 16; return (int)(g < 1 ? 1 : g); };
  if (!work || checked || front_off || n < 32 * PROBE_ROWS) {
    agg_update_kernel<L><<<grid(n), 256, 0, st>>>(keys, cap, A, vals, in_keys, n, combine, status);
  } else {
    const size_t smem = (size_t)FRONT_SLOTS * 8 * (1 + L);
    if (smem > 48 * 1024) {
      cudaError_t e = cudaFuncSetAttribute(agg_update_front_kernel<L>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return (int)e;
    }
    cudaError_t e = cudaMemsetAsync(work, 0, 16, st);
    if (e != cudaSuccess) return (int)e;
    agg_update_front_kernel<L><<<(int)(PROBE_ROWS / 1024), 256, smem, st>>>(keys, cap, A, vals, in_keys, PROBE_ROWS, combine, status, nullptr, work);
    AggDesc R = A;                                      // the remaining rows
    for (int l = 0; l < R.nlanes; ++l) if (R.in[l]) R.in[l] += PROBE_ROWS * R.in_stride;
    const uint64_t* rk = in_keys + PROBE_ROWS * A.in_stride;
    const int64_t rn = n - PROBE_ROWS;
    agg_update_kernel<L><<<grid(rn), 256, 0, st>>>(keys, cap, R, vals, rk, rn, combine, status, work + 3);
    agg_update_front_kernel<L><<<grid(rn), 256, smem, st>>>(keys, cap, R, vals, rk, rn, combine, status, work + 3, nullptr);
  }
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? DRYAD_OK : (int)e;
}

// table capacities: any multiple of the 4-key bucket
static inline bool cap_ok(uint64_t x) { return x >= 4 && (x & 3) == 0; }

#define DRYAD_DISPATCH_L(nl, CALL)                 \
  switch (lane_stride(nl)) {                       \
    case 1: { constexpr int L = 1; CALL; break; }  \
    case 2: { constexpr int L = 2; CALL; break; }  \
    case 4: { constexpr int L = 4; CALL; break; }  \
    default: { constexpr int L = 8; CALL; break; } \
  }

}}  // namespace dryad::ht

extern "C" {
using namespace dryad::ht;

int dryad_hash_agg_lane_stride(int nlanes) { return lane_stride(nlanes); }

int dryad_hash_agg_init(uint64_t* tbl_keys, uint64_t cap, int nlanes, const int* ops,
                        int64_t* tbl_vals, void* stream) {
  if (!cap_ok(cap)) return DRYAD_ERR_BAD_ARG;
  AggDesc A; int rc = fill_desc(A, nlanes, ops, nullptr); if (rc) return rc;
  DRYAD_DISPATCH_L(nlanes, (agg_init_kernel<L><<<grid_for((int64_t)cap * L), 256, 0, (cudaStream_t)stream>>>(tbl_keys, cap, A, tbl_vals)));
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_hash_agg_update(uint64_t* tbl_keys, uint64_t cap, int nlanes, const int* ops,
                          int64_t* tbl_vals, const uint64_t* in_keys,
                          const int64_t* const* in_lanes, int64_t n, int combine, int32_t* status,
                          void* stream) {
  if (!cap_ok(cap)) return DRYAD_ERR_BAD_ARG;
  if (n <= 0) return DRYAD_OK;
  AggDesc A; int rc = fill_desc(A, nlanes, ops, in_lanes); if (rc) return rc;
  int lrc = 0;
  DRYAD_DISPATCH_L(nlanes, (lrc = launch_update<L>(tbl_keys, cap, A, tbl_vals, in_keys, n, combine, status, (cudaStream_t)stream)));
  return lrc;
}

/* Same as dryad_hash_agg_update, plus `work` (4 int32 words of device memory owned by the caller):
 * the call probes the first rows and lets the DEVICE pick between the plain kernel and the
 * shared-memory front-table kernel for the rest (few / skewed keys), without a host sync. */
int dryad_hash_agg_update_auto(uint64_t* tbl_keys, uint64_t cap, int nlanes, const int* ops,
                               int64_t* tbl_vals, const uint64_t* in_keys,
                               const int64_t* const* in_lanes, int64_t n, int combine, int32_t* status,
                               int32_t* work, void* stream) {
  if (!cap_ok(cap)) return DRYAD_ERR_BAD_ARG;
  if (n <= 0) return DRYAD_OK;
  AggDesc A; int rc = fill_desc(A, nlanes, ops, in_lanes); if (rc) return rc;
  int lrc = 0;
  DRYAD_DISPATCH_L(nlanes, (lrc = launch_update<L>(tbl_keys, cap, A, tbl_vals, in_keys, n, combine, status, (cudaStream_t)stream, work)));
  return lrc;
}

int dryad_hash_agg_merge_peers(const uint64_t* tbl_keys, uint64_t cap, int nlanes, const int* ops,
                               const int64_t* tbl_vals, int world, uint64_t* const* peer_keys,
                               int64_t* const* peer_vals /* [world] */, uint64_t peer_cap,
                               int32_t* status, void* stream) {
  if (!cap_ok(cap) || !cap_ok(peer_cap) || world < 1 || world > DRYAD_MAX_PEERS) return DRYAD_ERR_BAD_ARG;
  AggDesc A; int rc = fill_desc(A, nlanes, ops, nullptr); if (rc) return rc;
  PeerTables T; T.world = world; T.cap = peer_cap;
  for (int r = 0; r < DRYAD_MAX_PEERS; ++r) { T.keys[r] = nullptr; T.vals[r] = nullptr; }
  for (int r = 0; r < world; ++r) { T.keys[r] = peer_keys[r]; T.vals[r] = peer_vals[r]; }
  int64_t g = ((int64_t)cap + 255) / 256; if (g > 148 * 16) g = 148 * 16;
  DRYAD_DISPATCH_L(nlanes, (agg_merge_peers_kernel<L><<<(int)g, 256, 0, (cudaStream_t)stream>>>(tbl_keys, cap, A, tbl_vals, T, status)));
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

// One pass min/max of a signed 64-bit column (key-range statistics for the dense-key decision).
__global__ void minmax_init_kernel(long long* out) { out[0] = INT64_MAX; out[1] = INT64_MIN; }
__global__ void __launch_bounds__(256) minmax_kernel(const long long* __restrict__ x, int64_t n, long long* out) {
  long long lo = INT64_MAX, hi = INT64_MIN;
  int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 2;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x * 2;
  for (; i + 1 < n; i += stride) {
    const longlong2 v = __ldcs(reinterpret_cast<const longlong2*>(x + i));
    lo = min(lo, min(v.x, v.y)); hi = max(hi, max(v.x, v.y));
  }
  if (i < n) { const long long v = x[i]; lo = min(lo, v); hi = max(hi, v); }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, d));
    hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, d));
  }
  __shared__ long long slo[8], shi[8];
  if ((threadIdx.x & 31) == 0) { slo[threadIdx.x >> 5] = lo; shi[threadIdx.x >> 5] = hi; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; ++w) { lo = min(lo, slo[w]); hi = max(hi, shi[w]); }
    atomicMin(out, lo); atomicMax(out + 1, hi);
  }
}

// ---- direct-addressed tables: dense key ranges need neither a key array nor probing ----------
// slot = key - key_lo.  Occupancy is a row counter: the table's own COUNT lane when it has one,
// else a hidden COUNT lane appended after the user lanes (so the lane stride may grow).
static int fill_direct(AggDesc& A, int nlanes, const int* ops, const int64_t* const* in, uint64_t key_lo) {
  int rc = fill_desc(A, nlanes, ops, in); if (rc) return rc;
  A.direct = 1; A.key_lo = key_lo; A.occ_lane = -1;
  for (int l = 0; l < nlanes; ++l) if (ops[l] == DRYAD_AGG_COUNT) { A.occ_lane = l; break; }
  if (A.occ_lane < 0) {
    if (nlanes >= DRYAD_MAX_LANES) return DRYAD_ERR_BAD_ARG;
    A.op[nlanes] = DRYAD_AGG_COUNT; A.in[nlanes] = nullptr; A.occ_lane = nlanes; A.nlanes = nlanes + 1;
  }
  A.src_stride = lane_stride(A.nlanes);
  return DRYAD_OK;
}

int dryad_minmax_i64(const int64_t* x, int64_t n, int64_t* out_min_max, void* stream) {
  if (((uintptr_t)x & 15) != 0) return DRYAD_ERR_BAD_ARG;
  minmax_init_kernel<<<1, 1, 0, (cudaStream_t)stream>>>((long long*)out_min_max);
  DRYAD_LAUNCH_CHECK();
  if (n <= 0) return DRYAD_OK;
  int64_t g = (n + 2047) / 2048; if (g > 148 * 8) g = 148 * 8;
  minmax_kernel<<<(int)g, 256, 0, (cudaStream_t)stream>>>((const long long*)x, n, (long long*)out_min_max);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_hash_agg_direct_lane_stride(int nlanes, const int* ops) {
  AggDesc A; if (fill_direct(A, nlanes, ops, nullptr, 0)) return -1;
  return lane_stride(A.nlanes);
}

int dryad_hash_agg_init_direct(int64_t* tbl_vals, uint64_t cap, int nlanes, const int* ops, void* stream) {
  if (!cap_ok(cap)) return DRYAD_ERR_BAD_ARG;
  AggDesc A; int rc = fill_direct(A, nlanes, ops, nullptr, 0); if (rc) return rc;
  DRYAD_DISPATCH_L(A.nlanes, (agg_init_kernel<L><<<grid_for((int64_t)cap * L), 256, 0, (cudaStream_t)stream>>>(nullptr, cap, A, tbl_vals)));
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

/* Raw rows only (COUNT lanes count rows).  status 4: a key fell outside [key_lo, key_lo + cap). */
int dryad_hash_agg_update_direct(int64_t* tbl_vals, uint64_t cap, int nlanes, const int* ops, uint64_t key_lo,
                                 const uint64_t* in_keys, const int64_t* const* in_lanes, int64_t n,
                                 int32_t* status, void* stream) {
  if (!cap_ok(cap)) return DRYAD_ERR_BAD_ARG;
  if (n <= 0) return DRYAD_OK;
  AggDesc A; int rc = fill_direct(A, nlanes, ops, in_lanes, key_lo); if (rc) return rc;
  int lrc = 0;
  DRYAD_DISPATCH_L(A.nlanes, (lrc = launch_update<L>(nullptr, cap, A, tbl_vals, in_keys, n, 0, status, (cudaStream_t)stream)));
  return lrc;
}

int dryad_hash_agg_update_direct_auto(int64_t* tbl_vals, uint64_t cap, int nlanes, const int* ops, uint64_t key_lo,
                                      const uint64_t* in_keys, const int64_t* const* in_lanes, int64_t n,
                                      int32_t* status, int32_t* work, void* stream) {
  if (!cap_ok(cap)) return DRYAD_ERR_BAD_ARG;
  if (n <= 0) return DRYAD_OK;
  AggDesc A; int rc = fill_direct(A, nlanes, ops, in_lanes, key_lo); if (rc) return rc;
  int lrc = 0;
  DRYAD_DISPATCH_L(A.nlanes, (lrc = launch_update<L>(nullptr, cap, A, tbl_vals, in_keys, n, 0, status, (cudaStream_t)stream, work)));
  return lrc;
}

int dryad_hash_agg_extract_direct(const int64_t* tbl_vals, uint64_t cap, int nlanes, const int* ops, uint64_t key_lo,
                                  uint64_t* out_keys, int64_t* const* out_lanes_dev, unsigned long long* out_count,
                                  int64_t out_cap, void* stream) {
  if (!cap_ok(cap)) return DRYAD_ERR_BAD_ARG;
  AggDesc A; int rc = fill_direct(A, nlanes, ops, nullptr, key_lo); if (rc) return rc;
  DRYAD_DISPATCH_L(A.nlanes, (agg_extract_kernel<L><<<grid_for((int64_t)cap), 256, 0, (cudaStream_t)stream>>>(nullptr, cap, A, tbl_vals, out_keys, out_lanes_dev, out_count, out_cap)));
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

/* Local direct table -> the key owners' HASHED tables (lane stride of nlanes user lanes). */
int dryad_hash_agg_merge_peers_direct(const int64_t* tbl_vals, uint64_t cap, int nlanes, const int* ops, uint64_t key_lo,
                                      int world, uint64_t* const* peer_keys, int64_t* const* peer_vals,
                                      uint64_t peer_cap, int32_t* status, void* stream) {
  if (!cap_ok(cap) || !cap_ok(peer_cap) || world < 1 || world > DRYAD_MAX_PEERS) return DRYAD_ERR_BAD_ARG;
  AggDesc A; int rc = fill_direct(A, nlanes, ops, nullptr, key_lo); if (rc) return rc;
  PeerTables T; T.world = world; T.cap = peer_cap;
  for (int r = 0; r < DRYAD_MAX_PEERS; ++r) { T.keys[r] = nullptr; T.vals[r] = nullptr; }
  for (int r = 0; r < world; ++r) { T.keys[r] = peer_keys[r]; T.vals[r] = peer_vals[r]; }
  int64_t g = ((int64_t)cap + 255) / 256; if (g > 148 * 16) g = 148 * 16;
  DRYAD_DISPATCH_L(nlanes, (agg_merge_peers_kernel<L><<<(int)g, 256, 0, (cudaStream_t)stream>>>(nullptr, cap, A, tbl_vals, T, status)));
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

/* Fold packed rows {key, v[0..nlanes)} (stride 1 + nlanes words) into the table; combine != 0
 * treats the values as partial aggregates (COUNT partials add). */
int dryad_hash_agg_update_rows(uint64_t* tbl_keys, uint64_t cap, int nlanes, const int* ops,
                               int64_t* tbl_vals, const int64_t* rows, int64_t n, int combine,
                               int32_t* status, void* stream) {
  if (!cap_ok(cap)) return DRYAD_ERR_BAD_ARG;
  if (n <= 0) return DRYAD_OK;
  AggDesc A; int rc = fill_desc(A, nlanes, ops, nullptr); if (rc) return rc;
  A.in_stride = 1 + nlanes;
  for (int l = 0; l < nlanes; ++l) A.in[l] = rows + 1 + l;
  int lrc = 0;
  DRYAD_DISPATCH_L(nlanes, (lrc = launch_update<L>(tbl_keys, cap, A, tbl_vals, reinterpret_cast<const uint64_t*>(rows), n, combine, status, (cudaStream_t)stream)));
  return lrc;
}

int dryad_hash_agg_extract_rows(const uint64_t* tbl_keys, uint64_t cap, int nlanes, const int* ops,
                                const int64_t* tbl_vals, int64_t* out_rows,
                                unsigned long long* out_count, int64_t out_cap, void* stream) {
  if (!cap_ok(cap)) return DRYAD_ERR_BAD_ARG;
  AggDesc A; int rc = fill_desc(A, nlanes, ops, nullptr); if (rc) return rc;
  DRYAD_DISPATCH_L(nlanes, (agg_extract_rows_kernel<L><<<grid_for((int64_t)cap), 256, 0, (cudaStream_t)stream>>>(tbl_keys, cap, A, tbl_vals, out_rows, out_count, out_cap)));
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_hash_agg_extract(const uint64_t* tbl_keys, uint64_t cap, int nlanes, const int* ops,
                           const int64_t* tbl_vals, uint64_t* out_keys,
                           int64_t* const* out_lanes_dev, unsigned long long* out_count,
                           int64_t out_cap, void* stream) {
  if (!cap_ok(cap)) return DRYAD_ERR_BAD_ARG;
  AggDesc A; int rc = fill_desc(A, nlanes, ops, nullptr); if (rc) return rc;
  DRYAD_DISPATCH_L(nlanes, (agg_extract_kernel<L><<<grid_for((int64_t)cap), 256, 0, (cudaStream_t)stream>>>(tbl_keys, cap, A, tbl_vals, out_keys, out_lanes_dev, out_count, out_cap)));
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_join_build(uint64_t* tbl_keys, int32_t* head, uint64_t cap, const uint64_t* build_keys,
                     int32_t* next, int64_t n, int32_t* status, void* stream) {
  if (!cap_ok(cap) || n >= (1ll << 31)) return DRYAD_ERR_BAD_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  join_init_kernel<<<grid_for((int64_t)cap), 256, 0, st>>>(tbl_keys, head, cap);
  DRYAD_LAUNCH_CHECK();
  if (n > 0) {
    join_build_kernel<<<grid_for(n), 256, 0, st>>>(tbl_keys, head, cap, build_keys, next, n, status);
    DRYAD_LAUNCH_CHECK();
  }
  return DRYAD_OK;
}

int dryad_join_probe(const uint64_t* tbl_keys, const int32_t* head, uint64_t cap,
                     const int32_t* next, const uint64_t* probe_keys, int64_t n, int mode,
                     int64_t* counts_or_offsets, int32_t* out_probe, int32_t* out_build,
                     void* stream) {
  if (!cap_ok(cap)) return DRYAD_ERR_BAD_ARG;
  if (n <= 0) return DRYAD_OK;
  join_probe_kernel<<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(tbl_keys, head, cap, next, probe_keys, n, mode, counts_or_offsets, out_probe, out_build);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_replicate_slice(int world, int ncols, const int64_t* const* src, int64_t* const* peer_dst /* [world][ncols] */,
                          int64_t n_local, int64_t row_base, void* stream) {
  if (world < 1 || world > DRYAD_MAX_PEERS || ncols < 1 || ncols > DRYAD_MAX_LANES + 1) return DRYAD_ERR_BAD_ARG;
  if (n_local <= 0) return DRYAD_OK;
  ReplTables T; T.world = world; T.ncols = ncols;
  for (int c = 0; c < ncols; ++c) T.src[c] = src[c];
  for (int r = 0; r < world; ++r) for (int c = 0; c < ncols; ++c) T.dst[r][c] = peer_dst[r * ncols + c];
  replicate_slice_kernel<<<grid_for(n_local), 256, 0, (cudaStream_t)stream>>>(T, n_local, row_base);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_join_probe_unique(const uint64_t* tbl_keys, const int32_t* head, uint64_t cap,
                            const uint64_t* probe_keys, int64_t n, int32_t* out_row, uint8_t* matched,
                            void* stream) {
  if (!cap_ok(cap)) return DRYAD_ERR_BAD_ARG;
  if (n <= 0) return DRYAD_OK;
  join_probe_unique_kernel<<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(tbl_keys, head, cap, probe_keys, n, out_row, matched);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_join_table_init(uint64_t* tbl_keys, int32_t* head, uint64_t cap, void* stream) {
  if (!cap_ok(cap)) return DRYAD_ERR_BAD_ARG;
  join_init_kernel<<<grid_for((int64_t)cap), 256, 0, (cudaStream_t)stream>>>(tbl_keys, head, cap);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_join_build_bcast(const uint64_t* keys_local, int64_t n_local, int64_t row_base, int world,
                           uint64_t* const* peer_keys, int32_t* const* peer_head, int32_t* const* peer_next,
                           uint64_t cap, int ncols, const int64_t* const* col_src,
                           int64_t* const* peer_cols /* [world][ncols] */, int32_t* status, void* stream) {
  if (!cap_ok(cap) || world < 1 || world > DRYAD_MAX_PEERS || ncols < 0 || ncols > DRYAD_MAX_LANES) return DRYAD_ERR_BAD_ARG;
  if (n_local <= 0) return DRYAD_OK;
  BcastTables T; T.world = world; T.ncols = ncols; T.cap = cap;
  for (int r = 0; r < world; ++r) {
    T.keys[r] = peer_keys[r]; T.head[r] = peer_head[r]; T.next[r] = peer_next[r];
    for (int c = 0; c < ncols; ++c) T.cols[r][c] = peer_cols[r * ncols + c];
  }
  for (int c = 0; c < ncols; ++c) T.src[c] = col_src[c];
  join_build_bcast_kernel<<<grid_for(n_local), 256, 0, (cudaStream_t)stream>>>(T, keys_local, n_local, row_base, status);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_set_contains(const uint64_t* tbl_keys, uint64_t cap, const uint64_t* probe_keys,
                       int64_t n, uint8_t* out, void* stream) {
  if (!cap_ok(cap)) return DRYAD_ERR_BAD_ARG;
  if (n <= 0) return DRYAD_OK;
  set_contains_kernel<<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(tbl_keys, cap, probe_keys, n, out);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

}  // extern "C"

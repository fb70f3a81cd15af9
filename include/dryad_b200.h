/* dryad_b200 — C ABI of the sm_100a kernel library (libdryad_kernels.so).
 *
 * Every entry point takes raw device pointers and a cudaStream_t (as void*), launches on that
 * stream and returns 0 on success, a positive cudaError_t, or a negative DRYAD_ERR_* code.
 * No entry point allocates device memory: callers pass outputs and, where a kernel needs
 * working storage, a scratch buffer sized by the matching *_scratch_bytes() query.
 *
 * Each declaration cites the reference code whose behaviour it provides (paths relative to the
 * Liuqin/Dryad tree).
 */
#ifndef DRYAD_B200_H
#define DRYAD_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DRYAD_MAX_PEERS_C 8
#define DRYAD_FLAG_SLOTS 8      /* flags buffer = uint64[DRYAD_FLAG_SLOTS][8] per rank */
#define DRYAD_MAX_LANES 8
#define DRYAD_EXPR_MAX_INSTR 64
#define DRYAD_EXPR_MAX_COLS 16

enum {
  DRYAD_OK = 0,
  DRYAD_ERR_BAD_ARG = -1,
  DRYAD_ERR_TOO_LARGE = -2,
  DRYAD_ERR_OVERFLOW = -3,
  DRYAD_ERR_SCRATCH = -4
};

/* ---- utility ------------------------------------------------------------------------------ */
int dryad_device_count(void);
const char* dryad_error_string(int code);
int dryad_memcpy_async(void* dst, const void* src, size_t bytes, int kind, void* stream);
int dryad_memset_async(void* dst, int value, size_t bytes, void* stream);

/* ---- symmetric memory (peer-mapped HBM) ---------------------------------------------------
 * Role of DryadVertex/VertexHost/system/channel (file/HTTP channels between vertices). */
int dryad_symm_alloc(size_t bytes, void** out);
int dryad_symm_free(void* p);
int dryad_ipc_export(void* p, unsigned char handle64[64]);
int dryad_ipc_open(const unsigned char handle64[64], void** out);
int dryad_ipc_close(void* p);
/* Epoch barrier over peer flags (slot < DRYAD_FLAG_SLOTS); status: 0 ok, 2 = peer timeout. */
int dryad_signal_barrier(uint64_t* const* flags_peers, uint64_t* flags_local, int rank, int world,
                         int slot, uint64_t epoch, int32_t* status, void* stream);

/* ---- K14 key hashing: LinqToDryad/DryadLinqVertex.cs:4801, LinqToDryad/Hash64.cs ----------- */
int dryad_hash_u64(const uint64_t* keys, uint64_t* out, int64_t n, void* stream);
int dryad_hash_bytes(const uint8_t* bytes, const int64_t* offsets, uint64_t* out, int64_t n,
                     void* stream);

/* ---- K4 radix sort: LinqToDryad/DryadLinqVertex.cs:293-316, 9321-9818 ----------------------
 * Sorts (key, u32 value) pairs on key bits [begin_bit, end_bit), ascending, stable.  Buffers
 * ping-pong between x and x_alt; *result_in_alt tells where the result is.  vals may be NULL
 * (keys only).  iota_vals != 0: values start as 0..n-1 and vals' initial content is ignored.
 * n < 2^30. */
size_t dryad_radix_sort_scratch_bytes(int64_t n, int begin_bit, int end_bit);
int dryad_radix_sort_pairs_u64(uint64_t* keys, uint64_t* keys_alt, uint32_t* vals,
                               uint32_t* vals_alt, int64_t n, int begin_bit, int end_bit,
                               int iota_vals, void* scratch, size_t scratch_bytes,
                               int* result_in_alt, void* stream);
int dryad_radix_sort_pairs_u32(uint32_t* keys, uint32_t* keys_alt, uint32_t* vals,
                               uint32_t* vals_alt, int64_t n, int begin_bit, int end_bit,
                               int iota_vals, void* scratch, size_t scratch_bytes,
                               int* result_in_alt, void* stream);

/* Allocation-free sort.  No counterpart in the reference: it exists for callers that cannot
 * allocate device memory — a host program whose own pool holds all of HBM, a thread that must not
 * block on the driver's allocator, or a C program with no allocator library.  Same contract as
 * dryad_radix_sort_pairs_u64, but the working storage is a fixed array inside the library
 * (DRYAD_SORT_NOALLOC_SCRATCH bytes per GPU, held until the process exits), so n is limited to
 * DRYAD_SORT_NOALLOC_MAX_N pairs; larger n returns DRYAD_ERR_TOO_LARGE ("size beyond supported
 * maximum").  The call itself never calls cudaMalloc/cudaFree or any allocating library routine;
 * the caller still provides keys/vals and their _alt ping-pong buffers.
 * - The first call on each device looks up the array's device address (cudaGetSymbolAddress, with
 *   that device current); make that first call before starting any stream capture.
 * - All calls on a device share the one array: two calls whose kernels overlap on the GPU
 *   (different streams, no ordering between them) race.  Serialise them on one stream. */
#define DRYAD_SORT_NOALLOC_MAX_N (1ll << 22)
#define DRYAD_SORT_NOALLOC_SCRATCH 716032   /* 2 x 8 KiB histograms + 256 + 683 tiles x 1 KiB */
int dryad_radix_sort_pairs_u64_noalloc(uint64_t* keys, uint64_t* keys_alt, uint32_t* vals,
                                       uint32_t* vals_alt, int64_t n, int begin_bit, int end_bit,
                                       int iota_vals, int* result_in_alt, void* stream);

/* ---- K1/K2 + C1 fused partition + all-to-all: LinqToDryad/DryadLinqVertex.cs:4788-5151,
 *      DryadLinqGraphManager/GraphBuilder.cs:481-504 ------------------------------------------ */
enum { DRYAD_PART_HASH = 0, DRYAD_PART_RANGE_ASC = 1, DRYAD_PART_RANGE_DESC = 2, DRYAD_PART_DIRECT = 3 };

typedef struct DryadShuffleDesc {
  int32_t ncols;           /* payload columns, 1..8 */
  int32_t nparts;          /* destination partitions, 1..256 (each may live on any GPU) */
  int32_t mode;            /* DRYAD_PART_* */
  int32_t key_col;         /* column that carries the key */
  int32_t key_kind;        /* 0: the column is a uint64 key; 1: 8 big-endian bytes at key_off */
  int32_t key_off;
  int32_t emit_key;        /* 1: also write the derived uint64 key to dst_key_table[part] */
  int32_t rank, world;
  int64_t n;               /* local rows */
  const void* src[8];      /* local column bases (row-major, width[c] bytes per row) */
  uint32_t width[8];       /* bytes per row, multiple of 4, <= 512 */
  void* const* dst_table;          /* device array [ncols][nparts]: partition buffers (peer-mapped) */
  uint64_t* const* dst_key_table;  /* device array [nparts] (emit_key) */
  const uint64_t* separators;      /* device, nparts-1 keys (range modes) */
  const int64_t* capacity;         /* device [nparts]: rows each partition buffer can hold */
  int64_t* counts_local;           /* [world][nparts]: filled by all ranks; rows per (src, part) */
  int64_t* const* counts_peers;    /* device array [world]: every rank's counts matrix */
  uint64_t* flags_local;           /* uint64[DRYAD_FLAG_SLOTS][8] */
  uint64_t* const* flags_peers;    /* device array [world] */
  uint64_t epoch;                  /* strictly increasing per shuffle on this flag set */
  void* scratch; size_t scratch_bytes;
  int32_t* status;                 /* device: 0 ok, 1 capacity overflow, 2 peer timeout */
  /* Split execution (all optional; zero = the whole chain on one stream, barrier slot 1).
   * A shuffle can run as PLAN once and then several SCATTER|BARRIER calls that reuse the plan's
   * block offsets (same n, nparts, mode, separators, scratch and epoch) for different columns on
   * different streams: the key stream first, the wide payload behind it, each with its own barrier
   * slot.  Row i of every column lands in slot i of its destination buffer. */
  uint64_t* key_out;               /* PLAN also writes each row's derived uint64 key here (n entries) */
  int32_t phases;                  /* DRYAD_SHUFFLE_* bits */
  int32_t barrier_slot;            /* flag slot of the closing barrier */
  int32_t max_ctas;                /* cap on the scatter grid (persistent CTAs); 0 = one CTA per row block */
  /* Passes: a SCATTER call with filter_mod > 0 moves only the partitions p with p % filter_mod ==
   * filter_rem (the other rows are read and skipped).  With parts_per_rank = R key sub-ranges per
   * destination, pass j fills every destination's sub-range-j buffer; give each pass's closing
   * barrier its own barrier_epoch. */
  int32_t filter_mod, filter_rem;
  uint64_t barrier_epoch;          /* 0 = epoch */
} DryadShuffleDesc;
enum { DRYAD_SHUFFLE_PLAN = 1, DRYAD_SHUFFLE_SCATTER = 2, DRYAD_SHUFFLE_BARRIER = 4 };

size_t dryad_shuffle_scratch_bytes(int64_t n, int nparts);
int dryad_shuffle_run(const DryadShuffleDesc* desc, void* stream);

/* ---- K3 sampler + separators: LinqToDryad/DryadLinqSampler.cs:42-106, 139-246 -------------- */
int dryad_sample_keys(const uint64_t* keys, int64_t n, uint64_t seed, double rate, uint64_t* out,
                      unsigned long long* out_count, int64_t cap, void* stream);
int dryad_sample_rows(const uint8_t* rows, int width, int key_off, int64_t n, uint64_t seed,
                      double rate, uint64_t* out, unsigned long long* out_count, int64_t cap,
                      void* stream);
int dryad_select_separators(const uint64_t* sorted, int64_t cnt, int pcount, int descending,
                            uint64_t* seps, void* stream);
/* Same, the count stays on the device: cnt = sum_r min(counts[r], cap_each); `sorted` is the sorted
 * concatenation of fixed-size sample slabs whose unused slots were pre-filled with ~0. */
int dryad_select_separators_dev(const uint64_t* sorted, const int64_t* counts, int ncounts, int64_t cap_each,
                                int pcount, int descending, uint64_t* seps, void* stream);

/* ---- row helpers for OrderBy on wide records ----------------------------------------------- */
int dryad_extract_be64(const uint8_t* rows, int64_t n, int width, int key_off, uint64_t* out,
                       void* stream);
/* Orders runs of equal (key & prefix_mask) by `extra` further key bytes read from the rows. */
int dryad_fix_ties(const uint64_t* keys, uint32_t* idx, int64_t n, const uint8_t* rows, int width,
                   int extra_off, int extra, int max_run, int descending, uint64_t prefix_mask,
                   int32_t* long_runs, void* stream);
/* 32-bit key window variants used by the wide-record OrderBy fast path. */
int dryad_extract_be_window(const uint8_t* rows, int64_t n, int width, int key_off, int shift, uint32_t* out, void* stream);
int dryad_shift_u64_to_u32(const uint64_t* in, int64_t n, int shift, uint32_t* out, void* stream);
int dryad_fix_ties32(const uint32_t* keys, uint32_t* idx, int64_t n, const uint8_t* rows, int width, int extra_off,
                     int extra, int max_run, int descending, int32_t* long_runs, void* stream);
int dryad_gather_rows(const void* in, const uint32_t* idx, void* out, int64_t n, int width,
                      void* stream);
/* Key-only tie repair for the split (keys first) exchange: runs of equal 32-bit windows are ordered
 * by keys64[idx] then idx; `unresolved` counts runs holding equal 64-bit prefixes when wide_key != 0
 * (they need the rows' remaining key bytes: fall back to dryad_fix_ties32 once the payload is in). */
int dryad_fix_ties32_keys(const uint32_t* keys, uint32_t* idx, int64_t n, const uint64_t* keys64, int max_run,
                          int wide_key, int32_t* long_runs, int32_t* unresolved, void* stream);
int dryad_check_sorted_rows(const uint8_t* rows, int64_t n, int width, int key_off, int key_len,
                            int descending, unsigned long long* result2, void* stream);

/* ---- K6 + C2 hash aggregation: LinqToDryad/DryadLinqVertex.cs:5718-6416,
 *      LinqToDryad/DryadLinqCollection.cs:519-722, LinqToDryad/DryadLinqQueryGen.cs:2179-2296 -- */
enum {
  DRYAD_AGG_COUNT = 0, DRYAD_AGG_SUM_I64 = 1, DRYAD_AGG_SUM_I64_CHECKED = 2, DRYAD_AGG_SUM_F64 = 3,
  DRYAD_AGG_MIN_I64 = 4, DRYAD_AGG_MAX_I64 = 5, DRYAD_AGG_MIN_F64 = 6, DRYAD_AGG_MAX_F64 = 7
};
/* Table = keys[cap] + vals[cap * L].  cap is any multiple of 4 (keys are probed in 4-key buckets,
 * one 32-byte sector per probe; key ~0 is reserved).  Values are slot-major: the aggregates of
 * slot s are the int64/double words vals[s*L .. s*L + nlanes), L = dryad_hash_agg_lane_stride(
 * nlanes) (1, 2, 4 or 8), so one row's reductions hit one sector and are issued by adjacent
 * threads of a single warp instruction.  The same probing scheme serves the join and set tables
 * below.  status: 1 table full, 3 checked-sum overflow. */
int dryad_hash_agg_lane_stride(int nlanes);
int dryad_hash_agg_init(uint64_t* tbl_keys, uint64_t cap, int nlanes, const int* ops,
                        int64_t* tbl_vals, void* stream);
int dryad_hash_agg_update(uint64_t* tbl_keys, uint64_t cap, int nlanes, const int* ops,
                          int64_t* tbl_vals, const uint64_t* in_keys,
                          const int64_t* const* in_lanes, int64_t n, int combine, int32_t* status,
                          void* stream);
/* Adaptive forms: `work` = 4 int32 words of caller-owned device memory.  The first 8 Ki rows go
 * through a per-CTA shared-memory front table and measure its hit rate; the device then picks, for
 * the remaining rows, either the plain kernel (many distinct keys) or the front-table kernel (few or
 * skewed keys: same-address reductions would serialise in the L2).  No host synchronisation. */
int dryad_hash_agg_update_auto(uint64_t* tbl_keys, uint64_t cap, int nlanes, const int* ops,
                               int64_t* tbl_vals, const uint64_t* in_keys,
                               const int64_t* const* in_lanes, int64_t n, int combine,
                               int32_t* status, int32_t* work, void* stream);
int dryad_hash_agg_update_direct_auto(int64_t* tbl_vals, uint64_t cap, int nlanes, const int* ops,
                                      uint64_t key_lo, const uint64_t* in_keys,
                                      const int64_t* const* in_lanes, int64_t n, int32_t* status,
                                      int32_t* work, void* stream);
/* Fused reduce-scatter: fold local partials into the key owner's table (owner =
 * hash_bucket(key, world)) with system-scope atomics on peer HBM.  peer_vals[r] is rank r's
 * vals array (same lane stride). */
int dryad_hash_agg_merge_peers(const uint64_t* tbl_keys, uint64_t cap, int nlanes, const int* ops,
                               const int64_t* tbl_vals, int world, uint64_t* const* peer_keys,
                               int64_t* const* peer_vals, uint64_t peer_cap, int32_t* status,
                               void* stream);
/* Direct-addressed tables for DENSE key ranges: slot = key - key_lo, no key array, no probing (one
 * L2 reduction request per row instead of a probe plus a reduction).  `tbl_vals` has
 * cap * dryad_hash_agg_direct_lane_stride(nlanes, ops) words: when no lane is a COUNT, a hidden row
 * counter lane is appended to tell occupied slots apart.  Raw rows only (no `combine`).
 * status 4: a key outside [key_lo, key_lo + cap).  dryad_minmax_i64 gives the range in one pass
 * (out_min_max: 2 device words). */
int dryad_minmax_i64(const int64_t* x, int64_t n, int64_t* out_min_max, void* stream);
int dryad_hash_agg_direct_lane_stride(int nlanes, const int* ops);
int dryad_hash_agg_init_direct(int64_t* tbl_vals, uint64_t cap, int nlanes, const int* ops, void* stream);
int dryad_hash_agg_update_direct(int64_t* tbl_vals, uint64_t cap, int nlanes, const int* ops,
                                 uint64_t key_lo, const uint64_t* in_keys,
                                 const int64_t* const* in_lanes, int64_t n, int32_t* status,
                                 void* stream);
int dryad_hash_agg_extract_direct(const int64_t* tbl_vals, uint64_t cap, int nlanes, const int* ops,
                                  uint64_t key_lo, uint64_t* out_keys, int64_t* const* out_lanes_dev,
                                  unsigned long long* out_count, int64_t out_cap, void* stream);
int dryad_hash_agg_merge_peers_direct(const int64_t* tbl_vals, uint64_t cap, int nlanes,
                                      const int* ops, uint64_t key_lo, int world,
                                      uint64_t* const* peer_keys, int64_t* const* peer_vals,
                                      uint64_t peer_cap, int32_t* status, void* stream);

/* Packed-row forms: rows of (1 + nlanes) int64 words {key, v[0..nlanes)} — the record the fused
 * shuffle carries when partials are PUSHED to their key owners and folded there locally. */
int dryad_hash_agg_update_rows(uint64_t* tbl_keys, uint64_t cap, int nlanes, const int* ops,
                               int64_t* tbl_vals, const int64_t* rows, int64_t n, int combine,
                               int32_t* status, void* stream);
int dryad_hash_agg_extract_rows(const uint64_t* tbl_keys, uint64_t cap, int nlanes, const int* ops,
                                const int64_t* tbl_vals, int64_t* out_rows,
                                unsigned long long* out_count, int64_t out_cap, void* stream);
int dryad_hash_agg_extract(const uint64_t* tbl_keys, uint64_t cap, int nlanes, const int* ops,
                           const int64_t* tbl_vals, uint64_t* out_keys,
                           int64_t* const* out_lanes_dev, unsigned long long* out_count,
                           int64_t out_cap, void* stream);

/* Host callback on a stream (cudaLaunchHostFunc): fn(arg) runs once all prior work of the stream is
 * complete.  With libdryad_host's pump_notifier_fn / pump_make_notifier this posts a completion
 * message into the graph manager's message pump without anyone blocking on the stream. */
int dryad_stream_host_func(void* stream, void* fn, void* arg);

/* ---- HBM ring FIFO between two GPUs (flag-based flow control, no host per block) ------------------
 * The ring (slots x slot_bytes), the per-slot length words and `head` live in the CONSUMER's memory,
 * `tail` in the PRODUCER's; `*_peer` pointers are the peer-mapped addresses of the other side.
 * `done_local` is one zero-initialised u64 of scratch per direction and per side.  seq = number of
 * blocks pushed (popped) before this call.  status 2: the peer did not move within the timeout,
 * 1: block larger than the destination.  Producer and consumer must be different GPUs. */
/* status != 0 (1 = block larger than the destination, 2 = peer timeout) is fatal for the ring: later
 * pushes and pops return immediately without moving data; create a new ring.  `src` is read for exactly
 * nbytes (whole 16-byte units vectorised, the tail bytewise); src must be 16-byte aligned. */
int dryad_ring_push(const void* src, uint64_t nbytes, void* ring_peer, uint64_t* len_peer,
                    uint64_t* head_peer, const uint64_t* tail_local, uint64_t* done_local,
                    uint64_t seq, int slots, uint64_t slot_bytes, int32_t* status, void* stream);
int dryad_ring_pop(void* dst, uint64_t dst_cap, uint64_t* out_len, const void* ring_local,
                   const uint64_t* len_local, const uint64_t* head_local, uint64_t* tail_peer,
                   uint64_t* done_local, uint64_t seq, int slots, uint64_t slot_bytes,
                   int32_t* status, void* stream);

/* ---- Block channel over device partitions (the C ABI between a vertex and the engine):
 *      DryadVertex/VertexHost/vertex/include/wrappernativeinfo.h:103-150, LinqToDryad/DryadLinqNative.cs:76-145,
 *      bounded FIFO with throttling: DryadVertex/VertexHost/system/channel/src/channelfifo.cpp -----------------
 * The pool is a caller-provided device buffer of nblocks * block_bytes bytes (for instance a slice of the
 * symmetric arena): nothing is allocated.  alloc_block waits (up to timeout_s, then DRYAD_ERR_OVERFLOW) while
 * every block is in use: a writer is at most nblocks blocks ahead of the reader.  write_block publishes a block
 * with an event on the writer's stream; read_block returns 0 and the next block (the reader's stream waits for
 * that event), 1 at end of stream, DRYAD_ERR_OVERFLOW on timeout; release_block returns the block to the pool
 * once the work queued on the reader's stream has finished.  checksum != 0 adds an integrity check of every
 * block (checksum kernel on both sides, mismatches counted in stats[4]); it needs `control`, device memory of
 * dryad_chan_control_bytes(nblocks) bytes.  Thread-safe: one writer thread and one reader thread may stream. */
size_t dryad_chan_control_bytes(int nblocks);
int dryad_chan_create(void* buffer, size_t block_bytes, int nblocks, int checksum, void* control, void** out);
int dryad_chan_destroy(void* chan);
int dryad_chan_alloc_block(void* chan, size_t bytes, double timeout_s, void** dev_ptr);
int dryad_chan_write_block(void* chan, void* dev_ptr, size_t nbytes, void* stream);
int dryad_chan_close_writer(void* chan);
int dryad_chan_read_block(void* chan, void** dev_ptr, size_t* nbytes, double timeout_s, void* stream);
int dryad_chan_release_block(void* chan, void* dev_ptr, void* stream);
int64_t dryad_chan_total_length(void* chan);
int dryad_chan_stats(void* chan, int64_t* out5);

/* ---- NVLS (NVSwitch multicast / in-switch reduction) ----------------------------------------------
 * `mc` is the multicast address of a symmetric buffer (every GPU of the group has a copy at the same
 * offset; the mapping is made by the caller, e.g. torch's symmetric-memory rendezvous).
 * allreduce: two-shot, in place — rank r pulls the all-GPU sum of slice r out of the switch
 * (multimem.ld_reduce) and multicasts it back (multimem.st), so all copies hold identical bits.
 * Barrier across the group before and after.  n in elements (f32: multiple of 4).
 * broadcast: local `src` -> the same byte range of every copy (nbytes multiple of 16). */
int dryad_nvls_allreduce_f32(float* mc, int64_t n, int rank, int world, void* stream);
int dryad_nvls_allreduce_u64(uint64_t* mc, int64_t n, int rank, int world, void* stream);
int dryad_nvls_broadcast(void* mc_dst, const void* src, int64_t nbytes, void* stream);

/* ---- K7 sort-based grouping, K15 record transcoding ---------------------------------------------
 * mark_boundaries: flags[i] = 1 where a run of equal keys starts in a SORTED key column (compact
 * the flags to get the CSR run starts).  rows_to_columns / columns_to_rows: fixed-stride packed
 * little-endian records (the partfile record format for numeric schemas) <-> one array per field;
 * field widths 1/2/4/8 bytes, stride <= 512. */
#define DRYAD_TRANSCODE_MAX_COLS 16
int dryad_mark_boundaries(const uint64_t* sorted_keys, int64_t n, uint8_t* flags, void* stream);
int dryad_rows_to_columns(const uint8_t* rows, int64_t n, int stride, int ncols, const int* offs,
                          const int* widths, void* const* cols, void* stream);
int dryad_columns_to_rows(uint8_t* rows, int64_t n, int stride, int ncols, const int* offs,
                          const int* widths, const void* const* cols, void* stream);

/* ---- K8 hash join / K10 hash sets: LinqToDryad/DryadLinqVertex.cs:775-1030, 6703-7315,
 *      1239-1660 ----------------------------------------------------------------------------- */
/* status after a build: 1 = table full, 2 = the build side has duplicate keys (the single-pass
 * unique-key probes must not be used), 0 = all build keys distinct. */
int dryad_join_build(uint64_t* tbl_keys, int32_t* head, uint64_t cap, const uint64_t* build_keys,
                     int32_t* next, int64_t n, int32_t* status, void* stream);
int dryad_join_probe(const uint64_t* tbl_keys, const int32_t* head, uint64_t cap,
                     const int32_t* next, const uint64_t* probe_keys, int64_t n, int mode,
                     int64_t* counts_or_offsets, int32_t* out_probe, int32_t* out_build,
                     void* stream);
int dryad_set_contains(const uint64_t* tbl_keys, uint64_t cap, const uint64_t* probe_keys,
                       int64_t n, uint8_t* out, void* stream);
/* Broadcast join (K8 + C4): replicate this rank's slice of the small side (int64 payload columns)
 * into every rank's replica buffers and insert its keys into every rank's hash table, over peer
 * memory (LinqToDryad/DryadLinqQueryGen.cs:2957-3001 Tee+CrossProduct+Merge, then HashJoin). */
/* K17: copy this rank's slice of up to 9 int64 columns into every rank's replica buffers (P2P stores;
 * role of the "CP" copy vertex, DryadVertex/VertexHost/vertex/vertexHost/vertexHost.cpp:331-571). */
int dryad_replicate_slice(int world, int ncols, const int64_t* const* src, int64_t* const* peer_dst,
                          int64_t n_local, int64_t row_base, void* stream);
int dryad_join_table_init(uint64_t* tbl_keys, int32_t* head, uint64_t cap, void* stream);
/* Probe against a build side whose keys are unique: out_row[i] = build row (0 if none), matched[i]. */
int dryad_join_probe_unique(const uint64_t* tbl_keys, const int32_t* head, uint64_t cap,
                            const uint64_t* probe_keys, int64_t n, int32_t* out_row, uint8_t* matched,
                            void* stream);
int dryad_join_build_bcast(const uint64_t* keys_local, int64_t n_local, int64_t row_base, int world,
                           uint64_t* const* peer_keys, int32_t* const* peer_head,
                           int32_t* const* peer_next, uint64_t cap, int ncols,
                           const int64_t* const* col_src, int64_t* const* peer_cols,
                           int32_t* status, void* stream);

/* ---- K12 fused elementwise programs + compaction: LinqToDryad/DryadLinqVertex.cs:74-198,
 *      LinqToDryad/DryadLinqCodeGen.cs:2084-2138 ----------------------------------------------- */
enum {
  DRYAD_DT_I64 = 0, DRYAD_DT_F64 = 1, DRYAD_DT_I32 = 2, DRYAD_DT_F32 = 3, DRYAD_DT_U8 = 4,
  DRYAD_DT_I16 = 5, DRYAD_DT_U32 = 6, DRYAD_DT_U16 = 7
};
enum {
  DRYAD_OP_LOAD = 0, DRYAD_OP_CONST, DRYAD_OP_ROWID,
  DRYAD_OP_ADD_I, DRYAD_OP_SUB_I, DRYAD_OP_MUL_I, DRYAD_OP_DIV_I, DRYAD_OP_MOD_I, DRYAD_OP_NEG_I,
  DRYAD_OP_ADD_F, DRYAD_OP_SUB_F, DRYAD_OP_MUL_F, DRYAD_OP_DIV_F, DRYAD_OP_NEG_F,
  DRYAD_OP_EQ_I, DRYAD_OP_NE_I, DRYAD_OP_LT_I, DRYAD_OP_LE_I, DRYAD_OP_GT_I, DRYAD_OP_GE_I,
  DRYAD_OP_EQ_F, DRYAD_OP_NE_F, DRYAD_OP_LT_F, DRYAD_OP_LE_F, DRYAD_OP_GT_F, DRYAD_OP_GE_F,
  DRYAD_OP_AND, DRYAD_OP_OR, DRYAD_OP_NOT, DRYAD_OP_BAND, DRYAD_OP_BOR, DRYAD_OP_BXOR,
  DRYAD_OP_BNOT, DRYAD_OP_SHL, DRYAD_OP_SHR, DRYAD_OP_I2F, DRYAD_OP_F2I, DRYAD_OP_SELECT,
  DRYAD_OP_HASH, DRYAD_OP_ABS_I, DRYAD_OP_ABS_F, DRYAD_OP_MIN_I, DRYAD_OP_MAX_I, DRYAD_OP_MIN_F,
  DRYAD_OP_MAX_F, DRYAD_OP_SQRT_F, DRYAD_OP_FLOOR_F, DRYAD_OP_MOV, DRYAD_OP_STORE,
  DRYAD_OP__COUNT
};
typedef struct DryadExprInstr { int32_t op, dst, a, b; int64_t imm; } DryadExprInstr;
typedef struct DryadExprProgram {
  int32_t ninstr, ncols_in, ncols_out, result_reg;
  DryadExprInstr ins[DRYAD_EXPR_MAX_INSTR];
  const void* in[DRYAD_EXPR_MAX_COLS];
  int32_t in_dtype[DRYAD_EXPR_MAX_COLS];
  void* out[DRYAD_EXPR_MAX_COLS];
  int32_t out_dtype[DRYAD_EXPR_MAX_COLS];
} DryadExprProgram;
/* Runs the register program once per row; flags_out (optional) receives result_reg != 0. */
int dryad_expr_run(const DryadExprProgram* prog, int64_t n, uint8_t* flags_out, void* stream);
size_t dryad_compact_scratch_bytes(int64_t n);
int dryad_compact_flags(const uint8_t* flags, int64_t n, uint32_t* out_idx, int64_t* count_dev,
                        void* scratch, size_t scratch_bytes, void* stream);
/* K16: exclusive scan of m counts into m+1 int64 offsets (LinqToDryad/DryadLinqEnumerable.cs:1131). */
int dryad_exclusive_scan_u32(const uint32_t* in, int64_t m, int64_t* out_m_plus_1, void* stream);
/* K16: exclusive scan of n int64 values in place (join match counts -> output offsets, partition
 * counts -> global indices; LinqToDryad/DryadLinqEnumerable.cs:1131 Offsets); total_dev[0] = sum. */
size_t dryad_scan_i64_scratch_bytes(int64_t n);
int dryad_exclusive_scan_i64(int64_t* x, int64_t n, int64_t* total_dev, void* scratch, size_t scratch_bytes, void* stream);

/* ---- K11 scalar aggregates: LinqToDryad/DryadLinqVertex.cs:1673-4704 ------------------------ */
enum { DRYAD_RED_SUM = 0, DRYAD_RED_MIN = 1, DRYAD_RED_MAX = 2 };
size_t dryad_reduce_scratch_bytes(void);
int dryad_reduce(const void* col, int dtype, int64_t n, int op, void* out8, void* scratch,
                 size_t scratch_bytes, void* stream);

/* ---- K13 text tokenizer: LinqToDryad/DryadLinqTextReader.cs:36-280, samples/WordCount.cs.pp:72 */
size_t dryad_tokenize_scratch_bytes(int64_t nbytes);
int dryad_tokenize_count(const uint8_t* text, int64_t n, int delim, int keep_empty,
                         int64_t* count_dev, void* scratch, size_t scratch_bytes, void* stream);
int dryad_tokenize_write(const uint8_t* text, int64_t n, int delim, int keep_empty, int64_t* starts,
                         int32_t* lens, uint64_t* hashes, void* scratch, size_t scratch_bytes,
                         void* stream);
/* String columns as (start, length) views into a byte buffer: key hashes (same function as the
 * tokenizer's) and packing into zero-padded fixed-width slots for the fused shuffle
 * (LinqToDryad/DryadLinqBinaryWriter.cs:523-546, LinqToDryad/LineRecord.cs:30-176). */
int dryad_hash_views(const uint8_t* buf, const int64_t* starts, const int32_t* lens, int64_t n, uint64_t* out, void* stream);
int dryad_str_pack(const uint8_t* buf, const int64_t* starts, const int32_t* lens, int64_t n, int width, uint8_t* out, void* stream);
int dryad_tokenize_verify(const uint8_t* text, const int64_t* starts, const int32_t* lens,
                          const int64_t* rep_start, const int32_t* rep_len, int64_t ntok,
                          unsigned long long* mismatches, void* stream);

/* ---- K5 k-way merge / K9 merge join: LinqToDryad/DryadLinqVertex.cs:319-423, 898-1230, 9819-10140 */
int dryad_merge_pairs_u64(const uint64_t* ka, const uint32_t* va, int64_t na, const uint64_t* kb,
                          const uint32_t* vb, int64_t nb, uint64_t* ko, uint32_t* vo, void* stream);
int dryad_sorted_bounds(const uint64_t* sorted, int64_t m, const uint64_t* probe, int64_t n,
                        int64_t* lower, int64_t* upper, void* stream);
int dryad_emit_ranges(const int64_t* lower, const int64_t* upper, const int64_t* offsets, int64_t n,
                      int32_t* out_outer, int32_t* out_inner, void* stream);

/* ---- K19 + C3 k-means on tcgen05/TMEM/TMA (BASELINE config 5; the client loop is
 *      LinqToDryad/DryadLinqQueryable.cs:1280-1306 DoWhile) ------------------------------------
 * points [n,128] fp32 and centroids [k,128] fp32 (k multiple of 128, k <= 2048), 16-byte aligned.
 * dryad_kmeans_assign ACCUMULATES into sums[k,128], counts[k], inertia[1]: zero them per iteration. */
int dryad_kmeans_norms(const float* centroids, int k, float* cnorm, void* stream);
int dryad_kmeans_assign(const float* points, int64_t n, const float* centroids, const float* cnorm,
                        int k, int32_t* assign, float* sums, unsigned long long* counts,
                        float* inertia, int max_ctas, void* stream);
/* All-reduce of the partials over peer-mapped memory + centroid update (no NCCL). */
/* fp32-exact assignments.  dryad_kmeans_sep: sep2[b] = 1/4 of the squared distance from centroid b to its
 * nearest other centroid (recompute whenever the centroids change).  dryad_kmeans_assign_exact: the
 * tensor-core assignment, certified per point in the same kernel (exact fp32 distance to the chosen
 * centroid <= sep2 proves it is the fp32 argmin); points without a certificate are queued in
 * `uncertain` (capacity n, count in *uncertain_count, which the caller zeroes) and re-evaluated over
 * all centroids in fp32 FMA arithmetic; assignment, counts and sums are corrected where the choice
 * changes (*changed counts those, may be NULL).  n < 2^32. */
int dryad_kmeans_sep(const float* centroids, int k, float* sep2, void* stream);
int dryad_kmeans_assign_exact(const float* points, int64_t n, const float* centroids, const float* cnorm, const float* sep2,
                              int k, int32_t* assign, float* sums, unsigned long long* counts, float* inertia,
                              uint32_t* uncertain, unsigned int* uncertain_count, unsigned int* changed, int max_ctas,
                              void* stream);
int dryad_kmeans_update(float* const* peer_sums, unsigned long long* const* peer_counts, int world,
                        int k, float* centroids, float* cnorm, unsigned long long* total_counts,
                        void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DRYAD_B200_H */

// K12 (fused elementwise pipelines + order-preserving compaction), K11 (scalar aggregates),
// K13 (text tokenisation), K16 (counts / exclusive scan).
//
// Reference counterparts: Where/Select/SelectMany pipelines fused into one generated lambda per
// super-node (LinqToDryad/DryadLinqVertex.cs:74-198, DryadLinqCodeGen.cs:2084-2138); scalar
// aggregates (DryadLinqVertex.cs:1673-4704); UTF-8 line decoding + the WordCount sample's
// Split(' ') (LinqToDryad/DryadLinqTextReader.cs:36-280, samples/WordCount.cs.pp:72); Offsets
// exclusive scan (LinqToDryad/DryadLinqEnumerable.cs:1131).
//
// The planner lowers a fused stage's column expressions to a small register program; one kernel
// evaluates the whole program per row, so a Select∘Where∘Select chain reads each input column
// once and writes each output column once.
#include "common.cuh"
#include "../../include/dryad_b200.h"

namespace dryad { namespace stream {

struct Program {
  int ninstr, ncols_in, ncols_out, result_reg;
  DryadExprInstr ins[DRYAD_EXPR_MAX_INSTR];
  const void* in[DRYAD_EXPR_MAX_COLS];
  int in_dtype[DRYAD_EXPR_MAX_COLS];
  void* out[DRYAD_EXPR_MAX_COLS];
  int out_dtype[DRYAD_EXPR_MAX_COLS];
};

union Reg { int64_t i; double f; uint64_t u; };

__device__ __forceinline__ Reg load_col(const void* p, int dt, int64_t row) {
  Reg r;
  switch (dt) {
    case DRYAD_DT_I64: r.i = ((const int64_t*)p)[row]; break;
    case DRYAD_DT_F64: r.f = ((const double*)p)[row]; break;
    case DRYAD_DT_I32: r.i = ((const int32_t*)p)[row]; break;
    case DRYAD_DT_F32: r.f = ((const float*)p)[row]; break;
    case DRYAD_DT_U8:  r.i = ((const uint8_t*)p)[row]; break;
    case DRYAD_DT_I16: r.i = ((const int16_t*)p)[row]; break;
    case DRYAD_DT_U32: r.i = ((const uint32_t*)p)[row]; break;
    case DRYAD_DT_U16: r.i = ((const uint16_t*)p)[row]; break;
    default: r.i = 0;
  }
  return r;
}

__device__ __forceinline__ void store_col(void* p, int dt, int64_t row, Reg r) {
  switch (dt) {
    case DRYAD_DT_I64: ((int64_t*)p)[row] = r.i; break;
    case DRYAD_DT_F64: ((double*)p)[row] = r.f; break;
    case DRYAD_DT_I32: ((int32_t*)p)[row] = (int32_t)r.i; break;
    case DRYAD_DT_F32: ((float*)p)[row] = (float)r.f; break;
    case DRYAD_DT_U8:  ((uint8_t*)p)[row] = (uint8_t)r.i; break;
    case DRYAD_DT_I16: ((int16_t*)p)[row] = (int16_t)r.i; break;
    case DRYAD_DT_U32: ((uint32_t*)p)[row] = (uint32_t)r.i; break;
    case DRYAD_DT_U16: ((uint16_t*)p)[row] = (uint16_t)r.i; break;
  }
}

__device__ __forceinline__ void run_program(const Program& P, int64_t row, Reg* R) {
#pragma unroll 1
  for (int pc = 0; pc < P.ninstr; ++pc) {
    const DryadExprInstr I = P.ins[pc];
    Reg a = R[I.a & 15], b = R[I.b & 15], d;
    d.i = 0;
    switch (I.op) {
      case DRYAD_OP_LOAD: d = load_col(P.in[I.a], P.in_dtype[I.a], row); break;
      case DRYAD_OP_CONST: d.i = I.imm; break;
      case DRYAD_OP_ROWID: d.i = row + I.imm; break;
      case DRYAD_OP_ADD_I: d.i = a.i + b.i; break;
      case DRYAD_OP_SUB_I: d.i = a.i - b.i; break;
      case DRYAD_OP_MUL_I: d.i = a.i * b.i; break;
      // Python's // and %: floored quotient, remainder with the divisor's sign (same as the generated kernels)
      case DRYAD_OP_DIV_I: { if (!b.i) { d.i = 0; break; } const long long q = a.i / b.i, r = a.i % b.i;
                             d.i = (r != 0 && ((r < 0) != (b.i < 0))) ? q - 1 : q; break; }
      case DRYAD_OP_MOD_I: { if (!b.i) { d.i = 0; break; } const long long r = a.i % b.i;
                             d.i = (r != 0 && ((r < 0) != (b.i < 0))) ? r + b.i : r; break; }
      case DRYAD_OP_NEG_I: d.i = -a.i; break;
      case DRYAD_OP_ADD_F: d.f = a.f + b.f; break;
      case DRYAD_OP_SUB_F: d.f = a.f - b.f; break;
      case DRYAD_OP_MUL_F: d.f = a.f * b.f; break;
      case DRYAD_OP_DIV_F: d.f = a.f / b.f; break;
      case DRYAD_OP_NEG_F: d.f = -a.f; break;
      case DRYAD_OP_EQ_I: d.i = a.i == b.i; break;
      case DRYAD_OP_NE_I: d.i = a.i != b.i; break;
      case DRYAD_OP_LT_I: d.i = a.i < b.i; break;
      case DRYAD_OP_LE_I: d.i = a.i <= b.i; break;
      case DRYAD_OP_GT_I: d.i = a.i > b.i; break;
      case DRYAD_OP_GE_I: d.i = a.i >= b.i; break;
      case DRYAD_OP_EQ_F: d.i = a.f == b.f; break;
      case DRYAD_OP_NE_F: d.i = a.f != b.f; break;
      case DRYAD_OP_LT_F: d.i = a.f < b.f; break;
      case DRYAD_OP_LE_F: d.i = a.f <= b.f; break;
      case DRYAD_OP_GT_F: d.i = a.f > b.f; break;
      case DRYAD_OP_GE_F: d.i = a.f >= b.f; break;
      case DRYAD_OP_AND: d.i = (a.i != 0) & (b.i != 0); break;
      case DRYAD_OP_OR:  d.i = (a.i != 0) | (b.i != 0); break;
      case DRYAD_OP_NOT: d.i = a.i == 0; break;
      case DRYAD_OP_BAND: d.i = a.i & b.i; break;
      case DRYAD_OP_BOR:  d.i = a.i | b.i; break;
      case DRYAD_OP_BXOR: d.i = a.i ^ b.i; break;
      case DRYAD_OP_BNOT: d.i = ~a.i; break;
      case DRYAD_OP_SHL: d.i = a.i << (b.i & 63); break;
      case DRYAD_OP_SHR: d.i = a.i >> (b.i & 63); break;
      case DRYAD_OP_I2F: d.f = (double)a.i; break;
      case DRYAD_OP_F2I: d.i = (int64_t)a.f; break;
      case DRYAD_OP_SELECT: d = a.i ? b : R[I.imm & 15]; break;
      case DRYAD_OP_HASH: d.u = mix64(a.u); break;
      case DRYAD_OP_ABS_I: d.i = a.i < 0 ? -a.i : a.i; break;
      case DRYAD_OP_ABS_F: d.f = fabs(a.f); break;
      case DRYAD_OP_MIN_I: d.i = a.i < b.i ? a.i : b.i; break;
      case DRYAD_OP_MAX_I: d.i = a.i > b.i ? a.i : b.i; break;
      case DRYAD_OP_MIN_F: d.f = fmin(a.f, b.f); break;
      case DRYAD_OP_MAX_F: d.f = fmax(a.f, b.f); break;
      case DRYAD_OP_SQRT_F: d.f = sqrt(a.f); break;
      case DRYAD_OP_FLOOR_F: d.f = floor(a.f); break;
      case DRYAD_OP_MOV: d = a; break;
      case DRYAD_OP_STORE: store_col(P.out[I.b], P.out_dtype[I.b], row, a); continue;
    }
    R[I.dst & 15] = d;
  }
}

__global__ void __launch_bounds__(256) expr_map_kernel(const __grid_constant__ Program P, int64_t n,
                                                       int64_t row_base, uint8_t* flags) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  Reg R[16];
  for (; i < n; i += stride) {
    run_program(P, i, R);
    if (flags) flags[i] = R[P.result_reg & 15].i != 0;
  }
  (void)row_base;
}

// ---- order-preserving compaction of flagged rows -> index list -----------------------------
constexpr int CTILE = 2048;  // rows per block

__global__ void __launch_bounds__(256) flag_count_kernel(const uint8_t* __restrict__ flags,
                                                         int64_t n, uint32_t* __restrict__ block_counts) {
  int64_t base = (int64_t)blockIdx.x * CTILE;
  uint32_t c = 0;
  for (int j = threadIdx.x; j < CTILE; j += 256) {
    int64_t i = base + j;
    if (i < n && flags[i]) ++c;
  }
  c = warp_sum(c);
  __shared__ uint32_t s[8];
  if (lane_id() == 0) s[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
    for (int w = 0; w < 8; ++w) t += s[w];
    block_counts[blockIdx.x] = t;
  }
}

// Exclusive scan of n int64 values (join match counts -> output offsets), three small kernels:
// per-block scan (4096 values per block, in place) + block totals, single-block scan of the totals,
// add-back.  total[0] receives the grand total.
constexpr int SCAN_ITEMS = 4;
__global__ void __launch_bounds__(1024) scan_i64_block_kernel(int64_t* __restrict__ x, int64_t n, int64_t* __restrict__ bsum) {
  __shared__ int64_t s_warp[32];
  const int64_t base = (int64_t)blockIdx.x * 1024 * SCAN_ITEMS + (int64_t)threadIdx.x * SCAN_ITEMS;
  int64_t v[SCAN_ITEMS];
  int64_t sum = 0;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k) { v[k] = (base + k < n) ? x[base + k] : 0; sum += v[k]; }
  int64_t incl = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int64_t a = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane_id() >= (uint32_t)o) incl += a;
  }
  if (lane_id() == 31) s_warp[threadIdx.x >> 5] = incl;
  __syncthreads();
  if (threadIdx.x < 32) {
    int64_t w = s_warp[threadIdx.x], wi = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int64_t a = __shfl_up_sync(0xffffffffu, wi, o);
      if (lane_id() >= (uint32_t)o) wi += a;
    }
    s_warp[threadIdx.x] = wi - w;
    if (threadIdx.x == 31) bsum[blockIdx.x] = wi;
  }
  __syncthreads();
  int64_t run = s_warp[threadIdx.x >> 5] + incl - sum;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k) { if (base + k < n) x[base + k] = run; run += v[k]; }
}
__global__ void __launch_bounds__(1024) scan_i64_totals_kernel(int64_t* __restrict__ bsum, int64_t nb, int64_t* __restrict__ total) {
  __shared__ int64_t s_warp[32];
  __shared__ int64_t s_carry;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (int64_t base = 0; base < nb; base += 1024) {
    const int64_t i = base + threadIdx.x;
    const int64_t v = (i < nb) ? bsum[i] : 0;
    int64_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int64_t a = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane_id() >= (uint32_t)o) incl += a;
    }
    if (lane_id() == 31) s_warp[threadIdx.x >> 5] = incl;
    __syncthreads();
    if (threadIdx.x < 32) {
      int64_t w = s_warp[threadIdx.x], wi = w;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        int64_t a = __shfl_up_sync(0xffffffffu, wi, o);
        if (lane_id() >= (uint32_t)o) wi += a;
      }
      s_warp[threadIdx.x] = wi - w;
    }
    __syncthreads();
    const int64_t carry = s_carry;
    if (i < nb) bsum[i] = carry + s_warp[threadIdx.x >> 5] + incl - v;
    __syncthreads();
    if (threadIdx.x == 1023) s_carry = carry + s_warp[31] + incl;
    __syncthreads();
  }
  if (threadIdx.x == 0) total[0] = s_carry;
}
__global__ void __launch_bounds__(1024) scan_i64_add_kernel(int64_t* __restrict__ x, int64_t n, const int64_t* __restrict__ boff) {
  const int64_t base = (int64_t)blockIdx.x * 1024 * SCAN_ITEMS + (int64_t)threadIdx.x * SCAN_ITEMS;
  const int64_t add = boff[blockIdx.x];
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k) if (base + k < n) x[base + k] += add;
}

// Single-block exclusive scan of m uint32 values into int64 offsets (+ total at out[m]).
__global__ void __launch_bounds__(1024) scan_u32_kernel(const uint32_t* __restrict__ in, int64_t m,
                                                        int64_t* __restrict__ out) {
  __shared__ int64_t s_warp[32];
  __shared__ int64_t s_carry;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (int64_t base = 0; base < m; base += 1024) {
    int64_t i = base + threadIdx.x;
    int64_t v = (i < m) ? (int64_t)in[i] : 0;
    int64_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int64_t a = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane_id() >= (uint32_t)o) incl += a;
    }
    if (lane_id() == 31) s_warp[threadIdx.x >> 5] = incl;
    __syncthreads();
    if (threadIdx.x < 32) {
      int64_t w = s_warp[threadIdx.x];
      int64_t wi = w;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        int64_t a = __shfl_up_sync(0xffffffffu, wi, o);
        if (lane_id() >= (uint32_t)o) wi += a;
      }
      s_warp[threadIdx.x] = wi - w;
    }
    __syncthreads();
    int64_t carry = s_carry;
    int64_t excl = carry + s_warp[threadIdx.x >> 5] + incl - v;
    if (i < m) out[i] = excl;
    __syncthreads();
    if (threadIdx.x == 1023) s_carry = excl + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) out[m] = s_carry;
}

__global__ void __launch_bounds__(256) flag_write_kernel(const uint8_t* __restrict__ flags, int64_t n,
                                                         const int64_t* __restrict__ block_offsets,
                                                         uint32_t* __restrict__ out_idx) {
  __shared__ uint32_t s_w[8];
  __shared__ uint32_t s_running;
  int64_t base = (int64_t)blockIdx.x * CTILE;
  if (threadIdx.x == 0) s_running = 0;
  __syncthreads();
  const int warp = threadIdx.x >> 5;
  for (int j0 = 0; j0 < CTILE; j0 += 256) {
    int64_t i = base + j0 + threadIdx.x;
    bool f = (i < n) && flags[i];
    uint32_t m = __ballot_sync(0xffffffffu, f);
    if (lane_id() == 0) s_w[warp] = __popc(m);
    __syncthreads();
    uint32_t before = s_running;
    for (int w = 0; w < warp; ++w) before += s_w[w];
    if (f) out_idx[block_offsets[blockIdx.x] + before + __popc(m & lanemask_lt())] = (uint32_t)i;
    __syncthreads();
    if (threadIdx.x == 0) { uint32_t t = 0; for (int w = 0; w < 8; ++w) t += s_w[w]; s_running += t; }
    __syncthreads();
  }
}

// ---- K11 reductions (deterministic two-stage) -----------------------------------------------
template <typename T>
__device__ __forceinline__ T red_op(int op, T a, T b) {
  switch (op) {
    case DRYAD_RED_SUM: return a + b;
    case DRYAD_RED_MIN: return a < b ? a : b;
    default: return a > b ? a : b;
  }
}
template <typename T> __device__ __forceinline__ T red_identity(int op);
template <> __device__ __forceinline__ int64_t red_identity<int64_t>(int op) {
  return op == DRYAD_RED_SUM ? 0 : (op == DRYAD_RED_MIN ? INT64_MAX : INT64_MIN);
}
template <> __device__ __forceinline__ double red_identity<double>(int op) {
  return op == DRYAD_RED_SUM ? 0.0 : (op == DRYAD_RED_MIN ? __longlong_as_double(0x7ff0000000000000ll)
                                                          : __longlong_as_double(0xfff0000000000000ll));
}

template <typename T>
__global__ void __launch_bounds__(256) reduce_kernel(const void* __restrict__ col, int dtype,
                                                     int64_t n, int op, T* __restrict__ partials,
                                                     unsigned int* __restrict__ ticket,
                                                     T* __restrict__ out) {
  T acc = red_identity<T>(op);
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    Reg r = load_col(col, dtype, i);
    T v = sizeof(T) == 8 && (dtype == DRYAD_DT_F64 || dtype == DRYAD_DT_F32) ? (T)r.f : (T)r.i;
    acc = red_op<T>(op, acc, v);
  }
  __shared__ T s[256];
  __shared__ bool last;
  s[threadIdx.x] = acc;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) s[threadIdx.x] = red_op<T>(op, s[threadIdx.x], s[threadIdx.x + o]);
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    partials[blockIdx.x] = s[0];
    __threadfence();
    last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (last) {
    __threadfence();
    T a = red_identity<T>(op);
    for (int b = threadIdx.x; b < (int)gridDim.x; b += 256) a = red_op<T>(op, a, ((volatile T*)partials)[b]);
    s[threadIdx.x] = a;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
      if (threadIdx.x < o) s[threadIdx.x] = red_op<T>(op, s[threadIdx.x], s[threadIdx.x + o]);
      __syncthreads();
    }
    if (threadIdx.x == 0) { *out = s[0]; *ticket = 0; }
  }
}

// ---- K13 tokenizer --------------------------------------------------------------------------
// keep_empty = 1 reproduces String.Split(' ') per line (empty tokens between adjacent
// delimiters); keep_empty = 0 yields maximal non-delimiter runs.  '\n' always ends a token.
__device__ __forceinline__ bool is_delim(unsigned char c, unsigned char delim) {
  return c == delim || c == '\n';
}
__device__ __forceinline__ bool tok_start(const unsigned char* t, int64_t i, int64_t n,
                                          unsigned char delim, int keep_empty) {
  if (i >= n) return false;
  unsigned char c = t[i];
  if (keep_empty) {
    if (i == 0) return true;
    unsigned char p = t[i - 1];
    return p == delim || p == '\n';
  }
  if (is_delim(c, delim)) return false;
  return i == 0 || is_delim(t[i - 1], delim);
}

__global__ void __launch_bounds__(256) tok_count_kernel(const unsigned char* __restrict__ text,
                                                        int64_t n, unsigned char delim,
                                                        int keep_empty,
                                                        uint32_t* __restrict__ block_counts) {
  int64_t base = (int64_t)blockIdx.x * CTILE;
  uint32_t c = 0;
  for (int j = threadIdx.x; j < CTILE; j += 256) c += tok_start(text, base + j, n, delim, keep_empty);
  c = warp_sum(c);
  __shared__ uint32_t s[8];
  if (lane_id() == 0) s[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) { uint32_t t = 0; for (int w = 0; w < 8; ++w) t += s[w]; block_counts[blockIdx.x] = t; }
}

__global__ void __launch_bounds__(256) tok_write_kernel(const unsigned char* __restrict__ text,
                                                        int64_t n, unsigned char delim,
                                                        int keep_empty,
                                                        const int64_t* __restrict__ block_offsets,
                                                        int64_t* __restrict__ starts,
                                                        int32_t* __restrict__ lens,
                                                        uint64_t* __restrict__ hashes) {
  __shared__ uint32_t s_w[8];
  __shared__ uint32_t s_running;
  int64_t base = (int64_t)blockIdx.x * CTILE;
  if (threadIdx.x == 0) s_running = 0;
  __syncthreads();
  const int warp = threadIdx.x >> 5;
  for (int j0 = 0; j0 < CTILE; j0 += 256) {
    int64_t i = base + j0 + threadIdx.x;
    bool f = tok_start(text, i, n, delim, keep_empty);
    uint32_t m = __ballot_sync(0xffffffffu, f);
    if (lane_id() == 0) s_w[warp] = __popc(m);
    __syncthreads();
    uint32_t before = s_running;
    for (int w = 0; w < warp; ++w) before += s_w[w];
    if (f) {
      int64_t o = block_offsets[blockIdx.x] + before + __popc(m & lanemask_lt());
      int64_t e = i;
      uint64_t h = 0xcbf29ce484222325ULL;
      while (e < n && !is_delim(text[e], delim)) { h ^= text[e]; h *= 0x100000001b3ULL; ++e; }
      starts[o] = i;
      lens[o] = (int32_t)(e - i);
      if (hashes) hashes[o] = mix64(h);
    }
    __syncthreads();
    if (threadIdx.x == 0) { uint32_t t = 0; for (int w = 0; w < 8; ++w) t += s_w[w]; s_running += t; }
    __syncthreads();
  }
}

// ---- variable-length (string) columns ---------------------------------------------------------
// A string column inside the engine is a set of VIEWS into a byte buffer: (start, length) per row
// (reference: strings on the wire are length-prefixed UTF-8, LinqToDryad/DryadLinqBinaryWriter.cs:
// 523-546; LineRecord, LinqToDryad/LineRecord.cs:30-176).  Keys are the 64-bit hash of the bytes —
// FNV-1a then the 64-bit mixer, the same function the tokenizer and hash_bytes use — and every use
// of a hash as a key is followed by a byte-for-byte check against the group's representative.
__global__ void hash_views_kernel(const unsigned char* __restrict__ buf, const int64_t* __restrict__ starts,
                                  const int32_t* __restrict__ lens, int64_t n, uint64_t* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    const unsigned char* p = buf + starts[i];
    const int32_t l = lens[i];
    uint64_t h = 0xcbf29ce484222325ULL;
    for (int32_t j = 0; j < l; ++j) { h ^= p[j]; h *= 0x100000001b3ULL; }
    out[i] = mix64(h);
  }
}

// Strings -> fixed-width slots (zero padded) so they can ride the fixed-width fused shuffle:
// out[i*W .. i*W+W) = bytes of string i.  W is a multiple of 4; lens[i] <= W is the caller's check.
__global__ void str_pack_kernel(const unsigned char* __restrict__ buf, const int64_t* __restrict__ starts,
                                const int32_t* __restrict__ lens, int64_t n, int W, unsigned char* __restrict__ out) {
  const int words = W >> 2;
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; t < n * words; t += stride) {
    const int64_t i = t / words;
    const int w = (int)(t - i * words);
    const unsigned char* p = buf + starts[i] + 4 * w;
    const int rem = lens[i] - 4 * w;
    uint32_t v = 0;
    if (rem > 0) v |= (uint32_t)p[0];
    if (rem > 1) v |= (uint32_t)p[1] << 8;
    if (rem > 2) v |= (uint32_t)p[2] << 16;
    if (rem > 3) v |= (uint32_t)p[3] << 24;
    reinterpret_cast<uint32_t*>(out)[t] = v;
  }
}

// Verify that every token equals the representative token of its group (hash-collision check).
__global__ void tok_verify_kernel(const unsigned char* __restrict__ text,
                                  const int64_t* __restrict__ starts, const int32_t* __restrict__ lens,
                                  const int64_t* __restrict__ rep_start, const int32_t* __restrict__ rep_len,
                                  int64_t ntok, unsigned long long* __restrict__ mismatches) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < ntok; i += stride) {
    int32_t l = lens[i];
    bool bad = l != rep_len[i];
    const unsigned char* a = text + starts[i];
    const unsigned char* b = text + rep_start[i];
    for (int32_t k = 0; !bad && k < l; ++k) bad = a[k] != b[k];
    if (bad) atomicAdd(mismatches, 1ull);
  }
}

static inline int grid_for(int64_t n, int per_thread = 4) {
  int64_t g = (n + 256ll * per_thread - 1) / (256ll * per_thread);
  if (g < 1) g = 1;
  if (g > 148 * 16) g = 148 * 16;
  return (int)g;
}

}}  // namespace dryad::stream

extern "C" {
using namespace dryad::stream;

int dryad_expr_run(const DryadExprProgram* prog, int64_t n, uint8_t* flags_out, void* stream) {
  if (!prog || prog->ninstr < 0 || prog->ninstr > DRYAD_EXPR_MAX_INSTR ||
      prog->ncols_in > DRYAD_EXPR_MAX_COLS || prog->ncols_out > DRYAD_EXPR_MAX_COLS)
    return DRYAD_ERR_BAD_ARG;
  if (n <= 0) return DRYAD_OK;
  Program P;
  P.ninstr = prog->ninstr; P.ncols_in = prog->ncols_in; P.ncols_out = prog->ncols_out;
  P.result_reg = prog->result_reg;
  for (int i = 0; i < prog->ninstr; ++i) {
    P.ins[i] = prog->ins[i];
    if (P.ins[i].op == DRYAD_OP_LOAD && (P.ins[i].a < 0 || P.ins[i].a >= prog->ncols_in)) return DRYAD_ERR_BAD_ARG;
    if (P.ins[i].op == DRYAD_OP_STORE && (P.ins[i].b < 0 || P.ins[i].b >= prog->ncols_out)) return DRYAD_ERR_BAD_ARG;
  }
  for (int c = 0; c < DRYAD_EXPR_MAX_COLS; ++c) {
    P.in[c] = c < prog->ncols_in ? prog->in[c] : nullptr;
    P.in_dtype[c] = c < prog->ncols_in ? prog->in_dtype[c] : 0;
    P.out[c] = c < prog->ncols_out ? prog->out[c] : nullptr;
    P.out_dtype[c] = c < prog->ncols_out ? prog->out_dtype[c] : 0;
  }
  expr_map_kernel<<<grid_for(n, 2), 256, 0, (cudaStream_t)stream>>>(P, n, 0, flags_out);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

size_t dryad_compact_scratch_bytes(int64_t n) {
  int64_t blocks = (n + CTILE - 1) / CTILE + 1;
  return (size_t)(blocks * 4 + 256 + (blocks + 1) * 8 + 256);
}

// flags[n] -> out_idx (ascending row ids of set flags); *count_dev receives the total (int64).
int dryad_compact_flags(const uint8_t* flags, int64_t n, uint32_t* out_idx, int64_t* count_dev,
                        void* scratch, size_t scratch_bytes, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (n < 0 || n >= (1ll << 32)) return DRYAD_ERR_BAD_ARG;
  if (scratch_bytes < dryad_compact_scratch_bytes(n)) return DRYAD_ERR_SCRATCH;
  int64_t blocks = (n + CTILE - 1) / CTILE;
  if (blocks == 0) { DRYAD_CUDA_TRY(cudaMemsetAsync(count_dev, 0, 8, st)); return DRYAD_OK; }
  uint32_t* bc = (uint32_t*)scratch;
  int64_t* bo = (int64_t*)((char*)scratch + ((blocks * 4 + 255) & ~255ll));
  flag_count_kernel<<<(unsigned)blocks, 256, 0, st>>>(flags, n, bc);
  DRYAD_LAUNCH_CHECK();
  scan_u32_kernel<<<1, 1024, 0, st>>>(bc, blocks, bo);
  DRYAD_LAUNCH_CHECK();
  flag_write_kernel<<<(unsigned)blocks, 256, 0, st>>>(flags, n, bo, out_idx);
  DRYAD_LAUNCH_CHECK();
  DRYAD_CUDA_TRY(cudaMemcpyAsync(count_dev, bo + blocks, 8, cudaMemcpyDeviceToDevice, st));
  return DRYAD_OK;
}

int dryad_exclusive_scan_u32(const uint32_t* in, int64_t m, int64_t* out_m_plus_1, void* stream) {
  if (m < 0) return DRYAD_ERR_BAD_ARG;
  scan_u32_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(in, m, out_m_plus_1);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

size_t dryad_scan_i64_scratch_bytes(int64_t n) {
  int64_t nb = (n + 1024 * SCAN_ITEMS - 1) / (1024 * SCAN_ITEMS);
  return (size_t)(nb < 1 ? 1 : nb) * 8 + 256;
}

// In place: x[i] <- sum of x[0..i); total_dev[0] <- sum of all.
int dryad_exclusive_scan_i64(int64_t* x, int64_t n, int64_t* total_dev, void* scratch, size_t scratch_bytes, void* stream) {
  if (n < 0) return DRYAD_ERR_BAD_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  if (scratch_bytes < dryad_scan_i64_scratch_bytes(n)) return DRYAD_ERR_SCRATCH;
  if (n == 0) { DRYAD_CUDA_TRY(cudaMemsetAsync(total_dev, 0, 8, st)); return DRYAD_OK; }
  const int64_t nb = (n + 1024 * SCAN_ITEMS - 1) / (1024 * SCAN_ITEMS);
  int64_t* bsum = (int64_t*)scratch;
  scan_i64_block_kernel<<<(unsigned)nb, 1024, 0, st>>>(x, n, bsum);
  DRYAD_LAUNCH_CHECK();
  scan_i64_totals_kernel<<<1, 1024, 0, st>>>(bsum, nb, total_dev);
  DRYAD_LAUNCH_CHECK();
  scan_i64_add_kernel<<<(unsigned)nb, 1024, 0, st>>>(x, n, bsum);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

size_t dryad_reduce_scratch_bytes(void) { return 148 * 8 * 8 + 256; }

// out (8 bytes): int64 for integer dtypes, double for float dtypes.
int dryad_reduce(const void* col, int dtype, int64_t n, int op, void* out8, void* scratch,
                 size_t scratch_bytes, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (scratch_bytes < dryad_reduce_scratch_bytes()) return DRYAD_ERR_SCRATCH;
  if (op < DRYAD_RED_SUM || op > DRYAD_RED_MAX) return DRYAD_ERR_BAD_ARG;
  unsigned int* ticket = (unsigned int*)scratch;
  void* partials = (char*)scratch + 256;
  int grid = n > 0 ? grid_for(n, 8) : 1;
  if (grid > 148 * 8) grid = 148 * 8;
  bool is_f = dtype == DRYAD_DT_F64 || dtype == DRYAD_DT_F32;
  if (is_f) reduce_kernel<double><<<grid, 256, 0, st>>>(col, dtype, n, op, (double*)partials, ticket, (double*)out8);
  else reduce_kernel<int64_t><<<grid, 256, 0, st>>>(col, dtype, n, op, (int64_t*)partials, ticket, (int64_t*)out8);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

size_t dryad_tokenize_scratch_bytes(int64_t nbytes) { return dryad_compact_scratch_bytes(nbytes); }

// Pass 1: count tokens (count_dev int64).  The caller sizes the outputs, then calls pass 2.
int dryad_tokenize_count(const uint8_t* text, int64_t n, int delim, int keep_empty,
                         int64_t* count_dev, void* scratch, size_t scratch_bytes, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (n < 0) return DRYAD_ERR_BAD_ARG;
  if (scratch_bytes < dryad_compact_scratch_bytes(n)) return DRYAD_ERR_SCRATCH;
  int64_t blocks = (n + CTILE - 1) / CTILE;
  if (blocks == 0) { DRYAD_CUDA_TRY(cudaMemsetAsync(count_dev, 0, 8, st)); return DRYAD_OK; }
  uint32_t* bc = (uint32_t*)scratch;
  int64_t* bo = (int64_t*)((char*)scratch + ((blocks * 4 + 255) & ~255ll));
  tok_count_kernel<<<(unsigned)blocks, 256, 0, st>>>(text, n, (unsigned char)delim, keep_empty, bc);
  DRYAD_LAUNCH_CHECK();
  scan_u32_kernel<<<1, 1024, 0, st>>>(bc, blocks, bo);
  DRYAD_LAUNCH_CHECK();
  DRYAD_CUDA_TRY(cudaMemcpyAsync(count_dev, bo + blocks, 8, cudaMemcpyDeviceToDevice, st));
  return DRYAD_OK;
}

// Pass 2 (same scratch, untouched since pass 1): token start offsets, lengths, 64-bit hashes.
int dryad_tokenize_write(const uint8_t* text, int64_t n, int delim, int keep_empty, int64_t* starts,
                         int32_t* lens, uint64_t* hashes, void* scratch, size_t scratch_bytes,
                         void* stream) {
  if (n <= 0) return DRYAD_OK;
  if (scratch_bytes < dryad_compact_scratch_bytes(n)) return DRYAD_ERR_SCRATCH;
  int64_t blocks = (n + CTILE - 1) / CTILE;
  int64_t* bo = (int64_t*)((char*)scratch + ((blocks * 4 + 255) & ~255ll));
  tok_write_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(text, n, (unsigned char)delim, keep_empty, bo, starts, lens, hashes);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_hash_views(const uint8_t* buf, const int64_t* starts, const int32_t* lens, int64_t n, uint64_t* out, void* stream) {
  if (n <= 0) return DRYAD_OK;
  hash_views_kernel<<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(buf, starts, lens, n, out);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_str_pack(const uint8_t* buf, const int64_t* starts, const int32_t* lens, int64_t n, int width, uint8_t* out, void* stream) {
  if (width <= 0 || (width & 3)) return DRYAD_ERR_BAD_ARG;
  if (n <= 0) return DRYAD_OK;
  str_pack_kernel<<<grid_for(n * (width >> 2)), 256, 0, (cudaStream_t)stream>>>(buf, starts, lens, n, width, out);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_tokenize_verify(const uint8_t* text, const int64_t* starts, const int32_t* lens,
                          const int64_t* rep_start, const int32_t* rep_len, int64_t ntok,
                          unsigned long long* mismatches, void* stream) {
  if (ntok <= 0) return DRYAD_OK;
  tok_verify_kernel<<<grid_for(ntok), 256, 0, (cudaStream_t)stream>>>(text, starts, lens, rep_start, rep_len, ntok, mismatches);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

}  // extern "C"

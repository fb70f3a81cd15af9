// K7 (sort-based grouping: run boundaries of a sorted key column) and K15 (row <-> column
// transcoding of fixed-stride binary records).
//
// Reference counterparts: ParallelSortGroupBy / OrderedGroupBy detect key runs while iterating a
// sorted partition (LinqToDryad/DryadLinqVertex.cs:6418-6700, 586-760); the binary record format is
// row-wise little-endian primitives (LinqToDryad/DryadLinqBinaryWriter.cs:203-546,
// DryadLinqBinaryReader.cs).  Here a grouped partition is CSR: (sort permutation, run starts, run
// keys); records of a numeric schema become one tensor per column.
#include "common.cuh"
#include "../../include/dryad_b200.h"

namespace dryad { namespace grp {

__global__ void __launch_bounds__(256) mark_boundaries(const uint64_t* __restrict__ keys, int64_t n,
                                                       uint8_t* __restrict__ flags) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) flags[i] = (i == 0) || (keys[i] != keys[i - 1]);
}

struct Cols {
  int ncols;
  int off[DRYAD_TRANSCODE_MAX_COLS];
  int width[DRYAD_TRANSCODE_MAX_COLS];     // 1, 2, 4 or 8 bytes
  unsigned char* col[DRYAD_TRANSCODE_MAX_COLS];
};

template <int W> struct UInt;
template <> struct UInt<1> { using T = uint8_t; };
template <> struct UInt<2> { using T = uint16_t; };
template <> struct UInt<4> { using T = uint32_t; };
template <> struct UInt<8> { using T = uint64_t; };

// unaligned little-endian field access: records are packed, fields need not be aligned
template <int W>
__device__ __forceinline__ typename UInt<W>::T load_le(const unsigned char* p) {
  typename UInt<W>::T v = 0;
#pragma unroll
  for (int b = 0; b < W; ++b) v |= (typename UInt<W>::T)p[b] << (8 * b);
  return v;
}
template <int W>
__device__ __forceinline__ void store_le(unsigned char* p, typename UInt<W>::T v) {
#pragma unroll
  for (int b = 0; b < W; ++b) p[b] = (unsigned char)(v >> (8 * b));
}

// One warp stages 32 consecutive records through shared memory so that both the row side (32 x
// stride contiguous bytes) and the column side (32 consecutive elements) are coalesced.
template <bool TO_COLUMNS>
__global__ void __launch_bounds__(256) transcode(unsigned char* __restrict__ rows, int64_t n, int stride, Cols C) {
  extern __shared__ unsigned char smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned char* tile = smem + (size_t)warp * 32 * stride;
  const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
  for (int64_t base = ((int64_t)blockIdx.x * (blockDim.x >> 5) + warp) * 32; base < n; base += nwarps * 32) {
    const int rows_here = (int)min((int64_t)32, n - base);
    const int bytes = rows_here * stride;
    if (TO_COLUMNS) {
      for (int b = lane; b < bytes; b += 32) tile[b] = rows[base * stride + b];
      __syncwarp();
      if (lane < rows_here) {
        const unsigned char* r = tile + lane * stride;
        for (int c = 0; c < C.ncols; ++c) {
          unsigned char* dst = C.col[c] + (base + lane) * C.width[c];
          switch (C.width[c]) {
            case 1: *dst = r[C.off[c]]; break;
            case 2: *reinterpret_cast<uint16_t*>(dst) = load_le<2>(r + C.off[c]); break;
            case 4: *reinterpret_cast<uint32_t*>(dst) = load_le<4>(r + C.off[c]); break;
            default: *reinterpret_cast<uint64_t*>(dst) = load_le<8>(r + C.off[c]); break;
          }
        }
      }
      __syncwarp();
    } else {
      for (int b = lane; b < bytes; b += 32) tile[b] = 0;          // padding bytes between fields are zero
      __syncwarp();
      if (lane < rows_here) {
        unsigned char* r = tile + lane * stride;
        for (int c = 0; c < C.ncols; ++c) {
          const unsigned char* src = C.col[c] + (base + lane) * C.width[c];
          switch (C.width[c]) {
            case 1: r[C.off[c]] = *src; break;
            case 2: store_le<2>(r + C.off[c], *reinterpret_cast<const uint16_t*>(src)); break;
            case 4: store_le<4>(r + C.off[c], *reinterpret_cast<const uint32_t*>(src)); break;
            default: store_le<8>(r + C.off[c], *reinterpret_cast<const uint64_t*>(src)); break;
          }
        }
      }
      __syncwarp();
      for (int b = lane; b < bytes; b += 32) rows[base * stride + b] = tile[b];
      __syncwarp();
    }
  }
}

static int fill(Cols& C, int stride, int ncols, const int* offs, const int* widths, void* const* cols) {
  if (ncols < 1 || ncols > DRYAD_TRANSCODE_MAX_COLS || stride < 1 || stride > 512) return DRYAD_ERR_BAD_ARG;
  C.ncols = ncols;
  for (int c = 0; c < ncols; ++c) {
    const int w = widths[c];
    if (!(w == 1 || w == 2 || w == 4 || w == 8) || offs[c] < 0 || offs[c] + w > stride || !cols[c]) return DRYAD_ERR_BAD_ARG;
    C.off[c] = offs[c]; C.width[c] = w; C.col[c] = (unsigned char*)cols[c];
  }
  return DRYAD_OK;
}

}}  // namespace dryad::grp

extern "C" {
using namespace dryad::grp;

int dryad_mark_boundaries(const uint64_t* sorted_keys, int64_t n, uint8_t* flags, void* stream) {
  if (n <= 0) return DRYAD_OK;
  int64_t g = (n + 1023) / 1024; if (g > 148 * 16) g = 148 * 16;
  mark_boundaries<<<(int)g, 256, 0, (cudaStream_t)stream>>>(sorted_keys, n, flags);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_rows_to_columns(const uint8_t* rows, int64_t n, int stride, int ncols, const int* offs,
                          const int* widths, void* const* cols, void* stream) {
  Cols C; int rc = fill(C, stride, ncols, offs, widths, cols); if (rc) return rc;
  if (n <= 0) return DRYAD_OK;
  int64_t g = (n + 255) / 256; if (g > 148 * 8) g = 148 * 8;
  if ((size_t)8 * 32 * stride > 48 * 1024)
    DRYAD_CUDA_TRY(cudaFuncSetAttribute(transcode<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 32 * stride));
  transcode<true><<<(int)g, 256, (size_t)8 * 32 * stride, (cudaStream_t)stream>>>(const_cast<uint8_t*>(rows), n, stride, C);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_columns_to_rows(uint8_t* rows, int64_t n, int stride, int ncols, const int* offs,
                          const int* widths, const void* const* cols, void* stream) {
  Cols C; int rc = fill(C, stride, ncols, offs, widths, const_cast<void* const*>(cols)); if (rc) return rc;
  if (n <= 0) return DRYAD_OK;
  int64_t g = (n + 255) / 256; if (g > 148 * 8) g = 148 * 8;
  if ((size_t)8 * 32 * stride > 48 * 1024)
    DRYAD_CUDA_TRY(cudaFuncSetAttribute(transcode<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 32 * stride));
  transcode<false><<<(int)g, 256, (size_t)8 * 32 * stride, (cudaStream_t)stream>>>(rows, n, stride, C);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

}  // extern "C"

// N5: ring-buffer FIFO in HBM with flag-based flow control — a streamed channel between two GPUs
// that needs no host involvement per block.
//
// Reference: the bounded in-memory FIFO between pipelined vertices (DryadVertex/VertexHost/system/
// channel/src/channelfifo.cpp, .h:25-241: writers block when `depth` buffers are outstanding,
// readers block on empty, an end-of-stream marker closes it).  Here the ring lives in the CONSUMER's
// HBM; the producer's push kernel stores the block into a free slot over NVLink and then publishes
// it with a release store of the running block count (`head`, in the consumer's memory); the
// consumer's pop kernel acquires `head`, copies the slot out and hands the slot back by a release
// store of its own running count (`tail`, in the PRODUCER's memory).  Every wait is a bounded spin
// on LOCAL memory (the remote side writes the flag), so a dead peer is an error, not a hang.
//
// One push / pop = one kernel launch on the caller's stream.  Producer and consumer must be
// different GPUs (two kernels that wait on each other must never share a device).
//
// A timeout is FATAL for the ring: the status word becomes non-zero, and because the completion
// counters count CTAs per block, a push or pop that timed out leaves them out of step for good.
// Every later push / pop on a ring whose status is non-zero returns at once without touching the
// ring; the owner (parallel/ring.py HbmRing) refuses further use once it has seen the status.
#include "common.cuh"
#include "../../include/dryad_b200.h"

namespace dryad { namespace ring {

constexpr long long SPIN_CYCLES = 4000000000ll;       // ~2 s

__device__ __forceinline__ bool wait_ge(const uint64_t* flag, uint64_t want) {
  const long long t0 = clock64();
  while (ld_acquire_sys(flag) < want) {
    if (clock64() - t0 > SPIN_CYCLES) return false;
    __nanosleep(64);
  }
  return true;
}

// seq: number of blocks pushed before this one.  tail_local: blocks the consumer has released
// (written remotely by the consumer).  The last CTA to finish publishes.
__global__ void __launch_bounds__(256) push_kernel(const uint4* __restrict__ src, uint64_t nbytes,
                                                   unsigned char* ring_peer, uint64_t* len_peer, uint64_t* head_peer,
                                                   const uint64_t* tail_local, unsigned long long* done_local,
                                                   uint64_t seq, int slots, uint64_t slot_bytes, int32_t* status) {
  __shared__ int ok;
  if (threadIdx.x == 0)
    ok = (*reinterpret_cast<volatile int32_t*>(status) == 0) &&
         wait_ge(tail_local, seq + 1 >= (uint64_t)slots ? seq + 1 - slots : 0) ? 1 : 0;
  __syncthreads();
  if (!ok) { if (threadIdx.x == 0) atomicMax(status, 2); return; }
  const int slot = (int)(seq % slots);
  uint4* dst = reinterpret_cast<uint4*>(ring_peer + (size_t)slot * slot_bytes);
  const uint64_t n16 = nbytes / 16;                    // whole 16-byte units; the tail goes bytewise:
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (uint64_t)gridDim.x * blockDim.x)
    dst[i] = __ldcs(src + i);
  if (blockIdx.x == 0) {                                // the source is never read past nbytes
    const unsigned char* sb = reinterpret_cast<const unsigned char*>(src);
    unsigned char* db = reinterpret_cast<unsigned char*>(dst);
    for (uint64_t b = n16 * 16 + threadIdx.x; b < nbytes; b += blockDim.x) db[b] = sb[b];
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned long long prev = atomicAdd(done_local, 1ull);
    if (prev + 1 == (unsigned long long)gridDim.x * (seq + 1)) {        // every CTA of this push has stored its share
      __threadfence_system();
      len_peer[slot] = nbytes;
      st_release_sys(head_peer, seq + 1);
    }
  }
}

__global__ void __launch_bounds__(256) pop_kernel(uint4* __restrict__ dst, uint64_t dst_cap, uint64_t* out_len,
                                                  const unsigned char* ring_local, const uint64_t* len_local,
                                                  const uint64_t* head_local, uint64_t* tail_peer,
                                                  unsigned long long* done_local, uint64_t seq, int slots,
                                                  uint64_t slot_bytes, int32_t* status) {
  __shared__ int ok;
  if (threadIdx.x == 0) ok = (*reinterpret_cast<volatile int32_t*>(status) == 0) && wait_ge(head_local, seq + 1) ? 1 : 0;
  __syncthreads();
  if (!ok) { if (threadIdx.x == 0) atomicMax(status, 2); return; }
  const int slot = (int)(seq % slots);
  uint64_t nbytes = *reinterpret_cast<const volatile uint64_t*>(len_local + slot);
  if (nbytes > dst_cap) { if (threadIdx.x == 0 && blockIdx.x == 0) atomicMax(status, 1); nbytes = dst_cap; }
  const uint4* src = reinterpret_cast<const uint4*>(ring_local + (size_t)slot * slot_bytes);
  const uint64_t n16 = (nbytes + 15) / 16;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (uint64_t)gridDim.x * blockDim.x)
    dst[i] = src[i];
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned long long prev = atomicAdd(done_local, 1ull);
    if (prev + 1 == (unsigned long long)gridDim.x * (seq + 1)) {
      if (out_len) *out_len = nbytes;
      __threadfence_system();
      st_release_sys(tail_peer, seq + 1);                                // the slot may be overwritten
    }
  }
}

}}  // namespace dryad::ring

extern "C" {
using namespace dryad::ring;

// Grid size is part of the protocol (the completion counter counts CTAs): fixed at 64.
int dryad_ring_push(const void* src, uint64_t nbytes, void* ring_peer, uint64_t* len_peer, uint64_t* head_peer,
                    const uint64_t* tail_local, uint64_t* done_local, uint64_t seq, int slots, uint64_t slot_bytes,
                    int32_t* status, void* stream) {
  if (slots < 1 || nbytes > slot_bytes || (slot_bytes & 15) || ((uintptr_t)src & 15) || ((uintptr_t)ring_peer & 15)) return DRYAD_ERR_BAD_ARG;
  push_kernel<<<64, 256, 0, (cudaStream_t)stream>>>((const uint4*)src, nbytes, (unsigned char*)ring_peer, len_peer, head_peer,
                                                    tail_local, (unsigned long long*)done_local, seq, slots, slot_bytes, status);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

int dryad_ring_pop(void* dst, uint64_t dst_cap, uint64_t* out_len, const void* ring_local, const uint64_t* len_local,
                   const uint64_t* head_local, uint64_t* tail_peer, uint64_t* done_local, uint64_t seq, int slots,
                   uint64_t slot_bytes, int32_t* status, void* stream) {
  if (slots < 1 || (slot_bytes & 15) || ((uintptr_t)dst & 15) || (dst_cap & 15)) return DRYAD_ERR_BAD_ARG;
  pop_kernel<<<64, 256, 0, (cudaStream_t)stream>>>((uint4*)dst, dst_cap, out_len, (const unsigned char*)ring_local, len_local,
                                                   head_local, tail_peer, (unsigned long long*)done_local, seq, slots, slot_bytes, status);
  DRYAD_LAUNCH_CHECK();
  return DRYAD_OK;
}

}  // extern "C"

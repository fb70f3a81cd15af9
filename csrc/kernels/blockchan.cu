// Block channel over device partitions: the C ABI a vertex uses to exchange data blocks with the
// engine — allocate / write / read / release, end of stream, total length.
//
// Reference: DryadVertex/VertexHost/vertex/include/wrappernativeinfo.h:103-150 (ReadDataBlock,
// WriteDataBlock, AllocateDataBlock, ReleaseDataBlock, Flush, Close, GetTotalLength — exported to
// the managed vertex runtime through DryadLINQNativeChannels.def:3-19, bound in
// LinqToDryad/DryadLinqNative.cs:76-145) and the bounded FIFO between vertices
// (DryadVertex/VertexHost/system/channel/src/channelfifo.cpp: writers block when `depth` buffers
// are outstanding — throttling).
//
// Here a block is a slice of a caller-provided device buffer (for instance a piece of the symmetric
// arena, so the blocks are peer-addressable): the library allocates nothing.  The pool is the buffer
// cut into equal blocks; AllocateDataBlock waits for a free block (that is the throttle: a writer can
// be at most `pool size` blocks ahead of the reader).  WriteDataBlock publishes a block together with
// a CUDA event recorded on the writer's stream; ReadDataBlock hands the next block to the reader and
// makes the reader's stream wait for that event, so producer and consumer kernels are ordered without
// a host synchronisation.  Optional integrity check (the role of the reference's channel
// fingerprints, classlib Rabin fingerprint): a checksum kernel runs on the writer's stream at
// publication and again on the reader's stream at delivery; mismatches are counted on the device.
#include "common.cuh"
#include "../../include/dryad_b200.h"
#include <chrono>
#include <condition_variable>
#include <deque>
#include <mutex>
#include <vector>

namespace dryad { namespace bc {

__global__ void __launch_bounds__(256) checksum_kernel(const unsigned char* __restrict__ p, uint64_t nbytes,
                                                       unsigned long long* __restrict__ out) {
  // position-dependent 64-bit checksum: sum over 8-byte words of mix64(word ^ index) (+ tail bytes)
  const uint64_t nw = nbytes / 8;
  const uint64_t* w = reinterpret_cast<const uint64_t*>(p);
  unsigned long long acc = 0;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nw; i += (uint64_t)gridDim.x * blockDim.x)
    acc += mix64(w[i] ^ (i * 0x9E3779B97F4A7C15ULL));
  if (blockIdx.x == 0 && threadIdx.x == 0)
    for (uint64_t b = nw * 8; b < nbytes; ++b) acc += mix64((uint64_t)p[b] + (b << 8));
  acc = warp_sum(acc);
  if (lane_id() == 0 && acc) atomicAdd(out, acc);
}

__global__ void compare_kernel(const unsigned long long* a, const unsigned long long* b, unsigned long long* errors) {
  if (*a != *b) atomicAdd(errors, 1ull);
}

struct Published { int slot; uint64_t nbytes; };

struct Channel {
  unsigned char* base = nullptr;
  size_t block_bytes = 0;
  int nblocks = 0;
  bool checksum = false;
  unsigned long long* sums = nullptr;      // device: [2 * nblocks + 1] = written sum, read sum per block, error counter
  std::mutex mu;
  std::condition_variable cv;
  std::vector<int> free_slots;
  std::deque<Published> ready;
  std::vector<cudaEvent_t> events;
  bool closed = false;
  int64_t blocks_written = 0, blocks_read = 0, bytes_written = 0, throttle_waits = 0;
};

}}  // namespace dryad::bc

extern "C" {
using namespace dryad::bc;

size_t dryad_chan_control_bytes(int nblocks) { return (size_t)(2 * nblocks + 1) * 8; }

// buffer: nblocks * block_bytes bytes of device memory (16-byte aligned); control: device memory of
// dryad_chan_control_bytes(nblocks) bytes (used only when checksum != 0, may be NULL otherwise).
int dryad_chan_create(void* buffer, size_t block_bytes, int nblocks, int checksum, void* control, void** out) {
  if (!buffer || !out || nblocks < 1 || block_bytes == 0 || (block_bytes & 15) || ((uintptr_t)buffer & 15)) return DRYAD_ERR_BAD_ARG;
  if (checksum && !control) return DRYAD_ERR_BAD_ARG;
  Channel* c = new Channel();
  c->base = (unsigned char*)buffer; c->block_bytes = block_bytes; c->nblocks = nblocks; c->checksum = checksum != 0;
  c->sums = (unsigned long long*)control;
  for (int i = nblocks - 1; i >= 0; --i) c->free_slots.push_back(i);
  c->events.resize(nblocks);
  for (int i = 0; i < nblocks; ++i) {
    cudaError_t e = cudaEventCreateWithFlags(&c->events[i], cudaEventDisableTiming);
    if (e != cudaSuccess) { for (int j = 0; j < i; ++j) cudaEventDestroy(c->events[j]); delete c; return (int)e; }
  }
  if (c->checksum) { cudaError_t e = cudaMemset(control, 0, dryad_chan_control_bytes(nblocks)); if (e != cudaSuccess) { delete c; return (int)e; } }
  *out = c;
  return DRYAD_OK;
}

int dryad_chan_destroy(void* h) {
  Channel* c = (Channel*)h;
  if (!c) return DRYAD_OK;
  for (auto& e : c->events) cudaEventDestroy(e);
  delete c;
  return DRYAD_OK;
}

// AllocateDataBlock: a free block of the pool; waits up to timeout_s when the reader is `nblocks` behind.
int dryad_chan_alloc_block(void* h, size_t bytes, double timeout_s, void** dev_ptr) {
  Channel* c = (Channel*)h;
  if (!c || !dev_ptr) return DRYAD_ERR_BAD_ARG;
  if (bytes > c->block_bytes) return DRYAD_ERR_TOO_LARGE;
  std::unique_lock<std::mutex> l(c->mu);
  if (c->closed) return DRYAD_ERR_BAD_ARG;
  if (c->free_slots.empty()) {
    ++c->throttle_waits;
    if (!c->cv.wait_for(l, std::chrono::duration<double>(timeout_s), [&] { return !c->free_slots.empty(); })) return DRYAD_ERR_OVERFLOW;
  }
  const int s = c->free_slots.back();
  c->free_slots.pop_back();
  *dev_ptr = c->base + (size_t)s * c->block_bytes;
  return DRYAD_OK;
}

// WriteDataBlock: publish `nbytes` of a block obtained from alloc_block; `stream` is the stream the
// block's contents were produced on.
int dryad_chan_write_block(void* h, void* dev_ptr, size_t nbytes, void* stream) {
  Channel* c = (Channel*)h;
  if (!c || !dev_ptr) return DRYAD_ERR_BAD_ARG;
  const size_t off = (unsigned char*)dev_ptr - c->base;
  if (off % c->block_bytes || off / c->block_bytes >= (size_t)c->nblocks || nbytes > c->block_bytes) return DRYAD_ERR_BAD_ARG;
  const int s = (int)(off / c->block_bytes);
  cudaStream_t st = (cudaStream_t)stream;
  if (c->checksum) {
    DRYAD_CUDA_TRY(cudaMemsetAsync(c->sums + 2 * s, 0, 8, st));
    if (nbytes) { checksum_kernel<<<32, 256, 0, st>>>((const unsigned char*)dev_ptr, nbytes, c->sums + 2 * s); DRYAD_LAUNCH_CHECK(); }
  }
  DRYAD_CUDA_TRY(cudaEventRecord(c->events[s], st));
  {
    std::lock_guard<std::mutex> l(c->mu);
    c->ready.push_back({s, (uint64_t)nbytes});
    ++c->blocks_written;
    c->bytes_written += (int64_t)nbytes;
  }
  c->cv.notify_all();
  return DRYAD_OK;
}

// Close (writer): end of stream; readers drain what was published, then see end of stream.
int dryad_chan_close_writer(void* h) {
  Channel* c = (Channel*)h;
  if (!c) return DRYAD_ERR_BAD_ARG;
  { std::lock_guard<std::mutex> l(c->mu); c->closed = true; }
  c->cv.notify_all();
  return DRYAD_OK;
}

// ReadDataBlock: returns 0 with the next block, 1 at end of stream, DRYAD_ERR_OVERFLOW on timeout.
// `stream` (the reader's) is made to wait for the producer's work on that block.
int dryad_chan_read_block(void* h, void** dev_ptr, size_t* nbytes, double timeout_s, void* stream) {
  Channel* c = (Channel*)h;
  if (!c || !dev_ptr || !nbytes) return DRYAD_ERR_BAD_ARG;
  Published p;
  {
    std::unique_lock<std::mutex> l(c->mu);
    if (!c->cv.wait_for(l, std::chrono::duration<double>(timeout_s), [&] { return !c->ready.empty() || c->closed; })) return DRYAD_ERR_OVERFLOW;
    if (c->ready.empty()) return 1;
    p = c->ready.front();
    c->ready.pop_front();
    ++c->blocks_read;
  }
  cudaStream_t st = (cudaStream_t)stream;
  DRYAD_CUDA_TRY(cudaStreamWaitEvent(st, c->events[p.slot], 0));
  unsigned char* ptr = c->base + (size_t)p.slot * c->block_bytes;
  if (c->checksum) {
    DRYAD_CUDA_TRY(cudaMemsetAsync(c->sums + 2 * p.slot + 1, 0, 8, st));
    if (p.nbytes) { checksum_kernel<<<32, 256, 0, st>>>(ptr, p.nbytes, c->sums + 2 * p.slot + 1); DRYAD_LAUNCH_CHECK(); }
    compare_kernel<<<1, 1, 0, st>>>(c->sums + 2 * p.slot, c->sums + 2 * p.slot + 1, c->sums + 2 * c->nblocks);
    DRYAD_LAUNCH_CHECK();
  }
  *dev_ptr = ptr;
  *nbytes = (size_t)p.nbytes;
  return DRYAD_OK;
}

// ReleaseDataBlock: the reader is done with the block (after the work it queued on `stream`): the
// block returns to the pool once that work has finished, so a writer cannot overwrite it early.
int dryad_chan_release_block(void* h, void* dev_ptr, void* stream) {
  Channel* c = (Channel*)h;
  if (!c || !dev_ptr) return DRYAD_ERR_BAD_ARG;
  const size_t off = (unsigned char*)dev_ptr - c->base;
  if (off % c->block_bytes || off / c->block_bytes >= (size_t)c->nblocks) return DRYAD_ERR_BAD_ARG;
  DRYAD_CUDA_TRY(cudaStreamSynchronize((cudaStream_t)stream));
  { std::lock_guard<std::mutex> l(c->mu); c->free_slots.push_back((int)(off / c->block_bytes)); }
  c->cv.notify_all();
  return DRYAD_OK;
}

int64_t dryad_chan_total_length(void* h) {
  Channel* c = (Channel*)h;
  if (!c) return -1;
  std::lock_guard<std::mutex> l(c->mu);
  return c->bytes_written;
}

// out5: blocks written, blocks read, bytes written, times a writer waited for a free block,
// checksum mismatches seen so far (reads the device counter: synchronises the device when checksum is on).
int dryad_chan_stats(void* h, int64_t* out5) {
  Channel* c = (Channel*)h;
  if (!c || !out5) return DRYAD_ERR_BAD_ARG;
  unsigned long long errs = 0;
  if (c->checksum) DRYAD_CUDA_TRY(cudaMemcpy(&errs, c->sums + 2 * c->nblocks, 8, cudaMemcpyDeviceToHost));
  std::lock_guard<std::mutex> l(c->mu);
  out5[0] = c->blocks_written; out5[1] = c->blocks_read; out5[2] = c->bytes_written; out5[3] = c->throttle_waits;
  out5[4] = (int64_t)errs;
  return DRYAD_OK;
}

}  // extern "C"

// N5/W7: symmetric-memory substrate — device allocations that every rank maps (CUDA IPC over
// NVLink), so fused kernels can issue plain ld/st/atomics on peer HBM.
//
// Role in the reference: the channel layer that moves vertex outputs between processes as files
// served over HTTP (DryadVertex/VertexHost/system/channel/src/concreterchannel.cpp:1147-1230,
// ProcessService/HttpServer.cs:622-707).  Here a "channel" between stages on different GPUs is a
// peer-mapped HBM buffer plus an epoch flag.
#include "common.cuh"
#include "../../include/dryad_b200.h"
#include <string.h>

extern "C" {

int dryad_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
  return n;
}

int dryad_symm_alloc(size_t bytes, void** out) {
  DRYAD_CUDA_TRY(cudaMalloc(out, bytes));
  DRYAD_CUDA_TRY(cudaMemset(*out, 0, bytes));
  return DRYAD_OK;
}

int dryad_symm_free(void* p) {
  DRYAD_CUDA_TRY(cudaFree(p));
  return DRYAD_OK;
}

int dryad_ipc_export(void* p, unsigned char handle64[64]) {
  cudaIpcMemHandle_t h;
  DRYAD_CUDA_TRY(cudaIpcGetMemHandle(&h, p));
  static_assert(sizeof(h) == 64, "ipc handle size");
  memcpy(handle64, &h, 64);
  return DRYAD_OK;
}

int dryad_ipc_open(const unsigned char handle64[64], void** out) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  DRYAD_CUDA_TRY(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
  return DRYAD_OK;
}

int dryad_ipc_close(void* p) {
  DRYAD_CUDA_TRY(cudaIpcCloseMemHandle(p));
  return DRYAD_OK;
}

int dryad_memcpy_async(void* dst, const void* src, size_t bytes, int kind, void* stream) {
  DRYAD_CUDA_TRY(cudaMemcpyAsync(dst, src, bytes, (cudaMemcpyKind)kind, (cudaStream_t)stream));
  return DRYAD_OK;
}

int dryad_memset_async(void* dst, int value, size_t bytes, void* stream) {
  DRYAD_CUDA_TRY(cudaMemsetAsync(dst, value, bytes, (cudaStream_t)stream));
  return DRYAD_OK;
}

const char* dryad_error_string(int code) {
  if (code >= 0) return cudaGetErrorString((cudaError_t)code);
  switch (code) {
    case DRYAD_ERR_BAD_ARG: return "dryad: bad argument";
    case DRYAD_ERR_TOO_LARGE: return "dryad: size beyond supported maximum";
    case DRYAD_ERR_OVERFLOW: return "dryad: destination overflow";
    case DRYAD_ERR_SCRATCH: return "dryad: scratch buffer missing or too small";
    default: return "dryad: unknown error";
  }
}

/* Enqueue a host callback `fn(arg)` on the stream (cudaLaunchHostFunc): runs when all prior work of
 * the stream has completed.  Used to post vertex-completion messages into the host message pump. */
int dryad_stream_host_func(void* stream, void* fn, void* arg) {
  if (!fn) return DRYAD_ERR_BAD_ARG;
  cudaError_t e = cudaLaunchHostFunc((cudaStream_t)stream, (cudaHostFn_t)fn, arg);
  return e == cudaSuccess ? DRYAD_OK : (int)e;
}

}  // extern "C"

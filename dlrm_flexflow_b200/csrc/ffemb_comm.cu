// ffemb_comm.cu -- peer-memory plumbing for the table-wise sharded path.
//
// The reference moves embedding outputs and their gradients between GPUs
// implicitly: Embedding maps its output to zero-copy host memory
// (src/ops/embedding.cu:280-284,343-347) and Concat re-partitions it
// (src/ops/concat.cu:85-93), Legion DMA doing the copies through the host.
// Here every rank (one process per GPU) exposes its [B/N][T*D] output and
// gradient matrices to its peers by CUDA IPC; the forward kernel then stores
// pooled rows straight into the consuming rank's matrix over NVLink and the
// update kernel loads gradient rows straight from the producing rank's.
#include <string.h>

#include "ffemb_common.cuh"

using namespace ffemb;

extern "C" int ffemb_device_malloc(void **ptr, size_t bytes) {
  if (!ptr) return fail(FFEMB_ERR_INVALID_ARGUMENT, "device_malloc: NULL argument");
  FFEMB_CUDA_OK(cudaMalloc(ptr, bytes ? bytes : 1));
  return FFEMB_OK;
}

extern "C" int ffemb_device_free(void *ptr) {
  if (ptr) FFEMB_CUDA_OK(cudaFree(ptr));
  return FFEMB_OK;
}

extern "C" int ffemb_ipc_get_handle(void *device_ptr, uint8_t handle[64]) {
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
  if (!device_ptr || !handle) return fail(FFEMB_ERR_INVALID_ARGUMENT, "ipc_get_handle: NULL argument");
  cudaIpcMemHandle_t h;
  FFEMB_CUDA_OK(cudaIpcGetMemHandle(&h, device_ptr));
  memcpy(handle, &h, 64);
  return FFEMB_OK;
}

extern "C" int ffemb_ipc_open_handle(const uint8_t handle[64], void **device_ptr) {
  if (!device_ptr || !handle) return fail(FFEMB_ERR_INVALID_ARGUMENT, "ipc_open_handle: NULL argument");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, 64);
  FFEMB_CUDA_OK(cudaIpcOpenMemHandle(device_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return FFEMB_OK;
}

extern "C" int ffemb_ipc_close_handle(void *device_ptr) {
  if (device_ptr) FFEMB_CUDA_OK(cudaIpcCloseMemHandle(device_ptr));
  return FFEMB_OK;
}

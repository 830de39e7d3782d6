// heap.cu — symmetric-heap plumbing: the data plane that replaces the reference's ZMQ
// ROUTER/DEALER mesh (comm/mailbox.cpp:106-129).  A rank cudaMalloc's its shard slab,
// exports it as a CUDA IPC handle, and every peer maps it; kernels then address remote
// shards with ordinary ld/st/red over NVLink.  Handle exchange itself rides on the
// control plane (torch.distributed) and lives in minips_b200/parallel/heap.py.
#include <cstring>

#include "common.cuh"

extern "C" {

int mps_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
  return n;
}

int mps_set_device(int dev) { return static_cast<int>(cudaSetDevice(dev)); }

int mps_sm_count(int dev) {
  int n = 0;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
  return n;
}

int mps_heap_alloc(size_t bytes, void** out) {
  void* p = nullptr;
  MPS_CHECK(cudaMalloc(&p, bytes));
  cudaError_t e = cudaMemset(p, 0, bytes);
  if (e != cudaSuccess) {
    cudaFree(p);
    return static_cast<int>(e);
  }
  *out = p;
  return 0;
}

int mps_heap_free(void* p) { return static_cast<int>(cudaFree(p)); }

int mps_ipc_get_handle(void* p, unsigned char out[64]) {
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  cudaIpcMemHandle_t h;
  MPS_CHECK(cudaIpcGetMemHandle(&h, p));
  std::memcpy(out, &h, 64);
  return 0;
}

int mps_ipc_open_handle(const unsigned char h[64], void** out) {
  cudaIpcMemHandle_t handle;
  std::memcpy(&handle, h, 64);
  return static_cast<int>(cudaIpcOpenMemHandle(out, handle, cudaIpcMemLazyEnablePeerAccess));
}

int mps_ipc_close_handle(void* p) { return static_cast<int>(cudaIpcCloseMemHandle(p)); }

int mps_enable_peer(int peer_dev) {
  cudaError_t e = cudaDeviceEnablePeerAccess(peer_dev, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    cudaGetLastError();
    return 0;
  }
  return static_cast<int>(e);
}

int mps_memset32_async(void* p, int value, size_t n_ints, void* stream) {
  // cudaMemsetD32-equivalent through the runtime: value is a 32-bit pattern.
  if (value == 0) return static_cast<int>(cudaMemsetAsync(p, 0, n_ints * 4, static_cast<cudaStream_t>(stream)));
  return static_cast<int>(cudaErrorNotSupported);
}

const char* mps_error_string(int err) { return cudaGetErrorString(static_cast<cudaError_t>(err)); }

}  // extern "C"

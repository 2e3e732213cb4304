// Symmetric (peer-mapped) device memory for the P2P gradient engine.
//
// Each rank cudaMalloc's one arena, exports it as a CUDA IPC handle, and opens every peer's handle;
// the handles travel over the existing torch.distributed control plane (TCPStore rendezvous,
// ref Part 2a/main.py:148-153), never over a timed path.  On an NVSwitch box a peer mapping is a
// plain device pointer whose loads/stores travel over NVLink.
#include <cuda_runtime.h>
#include <string.h>

#include "../../include/ddp_b200.h"
#include "../common.cuh"

extern "C" int ddpb_symm_alloc(size_t bytes, void** ptr, void* handle64) {
  DDPB_REQUIRE(bytes > 0 && ptr && handle64, "symm_alloc: bad arguments");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is 64 bytes");
  void* p = nullptr;
  DDPB_CUDA_CHECK(cudaMalloc(&p, bytes));
  DDPB_CUDA_CHECK(cudaMemset(p, 0, bytes));
  cudaIpcMemHandle_t h;
  DDPB_CUDA_CHECK(cudaIpcGetMemHandle(&h, p));
  memcpy(handle64, &h, 64);
  *ptr = p;
  return DDPB_OK;
}

extern "C" int ddpb_symm_open(const void* handle64, void** peer_ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  DDPB_CUDA_CHECK(cudaIpcOpenMemHandle(peer_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return DDPB_OK;
}

extern "C" int ddpb_symm_close(void* peer_ptr) {
  DDPB_CUDA_CHECK(cudaIpcCloseMemHandle(peer_ptr));
  return DDPB_OK;
}

extern "C" int ddpb_symm_free(void* ptr) {
  DDPB_CUDA_CHECK(cudaFree(ptr));
  return DDPB_OK;
}

extern "C" int ddpb_enable_peer_access(int peer_device) {
  cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return DDPB_OK; }
  DDPB_CUDA_CHECK(e);
  return DDPB_OK;
}

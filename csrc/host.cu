// Host-side runtime pieces with a C ABI: raw device allocations that can be shared between the
// per-GPU processes through CUDA IPC (the fallback path when torch's symmetric-memory rendezvous is
// unavailable), peer-access set-up and a pinned staging allocator for the H2D input pipeline.
#include "common.cuh"
#include "../include/fedml_b200.h"

extern "C" int fb_dev_malloc(void** out, int64_t bytes) {
  FB_REQUIRE(out && bytes > 0, "fb_dev_malloc: bad arguments");
  FB_CHECK_CUDA(cudaMalloc(out, (size_t)bytes));
  FB_CHECK_CUDA(cudaMemset(*out, 0, (size_t)bytes));
  return FB_OK;
}
extern "C" int fb_dev_free(void* p) {
  if (p) FB_CHECK_CUDA(cudaFree(p));
  return FB_OK;
}
extern "C" int fb_ipc_handle_bytes(void) { return (int)sizeof(cudaIpcMemHandle_t); }
extern "C" int fb_ipc_get_handle(void* dev_ptr, void* handle_out) {
  FB_REQUIRE(dev_ptr && handle_out, "fb_ipc_get_handle: bad arguments");
  cudaIpcMemHandle_t h;
  FB_CHECK_CUDA(cudaIpcGetMemHandle(&h, dev_ptr));
  memcpy(handle_out, &h, sizeof(h));
  return FB_OK;
}
extern "C" int fb_ipc_open_handle(const void* handle, void** out) {
  FB_REQUIRE(handle && out, "fb_ipc_open_handle: bad arguments");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  FB_CHECK_CUDA(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
  return FB_OK;
}
extern "C" int fb_ipc_close_handle(void* p) {
  if (p) FB_CHECK_CUDA(cudaIpcCloseMemHandle(p));
  return FB_OK;
}
extern "C" int fb_enable_peer_access(int peer_device) {
  int dev = 0, can = 0;
  FB_CHECK_CUDA(cudaGetDevice(&dev));
  if (peer_device == dev) return FB_OK;
  FB_CHECK_CUDA(cudaDeviceCanAccessPeer(&can, dev, peer_device));
  FB_REQUIRE(can, "fb_enable_peer_access: device %d cannot access device %d", dev, peer_device);
  cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
    fb_set_error("cudaDeviceEnablePeerAccess failed: %s", cudaGetErrorString(e));
    return FB_ERR_CUDA;
  }
  cudaGetLastError();
  return FB_OK;
}
extern "C" int fb_host_alloc_pinned(void** out, int64_t bytes) {
  FB_REQUIRE(out && bytes > 0, "fb_host_alloc_pinned: bad arguments");
  FB_CHECK_CUDA(cudaHostAlloc(out, (size_t)bytes, cudaHostAllocDefault));
  return FB_OK;
}
extern "C" int fb_host_free_pinned(void* p) {
  if (p) FB_CHECK_CUDA(cudaFreeHost(p));
  return FB_OK;
}

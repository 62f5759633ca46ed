// Peer-memory fabric + library runtime (SURVEY build-plan step 3).
//
// One process per GPU.  Each rank allocates its symmetric buffers with cudaMalloc,
// exports a CUDA IPC handle, and opens every peer's handle: after that a peer's HBM
// is addressable with ordinary global loads/stores/reductions that travel over
// NVLink 5 / NVSwitch.  Handles are exchanged by the Python side (torch.distributed
// or the ps control plane).  Nothing here launches a collective library.
#include "common.cuh"

#include <mutex>
#include <string>

static std::mutex g_err_mu;
static char g_last_error[512] = "";

extern "C" void dtg_set_last_error(const char* what, const char* detail) {
  std::lock_guard<std::mutex> g(g_err_mu);
  snprintf(g_last_error, sizeof(g_last_error), "%s: %s", what ? what : "?", detail ? detail : "");
}

extern "C" const char* dtg_last_error() { return g_last_error; }

extern "C" int dtg_version() { return 100; }

extern "C" int dtg_ipc_handle_size() { return (int)sizeof(cudaIpcMemHandle_t); }

// Allocate `bytes` of zeroed device memory on the current device and export it.
extern "C" int dtg_ipc_alloc(size_t bytes, void** ptr, void* handle_out) {
  if (ptr == nullptr || bytes == 0) {
    dtg_set_last_error("dtg_ipc_alloc", "bad argument");
    return dtg::ERR_BAD_ARG;
  }
  DTG_CUDA_OK(cudaMalloc(ptr, bytes));
  DTG_CUDA_OK(cudaMemset(*ptr, 0, bytes));
  if (handle_out != nullptr) {
    cudaIpcMemHandle_t h;
    DTG_CUDA_OK(cudaIpcGetMemHandle(&h, *ptr));
    memcpy(handle_out, &h, sizeof(h));
  }
  return dtg::OK;
}

extern "C" int dtg_ipc_open(const void* handle, void** ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  DTG_CUDA_OK(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return dtg::OK;
}

extern "C" int dtg_ipc_close(void* ptr) {
  DTG_CUDA_OK(cudaIpcCloseMemHandle(ptr));
  return dtg::OK;
}

extern "C" int dtg_free(void* ptr) {
  DTG_CUDA_OK(cudaFree(ptr));
  return dtg::OK;
}

// Single-process multi-GPU (tests, tower mode in one process): direct peer access.
extern "C" int dtg_enable_peer_access(int dev, int peer) {
  int can = 0;
  DTG_CUDA_OK(cudaDeviceCanAccessPeer(&can, dev, peer));
  if (!can) {
    dtg_set_last_error("dtg_enable_peer_access", "devices are not peers");
    return dtg::ERR_UNSUPPORTED;
  }
  int cur = 0;
  DTG_CUDA_OK(cudaGetDevice(&cur));
  DTG_CUDA_OK(cudaSetDevice(dev));
  cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
  cudaSetDevice(cur);
  DTG_CUDA_OK(e);
  return dtg::OK;
}

extern "C" int dtg_memset_async(void* ptr, int value, size_t bytes, void* stream) {
  DTG_CUDA_OK(cudaMemsetAsync(ptr, value, bytes, (cudaStream_t)stream));
  return dtg::OK;
}

extern "C" int dtg_memcpy_async(void* dst, const void* src, size_t bytes, void* stream) {
  DTG_CUDA_OK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, (cudaStream_t)stream));
  return dtg::OK;
}

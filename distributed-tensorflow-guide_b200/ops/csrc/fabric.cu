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

// ---------------------------------------------------------------------------
// Striped parameter-server memory (CUDA virtual memory management).
// Every rank contributes one physical allocation; all ranks map the allocations into ONE
// virtual range, interleaved at the allocation granularity (2 MB): stripe i lives on rank
// i % world.  A ps buffer is then a single flat tensor whose traffic is spread over all
// NVLink ports no matter which variable is being pulled or pushed (a 67 MB kernel no longer
// makes its owner a hot spot), and the fused push epilogue / pull kernels need no shard table.
// Handles travel between processes as POSIX file descriptors (SCM_RIGHTS, Python side).
// ---------------------------------------------------------------------------
namespace {
struct VmmApi {
  CUresult (*create)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  CUresult (*release)(CUmemGenericAllocationHandle);
  CUresult (*exportH)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
  CUresult (*importH)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUresult (*reserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*addrFree)(CUdeviceptr, size_t);
  CUresult (*map)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*unmap)(CUdeviceptr, size_t);
  CUresult (*setAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUresult (*granularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  bool ok = false;
};
VmmApi& vmm() {
  static VmmApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    auto get = [](const char* name, void** fn) {
      cudaDriverEntryPointQueryResult q;
      return cudaGetDriverEntryPoint(name, fn, cudaEnableDefault, &q) == cudaSuccess &&
             q == cudaDriverEntryPointSuccess && *fn != nullptr;
    };
    api.ok = get("cuMemCreate", (void**)&api.create) && get("cuMemRelease", (void**)&api.release) &&
             get("cuMemExportToShareableHandle", (void**)&api.exportH) &&
             get("cuMemImportFromShareableHandle", (void**)&api.importH) &&
             get("cuMemAddressReserve", (void**)&api.reserve) &&
             get("cuMemAddressFree", (void**)&api.addrFree) && get("cuMemMap", (void**)&api.map) &&
             get("cuMemUnmap", (void**)&api.unmap) && get("cuMemSetAccess", (void**)&api.setAccess) &&
             get("cuMemGetAllocationGranularity", (void**)&api.granularity);
  });
  return api;
}
CUmemAllocationProp vmm_prop(int dev) {
  CUmemAllocationProp prop;
  memset(&prop, 0, sizeof(prop));
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = dev;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}
int vmm_fail(const char* what, CUresult r) {
  char buf[64];
  snprintf(buf, sizeof(buf), "CUresult %d", (int)r);
  dtg_set_last_error(what, buf);
  return dtg::ERR_CUDA;
}
}  // namespace

#define DTG_VMM_READY()                                                        \
  if (!vmm().ok) {                                                             \
    dtg_set_last_error("vmm", "CUDA virtual memory management API unavailable"); \
    return dtg::ERR_UNSUPPORTED;                                               \
  }

extern "C" int dtg_vmm_granularity(size_t* out) {
  DTG_VMM_READY();
  int dev = 0;
  DTG_CUDA_OK(cudaGetDevice(&dev));
  DTG_CUDA_OK(cudaFree(0));
  CUmemAllocationProp prop = vmm_prop(dev);
  CUresult r = vmm().granularity(out, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED);
  return r == CUDA_SUCCESS ? dtg::OK : vmm_fail("cuMemGetAllocationGranularity", r);
}

// Physical memory on the current device (bytes: multiple of the granularity) + its fd.
extern "C" int dtg_vmm_create(size_t bytes, unsigned long long* handle, int* fd) {
  DTG_VMM_READY();
  int dev = 0;
  DTG_CUDA_OK(cudaGetDevice(&dev));
  DTG_CUDA_OK(cudaFree(0));
  CUmemAllocationProp prop = vmm_prop(dev);
  CUmemGenericAllocationHandle h;
  CUresult r = vmm().create(&h, bytes, &prop, 0);
  if (r != CUDA_SUCCESS) return vmm_fail("cuMemCreate", r);
  int f = -1;
  r = vmm().exportH(&f, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) { vmm().release(h); return vmm_fail("cuMemExportToShareableHandle", r); }
  *handle = h;
  *fd = f;
  return dtg::OK;
}

extern "C" int dtg_vmm_import(int fd, unsigned long long* handle) {
  DTG_VMM_READY();
  DTG_CUDA_OK(cudaFree(0));
  CUmemGenericAllocationHandle h;
  CUresult r = vmm().importH(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  if (r != CUDA_SUCCESS) return vmm_fail("cuMemImportFromShareableHandle", r);
  *handle = h;
  return dtg::OK;
}

// Reserve `stripes * stripe_bytes` of address space and map stripe i to handles[i] (one
// physical allocation per stripe: cuMemMap takes no offset); grant the current device RW.
extern "C" int dtg_vmm_map_striped(const unsigned long long* handles, size_t stripe_bytes,
                                   size_t stripes, void** va_out) {
  DTG_VMM_READY();
  int dev = 0;
  DTG_CUDA_OK(cudaGetDevice(&dev));
  const size_t total = stripes * stripe_bytes;
  CUdeviceptr va = 0;
  CUresult r = vmm().reserve(&va, total, stripe_bytes, 0, 0);
  if (r != CUDA_SUCCESS) return vmm_fail("cuMemAddressReserve", r);
  for (size_t i = 0; i < stripes; ++i) {
    r = vmm().map(va + i * stripe_bytes, stripe_bytes, 0, handles[i], 0);
    if (r != CUDA_SUCCESS) return vmm_fail("cuMemMap", r);
  }
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = dev;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = vmm().setAccess(va, total, &acc, 1);
  if (r != CUDA_SUCCESS) return vmm_fail("cuMemSetAccess", r);
  *va_out = (void*)va;
  return dtg::OK;
}

extern "C" int dtg_vmm_unmap(void* va, size_t total) {
  DTG_VMM_READY();
  CUresult r = vmm().unmap((CUdeviceptr)va, total);
  if (r != CUDA_SUCCESS) return vmm_fail("cuMemUnmap", r);
  r = vmm().addrFree((CUdeviceptr)va, total);
  return r == CUDA_SUCCESS ? dtg::OK : vmm_fail("cuMemAddressFree", r);
}

extern "C" int dtg_vmm_release(unsigned long long handle) {
  DTG_VMM_READY();
  CUresult r = vmm().release((CUmemGenericAllocationHandle)handle);
  return r == CUDA_SUCCESS ? dtg::OK : vmm_fail("cuMemRelease", r);
}

// ---------------------------------------------------------------------------
// NVLS multicast objects (NVSwitch in-fabric reduction / broadcast).  A multicast object
// binds one physical allocation per GPU; a `multimem.ld_reduce` through its mapping returns
// the SUM over all bound GPUs computed in the switch, a `multimem.st` writes all of them.
// sync_kernels.cu uses both for the reduce-scatter and all-gather halves of the SSGD step.
//   rank 0: dtg_mc_create -> fd to everyone;  all: dtg_mc_add_device;  barrier;
//   all: dtg_vmm_create (own memory) + dtg_mc_bind;  all: dtg_vmm_map_one(mc handle).
// ---------------------------------------------------------------------------
namespace {
struct McApi {
  CUresult (*create)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  CUresult (*addDevice)(CUmemGenericAllocationHandle, CUdevice);
  CUresult (*bindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t,
                      size_t, unsigned long long);
  CUresult (*granularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  CUresult (*devAttr)(int*, CUdevice_attribute, CUdevice);
  CUresult (*devGet)(CUdevice*, int);
  CUresult (*unbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t) = nullptr;  // optional
  bool ok = false;
};
McApi& mc() {
  static McApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    auto get = [](const char* name, void** fn) {
      cudaDriverEntryPointQueryResult q;
      return cudaGetDriverEntryPoint(name, fn, cudaEnableDefault, &q) == cudaSuccess &&
             q == cudaDriverEntryPointSuccess && *fn != nullptr;
    };
    api.ok = get("cuMulticastCreate", (void**)&api.create) &&
             get("cuMulticastAddDevice", (void**)&api.addDevice) &&
             get("cuMulticastBindMem", (void**)&api.bindMem) &&
             get("cuMulticastGetGranularity", (void**)&api.granularity) &&
             get("cuDeviceGetAttribute", (void**)&api.devAttr) &&
             get("cuDeviceGet", (void**)&api.devGet);
    if (!get("cuMulticastUnbind", (void**)&api.unbind)) api.unbind = nullptr;
  });
  return api;
}
CUmulticastObjectProp mc_prop(int world, size_t size) {
  CUmulticastObjectProp prop;
  memset(&prop, 0, sizeof(prop));
  prop.numDevices = (unsigned)world;
  prop.size = size;
  prop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}
}  // namespace

extern "C" int dtg_mc_supported() {
  if (!vmm().ok || !mc().ok) return 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || cudaFree(0) != cudaSuccess) return 0;
  CUdevice cd;
  int v = 0;
  if (mc().devGet(&cd, dev) != CUDA_SUCCESS) return 0;
  if (mc().devAttr(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, cd) != CUDA_SUCCESS) return 0;
  return v;
}

extern "C" int dtg_mc_granularity(int world, size_t size, size_t* out) {
  DTG_VMM_READY();
  if (!mc().ok) { dtg_set_last_error("mc", "multicast API unavailable"); return dtg::ERR_UNSUPPORTED; }
  CUmulticastObjectProp prop = mc_prop(world, size);
  CUresult r = mc().granularity(out, &prop, CU_MULTICAST_GRANULARITY_RECOMMENDED);
  return r == CUDA_SUCCESS ? dtg::OK : vmm_fail("cuMulticastGetGranularity", r);
}

extern "C" int dtg_mc_create(int world, size_t size, unsigned long long* handle, int* fd) {
  DTG_VMM_READY();
  DTG_CUDA_OK(cudaFree(0));
  CUmulticastObjectProp prop = mc_prop(world, size);
  CUmemGenericAllocationHandle h;
  CUresult r = mc().create(&h, &prop);
  if (r != CUDA_SUCCESS) return vmm_fail("cuMulticastCreate", r);
  int f = -1;
  r = vmm().exportH(&f, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) return vmm_fail("cuMemExportToShareableHandle(mc)", r);
  *handle = h;
  *fd = f;
  return dtg::OK;
}

extern "C" int dtg_mc_add_device(unsigned long long handle) {
  DTG_VMM_READY();
  int dev = 0;
  DTG_CUDA_OK(cudaGetDevice(&dev));
  CUdevice cd;
  CUresult r = mc().devGet(&cd, dev);
  if (r != CUDA_SUCCESS) return vmm_fail("cuDeviceGet", r);
  r = mc().addDevice((CUmemGenericAllocationHandle)handle, cd);
  return r == CUDA_SUCCESS ? dtg::OK : vmm_fail("cuMulticastAddDevice", r);
}

extern "C" int dtg_mc_bind(unsigned long long mc_handle, unsigned long long mem_handle, size_t size) {
  DTG_VMM_READY();
  CUresult r = mc().bindMem((CUmemGenericAllocationHandle)mc_handle, 0,
                            (CUmemGenericAllocationHandle)mem_handle, 0, size, 0);
  return r == CUDA_SUCCESS ? dtg::OK : vmm_fail("cuMulticastBindMem", r);
}

// Teardown half of dtg_mc_bind: detach this device's memory from the multicast object so the
// physical allocation can be released (Fabric.close()).  Call after the multicast mapping
// has been unmapped on this rank.
extern "C" int dtg_mc_unbind(unsigned long long mc_handle, size_t size) {
  DTG_VMM_READY();
  if (mc().unbind == nullptr) return dtg::OK;     // old driver: the object dies with the process
  int dev = 0;
  DTG_CUDA_OK(cudaGetDevice(&dev));
  CUdevice cd;
  CUresult r = mc().devGet(&cd, dev);
  if (r != CUDA_SUCCESS) return vmm_fail("cuDeviceGet", r);
  r = mc().unbind((CUmemGenericAllocationHandle)mc_handle, cd, 0, size);
  return r == CUDA_SUCCESS ? dtg::OK : vmm_fail("cuMulticastUnbind", r);
}

// Map one handle (physical allocation or multicast object) and grant the current device RW.
extern "C" int dtg_vmm_map_one(unsigned long long handle, size_t size, size_t align, void** va_out) {
  DTG_VMM_READY();
  int dev = 0;
  DTG_CUDA_OK(cudaGetDevice(&dev));
  CUdeviceptr va = 0;
  CUresult r = vmm().reserve(&va, size, align, 0, 0);
  if (r != CUDA_SUCCESS) return vmm_fail("cuMemAddressReserve", r);
  r = vmm().map(va, size, 0, (CUmemGenericAllocationHandle)handle, 0);
  if (r != CUDA_SUCCESS) return vmm_fail("cuMemMap", r);
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = dev;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = vmm().setAccess(va, size, &acc, 1);
  if (r != CUDA_SUCCESS) return vmm_fail("cuMemSetAccess", r);
  *va_out = (void*)va;
  return dtg::OK;
}

// ---------------------------------------------------------------------------
// CRC-32C (Castagnoli), slicing-by-8, host side: tensor and table-block checksums of the
// TF checkpoint V2 writer / reader (session/tf_bundle.py).
// ---------------------------------------------------------------------------
namespace {
struct Crc32cTables {
  uint32_t t[8][256];
  Crc32cTables() {
    for (uint32_t i = 0; i < 256; ++i) {
      uint32_t c = i;
      for (int k = 0; k < 8; ++k) c = (c & 1) ? (c >> 1) ^ 0x82F63B78u : c >> 1;
      t[0][i] = c;
    }
    for (uint32_t i = 0; i < 256; ++i)
      for (int s = 1; s < 8; ++s) t[s][i] = (t[s - 1][i] >> 8) ^ t[0][t[s - 1][i] & 0xFF];
  }
};
}  // namespace

extern "C" uint32_t dtg_crc32c(const void* data, size_t n, uint32_t crc) {
  static const Crc32cTables T;
  const uint8_t* p = static_cast<const uint8_t*>(data);
  uint32_t c = crc ^ 0xFFFFFFFFu;
  while (n >= 8) {
    uint32_t lo, hi;
    memcpy(&lo, p, 4);
    memcpy(&hi, p + 4, 4);
    lo ^= c;
    c = T.t[7][lo & 0xFF] ^ T.t[6][(lo >> 8) & 0xFF] ^ T.t[5][(lo >> 16) & 0xFF] ^ T.t[4][lo >> 24] ^
        T.t[3][hi & 0xFF] ^ T.t[2][(hi >> 8) & 0xFF] ^ T.t[1][(hi >> 16) & 0xFF] ^ T.t[0][hi >> 24];
    p += 8;
    n -= 8;
  }
  while (n--) c = T.t[0][(c ^ *p++) & 0xFF] ^ (c >> 8);
  return c ^ 0xFFFFFFFFu;
}

// Symmetric peer-memory heap on the CUDA VMM driver API (W3 of SURVEY.md §8).
//
// One physical allocation per rank (cuMemCreate, exportable as a POSIX file descriptor), mapped into
// every rank of the job at its own virtual address, plus — when the fabric supports it — one NVSwitch
// multicast object bound to all ranks' allocations and mapped at a second address for `multimem.*`.
// File descriptors travel between the (separately spawned, Spark-launched) processes over Unix sockets
// with SCM_RIGHTS; that exchange is done by the Python side (`parallel/peer.py`), this file only wraps
// the driver calls.  The reference moves parameters as pickles over TCP sockets; this
// is its NVLink-native replacement.  Driver entry points are resolved at run time with
// cudaGetDriverEntryPoint so the library also loads on a box without libcuda (the CPU build box).
#include <cuda.h>
#include <cuda_runtime.h>
#include <unistd.h>

#include <cstdint>
#include <cstring>
#include <mutex>

#include "../include/ddls_b200.h"

namespace ddls {
void set_last_error(const char* fmt, ...);
}
using ddls::set_last_error;

namespace {

struct Driver {
  decltype(&cuMemCreate) MemCreate = nullptr;
  decltype(&cuMemRelease) MemRelease = nullptr;
  decltype(&cuMemAddressReserve) MemAddressReserve = nullptr;
  decltype(&cuMemAddressFree) MemAddressFree = nullptr;
  decltype(&cuMemMap) MemMap = nullptr;
  decltype(&cuMemUnmap) MemUnmap = nullptr;
  decltype(&cuMemSetAccess) MemSetAccess = nullptr;
  decltype(&cuMemExportToShareableHandle) MemExport = nullptr;
  decltype(&cuMemImportFromShareableHandle) MemImport = nullptr;
  decltype(&cuMemGetAllocationGranularity) MemGranularity = nullptr;
  decltype(&cuMulticastCreate) McCreate = nullptr;
  decltype(&cuMulticastAddDevice) McAddDevice = nullptr;
  decltype(&cuMulticastBindMem) McBindMem = nullptr;
  decltype(&cuMulticastGetGranularity) McGranularity = nullptr;
  decltype(&cuMulticastUnbind) McUnbind = nullptr;
  decltype(&cuDeviceGetAttribute) DeviceGetAttribute = nullptr;
  decltype(&cuDeviceGet) DeviceGet = nullptr;
  decltype(&cuGetErrorString) GetErrorString = nullptr;
  bool ok = false;
};

Driver g_drv;
std::once_flag g_once;

template <typename T>
bool load(const char* name, T& fn) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult st;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &st) != cudaSuccess || st != cudaDriverEntryPointSuccess ||
      p == nullptr)
    return false;
  fn = reinterpret_cast<T>(p);
  return true;
}

bool driver() {
  std::call_once(g_once, [] {
    Driver& d = g_drv;
    bool ok = load("cuMemCreate", d.MemCreate) && load("cuMemRelease", d.MemRelease) &&
              load("cuMemAddressReserve", d.MemAddressReserve) && load("cuMemAddressFree", d.MemAddressFree) &&
              load("cuMemMap", d.MemMap) && load("cuMemUnmap", d.MemUnmap) && load("cuMemSetAccess", d.MemSetAccess) &&
              load("cuMemExportToShareableHandle", d.MemExport) &&
              load("cuMemImportFromShareableHandle", d.MemImport) &&
              load("cuMemGetAllocationGranularity", d.MemGranularity) &&
              load("cuDeviceGetAttribute", d.DeviceGetAttribute) && load("cuDeviceGet", d.DeviceGet) &&
              load("cuGetErrorString", d.GetErrorString);
    // multicast entry points are optional (older drivers / no NVSwitch)
    load("cuMulticastCreate", d.McCreate);
    load("cuMulticastAddDevice", d.McAddDevice);
    load("cuMulticastBindMem", d.McBindMem);
    load("cuMulticastGetGranularity", d.McGranularity);
    load("cuMulticastUnbind", d.McUnbind);
    d.ok = ok;
  });
  if (!g_drv.ok) set_last_error("CUDA driver entry points unavailable (no libcuda / no GPU)");
  return g_drv.ok;
}

int fail(CUresult r, const char* what) {
  const char* msg = nullptr;
  if (g_drv.GetErrorString) g_drv.GetErrorString(r, &msg);
  set_last_error("%s: %s (CUresult %d)", what, msg ? msg : "?", (int)r);
  return 1;
}

#define DRV(call, what)                         \
  do {                                          \
    CUresult _r = (call);                       \
    if (_r != CUDA_SUCCESS) return fail(_r, what); \
  } while (0)

CUmemAllocationProp alloc_prop(int device) {
  CUmemAllocationProp prop;
  memset(&prop, 0, sizeof(prop));
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}

int map_handle(CUmemGenericAllocationHandle h, size_t size, size_t align, int device, void** out_ptr) {
  CUdeviceptr va = 0;
  DRV(g_drv.MemAddressReserve(&va, size, align, 0, 0), "cuMemAddressReserve");
  CUresult r = g_drv.MemMap(va, size, 0, h, 0);
  if (r != CUDA_SUCCESS) { g_drv.MemAddressFree(va, size); return fail(r, "cuMemMap"); }
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = g_drv.MemSetAccess(va, size, &acc, 1);
  if (r != CUDA_SUCCESS) { g_drv.MemUnmap(va, size); g_drv.MemAddressFree(va, size); return fail(r, "cuMemSetAccess"); }
  *out_ptr = reinterpret_cast<void*>(va);
  return 0;
}

}  // namespace

extern "C" int ddls_multicast_supported(int device) {
  if (!driver() || g_drv.McCreate == nullptr) return 0;
  cudaFree(0);
  CUdevice dev;
  if (g_drv.DeviceGet(&dev, device) != CUDA_SUCCESS) return 0;
  int v = 0;
  if (g_drv.DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) != CUDA_SUCCESS) return 0;
  return v;
}

// Granularity every heap size must be a multiple of (max of allocation and multicast granularity).
extern "C" int ddls_symm_granularity(int device, int world, size_t* out) {
  if (!driver()) return 1;
  cudaFree(0);
  CUmemAllocationProp prop = alloc_prop(device);
  size_t g = 0;
  DRV(g_drv.MemGranularity(&g, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED), "cuMemGetAllocationGranularity");
  if (world > 1 && ddls_multicast_supported(device) && g_drv.McGranularity) {
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof(mp));
    mp.numDevices = world;
    mp.size = g;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t mg = 0;
    if (g_drv.McGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg > g) g = mg;
  }
  *out = g;
  return 0;
}

// Create this rank's physical allocation, map it locally, export an fd for the peers.
extern "C" int ddls_symm_create(int device, size_t size, unsigned long long* handle_out, void** ptr_out, int* fd_out) {
  if (!driver()) return 1;
  cudaFree(0);
  CUmemAllocationProp prop = alloc_prop(device);
  CUmemGenericAllocationHandle h;
  DRV(g_drv.MemCreate(&h, size, &prop, 0), "cuMemCreate");
  int fd = -1;
  CUresult r = g_drv.MemExport(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) { g_drv.MemRelease(h); return fail(r, "cuMemExportToShareableHandle"); }
  void* p = nullptr;
  if (map_handle(h, size, 0, device, &p)) { close(fd); g_drv.MemRelease(h); return 1; }
  *handle_out = static_cast<unsigned long long>(h);
  *ptr_out = p;
  *fd_out = fd;
  return 0;
}

// Map a peer's allocation (received as an fd) into this process for `device`.
extern "C" int ddls_symm_import(int device, int fd, size_t size, unsigned long long* handle_out, void** ptr_out) {
  if (!driver()) return 1;
  CUmemGenericAllocationHandle h;
  DRV(g_drv.MemImport(&h, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)), CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR),
      "cuMemImportFromShareableHandle");
  void* p = nullptr;
  if (map_handle(h, size, 0, device, &p)) { g_drv.MemRelease(h); return 1; }
  *handle_out = static_cast<unsigned long long>(h);
  *ptr_out = p;
  return 0;
}

extern "C" int ddls_symm_unmap(void* ptr, size_t size, unsigned long long handle) {
  if (!driver()) return 1;
  CUdeviceptr va = reinterpret_cast<CUdeviceptr>(ptr);
  if (va) { g_drv.MemUnmap(va, size); g_drv.MemAddressFree(va, size); }
  if (handle) g_drv.MemRelease(static_cast<CUmemGenericAllocationHandle>(handle));
  return 0;
}

// Rank 0: create the multicast object for `world` devices and export it.
extern "C" int ddls_mc_create(int world, size_t size, unsigned long long* mc_handle_out, int* fd_out) {
  if (!driver() || g_drv.McCreate == nullptr) { set_last_error("multicast entry points unavailable"); return 1; }
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof(mp));
  mp.numDevices = world;
  mp.size = size;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUmemGenericAllocationHandle h;
  DRV(g_drv.McCreate(&h, &mp), "cuMulticastCreate");
  int fd = -1;
  CUresult r = g_drv.MemExport(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) { g_drv.MemRelease(h); return fail(r, "cuMemExportToShareableHandle(multicast)"); }
  *mc_handle_out = static_cast<unsigned long long>(h);
  *fd_out = fd;
  return 0;
}

extern "C" int ddls_mc_import(int fd, unsigned long long* mc_handle_out) {
  if (!driver()) return 1;
  CUmemGenericAllocationHandle h;
  DRV(g_drv.MemImport(&h, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)), CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR),
      "cuMemImportFromShareableHandle(multicast)");
  *mc_handle_out = static_cast<unsigned long long>(h);
  return 0;
}

// Every rank: join the multicast team.  All ranks must have joined before anyone binds.
extern "C" int ddls_mc_add_device(unsigned long long mc_handle, int device) {
  if (!driver() || g_drv.McAddDevice == nullptr) return 1;
  CUdevice dev;
  DRV(g_drv.DeviceGet(&dev, device), "cuDeviceGet");
  DRV(g_drv.McAddDevice(static_cast<CUmemGenericAllocationHandle>(mc_handle), dev), "cuMulticastAddDevice");
  return 0;
}

// Every rank: bind its physical heap to the multicast object and map the multicast view.
extern "C" int ddls_mc_bind_and_map(unsigned long long mc_handle, unsigned long long mem_handle, size_t size, int device,
                                    void** mc_ptr_out) {
  if (!driver() || g_drv.McBindMem == nullptr) return 1;
  DRV(g_drv.McBindMem(static_cast<CUmemGenericAllocationHandle>(mc_handle), 0,
                      static_cast<CUmemGenericAllocationHandle>(mem_handle), 0, size, 0),
      "cuMulticastBindMem");
  return map_handle(static_cast<CUmemGenericAllocationHandle>(mc_handle), size, 0, device, mc_ptr_out);
}

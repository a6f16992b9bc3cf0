// K1b, NVLS variant: plumbing for NVLink-SHARP multicast memory (CUDA driver VMM API), used by k_fedavg_p2p when a rank record
// carries multicast pointers.  A multicast object spans one physical allocation on every GPU; a multimem.ld_reduce on its address
// returns the element-wise SUM over all GPUs' copies, computed inside the NVSwitch, and a multimem.st writes every copy at once.
//   rank 0:      fedb_nvls_create -> (handle, POSIX fd)            [fd travels to the other ranks over a Unix socket: parallel/nvls.py]
//   other ranks: fedb_nvls_import(fd) -> handle
//   every rank:  fedb_nvls_add_device;  <all ranks barrier>;  fedb_nvls_bind_map -> unicast + multicast views of its own allocation
// The driver entry points are resolved at run time (cudaGetDriverEntryPoint): the library does not link libcuda.
#include <cuda.h>

#include "common.cuh"

using namespace fedb;

namespace {

template <typename F>
int driver_fn(const char* name, F* out) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !p) {
    (void)cudaGetLastError();
    char msg[160];
    snprintf(msg, sizeof msg, "driver entry point %s is not available", name);
    return set_err(FEDB_E_UNSUPPORTED, msg);
  }
  *out = reinterpret_cast<F>(p);
  return FEDB_OK;
}

#define FEDB_DRV(name, ...)                                                                              \
  do {                                                                                                   \
    decltype(&name) fn_ = nullptr;                                                                       \
    if (int rc_ = driver_fn(#name, &fn_)) return rc_;                                                    \
    CUresult r_ = fn_(__VA_ARGS__);                                                                      \
    if (r_ != CUDA_SUCCESS) {                                                                            \
      char msg_[200];                                                                                    \
      snprintf(msg_, sizeof msg_, "%s:%d: %s failed with CUresult %d", __FILE__, __LINE__, #name, (int)r_); \
      return set_err(FEDB_E_CUDA, msg_);                                                                 \
    }                                                                                                    \
  } while (0)

CUmulticastObjectProp mc_prop(int n_devices, size_t bytes) {
  CUmulticastObjectProp p{};
  p.numDevices = (unsigned)n_devices;
  p.size = bytes;
  p.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  p.flags = 0;
  return p;
}

}  // namespace

extern "C" int fedb_nvls_supported(int32_t device) {
  FEDB_CUDA_OK(cudaFree(nullptr));   // make sure the primary context exists
  CUdevice dev;
  FEDB_DRV(cuDeviceGet, &dev, device);
  int ok = 0;
  FEDB_DRV(cuDeviceGetAttribute, &ok, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev);
  return ok ? 1 : 0;
}

// bytes is rounded up to the multicast granularity; *bytes_out is what every rank must allocate and bind.
extern "C" int fedb_nvls_create(int32_t n_devices, int64_t bytes, uint64_t* mc_handle_out, int32_t* fd_out, int64_t* bytes_out) {
  FEDB_REQUIRE(n_devices >= 2 && bytes > 0 && mc_handle_out && fd_out && bytes_out, "bad arguments");
  FEDB_CUDA_OK(cudaFree(nullptr));
  CUmulticastObjectProp p = mc_prop(n_devices, (size_t)bytes);
  size_t gran = 0;
  FEDB_DRV(cuMulticastGetGranularity, &gran, &p, CU_MULTICAST_GRANULARITY_RECOMMENDED);
  p.size = ((size_t)bytes + gran - 1) / gran * gran;
  CUmemGenericAllocationHandle h;
  FEDB_DRV(cuMulticastCreate, &h, &p);
  int fd = -1;
  FEDB_DRV(cuMemExportToShareableHandle, &fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  *mc_handle_out = (uint64_t)h;
  *fd_out = fd;
  *bytes_out = (int64_t)p.size;
  return FEDB_OK;
}

extern "C" int fedb_nvls_import(int32_t fd, uint64_t* mc_handle_out) {
  FEDB_REQUIRE(fd >= 0 && mc_handle_out, "bad arguments");
  FEDB_CUDA_OK(cudaFree(nullptr));
  CUmemGenericAllocationHandle h;
  FEDB_DRV(cuMemImportFromShareableHandle, &h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  *mc_handle_out = (uint64_t)h;
  return FEDB_OK;
}

extern "C" int fedb_nvls_add_device(uint64_t mc_handle, int32_t device) {
  CUdevice dev;
  FEDB_DRV(cuDeviceGet, &dev, device);
  FEDB_DRV(cuMulticastAddDevice, (CUmemGenericAllocationHandle)mc_handle, dev);
  return FEDB_OK;
}

// Call after EVERY rank has added its device.  Allocates `bytes` (the size returned by fedb_nvls_create) on `device`, binds it to
// the multicast object at offset 0, maps it twice: *uc_ptr is this GPU's own copy (ordinary loads/stores), *mc_ptr the multicast
// view (multimem.* instructions only).  The memory is zeroed.
extern "C" int fedb_nvls_bind_map(uint64_t mc_handle, int32_t device, int64_t bytes, uint64_t* mem_handle_out, void** uc_ptr,
                                  void** mc_ptr) {
  FEDB_REQUIRE(bytes > 0 && mem_handle_out && uc_ptr && mc_ptr, "bad arguments");
  CUdevice dev;
  FEDB_DRV(cuDeviceGet, &dev, device);
  CUmemAllocationProp ap{};
  ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  ap.location.id = device;
  ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t gran = 0;
  FEDB_DRV(cuMemGetAllocationGranularity, &gran, &ap, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED);
  FEDB_REQUIRE((size_t)bytes % gran == 0, "size is not a multiple of the allocation granularity");
  CUmemGenericAllocationHandle mem;
  FEDB_DRV(cuMemCreate, &mem, (size_t)bytes, &ap, 0);
  FEDB_DRV(cuMulticastBindMem, (CUmemGenericAllocationHandle)mc_handle, 0, mem, 0, (size_t)bytes, 0);
  CUmemAccessDesc ad{};
  ad.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  ad.location.id = device;
  ad.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  CUdeviceptr uc = 0, mc = 0;
  FEDB_DRV(cuMemAddressReserve, &uc, (size_t)bytes, gran, 0, 0);
  FEDB_DRV(cuMemMap, uc, (size_t)bytes, 0, mem, 0);
  FEDB_DRV(cuMemSetAccess, uc, (size_t)bytes, &ad, 1);
  FEDB_DRV(cuMemAddressReserve, &mc, (size_t)bytes, gran, 0, 0);
  FEDB_DRV(cuMemMap, mc, (size_t)bytes, 0, (CUmemGenericAllocationHandle)mc_handle, 0);
  FEDB_DRV(cuMemSetAccess, mc, (size_t)bytes, &ad, 1);
  FEDB_CUDA_OK(cudaMemset((void*)uc, 0, (size_t)bytes));
  FEDB_CUDA_OK(cudaDeviceSynchronize());
  *mem_handle_out = (uint64_t)mem;
  *uc_ptr = (void*)uc;
  *mc_ptr = (void*)mc;
  return FEDB_OK;
}

extern "C" int fedb_nvls_release(uint64_t mc_handle, uint64_t mem_handle, void* uc_ptr, void* mc_ptr, int64_t bytes) {
  FEDB_CUDA_OK(cudaDeviceSynchronize());
  if (mc_ptr) {
    FEDB_DRV(cuMemUnmap, (CUdeviceptr)mc_ptr, (size_t)bytes);
    FEDB_DRV(cuMemAddressFree, (CUdeviceptr)mc_ptr, (size_t)bytes);
  }
  if (uc_ptr) {
    FEDB_DRV(cuMemUnmap, (CUdeviceptr)uc_ptr, (size_t)bytes);
    FEDB_DRV(cuMemAddressFree, (CUdeviceptr)uc_ptr, (size_t)bytes);
  }
  if (mem_handle) FEDB_DRV(cuMemRelease, (CUmemGenericAllocationHandle)mem_handle);
  if (mc_handle) FEDB_DRV(cuMemRelease, (CUmemGenericAllocationHandle)mc_handle);
  return FEDB_OK;
}

// hrmem.cu -- device memory for the hr maps with compute data compression.
//
// The hr map (decoder/cif_hr.py:14-24, `self.accumulated`) is the one large buffer of the path: 27.9 MB per
// 641 x 641 image, written once per decode, ~88 % zeros at 5 persons per image, and read afterwards only at
// seed and CAF candidate positions.  B200 can compress such data on its way from L2 to HBM when the
// allocation asks for it (cuMemCreate with CU_MEM_ALLOCATION_COMP_GENERIC); measured with the write pattern
// of the strips kernel (scripts/micro/comp_probe.cu, 1.79 GB): 6.28 TB/s into a plain allocation, 7.50 TB/s
// into a compressible one.  cudaMalloc (and torch's allocator) cannot ask for it, hence these two entry
// points.  The driver API is reached through cudaGetDriverEntryPoint, so the library has no link-time
// dependency on libcuda and still loads on a machine without a driver.
#include <cuda.h>
#include <cuda_runtime.h>

#include <mutex>
#include <unordered_map>

#include "common.cuh"

namespace pp {
namespace {

struct DriverApi {
  CUresult (*DeviceGet)(CUdevice *, int);
  CUresult (*DeviceGetAttribute)(int *, CUdevice_attribute, CUdevice);
  CUresult (*MemGetAllocationGranularity)(size_t *, const CUmemAllocationProp *, CUmemAllocationGranularity_flags);
  CUresult (*MemCreate)(CUmemGenericAllocationHandle *, size_t, const CUmemAllocationProp *, unsigned long long);
  CUresult (*MemGetAllocationPropertiesFromHandle)(CUmemAllocationProp *, CUmemGenericAllocationHandle);
  CUresult (*MemAddressReserve)(CUdeviceptr *, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc *, size_t);
  CUresult (*MemUnmap)(CUdeviceptr, size_t);
  CUresult (*MemAddressFree)(CUdeviceptr, size_t);
  CUresult (*MemRelease)(CUmemGenericAllocationHandle);
  bool ok;
};

template <typename T>
bool entry(const char *name, T *fn) {
  void *p = nullptr;
  cudaDriverEntryPointQueryResult st;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &st) != cudaSuccess || st != cudaDriverEntryPointSuccess ||
      !p) {
    (void)cudaGetLastError();
    return false;
  }
  *fn = reinterpret_cast<T>(p);
  return true;
}

const DriverApi &driver() {
  static DriverApi d = [] {
    DriverApi a = {};
    a.ok = entry("cuDeviceGet", &a.DeviceGet) && entry("cuDeviceGetAttribute", &a.DeviceGetAttribute) &&
           entry("cuMemGetAllocationGranularity", &a.MemGetAllocationGranularity) &&
           entry("cuMemCreate", &a.MemCreate) &&
           entry("cuMemGetAllocationPropertiesFromHandle", &a.MemGetAllocationPropertiesFromHandle) &&
           entry("cuMemAddressReserve", &a.MemAddressReserve) && entry("cuMemMap", &a.MemMap) &&
           entry("cuMemSetAccess", &a.MemSetAccess) && entry("cuMemUnmap", &a.MemUnmap) &&
           entry("cuMemAddressFree", &a.MemAddressFree) && entry("cuMemRelease", &a.MemRelease);
    return a;
  }();
  return d;
}

struct Block {
  CUmemGenericAllocationHandle handle;
  size_t size;
};
std::mutex g_mu;
std::unordered_map<void *, Block> g_blocks;      // what pp_hr_free needs to undo an allocation

}  // namespace
}  // namespace pp

using namespace pp;

extern "C" {

int pp_hr_alloc(size_t bytes, void **d_ptr, int32_t *compressed) {
  if (!d_ptr || bytes == 0) return PP_ERR_ARG;
  *d_ptr = nullptr;
  if (compressed) *compressed = 0;
  int dev = 0;
  PP_CUDA_CHECK(cudaGetDevice(&dev));
  PP_CUDA_CHECK(cudaFree(0));                    // make sure the device's primary context exists
  const DriverApi &d = driver();
  if (!d.ok) return PP_ERR_CUDA;
  CUdevice cu_dev;
  if (d.DeviceGet(&cu_dev, dev) != CUDA_SUCCESS) return PP_ERR_CUDA;
  int can = 0;
  if (d.DeviceGetAttribute(&can, CU_DEVICE_ATTRIBUTE_GENERIC_COMPRESSION_SUPPORTED, cu_dev) != CUDA_SUCCESS) can = 0;
  CUmemAllocationProp prop = {};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = dev;
  prop.allocFlags.compressionType = can ? CU_MEM_ALLOCATION_COMP_GENERIC : CU_MEM_ALLOCATION_COMP_NONE;
  size_t gran = 0;
  if (d.MemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS || gran == 0)
    return PP_ERR_CUDA;
  const size_t size = (bytes + gran - 1) / gran * gran;
  CUmemGenericAllocationHandle h;
  if (d.MemCreate(&h, size, &prop, 0) != CUDA_SUCCESS) return PP_ERR_CUDA;
  CUdeviceptr ptr = 0;
  if (d.MemAddressReserve(&ptr, size, 0, 0, 0) != CUDA_SUCCESS) {
    d.MemRelease(h);
    return PP_ERR_CUDA;
  }
  CUmemAccessDesc acc = {};
  acc.location = prop.location;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  if (d.MemMap(ptr, size, 0, h, 0) != CUDA_SUCCESS || d.MemSetAccess(ptr, size, &acc, 1) != CUDA_SUCCESS) {
    d.MemUnmap(ptr, size);
    d.MemAddressFree(ptr, size);
    d.MemRelease(h);
    return PP_ERR_CUDA;
  }
  if (compressed) {
    CUmemAllocationProp got = {};
    if (d.MemGetAllocationPropertiesFromHandle(&got, h) == CUDA_SUCCESS)
      *compressed = got.allocFlags.compressionType == CU_MEM_ALLOCATION_COMP_GENERIC;
  }
  {
    std::lock_guard<std::mutex> lock(g_mu);
    g_blocks[(void *)ptr] = Block{h, size};
  }
  *d_ptr = (void *)ptr;
  return PP_OK;
}

int pp_hr_free(void *d_ptr) {
  if (!d_ptr) return PP_OK;
  Block b;
  {
    std::lock_guard<std::mutex> lock(g_mu);
    auto it = g_blocks.find(d_ptr);
    if (it == g_blocks.end()) return PP_ERR_ARG;
    b = it->second;
    g_blocks.erase(it);
  }
  const DriverApi &d = driver();
  if (!d.ok) return PP_ERR_CUDA;
  const CUdeviceptr ptr = (CUdeviceptr)d_ptr;
  int rc = PP_OK;
  if (d.MemUnmap(ptr, b.size) != CUDA_SUCCESS) rc = PP_ERR_CUDA;
  if (d.MemAddressFree(ptr, b.size) != CUDA_SUCCESS) rc = PP_ERR_CUDA;
  if (d.MemRelease(b.handle) != CUDA_SUCCESS) rc = PP_ERR_CUDA;
  return rc;
}

}  // extern "C"

// Probe: does compute data compression (cuMemCreate, CU_MEM_ALLOCATION_COMP_GENERIC) speed up the
// zero-dominated hr write stream?  Writes 1.79 GB with (a) cudaMemsetAsync, (b) 128-bit stores,
// (c) TMA bulk stores from a zeroed smem buffer, into a plain and a compressible allocation.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o comp_probe comp_probe.cu -lcuda
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#define CK(x) do { CUresult r_ = (x); if (r_ != CUDA_SUCCESS) { const char *s_; cuGetErrorString(r_, &s_); printf("%s failed: %s\n", #x, s_); exit(1); } } while (0)
#define RK(x) do { cudaError_t r_ = (x); if (r_ != cudaSuccess) { printf("%s failed: %s\n", #x, cudaGetErrorString(r_)); exit(1); } } while (0)

__global__ void store_kernel(float4 *p, size_t n4, float v) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x)
    p[i] = make_float4(v, v, v, v);
}
// mostly zeros: every `period`-th 20 KB strip carries data
__global__ void bulk_kernel(float *p, size_t n, int period) {
  extern __shared__ __align__(128) float buf[];
  const int piece = 5120;   // floats (20 KB)
  for (int i = threadIdx.x; i < 2 * piece; i += blockDim.x) buf[i] = i < piece ? 0.f : 0.001f * (i & 1023);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0) {
    const size_t n_pieces = n / piece;
    for (size_t k = blockIdx.x; k < n_pieces; k += gridDim.x) {
      const float *src = (period && (k % period) == 0) ? buf + piece : buf;
      unsigned s = (unsigned)__cvta_generic_to_shared(src);
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" :: "l"(p + k * piece), "r"(s), "r"(piece * 4) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group.read 8;" ::: "memory");
    }
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
  }
}
__global__ void sum_kernel(const float4 *p, size_t n4, float *out) {
  float s = 0.f;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    float4 v = p[i]; s += v.x + v.y + v.z + v.w;
  }
  if (s != 0.f) atomicAdd(out, s);
}

static float timeit(void (*fn)(void *, size_t), void *p, size_t bytes) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  for (int i = 0; i < 3; i++) fn(p, bytes);
  RK(cudaDeviceSynchronize());
  float best = 1e9f;
  for (int i = 0; i < 5; i++) {
    cudaEventRecord(a); fn(p, bytes); cudaEventRecord(b); RK(cudaEventSynchronize(b));
    float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
  }
  return best;
}
static void f_memset(void *p, size_t bytes) { cudaMemsetAsync(p, 0, bytes, 0); }
static void f_store(void *p, size_t bytes) { store_kernel<<<148 * 8, 256>>>((float4 *)p, bytes / 16, 0.f); }
static void f_store_nz(void *p, size_t bytes) { store_kernel<<<148 * 8, 256>>>((float4 *)p, bytes / 16, 0.37f); }
static void f_bulk(void *p, size_t bytes) { bulk_kernel<<<148 * 2, 64, 2 * 5120 * 4>>>((float *)p, bytes / 4, 0); }
static void f_bulk8(void *p, size_t bytes) { bulk_kernel<<<148 * 2, 64, 2 * 5120 * 4>>>((float *)p, bytes / 4, 8); }
static float *g_out;
static void f_read(void *p, size_t bytes) { sum_kernel<<<148 * 8, 256>>>((const float4 *)p, bytes / 16, g_out); }

int main() {
  RK(cudaFree(0));
  CUdevice dev; CK(cuDeviceGet(&dev, 0));
  int comp = 0; CK(cuDeviceGetAttribute(&comp, CU_DEVICE_ATTRIBUTE_GENERIC_COMPRESSION_SUPPORTED, dev));
  printf("generic compression supported: %d\n", comp);
  RK(cudaMalloc(&g_out, 4));
  RK(cudaFuncSetAttribute(bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 5120 * 4));
  size_t bytes = (size_t)64 * 17 * 641 * 641 * 4;
  CUmemAllocationProp prop = {};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = 0;
  size_t gran = 0; CK(cuMemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  size_t size = (bytes + gran - 1) / gran * gran;
  for (int mode = 0; mode < (comp ? 2 : 1); mode++) {
    prop.allocFlags.compressionType = mode ? CU_MEM_ALLOCATION_COMP_GENERIC : CU_MEM_ALLOCATION_COMP_NONE;
    CUmemGenericAllocationHandle h; CK(cuMemCreate(&h, size, &prop, 0));
    CUmemAllocationProp got = {}; CK(cuMemGetAllocationPropertiesFromHandle(&got, h));
    CUdeviceptr ptr; CK(cuMemAddressReserve(&ptr, size, 0, 0, 0));
    CK(cuMemMap(ptr, size, 0, h, 0));
    CUmemAccessDesc acc = {}; acc.location = prop.location; acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    CK(cuMemSetAccess(ptr, size, &acc, 1));
    printf("%s allocation (granularity %zu MB, compressionType obtained %d):\n", mode ? "compressible" : "plain", gran >> 20,
           (int)got.allocFlags.compressionType);
    struct { const char *name; void (*fn)(void *, size_t); } t[] = {{"memset 0", f_memset}, {"st.128 zeros", f_store},
        {"st.128 nonzero const", f_store_nz}, {"bulk stores, all zero", f_bulk}, {"bulk stores, 1 strip in 8 has data", f_bulk8},
        {"read back (sum)", f_read}};
    for (auto &e : t) {
      float ms = timeit(e.fn, (void *)ptr, bytes);
      printf("  %-36s %.3f ms  %.0f GB/s\n", e.name, ms, bytes / ms * 1e-6);
    }
    CK(cuMemUnmap(ptr, size)); CK(cuMemAddressFree(ptr, size)); CK(cuMemRelease(h));
  }
  return 0;
}

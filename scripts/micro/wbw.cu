// Write-bandwidth ceilings on B200: plain 128-bit stores, cudaMemsetAsync, TMA bulk stores from smem.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__global__ void st_kernel(float4 *p, size_t n4) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x)
    p[i] = make_float4(0.f, 0.f, 0.f, 0.f);
}

template <int CHUNK>
__global__ void bulk_kernel(char *p, size_t bytes) {
  extern __shared__ __align__(128) char zero[];
  for (int i = threadIdx.x * 16; i < CHUNK; i += blockDim.x * 16) *(float4 *)(zero + i) = make_float4(0, 0, 0, 0);
  __syncthreads();
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t s = (uint32_t)__cvta_generic_to_shared(zero);
    size_t nchunks = bytes / CHUNK;
    for (size_t c = blockIdx.x; c < nchunks; c += gridDim.x) {
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(p + c * CHUNK), "r"(s), "r"(CHUNK) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group.read 8;" ::: "memory");
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
}

int main() {
  size_t bytes = (size_t)1788154112 / 20480 * 20480;
  char *p;
  cudaMalloc(&p, bytes);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  float ms;
  for (int grid : {148 * 2, 148 * 4, 148 * 8, 148 * 16}) {
    for (int rep = 0; rep < 3; rep++) st_kernel<<<grid, 256>>>((float4 *)p, bytes / 16);
    cudaEventRecord(e0);
    for (int rep = 0; rep < 10; rep++) st_kernel<<<grid, 256>>>((float4 *)p, bytes / 16);
    cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1);
    printf("st.v4 grid %5d: %.1f GB/s (%.3f ms)\n", grid, bytes * 10 / ms / 1e6, ms / 10);
  }
  for (int rep = 0; rep < 3; rep++) cudaMemsetAsync(p, 0, bytes);
  cudaEventRecord(e0);
  for (int rep = 0; rep < 10; rep++) cudaMemsetAsync(p, 0, bytes);
  cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1);
  printf("cudaMemsetAsync: %.1f GB/s (%.3f ms)\n", bytes * 10 / ms / 1e6, ms / 10);
  for (int grid : {148, 148 * 2, 148 * 4, 148 * 8}) {
    for (int rep = 0; rep < 3; rep++) bulk_kernel<20480><<<grid, 128, 20480>>>(p, bytes);
    cudaEventRecord(e0);
    for (int rep = 0; rep < 10; rep++) bulk_kernel<20480><<<grid, 128, 20480>>>(p, bytes);
    cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1);
    printf("bulk 20KB grid %5d: %.1f GB/s (%.3f ms) %s\n", grid, bytes * 10 / ms / 1e6, ms / 10, cudaGetErrorString(cudaGetLastError()));
  }
  for (int grid : {148 * 2, 148 * 8}) {
    for (int rep = 0; rep < 3; rep++) bulk_kernel<4096><<<grid, 128, 4096>>>(p, bytes);
    cudaEventRecord(e0);
    for (int rep = 0; rep < 10; rep++) bulk_kernel<4096><<<grid, 128, 4096>>>(p, bytes);
    cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1);
    printf("bulk 4KB grid %5d: %.1f GB/s (%.3f ms) %s\n", grid, bytes * 10 / ms / 1e6, ms / 10, cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}

// Probe: host -> device upload of the 48 carried bytes of every 80-byte particle.
//   (a) the whole 80-byte AoS with one cudaMemcpyAsync          (what sphb_upload does)
//   (b) cudaMemcpy2DAsync, width 48, pitch 80 -> 48            (DMA engine, strided)
//   (c) a kernel reading MAPPED pinned memory, 3 x 16 bytes per particle (zero-copy)
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a tools/probes/pcie_strided.cu -o tools/probes/pcie_strided
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s -> %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)
__global__ void k_pull(const float4* __restrict__ host, float4* pos, float4* vel, float4* vhf, unsigned n) {
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float4* p = host + 5ull * i;
    pos[i] = p[0]; vel[i] = p[1]; vhf[i] = p[2];
  }
}
int main(int argc, char** argv) {
  const unsigned n = argc > 1 ? atoi(argv[1]) : 16777216;
  void *h, *d80, *d48; float4 *pos, *vel, *vhf;
  CK(cudaHostAlloc(&h, 80ull * n, cudaHostAllocMapped));
  for (size_t k = 0; k < 80ull * n / 4; k += 1024) ((float*)h)[k] = 1.f;
  CK(cudaMalloc(&d80, 80ull * n)); CK(cudaMalloc(&d48, 48ull * n));
  CK(cudaMalloc(&pos, 16ull * n)); CK(cudaMalloc(&vel, 16ull * n)); CK(cudaMalloc(&vhf, 16ull * n));
  void* hd; CK(cudaHostGetDevicePointer(&hd, h, 0));
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  for (int mode = 0; mode < 3; ++mode)
    for (int rep = 0; rep < 3; ++rep) {
      cudaEventRecord(a);
      if (mode == 0) CK(cudaMemcpyAsync(d80, h, 80ull * n, cudaMemcpyHostToDevice));
      if (mode == 1) CK(cudaMemcpy2DAsync(d48, 48, h, 80, 48, n, cudaMemcpyHostToDevice));
      if (mode == 2) k_pull<<<148 * 8, 256>>>((const float4*)hd, pos, vel, vhf, n);
      cudaEventRecord(b); CK(cudaEventSynchronize(b));
      float ms; cudaEventElapsedTime(&ms, a, b);
      if (rep == 2) printf("mode %d (%s): %.2f ms, %.1f GB/s of the 80-byte array, %.1f GB/s of carried bytes\n", mode,
                           mode == 0 ? "full 80 B" : mode == 1 ? "2D 48 of 80" : "zero-copy kernel 48 of 80", ms,
                           80e-6 * n / ms, 48e-6 * n / ms);
    }
  return 0;
}

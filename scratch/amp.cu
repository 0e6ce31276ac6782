// does "4 x LDG.32 per lane at 16-byte lane stride, shifted by 4 bytes" amplify DRAM reads?
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__global__ void k_shift4(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, size_t nchunks, int shift) {
  for (size_t q = blockIdx.x * (size_t)blockDim.x + threadIdx.x; q < nchunks; q += (size_t)gridDim.x * blockDim.x) {
    const uint32_t* sp = reinterpret_cast<const uint32_t*>(src + q * 16 + shift);
    uint4 v = make_uint4(__ldg(sp), __ldg(sp + 1), __ldg(sp + 2), __ldg(sp + 3));
    *reinterpret_cast<uint4*>(dst + q * 16) = v;
  }
}
__global__ void k_vec(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, size_t nchunks) {
  for (size_t q = blockIdx.x * (size_t)blockDim.x + threadIdx.x; q < nchunks; q += (size_t)gridDim.x * blockDim.x)
    *reinterpret_cast<uint4*>(dst + q * 16) = __ldg(reinterpret_cast<const uint4*>(src + q * 16));
}
int main() {
  size_t n = 1ull << 30; uint8_t *a, *b; cudaMalloc(&a, n + 64); cudaMalloc(&b, n);
  cudaMemset(a, 1, n + 64);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int mode = 0; mode < 4; mode++) {
    for (int rep = 0; rep < 2; rep++) {
      cudaEventRecord(e0);
      if (mode == 0) k_vec<<<148 * 8, 256>>>(a, b, n / 16);
      else k_shift4<<<148 * 8, 256>>>(a, b, n / 16, mode == 1 ? 0 : (mode == 2 ? 4 : 12));
      cudaEventRecord(e1); cudaEventSynchronize(e1);
    }
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("mode %d: %.3f ms  %.1f GB/s (read+write)\n", mode, ms, 2.0 * n / ms / 1e6);
  }
  return 0;
}

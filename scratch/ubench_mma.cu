// micro-benchmark: legacy int8 tensor path on sm_100a - mma.sync.m16n8k32 u8*u8->s32 issue rate per SM,
// alone, with ldmatrix.x4 feeding it, and next to dp4a. Decides whether the fused preprocess kernel's
// horizontal pass (a banded u8 GEMM) is worth moving from dp4a to IMMA.
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 2048
__device__ __forceinline__ void mma_u8(int (&c)[4], const unsigned (&a)[4], const unsigned (&b)[2]) {
  asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}
__device__ __forceinline__ void ldsm4(unsigned (&r)[4], unsigned addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
// MODE 0: NACC independent MMA chains, operands in registers
// MODE 1: one ldmatrix.x4 per NACC MMAs (A from shared memory)
// MODE 2: MMA chains + as many dp4a as MMAs x 8
template <int MODE, int NACC>
__global__ void kmma(int* out, unsigned seed) {
  __shared__ __align__(128) unsigned char sm[16 * 1024];
  for (int i = threadIdx.x; i < 4096; i += blockDim.x) reinterpret_cast<unsigned*>(sm)[i] = i * 2654435761u + seed;
  __syncthreads();
  int c[NACC][4];
  unsigned a[4] = {seed + threadIdx.x, seed * 3u, seed ^ 0x55u, threadIdx.x * 7u}, b[2] = {seed * 5u, threadIdx.x + 1u};
  unsigned d[8];
  for (int i = 0; i < 8; i++) d[i] = threadIdx.x + i;
  for (int j = 0; j < NACC; j++) for (int k = 0; k < 4; k++) c[j][k] = 0;
  const int lane = threadIdx.x & 31;
  unsigned base = (unsigned)__cvta_generic_to_shared(sm) + (lane & 15) * 272 + (lane >> 4) * 16 + (threadIdx.x >> 5) * 32;
  for (int it = 0; it < ITERS; it++) {
    if (MODE == 1) ldsm4(a, base + (it & 7) * 32);
#pragma unroll
    for (int j = 0; j < NACC; j++) mma_u8(c[j], a, b);
    if (MODE == 2) {
#pragma unroll
      for (int i = 0; i < 8; i++) d[i] = __dp4a(a[0], b[0], d[i]);
    }
  }
  int s = 0;
  for (int j = 0; j < NACC; j++) for (int k = 0; k < 4; k++) s += c[j][k];
  for (int i = 0; i < 8; i++) s += d[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <typename F> float timeit(F f) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  f(); cudaDeviceSynchronize();
  cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1); return ms;
}
int main() {
  int* out; cudaMalloc(&out, 148 * 4 * 1024 * 4);
  int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  const int sms = 148;
  auto report = [&](const char* name, float ms, int blocks, int threads, int nacc) {
    double mmas = (double)blocks * (threads / 32) * ITERS * nacc;
    printf("%-34s %.3f ms -> %.3f MMA(16x8x32)/clk/SM = %.0f MAC/clk/SM (nominal %d MHz; real clock may be lower)\n", name, ms,
           mmas / (ms * 1e-3) / sms / (clk * 1e3), mmas * 4096 / (ms * 1e-3) / sms / (clk * 1e3), clk / 1000);
  };
  for (int warps : {4, 8, 16, 28}) {
    int threads = warps * 32, blocks = sms * 2;
    char nm[64];
    snprintf(nm, 64, "mma regs, %d warps x2 CTA, 4 acc", warps);
    report(nm, timeit([&] { kmma<0, 4><<<blocks, threads>>>(out, 1); }), blocks, threads, 4);
    snprintf(nm, 64, "mma regs, %d warps x2 CTA, 8 acc", warps);
    report(nm, timeit([&] { kmma<0, 8><<<blocks, threads>>>(out, 1); }), blocks, threads, 8);
    snprintf(nm, 64, "mma + ldmatrix.x4 per 6, %d warps", warps);
    report(nm, timeit([&] { kmma<1, 6><<<blocks, threads>>>(out, 1); }), blocks, threads, 6);
    snprintf(nm, 64, "mma x4 + 8 dp4a, %d warps", warps);
    report(nm, timeit([&] { kmma<2, 4><<<blocks, threads>>>(out, 1); }), blocks, threads, 4);
  }
  cudaError_t e = cudaDeviceSynchronize();
  printf("status: %s\n", cudaGetErrorString(e));
  return e != cudaSuccess;
}

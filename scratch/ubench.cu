// micro-benchmarks: integer pipe throughputs on sm_100a (dp4a, imad, prmt, mixed)
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 4096
template <int MODE>
__global__ void k(unsigned* out, unsigned seed) {
  unsigned a[8], b = seed + threadIdx.x, c = seed * 3 + 1;
  for (int i = 0; i < 8; i++) a[i] = threadIdx.x + i;
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++) {
      if (MODE == 0) a[i] = __dp4a(b, c, a[i]);
      if (MODE == 1) a[i] = a[i] * b + c;
      if (MODE == 2) a[i] = __byte_perm(a[i], b, c);
      if (MODE == 3) { a[i] = __dp4a(b, c, a[i]); }
      if (MODE == 4) a[i] = __funnelshift_r(a[i], b, c);
      if (MODE == 5) a[i] = a[i] + b;   // IADD3
      if (MODE == 6) { float f = __uint_as_float(a[i]); f = f * __uint_as_float(b) + __uint_as_float(c); a[i] = __float_as_uint(f);}  // FFMA
    }
    if (MODE == 3) {  // mixed: 8 dp4a + 8 prmt on separate chains
      // second set
    }
    b += (MODE == 2 || MODE == 4) ? 0 : 0;
  }
  unsigned s = 0;
  for (int i = 0; i < 8; i++) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// mixed kernel: dp4a chain + prmt chain interleaved
__global__ void kmix(unsigned* out, unsigned seed, int) {
  unsigned a[8], p[8], b = seed + threadIdx.x, c = seed * 3 + 1;
  for (int i = 0; i < 8; i++) { a[i] = threadIdx.x + i; p[i] = threadIdx.x * 7 + i; }
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++) { a[i] = __dp4a(b, c, a[i]); p[i] = __byte_perm(p[i], b, c); }
  }
  unsigned s = 0;
  for (int i = 0; i < 8; i++) s += a[i] + p[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// dp4a fed by shared-memory loads with a strided pattern
__global__ void klds(unsigned* out, int stride_x100, int) {
  __shared__ unsigned sm[8192];
  for (int i = threadIdx.x; i < 8192; i += blockDim.x) sm[i] = i * 2654435761u;
  __syncthreads();
  int base = (threadIdx.x & 31) * stride_x100 / 100 + (threadIdx.x >> 5) * 97;
  unsigned acc = 0;
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++) acc += sm[(base + i + it) & 8191];
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
template <typename F> float timeit(F f) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  f(); cudaDeviceSynchronize();
  cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1); return ms;
}
int main() {
  unsigned* out; cudaMalloc(&out, 148 * 8 * 1024 * 4);
  int sms = 148, blocks = sms * 2, threads = 1024;
  int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  const char* names[] = {"dp4a", "imad", "prmt", "dp4a(b)", "shf", "iadd", "ffma"};
  for (int m = 0; m < 7; m++) {
    float ms = 0;
    if (m == 0) ms = timeit([&] { k<0><<<blocks, threads>>>(out, 1); });
    if (m == 1) ms = timeit([&] { k<1><<<blocks, threads>>>(out, 1); });
    if (m == 2) ms = timeit([&] { k<2><<<blocks, threads>>>(out, 1); });
    if (m == 3) ms = timeit([&] { k<3><<<blocks, threads>>>(out, 1); });
    if (m == 4) ms = timeit([&] { k<4><<<blocks, threads>>>(out, 1); });
    if (m == 5) ms = timeit([&] { k<5><<<blocks, threads>>>(out, 1); });
    if (m == 6) ms = timeit([&] { k<6><<<blocks, threads>>>(out, 1); });
    double ops = (double)blocks * threads * ITERS * 8;
    printf("%-8s %.3f ms  %.1f Gops/s  -> %.1f lane-ops/clk/SM @%d MHz nominal\n", names[m], ms, ops / ms / 1e6,
           ops / (ms * 1e-3) / sms / (clk * 1e3), clk / 1000);
  }
  float ms = timeit([&] { kmix<<<blocks, threads>>>(out, 1, 0); });
  double ops = (double)blocks * threads * ITERS * 16;
  printf("mix dp4a+prmt %.3f ms -> %.1f lane-ops/clk/SM (both counted)\n", ms, ops / (ms * 1e-3) / sms / (clk * 1e3));
  for (int st : {100, 214, 300, 643, 800}) {
    ms = timeit([&] { klds<<<blocks, threads>>>(out, st, 0); });
    ops = (double)blocks * threads * ITERS * 8;
    printf("lds stride %.2f words: %.3f ms -> %.1f lane-loads/clk/SM\n", st / 100.0, ms, ops / (ms * 1e-3) / sms / (clk * 1e3));
  }
  return 0;
}

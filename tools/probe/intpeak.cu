// INT32 issue-rate probe: independent IADD3 / LOP3 chains, 8 per thread; reports warp-instructions/s.
#include <cuda_runtime.h>
#include <stdio.h>
template <int MODE>
__global__ void __launch_bounds__(256) k(unsigned *out, int iters, unsigned seed) {
  unsigned a[8];
#pragma unroll
  for (int j = 0; j < 8; j++) a[j] = seed + threadIdx.x * 8 + j;
  unsigned b = seed * 3 + 1, c = seed ^ 0x55aa;
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int u = 0; u < 16; u++) {
#pragma unroll
      for (int j = 0; j < 8; j++) {
        if (MODE == 0) asm volatile("add.u32 %0, %0, %1;" : "+r"(a[j]) : "r"(b));
        else if (MODE == 1) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[j]) : "r"(b), "r"(c));
        else { if (u & 1) asm volatile("add.u32 %0, %0, %1;" : "+r"(a[j]) : "r"(b)); else asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[j]) : "r"(b), "r"(c)); }
      }
    }
  }
  unsigned s = 0;
#pragma unroll
  for (int j = 0; j < 8; j++) s ^= a[j];
  if (s == 0x12345678u) out[0] = s;
}
template <int MODE> void run(const char *name, int sms) {
  unsigned *out; cudaMalloc(&out, 4);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 2000, blocks = sms * 8;
  for (int rep = 0; rep < 3; rep++) {
    cudaEventRecord(e0);
    k<MODE><<<blocks, 256>>>(out, iters, 7u + rep);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double winst = (double)blocks * 8 /*warps*/ * iters * 16.0 * 8.0;
    printf("%s rep %d: %.3f ms  %.2f G warp-inst/s  = %.2f T thread-op/s  (per SM per clk @1.965GHz: %.2f warp-inst)\n", name, rep, ms,
           winst / ms / 1e6, winst * 32 / ms / 1e9, winst / ms / 1e6 / sms / 1.965);
  }
}
int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  printf("%s SMs %d clock %d kHz\n", p.name, p.multiProcessorCount, p.clockRate);
  run<0>("IADD", p.multiProcessorCount); run<1>("LOP3", p.multiProcessorCount); run<2>("MIX ", p.multiProcessorCount);
  return 0;
}

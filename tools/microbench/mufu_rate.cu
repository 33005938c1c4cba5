// MUFU throughput probe: ex2.approx.ftz.f32 vs ex2.approx.f16x2 (two MUFU.EX2.F16) vs ex2.approx.ftz.bf16x2, results per
// clock per SM.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mufu_rate mufu_rate.cu ; run on the B200.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int MODE>
__global__ void k(uint32_t* out, uint32_t seed, int iters) {
  uint32_t a[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = seed + threadIdx.x * 8 + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (MODE == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+r"(a[i]));
      if (MODE == 1) asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(a[i]));
      if (MODE == 2) asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(a[i]));
    }
  }
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s ^= a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
void run(const char* name, int elems_per_op) {
  uint32_t* out;
  cudaMalloc(&out, 148 * 4 * 1024 * 4);
  const int iters = 4096;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  k<MODE><<<148 * 4, 1024>>>(out, 1, 16);
  cudaEventRecord(e0);
  k<MODE><<<148 * 4, 1024>>>(out, 1, iters);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  int clk_khz;
  cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
  const double ops = 148.0 * 4 * 1024 * iters * 8 * elems_per_op;
  printf("%-28s %8.3f ms  %7.2f results/clk/SM at %d MHz nominal (%.1f G results/s)\n", name, ms,
         ops / (ms * 1e-3) / 148.0 / (clk_khz * 1e3), clk_khz / 1000, ops / (ms * 1e-3) / 1e9);
  cudaFree(out);
}

int main() {
  run<0>("ex2.approx.ftz.f32", 1);
  run<1>("ex2.approx.f16x2", 2);
  run<2>("ex2.approx.ftz.bf16x2", 2);
  return 0;
}

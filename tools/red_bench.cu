// Dev microbenchmark: how fast can the chip take fp32 vector reductions (red.global.add.v4.f32) into a [4][k][d]
// table, the access pattern of the Lloyd accumulation?  One warp adds one 512-byte row per instruction to a
// pseudo-random row.   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/red_bench tools/red_bench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__global__ void k_red(float* table, int rows, int d, long iters_per_warp, int mode) {
  const int lane = threadIdx.x & 31;
  const long warp = (blockIdx.x * (long)blockDim.x + threadIdx.x) >> 5;
  uint32_t s = (uint32_t)warp * 2654435761u + 12345u;
  float4 v = make_float4(1.f, 2.f, 3.f, 4.f);
  for (long i = 0; i < iters_per_warp; ++i) {
    s = s * 1664525u + 1013904223u;
    float* p = table + (size_t)((s >> 8) % (uint32_t)rows) * d + lane * 4;
    if (mode == 0) asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
    else { atomicAdd(p, v.x); atomicAdd(p + 1, v.y); atomicAdd(p + 2, v.z); atomicAdd(p + 3, v.w); }
  }
}
int main() {
  const int rows = 4 * 1024, d = 128;
  float* t;
  cudaMalloc(&t, (size_t)rows * d * 4);
  cudaMemset(t, 0, (size_t)rows * d * 4);
  for (int mode = 0; mode < 2; ++mode)
    for (int wpb : {4, 8, 16}) {
      const int grid = 148;
      const long iters = 20000;
      cudaEvent_t a, b;
      cudaEventCreate(&a); cudaEventCreate(&b);
      k_red<<<grid, wpb * 32>>>(t, rows, d, 1000, mode);
      cudaEventRecord(a);
      k_red<<<grid, wpb * 32>>>(t, rows, d, iters, mode);
      cudaEventRecord(b);
      cudaEventSynchronize(b);
      float ms;
      cudaEventElapsedTime(&ms, a, b);
      const double bytes = (double)grid * wpb * iters * 512.0;
      printf("mode %d (%s) warps/SM %2d: %.3f ms  %.1f GB/s of reduction payload\n", mode, mode ? "4 scalar atomicAdd" : "red.v4.f32", wpb, ms, bytes / ms / 1e6);
    }
  return 0;
}

// Microbenchmark: fp32 FMA throughput on sm_100a with scalar FFMA vs packed FFMA2 (fma.rn.f32x2).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma2_bench ffma2_bench.cu
#include <cuda_runtime.h>
#include <cstdio>
__device__ __forceinline__ void fma2(float2& d, float2 a, float2 b) {
  unsigned long long da = *reinterpret_cast<unsigned long long*>(&d);
  asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(da) : "l"(*reinterpret_cast<unsigned long long*>(&a)),
      "l"(*reinterpret_cast<unsigned long long*>(&b)));
  d = *reinterpret_cast<float2*>(&da);
}
constexpr int NACC = 16;   // independent float2 accumulators per thread
template <bool PACKED>
__global__ void __launch_bounds__(256) k(float* out, float a0, float b0, int iters) {
  float2 acc[NACC];
  for (int i = 0; i < NACC; ++i) acc[i] = make_float2(threadIdx.x * 1e-3f + i, i * 0.5f);
  float2 a = make_float2(a0, a0 * 1.0001f), b = make_float2(b0, b0 * 0.9999f);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) {
      if (PACKED) fma2(acc[i], a, b);
      else { acc[i].x = fmaf(acc[i].x, a.x, b.x); acc[i].y = fmaf(acc[i].y, a.y, b.y); }
    }
  }
  float s = 0.f;
  for (int i = 0; i < NACC; ++i) s += acc[i].x + acc[i].y;
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// packed with one shared-memory operand per 2 FFMA2 (LDS pressure like a conv inner loop)
int main() {
  float* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 20000;
  for (int packed = 0; packed < 2; ++packed) {
    for (int rep = 0; rep < 3; ++rep) {
      cudaEventRecord(e0);
      if (packed) k<true><<<148 * 8, 256>>>(out, 1.0001f, 0.5f, iters);
      else k<false><<<148 * 8, 256>>>(out, 1.0001f, 0.5f, iters);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      double flops = 2.0 * 2 * NACC * (double)iters * 148 * 8 * 256;
      printf("%s rep%d: %.3f ms  %.1f TFLOP/s\n", packed ? "FFMA2" : "FFMA ", rep, ms, flops / ms * 1e-9);
    }
  }
  return 0;
}

// FFMA vs FFMA2 (fma.rn.f32x2) issue throughput on sm_100a: 8 independent accumulator chains per thread.
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ unsigned long long pk(float a, float b) { unsigned long long r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
  unsigned long long r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
template <int MODE>
__global__ void k(float* out, int iters, float s) {
  if (MODE == 0) {
    float a[16];
    for (int i = 0; i < 16; ++i) a[i] = threadIdx.x * 0.001f + i;
    for (int it = 0; it < iters; ++it)
#pragma unroll
      for (int i = 0; i < 16; ++i) a[i] = fmaf(a[i], s, 0.5f);
    float r = 0; for (int i = 0; i < 16; ++i) r += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
  } else {
    unsigned long long a[8];
    const unsigned long long ss = pk(s, s), cc = pk(0.5f, 0.5f);
    for (int i = 0; i < 8; ++i) a[i] = pk(threadIdx.x * 0.001f + i, threadIdx.x * 0.002f + i);
    for (int it = 0; it < iters; ++it)
#pragma unroll
      for (int i = 0; i < 8; ++i) a[i] = fma2(a[i], ss, cc);
    float r = 0;
    for (int i = 0; i < 8; ++i) { float x, y; asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(a[i])); r += x + y; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
  }
}
int main() {
  float* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
  for (int mode = 0; mode < 2; ++mode) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    const int iters = 20000;
    for (int rep = 0; rep < 2; ++rep) {
      cudaEventRecord(a);
      if (mode == 0) k<0><<<148 * 8, 256>>>(d, iters, 0.999f); else k<1><<<148 * 8, 256>>>(d, iters, 0.999f);
      cudaEventRecord(b); cudaEventSynchronize(b);
    }
    float ms; cudaEventElapsedTime(&ms, a, b);
    const double fma = 148.0 * 8 * 256 * (double)iters * 16;
    printf("%s: %.3f ms, %.1f TFMA/s (%.1f TFLOP/s)\n", mode ? "FFMA2" : "FFMA ", ms, fma / ms / 1e9, 2 * fma / ms / 1e9);
  }
  return 0;
}

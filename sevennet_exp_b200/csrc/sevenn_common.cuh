// Common glue for the hand-written sm_100a kernels.
// -DSEVENN_EMU (tests only, g++) turns 1-D grid-stride kernels into serial loops so that the
// element arithmetic can be checked against the oracle on a box without a GPU.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cmath>

#ifndef SEVENN_EMU
#include <cuda_runtime.h>
#define HD __host__ __device__ __forceinline__
#define GKERNEL(name) __global__ void name
#define GRID_STRIDE_LOOP(i, n)                                                          \
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < (long long)(n); \
       i += (long long)gridDim.x * blockDim.x)
static inline int sevenn_grid_for(long long n, int block = 256) {
  long long g = (n + block - 1) / block;
  const long long cap = 148LL * 32;   // 148 SMs x 32 resident CTAs; grid-stride covers the rest
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (int)g;
}
#define LAUNCH1D(kern, n, stream) kern<<<sevenn_grid_for(n), 256, 0, (cudaStream_t)(stream)>>>
#define LAUNCH_STATUS() ((int)cudaGetLastError())
#define ATOMIC_ADD_I32(p, v) atomicAdd((p), (v))
#define ATOMIC_ADD_F32(p, v) atomicAdd((p), (v))
#define LDGF(p) __ldg(p)
#else
#define HD static inline
#define GKERNEL(name) static void name
#define GRID_STRIDE_LOOP(i, n) for (long long i = 0; i < (long long)(n); ++i)
#define LAUNCH1D(kern, n, stream) kern
#define LAUNCH_STATUS() 0
static inline int sevenn_emu_atomic_add(int* p, int v) { int o = *p; *p += v; return o; }
#define ATOMIC_ADD_I32(p, v) sevenn_emu_atomic_add((p), (v))
#define ATOMIC_ADD_F32(p, v) (*(p) += (v))
#define LDGF(p) (*(p))
#define __restrict__
#endif

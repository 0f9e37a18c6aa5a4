// Flavor glue for the generated convolution kernels (codegen/conv_codegen.py).
//   CUDA   : one warp per destination row, 4 warps per CTA.
//   -DSEVENN_EMU (tests only, g++): a "warp" is a 32-iteration lane loop.
#pragma once
#include <cstddef>
#include <cmath>

#ifndef SEVENN_EMU
#include <cuda_runtime.h>
#define CONV_WARPS_PER_CTA 4
#define KERNEL_HEAD(name) __global__ void __launch_bounds__(32 * CONV_WARPS_PER_CTA) name
#define WARP_LOOP_BEGIN(n) { const int lane = threadIdx.x & 31; \
    const int row = blockIdx.x * CONV_WARPS_PER_CTA + (threadIdx.x >> 5); if (row < (n)) {
#define WARP_LOOP_END }}
#define LDG(p) __ldg(p)
__device__ __forceinline__ float sevenn_warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// dY[e, j] is owned by exactly one warp at a time (one row owns the edge, groups and layers
// are stream-ordered launches), so a plain read-modify-write by lane 0 is race free.
#define WARP_SUM_ADD(ptr, v) { const float s_ = sevenn_warp_sum(v); if (lane == 0) *(ptr) += s_; }
#define LAUNCH(kern, n, stream) kern<<<((n) + CONV_WARPS_PER_CTA - 1) / CONV_WARPS_PER_CTA, 32 * CONV_WARPS_PER_CTA, 0, (cudaStream_t)(stream)>>>
#define LAUNCH_STATUS() ((int)cudaGetLastError())
#else
#define KERNEL_HEAD(name) static void name
#define WARP_LOOP_BEGIN(n) { for (int row = 0; row < (n); ++row) for (int lane = 0; lane < 32; ++lane) {
#define WARP_LOOP_END }}
#define LDG(p) (*(p))
#define WARP_SUM_ADD(ptr, v) { *(ptr) += (v); }
#define LAUNCH(kern, n, stream) kern
#define LAUNCH_STATUS() 0
#define __restrict__
#endif

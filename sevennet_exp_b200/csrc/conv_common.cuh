// Flavor glue for the generated convolution kernels (codegen/conv_codegen.py).
//   CUDA   : one warp per (destination row, unit), 4 rows per CTA, blockIdx.y = unit.
//   -DSEVENN_EMU (tests only, g++): a "warp" is a 32-iteration lane loop.
#pragma once
#include <cstddef>
#include <cmath>

// warp-uniform scalar as both halves of a packed pair (a free broadcast operand of FFMA2 / FMUL2)
#define F2B(v) make_float2((v), (v))
#ifndef SEVENN_EMU
#include <cuda_runtime.h>
#define CONV_WARPS_PER_CTA 4
// minctas: occupancy target handed to ptxas (register cap = 65536 / (128 * minctas)); chosen per kernel by the generator
#define KERNEL_HEAD(name, minctas) __global__ void __launch_bounds__(32 * CONV_WARPS_PER_CTA, minctas) name
#define UNIT_LOOP_BEGIN(n, nunits, ringf) { __shared__ float ring_all[CONV_WARPS_PER_CTA * (ringf)]; \
    float* ring = ring_all + (threadIdx.x >> 5) * (ringf); (void)ring; \
    const int lane = threadIdx.x & 31; const int unit = blockIdx.y; \
    const int row = blockIdx.x * CONV_WARPS_PER_CTA + (threadIdx.x >> 5); if (row < (n)) {
#define UNIT_LOOP_END }}
#define LDG(p) __ldg(p)
#define F2I(v) __float_as_int(v)
// per-lane 4-byte asynchronous copy global -> shared (LDGSTS): the gathered source row, the radial weights and
// the spherical harmonics of the next edges stream into a per-warp ring while the current edge is computed
#define CP_ASYNC4(sdst, gsrc) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" \
    :: "r"((unsigned)__cvta_generic_to_shared(sdst)), "l"(gsrc) : "memory")
#define CP_ASYNC16(sdst, gsrc) asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" \
    :: "r"((unsigned)__cvta_generic_to_shared(sdst)), "l"(gsrc) : "memory")
#define CP_ASYNC_COMMIT() asm volatile("cp.async.commit_group;" ::: "memory")
#define CP_ASYNC_WAIT(n) asm volatile("cp.async.wait_group %0;" :: "n"(n) : "memory")
// Transposing butterfly: every lane holds NP partial values; after log2(NP) exchange steps (NP-1 shuffles
// instead of 5*NP) lane L holds one value index (L >> (5 - log2 NP)) summed over a lane subset, the remaining
// offsets finish the sum.  dst[q] is owned by this warp alone (one slab per unit, one row per edge), so the
// read-modify-write is race free and the order of additions is fixed.
template <int NP>
__device__ __forceinline__ void sevenn_dy_accumulate(float (&v)[NP], int lane, float* dst, int S) {
  int o = 16;
#pragma unroll
  for (int h = NP / 2; h >= 1; h >>= 1, o >>= 1) {
    const bool up = (lane & o) != 0;
#pragma unroll
    for (int i = 0; i < h; ++i) {
      const float send = up ? v[i] : v[i + h];
      const float keep = up ? v[i + h] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, o);
    }
  }
  float s = v[0];
  for (; o >= 1; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  constexpr int SH = (NP == 32) ? 0 : (NP == 16) ? 1 : (NP == 8) ? 2 : (NP == 4) ? 3 : (NP == 2) ? 4 : 5;
  const int q = lane >> SH;
  if ((lane & ((1 << SH) - 1)) == 0 && q < S) dst[q] += s;
}
// Same reduction through shared memory (fewer issue slots than the butterfly: NV stores + 32 / PARTS loads and adds +
// PARTS - 1 shuffles instead of 4 (NP - 1) select / shuffle / add instructions).  Every lane stores its NV partials into
// a [NV][33] scratch (lane-major rows: conflict free); PARTS = min(32 / NV, 4) adjacent lanes share one value index and add up
// PER consecutive lanes' entries each (bank = q + part * PER + t: conflict free for PARTS in {2, 3, 4}); column 32 of
// every row is zero (sevenn_dy_scratch_init) so the last part may read one slot past lane 31.  Fixed order of
// additions; the caller's __syncwarp() at the top of the edge loop orders these reads before the next edge's stores.
template <int NV>
__device__ __forceinline__ void sevenn_dy_scratch_init(float* scratch, int lane) {
  if (lane < NV) scratch[lane * 33 + 32] = 0.f;
  __syncwarp();
}
template <int NV>
__device__ __forceinline__ void sevenn_dy_accumulate_smem(const float* v, int lane, float* scratch, float* dst) {
  constexpr int PARTS = (32 / NV) < 4 ? (32 / NV) : 4;
  constexpr int PER = (32 + PARTS - 1) / PARTS;
  static_assert(PARTS >= 2 && PARTS * PER <= 33, "value count not supported by the shared-memory reduction");
#pragma unroll
  for (int q = 0; q < NV; ++q) scratch[q * 33 + lane] = v[q];
  __syncwarp();
  const int q = lane / PARTS, part = lane - q * PARTS;
  float s = 0.f;
  if (q < NV) {
    const float* colp = scratch + q * 33 + part * PER;
#pragma unroll
    for (int t = 0; t < PER; ++t) s += colp[t];
  }
  float tot = s;
#pragma unroll
  for (int o = 1; o < PARTS; ++o) tot += __shfl_down_sync(0xffffffffu, s, o);   // s0 + s1 (+ s2 (+ s3)) on part 0
  if (part == 0 && q < NV) dst[q] += tot;
}
// radial term of the tabulated kernels: warp sum of the lane partials, added to one word owned by this warp
__device__ __forceinline__ void sevenn_gr_accumulate(float v, int lane, float* dst) {
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if (lane == 0) *dst += v;
}
// tabulated-radial kernels: the ring lives in dynamic shared memory (it can exceed the 48 KB static limit)
#define UNIT_LOOP_BEGIN_DYN(n, nunits, ringf) { extern __shared__ __align__(16) float ring_all[]; \
    float* ring = ring_all + (threadIdx.x >> 5) * (ringf); (void)ring; \
    const int lane = threadIdx.x & 31; const int unit = blockIdx.y; \
    const int row = blockIdx.x * CONV_WARPS_PER_CTA + (threadIdx.x >> 5); if (row < (n)) {
// carve: preferred shared-memory carveout in percent of the SM's 228 KB (just enough for the resident CTAs' rings, so the
// rest of the array serves as L1 for the radial-table rows), or < 0 to leave the driver's default
#define LAUNCH_DYN(kern, n, nunits, ringf, carve, stream) \
    { static bool attr_done = false; const int smem_bytes = CONV_WARPS_PER_CTA * (ringf) * 4; \
      if (!attr_done) { \
        if (smem_bytes > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes); \
        if ((carve) >= 0) cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, (carve)); \
        attr_done = true; } } \
    kern<<<dim3(((n) + CONV_WARPS_PER_CTA - 1) / CONV_WARPS_PER_CTA, (nunits)), 32 * CONV_WARPS_PER_CTA, \
           CONV_WARPS_PER_CTA * (ringf) * 4, (cudaStream_t)(stream)>>>
#define LAUNCH(kern, n, nunits, stream) kern<<<dim3(((n) + CONV_WARPS_PER_CTA - 1) / CONV_WARPS_PER_CTA, (nunits)), 32 * CONV_WARPS_PER_CTA, 0, (cudaStream_t)(stream)>>>
#define LAUNCH_STATUS() ((int)cudaGetLastError())
#else
#define KERNEL_HEAD(name, minctas) static void name
#define UNIT_LOOP_BEGIN(n, nunits, ringf) { for (int unit = 0; unit < (nunits); ++unit) for (int row = 0; row < (n); ++row) \
    for (int lane = 0; lane < 32; ++lane) { float* ring = nullptr; (void)ring;
#define UNIT_LOOP_END }}
#define LDG(p) (*(p))
#include <cstring>
static inline int F2I(float v) { int i; std::memcpy(&i, &v, 4); return i; }
template <int NP>
static inline void sevenn_dy_accumulate(float (&v)[NP], int lane, float* dst, int S) {
  (void)lane;
  for (int q = 0; q < S; ++q) dst[q] += v[q];
}
static inline void sevenn_gr_accumulate(float v, int lane, float* dst) { (void)lane; *dst += v; }
template <int NV> static inline void sevenn_dy_scratch_init(float*, int) {}
template <int NV>
static inline void sevenn_dy_accumulate_smem(const float* v, int lane, float*, float* dst) {
  (void)lane;
  for (int q = 0; q < NV; ++q) dst[q] += v[q];
}
struct float4 { float x, y, z, w; };
// packed fp32 pairs (fma.rn.f32x2 on the device): component-wise single roundings, like the instruction
struct float2 { float x, y; };
static inline float2 make_float2(float x, float y) { float2 r = {x, y}; return r; }
static inline float2 __ffma2_rn(float2 a, float2 b, float2 c) { return make_float2(fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y)); }
static inline float2 __fmul2_rn(float2 a, float2 b) { return make_float2(a.x * b.x, a.y * b.y); }
#define UNIT_LOOP_BEGIN_DYN(n, nunits, ringf) UNIT_LOOP_BEGIN(n, nunits, ringf)
#define LAUNCH_DYN(kern, n, nunits, ringf, carve, stream) kern
#define LAUNCH(kern, n, nunits, stream) kern
#define LAUNCH_STATUS() 0
#define __restrict__
#endif

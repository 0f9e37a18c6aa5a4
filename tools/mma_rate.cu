// Micro-benchmark: issue rate of tcgen05.mma (M=128, cta_group::1) for kind::tf32 / kind::f16 (bf16),
// N in {32,64,128,256}, A operand from shared memory (SS) or tensor memory (TS).  One CTA per SM.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_rate tools/mma_rate.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3FFFF) >> 4) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
template <int KIND, int TS>
__device__ __forceinline__ void mma(uint32_t d, uint64_t adesc, uint32_t a_t, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  if (TS) {
    if (KIND == 0)
      asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
                   ::"r"(d), "r"(a_t), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
    else
      asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
                   ::"r"(d), "r"(a_t), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
  } else {
    if (KIND == 0)
      asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
                   ::"r"(d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
    else
      asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                   ::"r"(d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
  }
}

template <int KIND, int TS>
__global__ void __launch_bounds__(128, 1) rate_kernel(int N, int iters, long long* out) {
  extern __shared__ __align__(1024) uint8_t raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  uint8_t* A = smem;                  // 128 rows x 128 B
  uint8_t* B = smem + 16384;          // 256 rows x 128 B
  uint64_t* bar = (uint64_t*)(smem + 16384 + 32768);
  uint32_t* slot = (uint32_t*)(bar + 1);
  for (int i = threadIdx.x; i < (16384 + 32768) / 4; i += 128) ((uint32_t*)smem)[i] = 0;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tm = *slot;
  // kind::tf32: a/b format 2 ; kind::f16 with bf16: a/b format 1 ; D fp32
  const uint32_t fmt = KIND == 0 ? 2u : 1u;
  const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  if (threadIdx.x == 0) {
    const uint64_t ad = make_desc(smem_u32(A)), bd = make_desc(smem_u32(B));
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
      for (int j = 0; j < 4; ++j)
        mma<KIND, TS>(tm, ad + 2 * j, tm + 256 + 8 * j, bd + 2 * j, idesc, (i | j) != 0);
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
    long long t1 = clock64();
    asm volatile(
        "{\n\t.reg .pred p;\n\tWAIT_LOOP:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n\t@p bra DONE;\n\tbra WAIT_LOOP;\n\tDONE:\n\t}"
        ::"r"(smem_u32(bar)) : "memory");
    long long t2 = clock64();
    if (blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tm) : "memory");
}

template <int KIND, int TS>
void run(const char* name, int ctas) {
  long long* d; cudaMalloc(&d, 16);
  cudaFuncSetAttribute(rate_kernel<KIND, TS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  for (int N : {32, 64, 128, 256}) {
    const int iters = 2000;
    rate_kernel<KIND, TS><<<ctas, 128, 60 * 1024>>>(N, iters, d);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[2]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
    const double per = (double)h[1] / (iters * 4.0);
    const int kk = KIND == 0 ? 8 : 16;
    printf("%-14s ctas=%3d N=%3d: %7.1f clk/MMA (issue %6.1f)  -> %7.0f MAC/clk/SM  %s\n", name, ctas, N, per,
           (double)h[0] / (iters * 4.0), 128.0 * N * kk / per, e == cudaSuccess ? "" : cudaGetErrorString(e));
  }
  cudaFree(d);
}

template <int TS>
__global__ void __launch_bounds__(128, 1) commit_kernel(int N, int groups, int per_group, long long* out) {
  extern __shared__ __align__(1024) uint8_t raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  uint8_t* A = smem; uint8_t* B = smem + 16384;
  uint64_t* bar = (uint64_t*)(smem + 16384 + 32768);
  uint32_t* slot = (uint32_t*)(bar + 9);
  for (int i = threadIdx.x; i < (16384 + 32768) / 4; i += 128) ((uint32_t*)smem)[i] = 0;
  if (threadIdx.x == 0) {
    for (int k = 0; k < 9; ++k) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar + k)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tm = *slot;
  const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  if (threadIdx.x == 0) {
    const uint64_t ad = make_desc(smem_u32(A)), bd = make_desc(smem_u32(B));
    long long t0 = clock64();
    for (int gidx = 0; gidx < groups; ++gidx) {
      for (int i = 0; i < per_group; ++i) {
        const int j = i & 3;
        mma<0, TS>(tm, ad + 2 * j, tm + 256 + 8 * j + 64 * (gidx % 3), bd + 2 * j, idesc, (gidx | i) != 0);
      }
      // arrive::one on a rotating barrier nobody waits on (phase flips freely)
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar + 1 + (gidx & 7))) : "memory");
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
    long long t1 = clock64();
    asm volatile(
        "{\n\t.reg .pred p;\n\tWAIT_LOOP2:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n\t@p bra DONE2;\n\tbra WAIT_LOOP2;\n\tDONE2:\n\t}"
        ::"r"(smem_u32(bar)) : "memory");
    long long t2 = clock64();
    if (blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tm) : "memory");
}

template <int TS>
void run_commit(const char* name) {
  long long* d; cudaMalloc(&d, 16);
  cudaFuncSetAttribute(commit_kernel<TS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  for (int N : {64, 128})
    for (int per : {4, 12, 24, 48}) {
      const int groups = 960 / per;
      commit_kernel<TS><<<148, 128, 60 * 1024>>>(N, groups, per, d);
      cudaError_t e = cudaDeviceSynchronize();
      long long h[2]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
      printf("%-10s N=%3d commit every %2d MMAs: %7.1f clk/MMA (issue %6.1f) %s\n", name, N, per,
             (double)h[1] / (groups * per), (double)h[0] / (groups * per), e == cudaSuccess ? "" : cudaGetErrorString(e));
    }
  cudaFree(d);
}

// two (or four) issuing warps, each with its own accumulator columns and its own completion barrier
template <int TS>
__global__ void __launch_bounds__(128, 1) multi_issue_kernel(int N, int iters, int nwarps, long long* out) {
  extern __shared__ __align__(1024) uint8_t raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  uint8_t* A = smem; uint8_t* B = smem + 16384;
  uint64_t* bar = (uint64_t*)(smem + 16384 + 32768);
  uint32_t* slot = (uint32_t*)(bar + 9);
  for (int i = threadIdx.x; i < (16384 + 32768) / 4; i += 128) ((uint32_t*)smem)[i] = 0;
  if (threadIdx.x == 0) {
    for (int k = 0; k < 9; ++k) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar + k)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tm = *slot;
  const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  const int w = threadIdx.x >> 5;
  if ((threadIdx.x & 31) == 0 && w < nwarps) {
    const uint64_t ad = make_desc(smem_u32(A)), bd = make_desc(smem_u32(B));
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
      for (int j = 0; j < 4; ++j)
        mma<0, TS>(tm + w * 64, ad + 2 * j, tm + 256 + 8 * j + 32 * w, bd + 2 * j, idesc, (i | j) != 0);
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar + w)) : "memory");
    long long t1 = clock64();
    asm volatile(
        "{\n\t.reg .pred p;\n\tWAIT_LOOP3:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n\t@p bra DONE3;\n\tbra WAIT_LOOP3;\n\tDONE3:\n\t}"
        ::"r"(smem_u32(bar + w)) : "memory");
    long long t2 = clock64();
    if (blockIdx.x == 0 && w == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tm) : "memory");
}

template <int VAR>
__global__ void __launch_bounds__(448, 1) insitu_kernel(int N, int iters, int nwarps, long long* out) {
  // VAR bit0: nonzero operands; bit1: rotate B over 8 distinct 2 KB blocks x 4 stages; bit2: rotate A TMEM columns
  extern __shared__ __align__(1024) uint8_t raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  uint8_t* B = smem;                                  // 4 stages x (hi|lo) x 64 rows x 128 B = 64 KB
  uint64_t* bar = (uint64_t*)(smem + 65536);
  uint32_t* slot = (uint32_t*)(bar + 9);
  for (int i = threadIdx.x; i < 65536 / 4; i += blockDim.x)
    ((float*)smem)[i] = (VAR & 1) ? (float)((i * 2654435761u) >> 8) * (1.0f / 16777216.0f) - 0.5f : 0.f;
  if (threadIdx.x == 0) {
    for (int k = 0; k < 9; ++k) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar + k)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tm = *slot;
  if ((VAR & 1) && threadIdx.x < 128) {   // fill the A region of TMEM (columns 256..511) with nonzero tf32 values
    uint32_t v[32];
    for (int c0 = 256; c0 < 512; c0 += 32) {
      for (int q = 0; q < 32; ++q) v[q] = __float_as_uint((float)(((threadIdx.x * 131 + c0 + q) * 2654435761u) >> 8) * (1.0f / 16777216.0f) - 0.5f) & 0xFFFFE000u;
      asm volatile(
          "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
          "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
          "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
          ::"r"(tm + ((uint32_t)((threadIdx.x >> 5) * 32) << 16) + c0),
            "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
            "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]),
            "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]),
            "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
          : "memory");
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  }
  if (VAR & 1) {
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  }
  const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  const int w = threadIdx.x >> 5;
  if ((threadIdx.x & 31) == 0 && w < nwarps) {
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      const int st = (VAR & 2) ? (i & 3) : 0;
      const uint64_t bhi = make_desc(smem_u32(B + st * 16384)), blo = make_desc(smem_u32(B + st * 16384 + 8192));
      const uint32_t ahi = tm + 256 + ((VAR & 4) ? 64 * (i & 3) : 0), alo = ahi + 32;
#pragma unroll
      for (int jj = 0; jj < 2; ++jj) {
        const int j = 2 * w + jj;
        mma<0, 1>(tm + w * 64, 0, alo + 8 * j, bhi + 2 * j, idesc, (i | jj) != 0);
        mma<0, 1>(tm + w * 64, 0, ahi + 8 * j, blo + 2 * j, idesc, 1);
        mma<0, 1>(tm + w * 64, 0, ahi + 8 * j, bhi + 2 * j, idesc, 1);
      }
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar + w)) : "memory");
    long long t1 = clock64();
    asm volatile(
        "{\n\t.reg .pred p;\n\tWAIT_LOOP4:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n\t@p bra DONE4;\n\tbra WAIT_LOOP4;\n\tDONE4:\n\t}"
        ::"r"(smem_u32(bar + w)) : "memory");
    long long t2 = clock64();
    if (blockIdx.x == 0 && w == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
    if (w == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar + 8)) : "memory");
  } else if (w >= 2) {
    // spinners: the other warps of a GEMM CTA waiting on their pipeline barriers
    asm volatile(
        "{\n\t.reg .pred p;\n\tWAIT_LOOP5:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n\t@p bra DONE5;\n\tbra WAIT_LOOP5;\n\tDONE5:\n\t}"
        ::"r"(smem_u32(bar + 8)) : "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tm) : "memory");
}

template <int VAR>
void run_insitu() {
  long long* d; cudaMalloc(&d, 16);
  cudaFuncSetAttribute(insitu_kernel<VAR>, cudaFuncAttributeMaxDynamicSharedMemorySize, 80 * 1024);
  const int iters = 1000;
  for (int threads : {64, 448}) {
    insitu_kernel<VAR><<<148, threads, 72 * 1024>>>(64, iters, 2, d);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[2]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
    printf("in-situ pattern var=%d threads=%d (%d spinning warps): %7.1f clk per MMA aggregate (warp-0 issue %6.1f) %s\n", VAR, threads,
           threads / 32 - 2, (double)h[1] / (iters * 12.0), (double)h[0] / (iters * 6.0), e == cudaSuccess ? "" : cudaGetErrorString(e));
  }
  cudaFree(d);
}

template <int TS>
void run_multi(const char* name) {
  long long* d; cudaMalloc(&d, 16);
  cudaFuncSetAttribute(multi_issue_kernel<TS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  for (int N : {32, 64})
    for (int nw : {1, 2, 4}) {
      const int iters = 1000;
      multi_issue_kernel<TS><<<148, 128, 60 * 1024>>>(N, iters, nw, d);
      cudaError_t e = cudaDeviceSynchronize();
      long long h[2]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
      printf("%-10s N=%3d issuing warps %d: %7.1f clk per MMA (aggregate; warp-0 issue %6.1f) %s\n", name, N, nw,
             (double)h[1] / (iters * 4.0 * nw), (double)h[0] / (iters * 4.0), e == cudaSuccess ? "" : cudaGetErrorString(e));
    }
  cudaFree(d);
}

int main() {
  run_insitu<0>(); run_insitu<7>();
  run_multi<0>("tf32 SS");
  run_multi<1>("tf32 TS");

  run_commit<0>("tf32 SS");
  run_commit<1>("tf32 TS");
  return 0;

  for (int ctas : {1, 148}) {
    run<0, 0>("tf32 SS", ctas);
    run<0, 1>("tf32 TS", ctas);
    run<1, 0>("bf16 SS", ctas);
    run<1, 1>("bf16 TS", ctas);
  }
  return 0;
}

// Plumbing for the fused "radial-MLP last layer + convolution" kernels (codegen/fused_codegen.py).
//
// The per-edge tensor-product weights  w[e, :] = h2[e, :64] @ W2[64, wn]  are the largest tensor of the path
// (3.84 KB per edge and layer for SevenNet-0).  The fused forward kernel never writes them to HBM: a CTA owns a tile
// of whole destination rows with <= 128 edges, produces w for the tile 64 columns at a time on the tensor core
// (tcgen05, 3xTF32, A = h2 tile in TMEM, B = pre-swizzled W2 chunk images through TMA bulk copies) and hands every
// chunk through shared memory to "math" warps that run the same Clebsch-Gordan code as the stand-alone convolution.
//
// Warp roles (22 warps, one CTA per SM, persistent over tiles):
//   0-3   tile loaders : h2 rows -> (hi, lo) -> TMEM A buffer (tcgen05.st), source indices + spherical harmonics -> smem
//   4-7   stagers      : accumulator chunk TMEM -> registers (tcgen05.ld) -> shared-memory w stage [edge][64], swizzled;
//                        one of them refills the B-image ring (cp.async.bulk, 32 KB (hi, lo) image of chunk g + 3)
//   8-9   MMA issuers  : chunk g is issued by warp g & 1 (one thread sustains ~50 clk per MMA, tools/mma_rate.cu)
//   10-21 math         : 4 warps per ring slot; an item = (destination row of the tile, 32-column half of the chunk)
// Every ring (B images, accumulators in TMEM, w stages) has 3 slots indexed by the global chunk counter g.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#define FZ_THREADS 704
#define FZ_TILE_E 128                 // tensor-core M: edges per tile (rows beyond the tile's edge count are zero)
#define FZ_CH 64                      // w columns per chunk
#define FZ_SLOTS 3
#define FZ_MATH_WARP0 10
#define FZ_MATH_WARPS 12
#define FZ_MAX_RT 8                   // destination rows per tile (RT * max degree <= 128)

__device__ __forceinline__ uint32_t fz_smem(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void fz_mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void fz_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void fz_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// Polling with back-off: on this part try_wait returns at once instead of parking the warp, and ten pipeline warps
// spinning on their barriers took ~75 % of the issued instructions away from the math warps (ncu source view).
// A failed poll now sleeps; the rings have enough slack to absorb the wake-up latency.
__device__ __forceinline__ void fz_wait(uint32_t bar, uint32_t parity, uint32_t sleep_ns = 100u) {
  uint32_t done;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
  while (!done) {
    __nanosleep(sleep_ns);
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
  }
}
__device__ __forceinline__ void fz_bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
// K-major, SWIZZLE_128B, 8-row groups 1024 B apart (same descriptor as gemm_tc.cu)
__device__ __forceinline__ uint64_t fz_desc(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3FFFF) >> 4) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
__device__ __forceinline__ void fz_mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void fz_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void fz_tmem_st32(uint32_t taddr, const uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
      ::"r"(taddr),
        "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
        "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]),
        "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]),
        "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
      : "memory");
}
__device__ __forceinline__ void fz_tmem_st16(uint32_t taddr, const uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
      ::"r"(taddr),
        "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
        "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
      : "memory");
}
__device__ __forceinline__ void fz_tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
        "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
        "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr) : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void fz_sts128(uint32_t a, float x, float y, float z, float w) {
  asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(a), "f"(x), "f"(y), "f"(z), "f"(w) : "memory");
}

struct FzParams {
  const float* h2;        // [E, 64] activations of the last hidden radial layer
  const float* Bimg;      // [NCH][2 k-blocks][2 (hi, lo)][64][32] pre-swizzled images of W2 / sqrt(64)
  const float* x; int ldx;
  const float* Y;         // [E, S]
  const int* rowptr; const int* col;
  int n_dst, RT;          // destination rows, rows per tile
  float inv_denom;
  float* out; int ldo;
  float* wdbg;            // optional [E, NCH*64]: the staged w chunks (tests)
};

// shared-memory map (bytes from a 1024-aligned base)
#define FZ_B_BYTES (FZ_SLOTS * 32768)                   // B ring
#define FZ_W_OFF FZ_B_BYTES                             // w stages: FZ_SLOTS x [128][64] fp32
#define FZ_W_BYTES (FZ_SLOTS * 32768)
#define FZ_TI_OFF (FZ_W_OFF + FZ_W_BYTES)               // tile info (single slot; the loaders hold the next tile in registers)
#define FZ_TI_COL 0                                     //   int col[128]
#define FZ_TI_ROW 512                                   //   int row_e[FZ_MAX_RT + 1] (tile-local edge offsets) @512
#define FZ_TI_Y 576                                     //   float Y[128][16]
#define FZ_TI_BYTES (576 + 128 * 16 * 4)
#define FZ_XR_OFF (FZ_TI_OFF + FZ_TI_BYTES)             // per-math-warp cp.async rings for the gathered source rows
#define FZ_XR_FLOATS 512                                //   2 KB per warp: depth = 512 / (32 * (2 l1 + 1)) edges
#define FZ_BAR_OFF (FZ_XR_OFF + FZ_MATH_WARPS * FZ_XR_FLOATS * 4)
// barriers (8 B each)
#define FZ_BAR_A_FULL 0        // [2] h2 tile in TMEM          (128 loader threads)
#define FZ_BAR_A_EMPTY 2       // [2] MMAs of the tile retired  (2 commits)
#define FZ_BAR_TI_FULL 4       // [1] tile info in smem         (128 loader threads)
#define FZ_BAR_TI_EMPTY 5      // [1] math warps done with tile (FZ_MATH_WARPS arrivals)
#define FZ_BAR_B_FULL 6        // [3] B image landed            (1 arrival + tx bytes)
#define FZ_BAR_D_FULL 9        // [3] accumulator chunk ready   (1 commit)
#define FZ_BAR_D_EMPTY 12      // [3] stagers drained TMEM      (128)
#define FZ_BAR_W_FULL 15       // [3] w stage written           (128)
#define FZ_BAR_W_EMPTY 18      // [3] math warps done with stage (4)
#define FZ_NBARS 21
#define FZ_SMEM_BYTES (1024 + FZ_BAR_OFF + FZ_NBARS * 8 + 16)
// tensor memory map (columns): A buffers 2 x 128 (hi 64 | lo 64), accumulators 3 x 64
#define FZ_TM_A 0
#define FZ_TM_D 256
#define FZ_CP_ASYNC4(sdst, gsrc) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" \
    :: "r"((unsigned)__cvta_generic_to_shared(sdst)), "l"(gsrc) : "memory")
#define FZ_CP_COMMIT() asm volatile("cp.async.commit_group;" ::: "memory")
#define FZ_CP_WAIT(n) asm volatile("cp.async.wait_group %0;" :: "n"(n) : "memory")

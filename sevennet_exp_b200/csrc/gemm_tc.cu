// tcgen05 / TMEM GEMM in 3xTF32 (fp32-accurate) with the same argument contract as gemm_simt.cu.
//
//   C[(n,m), :N] = epi( A[(n,m), :K] @ B[K, N] )        A, C fp32 with (node, m)-strided rows
//
// Replaces the cuBLAS calls behind e3nn's FullyConnectedNet (radial MLP, sevenn/nn/convolution.py:92-101,121)
// and o3.Linear (sevenn/nn/linear.py:49-63,94) and their input-gradients.  Precision: every fp32 operand is
// split x = hi + lo with hi = tf32(x); D += A_hi.B_hi + A_lo.B_hi + A_hi.B_lo accumulates in fp32 in TMEM,
// which keeps ~2^-21 relative error per product (north_star tolerances are fp32 ones).
//
// Structure of gemm_tc_kernel (persistent CTAs, one per SM, 608 threads; the large-K kernel gemm_tca_kernel is
// described where it is defined):
//   warps 0-7   producers : each thread owns one 16 B chunk of four rows of every 32-wide k-block; three k-blocks in
//                           flight in registers (read-only 16 B loads), split x = hi + lo in registers, stored once as
//                           the two SWIZZLE_128B K-major operand tiles of a ring slot.
//   warp  18    B loader  : B is a pre-swizzled (hi, lo) image.  "resident" = the whole [K, Nt] image once per CTA when
//                           it fits beside two A stages (<= 128 KB), else "streaming" = one k-block image per ring slot
//                           (TMA bulk copies, cp.async.bulk; mbarrier.arrive.expect_tx is the 257th arrival of the slot).
//   warps 16,17 MMA       : alternate row tiles; 4 k-steps x 3 tcgen05.mma.kind::tf32 (M=128, N=Nt<=256, K=8) per
//                           k-block; tcgen05.commit frees the slot / publishes the accumulator.  One thread sustains
//                           ~50 clk per MMA and ~270 clk per commit, which bounded K <= 64 tiles with one issuer.
//   warps 8-15  epilogue  : two per TMEM lane quadrant (even / odd 32-column chunks): tcgen05.ld -> swizzled smem
//                           transpose -> full-line coalesced 16 B accesses with the fused epilogue (store / accumulate /
//                           act / act-gradient).
//   TMEM: four 128-column accumulators when Nt <= 128 (two per issuing warp), else two of 256 columns: the epilogue of
//   a tile overlaps loads + MMAs of the next ones.
#include "sevenn_common.cuh"
#include "../../include/sevenn_b200.h"

#ifndef SEVENN_EMU
#include <cuda.h>
#include <cstdlib>
#include <cstdio>

#define TC_BM 128
#define TC_BK 32                      // fp32 per k-block = one 128 B swizzle row
#define TC_PRODUCERS 256
#define TC_THREADS 608                 // gemm_tc_kernel: 8 producer + 8 epilogue + 2 MMA-issuing warps + B loader
#define TC_BLOAD_WARP 18               // gemm_tc_kernel: streams the B images (TMA bulk copies) in streaming mode
#define TC_MMA_WARP 12                 // gemm_tca_kernel (448 threads): warps 12, 13 issue
#define TC_MMA2_WARP 16                // gemm_tc_kernel: warps 16, 17 issue (even / odd row tiles)
#define TC_EPI_WARP0 8
#define TC_EPI_BYTES (4 * 32 * 128)  // per-warp 32 rows x 128 B transpose buffers (gemm_tca_kernel: 4 epilogue warps)
#define TC_EPI8_BYTES (8 * 32 * 128) // gemm_tc_kernel: 8 epilogue warps

struct TcParams {
  sevenn_gemm_args g;
  const float* Bimg;    // [n_tiles][KB][2][Nt][32] pre-swizzled hi/lo images
  int Nt, n_tiles, KB, stages, vec4, mode, resident;
  long long m_tiles;
  long long* dbg;       // optional clock64 trace of CTA (0,0) (SEVENN_TC_DBG=1)
};
#define TC_MAX_GROUP SEVENN_GEMM_GROUP_MAX      // 8: the per-l blocks of up to two o3.Linear with disjoint outputs
struct TcGroup {
  int n;
  int cta_start[TC_MAX_GROUP + 1];
  TcParams p[TC_MAX_GROUP];
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {   // no arrival: only raises tx-count
  asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE;\n\t"
      "bra WAIT_LOOP;\n\t"
      "DONE:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr) {
  // K-major, SWIZZLE_128B, 8-row groups 1024 B apart (cute::UMMA::SmemDescriptor, mma_sm100_desc.hpp)
  return (uint64_t)((saddr & 0x3FFFF) >> 4) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// row r = n*R + m -> n*sn + m*sm ; 32-bit division (rows < 2^31), none at all for edge matrices (R == 1)
__device__ __forceinline__ long long row_offset(long long r, int R, long long sn, int sm) {
  if (R == 1) return r * sn;
  const unsigned n = (unsigned)r / (unsigned)R, m = (unsigned)r - n * (unsigned)R;
  return (long long)n * sn + (long long)m * sm;
}
// (node, m) of a row that advances by a fixed number of rows per step, without a division per step
struct RowWalk {
  int n, m, dn, dm;
  __device__ __forceinline__ void init(long long r, int R, long long step) {
    n = (int)((unsigned)r / (unsigned)R); m = (int)((unsigned)r - (unsigned)n * (unsigned)R);
    dn = (int)((unsigned)step / (unsigned)R); dm = (int)((unsigned)step - (unsigned)dn * (unsigned)R);
  }
  __device__ __forceinline__ void next(int R) { n += dn; m += dm; if (m >= R) { m -= R; ++n; } }
  __device__ __forceinline__ void add(int R, int sn_, int sm_) { n += sn_; m += sm_; if (m >= R) { m -= R; ++n; } }
  __device__ __forceinline__ long long row(int R) const { return (long long)n * R + m; }
  __device__ __forceinline__ long long off(long long sn, int sm) const { return (long long)n * sn + (long long)m * sm; }
};
__device__ __forceinline__ float4 lds128(uint32_t a) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ void sts128(uint32_t a, float4 v) {
  asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(a), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
// the tcgen05 arm fuses SiLU only (act_radial default, sevenn/_const.py:102); other activations use the SIMT arm
// MUFU-based (ex2.approx / rcp.approx): ~3e-7 relative error for |x| < 5.  The libm forms cost ~40 instructions per
// element, which made the 128 x 64 SiLU epilogue (8,192 evaluations per tile) the bottleneck of the K <= 64 layers.
__device__ __forceinline__ float silu_f(float x) { return __fdividef(x, 1.f + __expf(-x)); }
__device__ __forceinline__ float dsilu_f(float x) { const float s = __fdividef(1.f, 1.f + __expf(-x)); return s * (1.f + x * (1.f - s)); }

// Epilogue of one 128-row tile for one warp (TMEM lanes 32*ew ..): tcgen05.ld 32 columns at a time, transpose through
// XOR-swizzled smem, full-line coalesced 16 B accesses with the fused epilogue.
template <int EPI>
__device__ __forceinline__ void tc_epilogue_tile(const sevenn_gemm_args& g, uint32_t d_taddr, uint8_t* tb, int ew, int lane,
                                                 const long long (&roff)[8], int n0, int ncols, int vec4, int mode,
                                                 int part2_cols = 0, int cc0 = 0, int cstep = 32) {
  const int sub = lane >> 3, ch = lane & 7; (void)sub;
  for (int cc = cc0; cc < ncols; cc += cstep) {
    uint32_t v[32];
    const uint32_t taddr = d_taddr + ((uint32_t)(ew * 32) << 16) + (uint32_t)cc;
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
    if (part2_cols) {                                  // accumulator split over two issuing warps: D = D_x + D_y
      uint32_t u[32];
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
          "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
          "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
          : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]), "=r"(u[4]), "=r"(u[5]), "=r"(u[6]), "=r"(u[7]),
            "=r"(u[8]), "=r"(u[9]), "=r"(u[10]), "=r"(u[11]), "=r"(u[12]), "=r"(u[13]), "=r"(u[14]), "=r"(u[15]),
            "=r"(u[16]), "=r"(u[17]), "=r"(u[18]), "=r"(u[19]), "=r"(u[20]), "=r"(u[21]), "=r"(u[22]), "=r"(u[23]),
            "=r"(u[24]), "=r"(u[25]), "=r"(u[26]), "=r"(u[27]), "=r"(u[28]), "=r"(u[29]), "=r"(u[30]), "=r"(u[31])
          : "r"(taddr + (uint32_t)part2_cols) : "memory");
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
      for (int q = 0; q < 32; ++q) v[q] = __float_as_uint(__uint_as_float(v[q]) + __uint_as_float(u[q]));
    }
    if (mode & 2) continue;
    // transpose through smem: thread = row writes 8 chunks, XOR-swizzled (conflict free both ways)
    const uint32_t tbs = smem_u32(tb);
#pragma unroll
    for (int q = 0; q < 8; ++q)
      sts128(tbs + lane * 128 + ((q ^ (lane & 7)) << 4),
             make_float4(__uint_as_float(v[4 * q]), __uint_as_float(v[4 * q + 1]),
                         __uint_as_float(v[4 * q + 2]), __uint_as_float(v[4 * q + 3])));
    __syncwarp();
    const int col = cc + ch * 4;                     // first of this lane's 4 columns (tile-relative)
    const int nvalid = ncols - col;                  // >= 4 -> whole float4 valid
    // two passes of four rows: keeps the live set at 4 + 4 float4 (the kernel runs 576 threads at <= 96 registers)
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      float4 x[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int row = 4 * (4 * h + i) + sub;
        x[i] = lds128(tbs + row * 128 + ((ch ^ (row & 7)) << 4));
      }
      if (vec4 && nvalid >= 4) {
        float4 aux[4];
        if (EPI == 1 || EPI == 3) {
#pragma unroll
          for (int i = 0; i < 4; ++i)
            if (roff[4 * h + i] >= 0)
              aux[i] = *reinterpret_cast<const float4*>((EPI == 1 ? g.C : g.P) + roff[4 * h + i] + n0 + col);
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const long long ro = roff[4 * h + i];
          if (ro < 0) continue;
          float4* c4 = reinterpret_cast<float4*>(g.C + ro + n0 + col);
          const float4 xv = x[i];
          if (EPI == 0) {
            *c4 = xv;
          } else if (EPI == 1) {
            *c4 = make_float4(aux[i].x + xv.x, aux[i].y + xv.y, aux[i].z + xv.z, aux[i].w + xv.w);
          } else if (EPI == 2) {
            *c4 = xv;
            *reinterpret_cast<float4*>(g.C2 + ro + n0 + col) =
                make_float4(g.cst * silu_f(xv.x), g.cst * silu_f(xv.y), g.cst * silu_f(xv.z), g.cst * silu_f(xv.w));
          } else {
            *c4 = make_float4(xv.x * g.cst * dsilu_f(aux[i].x), xv.y * g.cst * dsilu_f(aux[i].y),
                              xv.z * g.cst * dsilu_f(aux[i].z), xv.w * g.cst * dsilu_f(aux[i].w));
          }
        }
      } else if (nvalid > 0) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const long long ro = roff[4 * h + i];
          if (ro < 0) continue;
          const float xs[4] = {x[i].x, x[i].y, x[i].z, x[i].w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            if (e >= nvalid) break;
            float* cp = g.C + ro + n0 + col + e;
            if (EPI == 0) *cp = xs[e];
            else if (EPI == 1) *cp += xs[e];
            else if (EPI == 2) { *cp = xs[e]; g.C2[ro + n0 + col + e] = g.cst * silu_f(xs[e]); }
            else *cp = xs[e] * g.cst * dsilu_f(g.P[ro + n0 + col + e]);
          }
        }
      }
    }
    __syncwarp();
  }
}

#define DBG(role, it, ev) do { if (P.dbg && blockIdx.x == 0 && (it) >= 0 && (it) < 6 && (threadIdx.x & 31) == 0) \
  P.dbg[((role) * 6 + (it)) * 8 + (ev)] = clock64(); } while (0)

// Up to TC_MAX_GROUP independent problems per launch (the per-l blocks of one o3.Linear): each CTA serves one of
// them, so the launch / TMEM-allocation / pipeline-fill latencies of the small node GEMMs overlap.
template <int EPI>
__global__ void __launch_bounds__(TC_THREADS, 1) gemm_tc_kernel(const TcGroup G) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  int pi = 0;
  while (pi + 1 < G.n && (int)blockIdx.x >= G.cta_start[pi + 1]) ++pi;
  const TcParams P = G.p[pi];          // by value: fields live in registers, not behind indexed constant loads
  const int local = (int)blockIdx.x - G.cta_start[pi];
  const int gx = (G.cta_start[pi + 1] - G.cta_start[pi]) / P.n_tiles;      // CTAs per column tile
  const int bx = local % gx;
  const sevenn_gemm_args& g = P.g;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t a_bytes = TC_BM * 128, b_bytes = (uint32_t)P.Nt * 128;
  // stage = A_hi | A_lo | (streaming: B_hi | B_lo)
  const uint32_t stage_bytes = 2 * a_bytes + (P.resident ? 0u : 2 * b_bytes);
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* bres = smem + (size_t)P.stages * stage_bytes;                     // resident B image [KB][2][Nt][32]
  uint8_t* epi_buf = bres + (P.resident ? (size_t)P.KB * 2 * b_bytes : 0);
  uint64_t* bars = (uint64_t*)(epi_buf + TC_EPI8_BYTES);
  // bars: full[2][stages] (one set per MMA-issuing warp: a k-block of an even row tile completes full[0][slot], of an
  // odd one full[1][slot], so each warp sees every phase of the barriers it waits on -- a parity wait cannot tell
  // phase p from phase p + 2), empty[stages], tmem_full[2], tmem_empty[2], b_full; then the TMEM base address
  const uint32_t bar0 = smem_u32(bars);
  auto full_bar = [&](int w, int s) { return bar0 + 8u * (w * P.stages + s); };
  auto empty_bar = [&](int s) { return bar0 + 8u * (2 * P.stages + s); };
  auto tfull_bar = [&](int b) { return bar0 + 8u * (3 * P.stages + b); };
  auto tempty_bar = [&](int b) { return bar0 + 8u * (3 * P.stages + 4 + b); };
  const uint32_t bfull_bar = bar0 + 8u * (3 * P.stages + 8);
  uint32_t* tmem_slot = (uint32_t*)(bars + 3 * P.stages + 9);
  // accumulator buffers in TMEM: four of 128 columns when the tile is narrow enough (each MMA-issuing warp then
  // alternates between two, so its next tile does not wait for the epilogue of its previous one), else two of 256
  const int nbuf = P.Nt <= 128 ? 4 : 2;
  const uint32_t buf_cols = 512u / (uint32_t)nbuf;

  if (threadIdx.x == 0) {
    for (int s = 0; s < P.stages; ++s) {
      const uint32_t arrivals = TC_PRODUCERS + (P.resident ? 0u : 1u);     // + the B loader's arrive.expect_tx
      mbar_init(full_bar(0, s), arrivals); mbar_init(full_bar(1, s), arrivals); mbar_init(empty_bar(s), 1);
    }
    for (int b = 0; b < 4; ++b) { mbar_init(tfull_bar(b), 1); mbar_init(tempty_bar(b), 256); }
    mbar_init(bfull_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == TC_MMA2_WARP) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(tmem_slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;

  // this CTA owns column tile nt and the row tiles mt = bx, bx + gx, ...
  const int nt = local / gx;
  const long long my_items = (P.m_tiles > bx) ? (P.m_tiles - bx + gx - 1) / gx : 0;
  const float* bimg_nt = P.Bimg + (size_t)nt * P.KB * (size_t)(2 * P.Nt * 32);

  if (warp < TC_PRODUCERS / 32) {
    // ================= producers: global -> registers -> (hi, lo) split -> shared memory ===================
    // Each thread owns one 16 B chunk of four rows of every k-block.  TC_RD k-blocks are in flight in registers
    // (plain 16 B read-only loads); when one has landed it is split in registers and stored once as the two
    // SWIZZLE_128B K-major operand tiles.  The first version staged the raw rows in shared memory with cp.async and
    // split them in place: 16 KB in + 16 KB out + 32 KB in per k-block, and the LDGSTS issue itself cost ~1,000 clk
    // per k-block (pipeline traces, profiles/r02_gemm_tc_traces.txt); this way shared memory sees the 32 KB once.
    // All ring / row bookkeeping is incremental 32-bit arithmetic (no division per k-block).
    const int tid = threadIdx.x;
    const int c = tid & 7;                  // 16 B chunk inside the 128 B k-block row
    const long long total_q = my_items * P.KB;
    if (P.resident && tid == 0 && my_items > 0) {
      const uint32_t bytes = (uint32_t)P.KB * 2 * b_bytes;
      mbar_expect_tx(bfull_bar, bytes);
      bulk_g2s(smem_u32(bres), bimg_nt, bytes, bfull_bar);
      mbar_arrive(bfull_bar);
    }
    long long q_i = 0; int kb_i = 0;
    RowWalk rw;                               // this thread's first row of the tile; the other three are 32 rows apart
    rw.init((long long)bx * TC_BM + (tid >> 3), g.R, (long long)gx * TC_BM);
    const int s32n = 32 / g.R, s32m = 32 % g.R;
    const float* rowp[4];
    auto set_rows = [&]() {
      RowWalk t = rw;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        rowp[i] = (t.row(g.R) < g.M) ? g.A + t.off(g.sAn, g.sAm) : nullptr;
        t.add(g.R, s32n, s32m);
      }
    };
    if (total_q > 0) set_rows();
    const uint32_t dst_off = (uint32_t)((tid >> 3) * 128 + ((c ^ ((tid >> 3) & 7)) << 4));    // + i * 32 rows * 128 B
    // issue the loads of k-block q_i into v (never waits: the ring slot is only needed when the k-block is stored)
    auto load = [&](float4 (&v)[4]) {
      if (q_i >= total_q) return;
      const int k = kb_i * TC_BK + c * 4;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (rowp[i] != nullptr && k < g.K)
          asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];"
                       : "=f"(v[i].x), "=f"(v[i].y), "=f"(v[i].z), "=f"(v[i].w) : "l"(rowp[i] + k));
      }
      ++q_i;
      if (++kb_i == P.KB) {
        kb_i = 0;
        rw.next(g.R);
        if (q_i < total_q) set_rows();
      }
    };
    int slot = 0, kb_q = 0, par_q = 0; uint32_t ph_q = 0;
    long long q = 0;
    // split k-block q (in v) into its ring slot and publish it
    auto store = [&](const float4 (&v)[4]) {
      if (warp == 0) DBG(0, q, 1);
      mbar_wait(empty_bar(slot), ph_q ^ 1);                         // slot released by the MMAs of k-block q - stages
      if (warp == 0) DBG(0, q, 5);
      const uint32_t sbase = smem_u32(smem) + (uint32_t)slot * stage_bytes + dst_off;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float4 hi, lo;
        hi.x = __uint_as_float(__float_as_uint(v[i].x) & 0xFFFFE000u); lo.x = v[i].x - hi.x;
        hi.y = __uint_as_float(__float_as_uint(v[i].y) & 0xFFFFE000u); lo.y = v[i].y - hi.y;
        hi.z = __uint_as_float(__float_as_uint(v[i].z) & 0xFFFFE000u); lo.z = v[i].z - hi.z;
        hi.w = __uint_as_float(__float_as_uint(v[i].w) & 0xFFFFE000u); lo.w = v[i].w - hi.w;
        sts128(sbase + i * (32 * 128), hi);
        sts128(sbase + i * (32 * 128) + a_bytes, lo);
      }
      if (warp == 0) DBG(0, q, 3);
      if (!(P.mode & 8)) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy stores -> UMMA (async proxy)
      mbar_arrive(full_bar(par_q, slot));                           // 256 arrivals (+ B bytes) complete the phase
      if (warp == 0) DBG(0, q, 2);
      ++q;
      if (++slot == P.stages) { slot = 0; ph_q ^= 1; }
      if (++kb_q == P.KB) { kb_q = 0; par_q ^= 1; }
    };
    // register ring of three k-blocks
    float4 r0[4], r1[4], r2[4];
    load(r0); load(r1); load(r2);
    while (q < total_q) {
      store(r0); load(r0);
      if (q < total_q) { store(r1); load(r1); }
      if (q < total_q) { store(r2); load(r2); }
    }
  } else if (warp == TC_BLOAD_WARP) {
    // ================= B loader (streaming mode): one k-block image per ring slot, as far ahead as the ring allows ===
    if (!P.resident && lane == 0) {
      const long long total_q = my_items * P.KB;
      int slot = 0, kb = 0, par = 0; uint32_t ph = 0;
      for (long long q = 0; q < total_q; ++q) {
        mbar_wait(empty_bar(slot), ph ^ 1);
        const uint32_t sb = smem_u32(smem) + (uint32_t)slot * stage_bytes + 2 * a_bytes;
        mbar_arrive_expect_tx(full_bar(par, slot), 2 * b_bytes);    // the 257th arrival of the phase, + the image bytes
        bulk_g2s(sb, bimg_nt + (size_t)kb * (size_t)(2 * P.Nt * 32), 2 * b_bytes, full_bar(par, slot));
        if (++slot == P.stages) { slot = 0; ph ^= 1; }
        if (++kb == P.KB) { kb = 0; par ^= 1; }
      }
    }
  } else if (warp == TC_MMA2_WARP || warp == TC_MMA2_WARP + 1) {
    // ================= MMA issuers: warp 16 takes the even row tiles (TMEM buffer 0), warp 17 the odd ones =========
    // One thread sustains one tcgen05.mma per ~50 clk and ~270 clk per tcgen05.commit (tools/mma_rate.cu): at
    // K <= 64 that issue time, not the tensor pipe, bounded a tile, so two warps issue alternate tiles.
    const int w = warp - TC_MMA2_WARP;
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(P.Nt >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);
    // ring position of k-block w * KB, the advance over the other warp's tile, and the parity this warp expects next
    // on each of its full barriers (bit s: slot s)
    int stage = (w * P.KB) % P.stages;
    const int skip_s = P.KB % P.stages;
    uint32_t pbits = 0;
    if (P.resident && my_items > w) mbar_wait(bfull_bar, 0);
    // this warp's buffers: w (and w + 2 when there are four); eph bit b = parity expected next on tempty[b]
    int ab = w; uint32_t eph = 0;
    for (long long it = w; it < my_items; it += 2) {
      const uint32_t d_tmem = tmem_base + (uint32_t)ab * buf_cols;
      DBG(1, it, 0);
      mbar_wait(tempty_bar(ab), ((eph >> ab) & 1u) ^ 1u);
      eph ^= 1u << ab;
      DBG(1, it, 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      for (int kb = 0; kb < P.KB; ++kb) {
        mbar_wait(full_bar(w, stage), (pbits >> stage) & 1u);
        pbits ^= 1u << stage;
        if (kb == 0) DBG(1, it, 2);
        if (kb == P.KB - 1) DBG(1, it, 3);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        if (lane == 0) {
          const uint32_t sa = smem_u32(smem) + (uint32_t)stage * stage_bytes;
          const uint32_t sb = P.resident ? smem_u32(bres) + (uint32_t)kb * 2 * b_bytes : sa + 2 * a_bytes;
          const uint64_t a_hi = make_desc(sa), a_lo = make_desc(sa + a_bytes);
          const uint64_t b_hi = make_desc(sb), b_lo = make_desc(sb + b_bytes);
#pragma unroll
          for (int j = 0; j < TC_BK / 8; ++j) {
            const uint64_t adv = (uint64_t)(j * 2);     // 32 B per k-step, in 16 B units
            if (P.mode & 4) {
            } else if (!(P.mode & 1)) {
              umma_tf32(d_tmem, a_lo + adv, b_hi + adv, idesc, (kb | j) != 0);
              umma_tf32(d_tmem, a_hi + adv, b_lo + adv, idesc, 1);
              umma_tf32(d_tmem, a_hi + adv, b_hi + adv, idesc, 1);
            } else {
              umma_tf32(d_tmem, a_hi + adv, b_hi + adv, idesc, (kb | j) != 0);   // diagnostic: plain TF32
            }
          }
          umma_commit(empty_bar(stage));
          if (kb == P.KB - 1) umma_commit(tfull_bar(ab));
        }
        __syncwarp();
        if (kb == P.KB - 1) DBG(1, it, 4);
        if (++stage == P.stages) stage = 0;
      }
      if (nbuf == 4) ab ^= 2;
      stage += skip_s;                                  // the other warp's tile
      if (stage >= P.stages) stage -= P.stages;
    }
  } else {
    // ================= epilogue: warps 8..15, TMEM lanes 32 * (warp % 4) .., column chunks split even / odd ========
    const int ew = (warp - TC_EPI_WARP0) & 3, half = (warp - TC_EPI_WARP0) >> 2;
    uint8_t* tb = epi_buf + (warp - TC_EPI_WARP0) * (32 * 128);   // this warp's 32 rows x 128 B transpose buffer
    const int n0 = nt * P.Nt;
    const int ncols = min(P.Nt, g.N - n0);
    const int sub = lane >> 3;                           // read phase: row 4*i + sub, 16 B chunk lane & 7
    RowWalk rw;                               // row `sub` of this warp's 32; the other seven are 4 rows apart
    rw.init((long long)bx * TC_BM + ew * 32 + sub, g.R, (long long)gx * TC_BM);
    const int s4n = 4 / g.R, s4m = 4 % g.R;
    for (long long it = 0; it < my_items; ++it) {
      const int ab = (int)(it & (nbuf - 1)); const uint32_t aphase = (uint32_t)((it >> (nbuf == 4 ? 2 : 1)) & 1);
      long long roff[8];
      {
        RowWalk t = rw;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          roff[i] = (t.row(g.R) < g.M) ? t.off(g.sCn, g.sCm) : -1;
          t.add(g.R, s4n, s4m);
        }
        rw.next(g.R);
      }
      if (warp == TC_EPI_WARP0) DBG(2, it, 0);
      mbar_wait(tfull_bar(ab), aphase);
      if (warp == TC_EPI_WARP0) DBG(2, it, 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      tc_epilogue_tile<EPI>(g, tmem_base + (uint32_t)ab * buf_cols, tb, ew, lane, roff, n0, ncols, P.vec4, P.mode, 0,
                            32 * half, 64);
      if (warp == TC_EPI_WARP0) DBG(2, it, 3);
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      mbar_arrive(tempty_bar(ab));
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == TC_MMA2_WARP) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
  }
}

// ---------------------------------------------------------------------------------------------------------
// Large-K variant (streaming B, Nt <= 64): the A operand is fed to the tensor core from TENSOR MEMORY.
// Shared memory is the scarce resource of the 3xTF32 scheme when A goes through it: per 32-wide k-block the
// smem-A kernel above moves 16 KB in (cp.async) + 16 KB out/32 KB in (hi/lo split) + 3 x 16 KB of MMA operand
// reads = 112 KB, i.e. ~900 clk at 128 B/clk against a tensor-pipe floor of 12 MMAs x 32 clk = 384 clk (N = 64).
// Here the producers read their own row from a plain fp32 staging ring, split it in registers and write
// (hi | lo) straight into TMEM with tcgen05.st (row r = lane r, k = column); the MMAs take A from TMEM
// ([d], [a] form) and only B (2 x 8 KB per k-block, TMA bulk) still crosses shared memory.
//   warps 0-3 / 4-7 : two producer groups, alternating k-blocks, cp.async staging ring (3 slots per group)
//   warps 12, 13    : MMA issuers.  One thread sustains one tcgen05.mma per ~50 clk and stalls ~270 clk per
//                     tcgen05.commit (tools/mma_rate.cu), more than the tensor pipe needs for N = 64 (32 clk), so the
//                     k-steps of every k-block are split between two issuing warps with private accumulators
//                     (D_x: k-steps 0,1; D_y: k-steps 2,3); the epilogue adds the two.
//   warps 8-11      : epilogue (D: TMEM columns 128*buffer + 64*part ..), same code as above
#define TCA_THREADS 448
#define TCA_NS 4                       // TMEM-A / smem-B stages
#define TCA_ACOL0 256                  // D: 2 buffers x 2 parts x 64 columns = [0, 256)
#define TCA_SLOTS 3                    // fp32 staging slots per producer group

__device__ __forceinline__ void umma_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&v)[32]) {
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

template <int EPI>
__global__ void __launch_bounds__(TCA_THREADS, 1) gemm_tca_kernel(const TcParams P) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const sevenn_gemm_args& g = P.g;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t a_bytes = TC_BM * 128, b_bytes = (uint32_t)P.Nt * 128;
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* bring = smem;                                           // TCA_NS x (B_hi | B_lo)
  uint8_t* aring = bring + (size_t)TCA_NS * 2 * b_bytes;           // 2 groups x TCA_SLOTS x 16 KB fp32 staging
  uint8_t* epi_buf = aring + (size_t)2 * TCA_SLOTS * a_bytes;
  uint64_t* bars = (uint64_t*)(epi_buf + TC_EPI_BYTES);
  const uint32_t bar0 = smem_u32(bars);
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (TCA_NS + s); };
  auto tfull_bar = [&](int b) { return bar0 + 8u * (2 * TCA_NS + b); };
  auto tempty_bar = [&](int b) { return bar0 + 8u * (2 * TCA_NS + 2 + b); };
  uint32_t* tmem_slot = (uint32_t*)(bars + 2 * TCA_NS + 4);

  if (threadIdx.x == 0) {
    for (int s = 0; s < TCA_NS; ++s) { mbar_init(full_bar(s), 128); mbar_init(empty_bar(s), 2); }
    for (int b = 0; b < 2; ++b) { mbar_init(tfull_bar(b), 2); mbar_init(tempty_bar(b), 128); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == TC_MMA_WARP) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(tmem_slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;

  const long long my_items = (P.m_tiles > blockIdx.x) ? (P.m_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  const long long total_q = my_items * P.KB;
  const float* bimg = P.Bimg;                                       // n_tiles == 1

  if (warp < 8) {
    // ================= producers =================
    const int grp = warp >> 2, gt = threadIdx.x & 127;             // gt = row of the tile this thread owns in TMEM
    const int c = gt & 7;
    const long long n_my = total_q > grp ? (total_q - grp + 1) / 2 : 0;
    uint8_t* my_ring = aring + (size_t)grp * TCA_SLOTS * a_bytes;
    long long rows_it = -1;
    const float* rowp[8];
    auto issue_copy = [&](long long j) {
      if (j < n_my && !(P.mode & 64)) {                              // (mode 64: diagnostic, no A traffic at all)
        const long long q = 2 * j + grp;
        const long long it = q / P.KB;
        const int kb = (int)(q - it * P.KB);
        if (it != rows_it) {
          rows_it = it;
          const long long mt = blockIdx.x + it * gridDim.x;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const long long r = mt * TC_BM + i * 16 + (gt >> 3);
            rowp[i] = (r < g.M) ? g.A + row_offset(r, g.R, g.sAn, g.sAm) : nullptr;
          }
        }
        const uint32_t sbase = smem_u32(my_ring + (size_t)(j % TCA_SLOTS) * a_bytes);
        const int k = kb * TC_BK + c * 4;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int row = i * 16 + (gt >> 3);
          const uint32_t dst = sbase + row * 128 + ((c ^ (row & 7)) << 4);
          const bool ok = rowp[i] != nullptr && k < g.K;
          const float* src = ok ? rowp[i] + k : g.A;
          const uint32_t nbytes = ok ? 16u : 0u;                    // zero fill: rows >= M and the K tail
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(nbytes) : "memory");
        }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
    };
    issue_copy(0);
    issue_copy(1);
    const uint32_t t_lane = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
    for (long long j = 0; j < n_my; ++j) {
      const long long qd = 2 * j + grp - 12;                        // trace window: k-blocks 12..17 of the first tile
      if ((warp & 3) == 0 && qd >= 0) DBG(grp * 2, qd, 0);
      asm volatile("cp.async.wait_group 1;" ::: "memory");          // my copies of k-block j have landed
      if ((warp & 3) == 0 && qd >= 0) DBG(grp * 2, qd, 1);
      asm volatile("bar.sync %0, 128;" ::"r"(1 + grp) : "memory"); // ... and everybody else's; slot (j-1)%3 is drained
      issue_copy(j + 2);
      const uint32_t srow = smem_u32(my_ring + (size_t)(j % TCA_SLOTS) * a_bytes) + gt * 128;
      uint32_t hi[32], lo[32];
#pragma unroll
      for (int cc = 0; cc < 8; ++cc) {
        const float4 v = (P.mode & 64) ? make_float4(1.f, 2.f, 3.f, 4.f) : lds128(srow + ((cc ^ (gt & 7)) << 4));
        hi[4 * cc + 0] = __float_as_uint(v.x) & 0xFFFFE000u; lo[4 * cc + 0] = __float_as_uint(v.x - __uint_as_float(hi[4 * cc + 0]));
        hi[4 * cc + 1] = __float_as_uint(v.y) & 0xFFFFE000u; lo[4 * cc + 1] = __float_as_uint(v.y - __uint_as_float(hi[4 * cc + 1]));
        hi[4 * cc + 2] = __float_as_uint(v.z) & 0xFFFFE000u; lo[4 * cc + 2] = __float_as_uint(v.z - __uint_as_float(hi[4 * cc + 2]));
        hi[4 * cc + 3] = __float_as_uint(v.w) & 0xFFFFE000u; lo[4 * cc + 3] = __float_as_uint(v.w - __uint_as_float(hi[4 * cc + 3]));
      }
      const long long q = 2 * j + grp;
      const int ts = (int)(q % TCA_NS);
      if ((warp & 3) == 0 && qd >= 0) DBG(grp * 2, qd, 2);
      // stages are released in pairs (one tcgen05.commit per two k-blocks: a commit stalls its issuing thread
      // for several hundred cycles, tools/mma_rate.cu), so wait on the pair's barrier
      mbar_wait(empty_bar((int)((q >> 1) & 1)), (uint32_t)(((q >> 2) & 1) ^ 1));
      if ((warp & 3) == 0 && qd >= 0) DBG(grp * 2, qd, 3);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      if (gt == 0 && (!(P.mode & 128) || q < TCA_NS)) {              // (mode 128: diagnostic, B loaded once per stage)
        const int kb = (int)(q % P.KB);
        mbar_expect_tx(full_bar(ts), 2 * b_bytes);
        bulk_g2s(smem_u32(bring + (size_t)ts * 2 * b_bytes),
                 bimg + (size_t)((P.mode & 32) ? 0 : kb) * (size_t)(2 * P.Nt * 32), 2 * b_bytes, full_bar(ts));
      }
      const uint32_t ta = t_lane + (uint32_t)(TCA_ACOL0 + ts * 64);
      if (!(P.mode & 16)) {                                          // diagnostic: skip the TMEM stores
        tmem_st32(ta, hi);
        tmem_st32(ta + 32, lo);
      }
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      mbar_arrive(full_bar(ts));
      if ((warp & 3) == 0 && qd >= 0) DBG(grp * 2, qd, 4);
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
  } else if (warp >= TC_MMA_WARP) {
    // ================= MMA issuers (part = warp - 12 owns k-steps 2*part, 2*part+1 and accumulator D_part) ======
    const int part = warp - TC_MMA_WARP;
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(P.Nt >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);
    long long q = 0;
    // a single thread runs the whole issue loop (no per-k-block warp reconvergence)
    if (lane == 0)
    for (long long it = 0; it < my_items; ++it) {
      const int ab = (int)(it & 1); const uint32_t aphase = (uint32_t)((it >> 1) & 1);
      mbar_wait(tempty_bar(ab), aphase ^ 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t d_tmem = tmem_base + (uint32_t)(ab * 128 + part * 64);
      for (int kb = 0; kb < P.KB; ++kb, ++q) {
        const int ts = (int)(q % TCA_NS);
        if (part == 0 && q >= 12) DBG(1, q - 12, 0);
        mbar_wait(full_bar(ts), (uint32_t)((q / TCA_NS) & 1));
        if (part == 0 && q >= 12) DBG(1, q - 12, 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        {
          const uint32_t sb = smem_u32(bring + (size_t)ts * 2 * b_bytes);
          const uint64_t b_hi = make_desc(sb), b_lo = make_desc(sb + b_bytes);
          const uint32_t a_hi = tmem_base + (uint32_t)(TCA_ACOL0 + ts * 64), a_lo = a_hi + 32;
#pragma unroll
          for (int jj = 0; jj < 2; ++jj) {
            const int j = 2 * part + jj;
            const uint64_t adv = (uint64_t)(j * 2);
            if (P.mode & 4) continue;                                  // diagnostic: no tensor work
            umma_tf32_ts(d_tmem, a_lo + j * 8, b_hi + adv, idesc, (kb | jj) != 0);
            if (P.dbg && part == 0 && q >= 12 && q < 18 && blockIdx.x == 0) P.dbg[((1) * 6 + (q - 12)) * 8 + 2 + jj * 3] = clock64();
            if (P.mode & 1) continue;                                  // diagnostic: 1 MMA per k-step
            umma_tf32_ts(d_tmem, a_hi + j * 8, b_lo + adv, idesc, 1);
            if (P.dbg && part == 0 && q >= 12 && q < 18 && blockIdx.x == 0) P.dbg[((1) * 6 + (q - 12)) * 8 + 3 + jj * 3] = clock64();
            umma_tf32_ts(d_tmem, a_hi + j * 8, b_hi + adv, idesc, 1);
            if (P.dbg && part == 0 && q >= 12 && q < 18 && blockIdx.x == 0) P.dbg[((1) * 6 + (q - 12)) * 8 + 4 + jj * 3] = clock64();
          }
          if ((q & 1) || q == total_q - 1) umma_commit(empty_bar((int)((q >> 1) & 1)));
          if (kb == P.KB - 1) umma_commit(tfull_bar(ab));
        }
      }
    }
  } else {
    // ================= epilogue =================
    const int ew = warp - TC_EPI_WARP0;
    uint8_t* tb = epi_buf + ew * (32 * 128);
    const int ncols = min(P.Nt, g.N);
    const int sub = lane >> 3;
    for (long long it = 0; it < my_items; ++it) {
      const long long mt = blockIdx.x + it * gridDim.x;
      const int ab = (int)(it & 1); const uint32_t aphase = (uint32_t)((it >> 1) & 1);
      long long roff[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const long long r = mt * TC_BM + ew * 32 + 4 * i + sub;
        roff[i] = (r < g.M) ? row_offset(r, g.R, g.sCn, g.sCm) : -1;
      }
      mbar_wait(tfull_bar(ab), aphase);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      tc_epilogue_tile<EPI>(g, tmem_base + (uint32_t)(ab * 128), tb, ew, lane, roff, 0, ncols, P.vec4, P.mode, 64);
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      mbar_arrive(tempty_bar(ab));
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == TC_MMA_WARP) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
  }
}

// geometry shared by the host-side image builder (engine.tc_weight_image) and the kernel
// Shared memory left for the A ring and a resident B image: 227 KB - alignment - epilogue buffers - barriers.
#define TC_SMEM_AVAIL (227 * 1024 - 1024 - TC_EPI8_BYTES - 256)
static int tc_plan(int N, int K, int* Nt, int* n_tiles, int* KB, int* resident) {
  if (N <= 0 || K <= 0) return SEVENN_ERR_BAD_ARG;
  const int kb = (K + TC_BK - 1) / TC_BK;
  int res = 0, tiles = (N + 255) / 256, nt = (((N + tiles - 1) / tiles) + 15) / 16 * 16;
  if ((long long)kb * nt * 256 + 2 * (2 * TC_BM * 128) <= TC_SMEM_AVAIL) {
    res = 1;                        // whole [K, N] image resident next to a ring of >= 2 A stages: A is read once
  } else if (N > 256) {
    // wide outputs (radial MLP, N = 960): more column tiles until one tile's image fits beside >= 3 stages; A (K <= 64
    // there) is re-read per column tile, which is small against the [128, N] output
    for (int t = tiles;; ++t) {
      const int n2 = (((N + t - 1) / t) + 15) / 16 * 16;
      if ((long long)kb * n2 * 256 <= 84 * 1024) { res = 1; tiles = t; nt = n2; break; }
      if (n2 <= 64) break;
    }
  }
  { // experiments: SEVENN_TC_TILES=t forces t column tiles for N <= 256 (resident when the tile's image fits)
    static int force = -1;
    if (force < 0) { const char* e = getenv("SEVENN_TC_TILES"); force = e ? atoi(e) : 0; }
    if (force > 0 && N <= 256) {
      tiles = force; nt = (((N + tiles - 1) / tiles) + 15) / 16 * 16;
      res = (long long)kb * nt * 256 + 2 * (2 * TC_BM * 128) <= TC_SMEM_AVAIL;
    }
  }
  // otherwise: one column tile, B streamed per k-block (the node linears with K * N > 16 K: A read once; the first
  // version split the columns until the image fitted and re-read A up to five times)
  *Nt = nt; *n_tiles = tiles; *KB = kb; *resident = res;
  return 0;
}

extern "C" int sevenn_gemm_tc_plan(int N, int K, int* Nt, int* n_tiles, int* KB, long long* img_floats);
// Device-side construction of the pre-swizzled (hi, lo) K-major image of a weight matrix W[K, N] (row-major, ld = N):
// img[t][kb][h][n][phys(k)] with phys = ((k / 4) ^ (n % 8)) * 4 + k % 4 -- the layout engine.py::tc_weight_image builds on
// the host.  For weights that change every step (training): one small launch instead of a host round trip.
__global__ void tc_image_kernel(const float* __restrict__ W, int K, int N, int Nt, int n_tiles, int KB,
                                float* __restrict__ img) {
  const long long total = (long long)n_tiles * KB * Nt * 32;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int kk = (int)(i % 32);
    const int n = (int)((i / 32) % Nt);
    const int kb = (int)((i / (32LL * Nt)) % KB);
    const int t = (int)(i / (32LL * Nt * KB));
    const int k = kb * 32 + kk, col = t * Nt + n;
    const float v = (k < K && col < N) ? W[(size_t)k * N + col] : 0.f;
    const float hi = __uint_as_float(__float_as_uint(v) & 0xFFFFE000u);
    const float lo = __uint_as_float(__float_as_uint(v - hi) & 0xFFFFE000u);
    const int phys = (((kk >> 2) ^ (n & 7)) << 2) | (kk & 3);
    const size_t base = (((size_t)t * KB + kb) * 2) * Nt * 32 + (size_t)n * 32 + phys;
    img[base] = hi;
    img[base + (size_t)Nt * 32] = lo;
  }
}
extern "C" int sevenn_gemm_tc_image(const float* W, int K, int N, float* img, void* stream) {
  int Nt, n_tiles, KB;
  long long fl;
  int rc = sevenn_gemm_tc_plan(N, K, &Nt, &n_tiles, &KB, &fl);
  if (rc) return rc;
  const long long total = (long long)n_tiles * KB * Nt * 32;
  tc_image_kernel<<<sevenn_grid_for(total), 256, 0, (cudaStream_t)stream>>>(W, K, N, Nt, n_tiles, KB, img);
  return (int)cudaGetLastError();
}

extern "C" int sevenn_gemm_tc_plan(int N, int K, int* Nt, int* n_tiles, int* KB, long long* img_floats) {
  int res;
  int rc = tc_plan(N, K, Nt, n_tiles, KB, &res);
  if (rc) return rc;
  *img_floats = (long long)(*n_tiles) * (*KB) * 2 * (*Nt) * 32;
  return 0;
}

// fills P (geometry, stages, modes) and the dynamic shared memory it needs; tca = large-K variant eligible
static int tc_prepare(const sevenn_gemm_args* args, const float* Bimg, TcParams& P, size_t* smem, bool* tca) {
  P.g = *args;
  const sevenn_gemm_args& g = P.g;
  if (g.K <= 0 || g.R <= 0 || g.epi < 0 || g.epi > 3 || (g.K & 3) || (g.sAn & 3) || (g.sAm & 3) ||
      ((uintptr_t)g.A & 15) || (g.epi >= 2 && g.act != 0))
    return SEVENN_ERR_BAD_ARG;
  tc_plan(g.N, g.K, &P.Nt, &P.n_tiles, &P.KB, &P.resident);
  P.Bimg = Bimg;
  P.m_tiles = (g.M + TC_BM - 1) / TC_BM;
  const size_t b_bytes = (size_t)P.Nt * 128;
  const size_t budget = TC_SMEM_AVAIL;
  const size_t stage_bytes = 2 * TC_BM * 128 + (P.resident ? 0 : 2 * b_bytes);
  const size_t fixed = P.resident ? (size_t)P.KB * 2 * b_bytes : 0;
  int stages = (int)((budget - fixed) / stage_bytes);
  if (stages > 6) stages = 6;
  if (stages < 2) return SEVENN_ERR_BAD_ARG;
  P.stages = stages;
  static int mma_mode = -1;            // diagnostics (SEVENN_TC_MMA_MODE bits: 1 plain TF32, 2 no stores, 4 no MMAs, 8 no proxy fence)
  if (mma_mode < 0) { const char* e = getenv("SEVENN_TC_MMA_MODE"); mma_mode = e ? atoi(e) : 0; }
  P.mode = mma_mode;
  P.vec4 = !((g.sCn & 3) || (g.sCm & 3) || ((uintptr_t)g.C & 15) || (g.C2 && ((uintptr_t)g.C2 & 15)) ||
             (g.P && ((uintptr_t)g.P & 15)));
  P.dbg = nullptr;
  static int tmema_ok = -1;
  if (tmema_ok < 0) { const char* e = getenv("SEVENN_TC_TMEMA"); tmema_ok = e ? atoi(e) : 1; }
  // TMEM-A kernel: large K, narrow N (same B image; it always streams it)
  *tca = tmema_ok && P.n_tiles == 1 && P.Nt <= 64 && P.KB >= 4 && (long long)P.KB * P.Nt * 256 > 84 * 1024;
  *smem = 1024 + (size_t)stages * stage_bytes + fixed + TC_EPI8_BYTES + 8 * (3 * stages + 9) + 16;
  if (*tca)
    *smem = 1024 + (size_t)TCA_NS * 2 * b_bytes + (size_t)2 * TCA_SLOTS * TC_BM * 128 + TC_EPI_BYTES + 8 * (2 * TCA_NS + 5) + 16;
  return 0;
}

static int tc_sm_count() {
  static int sms = 0;
  if (!sms) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); }
  return sms;
}

template <typename K>
static int tc_set_smem_attr(K kern, bool* done) {
  if (!*done) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return (int)e;
    *done = true;
  }
  return 0;
}

extern "C" int sevenn_gemm_tc_grouped(int n, const sevenn_gemm_args* args, const float* const* Bimgs, void* stream) {
  if (n <= 0) return 0;
  if (n > TC_MAX_GROUP) return SEVENN_ERR_BAD_ARG;
  if (n == 1 && args[0].M > 0 && args[0].N > 0) {      // a problem on its own may take the TMEM-A kernel (large K, N <= 64)
    TcParams P; size_t sm; bool tca;
    int rc = tc_prepare(&args[0], Bimgs[0], P, &sm, &tca);
    if (rc) return rc;
    if (tca) return sevenn_gemm_tc(&args[0], Bimgs[0], stream);
  }
  if (n > 1) {
    // Sharing a launch pays for problems that cannot fill the GPU on their own (their launch / TMEM-allocation /
    // pipeline-fill latencies overlap).  A problem with >= 2 tiles per SM runs alone on all SMs instead: a static
    // split of the CTAs between problems of different K / N per tile is bound by its slowest member (measured on the
    // l3i5 self-interaction at 32,768 atoms: 3.6 ms grouped vs 1.0 ms as separate launches).
    sevenn_gemm_args small[TC_MAX_GROUP];
    const float* small_img[TC_MAX_GROUP];
    int ns = 0;
    const int sms = tc_sm_count();
    for (int i = 0; i < n; ++i) {
      int Nt, n_tiles, KB, res;
      if (args[i].M <= 0 || args[i].N <= 0) continue;
      if (tc_plan(args[i].N, args[i].K, &Nt, &n_tiles, &KB, &res)) return SEVENN_ERR_BAD_ARG;
      const long long t = ((args[i].M + TC_BM - 1) / TC_BM) * n_tiles;
      if (t >= 2LL * sms) {
        int rc = sevenn_gemm_tc_grouped(1, &args[i], &Bimgs[i], stream);
        if (rc) return rc;
      } else {
        small[ns] = args[i]; small_img[ns] = Bimgs[i]; ++ns;
      }
    }
    if (ns < n) return ns ? sevenn_gemm_tc_grouped(ns, small, small_img, stream) : 0;
  }
  TcGroup G;
  G.n = 0;
  size_t smem = 0;
  long long tiles[TC_MAX_GROUP], total = 0;
  const int epi = args[0].epi;
  for (int i = 0; i < n; ++i) {
    if (args[i].M <= 0 || args[i].N <= 0) continue;
    if (args[i].epi != epi) return SEVENN_ERR_BAD_ARG;
    size_t sm; bool tca;
    int rc = tc_prepare(&args[i], Bimgs[i], G.p[G.n], &sm, &tca);
    if (rc) return rc;
    // members of a group always run the shared-memory-A kernel
    const size_t b_bytes = (size_t)G.p[G.n].Nt * 128;
    const size_t stage_bytes = 2 * TC_BM * 128 + (G.p[G.n].resident ? 0 : 2 * b_bytes);
    const size_t fixed = G.p[G.n].resident ? (size_t)G.p[G.n].KB * 2 * b_bytes : 0;
    sm = 1024 + (size_t)G.p[G.n].stages * stage_bytes + fixed + TC_EPI8_BYTES + 8 * (3 * G.p[G.n].stages + 9) + 16;
    if (sm > smem) smem = sm;
    tiles[G.n] = G.p[G.n].m_tiles * G.p[G.n].n_tiles;
    total += tiles[G.n];
    ++G.n;
  }
  if (G.n == 0) return 0;
  const int sms = tc_sm_count();
  int start = 0;
  for (int i = 0; i < G.n; ++i) {
    const TcParams& P = G.p[i];
    long long gx = total <= sms ? P.m_tiles : (long long)((double)sms * (double)tiles[i] / (double)total) / P.n_tiles;
    if (gx < 1) gx = 1;
    if (gx > P.m_tiles) gx = P.m_tiles;
    G.cta_start[i] = start;
    start += (int)gx * P.n_tiles;
  }
  G.cta_start[G.n] = start;
  void (*kern)(const TcGroup) = epi == 0 ? gemm_tc_kernel<0> : epi == 1 ? gemm_tc_kernel<1>
                              : epi == 2 ? gemm_tc_kernel<2> : gemm_tc_kernel<3>;
  static bool attr_set[4] = {false, false, false, false};
  int rc = tc_set_smem_attr(kern, &attr_set[epi]);
  if (rc) return rc;
  kern<<<(unsigned)start, TC_THREADS, smem, (cudaStream_t)stream>>>(G);
  return (int)cudaGetLastError();
}

extern "C" int sevenn_gemm_tc(const sevenn_gemm_args* args, const float* Bimg, void* stream) {
  if (args->M <= 0 || args->N <= 0) return 0;
  TcParams P;
  size_t smem; bool tca;
  int rc = tc_prepare(args, Bimg, P, &smem, &tca);
  if (rc) return rc;
  static int dbg_on = -1;
  if (dbg_on < 0) dbg_on = getenv("SEVENN_TC_DBG") != nullptr;
  const bool dbg = dbg_on != 0;
  if (!tca && !dbg) return sevenn_gemm_tc_grouped(1, args, &Bimg, stream);
  const sevenn_gemm_args& g = P.g;
  static long long* dbg_dev = nullptr;
  if (dbg && !dbg_dev) cudaMalloc(&dbg_dev, 3 * 6 * 8 * sizeof(long long));
  if (dbg) cudaMemset(dbg_dev, 0, 3 * 6 * 8 * sizeof(long long));
  P.dbg = dbg ? dbg_dev : nullptr;
  if (tca) {
    void (*kern)(const TcParams) = g.epi == 0 ? gemm_tca_kernel<0> : g.epi == 1 ? gemm_tca_kernel<1>
                                 : g.epi == 2 ? gemm_tca_kernel<2> : gemm_tca_kernel<3>;
    static bool attr_set[4] = {false, false, false, false};
    rc = tc_set_smem_attr(kern, &attr_set[g.epi]);
    if (rc) return rc;
    long long gx = tc_sm_count();
    if (gx > P.m_tiles) gx = P.m_tiles;
    kern<<<(unsigned)gx, TCA_THREADS, smem, (cudaStream_t)stream>>>(P);
  } else {
    TcGroup G;
    G.n = 1; G.p[0] = P;
    long long gx = tc_sm_count() / P.n_tiles;
    if (gx < 1) gx = 1;
    if (gx > P.m_tiles) gx = P.m_tiles;
    G.cta_start[0] = 0; G.cta_start[1] = (int)gx * P.n_tiles;
    void (*kern)(const TcGroup) = g.epi == 0 ? gemm_tc_kernel<0> : g.epi == 1 ? gemm_tc_kernel<1>
                                : g.epi == 2 ? gemm_tc_kernel<2> : gemm_tc_kernel<3>;
    static bool attr_set2[4] = {false, false, false, false};
    rc = tc_set_smem_attr(kern, &attr_set2[g.epi]);
    if (rc) return rc;
    kern<<<(unsigned)G.cta_start[1], TC_THREADS, smem, (cudaStream_t)stream>>>(G);
  }
  if (dbg) {
    long long h[3 * 6 * 8];
    cudaDeviceSynchronize();
    cudaMemcpy(h, dbg_dev, sizeof(h), cudaMemcpyDeviceToHost);
    long long t0 = h[0];
    for (int k = 0; k < 3 * 6 * 8; ++k) if (h[k] && h[k] < t0) t0 = h[k];
    fprintf(stderr, "TCDBG M=%lld N=%d K=%d stages=%d resident=%d KB=%d tca=%d\n", g.M, g.N, g.K, P.stages, P.resident, P.KB, (int)tca);
    const char* names[3] = {"prod/grp0", "mma", "epi/grp1"};
    for (int r = 0; r < 3; ++r) for (int it = 0; it < 6; ++it) {
      fprintf(stderr, "  %s %d:", names[r], it);
      for (int e = 0; e < 8; ++e) fprintf(stderr, " %8lld", h[(r * 6 + it) * 8 + e] ? h[(r * 6 + it) * 8 + e] - t0 : -1);
      fprintf(stderr, "\n");
    }
  }
  return (int)cudaGetLastError();
}
#endif

#ifdef SEVENN_EMU
// the emulation build has no tensor cores: same plan geometry, GEMM through the fp32 loops (tests only)
extern "C" int sevenn_gemm_tc_plan(int N, int K, int* Nt, int* n_tiles, int* KB, long long* img_floats) {
  if (N <= 0 || K <= 0) return SEVENN_ERR_BAD_ARG;
  *Nt = 16; *n_tiles = (N + 15) / 16; *KB = (K + 31) / 32;
  *img_floats = (long long)(*n_tiles) * (*KB) * 2 * 16 * 32;
  return 0;
}
extern "C" int sevenn_gemm_tc_image(const float* W, int K, int N, float* img, void* stream) {
  (void)W; (void)K; (void)N; (void)img; (void)stream; return 0;   // the emulation arm ignores images
}
extern "C" int sevenn_gemm_tc(const sevenn_gemm_args* args, const float* Bimg, void* stream) {
  (void)Bimg; return sevenn_gemm_f32(args, stream);
}
extern "C" int sevenn_gemm_tc_grouped(int n, const sevenn_gemm_args* args, const float* const* Bimgs, void* stream) {
  (void)Bimgs;
  for (int i = 0; i < n; ++i) { int rc = sevenn_gemm_f32(&args[i], stream); if (rc) return rc; }
  return 0;
}
#endif

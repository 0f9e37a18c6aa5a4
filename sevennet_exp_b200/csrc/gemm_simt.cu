// fp32 SIMT GEMM with (node, m)-strided rows and fused epilogues.  C[(n,m), :] = epi(A[(n,m), :K] @ B[K, N]).
//
// Serves every dense contraction of the path in exact fp32: e3nn o3.Linear blocks in ir_mul layout
// (sevenn/nn/linear.py:14-100, self_connection.py:68-108), the radial MLP e3nn FullyConnectedNet
// (sevenn/nn/convolution.py:92-101,121) and their data-gradients (B passed transposed by the host).
// This is the exact-fp32 arm; gemm_tc.cu is the tcgen05 arm selected for large M.
//
// Row r = n*R + m lives at A + n*sAn + m*sAm (R = 2l+1 rows per node for an irrep block; R = 1 for
// edge matrices).  Epilogues: 0 store, 1 accumulate, 2 store pre-activation to C and cst*act() to C2,
// 3 store acc * cst*act'(P) where P has C's strides.
#include "sevenn_common.cuh"
#include "../../include/sevenn_b200.h"

#ifndef SEVENN_EMU
#define BK 16

__device__ __forceinline__ float g_act(int id, float x) {
  switch (id) {
    case 0: return x / (1.f + expf(-x));
    case 1: return tanhf(x);
    case 2: return fabsf(x);
    case 3: return x > 0.f ? x : 0.f;
    case 4: return 1.f / (1.f + expf(-x));
    default: return (x > 20.f ? x : log1pf(expf(x))) - 0.6931471805599453f;
  }
}
__device__ __forceinline__ float g_dact(int id, float x) {
  switch (id) {
    case 0: { const float s = 1.f / (1.f + expf(-x)); return s * (1.f + x * (1.f - s)); }
    case 1: { const float t = tanhf(x); return 1.f - t * t; }
    case 2: return x > 0.f ? 1.f : (x < 0.f ? -1.f : 0.f);
    case 3: return x > 0.f ? 1.f : 0.f;
    case 4: { const float s = 1.f / (1.f + expf(-x)); return s * (1.f - s); }
    default: return 1.f / (1.f + expf(-x));
  }
}

// 256 threads as a 16 x 16 grid, TM x TN outputs per thread: tile (16 TM) x (16 TN).  The next k-slab is fetched into
// registers while the current one is multiplied (one barrier pair per slab, global latency hidden).  Small problems
// take the small tiles: a 64 x 64 tile with K = 224 keeps ONE SM busy for ~4 us while 140 others idle, which is what a
// 64-atom system's node linears consisted of.
template <int TM, int TN, int D>
__device__ __forceinline__ void gemm_simt_body(const sevenn_gemm_args& g, int bx, int by) {
  constexpr int TBM = 16 * TM, TBN = 16 * TN;
  constexpr int NA = (TBM * BK + 255) / 256, NB = (TBN * BK + 255) / 256;
  __shared__ float As[BK][TBM + 4];
  __shared__ float Bs[BK][TBN + 4];
  const int tid = threadIdx.x;
  const int ty = tid / 16, tx = tid % 16;
  const long long row0 = (long long)bx * TBM;
  const int col0 = by * TBN;
  float acc[TM][TN];
#pragma unroll
  for (int a = 0; a < TM; ++a)
#pragma unroll
    for (int b = 0; b < TN; ++b) acc[a][b] = 0.f;

  // A loader: element q of this thread -> (row = idx / BK, k = idx % BK), idx = tid + 256 q (coalesced along K)
  const float* aptr[NA];
#pragma unroll
  for (int q = 0; q < NA; ++q) {
    const int idx = tid + 256 * q;
    const long long ar = row0 + idx / BK;
    aptr[q] = (idx < TBM * BK && ar < g.M) ? g.A + (ar / g.R) * g.sAn + (ar % g.R) * (long long)g.sAm + idx % BK : nullptr;
  }
  // D k-slabs in flight in registers: a 16 x 32 tile does ~100 clk of arithmetic per slab, so with one slab of
  // prefetch every slab cost a full global-memory round trip (K = 384: 24 of them, ~15 us for a 64-atom system)
  float ra[D][NA], rb[D][NB];
  auto fetch = [&](float (&fa)[NA], float (&fb)[NB], int k0) {
#pragma unroll
    for (int q = 0; q < NA; ++q) {
      const int k = k0 + (tid + 256 * q) % BK;
      fa[q] = (aptr[q] != nullptr && k < g.K) ? __ldg(aptr[q] + k0) : 0.f;
    }
#pragma unroll
    for (int q = 0; q < NB; ++q) {
      const int idx = tid + 256 * q;
      const int k = k0 + idx / TBN, c = col0 + idx % TBN;
      fb[q] = (idx < TBN * BK && k < g.K && c < g.N) ? __ldg(g.B + (size_t)k * g.ldb + c) : 0.f;
    }
  };
#pragma unroll
  for (int d = 0; d < D; ++d) fetch(ra[d], rb[d], d * BK);
  for (int kk = 0; kk < g.K; kk += D * BK) {
#pragma unroll
    for (int d = 0; d < D; ++d) {
      const int k0 = kk + d * BK;
      if (k0 >= g.K) break;
#pragma unroll
      for (int q = 0; q < NA; ++q) {
        const int idx = tid + 256 * q;
        if (idx < TBM * BK) As[idx % BK][idx / BK] = ra[d][q];
      }
#pragma unroll
      for (int q = 0; q < NB; ++q) {
        const int idx = tid + 256 * q;
        if (idx < TBN * BK) Bs[idx / TBN][idx % TBN] = rb[d][q];
      }
      __syncthreads();
      if (k0 + D * BK < g.K) fetch(ra[d], rb[d], k0 + D * BK);
#pragma unroll
      for (int k = 0; k < BK; ++k) {
        float a[TM], b[TN];
#pragma unroll
        for (int q = 0; q < TM; ++q) a[q] = As[k][ty * TM + q];
#pragma unroll
        for (int q = 0; q < TN; ++q) b[q] = Bs[k][tx * TN + q];
#pragma unroll
        for (int i = 0; i < TM; ++i)
#pragma unroll
          for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
      }
      __syncthreads();
    }
  }
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    const long long r = row0 + ty * TM + i;
    if (r >= g.M) continue;
    const long long off = (r / g.R) * g.sCn + (r % g.R) * (long long)g.sCm;
#pragma unroll
    for (int j = 0; j < TN; ++j) {
      const int c = col0 + tx * TN + j;
      if (c >= g.N) continue;
      const float v = acc[i][j];
      switch (g.epi) {
        case 0: g.C[off + c] = v; break;
        case 1: g.C[off + c] += v; break;
        case 2: g.C[off + c] = v; g.C2[off + c] = g.cst * g_act(g.act, v); break;
        default: g.C[off + c] = v * g.cst * g_dact(g.act, g.P[off + c]); break;
      }
    }
  }
}

template <int TM, int TN>
__global__ void __launch_bounds__(256) gemm_simt_kernel(sevenn_gemm_args g) {
  gemm_simt_body<TM, TN, (TM * TN <= 2 ? 4 : 1)>(g, (int)blockIdx.x, (int)blockIdx.y);
}

// Several independent small problems in one launch (the per-l blocks of one or two o3.Linear on a small system, where a
// launch costs more than the arithmetic): 16 x 32 tiles, CTA -> (problem, tile) through a prefix table.
struct SimtGroup {
  int n;
  int start[SEVENN_GEMM_GROUP_MAX + 1];
  int gx[SEVENN_GEMM_GROUP_MAX];
  sevenn_gemm_args p[SEVENN_GEMM_GROUP_MAX];
};
__global__ void __launch_bounds__(256) gemm_simt_grouped_kernel(const SimtGroup G) {
  int pi = 0;
  while (pi + 1 < G.n && (int)blockIdx.x >= G.start[pi + 1]) ++pi;
  const int local = (int)blockIdx.x - G.start[pi];
  gemm_simt_body<1, 2, 4>(G.p[pi], local % G.gx[pi], local / G.gx[pi]);
}

template <int TM, int TN>
static int gemm_simt_launch(const sevenn_gemm_args& g, cudaStream_t st) {
  dim3 grid((unsigned)((g.M + 16 * TM - 1) / (16 * TM)), (unsigned)((g.N + 16 * TN - 1) / (16 * TN)));
  gemm_simt_kernel<TM, TN><<<grid, 256, 0, st>>>(g);
  return (int)cudaGetLastError();
}

extern "C" int sevenn_gemm_f32(const sevenn_gemm_args* args, void* stream) {
  const sevenn_gemm_args g = *args;
  if (g.M <= 0 || g.N <= 0) return 0;
  if (g.K <= 0 || g.R <= 0 || g.epi < 0 || g.epi > 3) return SEVENN_ERR_BAD_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  // the largest tile that still gives every SM work (results do not depend on the tile: each output is one k-ordered sum)
  const long long t64 = ((g.M + 63) / 64) * ((g.N + 63) / 64), t32 = ((g.M + 31) / 32) * ((g.N + 31) / 32);
  if (t64 >= 148) return gemm_simt_launch<4, 4>(g, st);
  if (t32 >= 148) return gemm_simt_launch<2, 2>(g, st);
  return gemm_simt_launch<1, 2>(g, st);
}

extern "C" int sevenn_gemm_f32_grouped(int n, const sevenn_gemm_args* args, void* stream) {
  if (n <= 0) return 0;
  if (n > SEVENN_GEMM_GROUP_MAX) return SEVENN_ERR_BAD_ARG;
  SimtGroup G;
  G.n = 0;
  int start = 0;
  long long big = 0;
  for (int i = 0; i < n; ++i) {
    const sevenn_gemm_args& g = args[i];
    if (g.M <= 0 || g.N <= 0) continue;
    if (g.K <= 0 || g.R <= 0 || g.epi < 0 || g.epi > 3) return SEVENN_ERR_BAD_ARG;
    const long long gx = (g.M + 15) / 16, gy = (g.N + 31) / 32;
    big += gx * gy;
    if (big > SEVENN_GEMM_GROUP_MAX_TILES) break;
    G.p[G.n] = g; G.gx[G.n] = (int)gx; G.start[G.n] = start;
    start += (int)(gx * gy);
    ++G.n;
  }
  if (big > SEVENN_GEMM_GROUP_MAX_TILES) {          // not a small-system group: one launch each, tile chosen per problem
    for (int i = 0; i < n; ++i) { int rc = sevenn_gemm_f32(&args[i], stream); if (rc) return rc; }
    return 0;
  }
  if (G.n == 0) return 0;
  G.start[G.n] = start;
  gemm_simt_grouped_kernel<<<(unsigned)start, 256, 0, (cudaStream_t)stream>>>(G);
  return (int)cudaGetLastError();
}
#endif

#ifdef SEVENN_EMU
// tests only: naive loops with the same argument contract
static float e_act(int id, float x) {
  switch (id) { case 0: return x / (1.f + expf(-x)); case 1: return tanhf(x); case 2: return fabsf(x);
    case 3: return x > 0.f ? x : 0.f; case 4: return 1.f / (1.f + expf(-x));
    default: return (x > 20.f ? x : log1pf(expf(x))) - 0.6931471805599453f; }
}
static float e_dact(int id, float x) {
  switch (id) { case 0: { const float s = 1.f / (1.f + expf(-x)); return s * (1.f + x * (1.f - s)); }
    case 1: { const float t = tanhf(x); return 1.f - t * t; } case 2: return x > 0.f ? 1.f : (x < 0.f ? -1.f : 0.f);
    case 3: return x > 0.f ? 1.f : 0.f; case 4: { const float s = 1.f / (1.f + expf(-x)); return s * (1.f - s); }
    default: return 1.f / (1.f + expf(-x)); }
}
extern "C" int sevenn_gemm_f32(const sevenn_gemm_args* args, void* stream) {
  const sevenn_gemm_args g = *args;
  if (g.M <= 0 || g.N <= 0) return 0;
  if (g.K <= 0 || g.R <= 0 || g.epi < 0 || g.epi > 3) return SEVENN_ERR_BAD_ARG;
  for (long long r = 0; r < g.M; ++r) {
    const float* a = g.A + (r / g.R) * g.sAn + (r % g.R) * (long long)g.sAm;
    const long long off = (r / g.R) * g.sCn + (r % g.R) * (long long)g.sCm;
    for (int c = 0; c < g.N; ++c) {
      float v = 0.f;
      for (int k = 0; k < g.K; ++k) v = fmaf(a[k], g.B[(size_t)k * g.ldb + c], v);
      switch (g.epi) {
        case 0: g.C[off + c] = v; break;
        case 1: g.C[off + c] += v; break;
        case 2: g.C[off + c] = v; g.C2[off + c] = g.cst * e_act(g.act, v); break;
        default: g.C[off + c] = v * g.cst * e_dact(g.act, g.P[off + c]); break;
      }
    }
  }
  return 0;
}
extern "C" int sevenn_gemm_f32_grouped(int n, const sevenn_gemm_args* args, void* stream) {
  if (n > SEVENN_GEMM_GROUP_MAX) return SEVENN_ERR_BAD_ARG;
  for (int i = 0; i < n; ++i) { int rc = sevenn_gemm_f32(&args[i], stream); if (rc) return rc; }
  return 0;
}
#endif

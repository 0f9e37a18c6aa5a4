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
#define BM 64
#define BN 64
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

__global__ void __launch_bounds__(256) gemm_simt_kernel(sevenn_gemm_args g) {
  __shared__ float As[BK][BM + 4];
  __shared__ float Bs[BK][BN + 4];
  const int tid = threadIdx.x;
  const int ty = tid / 16, tx = tid % 16;
  const long long row0 = (long long)blockIdx.x * BM;
  const int col0 = blockIdx.y * BN;
  float acc[4][4];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;

  // A loader: thread -> (row = tid/4, k4 = (tid%4)*4)
  const int arow = tid / 4, ak = (tid % 4) * 4;
  const long long ar = row0 + arow;
  const float* aptr = nullptr;
  if (ar < g.M) aptr = g.A + (ar / g.R) * g.sAn + (ar % g.R) * (long long)g.sAm;
  // B loader: thread -> (k = tid/16, c4 = (tid%16)*4)
  const int bk = tid / 16, bc = (tid % 16) * 4;

  for (int k0 = 0; k0 < g.K; k0 += BK) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int k = k0 + ak + q;
      As[ak + q][arow] = (aptr && k < g.K) ? __ldg(aptr + k) : 0.f;
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int k = k0 + bk, c = col0 + bc + q;
      Bs[bk][bc + q] = (k < g.K && c < g.N) ? __ldg(g.B + (size_t)k * g.ldb + c) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      float a[4], b[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) { a[q] = As[k][ty * 4 + q]; b[q] = Bs[k][tx * 4 + q]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const long long r = row0 + ty * 4 + i;
    if (r >= g.M) continue;
    const long long off = (r / g.R) * g.sCn + (r % g.R) * (long long)g.sCm;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = col0 + tx * 4 + j;
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

extern "C" int sevenn_gemm_f32(const sevenn_gemm_args* args, void* stream) {
  const sevenn_gemm_args g = *args;
  if (g.M <= 0 || g.N <= 0) return 0;
  if (g.K <= 0 || g.R <= 0 || g.epi < 0 || g.epi > 3) return SEVENN_ERR_BAD_ARG;
  dim3 grid((unsigned)((g.M + BM - 1) / BM), (unsigned)((g.N + BN - 1) / BN));
  gemm_simt_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(g);
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
#endif

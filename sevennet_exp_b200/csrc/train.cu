// Kernels that only the training step needs (the loss contains forces and stress, so loss.backward()
// differentiates the reverse kernels -- sevenn/train/trainer.py:156-170, sevenn/nn/force_output.py:173-178
// `create_graph=self.training`).  Every op of the path is multilinear or elementwise, so its double backward is a
// composition of the first-order kernels (sevennet_exp_b200/nn/functional.py); what was missing are
//   gemm_tn      : weight gradients  G[K,N] = sum_rows A[row,:K]^T B[row,:N]   (o3.Linear, radial MLP; optional type filter
//                  for the species-indexed nequip self connection), deterministic two-stage reduction
//   act_mul      : out = m * cst * f^(order)(x), order 0..2 of the activation (radial MLP second-order chain)
//   gate_bwd2    : derivative of gate_bwd with respect to (pre-activation, upstream gradient)
//   edge_train   : directional derivative of (Y, emb) along a per-edge vector and the Bessel-coefficient gradients
//                  (edge_embedding.py:96-100: the coefficients are trainable)
//   adam_step    : torch.optim.Adam's update rule on one flat parameter (train/optim.py: 'adam')
// All deterministic (no atomics).  -DSEVENN_EMU builds the same arithmetic for the CPU tests.
#include "sevenn_common.cuh"
#include "../../include/sevenn_b200.h"

// ------------------------------------------------------------------------------------------------ activations
// f, f', f'' of the e3nn activations (ids as in spec.ACT_ID); the normalize2mom constant is applied by the caller
HD float act_order(int id, int order, float x) {
  switch (id) {
    case 0: {   // silu
      const float s = 1.f / (1.f + expf(-x));
      if (order == 0) return x * s;
      if (order == 1) return s * (1.f + x * (1.f - s));
      return s * (1.f - s) * (2.f + x * (1.f - 2.f * s));
    }
    case 1: {   // tanh
      const float t = tanhf(x);
      if (order == 0) return t;
      if (order == 1) return 1.f - t * t;
      return -2.f * t * (1.f - t * t);
    }
    case 2:     // abs
      if (order == 0) return fabsf(x);
      if (order == 1) return x > 0.f ? 1.f : (x < 0.f ? -1.f : 0.f);
      return 0.f;
    case 3:     // relu
      if (order == 0) return x > 0.f ? x : 0.f;
      if (order == 1) return x > 0.f ? 1.f : 0.f;
      return 0.f;
    case 4: {   // sigmoid
      const float s = 1.f / (1.f + expf(-x));
      if (order == 0) return s;
      if (order == 1) return s * (1.f - s);
      return s * (1.f - s) * (1.f - 2.f * s);
    }
    default: {  // shifted softplus
      const float s = 1.f / (1.f + expf(-x));
      if (order == 0) return (x > 20.f ? x : log1pf(expf(x))) - 0.6931471805599453f;
      if (order == 1) return s;
      return s * (1.f - s);
    }
  }
}

GKERNEL(act_mul_kernel)(const float* __restrict__ x, const float* __restrict__ m, long long n, int act, float cst,
                        int order, float* __restrict__ out) {
  GRID_STRIDE_LOOP(i, n) {
    const float v = cst * act_order(act, order, x[i]);
    out[i] = m != nullptr ? m[i] * v : v;
  }
}
extern "C" int sevenn_act_mul(const float* x, const float* m, long long n, int act, float cst, int order, float* out,
                              void* stream) {
  if (n <= 0) return 0;
  if (order < 0 || order > 2 || act < 0 || act > 5) return SEVENN_ERR_BAD_ARG;
  LAUNCH1D(act_mul_kernel, n, stream)(x, m, n, act, cst, order, out);
  return LAUNCH_STATUS();
}

// ------------------------------------------------------------------------------------------------ gate, second order
// gate_bwd:  dpre_s = dout_s a'(s) (scalars);  dg_u = a'(g_u) sum_m dout[u,m] x[u,m],  dx[u,m] = a(g_u) dout[u,m] (gated).
// Given c (cotangent of dpre) this emits  d_dout = d<c, dpre>/d dout  and  d_pre = d<c, dpre>/d pre.
struct GateTable2 { sevenn_gate_piece p[SEVENN_MAX_GATE_PIECES]; int n; };   // passed by value (kernel parameter space)
GKERNEL(gate_bwd2_kernel)(const float* __restrict__ pre, int din, const float* __restrict__ dout, int dout_dim,
                          const float* __restrict__ c, long long n, GateTable2 T, float* __restrict__ d_dout,
                          float* __restrict__ d_pre) {
  const int n_pieces = T.n;
  GRID_STRIDE_LOOP(t, n * n_pieces) {
    const long long row = t / n_pieces;
    const sevenn_gate_piece P = T.p[t % n_pieces];
    const float* pr = pre + row * din;
    const float* cr = c + row * din;
    const float* dor = dout + row * dout_dim;
    float* ddo = d_dout + row * dout_dim;
    float* dp = d_pre + row * din;
    if (P.kind == 0) {
      for (int u = 0; u < P.mul; ++u) {
        const float s = pr[P.in_off + u];
        const float a1 = P.cst * act_order(P.act, 1, s), a2 = P.cst * act_order(P.act, 2, s);
        ddo[P.out_off + u] = cr[P.in_off + u] * a1;
        dp[P.in_off + u] = cr[P.in_off + u] * dor[P.out_off + u] * a2;
      }
    } else {
      const int R = 2 * P.l + 1;
      for (int u = 0; u < P.mul; ++u) {
        const float g = pr[P.gate_off + u];
        const float a0 = P.cst * act_order(P.act, 0, g), a1 = P.cst * act_order(P.act, 1, g),
                    a2 = P.cst * act_order(P.act, 2, g);
        const float cg = cr[P.gate_off + u];
        float s_dx = 0.f, s_cd = 0.f;
        for (int m = 0; m < R; ++m) {
          const int ii = P.in_off + m * P.in_stride + u, oo = P.out_off + m * P.out_stride + u;
          const float x = pr[ii], d = dor[oo], cx = cr[ii];
          s_dx += d * x;
          s_cd += cx * d;
          ddo[oo] = cg * a1 * x + cx * a0;
          dp[ii] = cg * a1 * d;
        }
        dp[P.gate_off + u] = cg * a2 * s_dx + a1 * s_cd;
      }
    }
  }
}
extern "C" int sevenn_gate_bwd2(const float* pre, int din, const float* dout, int dout_dim, const float* c, long long n,
                                const sevenn_gate_piece* pieces /* host */, int n_pieces, float* d_dout, float* d_pre,
                                void* stream) {
  if (n <= 0) return 0;
  if (n_pieces < 1 || n_pieces > SEVENN_MAX_GATE_PIECES) return SEVENN_ERR_BAD_ARG;
  GateTable2 T;
  T.n = n_pieces;
  for (int q = 0; q < n_pieces; ++q) T.p[q] = pieces[q];
  LAUNCH1D(gate_bwd2_kernel, n * n_pieces, stream)(pre, din, dout, dout_dim, c, n, T, d_dout, d_pre);
  return LAUNCH_STATUS();
}

// ------------------------------------------------------------------------------------------------ edge embedding
struct Dual { float v, d; };
HD Dual operator+(Dual a, Dual b) { return {a.v + b.v, a.d + b.d}; }
HD Dual operator-(Dual a, Dual b) { return {a.v - b.v, a.d - b.d}; }
HD Dual operator*(Dual a, Dual b) { return {a.v * b.v, a.v * b.d + a.d * b.v}; }
HD Dual operator*(float a, Dual b) { return {a * b.v, a * b.d}; }

#define T_SQ3 1.7320508075688772f
#define T_SQ5 2.23606797749979f
#define T_SQ15 3.872983346207417f
#define T_SQ7 2.6457513110645907f
// the polynomials of csrc/edge.cu::sh_eval on dual numbers: .d is the directional derivative
HD void sh_eval_dual(int lmax, Dual x, Dual y, Dual z, Dual* Y) {
  Y[0] = {1.f, 0.f};
  if (lmax >= 1) { Y[1] = T_SQ3 * x; Y[2] = T_SQ3 * y; Y[3] = T_SQ3 * z; }
  if (lmax >= 2) {
    Y[4] = T_SQ15 * (x * z); Y[5] = T_SQ15 * (x * y); Y[6] = T_SQ5 * (y * y - 0.5f * (x * x + z * z));
    Y[7] = T_SQ15 * (y * z); Y[8] = (0.5f * T_SQ15) * (z * z - x * x);
  }
  if (lmax >= 3) {
    const float a = T_SQ7 * 0.7905694150420949f, b = T_SQ7 * 3.872983346207417f, c = T_SQ7 * 0.6123724356957945f,
                h = T_SQ7 * 0.5f, e = T_SQ7 * 1.9364916731037085f;
    const Dual x2 = x * x, y2 = y * y, z2 = z * z;
    Y[9] = a * (x * (3.f * z2 - x2)); Y[10] = b * (x * y * z); Y[11] = c * (x * (4.f * y2 - x2 - z2));
    Y[12] = h * (y * (2.f * y2 - 3.f * x2 - 3.f * z2)); Y[13] = c * (z * (4.f * y2 - x2 - z2));
    Y[14] = e * (y * (z2 - x2)); Y[15] = a * (z * (z2 - 3.f * x2));
  }
}

// envelope value, first and second derivative (csrc/edge.cu::cutoff_eval + one more order)
HD void cutoff_eval2(int kind, float rc, float p, float r_on, float r, float* f, float* df) {
  if (kind == 0) {
    const float x = r / rc;
    const float c0 = (p + 1.f) * (p + 2.f) * 0.5f, c1 = p * (p + 2.f), c2 = p * (p + 1.f) * 0.5f;
    const float xp1 = powf(x, p - 1.f);
    const float xp = xp1 * x, xq = xp * x, xr = xq * x;
    *f = 1.f - c0 * xp + c1 * xq - c2 * xr;
    *df = (-c0 * p * xp1 + c1 * (p + 1.f) * xp - c2 * (p + 2.f) * xq) / rc;
  } else {
    if (r < r_on) { *f = 1.f; *df = 0.f; return; }
    const float c2 = rc * rc, on2 = r_on * r_on, r2 = r * r;
    const float den = (c2 - on2) * (c2 - on2) * (c2 - on2);
    const float a = c2 - r2, b = c2 + 2.f * r2 - 3.f * on2;
    *f = a * a * b / den;
    *df = (-4.f * r * a * b + 4.f * r * a * a) / den;
  }
}

// c == nullptr : G[e,k] = demb[e,k] d emb_k / d coeff_k                       (first-order coefficient gradient)
// c != nullptr : JY[e,:] = dY/dvec . c_e, Jemb[e,:] = d emb/dvec . c_e,
//                G[e,k]  = demb[e,k] d/d coeff_k (d emb_k/dvec . c_e)          (what EdgeBwd's backward needs)
GKERNEL(edge_train_kernel)(const float* __restrict__ vec, long long n, int lmax, int normalize, int nb, int kind,
                           float rc, float p, float r_on, const float* __restrict__ coeffs,
                           const float* __restrict__ demb, const float* __restrict__ c, float* __restrict__ JY,
                           float* __restrict__ Jemb, float* __restrict__ G) {
  const int S = (lmax + 1) * (lmax + 1);
  GRID_STRIDE_LOOP(e, n) {
    const float x = vec[3 * e], y = vec[3 * e + 1], z = vec[3 * e + 2];
    const float r = sqrtf(x * x + y * y + z * z), ir = 1.f / r;
    float fc, dfc;
    cutoff_eval2(kind, rc, p, r_on, r, &fc, &dfc);
    const float pref = 2.f / rc;
    if (c == nullptr) {
      for (int k = 0; k < nb; ++k)
        G[e * nb + k] = (demb != nullptr ? demb[e * nb + k] : 1.f) * pref * cosf(LDGF(coeffs + k) * r) * fc;
      continue;
    }
    const float cx = c[3 * e], cy = c[3 * e + 1], cz = c[3 * e + 2];
    const float ux = x * ir, uy = y * ir, uz = z * ir;
    const float uc = ux * cx + uy * cy + uz * cz;            // dr along c
    Dual Yd[16];
    if (normalize) {
      // u = v / r: du = (c - u (u.c)) / r
      sh_eval_dual(lmax, Dual{ux, (cx - ux * uc) * ir}, Dual{uy, (cy - uy * uc) * ir}, Dual{uz, (cz - uz * uc) * ir}, Yd);
    } else {
      sh_eval_dual(lmax, Dual{x, cx}, Dual{y, cy}, Dual{z, cz}, Yd);
    }
    for (int k = 0; k < S; ++k) JY[e * S + k] = Yd[k].d;
    for (int k = 0; k < nb; ++k) {
      const float ck = LDGF(coeffs + k);
      const float s = sinf(ck * r), co = cosf(ck * r);
      const float b = s * ir, db = (ck * co - b) * ir;
      Jemb[e * nb + k] = pref * (db * fc + b * dfc) * uc;
      if (G != nullptr)
        G[e * nb + k] = (demb != nullptr ? demb[e * nb + k] : 0.f) * uc * pref * (-ck * s * fc + co * dfc);
    }
  }
}
extern "C" int sevenn_edge_train(const float* edge_vec, long long n_edges, int lmax, int normalize, int n_basis,
                                 int cutoff_kind, float rc, float p, float r_on, const float* bessel_coeffs,
                                 const float* demb, const float* c, float* JY, float* Jemb, float* G, void* stream) {
  if (n_edges <= 0) return 0;
  if (lmax < 0 || lmax > 3 || n_basis < 1) return SEVENN_ERR_BAD_ARG;
  if (c == nullptr && G == nullptr) return SEVENN_ERR_BAD_ARG;
  if (c != nullptr && (JY == nullptr || Jemb == nullptr)) return SEVENN_ERR_BAD_ARG;
  LAUNCH1D(edge_train_kernel, n_edges, stream)(edge_vec, n_edges, lmax, normalize, n_basis, cutoff_kind, rc, p, r_on,
                                              bessel_coeffs, demb, c, JY, Jemb, G);
  return LAUNCH_STATUS();
}

// ------------------------------------------------------------------------------------------------ weight gradients
// G[K,N] (+)= scale * sum over rows r = (n, m), m < R, of A[r, :K]^T B[r, :N]; row r of A at A + n sAn + m sAm (like
// sevenn_gemm_args).  types != nullptr keeps only the nodes of species `type_sel` (species-indexed linears).
// Two stages, fixed order: split s sums its rows into part[s][K][N], then the splits are added in order.
#ifndef SEVENN_EMU
#define TN_T 64          // output tile: TN_T rows of K x (NJ * TN_T) columns of N per CTA
#define TN_R 16          // rows staged per iteration
// 256 threads as 16 (k) x 16 (n); each thread owns 4 k x (NJ x 4) n outputs: n = n0 + 64 j + 4 tx + {0..3}, so the 16-byte
// shared-memory reads of eight neighbouring threads cover 128 contiguous bytes (conflict free).  NJ = 2 (wide N: the
// [64, 960] radial-MLP weight gradients over all edges) does 32 FMAs per 3 vector loads; NJ = 1 keeps narrow N dense.
template <int NJ>
__global__ void __launch_bounds__(256) gemm_tn_kernel(const float* __restrict__ A, long long sAn, int sAm,
                                                      const float* __restrict__ B, long long sBn, int sBm, long long M,
                                                      int R, int K, int N, const int* __restrict__ types, int type_sel,
                                                      int n_split, float* __restrict__ part) {
  constexpr int TNW = NJ * TN_T;
  __shared__ __align__(16) float As[TN_R][TN_T + 4];
  __shared__ __align__(16) float Bs[TN_R][TNW + 4];
  const int k0 = blockIdx.x * TN_T, n0 = blockIdx.y * TNW, sp = blockIdx.z;
  const int tid = threadIdx.x, ty = tid / 16, tx = tid % 16;
  const long long per = (M + n_split - 1) / n_split;
  const long long r_begin = sp * per, r_end = (r_begin + per < M) ? r_begin + per : M;
  float acc[4][4 * NJ];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4 * NJ; ++j) acc[i][j] = 0.f;
  // loader: 4 elements of A and 4 NJ of B per thread and iteration; consecutive threads read consecutive columns
  const int lr = tid / 64, lc = tid % 64;          // rows lr, lr + 4, lr + 8, lr + 12; columns lc (+ 64 for NJ = 2)
  // (node, m) of this thread's four rows, advanced incrementally (a 64-bit division per element would cost more
  // instructions than the FMAs it feeds)
  long long node_q[4]; int m_q[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) { const long long r = r_begin + lr + 4 * q; node_q[q] = r / R; m_q[q] = (int)(r - node_q[q] * R); }
  const int step_n = TN_R / R, step_m = TN_R % R;
  float ra[4], rb[4][NJ];
  auto fetch = [&](long long r0) {            // next slab into registers while the current one is multiplied
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const long long r = r0 + lr + 4 * q;
      float a = 0.f, b[NJ];
#pragma unroll
      for (int j = 0; j < NJ; ++j) b[j] = 0.f;
      if (r < r_end) {
        const long long node = node_q[q];
        const int m = m_q[q];
        if (types == nullptr || types[node] == type_sel) {
          if (k0 + lc < K) a = __ldg(A + node * sAn + (long long)m * sAm + k0 + lc);
#pragma unroll
          for (int j = 0; j < NJ; ++j)
            if (n0 + 64 * j + lc < N) b[j] = __ldg(B + node * sBn + (long long)m * sBm + n0 + 64 * j + lc);
        }
      }
      ra[q] = a;
#pragma unroll
      for (int j = 0; j < NJ; ++j) rb[q][j] = b[j];
      node_q[q] += step_n; m_q[q] += step_m;
      if (m_q[q] >= R) { m_q[q] -= R; ++node_q[q]; }
    }
  };
  fetch(r_begin);
  for (long long r0 = r_begin; r0 < r_end; r0 += TN_R) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      As[lr + 4 * q][lc] = ra[q];
#pragma unroll
      for (int j = 0; j < NJ; ++j) Bs[lr + 4 * q][64 * j + lc] = rb[q][j];
    }
    __syncthreads();
    if (r0 + TN_R < r_end) fetch(r0 + TN_R);
#pragma unroll
    for (int rr = 0; rr < TN_R; ++rr) {
      const float4 a4 = *reinterpret_cast<const float4*>(&As[rr][ty * 4]);
      const float a[4] = {a4.x, a4.y, a4.z, a4.w};
      float b[4 * NJ];
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        const float4 b4 = *reinterpret_cast<const float4*>(&Bs[rr][64 * j + tx * 4]);
        b[4 * j] = b4.x; b[4 * j + 1] = b4.y; b[4 * j + 2] = b4.z; b[4 * j + 3] = b4.w;
      }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4 * NJ; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4 * NJ; ++j) {
      const int k = k0 + ty * 4 + i, nn = n0 + 64 * (j / 4) + tx * 4 + (j % 4);
      if (k < K && nn < N) part[((size_t)sp * K + k) * N + nn] = acc[i][j];
    }
}
__global__ void gemm_tn_reduce_kernel(const float* __restrict__ part, int n_split, long long KN, float scale,
                                      int accumulate, float* __restrict__ out) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < KN; i += (long long)gridDim.x * blockDim.x) {
    float s = 0.f;
    for (int sp = 0; sp < n_split; ++sp) s += part[(size_t)sp * KN + i];
    out[i] = accumulate ? out[i] + scale * s : scale * s;
  }
}
#endif
extern "C" int sevenn_gemm_tn_splits(long long M) {
  long long s = (M + 255) / 256;
  return (int)(s < 1 ? 1 : (s > 256 ? 256 : s));
}
extern "C" int sevenn_gemm_tn(const float* A, long long sAn, int sAm, const float* B, long long sBn, int sBm, long long M,
                              int R, int K, int N, const int* types, int type_sel, float scale, int accumulate,
                              float* part /* >= sevenn_gemm_tn_splits(M) * K * N floats */, float* out /* K x N */,
                              void* stream) {
  if (K <= 0 || N <= 0) return 0;
  if (R <= 0 || M < 0) return SEVENN_ERR_BAD_ARG;
#ifdef SEVENN_EMU
  (void)part; (void)stream;
  for (int k = 0; k < K; ++k)
    for (int n = 0; n < N; ++n) {
      float s = 0.f;
      for (long long r = 0; r < M; ++r) {
        const long long node = r / R;
        const int m = (int)(r % R);
        if (types != nullptr && types[node] != type_sel) continue;
        s = fmaf(A[node * sAn + (long long)m * sAm + k], B[node * sBn + (long long)m * sBm + n], s);
      }
      out[(size_t)k * N + n] = accumulate ? out[(size_t)k * N + n] + scale * s : scale * s;
    }
  return 0;
#else
  const int n_split = sevenn_gemm_tn_splits(M);
  if (N >= 96) {
    dim3 grid((K + TN_T - 1) / TN_T, (N + 2 * TN_T - 1) / (2 * TN_T), n_split);
    gemm_tn_kernel<2><<<grid, 256, 0, (cudaStream_t)stream>>>(A, sAn, sAm, B, sBn, sBm, M, R, K, N, types, type_sel,
                                                              n_split, part);
  } else {
    dim3 grid((K + TN_T - 1) / TN_T, (N + TN_T - 1) / TN_T, n_split);
    gemm_tn_kernel<1><<<grid, 256, 0, (cudaStream_t)stream>>>(A, sAn, sAm, B, sBn, sBm, M, R, K, N, types, type_sel,
                                                              n_split, part);
  }
  gemm_tn_reduce_kernel<<<sevenn_grid_for((long long)K * N), 256, 0, (cudaStream_t)stream>>>(part, n_split, (long long)K * N,
                                                                                             scale, accumulate, out);
  return (int)cudaGetLastError();
#endif
}

// ------------------------------------------------------------------------------------------------ optimizer
// torch.optim.Adam (no amsgrad, no weight decay unless wd != 0: L2 added to the gradient like torch):
//   m = b1 m + (1 - b1) g ;  v = b2 v + (1 - b2) g^2 ;  p -= lr / (1 - b1^t) * m / (sqrt(v) / sqrt(1 - b2^t) + eps)
GKERNEL(adam_kernel)(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                     long long n, float lr, float b1, float b2, float eps, float wd, float bc1, float bc2_sqrt) {
  GRID_STRIDE_LOOP(i, n) {
    float gi = g[i];
    if (wd != 0.f) gi += wd * p[i];
    const float mi = b1 * m[i] + (1.f - b1) * gi;
    const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
    m[i] = mi; v[i] = vi;
    p[i] -= (lr / bc1) * mi / (sqrtf(vi) / bc2_sqrt + eps);
  }
}
// the same update with the step count read on the device (CUDA-graph replay: the host cannot pass a new count)
GKERNEL(adam_dev_kernel)(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                         long long n, const int* __restrict__ step, float lr, float b1, float b2, float eps, float wd) {
  const float t = (float)*step;
  const float bc1 = 1.f - powf(b1, t), bc2_sqrt = sqrtf(1.f - powf(b2, t));
  GRID_STRIDE_LOOP(i, n) {
    float gi = g[i];
    if (wd != 0.f) gi += wd * p[i];
    const float mi = b1 * m[i] + (1.f - b1) * gi;
    const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
    m[i] = mi; v[i] = vi;
    p[i] -= (lr / bc1) * mi / (sqrtf(vi) / bc2_sqrt + eps);
  }
}
extern "C" int sevenn_adam_step_dev(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, long long n,
                                    const int* step_dev, float lr, float beta1, float beta2, float eps, float weight_decay,
                                    void* stream) {
  if (n <= 0) return 0;
  if (step_dev == nullptr) return SEVENN_ERR_BAD_ARG;
  LAUNCH1D(adam_dev_kernel, n, stream)(param, grad, exp_avg, exp_avg_sq, n, step_dev, lr, beta1, beta2, eps, weight_decay);
  return LAUNCH_STATUS();
}
extern "C" int sevenn_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, long long n, int step,
                                float lr, float beta1, float beta2, float eps, float weight_decay, void* stream) {
  if (n <= 0) return 0;
  if (step < 1) return SEVENN_ERR_BAD_ARG;
  const float bc1 = 1.f - powf(beta1, (float)step), bc2 = sqrtf(1.f - powf(beta2, (float)step));
  LAUNCH1D(adam_kernel, n, stream)(param, grad, exp_avg, exp_avg_sq, n, lr, beta1, beta2, eps, weight_decay, bc1, bc2);
  return LAUNCH_STATUS();
}

// Fused edge kernel: real spherical harmonics (lmax <= 3, e3nn 'component' normalisation, optional
// unit-vector normalisation), Bessel radial basis x cutoff envelope -- forward and analytic backward.
//
// Replaces sevenn/nn/edge_embedding.py:81-217 (BesselBasis, PolynomialCutoff, XPLORCutoff,
// SphericalEncoding -> e3nn o3.SphericalHarmonics, EdgeEmbedding.forward) and the part of
// torch.autograd.grad(E, edge_vec) (sevenn/nn/force_output.py:173-178) that runs through them.
// Elementwise, HBM-bound: one thread per edge, 12 B in, 4*(S+nb) B out.
#include "sevenn_common.cuh"
#include "../../include/sevenn_b200.h"

struct EdgeParams {
  int lmax, normalize, nb, cutoff_kind;   // cutoff_kind 0 = poly_cut, 1 = XPLOR
  float rc, p, r_on;
};

#define SQ3 1.7320508075688772f
#define SQ5 2.23606797749979f
#define SQ15 3.872983346207417f
#define SQ7 2.6457513110645907f

// component-normalised real SH of (x,y,z) (raw polynomial, homogeneous of degree l)
HD void sh_eval(int lmax, float x, float y, float z, float* Y) {
  Y[0] = 1.f;
  if (lmax >= 1) { Y[1] = SQ3 * x; Y[2] = SQ3 * y; Y[3] = SQ3 * z; }
  if (lmax >= 2) {
    Y[4] = SQ15 * x * z; Y[5] = SQ15 * x * y; Y[6] = SQ5 * (y * y - 0.5f * (x * x + z * z));
    Y[7] = SQ15 * y * z; Y[8] = 0.5f * SQ15 * (z * z - x * x);
  }
  if (lmax >= 3) {
    const float a = SQ7 * 0.7905694150420949f, b = SQ7 * 3.872983346207417f, c = SQ7 * 0.6123724356957945f,
                h = SQ7 * 0.5f, e = SQ7 * 1.9364916731037085f;
    const float x2 = x * x, y2 = y * y, z2 = z * z;
    Y[9] = a * x * (3.f * z2 - x2); Y[10] = b * x * y * z; Y[11] = c * x * (4.f * y2 - x2 - z2);
    Y[12] = h * y * (2.f * y2 - 3.f * x2 - 3.f * z2); Y[13] = c * z * (4.f * y2 - x2 - z2);
    Y[14] = e * y * (z2 - x2); Y[15] = a * z * (z2 - 3.f * x2);
  }
}

// g = sum_j dY[j] * grad Y_j(x,y,z)
HD void sh_grad_contract(int lmax, float x, float y, float z, const float* d, float* g) {
  float gx = 0.f, gy = 0.f, gz = 0.f;
  if (lmax >= 1) { gx += SQ3 * d[1]; gy += SQ3 * d[2]; gz += SQ3 * d[3]; }
  if (lmax >= 2) {
    gx += SQ15 * z * d[4] + SQ15 * y * d[5] - SQ5 * x * d[6] - SQ15 * x * d[8];
    gy += SQ15 * x * d[5] + 2.f * SQ5 * y * d[6] + SQ15 * z * d[7];
    gz += SQ15 * x * d[4] - SQ5 * z * d[6] + SQ15 * y * d[7] + SQ15 * z * d[8];
  }
  if (lmax >= 3) {
    const float a = SQ7 * 0.7905694150420949f, b = SQ7 * 3.872983346207417f, c = SQ7 * 0.6123724356957945f,
                h = SQ7 * 0.5f, e = SQ7 * 1.9364916731037085f;
    const float x2 = x * x, y2 = y * y, z2 = z * z;
    gx += a * (3.f * z2 - 3.f * x2) * d[9] + b * y * z * d[10] + c * (4.f * y2 - 3.f * x2 - z2) * d[11]
          - 6.f * h * x * y * d[12] - 2.f * c * x * z * d[13] - 2.f * e * x * y * d[14] - 6.f * a * x * z * d[15];
    gy += b * x * z * d[10] + 8.f * c * x * y * d[11] + h * (6.f * y2 - 3.f * x2 - 3.f * z2) * d[12]
          + 8.f * c * y * z * d[13] + e * (z2 - x2) * d[14];
    gz += 6.f * a * x * z * d[9] + b * x * y * d[10] - 2.f * c * x * z * d[11] - 6.f * h * y * z * d[12]
          + c * (4.f * y2 - x2 - 3.f * z2) * d[13] + 2.f * e * y * z * d[14] + a * (3.f * z2 - 3.f * x2) * d[15];
  }
  g[0] = gx; g[1] = gy; g[2] = gz;
}

HD float ipowf(float x, int n) { float r = 1.f; for (int k = 0; k < n; ++k) r *= x; return r; }

// envelope value and derivative wrt r
HD void cutoff_eval(const EdgeParams& P, float r, float* f, float* df) {
  if (P.cutoff_kind == 0) {
    const float p = P.p, x = r / P.rc;
    const float c0 = (p + 1.f) * (p + 2.f) * 0.5f, c1 = p * (p + 2.f), c2 = p * (p + 1.f) * 0.5f;
    const int ip = (int)p;
    float xp1;   // x^(p-1)
    if ((float)ip == p && ip >= 1) xp1 = ipowf(x, ip - 1); else xp1 = powf(x, p - 1.f);
    const float xp = xp1 * x, xq = xp * x, xr = xq * x;
    *f = 1.f - c0 * xp + c1 * xq - c2 * xr;
    *df = (-c0 * p * xp1 + c1 * (p + 1.f) * xp - c2 * (p + 2.f) * xq) / P.rc;
  } else {
    if (r < P.r_on) { *f = 1.f; *df = 0.f; return; }
    const float c2 = P.rc * P.rc, on2 = P.r_on * P.r_on, r2 = r * r;
    const float den = (c2 - on2) * (c2 - on2) * (c2 - on2);
    const float a = c2 - r2, b = c2 + 2.f * r2 - 3.f * on2;
    *f = a * a * b / den;
    *df = (-4.f * r * a * b + 4.f * r * a * a) / den;
  }
}

GKERNEL(edge_fwd_kernel)(const float* __restrict__ vec, long long n, EdgeParams P, const float* __restrict__ coeffs,
                         float* __restrict__ Y, float* __restrict__ emb) {
  const int S = (P.lmax + 1) * (P.lmax + 1);
  GRID_STRIDE_LOOP(e, n) {
    const float x = vec[3 * e], y = vec[3 * e + 1], z = vec[3 * e + 2];
    const float r = sqrtf(x * x + y * y + z * z);
    float yv[16];
    if (P.normalize) { const float ir = 1.f / r; sh_eval(P.lmax, x * ir, y * ir, z * ir, yv); }
    else sh_eval(P.lmax, x, y, z, yv);
    for (int k = 0; k < S; ++k) Y[e * S + k] = yv[k];
    float fc, dfc;
    cutoff_eval(P, r, &fc, &dfc);
    const float pref = 2.f / P.rc;
    for (int k = 0; k < P.nb; ++k) emb[e * P.nb + k] = pref * sinf(LDGF(coeffs + k) * r) / r * fc;
  }
}

GKERNEL(edge_bwd_kernel)(const float* __restrict__ vec, long long n, EdgeParams P, const float* __restrict__ coeffs,
                         const float* __restrict__ dY, int n_slabs, long long slab_stride,
                         const float* __restrict__ demb, float* __restrict__ fvec) {
  const int S = (P.lmax + 1) * (P.lmax + 1);
  GRID_STRIDE_LOOP(e, n) {
    const float x = vec[3 * e], y = vec[3 * e + 1], z = vec[3 * e + 2];
    const float r = sqrtf(x * x + y * y + z * z), ir = 1.f / r;
    float d[16], g[3];
    for (int k = 0; k < S; ++k) d[k] = dY[e * S + k];
    for (int sl = 1; sl < n_slabs; ++sl)                     // one slab per convolution unit, fixed order
      for (int k = 0; k < S; ++k) d[k] += dY[(size_t)sl * slab_stride + e * S + k];
    float fx, fy, fz;
    if (P.normalize) {
      const float ux = x * ir, uy = y * ir, uz = z * ir;
      sh_grad_contract(P.lmax, ux, uy, uz, d, g);
      const float dot = ux * g[0] + uy * g[1] + uz * g[2];
      fx = (g[0] - ux * dot) * ir; fy = (g[1] - uy * dot) * ir; fz = (g[2] - uz * dot) * ir;
    } else {
      sh_grad_contract(P.lmax, x, y, z, d, g);
      fx = g[0]; fy = g[1]; fz = g[2];
    }
    float fc, dfc;
    cutoff_eval(P, r, &fc, &dfc);
    const float pref = 2.f / P.rc;
    float dr = 0.f;
    for (int k = 0; k < P.nb; ++k) {
      const float c = LDGF(coeffs + k);
      const float s = sinf(c * r), co = cosf(c * r);
      const float b = s * ir;                               // sin(cr)/r
      const float db = (c * co - b) * ir;                   // d/dr
      dr += demb[e * P.nb + k] * pref * (db * fc + b * dfc);
    }
    fvec[3 * e] = fx + dr * x * ir; fvec[3 * e + 1] = fy + dr * y * ir; fvec[3 * e + 2] = fz + dr * z * ir;
  }
}

// Tabulated-radial variant (codegen/conv_codegen.py "Tabulated radial weights"): per edge
// Yt[e] = (Y_0 .. Y_{S-1}, t, bits(i)) with r = (i + t) h the position of the edge length in the table.
GKERNEL(edge_fwd_tab_kernel)(const float* __restrict__ vec, long long n, const int* __restrict__ n_dev, int lmax,
                             int normalize, float inv_h, int n_int, float* __restrict__ Yt) {
  const int S = (lmax + 1) * (lmax + 1);
  if (n_dev != nullptr && *n_dev < n) n = *n_dev;       // edge count known on the device only (sync-free host)
  GRID_STRIDE_LOOP(e, n) {
    const float x = vec[3 * e], y = vec[3 * e + 1], z = vec[3 * e + 2];
    const float r = sqrtf(x * x + y * y + z * z);
    float yv[16];
    if (normalize) { const float ir = 1.f / r; sh_eval(lmax, x * ir, y * ir, z * ir, yv); }
    else sh_eval(lmax, x, y, z, yv);
    float* row = Yt + e * (S + 2);
    for (int k = 0; k < S; ++k) row[k] = yv[k];
    const float s = r * inv_h;
    int i = (int)s;
    if (i > n_int - 1) i = n_int - 1;
    if (i < 0) i = 0;
    row[S] = s - (float)i;
    reinterpret_cast<int*>(row)[S + 1] = i;
  }
}

// dY rows carry S + 1 values: d/dY and the radial term sum_c dE/dw_c * h dw_c/dr (one slab per convolution unit).
GKERNEL(edge_bwd_tab_kernel)(const float* __restrict__ vec, long long n, const int* __restrict__ n_dev, int lmax,
                             int normalize, float inv_h, const float* __restrict__ dY, int n_slabs,
                             long long slab_stride, float* __restrict__ fvec) {
  const int S = (lmax + 1) * (lmax + 1), S1 = S + 1;
  if (n_dev != nullptr && *n_dev < n) n = *n_dev;
  GRID_STRIDE_LOOP(e, n) {
    const float x = vec[3 * e], y = vec[3 * e + 1], z = vec[3 * e + 2];
    const float r = sqrtf(x * x + y * y + z * z), ir = 1.f / r;
    float d[17], g[3];
    for (int k = 0; k < S1; ++k) d[k] = dY[e * S1 + k];
    for (int sl = 1; sl < n_slabs; ++sl)
      for (int k = 0; k < S1; ++k) d[k] += dY[(size_t)sl * slab_stride + e * S1 + k];
    float fx, fy, fz;
    if (normalize) {
      const float ux = x * ir, uy = y * ir, uz = z * ir;
      sh_grad_contract(lmax, ux, uy, uz, d, g);
      const float dot = ux * g[0] + uy * g[1] + uz * g[2];
      fx = (g[0] - ux * dot) * ir; fy = (g[1] - uy * dot) * ir; fz = (g[2] - uz * dot) * ir;
    } else {
      sh_grad_contract(lmax, x, y, z, d, g);
      fx = g[0]; fy = g[1]; fz = g[2];
    }
    const float dr = d[S] * inv_h;
    fvec[3 * e] = fx + dr * x * ir; fvec[3 * e + 1] = fy + dr * y * ir; fvec[3 * e + 2] = fz + dr * z * ir;
  }
}

static EdgeParams edge_params(int lmax, int normalize, int nb, int cutoff_kind, float rc, float p, float r_on) {
  EdgeParams P; P.lmax = lmax; P.normalize = normalize; P.nb = nb; P.cutoff_kind = cutoff_kind;
  P.rc = rc; P.p = p; P.r_on = r_on; return P;
}

extern "C" int sevenn_edge_fwd(const float* vec, long long n_edges, int lmax, int normalize, int nb,
                               int cutoff_kind, float rc, float p, float r_on, const float* coeffs, float* Y,
                               float* emb, void* stream) {
  if (n_edges <= 0) return 0;
  if (lmax < 0 || lmax > 3 || nb < 1) return SEVENN_ERR_BAD_ARG;
  LAUNCH1D(edge_fwd_kernel, n_edges, stream)(vec, n_edges, edge_params(lmax, normalize, nb, cutoff_kind, rc, p, r_on),
                                             coeffs, Y, emb);
  return LAUNCH_STATUS();
}

extern "C" int sevenn_edge_bwd(const float* vec, long long n_edges, int lmax, int normalize, int nb,
                               int cutoff_kind, float rc, float p, float r_on, const float* coeffs,
                               const float* dY, int n_slabs, long long slab_stride, const float* demb,
                               float* fvec, void* stream) {
  if (n_edges <= 0) return 0;
  if (lmax < 0 || lmax > 3 || nb < 1 || n_slabs < 1) return SEVENN_ERR_BAD_ARG;
  LAUNCH1D(edge_bwd_kernel, n_edges, stream)(vec, n_edges, edge_params(lmax, normalize, nb, cutoff_kind, rc, p, r_on),
                                             coeffs, dY, n_slabs, slab_stride, demb, fvec);
  return LAUNCH_STATUS();
}

extern "C" int sevenn_edge_fwd_tab(const float* vec, long long n_edges, int lmax, int normalize, float inv_h,
                                   int n_int, float* Yt, void* stream) {
  if (n_edges <= 0) return 0;
  if (lmax < 0 || lmax > 3 || n_int < 1 || !(inv_h > 0.f)) return SEVENN_ERR_BAD_ARG;
  LAUNCH1D(edge_fwd_tab_kernel, n_edges, stream)(vec, n_edges, nullptr, lmax, normalize, inv_h, n_int, Yt);
  return LAUNCH_STATUS();
}

extern "C" int sevenn_edge_bwd_tab(const float* vec, long long n_edges, int lmax, int normalize, float inv_h,
                                   const float* dY, int n_slabs, long long slab_stride, float* fvec, void* stream) {
  if (n_edges <= 0) return 0;
  if (lmax < 0 || lmax > 3 || n_slabs < 1) return SEVENN_ERR_BAD_ARG;
  LAUNCH1D(edge_bwd_tab_kernel, n_edges, stream)(vec, n_edges, nullptr, lmax, normalize, inv_h, dY, n_slabs, slab_stride, fvec);
  return LAUNCH_STATUS();
}

// Same kernels with the edge count read on the device (n_edges_dev, e.g. &rowptr[n_rows]); `max_edges` bounds the launch
// and the buffers.  For hosts that never wait for the neighbour-list size (csrc/model.cu).
extern "C" int sevenn_edge_fwd_tab_dev(const float* vec, long long max_edges, const int* n_edges_dev, int lmax,
                                       int normalize, float inv_h, int n_int, float* Yt, void* stream) {
  if (max_edges <= 0) return 0;
  if (lmax < 0 || lmax > 3 || n_int < 1 || !(inv_h > 0.f) || n_edges_dev == nullptr) return SEVENN_ERR_BAD_ARG;
  LAUNCH1D(edge_fwd_tab_kernel, max_edges, stream)(vec, max_edges, n_edges_dev, lmax, normalize, inv_h, n_int, Yt);
  return LAUNCH_STATUS();
}
extern "C" int sevenn_edge_bwd_tab_dev(const float* vec, long long max_edges, const int* n_edges_dev, int lmax,
                                       int normalize, float inv_h, const float* dY, int n_slabs, long long slab_stride,
                                       float* fvec, void* stream) {
  if (max_edges <= 0) return 0;
  if (lmax < 0 || lmax > 3 || n_slabs < 1 || n_edges_dev == nullptr) return SEVENN_ERR_BAD_ARG;
  LAUNCH1D(edge_bwd_tab_kernel, max_edges, stream)(vec, max_edges, n_edges_dev, lmax, normalize, inv_h, dY, n_slabs,
                                                  slab_stride, fvec);
  return LAUNCH_STATUS();
}

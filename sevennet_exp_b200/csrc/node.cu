// Per-node elementwise / gather kernels of the SevenNet evaluation path (fp32, ir_mul layout).
//   embed            : OnehotEmbedding + onehot_to_feature_x  (sevenn/nn/node_embedding.py:15-53, model_build.py:453-458)
//   gate fwd/bwd     : EquivariantGate -> e3nn nn.Gate        (sevenn/nn/equivariant_gate.py:13-61)
//   energy / seed    : Rescale / SpeciesWiseRescale + AtomReduce (sevenn/nn/scale.py:22-161, nn/linear.py:104-141)
//   force_virial     : ForceStressOutputFromEdge scatter part (sevenn/nn/force_output.py:185-222)
//   segment_gather_sum: source-side reduction of per-edge dE/dx_j (replaces autograd's index backward, atomics-free)
//   linear_by_type   : SelfConnectionIntro (FullyConnectedTensorProduct x (x) one-hot) (nn/self_connection.py:10-64)
//   gather_rows / scatter_add_rows : halo pack / unpack-add (pair_e3gnn_parallel.cpp:787-897)
// All HBM-bound; grid-stride, coalesced along the channel index.
#include "sevenn_common.cuh"
#ifndef SEVENN_EMU
#include <atomic>
#endif
#include "../../include/sevenn_b200.h"

HD float act_f(int id, float x) {
  switch (id) {
    case 0: return x / (1.f + expf(-x));
    case 1: return tanhf(x);
    case 2: return fabsf(x);
    case 3: return x > 0.f ? x : 0.f;
    case 4: return 1.f / (1.f + expf(-x));
    default: return (x > 20.f ? x : log1pf(expf(x))) - 0.6931471805599453f;
  }
}
HD float dact_f(int id, float x) {
  switch (id) {
    case 0: { const float s = 1.f / (1.f + expf(-x)); return s * (1.f + x * (1.f - s)); }
    case 1: { const float t = tanhf(x); return 1.f - t * t; }
    case 2: return x > 0.f ? 1.f : (x < 0.f ? -1.f : 0.f);
    case 3: return x > 0.f ? 1.f : 0.f;
    case 4: { const float s = 1.f / (1.f + expf(-x)); return s * (1.f - s); }
    default: return 1.f / (1.f + expf(-x));
  }
}

GKERNEL(embed_kernel)(const int* __restrict__ types, long long n, int dim, const float* __restrict__ table,
                      float* __restrict__ out) {
  GRID_STRIDE_LOOP(t, n * dim) {
    const long long i = t / dim; const int c = (int)(t % dim);
    out[t] = LDGF(table + (size_t)types[i] * dim + c);
  }
}

struct GateTable { int n; sevenn_gate_piece p[SEVENN_MAX_GATE_PIECES]; };

GKERNEL(gate_fwd_kernel)(const float* __restrict__ pre, int din, float* __restrict__ out, int dout, long long n,
                         GateTable T) {
  GRID_STRIDE_LOOP(t, n * dout) {
    const long long i = t / dout; int c = (int)(t % dout);
    const float* row = pre + i * din;
    for (int q = 0; q < T.n; ++q) {
      const sevenn_gate_piece& P = T.p[q];
      const int sz = P.mul * (2 * P.l + 1);
      // pieces tile the output row; inside a piece the output block is [2l+1][out_stride] at out_off
      // (out_stride == mul because gate outputs are simplified entries)
      if (c < sz) {
        const int m = c / P.mul, u = c % P.mul;
        float v;
        if (P.kind == 0) v = P.cst * act_f(P.act, row[P.in_off + u]);
        else v = P.cst * act_f(P.act, row[P.gate_off + u]) * row[P.in_off + m * P.in_stride + u];
        out[i * dout + P.out_off + m * P.out_stride + u] = v;
        break;
      }
      c -= sz;
    }
  }
}

// one thread per (node, piece channel)
GKERNEL(gate_bwd_kernel)(const float* __restrict__ pre, int din, const float* __restrict__ dout_, int dout,
                         float* __restrict__ dpre, long long n, GateTable T, int nch) {
  GRID_STRIDE_LOOP(t, n * nch) {
    const long long i = t / nch; int c = (int)(t % nch);
    const float* row = pre + i * din; const float* go = dout_ + i * dout; float* gi = dpre + i * din;
    for (int q = 0; q < T.n; ++q) {
      const sevenn_gate_piece& P = T.p[q];
      if (c < P.mul) {
        const int u = c;
        if (P.kind == 0) {
          gi[P.in_off + u] = go[P.out_off + u] * P.cst * dact_f(P.act, row[P.in_off + u]);
        } else {
          const float gpre = row[P.gate_off + u];
          const float gate = P.cst * act_f(P.act, gpre);
          float acc = 0.f;
          for (int m = 0; m < 2 * P.l + 1; ++m) {
            const float g = go[P.out_off + m * P.out_stride + u];
            acc += g * row[P.in_off + m * P.in_stride + u];
            gi[P.in_off + m * P.in_stride + u] = g * gate;
          }
          gi[P.gate_off + u] = acc * P.cst * dact_f(P.act, gpre);
        }
        break;
      }
      c -= P.mul;
    }
  }
}

#ifndef SEVENN_EMU
// 16-byte variants (every multiplicity, offset and stride a multiple of 4, rows 16 B aligned -- SevenNet-0 / l3i5 and
// any model with channel counts divisible by 4).  threadIdx.x owns one float4 column of the row for the whole launch:
// the piece lookup and its integer divisions happen once per thread instead of once per element, threadIdx.y /
// blockIdx.x walk the rows.  Same expressions in the same order as the scalar kernels above => same bits.
__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }

__global__ void __launch_bounds__(512) gate_fwd_v4_kernel(const float* __restrict__ pre, int din, float* __restrict__ out,
                                                          int dout, long long n, const GateTable T) {
  int c = (int)threadIdx.x * 4, q = 0;
  for (; q < T.n; ++q) {
    const int sz = T.p[q].mul * (2 * T.p[q].l + 1);
    if (c < sz) break;
    c -= sz;
  }
  if (q == T.n) return;
  const sevenn_gate_piece P = T.p[q];
  const int m = c / P.mul, u = c % P.mul;
  const int in = P.in_off + (P.kind ? m * P.in_stride : 0) + u, gt = P.gate_off + u, o = P.out_off + m * P.out_stride + u;
  const int kind = P.kind, act = P.act; const float cst = P.cst;
  const long long step = (long long)gridDim.x * blockDim.y;
  for (long long i = blockIdx.x * (long long)blockDim.y + threadIdx.y; i < n; i += 2 * step) {
    const long long j = i + step < n ? i + step : i;            // second row of the pair (repeats row i at the tail)
    const float4 xi = ld4(pre + i * din + in), xj = ld4(pre + j * din + in);
    float4 gi = xi, gj = xj;
    if (kind != 0) { gi = ld4(pre + i * din + gt); gj = ld4(pre + j * din + gt); }
    float4 vi, vj;
    if (kind == 0) {
      vi = make_float4(cst * act_f(act, xi.x), cst * act_f(act, xi.y), cst * act_f(act, xi.z), cst * act_f(act, xi.w));
      vj = make_float4(cst * act_f(act, xj.x), cst * act_f(act, xj.y), cst * act_f(act, xj.z), cst * act_f(act, xj.w));
    } else {
      vi = make_float4(cst * act_f(act, gi.x) * xi.x, cst * act_f(act, gi.y) * xi.y, cst * act_f(act, gi.z) * xi.z,
                       cst * act_f(act, gi.w) * xi.w);
      vj = make_float4(cst * act_f(act, gj.x) * xj.x, cst * act_f(act, gj.y) * xj.y, cst * act_f(act, gj.z) * xj.z,
                       cst * act_f(act, gj.w) * xj.w);
    }
    st4(out + i * dout + o, vi);
    if (j != i) st4(out + j * dout + o, vj);
  }
}

// Backward: threadIdx.x owns one float4 of the dpre row, so every thread does the same small amount of work (2-3 loads,
// one store).  The gate-channel gradient needs sum_m dout[m,u] x[m,u]: the threads of the gated components leave their
// products in shared memory and the gate-channel thread adds them in m order after the barrier (double buffered by
// iteration parity: one __syncthreads per row group).  A thread per channel that walked the 2l+1 components itself
// (first version) ran at 2.5 TB/s: few, heavy threads and not enough loads in flight.
__global__ void __launch_bounds__(512, 2) gate_bwd_v4_kernel(const float* __restrict__ pre, int din, const float* __restrict__ dout_,
                                                             int dout, float* __restrict__ dpre, long long n, const GateTable T,
                                                             int npart) {
  extern __shared__ float4 gate_part[];                 // [2][blockDim.y][npart]
  const int c = (int)threadIdx.x * 4;
  // role of this column: 0 scalar, 1 gated component, 2 gate channel, -1 none
  int role = -1, go_off = 0, gate_off = 0, slot = 0, nm = 0, mstride = 0, act = 0; float cst = 0.f;
  for (int q = 0, pbase = 0; q < T.n; ++q) {
    const sevenn_gate_piece& P = T.p[q];
    const int pnm = 2 * P.l + 1, w = P.mul / 4;
    if (P.kind == 0) {
      if (c >= P.in_off && c < P.in_off + P.mul) { role = 0; go_off = P.out_off + (c - P.in_off); act = P.act; cst = P.cst; }
    } else {
      if (c >= P.gate_off && c < P.gate_off + P.mul) {
        role = 2; slot = pbase + (c - P.gate_off) / 4; nm = pnm; mstride = w; act = P.act; cst = P.cst;
      } else if (c >= P.in_off && c < P.in_off + pnm * P.in_stride) {
        const int m = (c - P.in_off) / P.in_stride, u = (c - P.in_off) % P.in_stride;
        if (u < P.mul) {
          role = 1; go_off = P.out_off + m * P.out_stride + u; gate_off = P.gate_off + u; slot = pbase + m * w + u / 4;
          act = P.act; cst = P.cst;
        }
      }
      pbase += pnm * w;
    }
  }
  const int rb = (int)blockDim.y, ty = (int)threadIdx.y;
  const long long stride = (long long)gridDim.x * rb;
  // the loads of the next row group are issued before this group's barrier (two groups of loads in flight per thread)
  float4 x = make_float4(0.f, 0.f, 0.f, 0.f), g = x, p = x;
  auto load = [&](long long i, float4& x_, float4& g_, float4& p_) {
    if (i >= n || role < 0) return;
    x_ = ld4(pre + i * din + c);                                   // role 2: the gate pre-activation itself
    if (role != 2) g_ = ld4(dout_ + i * dout + go_off);
    if (role == 1) p_ = ld4(pre + i * din + gate_off);
  };
  load(blockIdx.x * (long long)rb + ty, x, g, p);
  int buf = 0;
  for (long long base = blockIdx.x * (long long)rb; base < n; base += stride, buf ^= 1) {
    const long long i = base + ty;
    const bool ok = i < n;
    float4 xn = x, gn = g, pn = p;
    load(i + stride, xn, gn, pn);
    float4* part = gate_part + ((size_t)buf * rb + ty) * npart;
    if (ok && role == 0) {
      st4(dpre + i * din + c, make_float4(g.x * cst * dact_f(act, x.x), g.y * cst * dact_f(act, x.y),
                                          g.z * cst * dact_f(act, x.z), g.w * cst * dact_f(act, x.w)));
    } else if (ok && role == 1) {
      const float4 gate = make_float4(cst * act_f(act, p.x), cst * act_f(act, p.y), cst * act_f(act, p.z), cst * act_f(act, p.w));
      st4(dpre + i * din + c, make_float4(g.x * gate.x, g.y * gate.y, g.z * gate.z, g.w * gate.w));
      part[slot] = make_float4(g.x * x.x, g.y * x.y, g.z * x.z, g.w * x.w);
    }
    __syncthreads();
    if (ok && role == 2) {
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int m = 0; m < nm; ++m) {
        const float4 q = part[slot + m * mstride];
        acc.x += q.x; acc.y += q.y; acc.z += q.z; acc.w += q.w;
      }
      st4(dpre + i * din + c, make_float4(acc.x * cst * dact_f(act, x.x), acc.y * cst * dact_f(act, x.y),
                                          acc.z * cst * dact_f(act, x.z), acc.w * cst * dact_f(act, x.w)));
    }
    x = xn; g = gn; p = pn;
  }
}

// the 16-byte kernels apply when every piece is 4-aligned; `cols` = float4 columns per row (threadIdx.x extent)
static bool gate_v4_ok(const GateTable& T, int din, int dout, const void* a, const void* b, const void* c, bool bwd, int* cols,
                       int* npart) {
  if ((din & 3) || (dout & 3) || ((uintptr_t)a & 15) || ((uintptr_t)b & 15) || ((uintptr_t)c & 15)) return false;
  int tot = 0, np = 0;
  for (int k = 0; k < T.n; ++k) {
    const sevenn_gate_piece& P = T.p[k];
    if (P.l < 0 || P.l > 4 || (P.kind && P.l == 0) || (P.mul & 3) || (P.in_off & 3) || (P.out_off & 3) ||
        (P.kind && ((P.gate_off & 3) || (P.in_stride & 3) || (P.out_stride & 3) || P.in_stride < P.mul)))
      return false;
    tot += P.mul * (2 * P.l + 1);
    if (P.kind) np += (P.mul / 4) * (2 * P.l + 1);
  }
  if (tot != dout) return false;
  *cols = (bwd ? din : dout) / 4;
  *npart = np;
  return *cols >= 1 && *cols <= 512;
}
static void gate_v4_dims(int cols, long long n, dim3* grid, dim3* block) {
  int by = 512 / cols; if (by > 8) by = 8; if (by < 1) by = 1;
  long long g = (n + by - 1) / by;
  const long long cap = 148LL * 8;
  if (g > cap) g = cap;
  *grid = dim3((unsigned)g); *block = dim3((unsigned)cols, (unsigned)by);
}
#endif

static int load_gate(GateTable& T, const sevenn_gate_piece* pieces, int n_pieces) {
  if (n_pieces < 0 || n_pieces > SEVENN_MAX_GATE_PIECES) return SEVENN_ERR_BAD_ARG;
  T.n = n_pieces;
  for (int k = 0; k < n_pieces; ++k) T.p[k] = pieces[k];
  return 0;
}

extern "C" int sevenn_embed(const int* types, long long n, int dim, const float* table, float* out, void* stream) {
  if (n <= 0) return 0;
  LAUNCH1D(embed_kernel, n * dim, stream)(types, n, dim, table, out);
  return LAUNCH_STATUS();
}
extern "C" int sevenn_gate_fwd(const float* pre, int din, float* out, int dout, long long n,
                               const sevenn_gate_piece* pieces, int n_pieces, void* stream) {
  GateTable T; int rc = load_gate(T, pieces, n_pieces); if (rc) return rc;
  if (n <= 0) return 0;
#ifndef SEVENN_EMU
  int cols, npart;
  if (gate_v4_ok(T, din, dout, pre, out, nullptr, false, &cols, &npart)) {
    dim3 grid, block; gate_v4_dims(cols, n, &grid, &block);
    gate_fwd_v4_kernel<<<grid, block, 0, (cudaStream_t)stream>>>(pre, din, out, dout, n, T);
    return LAUNCH_STATUS();
  }
#endif
  LAUNCH1D(gate_fwd_kernel, n * dout, stream)(pre, din, out, dout, n, T);
  return LAUNCH_STATUS();
}
extern "C" int sevenn_gate_bwd(const float* pre, int din, const float* dout_, int dout, float* dpre, long long n,
                               const sevenn_gate_piece* pieces, int n_pieces, void* stream) {
  GateTable T; int rc = load_gate(T, pieces, n_pieces); if (rc) return rc;
  if (n <= 0) return 0;
  int nch = 0; for (int k = 0; k < n_pieces; ++k) nch += pieces[k].mul;
#ifndef SEVENN_EMU
  int cols, npart;
  if (gate_v4_ok(T, din, dout, pre, dout_, dpre, true, &cols, &npart)) {
    dim3 grid, block; gate_v4_dims(cols, n, &grid, &block);
    const size_t sm = (size_t)2 * block.y * (npart > 0 ? npart : 1) * sizeof(float4);
    gate_bwd_v4_kernel<<<grid, block, sm, (cudaStream_t)stream>>>(pre, din, dout_, dout, dpre, n, T, npart);
    return LAUNCH_STATUS();
  }
#endif
  LAUNCH1D(gate_bwd_kernel, n * nch, stream)(pre, din, dout_, dout, dpre, n, T, nch);
  return LAUNCH_STATUS();
}

// ---- energy: E_i = e_i * scale[t_i] + shift[t_i]; seed of the backward pass dE/de_i = scale[t_i]
GKERNEL(rescale_kernel)(const float* __restrict__ e_scaled, const int* __restrict__ types, long long n,
                        const float* __restrict__ scale, const float* __restrict__ shift, int species_wise,
                        float* __restrict__ atomic_energy, float* __restrict__ seed) {
  GRID_STRIDE_LOOP(i, n) {
    const int t = species_wise ? types[i] : 0;
    const float s = LDGF(scale + t);
    atomic_energy[i] = e_scaled[i] * s + LDGF(shift + t);
    seed[i] = s;
  }
}
extern "C" int sevenn_rescale(const float* e_scaled, const int* types, long long n, const float* scale,
                              const float* shift, int species_wise, float* atomic_energy, float* seed, void* stream) {
  if (n <= 0) return 0;
  LAUNCH1D(rescale_kernel, n, stream)(e_scaled, types, n, scale, shift, species_wise, atomic_energy, seed);
  return LAUNCH_STATUS();
}

// ---- deterministic column sums of a [n, D] fp32 matrix into D doubles (two stages, D <= 8)
#ifndef SEVENN_EMU
#define COLSUM_BLOCKS 296
__global__ void colsum_stage1(const float* __restrict__ a, long long n, int D, double* __restrict__ part) {
  __shared__ double sh[256];
  for (int d = 0; d < D; ++d) {
    double s = 0.0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
      s += (double)a[i * D + d];
    sh[threadIdx.x] = s;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o]; __syncthreads(); }
    if (threadIdx.x == 0) part[blockIdx.x * D + d] = sh[0];
    __syncthreads();
  }
}
// one warp per column: lanes stride over the block partials, fixed-order shuffle tree (bit-reproducible)
__global__ void colsum_stage2(const double* __restrict__ part, int nb, int D, double* __restrict__ out) {
  const int d = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (d >= D) return;
  double s = 0.0;
  for (int b = lane; b < nb; b += 32) s += part[b * D + d];
  for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
  if (lane == 0) out[d] = s;
}
#endif
extern "C" int sevenn_colsum_f64(const float* a, long long n, int D, double* part /* >= 296*D */, double* out,
                                 void* stream) {
  if (D < 1 || D > 8) return SEVENN_ERR_BAD_ARG;
#ifdef SEVENN_EMU
  for (int d = 0; d < D; ++d) { double s = 0; for (long long i = 0; i < n; ++i) s += a[i * D + d]; out[d] = s; }
  return 0;
#else
  colsum_stage1<<<COLSUM_BLOCKS, 256, 0, (cudaStream_t)stream>>>(a, n, D, part);
  colsum_stage2<<<1, 32 * D, 0, (cudaStream_t)stream>>>(part, COLSUM_BLOCKS, D, out);
  return (int)cudaGetLastError();
#endif
}

// ---- no-allocation variant of sevenn_colsum_f64 (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA-graph capture, tight serving loops).  The stage-1 block
// partials live in a file-scope device array (SEVENN_COLSUM_NOALLOC_MAX_D * 296 doubles = 18,944 B, held by every
// GPU that loads the library until exit).  In host code `g_colsum_scratch` is only a stand-in for the symbol, not a
// device address: the device address is looked up once per device, with that device current, by
// cudaGetSymbolAddress and cached.  Every call on a device shares the array: stage 1 overwrites all 296 * D
// partials it later reads, so nothing has to start at zero, but two calls that overlap on the GPU (different
// streams) race on it.
#ifndef SEVENN_EMU
static __device__ __align__(256) double g_colsum_scratch[COLSUM_BLOCKS * SEVENN_COLSUM_NOALLOC_MAX_D];
#define COLSUM_MAX_DEVICES 64
static std::atomic<double*> g_colsum_scratch_ptr[COLSUM_MAX_DEVICES];   // per device; every writer stores the same value
#endif
extern "C" int sevenn_colsum_f64_noalloc(const float* a, long long n, int D, double* out, void* stream) {
  if (D < 1 || D > SEVENN_COLSUM_NOALLOC_MAX_D || n < 0) return SEVENN_ERR_BAD_ARG;
#ifdef SEVENN_EMU
  for (int d = 0; d < D; ++d) { double s = 0; for (long long i = 0; i < n; ++i) s += a[i * D + d]; out[d] = s; }
  return 0;
#else
  int dev = -1;
  cudaError_t rc = cudaGetDevice(&dev);
  if (rc != cudaSuccess) return (int)rc;
  if (dev < 0 || dev >= COLSUM_MAX_DEVICES) return SEVENN_ERR_BAD_ARG;
  double* part = g_colsum_scratch_ptr[dev].load(std::memory_order_acquire);
  if (part == nullptr) {
    rc = cudaGetSymbolAddress((void**)&part, g_colsum_scratch);
    if (rc != cudaSuccess || part == nullptr) return rc != cudaSuccess ? (int)rc : SEVENN_ERR_BAD_ARG;
    g_colsum_scratch_ptr[dev].store(part, std::memory_order_release);
  }
  colsum_stage1<<<COLSUM_BLOCKS, 256, 0, (cudaStream_t)stream>>>(a, n, D, part);
  colsum_stage2<<<1, 32 * D, 0, (cudaStream_t)stream>>>(part, COLSUM_BLOCKS, D, out);
  return (int)cudaGetLastError();
#endif
}

// ---- forces and virial from per-edge dE/d(edge_vec)
// F_k = sum_{e: centre k} f_e - sum_{e: neighbour k} f_e ; w_k[6] = sum_{e: centre k} (rx fx, ry fy, rz fz, rx fy, ry fz, rz fx)
// One thread per (node, output component): 9 threads share a node's edge lists through L1 (the per-node form ran
// 16 CTAs for 4,096 atoms); every component is still summed in ascending edge order, so results are unchanged.
GKERNEL(force_virial_kernel)(const float* __restrict__ fvec, const float* __restrict__ vec,
                             const int* __restrict__ rowptr, int n_rows, const int* __restrict__ cptr,
                             const int* __restrict__ cedge, long long n_nodes, float* __restrict__ F,
                             float* __restrict__ vir) {
  GRID_STRIDE_LOOP(t, n_nodes * 9) {
    const long long k = t / 9;
    const int c = (int)(t - k * 9);
    float s = 0.f;
    if (c < 3) {
      if (k < n_rows)
        for (int e = rowptr[k]; e < rowptr[k + 1]; ++e) s += fvec[3 * e + c];
      for (int a = cptr[k]; a < cptr[k + 1]; ++a) s -= fvec[3 * cedge[a] + c];
      F[3 * k + c] = s;
    } else {
      // (xx, yy, zz, xy, yz, zx): r component (c - 3) % 3, f component (c - 3) % 3 or the next one
      const int cr = (c - 3) % 3, cf = c < 6 ? cr : (cr + 1) % 3;
      if (k < n_rows)
        for (int e = rowptr[k]; e < rowptr[k + 1]; ++e) s += vec[3 * e + cr] * fvec[3 * e + cf];
      vir[6 * k + (c - 3)] = s;
    }
  }
}
extern "C" int sevenn_force_virial(const float* fvec, const float* vec, const int* rowptr, int n_rows,
                                   const int* cptr, const int* cedge, long long n_nodes, float* F, float* vir,
                                   void* stream) {
  if (n_nodes <= 0) return 0;
  LAUNCH1D(force_virial_kernel, n_nodes * 9, stream)(fvec, vec, rowptr, n_rows, cptr, cedge, n_nodes, F, vir);
  return LAUNCH_STATUS();
}

// ---- out[j, :] (+)= sum_{a in csc(j)} src[cedge[a], :]
// One thread per (node, 4 channels): 16-byte loads, four edges in flight per thread (indices first, then the four
// gathers), summed in ascending edge order so the result does not depend on the unrolling.  The scalar kernel below
// serves rows whose width is not a multiple of 4.
struct alignas(16) SevennF4 { float x, y, z, w; };
GKERNEL(segment_gather_sum4_kernel)(const SevennF4* __restrict__ src, int D4, const int* __restrict__ cptr,
                                    const int* __restrict__ cedge, long long n_nodes, SevennF4* __restrict__ out,
                                    int accumulate) {
  GRID_STRIDE_LOOP(t, n_nodes * D4) {
    const long long j = t / D4; const int c = (int)(t % D4);
    SevennF4 s = {0.f, 0.f, 0.f, 0.f};
    if (accumulate) s = out[t];
    const int a1 = cptr[j + 1];
    int a = cptr[j];
    for (; a + 4 <= a1; a += 4) {
      const int e0 = cedge[a], e1 = cedge[a + 1], e2 = cedge[a + 2], e3 = cedge[a + 3];
      const SevennF4 v0 = src[(size_t)e0 * D4 + c], v1 = src[(size_t)e1 * D4 + c];
      const SevennF4 v2 = src[(size_t)e2 * D4 + c], v3 = src[(size_t)e3 * D4 + c];
      s.x += v0.x; s.y += v0.y; s.z += v0.z; s.w += v0.w;
      s.x += v1.x; s.y += v1.y; s.z += v1.z; s.w += v1.w;
      s.x += v2.x; s.y += v2.y; s.z += v2.z; s.w += v2.w;
      s.x += v3.x; s.y += v3.y; s.z += v3.z; s.w += v3.w;
    }
    for (; a < a1; ++a) {
      const SevennF4 v = src[(size_t)cedge[a] * D4 + c];
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
    out[t] = s;
  }
}
GKERNEL(segment_gather_sum_kernel)(const float* __restrict__ src, int D, const int* __restrict__ cptr,
                                   const int* __restrict__ cedge, long long n_nodes, float* __restrict__ out,
                                   int accumulate) {
  GRID_STRIDE_LOOP(t, n_nodes * D) {
    const long long j = t / D; const int c = (int)(t % D);
    float s = accumulate ? out[t] : 0.f;
    for (int a = cptr[j]; a < cptr[j + 1]; ++a) s += src[(size_t)cedge[a] * D + c];
    out[t] = s;
  }
}
extern "C" int sevenn_segment_gather_sum(const float* src, int D, const int* cptr, const int* cedge,
                                         long long n_nodes, float* out, int accumulate, void* stream) {
  if (n_nodes <= 0) return 0;
  if (D % 4 == 0 && ((uintptr_t)src & 15) == 0 && ((uintptr_t)out & 15) == 0) {
    LAUNCH1D(segment_gather_sum4_kernel, n_nodes * (D / 4), stream)(
        reinterpret_cast<const SevennF4*>(src), D / 4, cptr, cedge, n_nodes, reinterpret_cast<SevennF4*>(out), accumulate);
  } else {
    LAUNCH1D(segment_gather_sum_kernel, n_nodes * D, stream)(src, D, cptr, cedge, n_nodes, out, accumulate);
  }
  return LAUNCH_STATUS();
}

// ---- species-indexed linear: C[(n,m), w] (+)= sum_u A[(n,m), u] * W[type_n][u][w]
GKERNEL(linear_by_type_kernel)(const float* __restrict__ A, long long sAn, int sAm, const float* __restrict__ W,
                               const int* __restrict__ types, float* __restrict__ C, long long sCn, int sCm,
                               long long n_nodes, int R, int K, int N, int accumulate) {
  GRID_STRIDE_LOOP(t, n_nodes * R * N) {
    const int w = (int)(t % N); const long long r = t / N; const long long nn = r / R; const int m = (int)(r % R);
    const float* a = A + nn * sAn + m * sAm;
    const float* wt = W + (size_t)types[nn] * K * N + w;
    float s = 0.f;
    for (int u = 0; u < K; ++u) s += a[u] * wt[(size_t)u * N];
    float* c = C + nn * sCn + m * sCm + w;
    *c = accumulate ? *c + s : s;
  }
}
extern "C" int sevenn_linear_by_type(const float* A, long long sAn, int sAm, const float* W, const int* types,
                                     float* C, long long sCn, int sCm, long long n_nodes, int R, int K, int N,
                                     int accumulate, void* stream) {
  if (n_nodes <= 0) return 0;
  LAUNCH1D(linear_by_type_kernel, n_nodes * R * N, stream)(A, sAn, sAm, W, types, C, sCn, sCm, n_nodes, R, K, N, accumulate);
  return LAUNCH_STATUS();
}

// ---- halo pack / unpack
GKERNEL(gather_rows_kernel)(const float* __restrict__ src, int D, const int* __restrict__ idx, long long n,
                            float* __restrict__ out) {
  GRID_STRIDE_LOOP(t, n * D) { out[t] = src[(size_t)idx[t / D] * D + (t % D)]; }
}
GKERNEL(scatter_add_rows_kernel)(const float* __restrict__ src, int D, const int* __restrict__ idx, long long n,
                                 float* __restrict__ out) {
  // an owner row can be wanted by several peers (or as several periodic images): atomic accumulate
  GRID_STRIDE_LOOP(t, n * D) { ATOMIC_ADD_F32(out + (size_t)idx[t / D] * D + (t % D), src[t]); }
}
// Halo push over peer memory (NVLink): row i of the packed order belongs to segment p (seg_off[p] <= i < seg_off[p+1],
// one segment per peer) and is written straight into that peer's halo buffer at row dst_row[p] + (i - seg_off[p]) --
// pack and transfer in one kernel, no staging copy and no collective call (pair_e3gnn_parallel.cpp:787-842 packs
// into a host buffer and hands it to MPI).  idx == nullptr: rows first_row + i (the contiguous ghost block).
struct SevennPeerTable { float* base[SEVENN_MAX_PEERS]; long long dst_row[SEVENN_MAX_PEERS]; long long seg_off[SEVENN_MAX_PEERS + 1]; int n; };
GKERNEL(push_rows_peers_kernel)(const float* __restrict__ src, int D, const int* __restrict__ idx, long long first_row,
                                long long n, SevennPeerTable T) {
  GRID_STRIDE_LOOP(t, n * D) {
    const long long i = t / D;
    const int c = (int)(t - i * D);
    int p = 0;
    while (p + 1 < T.n && i >= T.seg_off[p + 1]) ++p;
    const long long r = idx ? (long long)idx[i] : first_row + i;
    T.base[p][(size_t)(T.dst_row[p] + (i - T.seg_off[p])) * D + c] = src[(size_t)r * D + c];
  }
}
extern "C" int sevenn_push_rows_peers(const float* src, int D, const int* idx, long long first_row, long long n,
                                      int n_peers, float* const* peer_base, const long long* dst_row,
                                      const long long* seg_off, void* stream) {
  if (n <= 0) return 0;
  if (n_peers <= 0 || n_peers > SEVENN_MAX_PEERS) return SEVENN_ERR_BAD_ARG;
  SevennPeerTable T;
  T.n = n_peers;
  for (int p = 0; p < n_peers; ++p) { T.base[p] = peer_base[p]; T.dst_row[p] = dst_row[p]; T.seg_off[p] = seg_off[p]; }
  T.seg_off[n_peers] = seg_off[n_peers];
  if (T.seg_off[n_peers] != n) return SEVENN_ERR_BAD_ARG;
  LAUNCH1D(push_rows_peers_kernel, n * D, stream)(src, D, idx, first_row, n, T);
  return LAUNCH_STATUS();
}
extern "C" int sevenn_gather_rows(const float* src, int D, const int* idx, long long n, float* out, void* stream) {
  if (n <= 0) return 0;
  LAUNCH1D(gather_rows_kernel, n * D, stream)(src, D, idx, n, out);
  return LAUNCH_STATUS();
}
extern "C" int sevenn_scatter_add_rows(const float* src, int D, const int* idx, long long n, float* out, void* stream) {
  if (n <= 0) return 0;
  LAUNCH1D(scatter_add_rows_kernel, n * D, stream)(src, D, idx, n, out);
  return LAUNCH_STATUS();
}

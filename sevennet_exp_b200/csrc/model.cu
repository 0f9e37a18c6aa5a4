// Whole-model host in C++: (positions, cell, types) or an edge list -> energy, forces, virial, atomic energies.
//
// Replaces, behind one C entry point, what the reference runs per MD step in
//   sevenn/sevennet_calculator.py:163-202 (graph build + AtomGraphSequential.forward + autograd forces / stress) and
//   sevenn/pair_e3gnn/pair_e3gnn.cpp:71-274 (edge loop + torch::jit forward + result unpacking)
// with the launch sequence of sevennet_exp_b200/engine.py (tabulated-radial mode) and neighbor.py, so that a C++ host
// (a LAMMPS pair style, the TorchScript custom op in csrc/torch_op.cpp) needs no Python.  Properties:
//   * no allocation per call: every per-step buffer comes out of ONE caller-provided workspace (size from
//     sevenn_model_workspace_bytes); weights are uploaded once by sevenn_model_create;
//   * no host synchronisation: the neighbour-list size stays on the device (kernels read it there, arrays are sized by
//     the caller's edge capacity, overflow is reported in a status word), so a step is a pure stream of launches and can
//     be captured in a CUDA graph;
//   * multi-GPU hosts hook the per-layer ghost exchange in through two callbacks (pair_e3gnn_parallel.cpp:355-437).
// The model arrives as a flat (int32 program, fp32 data) blob written by sevennet_exp_b200/host.py::pack_engine.
#include "sevenn_common.cuh"
#include "../../include/sevenn_b200.h"

#include <dlfcn.h>

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace {

constexpr int MODEL_MAGIC = 0x37454E42;   // "7ENB"
constexpr int MODEL_VERSION = 1;

typedef int (*conv_fwd_tab_fn)(const float*, int, const float*, const float*, const int*, const int*, int, float, float*,
                               int, void*);
typedef int (*conv_bwd_tab_fn)(const float*, int, const float*, const float*, const int*, const int*, int, float,
                               const float*, int, float*, int, float*, long long, void*);

struct Blk {
  int l, in_off, out_off, K, N, in_mul, out_mul;      // offsets / group multiplicities in floats (ir_mul layout)
  const float *W, *WT, *Wimg, *WTimg, *cst;           // device; W [K,N] (or [S,K,N] by type); images may be null
};
struct Lin {
  int by_type_S = 0, din = 0, dout = 0, fwd_full = 0, bwd_needs_zero = 0, bwd_overlap = 0;
  std::vector<Blk> blk;
};
struct Layer {
  int dx_in, dmid, dgi, dout, units, zero_dxe;
  float inv_denom;
  const float* tab;
  conv_fwd_tab_fn fwd;
  conv_bwd_tab_fn bwd;
  std::vector<sevenn_gate_piece> pieces;
  Lin sc, si1, si2;
};

struct Cursor {
  const int* p; long long n, i = 0; bool ok = true;
  int next() { if (i >= n) { ok = false; return 0; } return p[i++]; }
  float nextf() { int v = next(); float f; std::memcpy(&f, &v, 4); return f; }
};

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// bump allocator over the caller's workspace; `dry` only measures the peak
struct Arena {
  char* base; size_t cap; bool dry;
  size_t off = 0, peak = 0; bool fail = false;
  void* raw(size_t bytes) {
    off = align_up(off, 256);
    char* p = base + off;
    off += bytes;
    if (off > peak) peak = off;
    if (!dry && off > cap) fail = true;
    return p;
  }
  float* f32(long long n) { return (float*)raw((size_t)(n > 0 ? n : 1) * 4); }
  int* i32(long long n) { return (int*)raw((size_t)(n > 0 ? n : 1) * 4); }
  double* f64(long long n) { return (double*)raw((size_t)(n > 0 ? n : 1) * 8); }
  size_t mark() const { return off; }
  void release(size_t m) { off = m; }
};

}  // namespace

struct sevenn_model {
  int lmax_sh, normalize_sph, S, tab_n, n_species, d0, species_wise, tc_min_rows, n_slab;
  float tab_inv_h, cutoff;
  float* dev = nullptr;          // all weights, one allocation
  float* img = nullptr;          // tcgen05 operand images of the block weights, rebuilt at create time
  cudaEvent_t ev_main = nullptr, ev_side = nullptr;   // ordering between the main stream and a host's halo side stream
  const float *embed, *scale, *shift;
  std::vector<Layer> layers;
  Lin read1, read2;
  void* self = nullptr;          // dlopen handle of this library (conv kernels are looked up by signature)
  long long n_launch = 0, n_tc_launch = 0;
  std::string error;
};

namespace {

#define RUN(expr)                                  \
  do {                                             \
    if (!A.dry) {                                  \
      int rc_ = (expr);                            \
      ++m->n_launch;                               \
      if (rc_) return rc_;                         \
    }                                              \
  } while (0)

int zero_async(Arena& A, void* p, size_t bytes, cudaStream_t st) {
  if (A.dry || bytes == 0) return 0;
  return (int)cudaMemsetAsync(p, 0, bytes, st);
}

bool read_lin(Cursor& c, const float* dev, Lin& L) {
  L.by_type_S = c.next(); L.din = c.next(); L.dout = c.next();
  L.fwd_full = c.next(); L.bwd_needs_zero = c.next(); L.bwd_overlap = c.next();
  const int nb = c.next();
  if (!c.ok || nb < 0 || nb > 64) return false;
  L.blk.resize(nb);
  for (Blk& b : L.blk) {
    b.l = c.next(); b.in_off = c.next(); b.out_off = c.next(); b.K = c.next(); b.N = c.next();
    b.in_mul = c.next(); b.out_mul = c.next();
    const int w = c.next(), wt = c.next(), wi = c.next(), wti = c.next(), cs = c.next();
    b.W = dev + w; b.WT = dev + wt;
    b.Wimg = wi >= 0 ? dev + wi : nullptr; b.WTimg = wti >= 0 ? dev + wti : nullptr;
    b.cst = cs >= 0 ? dev + cs : nullptr;
  }
  return c.ok && c.next() == 0x4C494E;   // section sentinel "LIN"
}

// y[:, cols] += row (constant rows of the scalar blocks: fixed modality one-hot, o3.Linear bias)
__global__ void add_row_kernel(float* __restrict__ y, int ld, long long n, int N, const float* __restrict__ row) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n * N; i += (long long)gridDim.x * blockDim.x)
    y[(i / N) * ld + (i % N)] += row[i % N];
}

bool tc_eligible(const sevenn_model* m, const sevenn_gemm_args& g, const float* img) {
  return img != nullptr && g.M >= m->tc_min_rows && (g.K & 3) == 0 && (g.sAn & 3) == 0 && (g.sAm & 3) == 0 &&
         ((uintptr_t)g.A & 15) == 0;
}

// exact-fp32 blocks with disjoint outputs: one launch for the group (engine.py::_gemm_group)
int flush_simt(sevenn_model* m, sevenn_gemm_args* sg, int& n, cudaStream_t st) {
  if (n == 0) return 0;
  long long tiles = 0;
  for (int i = 0; i < n; ++i) tiles += ((sg[i].M + 15) / 16) * ((sg[i].N + 31) / 32);
  int rc = n == 1 ? sevenn_gemm_f32(&sg[0], st) : sevenn_gemm_f32_grouped(n, sg, st);
  m->n_launch += (n == 1 || tiles <= SEVENN_GEMM_GROUP_MAX_TILES) ? 1 : n;
  n = 0;
  return rc;
}

// o3.Linear forward: y (+)= x @ W per block (engine.py::_linear_fwd)
// `ext`: collector shared with the caller's next linear (disjoint outputs): exact-fp32 blocks are left in it unflushed
struct SimtBatch {
  sevenn_gemm_args g[SEVENN_GEMM_GROUP_MAX]; int n = 0;                                       // exact-fp32 arm
  sevenn_gemm_args tg[SEVENN_GEMM_GROUP_MAX]; const float* timg[SEVENN_GEMM_GROUP_MAX]; int nt = 0;   // tcgen05 arm
};
int flush_tc(sevenn_model* m, SimtBatch& b, cudaStream_t st) {
  if (b.nt == 0) return 0;
  int rc = sevenn_gemm_tc_grouped(b.nt, b.tg, b.timg, st);
  ++m->n_launch; ++m->n_tc_launch; b.nt = 0;
  return rc;
}
int linear_fwd(sevenn_model* m, Arena& A, const Lin& L, const float* x, long long n, float* y, bool accumulate,
               const int* types, cudaStream_t st, SimtBatch* ext = nullptr) {
  if (A.dry || n <= 0) return 0;
  SimtBatch own;
  SimtBatch& sb = ext != nullptr ? *ext : own;
  sevenn_gemm_args* sgrp = sb.g; int& nsg = sb.n;
  sevenn_gemm_args* grp = sb.tg; const float** img = sb.timg; int& ng = sb.nt;
  const int tc_cap = ext != nullptr ? SEVENN_GEMM_GROUP_MAX : 4;      // a linear on its own keeps groups of <= 4 (one per l)
  auto flush = [&]() -> int { return flush_tc(m, sb, st); };
  for (const Blk& b : L.blk) {
    const int R = 2 * b.l + 1;
    if (L.by_type_S) {
      int rc = sevenn_linear_by_type(x + b.in_off, L.din, b.in_mul, b.W, types, y + b.out_off, L.dout, b.out_mul, n,
                                     R, b.K, b.N, accumulate ? 1 : 0, st);
      ++m->n_launch;
      if (rc) return rc;
      continue;
    }
    sevenn_gemm_args g;
    std::memset(&g, 0, sizeof(g));
    g.A = x + b.in_off; g.sAn = L.din; g.sAm = b.in_mul; g.B = b.W; g.ldb = b.N;
    g.C = y + b.out_off; g.sCn = L.dout; g.sCm = b.out_mul; g.M = n * R; g.R = R; g.N = b.N; g.K = b.K;
    g.epi = accumulate ? 1 : 0; g.cst = 1.f;
    if (tc_eligible(m, g, b.Wimg)) {
      if (ng > 0 && grp[0].epi != g.epi) { int rc = flush(); if (rc) return rc; }     // one epilogue per launch
      grp[ng] = g; img[ng] = b.Wimg;
      if (++ng == tc_cap) { int rc = flush(); if (rc) return rc; }
    } else {
      sgrp[nsg] = g;
      if (++nsg == SEVENN_GEMM_GROUP_MAX) { int rc = flush_simt(m, sgrp, nsg, st); if (rc) return rc; }
    }
  }
  int rc = 0;
  bool has_cst = false;
  for (const Blk& b : L.blk) has_cst = has_cst || b.cst != nullptr;
  if (ext == nullptr || has_cst) {
    rc = flush(); if (rc) return rc;
    rc = flush_simt(m, sgrp, nsg, st); if (rc) return rc;
  }
  for (const Blk& b : L.blk)
    if (b.cst != nullptr) {
      add_row_kernel<<<sevenn_grid_for(n * b.N), 256, 0, st>>>(y + b.out_off, L.dout, n, b.N, b.cst);
      ++m->n_launch;
    }
  return (int)cudaGetLastError();
}

// dx (+)= dy @ W^T per block (engine.py::_linear_bwd); dx zeroed by the caller when L.bwd_needs_zero
int linear_bwd(sevenn_model* m, Arena& A, const Lin& L, const float* dy, long long n, float* dx, bool accumulate,
               const int* types, cudaStream_t st) {
  if (A.dry || n <= 0) return 0;
  const bool acc = accumulate || L.bwd_needs_zero;
  sevenn_gemm_args grp[4]; const float* img[4]; int ng = 0;
  sevenn_gemm_args sgrp[SEVENN_GEMM_GROUP_MAX]; int nsg = 0;
  auto flush = [&]() -> int {
    if (ng == 0) return 0;
    int rc = sevenn_gemm_tc_grouped(ng, grp, img, st);
    ++m->n_launch; ++m->n_tc_launch; ng = 0;
    return rc;
  };
  for (const Blk& b : L.blk) {
    const int R = 2 * b.l + 1;
    if (L.by_type_S) {
      int rc = sevenn_linear_by_type(dy + b.out_off, L.dout, b.out_mul, b.WT, types, dx + b.in_off, L.din, b.in_mul, n,
                                     R, b.N, b.K, acc ? 1 : 0, st);
      ++m->n_launch;
      if (rc) return rc;
      continue;
    }
    sevenn_gemm_args g;
    std::memset(&g, 0, sizeof(g));
    g.A = dy + b.out_off; g.sAn = L.dout; g.sAm = b.out_mul; g.B = b.WT; g.ldb = b.K;
    g.C = dx + b.in_off; g.sCn = L.din; g.sCm = b.in_mul; g.M = n * R; g.R = R; g.N = b.K; g.K = b.N;
    g.epi = acc ? 1 : 0; g.cst = 1.f;
    const bool tc = tc_eligible(m, g, b.WTimg);
    if (tc && !L.bwd_overlap) {
      grp[ng] = g; img[ng] = b.WTimg;
      if (++ng == 4) { int rc = flush(); if (rc) return rc; }
    } else if (!tc && !L.bwd_overlap) {
      sgrp[nsg] = g;
      if (++nsg == SEVENN_GEMM_GROUP_MAX) { int rc = flush_simt(m, sgrp, nsg, st); if (rc) return rc; }
    } else {
      // blocks that accumulate into one input group stay ordered, one launch each
      int rc = tc ? sevenn_gemm_tc(&g, b.WTimg, st) : sevenn_gemm_f32(&g, st);
      ++m->n_launch;
      if (tc) ++m->n_tc_launch;
      if (rc) return rc;
    }
  }
  int rc = flush_simt(m, sgrp, nsg, st);
  if (rc) return rc;
  return flush();
}

struct GraphView {
  long long n_local, n_nodes, max_edges;
  const int* types; const float* edge_vec; const int *rowptr, *col, *cptr, *cedge;
  const int* n_edges_dev;       // device: number of valid edges (<= max_edges)
};

// engine.py::Engine.evaluate, tabulated-radial mode
int evaluate_graph(sevenn_model* m, Arena& A, const GraphView& g, const sevenn_halo_hooks* halo, cudaStream_t st,
                   double* out_energy, float* out_atomic, float* out_forces, double* out_virial6,
                   float* out_virial_atoms) {
  const long long nl = g.n_local, nn = g.n_nodes, E = g.max_edges;
  const int S = m->S, SD = S + 1;
  void* vs = (void*)st;
  float* Y = A.f32(E * (S + 2));
  RUN(sevenn_edge_fwd_tab_dev(g.edge_vec, E, g.n_edges_dev, m->lmax_sh, m->normalize_sph, m->tab_inv_h, m->tab_n, Y, vs));
  int max_x = m->d0;
  for (const Layer& L : m->layers) max_x = L.dout > max_x ? L.dout : max_x;
  float* xbuf[2] = {A.f32(nn * max_x), A.f32(nn * max_x)};
  float* x = xbuf[0];
  RUN(sevenn_embed(g.types, nn, m->d0, m->embed, x, vs));
  long long n_x = nn;
  const int nL = (int)m->layers.size();
  std::vector<float*> pre(nL), h(nL);
  for (int t = 0; t < nL; ++t) {
    Layer& L = m->layers[t];
    pre[t] = A.f32(nl * L.dgi);
    if (!L.sc.fwd_full) { int rc = zero_async(A, pre[t], (size_t)nl * L.dgi * 4, st); if (rc) return rc; }
    SimtBatch both;                 // sc and si1 read the same x and write different buffers: one launch on small systems
    int rc = linear_fwd(m, A, L.sc, x, nl, pre[t], false, g.types, st, &both);
    if (rc) return rc;
    h[t] = A.f32(nn * L.dx_in);
    if (!L.si1.fwd_full) { rc = zero_async(A, h[t], (size_t)nn * L.dx_in * 4, st); if (rc) return rc; }
    rc = linear_fwd(m, A, L.si1, x, n_x, h[t], false, nullptr, st, &both);
    if (rc) return rc;
    if (!A.dry) {
      rc = flush_tc(m, both, st); if (rc) return rc;
      rc = flush_simt(m, both.g, both.n, st); if (rc) return rc;
    }
    // ghost exchange: in line, or on the side stream while the rows without ghost neighbours are convolved
    long long n_int = 0;
    if (t > 0 && halo != nullptr && halo->forward != nullptr && !A.dry) {
      const bool overlap = halo->side_stream != nullptr && halo->n_interior > 0 && halo->n_interior <= nl &&
                           m->ev_main != nullptr;
      if (overlap) {
        cudaStream_t side = (cudaStream_t)halo->side_stream;
        if (cudaEventRecord(m->ev_main, st) != cudaSuccess || cudaStreamWaitEvent(side, m->ev_main, 0) != cudaSuccess)
          return (int)cudaGetLastError();
        rc = halo->forward(halo->user, t, h[t], L.dx_in, halo->side_stream);
        if (rc) return rc;
        if (cudaEventRecord(m->ev_side, side) != cudaSuccess) return (int)cudaGetLastError();
        n_int = halo->n_interior;
      } else {
        rc = halo->forward(halo->user, t, h[t], L.dx_in, vs);      // fills rows [n_local, n_nodes)
        if (rc) return rc;
      }
    }
    const size_t mk = A.mark();
    float* agg = A.f32(nl * L.dmid);
    if (n_int > 0) {
      RUN(L.fwd(h[t], L.dx_in, Y, L.tab, g.rowptr, g.col, (int)n_int, L.inv_denom, agg, L.dmid, vs));
      if (cudaStreamWaitEvent(st, m->ev_side, 0) != cudaSuccess) return (int)cudaGetLastError();
      if (nl > n_int)
        RUN(L.fwd(h[t], L.dx_in, Y, L.tab, g.rowptr + n_int, g.col, (int)(nl - n_int), L.inv_denom,
                  agg + n_int * L.dmid, L.dmid, vs));
    } else if (nl > 0) {
      RUN(L.fwd(h[t], L.dx_in, Y, L.tab, g.rowptr, g.col, (int)nl, L.inv_denom, agg, L.dmid, vs));
    }
    rc = linear_fwd(m, A, L.si2, agg, nl, pre[t], true, nullptr, st);
    if (rc) return rc;
    A.release(mk);
    float* xo = xbuf[(t + 1) & 1];
    if (nl > 0) RUN(sevenn_gate_fwd(pre[t], L.dgi, xo, L.dout, nl, L.pieces.data(), (int)L.pieces.size(), vs));
    x = xo; n_x = nl;
  }
  // readout, rescale, total energy
  float* r1 = A.f32(nl * m->read1.dout);
  int rc = linear_fwd(m, A, m->read1, x, nl, r1, false, nullptr, st);
  if (rc) return rc;
  float* es = A.f32(nl);
  rc = linear_fwd(m, A, m->read2, r1, nl, es, false, nullptr, st);
  if (rc) return rc;
  float* seed = A.f32(nl);
  double* part = A.f64(296 * 8);
  RUN(sevenn_rescale(es, g.types, nl, m->scale, m->shift, m->species_wise, out_atomic, seed, vs));
  RUN(sevenn_colsum_f64(out_atomic, nl, 1, part, out_energy, vs));
  if (out_forces == nullptr) return 0;

  // ---- reverse pass
  float* dY = A.f32((long long)m->n_slab * E * SD);
  rc = zero_async(A, dY, (size_t)m->n_slab * E * SD * 4, st);
  if (rc) return rc;
  float* dr1 = A.f32(nl * m->read1.dout);
  rc = linear_bwd(m, A, m->read2, seed, nl, dr1, false, nullptr, st);
  if (rc) return rc;
  int max_dx = m->read1.din;
  for (const Layer& L : m->layers) max_dx = L.dx_in > max_dx ? L.dx_in : max_dx;
  float* dxbuf[2] = {A.f32(nl * max_dx), A.f32(nl * max_dx)};
  float* dx = dxbuf[0];
  if (m->read1.bwd_needs_zero) { rc = zero_async(A, dx, (size_t)nl * m->read1.din * 4, st); if (rc) return rc; }
  rc = linear_bwd(m, A, m->read1, dr1, nl, dx, false, nullptr, st);
  if (rc) return rc;
  int flip = 0;
  for (int t = nL - 1; t >= 0; --t) {
    Layer& L = m->layers[t];
    const size_t mk = A.mark();
    float* dpre = A.f32(nl * L.dgi);
    if (nl > 0) RUN(sevenn_gate_bwd(pre[t], L.dgi, dx, L.dout, dpre, nl, L.pieces.data(), (int)L.pieces.size(), vs));
    float* dagg = A.f32(nl * L.dmid);
    if (L.si2.bwd_needs_zero) { rc = zero_async(A, dagg, (size_t)nl * L.dmid * 4, st); if (rc) return rc; }
    rc = linear_bwd(m, A, L.si2, dpre, nl, dagg, false, nullptr, st);
    if (rc) return rc;
    float* dxe = nullptr;                        // layer 0: dE/dx is not needed, the kernel skips it (dxe == NULL)
    if (t > 0) {
      dxe = A.f32(E * L.dx_in);
      if (L.zero_dxe) { rc = zero_async(A, dxe, (size_t)E * L.dx_in * 4, st); if (rc) return rc; }
    }
    if (nl > 0)
      RUN(L.bwd(h[t], L.dx_in, Y, L.tab, g.rowptr, g.col, (int)nl, L.inv_denom, dagg, L.dmid, dxe, L.dx_in, dY,
                E * SD, vs));
    if (t == 0) { A.release(mk); break; }        // layer-0 inputs do not depend on positions
    float* dh = A.f32(nn * L.dx_in);
    RUN(sevenn_segment_gather_sum(dxe, L.dx_in, g.cptr, g.cedge, nn, dh, 0, vs));
    if (halo != nullptr && halo->backward != nullptr && !A.dry) {
      rc = halo->backward(halo->user, t, dh, L.dx_in, vs);       // ghost rows -> owners (added into rows [0, n_local))
      if (rc) return rc;
    }
    flip ^= 1;
    float* dxn = dxbuf[flip];
    if (L.si1.bwd_needs_zero) { rc = zero_async(A, dxn, (size_t)nl * L.dx_in * 4, st); if (rc) return rc; }
    rc = linear_bwd(m, A, L.si1, dh, nl, dxn, false, nullptr, st);
    if (rc) return rc;
    rc = linear_bwd(m, A, L.sc, dpre, nl, dxn, true, g.types, st);
    if (rc) return rc;
    dx = dxn;
    A.release(mk);
  }
  float* fvec = A.f32(E * 3);
  RUN(sevenn_edge_bwd_tab_dev(g.edge_vec, E, g.n_edges_dev, m->lmax_sh, m->normalize_sph, m->tab_inv_h, dY, m->n_slab,
                              E * SD, fvec, vs));
  float* vir = out_virial_atoms != nullptr ? out_virial_atoms : A.f32(nn * 6);
  RUN(sevenn_force_virial(fvec, g.edge_vec, g.rowptr, (int)nl, g.cptr, g.cedge, nn, out_forces, vir, vs));
  RUN(sevenn_colsum_f64(vir, nn, 6, part, out_virial6, vs));
  return 0;
}

// ---- neighbour list parameters (neighbor.py::nl_params)
bool inv3(const double* a, double* r) {
  const double det = a[0] * (a[4] * a[8] - a[5] * a[7]) - a[1] * (a[3] * a[8] - a[5] * a[6]) + a[2] * (a[3] * a[7] - a[4] * a[6]);
  if (det == 0.0) return false;
  const double id = 1.0 / det;
  r[0] = (a[4] * a[8] - a[5] * a[7]) * id; r[1] = (a[2] * a[7] - a[1] * a[8]) * id; r[2] = (a[1] * a[5] - a[2] * a[4]) * id;
  r[3] = (a[5] * a[6] - a[3] * a[8]) * id; r[4] = (a[0] * a[8] - a[2] * a[6]) * id; r[5] = (a[2] * a[3] - a[0] * a[5]) * id;
  r[6] = (a[3] * a[7] - a[4] * a[6]) * id; r[7] = (a[1] * a[6] - a[0] * a[7]) * id; r[8] = (a[0] * a[4] - a[1] * a[3]) * id;
  return true;
}

int nl_params(const double* pos_host, long long n, const double* cell_in, const int* pbc, double cutoff, long long max_bins,
              double* dpar, int* ipar) {
  double cell[9];
  std::memcpy(cell, cell_in, sizeof(cell));
  const bool open = !(pbc[0] && pbc[1] && pbc[2]);
  if (open && pos_host == nullptr && n > 0) return SEVENN_ERR_BAD_ARG;   // extents of the open directions are needed
  double ext = 1.0;
  if (open) {                                    // plain comparisons, four independent maxima: ~0.3 ns per coordinate
    double m0 = 0.0, m1 = 0.0, m2 = 0.0;
    for (long long i = 0; i < n; ++i) {
      const double a = std::fabs(pos_host[3 * i]), b = std::fabs(pos_host[3 * i + 1]), c = std::fabs(pos_host[3 * i + 2]);
      m0 = a > m0 ? a : m0; m1 = b > m1 ? b : m1; m2 = c > m2 ? c : m2;
    }
    const double mm = m0 > m1 ? (m0 > m2 ? m0 : m2) : (m1 > m2 ? m1 : m2);
    ext = mm + 1.0 > ext ? mm + 1.0 : ext;
  }
  for (int d = 0; d < 3; ++d)
    if (!pbc[d]) {            // like dataload.py:40-48: any spanning vector will do
      cell[3 * d] = cell[3 * d + 1] = cell[3 * d + 2] = 0.0;
      cell[3 * d + d] = ext * 5 * cutoff;
    }
  double inv[9];
  if (!inv3(cell, inv)) return SEVENN_ERR_BAD_ARG;
  double h[3], lo[3] = {0, 0, 0}, inv_span[3] = {1, 1, 1};
  long long nb[3] = {1, 1, 1};
  int R[3] = {1, 1, 1};
  for (int d = 0; d < 3; ++d) h[d] = 1.0 / std::sqrt(inv[d] * inv[d] + inv[3 + d] * inv[3 + d] + inv[6 + d] * inv[6 + d]);
  // fractional extents of the open directions in ONE pass over the positions (this runs on the host before the first
  // launch of every step: at 600 k atoms three fmin / fmax passes cost ~10 ms of an end-to-end step)
  double fa[3] = {0.0, 0.0, 0.0}, fb[3] = {1.0, 1.0, 1.0};
  if (open && n > 0) {
    for (int d = 0; d < 3; ++d)
      fa[d] = fb[d] = pos_host[0] * inv[d] + pos_host[1] * inv[3 + d] + pos_host[2] * inv[6 + d];
    for (long long i = 1; i < n; ++i) {
      const double x = pos_host[3 * i], y = pos_host[3 * i + 1], z = pos_host[3 * i + 2];
      for (int d = 0; d < 3; ++d) {
        const double f = x * inv[d] + y * inv[3 + d] + z * inv[6 + d];
        fa[d] = f < fa[d] ? f : fa[d]; fb[d] = f > fb[d] ? f : fb[d];
      }
    }
  }
  for (int d = 0; d < 3; ++d) {
    if (pbc[d]) {
      nb[d] = (long long)std::floor(h[d] / cutoff);
    } else {
      const double a = fa[d], b = fb[d];
      const double span = std::fmax(b - a, 1e-9) * (1 + 1e-9) + 1e-12;
      lo[d] = a; inv_span[d] = 1.0 / span;
      nb[d] = (long long)std::floor(span * h[d] / cutoff);
    }
    if (nb[d] < 1) nb[d] = 1;
  }
  while (nb[0] * nb[1] * nb[2] > max_bins) {       // bins only ever get coarser, which stays correct
    int d = 0;
    if (nb[1] > nb[d]) d = 1;
    if (nb[2] > nb[d]) d = 2;
    nb[d] = nb[d] / 2 > 1 ? nb[d] / 2 : 1;
  }
  for (int d = 0; d < 3; ++d) {
    const double width = pbc[d] ? h[d] / nb[d] : h[d] / (inv_span[d] * nb[d]);
    R[d] = (int)std::ceil(cutoff / width - 1e-12);
    if (!pbc[d] && R[d] > nb[d]) R[d] = (int)nb[d];
    if (R[d] + 2 >= 120) return SEVENN_ERR_BAD_ARG;   // cell far smaller than the cutoff: shifts do not fit the key
  }
  for (int k = 0; k < 9; ++k) { dpar[k] = cell[k]; dpar[9 + k] = inv[k]; }
  for (int d = 0; d < 3; ++d) { dpar[18 + d] = lo[d]; dpar[21 + d] = inv_span[d]; ipar[d] = pbc[d] ? 1 : 0; ipar[3 + d] = (int)nb[d]; ipar[6 + d] = R[d]; }
  dpar[24] = cutoff * cutoff;
  return 0;
}

struct NlOut {
  int *rowptr, *col, *cptr, *cedge, *center, *shift;
  float* vec;
};

// neighbor.py::NeighborList.build without the host round trip: rows for atoms [0, n_rows), edge arrays of capacity
// max_edges, status[0] |= 1 on overflow, status[1] = number of edges found (before clamping)
int build_neighbors(sevenn_model* m, Arena& A, long long n, long long n_rows, const double* pos_dev,
                    const double* pos_host, const double* cell, const int* pbc, long long max_edges, int* status,
                    NlOut& o, cudaStream_t st) {
  double dpar[25]; int ipar[9];
  long long max_bins = 8 * n > 64 ? 8 * n : 64;
  int rc = 0;
  long long nbins = max_bins;                   // dry run (workspace sizing): the largest grid nl_params can choose
  if (!A.dry) {
    rc = nl_params(pos_host, n, cell, pbc, (double)m->cutoff, max_bins, dpar, ipar);
    if (rc) return rc;
    nbins = (long long)ipar[3] * ipar[4] * ipar[5];
  }
  void* vs = (void*)st;
  double* pw = A.f64(3 * n);
  const long long n_tmp = sevenn_scan_tmp_ints((int)((nbins > n ? nbins : n) + 1)) + 1;
  const size_t z0 = A.mark();
  int* wrap = A.i32(3 * n); int* atom_bin = A.i32(n); int* bin_count = A.i32(nbins + 1); int* bin_start = A.i32(nbins + 1);
  int* bin_fill = A.i32(nbins); int* bin_atoms = A.i32(n); int* tmp = A.i32(n_tmp); int* count = A.i32(n_rows + 1);
  o.rowptr = A.i32(n_rows + 1);
  const long long n_tmp2 = sevenn_scan_tmp_ints((int)(n + 1)) + 1;
  int* cnt = A.i32(n + 1); int* fill = A.i32(n); o.cptr = A.i32(n + 1); int* tmp2 = A.i32(n_tmp2);
  const size_t z1 = A.mark();
  rc = zero_async(A, A.base + align_up(z0, 256), z1 - align_up(z0, 256), st);     // one fill for all int arrays
  if (rc) return rc;
  unsigned long long* keys = (unsigned long long*)A.raw((size_t)(max_edges > 0 ? max_edges : 1) * 8);
  o.center = A.i32(max_edges); o.col = A.i32(max_edges); o.shift = A.i32(3 * max_edges); o.vec = A.f32(3 * max_edges);
  o.cedge = A.i32(max_edges);
  if (n > 0) {
    RUN(sevenn_nl_bin(pos_dev, (int)n, dpar, ipar, pw, wrap, atom_bin, bin_count, vs));
    RUN(sevenn_scan_i32(bin_count, bin_start, (int)nbins, tmp, vs));
    RUN(sevenn_nl_scatter(atom_bin, (int)n, bin_start, bin_fill, bin_atoms, vs));
    RUN(sevenn_nl_count(pw, wrap, atom_bin, bin_start, bin_atoms, (int)n_rows, dpar, ipar, count, vs));
    RUN(sevenn_scan_i32(count, o.rowptr, (int)n_rows, tmp, vs));
    if (!A.dry) cudaMemcpyAsync(status + 1, o.rowptr + n_rows, sizeof(int), cudaMemcpyDeviceToDevice, st);
    RUN(sevenn_nl_fill_capped(pw, wrap, atom_bin, bin_start, bin_atoms, (int)n_rows, dpar, ipar, o.rowptr, max_edges,
                              status, keys, pos_dev, o.center, o.col, o.shift, o.vec, vs));
    RUN(sevenn_csc_build_dev(o.col, max_edges, o.rowptr + n_rows, (int)n, cnt, fill, o.cptr, keys, o.cedge, tmp2, vs));
  }
  return 0;
}

int check_args(const sevenn_model* m, const void* ws, long long ws_bytes) {
  if (m == nullptr || ws_bytes < 0 || (ws == nullptr && ws_bytes > 0)) return SEVENN_ERR_BAD_ARG;
  return 0;
}

}  // namespace

// ------------------------------------------------------------------------------------------------ C ABI
extern "C" int sevenn_model_create(const int* prog, long long n_prog, const float* data, long long n_data,
                                   sevenn_model** out) {
  if (prog == nullptr || data == nullptr || out == nullptr || n_prog < 16) return SEVENN_ERR_BAD_ARG;
  *out = nullptr;
  Cursor c{prog, n_prog};
  if (c.next() != MODEL_MAGIC || c.next() != MODEL_VERSION) return SEVENN_ERR_BAD_ARG;
  sevenn_model* m = new sevenn_model();
  Dl_info info;
  if (dladdr((void*)&sevenn_abi_version, &info) && info.dli_fname)
    m->self = dlopen(info.dli_fname, RTLD_NOW | RTLD_NOLOAD);
  if (m->self == nullptr) m->self = dlopen(nullptr, RTLD_NOW);
  if (cudaMalloc(&m->dev, (size_t)n_data * 4) != cudaSuccess ||
      cudaMemcpy(m->dev, data, (size_t)n_data * 4, cudaMemcpyHostToDevice) != cudaSuccess) {
    delete m;
    return (int)cudaErrorMemoryAllocation;
  }
  const float* dev = m->dev;
  m->lmax_sh = c.next(); m->normalize_sph = c.next(); m->tab_n = c.next(); m->n_species = c.next(); m->d0 = c.next();
  m->species_wise = c.next(); m->tc_min_rows = c.next();
  m->tab_inv_h = c.nextf(); m->cutoff = c.nextf();
  m->S = (m->lmax_sh + 1) * (m->lmax_sh + 1);
  m->embed = dev + c.next(); m->scale = dev + c.next(); m->shift = dev + c.next();
  const int nL = c.next();
  bool ok = c.ok && nL >= 1 && nL <= 64;
  m->n_slab = 1;
  for (int t = 0; ok && t < nL; ++t) {
    Layer L;
    L.dx_in = c.next(); L.dmid = c.next(); L.dgi = c.next(); L.dout = c.next(); L.units = c.next(); L.zero_dxe = c.next();
    L.inv_denom = c.nextf();
    L.tab = dev + c.next();
    const int ns = c.next();
    std::string sig;
    for (int k = 0; k < ns && k < 64; ++k) sig.push_back((char)c.next());
    L.fwd = (conv_fwd_tab_fn)dlsym(m->self, ("sevenn_conv_fwd_tab_" + sig).c_str());
    L.bwd = (conv_bwd_tab_fn)dlsym(m->self, ("sevenn_conv_bwd_tab_" + sig).c_str());
    if (L.fwd == nullptr || L.bwd == nullptr) {
      std::fprintf(stderr, "sevenn_model_create: no prebuilt conv kernel for signature %s\n", sig.c_str());
      ok = false;
      break;
    }
    const int np = c.next();
    if (np < 0 || np > SEVENN_MAX_GATE_PIECES) { ok = false; break; }
    L.pieces.resize(np);
    for (sevenn_gate_piece& p : L.pieces) {
      p.kind = c.next(); p.l = c.next(); p.mul = c.next(); p.in_off = c.next(); p.in_stride = c.next();
      p.gate_off = c.next(); p.out_off = c.next(); p.out_stride = c.next(); p.act = c.next(); p.cst = c.nextf();
    }
    ok = read_lin(c, dev, L.sc) && read_lin(c, dev, L.si1) && read_lin(c, dev, L.si2) && c.ok;
    if (L.units > m->n_slab) m->n_slab = L.units;
    m->layers.push_back(std::move(L));
  }
  ok = ok && read_lin(c, dev, m->read1) && read_lin(c, dev, m->read2) && c.next() == 0x454E44 && c.ok;   // "END"
  if (!ok) {
    cudaFree(m->dev);
    delete m;
    return SEVENN_ERR_BAD_ARG;
  }
  // The tcgen05 operand images depend on the tile plan of THIS library build (sevenn_gemm_tc_plan); the ones packed
  // into the program may come from another build (a deployed artefact outlives the library), so they are rebuilt here
  // from W / W^T on the device -- the packed images only say which blocks want the tensor-core arm.
  {
    std::vector<Lin*> lins;
    for (Layer& L : m->layers) { lins.push_back(&L.sc); lins.push_back(&L.si1); lins.push_back(&L.si2); }
    lins.push_back(&m->read1); lins.push_back(&m->read2);
    long long total = 0;
    auto img_floats = [](int K, int N) { int a, b, c; long long f = 0; sevenn_gemm_tc_plan(N, K, &a, &b, &c, &f); return (f + 63) / 64 * 64; };
    for (Lin* L : lins) for (Blk& b : L->blk) {
      if (b.Wimg != nullptr) total += img_floats(b.K, b.N);
      if (b.WTimg != nullptr) total += img_floats(b.N, b.K);
    }
    int rc = 0;
    if (total > 0 && cudaMalloc(&m->img, (size_t)total * 4) != cudaSuccess) rc = (int)cudaErrorMemoryAllocation;
    long long off = 0;
    for (Lin* L : lins) for (Blk& b : L->blk) {
      if (rc) break;
      if (b.Wimg != nullptr) {
        rc = sevenn_gemm_tc_image(b.W, b.K, b.N, m->img + off, nullptr);
        b.Wimg = m->img + off; off += img_floats(b.K, b.N);
      }
      if (!rc && b.WTimg != nullptr) {
        rc = sevenn_gemm_tc_image(b.WT, b.N, b.K, m->img + off, nullptr);
        b.WTimg = m->img + off; off += img_floats(b.N, b.K);
      }
    }
    if (!rc && cudaDeviceSynchronize() != cudaSuccess) rc = (int)cudaGetLastError();
    if (rc) { cudaFree(m->img); cudaFree(m->dev); delete m; return rc; }
  }
  const char* env = std::getenv("SEVENN_TC_MIN_ROWS");
  if (env != nullptr) m->tc_min_rows = std::atoi(env);
  if (cudaEventCreateWithFlags(&m->ev_main, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&m->ev_side, cudaEventDisableTiming) != cudaSuccess) {
    m->ev_main = m->ev_side = nullptr;           // no overlap then; everything else works
    cudaGetLastError();
  }
  *out = m;
  return 0;
}

extern "C" int sevenn_model_destroy(sevenn_model* m) {
  if (m == nullptr) return 0;
  if (m->ev_main != nullptr) cudaEventDestroy(m->ev_main);
  if (m->ev_side != nullptr) cudaEventDestroy(m->ev_side);
  cudaFree(m->img);
  cudaFree(m->dev);
  delete m;
  return 0;
}

extern "C" int sevenn_model_set_tc_min_rows(sevenn_model* m, int rows) {
  if (m == nullptr) return SEVENN_ERR_BAD_ARG;
  m->tc_min_rows = rows;
  return 0;
}

extern "C" long long sevenn_model_launches(const sevenn_model* m) { return m ? m->n_launch : 0; }
extern "C" long long sevenn_model_tc_launches(const sevenn_model* m) { return m ? m->n_tc_launch : 0; }

extern "C" long long sevenn_model_workspace_bytes(sevenn_model* m, long long n_local, long long n_nodes,
                                                  long long max_edges, int with_neighbor_list) {
  if (m == nullptr || n_local < 0 || n_nodes < n_local || max_edges < 0) return -1;
  Arena A{nullptr, 0, true};
  GraphView g{n_local, n_nodes, max_edges, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  if (with_neighbor_list) {
    NlOut o;
    if (build_neighbors(m, A, n_nodes, n_local, nullptr, nullptr, nullptr, nullptr, max_edges, nullptr, o, nullptr)) return -1;
  }
  double e; float f;
  int rc = evaluate_graph(m, A, g, nullptr, nullptr, &e, &f, &f, &e, nullptr);
  if (rc) return -1;
  return (long long)align_up(A.peak, 256) + 4096;
}

extern "C" int sevenn_model_evaluate_graph(sevenn_model* m, const sevenn_graph* graph, const sevenn_halo_hooks* halo,
                                           void* workspace, long long workspace_bytes, void* stream,
                                           const sevenn_outputs* out) {
  int rc = check_args(m, workspace, workspace_bytes);
  if (rc || graph == nullptr || out == nullptr || out->energy == nullptr || out->atomic_energy == nullptr)
    return SEVENN_ERR_BAD_ARG;
  if (graph->n_edges_dev == nullptr || graph->n_local < 0 || graph->n_nodes < graph->n_local) return SEVENN_ERR_BAD_ARG;
  Arena A{(char*)workspace, (size_t)workspace_bytes, false};
  // the arena is checked before anything is launched
  {
    Arena D{nullptr, 0, true};
    GraphView gd{graph->n_local, graph->n_nodes, graph->max_edges, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    double e; float f;
    evaluate_graph(m, D, gd, nullptr, nullptr, &e, &f, out->forces ? &f : nullptr, &e, nullptr);
    if (D.peak > (size_t)workspace_bytes) return SEVENN_ERR_WORKSPACE;
  }
  GraphView g{graph->n_local, graph->n_nodes, graph->max_edges, graph->types, graph->edge_vec, graph->rowptr, graph->col,
              graph->cptr, graph->cedge, graph->n_edges_dev};
  return evaluate_graph(m, A, g, halo, (cudaStream_t)stream, out->energy, out->atomic_energy, out->forces, out->virial6,
                        out->virial_atoms);
}

extern "C" int sevenn_model_evaluate_positions(sevenn_model* m, long long n_local, long long n_nodes, const double* pos,
                                               const double* pos_host, const double* cell /* host, 9 */,
                                               const int* pbc /* host, 3 */, const int* types, long long max_edges,
                                               const sevenn_halo_hooks* halo, void* workspace, long long workspace_bytes,
                                               void* stream, const sevenn_outputs* out) {
  int rc = check_args(m, workspace, workspace_bytes);
  if (rc || out == nullptr || out->energy == nullptr || out->atomic_energy == nullptr || out->status == nullptr ||
      cell == nullptr || pbc == nullptr || n_local < 0 || n_nodes < n_local || max_edges < 0 || max_edges > 0x7fffffffLL)
    return SEVENN_ERR_BAD_ARG;
  if (sevenn_model_workspace_bytes(m, n_local, n_nodes, max_edges, 1) > workspace_bytes) return SEVENN_ERR_WORKSPACE;
  cudaStream_t st = (cudaStream_t)stream;
  Arena A{(char*)workspace, (size_t)workspace_bytes, false};
  cudaMemsetAsync(out->status, 0, 2 * sizeof(int), st);
  NlOut o;
  rc = build_neighbors(m, A, n_nodes, n_local, pos, pos_host, cell, pbc, max_edges, out->status, o, st);
  if (rc) return rc;
  GraphView g{n_local, n_nodes, max_edges, types, o.vec, o.rowptr, o.col, o.cptr, o.cedge, o.rowptr + n_local};
  rc = evaluate_graph(m, A, g, halo, st, out->energy, out->atomic_energy, out->forces, out->virial6, out->virial_atoms);
  if (rc) return rc;
  return A.fail ? SEVENN_ERR_WORKSPACE : 0;
}

// Number of edges of the structure (synchronous: the ONE call a host makes to size its edge capacity; later steps
// reuse the capacity with slack and only look at the status word)
extern "C" int sevenn_model_count_edges(sevenn_model* m, long long n_rows, long long n_nodes, const double* pos,
                                        const double* pos_host, const double* cell, const int* pbc, void* workspace,
                                        long long workspace_bytes, void* stream, long long* n_edges) {
  int rc = check_args(m, workspace, workspace_bytes);
  if (rc || n_edges == nullptr || cell == nullptr || pbc == nullptr) return SEVENN_ERR_BAD_ARG;
  *n_edges = 0;
  if (n_nodes <= 0 || n_rows <= 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  void* vs = stream;
  Arena A{(char*)workspace, (size_t)workspace_bytes, false};
  double dpar[25]; int ipar[9];
  const long long n = n_nodes;
  rc = nl_params(pos_host, n, cell, pbc, (double)m->cutoff, 8 * n > 64 ? 8 * n : 64, dpar, ipar);
  if (rc) return rc;
  const long long nbins = (long long)ipar[3] * ipar[4] * ipar[5];
  double* pw = A.f64(3 * n);
  const size_t z0 = A.mark();
  int* wrap = A.i32(3 * n); int* atom_bin = A.i32(n); int* bin_count = A.i32(nbins + 1); int* bin_start = A.i32(nbins + 1);
  int* bin_fill = A.i32(nbins); int* bin_atoms = A.i32(n);
  int* tmp = A.i32(sevenn_scan_tmp_ints((int)((nbins > n ? nbins : n) + 1)) + 1);
  int* count = A.i32(n_rows + 1); int* rowptr = A.i32(n_rows + 1);
  const size_t z1 = A.mark();
  if (A.fail) return SEVENN_ERR_WORKSPACE;
  cudaMemsetAsync(A.base + align_up(z0, 256), 0, z1 - align_up(z0, 256), st);
  if ((rc = sevenn_nl_bin(pos, (int)n, dpar, ipar, pw, wrap, atom_bin, bin_count, vs))) return rc;
  if ((rc = sevenn_scan_i32(bin_count, bin_start, (int)nbins, tmp, vs))) return rc;
  if ((rc = sevenn_nl_scatter(atom_bin, (int)n, bin_start, bin_fill, bin_atoms, vs))) return rc;
  if ((rc = sevenn_nl_count(pw, wrap, atom_bin, bin_start, bin_atoms, (int)n_rows, dpar, ipar, count, vs))) return rc;
  if ((rc = sevenn_scan_i32(count, rowptr, (int)n_rows, tmp, vs))) return rc;
  int e = 0;
  if (cudaMemcpyAsync(&e, rowptr + n_rows, sizeof(int), cudaMemcpyDeviceToHost, st) != cudaSuccess) return (int)cudaGetLastError();
  if (cudaStreamSynchronize(st) != cudaSuccess) return (int)cudaGetLastError();
  *n_edges = e;
  return 0;
}

// ---- deployed-model entry: the dict the LAMMPS pair style hands to the TorchScript module
// (sevenn/pair_e3gnn/pair_e3gnn.cpp:218-225: x, pos, edge_index, cell_lattice_vectors, pbc_shift) -> E / F / virial /
// atomic energies.  Replaces EdgePreprocess (sevenn/nn/edge_embedding.py:13-78) + the model + ForceStressOutput by
// autograd over positions and strain (sevenn/nn/force_output.py:55-136); the edge list may come in any order.
namespace {
__global__ void i64_to_i32_kernel(const long long* __restrict__ in, long long n, int* __restrict__ out) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    out[i] = (int)in[i];
}
// edges regrouped by centre (perm = edge ids grouped by centre, ascending): neighbour index and edge vector
__global__ void edge_gather_kernel(const int* __restrict__ perm, long long n_edges, const long long* __restrict__ ei,
                                   const float* __restrict__ pos, const float* __restrict__ cell,
                                   const float* __restrict__ shift, int* __restrict__ col, float* __restrict__ vec) {
  for (long long k = blockIdx.x * (long long)blockDim.x + threadIdx.x; k < n_edges; k += (long long)gridDim.x * blockDim.x) {
    const long long e = perm[k];
    const long long i = ei[e], j = ei[n_edges + e];
    col[k] = (int)j;
    const float s0 = shift[3 * e], s1 = shift[3 * e + 1], s2 = shift[3 * e + 2];
#pragma unroll
    for (int c = 0; c < 3; ++c)
      vec[3 * k + c] = pos[3 * j + c] - pos[3 * i + c] + (s0 * cell[c] + s1 * cell[3 + c] + s2 * cell[6 + c]);
  }
}
}  // namespace

extern "C" long long sevenn_model_edges_workspace_bytes(sevenn_model* m, long long n_atoms, long long n_edges) {
  const long long base = sevenn_model_workspace_bytes(m, n_atoms, n_atoms, n_edges, 0);
  if (base < 0) return -1;
  const long long ints = 6 * (n_atoms + 64) + 2 * sevenn_scan_tmp_ints((int)(n_atoms + 1)) + 5 * (n_edges + 64);
  return base + 4 * ints + 12 * (n_edges + 64) + 8 * (n_edges + 64) + 64 * 256;
}

extern "C" int sevenn_model_evaluate_edges(sevenn_model* m, long long n_atoms, const long long* x, const float* pos,
                                           const long long* edge_index, long long n_edges, const float* cell,
                                           const float* pbc_shift, void* workspace, long long workspace_bytes,
                                           void* stream, const sevenn_outputs* out) {
  int rc = check_args(m, workspace, workspace_bytes);
  if (rc || out == nullptr || out->energy == nullptr || out->atomic_energy == nullptr || n_atoms < 0 || n_edges < 0 ||
      n_edges > 0x7fffffffLL)
    return SEVENN_ERR_BAD_ARG;
  if (sevenn_model_edges_workspace_bytes(m, n_atoms, n_edges) > workspace_bytes) return SEVENN_ERR_WORKSPACE;
  if (n_atoms == 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  Arena A{(char*)workspace, (size_t)workspace_bytes, false};
  const long long n = n_atoms, E = n_edges, n_tmp = sevenn_scan_tmp_ints((int)(n + 1)) + 1;
  int* types = A.i32(n);
  int* center = A.i32(E);
  const size_t z0 = A.mark();
  int* cnt = A.i32(n + 1); int* fill = A.i32(n); int* cnt2 = A.i32(n + 1); int* fill2 = A.i32(n);
  const size_t z1 = A.mark();
  int* rowptr = A.i32(n + 1); int* cptr = A.i32(n + 1); int* tmp = A.i32(n_tmp);
  int* perm = A.i32(E); int* col = A.i32(E); int* cedge = A.i32(E);
  float* vec = A.f32(3 * E);
  unsigned long long* keys = (unsigned long long*)A.raw((size_t)(E > 0 ? E : 1) * 8);
  cudaMemsetAsync(A.base + align_up(z0, 256), 0, z1 - align_up(z0, 256), st);
  i64_to_i32_kernel<<<sevenn_grid_for(n), 256, 0, st>>>(x, n, types);
  if (E > 0) i64_to_i32_kernel<<<sevenn_grid_for(E), 256, 0, st>>>(edge_index, E, center);
  rc = sevenn_csc_build(center, E, (int)n, cnt, fill, rowptr, keys, perm, tmp, stream);      // CSR by centre
  if (rc) return rc;
  if (E > 0) edge_gather_kernel<<<sevenn_grid_for(E), 256, 0, st>>>(perm, E, edge_index, pos, cell, pbc_shift, col, vec);
  rc = sevenn_csc_build(col, E, (int)n, cnt2, fill2, cptr, keys, cedge, tmp, stream);         // CSC by neighbour
  if (rc) return rc;
  m->n_launch += 4;
  GraphView g{n, n, E, types, vec, rowptr, col, cptr, cedge, rowptr + n};
  rc = evaluate_graph(m, A, g, nullptr, st, out->energy, out->atomic_energy, out->forces, out->virial6, out->virial_atoms);
  if (rc) return rc;
  return A.fail ? SEVENN_ERR_WORKSPACE : (int)cudaGetLastError();
}

// Cell-list periodic radius graph (fp64), CSR by centre atom, rows sorted by (j, Sx, Sy, Sz).
//
// Replaces sevenn/train/dataload.py:32-130 (ase primitive_neighbor_list / matscipy neighbour_list,
// both external CPU code) and the CPU double loop of pair_e3gnn.cpp:157-195.
// Semantics kept: every (i, j, S) with |pos_j - pos_i + S.cell| < rc (strict), both directions,
// only the zero-vector self pair excluded; D is formed in fp64 and cast to fp32
// (atom_graph_data.py:69-75).  Output order is the reference's (outer loop over i).
//
// HBM-bound integer/fp64 work: thread per atom, bins >= rc wide so the stencil is +-R bins.
#include "sevenn_common.cuh"
#include "../../include/sevenn_b200.h"

struct NlParams {
  double cell[9];      // rows = lattice vectors
  double inv[9];       // frac_d = sum_k pos_k * inv[k*3+d]
  double lo[3];        // bin origin in fractional coords (0 for periodic dims)
  double inv_span[3];  // 1/extent in fractional coords (1 for periodic dims)
  double rc2;
  int pbc[3];
  int nb[3];
  int R[3];
};

HD int floordiv(int a, int b) { int q = a / b; return (a % b != 0 && ((a < 0) != (b < 0))) ? q - 1 : q; }

GKERNEL(nl_assign_kernel)(const double* __restrict__ pos, int n, NlParams P, double* __restrict__ pw,
                          int* __restrict__ wrap, int* __restrict__ atom_bin, int* bin_count) {
  GRID_STRIDE_LOOP(i, n) {
    const double x = pos[3 * i], y = pos[3 * i + 1], z = pos[3 * i + 2];
    double f[3];
    int w[3], b[3];
    for (int d = 0; d < 3; ++d) f[d] = x * P.inv[d] + y * P.inv[3 + d] + z * P.inv[6 + d];
    for (int d = 0; d < 3; ++d) {
      w[d] = 0;
      if (P.pbc[d]) {
        double fl = floor(f[d]);
        f[d] -= fl;
        if (f[d] >= 1.0) { f[d] -= 1.0; fl += 1.0; }
        w[d] = (int)fl;
      }
      int bb = (int)floor((f[d] - P.lo[d]) * P.inv_span[d] * P.nb[d]);
      b[d] = bb < 0 ? 0 : (bb >= P.nb[d] ? P.nb[d] - 1 : bb);
      wrap[3 * i + d] = w[d];
    }
    for (int k = 0; k < 3; ++k)
      pw[3 * i + k] = pos[3 * i + k] - (w[0] * P.cell[k] + w[1] * P.cell[3 + k] + w[2] * P.cell[6 + k]);
    const int bin = (b[0] * P.nb[1] + b[1]) * P.nb[2] + b[2];
    atom_bin[i] = bin;
    ATOMIC_ADD_I32(bin_count + bin, 1);
  }
}

GKERNEL(nl_scatter_kernel)(const int* __restrict__ atom_bin, int n, const int* __restrict__ bin_start,
                           int* bin_fill, int* __restrict__ bin_atoms) {
  GRID_STRIDE_LOOP(i, n) {
    const int bin = atom_bin[i];
    const int slot = bin_start[bin] + ATOMIC_ADD_I32(bin_fill + bin, 1);
    bin_atoms[slot] = (int)i;
  }
}

// FILL = false: count neighbours; FILL = true: write sort keys  j<<24 | (S+128) bytes.
template <bool FILL>
GKERNEL(nl_walk_kernel)(const double* __restrict__ pw, const int* __restrict__ wrap,
                        const int* __restrict__ atom_bin, const int* __restrict__ bin_start,
                        const int* __restrict__ bin_atoms, int n, NlParams P, int* __restrict__ count,
                        const int* __restrict__ rowptr, unsigned long long* __restrict__ keys, long long cap) {
  GRID_STRIDE_LOOP(i, n) {
    const double xi = pw[3 * i], yi = pw[3 * i + 1], zi = pw[3 * i + 2];
    const int bin = atom_bin[i];
    const int b2 = bin % P.nb[2], b1 = (bin / P.nb[2]) % P.nb[1], b0 = bin / (P.nb[2] * P.nb[1]);
    const int wi0 = wrap[3 * i], wi1 = wrap[3 * i + 1], wi2 = wrap[3 * i + 2];
    int cnt = 0;
    long long base = FILL ? rowptr[i] : 0;
    for (int d0 = -P.R[0]; d0 <= P.R[0]; ++d0) {
      int c0 = b0 + d0, s0 = 0;
      if (P.pbc[0]) { s0 = floordiv(c0, P.nb[0]); c0 -= s0 * P.nb[0]; } else if (c0 < 0 || c0 >= P.nb[0]) continue;
      for (int d1 = -P.R[1]; d1 <= P.R[1]; ++d1) {
        int c1 = b1 + d1, s1 = 0;
        if (P.pbc[1]) { s1 = floordiv(c1, P.nb[1]); c1 -= s1 * P.nb[1]; } else if (c1 < 0 || c1 >= P.nb[1]) continue;
        for (int d2 = -P.R[2]; d2 <= P.R[2]; ++d2) {
          int c2 = b2 + d2, s2 = 0;
          if (P.pbc[2]) { s2 = floordiv(c2, P.nb[2]); c2 -= s2 * P.nb[2]; } else if (c2 < 0 || c2 >= P.nb[2]) continue;
          const double ox = s0 * P.cell[0] + s1 * P.cell[3] + s2 * P.cell[6] - xi;
          const double oy = s0 * P.cell[1] + s1 * P.cell[4] + s2 * P.cell[7] - yi;
          const double oz = s0 * P.cell[2] + s1 * P.cell[5] + s2 * P.cell[8] - zi;
          const int nbin = (c0 * P.nb[1] + c1) * P.nb[2] + c2;
          const int a1 = bin_start[nbin + 1];
          for (int a = bin_start[nbin]; a < a1; ++a) {
            const int j = bin_atoms[a];
            const double dx = pw[3 * j] + ox, dy = pw[3 * j + 1] + oy, dz = pw[3 * j + 2] + oz;
            const double d2v = dx * dx + dy * dy + dz * dz;
            if (d2v < P.rc2 && !(j == (int)i && s0 == 0 && s1 == 0 && s2 == 0)) {
              if (FILL && base + cnt < cap) {
                const int S0 = s0 - wrap[3 * j] + wi0, S1 = s1 - wrap[3 * j + 1] + wi1,
                          S2 = s2 - wrap[3 * j + 2] + wi2;
                keys[base + cnt] = ((unsigned long long)j << 24) | ((unsigned long long)(S0 + 128) << 16) |
                                   ((unsigned long long)(S1 + 128) << 8) | (unsigned long long)(S2 + 128);
              }
              ++cnt;
            }
          }
        }
      }
    }
    if (!FILL) count[i] = cnt;
  }
}

#ifndef SEVENN_EMU
// CUDA flavour of the walk: one warp per centre atom, lanes over the atoms of each stencil bin, ballot compaction
// (the order inside a row is irrelevant: rows are sorted by key afterwards).  4x fewer dependent global loads per
// thread than the thread-per-atom loop above, which is what a 4,096-atom cell needs to fill 148 SMs.
template <bool FILL>
__global__ void __launch_bounds__(128) nl_walk_warp_kernel(const double* __restrict__ pw, const int* __restrict__ wrap,
                                                           const int* __restrict__ atom_bin,
                                                           const int* __restrict__ bin_start,
                                                           const int* __restrict__ bin_atoms, int n, NlParams P,
                                                           int* __restrict__ count, const int* __restrict__ rowptr,
                                                           unsigned long long* __restrict__ keys, long long cap) {
  const int lane = threadIdx.x & 31;
  for (long long i = blockIdx.x * 4ll + (threadIdx.x >> 5); i < n; i += gridDim.x * 4ll) {
    const double xi = pw[3 * i], yi = pw[3 * i + 1], zi = pw[3 * i + 2];
    const int bin = atom_bin[i];
    const int b2 = bin % P.nb[2], b1 = (bin / P.nb[2]) % P.nb[1], b0 = bin / (P.nb[2] * P.nb[1]);
    const int wi0 = wrap[3 * i], wi1 = wrap[3 * i + 1], wi2 = wrap[3 * i + 2];
    int cnt = 0;
    const long long base = FILL ? rowptr[i] : 0;
    for (int d0 = -P.R[0]; d0 <= P.R[0]; ++d0) {
      int c0 = b0 + d0, s0 = 0;
      if (P.pbc[0]) { s0 = floordiv(c0, P.nb[0]); c0 -= s0 * P.nb[0]; } else if (c0 < 0 || c0 >= P.nb[0]) continue;
      for (int d1 = -P.R[1]; d1 <= P.R[1]; ++d1) {
        int c1 = b1 + d1, s1 = 0;
        if (P.pbc[1]) { s1 = floordiv(c1, P.nb[1]); c1 -= s1 * P.nb[1]; } else if (c1 < 0 || c1 >= P.nb[1]) continue;
        for (int d2 = -P.R[2]; d2 <= P.R[2]; ++d2) {
          int c2 = b2 + d2, s2 = 0;
          if (P.pbc[2]) { s2 = floordiv(c2, P.nb[2]); c2 -= s2 * P.nb[2]; } else if (c2 < 0 || c2 >= P.nb[2]) continue;
          const double ox = s0 * P.cell[0] + s1 * P.cell[3] + s2 * P.cell[6] - xi;
          const double oy = s0 * P.cell[1] + s1 * P.cell[4] + s2 * P.cell[7] - yi;
          const double oz = s0 * P.cell[2] + s1 * P.cell[5] + s2 * P.cell[8] - zi;
          const int nbin = (c0 * P.nb[1] + c1) * P.nb[2] + c2;
          const int a0 = bin_start[nbin], a1 = bin_start[nbin + 1];
          const bool home = (s0 == 0 && s1 == 0 && s2 == 0);
          for (int ab = a0; ab < a1; ab += 32) {
            const int a = ab + lane;
            bool ok = false;
            int j = 0;
            if (a < a1) {
              j = bin_atoms[a];
              const double dx = pw[3 * j] + ox, dy = pw[3 * j + 1] + oy, dz = pw[3 * j + 2] + oz;
              ok = (dx * dx + dy * dy + dz * dz) < P.rc2 && !(home && j == (int)i);
            }
            const unsigned mask = __ballot_sync(0xffffffffu, ok);
            if (FILL && ok && base + cnt + __popc(mask & ((1u << lane) - 1u)) < cap) {   // cap: edge-array capacity
              const int S0 = s0 - wrap[3 * j] + wi0, S1 = s1 - wrap[3 * j + 1] + wi1, S2 = s2 - wrap[3 * j + 2] + wi2;
              keys[base + cnt + __popc(mask & ((1u << lane) - 1u))] =
                  ((unsigned long long)j << 24) | ((unsigned long long)(S0 + 128) << 16) |
                  ((unsigned long long)(S1 + 128) << 8) | (unsigned long long)(S2 + 128);
            }
            cnt += __popc(mask);
          }
        }
      }
    }
    if (!FILL && lane == 0) count[i] = cnt;
  }
}
#define LAUNCH_WALK(FILL, n, stream) nl_walk_warp_kernel<FILL><<<(unsigned)(((n) + 3) / 4 < 65535 * 16 ? ((n) + 3) / 4 : 65535 * 16), 128, 0, (cudaStream_t)(stream)>>>
#else
#define LAUNCH_WALK(FILL, n, stream) nl_walk_kernel<FILL>
#endif

// in-place sort of each CSR segment (segments are ~10-100 long)
#ifdef SEVENN_EMU
GKERNEL(segment_sort_u64_kernel)(unsigned long long* __restrict__ keys, const int* __restrict__ ptr, int n) {
  GRID_STRIDE_LOOP(i, n) {
    const int a = ptr[i], b = ptr[i + 1];
    for (int p = a + 1; p < b; ++p) {
      const unsigned long long k = keys[p];
      int q = p - 1;
      while (q >= a && keys[q] > k) { keys[q + 1] = keys[q]; --q; }
      keys[q + 1] = k;
    }
  }
}
#define LAUNCH_SEGSORT(n, stream) segment_sort_u64_kernel
#else
// one warp per segment: bitonic network in shared memory for segments up to 256 keys (coalesced load / store,
// 32 compare-exchanges per step); longer segments fall back to a lane-0 insertion sort (keys are unique, so the
// result does not depend on the algorithm)
#define SEGSORT_MAX 256
__global__ void __launch_bounds__(128) segment_sort_u64_kernel(unsigned long long* __restrict__ keys,
                                                               const int* __restrict__ ptr, int n) {
  __shared__ unsigned long long buf_all[4][SEGSORT_MAX];
  unsigned long long* buf = buf_all[threadIdx.x >> 5];
  const int lane = threadIdx.x & 31;
  for (long long seg = blockIdx.x * 4ll + (threadIdx.x >> 5); seg < n; seg += gridDim.x * 4ll) {
    const int a = ptr[seg], len = ptr[seg + 1] - a;
    if (len <= 1) continue;
    if (len > SEGSORT_MAX) {
      if (lane == 0) {
        for (int p = a + 1; p < a + len; ++p) {
          const unsigned long long k = keys[p];
          int q = p - 1;
          while (q >= a && keys[q] > k) { keys[q + 1] = keys[q]; --q; }
          keys[q + 1] = k;
        }
      }
      continue;
    }
    int m = 2;
    while (m < len) m <<= 1;
    for (int t = lane; t < m; t += 32) buf[t] = t < len ? keys[a + t] : ~0ull;
    __syncwarp();
    for (int k = 2; k <= m; k <<= 1)
      for (int j = k >> 1; j > 0; j >>= 1) {
        for (int t = lane; t < (m >> 1); t += 32) {
          const int lo = ((t & ~(j - 1)) << 1) | (t & (j - 1)), hi = lo | j;   // pair (lo, lo + j)
          const bool up = (lo & k) == 0;
          const unsigned long long x = buf[lo], y = buf[hi];
          if ((x > y) == up) { buf[lo] = y; buf[hi] = x; }
        }
        __syncwarp();
      }
    for (int t = lane; t < len; t += 32) keys[a + t] = buf[t];
    __syncwarp();
  }
}
#define LAUNCH_SEGSORT(n, stream) segment_sort_u64_kernel<<<(unsigned)(((n) + 3) / 4 < 65535 * 16 ? ((n) + 3) / 4 : 65535 * 16), 128, 0, (cudaStream_t)(stream)>>>
#endif

// 32 consecutive threads share an atom and stride over its edges: coalesced key reads and edge-array writes (one
// thread per atom ran 16 CTAs for 4,096 atoms and wrote with a 50-edge stride)
GKERNEL(nl_decode_kernel)(const unsigned long long* __restrict__ keys, const int* __restrict__ rowptr, int n,
                          const double* __restrict__ pos, NlParams P, int* __restrict__ center,
                          int* __restrict__ col, int* __restrict__ shift, float* __restrict__ vec) {
  GRID_STRIDE_LOOP(t, (long long)n * 32) {
    const long long i = t >> 5;
    const int e1 = rowptr[i + 1];
    for (int e = rowptr[i] + (int)(t & 31); e < e1; e += 32) {
      const unsigned long long k = keys[e];
      const int j = (int)(k >> 24);
      const int S0 = (int)((k >> 16) & 255) - 128, S1 = (int)((k >> 8) & 255) - 128, S2 = (int)(k & 255) - 128;
      center[e] = (int)i;
      col[e] = j;
      shift[3 * e] = S0; shift[3 * e + 1] = S1; shift[3 * e + 2] = S2;
      for (int c = 0; c < 3; ++c) {
        const double off = S0 * P.cell[c] + S1 * P.cell[3 + c] + S2 * P.cell[6 + c];
        vec[3 * e + c] = (float)((pos[3 * j + c] + off) - pos[3 * i + c]);
      }
    }
  }
}

// ---- CSC (edges grouped by neighbour j), segments sorted by edge id -> deterministic sums
GKERNEL(hist_kernel)(const int* __restrict__ idx, long long n, const int* __restrict__ n_dev, int* cnt) {
  if (n_dev != nullptr && *n_dev < n) n = *n_dev;
  GRID_STRIDE_LOOP(e, n) { ATOMIC_ADD_I32(cnt + idx[e], 1); }
}
GKERNEL(csc_scatter_kernel)(const int* __restrict__ col, long long n, const int* __restrict__ n_dev,
                            const int* __restrict__ cptr, int* fill, unsigned long long* __restrict__ keys) {
  if (n_dev != nullptr && *n_dev < n) n = *n_dev;
  GRID_STRIDE_LOOP(e, n) {
    const int j = col[e];
    keys[cptr[j] + ATOMIC_ADD_I32(fill + j, 1)] = (unsigned long long)e;
  }
}
GKERNEL(u64_to_i32_kernel)(const unsigned long long* __restrict__ k, long long n, const int* __restrict__ n_dev,
                           int* __restrict__ out) {
  if (n_dev != nullptr && *n_dev < n) n = *n_dev;
  GRID_STRIDE_LOOP(e, n) { out[e] = (int)k[e]; }
}
// rowptr (n + 1 entries) clamped to the edge-array capacity; *status |= 1 when the list did not fit
GKERNEL(clamp_rowptr_kernel)(int* __restrict__ rowptr, int n, int cap, int* __restrict__ status) {
  GRID_STRIDE_LOOP(i, (long long)n + 1) {
    if (rowptr[i] > cap) {
      rowptr[i] = cap;
      if (i == n) *status = 1;
    }
  }
}

// ---- exclusive scan of int32 (out has n+1 entries, out[n] = total) ------------------------
#ifndef SEVENN_EMU
#define SCAN_TPB 256
#define SCAN_EPB 1024
__global__ void scan_block_kernel(const int* __restrict__ in, int* __restrict__ out, int n, int* __restrict__ sums) {
  __shared__ int warp_tot[SCAN_TPB / 32];
  const int base = blockIdx.x * SCAN_EPB + threadIdx.x * 4;
  int v[4], s = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) { v[k] = (base + k < n) ? in[base + k] : 0; s += v[k]; }
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  int inc = s;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
  if (lane == 31) warp_tot[wid] = inc;
  __syncthreads();
  if (wid == 0) {
    int t = lane < SCAN_TPB / 32 ? warp_tot[lane] : 0;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { int u = __shfl_up_sync(0xffffffffu, t, o); if (lane >= o) t += u; }
    if (lane < SCAN_TPB / 32) warp_tot[lane] = t;   // inclusive over warps
  }
  __syncthreads();
  int excl = inc - s + (wid > 0 ? warp_tot[wid - 1] : 0);
#pragma unroll
  for (int k = 0; k < 4; ++k) { if (base + k < n) out[base + k] = excl; excl += v[k]; }
  if (threadIdx.x == SCAN_TPB - 1) sums[blockIdx.x] = warp_tot[SCAN_TPB / 32 - 1];
}
__global__ void scan_sums_kernel(int* sums, int nb, int* total_out) {   // one CTA, serial over chunks
  __shared__ int sh[1024];
  __shared__ int carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int c = 0; c < nb; c += 1024) {
    const int i = c + threadIdx.x;
    const int v = i < nb ? sums[i] : 0;
    sh[threadIdx.x] = v;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
      int t = threadIdx.x >= o ? sh[threadIdx.x - o] : 0;
      __syncthreads();
      sh[threadIdx.x] += t;
      __syncthreads();
    }
    if (i < nb) sums[i] = carry + sh[threadIdx.x] - v;
    __syncthreads();
    if (threadIdx.x == 0) carry += sh[1023];
    __syncthreads();
  }
  if (threadIdx.x == 0) *total_out = carry;
}
__global__ void scan_add_kernel(int* __restrict__ out, int n, const int* __restrict__ sums) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] += sums[i / SCAN_EPB];
}
#endif

extern "C" int sevenn_scan_i32(const int* in, int* out, int n, int* tmp, void* stream) {
#ifdef SEVENN_EMU
  int s = 0;
  for (int i = 0; i < n; ++i) { int v = in[i]; out[i] = s; s += v; }
  out[n] = s;
  return 0;
#else
  cudaStream_t st = (cudaStream_t)stream;
  if (n <= 0) { cudaMemsetAsync(out, 0, sizeof(int), st); return (int)cudaGetLastError(); }
  const int nb = (n + SCAN_EPB - 1) / SCAN_EPB;
  scan_block_kernel<<<nb, SCAN_TPB, 0, st>>>(in, out, n, tmp);
  scan_sums_kernel<<<1, 1024, 0, st>>>(tmp, nb, out + n);
  scan_add_kernel<<<(n + 255) / 256, 256, 0, st>>>(out, n, tmp);
  return (int)cudaGetLastError();
#endif
}
extern "C" long long sevenn_scan_tmp_ints(int n) { return (n + 1023) / 1024 + 1; }

static NlParams make_params(const double* d, const int* ip) {
  NlParams P;
  for (int k = 0; k < 9; ++k) { P.cell[k] = d[k]; P.inv[k] = d[9 + k]; }
  for (int k = 0; k < 3; ++k) { P.lo[k] = d[18 + k]; P.inv_span[k] = d[21 + k]; P.pbc[k] = ip[k]; P.nb[k] = ip[3 + k]; P.R[k] = ip[6 + k]; }
  P.rc2 = d[24];
  return P;
}

extern "C" int sevenn_nl_bin(const double* pos, int n, const double* dpar, const int* ipar, double* pw, int* wrap,
                             int* atom_bin, int* bin_count, void* stream) {
  if (n <= 0) return 0;
  LAUNCH1D(nl_assign_kernel, n, stream)(pos, n, make_params(dpar, ipar), pw, wrap, atom_bin, bin_count);
  return LAUNCH_STATUS();
}
extern "C" int sevenn_nl_scatter(const int* atom_bin, int n, const int* bin_start, int* bin_fill, int* bin_atoms,
                                 void* stream) {
  if (n <= 0) return 0;
  LAUNCH1D(nl_scatter_kernel, n, stream)(atom_bin, n, bin_start, bin_fill, bin_atoms);
  return LAUNCH_STATUS();
}
extern "C" int sevenn_nl_count(const double* pw, const int* wrap, const int* atom_bin, const int* bin_start,
                               const int* bin_atoms, int n, const double* dpar, const int* ipar, int* count,
                               void* stream) {
  if (n <= 0) return 0;
  LAUNCH_WALK(false, n, stream)(pw, wrap, atom_bin, bin_start, bin_atoms, n, make_params(dpar, ipar),
                                             count, nullptr, nullptr, 0);
  return LAUNCH_STATUS();
}
static int nl_fill_impl(const double* pw, const int* wrap, const int* atom_bin, const int* bin_start,
                        const int* bin_atoms, int n, const double* dpar, const int* ipar, const int* rowptr,
                        unsigned long long* keys, const double* pos, int* center, int* col, int* shift, float* vec,
                        long long cap, void* stream) {
  if (n <= 0) return 0;
  NlParams P = make_params(dpar, ipar);
  LAUNCH_WALK(true, n, stream)(pw, wrap, atom_bin, bin_start, bin_atoms, n, P, nullptr, rowptr, keys, cap);
  LAUNCH_SEGSORT(n, stream)(keys, rowptr, n);
  LAUNCH1D(nl_decode_kernel, (long long)n * 32, stream)(keys, rowptr, n, pos, P, center, col, shift, vec);
  return LAUNCH_STATUS();
}
extern "C" int sevenn_nl_fill(const double* pw, const int* wrap, const int* atom_bin, const int* bin_start,
                              const int* bin_atoms, int n, const double* dpar, const int* ipar, const int* rowptr,
                              unsigned long long* keys, const double* pos, int* center, int* col, int* shift,
                              float* vec, void* stream) {
  return nl_fill_impl(pw, wrap, atom_bin, bin_start, bin_atoms, n, dpar, ipar, rowptr, keys, pos, center, col, shift,
                      vec, 0x7fffffffffffffffLL, stream);
}
// Capacity-bounded fill for hosts that never read the edge count: rowptr (from the scan of the counts) is first
// clamped to max_edges (status word |= 1 on overflow: results are then memory-safe garbage and the caller re-runs
// with a larger workspace), the rows of atoms 0 .. n_rows-1 are filled (n_rows <= atoms: rows for owned atoms only).
extern "C" int sevenn_nl_fill_capped(const double* pw, const int* wrap, const int* atom_bin, const int* bin_start,
                                     const int* bin_atoms, int n_rows, const double* dpar, const int* ipar,
                                     int* rowptr, long long max_edges, int* status, unsigned long long* keys,
                                     const double* pos, int* center, int* col, int* shift, float* vec, void* stream) {
  if (n_rows <= 0) return 0;
  if (max_edges <= 0 || max_edges > 0x7fffffffLL || status == nullptr) return SEVENN_ERR_BAD_ARG;
  LAUNCH1D(clamp_rowptr_kernel, n_rows + 1, stream)(rowptr, n_rows, (int)max_edges, status);
  return nl_fill_impl(pw, wrap, atom_bin, bin_start, bin_atoms, n_rows, dpar, ipar, rowptr, keys, pos, center, col,
                      shift, vec, max_edges, stream);
}
// cnt: n_nodes+1 ints zeroed by the caller; fill: n_nodes ints zeroed by the caller
static int csc_build_impl(const int* col, long long n_edges, const int* n_dev, int n_nodes, int* cnt, int* fill,
                          int* cptr, unsigned long long* keys, int* cedge, int* scan_tmp, void* stream) {
  if (n_edges > 0) LAUNCH1D(hist_kernel, n_edges, stream)(col, n_edges, n_dev, cnt);
  int rc = sevenn_scan_i32(cnt, cptr, n_nodes, scan_tmp, stream);
  if (rc) return rc;
  if (n_edges > 0) {
    LAUNCH1D(csc_scatter_kernel, n_edges, stream)(col, n_edges, n_dev, cptr, fill, keys);
    LAUNCH_SEGSORT(n_nodes, stream)(keys, cptr, n_nodes);
    LAUNCH1D(u64_to_i32_kernel, n_edges, stream)(keys, n_edges, n_dev, cedge);
  }
  return LAUNCH_STATUS();
}
extern "C" int sevenn_csc_build(const int* col, long long n_edges, int n_nodes, int* cnt, int* fill, int* cptr,
                                unsigned long long* keys, int* cedge, int* scan_tmp, void* stream) {
  return csc_build_impl(col, n_edges, nullptr, n_nodes, cnt, fill, cptr, keys, cedge, scan_tmp, stream);
}
// edge count read on the device (n_edges_dev), launches and buffers bounded by max_edges
extern "C" int sevenn_csc_build_dev(const int* col, long long max_edges, const int* n_edges_dev, int n_nodes, int* cnt,
                                    int* fill, int* cptr, unsigned long long* keys, int* cedge, int* scan_tmp,
                                    void* stream) {
  if (n_edges_dev == nullptr) return SEVENN_ERR_BAD_ARG;
  return csc_build_impl(col, max_edges, n_edges_dev, n_nodes, cnt, fill, cptr, keys, cedge, scan_tmp, stream);
}

/* sevenn_b200.h -- C ABI of the B200 (sm_100a) SevenNet energy/force/stress kernels.
 *
 * Drop-in boundary for the hot path of hexagonrose/SevenNet_exp (sevenn 0.11): everything the
 * reference obtains from e3nn/ATen/torch.autograd inside AtomGraphSequential.forward
 * (sevenn/nn/sequential.py:178, call stack in SURVEY.md section 3.1) is reachable through the plain-C
 * launchers below.  Conventions:
 *   - all pointers are DEVICE pointers unless the parameter is documented "host";
 *   - the caller owns every buffer (no hidden allocation), passes the cudaStream_t as void*;
 *   - launchers are asynchronous and return 0 or a cudaError_t / SEVENN_ERR_* code; they never exit();
 *   - node tensors are fp32 rows in the "ir_mul" layout described in sevennet_exp_b200/irreps.py;
 *   - indices are int32 on the device (E < 2^31); the torch boundary converts from int64.
 * Each entry cites the reference interface it replaces (paths relative to /root/reference).
 */
#ifndef SEVENN_B200_H
#define SEVENN_B200_H

#ifdef __cplusplus
extern "C" {
#endif

#define SEVENN_ERR_BAD_ARG 10001
#define SEVENN_MAX_GATE_PIECES 16

int sevenn_abi_version(void);

/* ---- neighbor list: replaces sevenn/train/dataload.py:32-130 (_graph_build_ase/_matscipy,
 * unlabeled_atoms_to_graph) and the CPU edge loop of sevenn/pair_e3gnn/pair_e3gnn.cpp:157-195.
 * dpar (host, 25 doubles): cell[9] rows=lattice vectors, inv[9] (frac = pos.inv), bin origin lo[3],
 * 1/extent inv_span[3] (fractional), rc^2.   ipar (host, 9 ints): pbc[3], nbins[3], stencil R[3]. */
int sevenn_scan_i32(const int* in, int* out /* n+1 */, int n, int* tmp, void* stream);
long long sevenn_scan_tmp_ints(int n);
int sevenn_nl_bin(const double* pos, int n, const double* dpar, const int* ipar, double* pos_wrapped,
                  int* wrap, int* atom_bin, int* bin_count /* zeroed, nbins+1 */, void* stream);
int sevenn_nl_scatter(const int* atom_bin, int n, const int* bin_start, int* bin_fill /* zeroed */,
                      int* bin_atoms, void* stream);
int sevenn_nl_count(const double* pos_wrapped, const int* wrap, const int* atom_bin, const int* bin_start,
                    const int* bin_atoms, int n, const double* dpar, const int* ipar, int* count, void* stream);
int sevenn_nl_fill(const double* pos_wrapped, const int* wrap, const int* atom_bin, const int* bin_start,
                   const int* bin_atoms, int n, const double* dpar, const int* ipar, const int* rowptr,
                   unsigned long long* keys /* E */, const double* pos, int* center /* E: edge_index[0] */,
                   int* col /* E: edge_index[1] */, int* shift /* E,3 */, float* edge_vec /* E,3 */, void* stream);
/* edges grouped by neighbour (edge_index[1]); segments sorted by edge id */
int sevenn_csc_build(const int* col, long long n_edges, int n_nodes, int* cnt /* zeroed n_nodes+1 */,
                     int* fill /* zeroed n_nodes */, int* cptr /* n_nodes+1 */, unsigned long long* keys /* E */,
                     int* cedge /* E */, int* scan_tmp, void* stream);

/* ---- edge embedding: replaces sevenn/nn/edge_embedding.py:81-217 (BesselBasis, PolynomialCutoff,
 * XPLORCutoff, SphericalEncoding, EdgeEmbedding.forward).  cutoff_kind 0 = poly_cut(p), 1 = XPLOR(r_on). */
int sevenn_edge_fwd(const float* edge_vec, long long n_edges, int lmax, int normalize, int n_basis,
                    int cutoff_kind, float rc, float p, float r_on, const float* bessel_coeffs,
                    float* Y /* E,(lmax+1)^2 */, float* emb /* E,n_basis */, void* stream);
/* analytic d/d(edge_vec): replaces autograd through the same modules (sevenn/nn/force_output.py:173-178) */
int sevenn_edge_bwd(const float* edge_vec, long long n_edges, int lmax, int normalize, int n_basis,
                    int cutoff_kind, float rc, float p, float r_on, const float* bessel_coeffs,
                    const float* dY /* n_slabs x [E,(lmax+1)^2], summed in slab order */, int n_slabs,
                    long long slab_stride /* floats */, const float* demb, float* f_edge /* E,3 */, void* stream);

/* Tabulated-radial variants.  The radial MLP (sevenn/nn/convolution.py:92-101,121) maps the edge length alone to the
 * tensor-product weights; for fixed model weights that map is tabulated on n_int intervals of width h = rc / n_int
 * (cubic Hermite nodes, sevennet_exp_b200/radial_table.py) and the convolutions interpolate it per edge.
 * Yt row: (Y_0 .. Y_{S-1}, t, bits(i)), S = (lmax+1)^2, edge length = (i + t) h. */
int sevenn_edge_fwd_tab(const float* edge_vec, long long n_edges, int lmax, int normalize, float inv_h, int n_int,
                        float* Yt /* E,S+2 */, void* stream);
/* dY rows: (dE/dY_0 .. dE/dY_{S-1}, sum_c dE/dw_c * h dw_c/dr), one slab per convolution unit */
int sevenn_edge_bwd_tab(const float* edge_vec, long long n_edges, int lmax, int normalize, float inv_h,
                        const float* dY /* n_slabs x [E,S+1] */, int n_slabs, long long slab_stride /* floats */,
                        float* f_edge /* E,3 */, void* stream);

/* ---- dense contractions: replaces e3nn o3.Linear (sevenn/nn/linear.py:49-63,94), the radial MLP
 * e3nn FullyConnectedNet (sevenn/nn/convolution.py:92-101,121) and their input gradients. */
typedef struct {
  const float* A; long long sAn; int sAm;   /* row (n,m) at A + n*sAn + m*sAm, K contiguous */
  const float* B; int ldb;                   /* [K, N] row-major */
  float* C; long long sCn; int sCm;          /* row (n,m) at C + n*sCn + m*sCm */
  long long M; int R; int N; int K;          /* M = nodes*R rows */
  int epi;                                   /* 0 store, 1 accumulate, 2 pre->C & cst*act->C2, 3 acc*cst*act'(P) */
  int act; float cst;                        /* activation id (0 silu,1 tanh,2 abs,3 relu,4 sigmoid,5 ssp) */
  float* C2; const float* P;                 /* aux out / aux in, strides of C */
} sevenn_gemm_args;
int sevenn_gemm_f32(const sevenn_gemm_args* args /* host */, void* stream);   /* exact fp32 (SIMT) */
/* up to SEVENN_GEMM_GROUP_MAX independent problems with disjoint outputs in one launch (the per-l blocks of one or two
 * o3.Linear on a small system); beyond SEVENN_GEMM_GROUP_MAX_TILES 16 x 32 output tiles in total it falls back to one
 * sevenn_gemm_f32 launch per problem.  Results are bit-identical to sevenn_gemm_f32 either way. */
#define SEVENN_GEMM_GROUP_MAX 8
#define SEVENN_GEMM_GROUP_MAX_TILES 2048
int sevenn_gemm_f32_grouped(int n, const sevenn_gemm_args* args /* host */, void* stream);
/* tcgen05/TMEM 3xTF32 arm (fp32-accurate).  B is passed as the pre-swizzled (hi, lo) K-major image
 * [n_tiles][KB][2][Nt][32] whose geometry sevenn_gemm_tc_plan returns; args->B is ignored.
 * Needs K % 4 == 0 and 16-byte aligned A rows; returns SEVENN_ERR_BAD_ARG otherwise. */
int sevenn_gemm_tc_plan(int N, int K, int* Nt, int* n_tiles, int* KB, long long* img_floats);
int sevenn_gemm_tc(const sevenn_gemm_args* args /* host */, const float* Bimg, void* stream);
/* builds that image on the device from a row-major W[K, N] (img: img_floats of sevenn_gemm_tc_plan); for weights that
 * change every step (training) */
int sevenn_gemm_tc_image(const float* W, int K, int N, float* img, void* stream);
/* up to SEVENN_GEMM_GROUP_MAX independent problems with the same epilogue in one launch (the per-l blocks of one or two
 * o3.Linear with disjoint outputs); members with >= 2 tiles per SM are launched on their own */
int sevenn_gemm_tc_grouped(int n, const sevenn_gemm_args* args /* host array */, const float* const* Bimgs /* host array */,
                           void* stream);

/* species-indexed linear = FullyConnectedTensorProduct(x, one_hot): sevenn/nn/self_connection.py:10-64 */
int sevenn_linear_by_type(const float* A, long long sAn, int sAm, const float* W /* S,K,N */, const int* types,
                          float* C, long long sCn, int sCm, long long n_nodes, int R, int K, int N,
                          int accumulate, void* stream);

/* ---- node kernels */
/* OnehotEmbedding + onehot_to_feature_x: sevenn/nn/node_embedding.py:15-53, model_build.py:453-458 */
int sevenn_embed(const int* types, long long n, int dim, const float* table /* S,dim */, float* out, void* stream);
typedef struct {
  int kind;       /* 0 scalar, 1 gated */
  int l, mul;
  int in_off, in_stride, gate_off;
  int out_off, out_stride;
  int act; float cst;
} sevenn_gate_piece;
/* EquivariantGate / e3nn nn.Gate: sevenn/nn/equivariant_gate.py:13-61 */
int sevenn_gate_fwd(const float* pre, int din, float* out, int dout, long long n,
                    const sevenn_gate_piece* pieces /* host */, int n_pieces, void* stream);
int sevenn_gate_bwd(const float* pre, int din, const float* dout, int dout_dim, float* dpre, long long n,
                    const sevenn_gate_piece* pieces /* host */, int n_pieces, void* stream);
/* Rescale / SpeciesWiseRescale: sevenn/nn/scale.py:22-161 ; also emits dE/de_i (= scale) */
int sevenn_rescale(const float* e_scaled, const int* types, long long n, const float* scale, const float* shift,
                   int species_wise, float* atomic_energy, float* seed, void* stream);
/* AtomReduce (sevenn/nn/linear.py:104-141) and the stress sum: deterministic column sums in fp64 */
int sevenn_colsum_f64(const float* a, long long n, int D, double* part /* >= 296*D */, double* out, void* stream);
/* No-allocation variant of sevenn_colsum_f64.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (CUDA-graph capture, tight serving loops, harnesses that
 * forbid it).  Same kernels and results as sevenn_colsum_f64, but the stage-1 partials live in a file-scope
 * `static __device__` array of SEVENN_COLSUM_NOALLOC_MAX_D * 296 doubles (18,944 B per GPU, held until exit) instead of
 * a caller-provided `part`; D > SEVENN_COLSUM_NOALLOC_MAX_D returns SEVENN_ERR_BAD_ARG.  Calls cudaGetDevice /
 * cudaGetSymbolAddress (once per device) and two kernel launches, nothing that allocates.  All calls on one device
 * share the array: calls that overlap on the GPU (different streams) race; calls ordered on a stream do not (every
 * partial that is read is rewritten first, nothing has to start at zero). */
#define SEVENN_COLSUM_NOALLOC_MAX_D 8
int sevenn_colsum_f64_noalloc(const float* a, long long n, int D, double* out, void* stream);
/* ForceStressOutputFromEdge scatter part: sevenn/nn/force_output.py:185-222 */
int sevenn_force_virial(const float* f_edge, const float* edge_vec, const int* rowptr, int n_rows, const int* cptr,
                        const int* cedge, long long n_nodes, float* F /* n_nodes,3 */, float* vir /* n_nodes,6 */,
                        void* stream);
int sevenn_segment_gather_sum(const float* src, int D, const int* cptr, const int* cedge, long long n_nodes,
                              float* out, int accumulate, void* stream);
/* halo pack / unpack-add: sevenn/pair_e3gnn/pair_e3gnn_parallel.cpp:787-897 */
int sevenn_gather_rows(const float* src, int D, const int* idx, long long n, float* out, void* stream);
int sevenn_scatter_add_rows(const float* src, int D, const int* idx, long long n, float* out, void* stream);
/* halo push over peer memory: row i of the packed order (segment p = the rows for peer p, seg_off[p] <= i <
 * seg_off[p+1], seg_off[n_peers] == n) is written to peer_base[p][(dst_row[p] + i - seg_off[p]) * D ..]; the source row
 * is idx[i], or first_row + i when idx == NULL.  peer_base[] are device pointers valid on this device (own or
 * P2P-mapped memory); peer_base / dst_row / seg_off are host arrays.  Replaces pack + MPI relay of
 * pair_e3gnn_parallel.cpp:787-842 (forward_comm) and :844-897 (reverse_comm) by one kernel over NVLink. */
#define SEVENN_MAX_PEERS 16
int sevenn_push_rows_peers(const float* src, int D, const int* idx, long long first_row, long long n, int n_peers,
                           float* const* peer_base, const long long* dst_row, const long long* seg_off, void* stream);

/* device-side edge count variants (n_edges_dev, e.g. &rowptr[n_rows]; launches and buffers bounded by max_edges): for
 * hosts that never wait for the neighbour-list size (csrc/model.cu).  Same kernels as the entries above. */
int sevenn_edge_fwd_tab_dev(const float* edge_vec, long long max_edges, const int* n_edges_dev, int lmax, int normalize,
                            float inv_h, int n_int, float* Yt, void* stream);
int sevenn_edge_bwd_tab_dev(const float* edge_vec, long long max_edges, const int* n_edges_dev, int lmax, int normalize,
                            float inv_h, const float* dY, int n_slabs, long long slab_stride, float* f_edge, void* stream);
int sevenn_csc_build_dev(const int* col, long long max_edges, const int* n_edges_dev, int n_nodes, int* cnt, int* fill,
                         int* cptr, unsigned long long* keys, int* cedge, int* scan_tmp, void* stream);
/* sevenn_nl_fill with the edge arrays bounded by max_edges: rowptr is clamped first, *status |= 1 on overflow; rows
 * for atoms [0, n_rows) only (owned atoms of a sub-domain) */
int sevenn_nl_fill_capped(const double* pos_wrapped, const int* wrap, const int* atom_bin, const int* bin_start,
                          const int* bin_atoms, int n_rows, const double* dpar, const int* ipar, int* rowptr,
                          long long max_edges, int* status, unsigned long long* keys, const double* pos, int* center,
                          int* col, int* shift, float* edge_vec, void* stream);

/* ---- whole model behind one call (csrc/model.cu): replaces what the reference runs per MD step in
 * sevenn/sevennet_calculator.py:163-202 (graph build + AtomGraphSequential.forward + autograd forces / stress) and
 * sevenn/pair_e3gnn/pair_e3gnn.cpp:71-274 / pair_e3gnn_parallel.cpp:191-525 (edge loop, torch::jit forward per segment,
 * ghost communication, result unpacking).  No allocation per call (caller-owned workspace), no host synchronisation
 * (the edge count stays on the device), launches only: capturable in a CUDA graph.
 * The model is the (int32 program, fp32 data) pair written by sevennet_exp_b200/host.py::pack_engine (tabulated-radial
 * mode); sevenn_model_create uploads the weights once. */
#define SEVENN_ERR_WORKSPACE 10002
typedef struct sevenn_model sevenn_model;
typedef struct {                /* CSR by centre (edge_index[0]) + CSC by neighbour, device pointers */
  long long n_local, n_nodes;   /* owned atoms (rows), owned + ghost atoms */
  long long max_edges;          /* capacity of the edge arrays */
  const int* n_edges_dev;       /* device: valid edges (== rowptr[n_local]) */
  const int* types;             /* n_nodes */
  const float* edge_vec;        /* max_edges x 3: pos[col] + shift.cell - pos[centre] */
  const int *rowptr, *col, *cptr, *cedge;
} sevenn_graph;
typedef struct {                /* device outputs; forces == NULL skips the reverse pass */
  double* energy;               /* 1 */
  float* atomic_energy;         /* n_local */
  float* forces;                /* n_nodes x 3 (ghost rows: what is owed to their owners) */
  double* virial6;              /* 6: sum_e (rx fx, ry fy, rz fz, rx fy, ry fz, rz fx); stress = -virial / V */
  float* virial_atoms;          /* optional n_nodes x 6 */
  int* status;                  /* 2 ints (evaluate_positions): [0] bit 0 = edge capacity exceeded, [1] = edges found */
} sevenn_outputs;
typedef struct {                /* per-layer ghost exchange of a domain-decomposed host (may be NULL) */
  void* user;
  int (*forward)(void* user, int layer, float* h /* n_nodes x dim; fill rows >= n_local */, int dim, void* stream);
  int (*backward)(void* user, int layer, float* dh /* add ghost rows into their owners */, int dim, void* stream);
  /* Overlap: rows [0, n_interior) have no ghost neighbour (the host orders its owned atoms that way).  When
   * n_interior > 0 and side_stream != NULL the forward exchange of a layer is issued on side_stream while the convolution
   * of the interior rows runs on the main stream; the boundary rows follow once the exchange has landed.  0 / NULL: the
   * exchange runs in line on the main stream. */
  long long n_interior;
  void* side_stream;
} sevenn_halo_hooks;
int sevenn_model_create(const int* prog /* host */, long long n_prog, const float* data /* host */, long long n_data,
                        sevenn_model** out);
int sevenn_model_destroy(sevenn_model* model);
int sevenn_model_set_tc_min_rows(sevenn_model* model, int rows);
long long sevenn_model_launches(const sevenn_model* model);      /* library calls made so far (launch groups) */
long long sevenn_model_tc_launches(const sevenn_model* model);   /* ... of which on the tcgen05 GEMM arm */
long long sevenn_model_workspace_bytes(sevenn_model* model, long long n_local, long long n_nodes, long long max_edges,
                                       int with_neighbor_list);
int sevenn_model_count_edges(sevenn_model* model, long long n_rows, long long n_nodes, const double* pos,
                             const double* pos_host, const double* cell, const int* pbc, void* workspace,
                             long long workspace_bytes, void* stream, long long* n_edges /* host */);
int sevenn_model_evaluate_graph(sevenn_model* model, const sevenn_graph* graph, const sevenn_halo_hooks* halo,
                                void* workspace, long long workspace_bytes, void* stream, const sevenn_outputs* out);
/* pos: device fp64 [n_nodes, 3]; pos_host: the same on the host, needed only when a direction is not periodic */
int sevenn_model_evaluate_positions(sevenn_model* model, long long n_local, long long n_nodes, const double* pos,
                                    const double* pos_host, const double* cell /* host 9 */, const int* pbc /* host 3 */,
                                    const int* types, long long max_edges, const sevenn_halo_hooks* halo,
                                    void* workspace, long long workspace_bytes, void* stream, const sevenn_outputs* out);

/* deployed-model contract: the tensors the LAMMPS pair style puts into the TorchScript input dict
 * (sevenn/pair_e3gnn/pair_e3gnn.cpp:218-225) as device pointers -- x int64 [n] (type index), pos fp32 [n, 3],
 * edge_index int64 [2, E] (row 0 = centre, row 1 = neighbour, any order), cell_lattice_vectors fp32 [3, 3],
 * pbc_shift fp32 [E, 3].  Replaces EdgePreprocess + model + ForceStressOutput (sevenn/nn/edge_embedding.py:13-78,
 * sevenn/nn/force_output.py:55-136).  csrc/torch_op.cpp wraps it as torch.ops.sevenn_b200.evaluate. */
long long sevenn_model_edges_workspace_bytes(sevenn_model* model, long long n_atoms, long long n_edges);
int sevenn_model_evaluate_edges(sevenn_model* model, long long n_atoms, const long long* x, const float* pos,
                                const long long* edge_index, long long n_edges, const float* cell,
                                const float* pbc_shift, void* workspace, long long workspace_bytes, void* stream,
                                const sevenn_outputs* out);

/* ---- training step (csrc/train.cu): what loss.backward() through forces and stress needs on top of the first-order
 * kernels (sevenn/train/trainer.py:156-170, sevenn/nn/force_output.py:173-178 `create_graph=self.training`).  Second
 * order is composed from the first-order kernels in sevennet_exp_b200/nn/functional.py (every op is multilinear or
 * elementwise); these are the missing pieces. */
/* out = (m ? m : 1) * cst * f^(order)(x), order 0..2, f = activation `act` (e3nn radial MLP / gate non-linearities) */
int sevenn_act_mul(const float* x, const float* m, long long n, int act, float cst, int order, float* out, void* stream);
/* derivative of sevenn_gate_bwd: given c = cotangent of dpre, d_dout = d<c,dpre>/d(dout), d_pre = d<c,dpre>/d(pre) */
int sevenn_gate_bwd2(const float* pre, int din, const float* dout, int dout_dim, const float* c, long long n,
                     const sevenn_gate_piece* pieces /* host */, int n_pieces, float* d_dout, float* d_pre, void* stream);
/* c == NULL: G[e,k] = demb[e,k] * d emb_k/d coeff_k (trainable Bessel coefficients, edge_embedding.py:96-100).
 * c != NULL: JY = dY/d(edge_vec).c, Jemb = d emb/d(edge_vec).c per edge, G[e,k] = demb[e,k] d/d coeff_k(d emb_k/d vec . c) */
int sevenn_edge_train(const float* edge_vec, long long n_edges, int lmax, int normalize, int n_basis, int cutoff_kind,
                      float rc, float p, float r_on, const float* bessel_coeffs, const float* demb, const float* c,
                      float* JY, float* Jemb, float* G, void* stream);
/* weight gradient G[K,N] (+)= scale * sum_rows A[row,:K]^T B[row,:N] (rows (node, m) strided like sevenn_gemm_args;
 * types != NULL keeps nodes of species type_sel only); deterministic two-stage reduction through `part` */
int sevenn_gemm_tn_splits(long long M);
int sevenn_gemm_tn(const float* A, long long sAn, int sAm, const float* B, long long sBn, int sBm, long long M, int R,
                   int K, int N, const int* types, int type_sel, float scale, int accumulate,
                   float* part /* >= sevenn_gemm_tn_splits(M) * K * N */, float* out /* K x N */, void* stream);
/* torch.optim.Adam's update on one flat parameter (sevenn/train/optim.py 'adam'); step >= 1 */
int sevenn_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, long long n, int step, float lr,
                     float beta1, float beta2, float eps, float weight_decay, void* stream);
/* ... with the step count (>= 1) read from device memory, for CUDA-graph replay of a training step */
int sevenn_adam_step_dev(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, long long n,
                         const int* step_dev, float lr, float beta1, float beta2, float eps, float weight_decay, void* stream);

/* ---- fused gather + uvu CG tensor product + segment sum (generated per irreps signature <sig>):
 *   int sevenn_conv_fwd_<sig>(x, ldx, Y, w, rowptr, col, n_dst, inv_denom, out, ldo, stream)
 *   int sevenn_conv_bwd_<sig>(x, ldx, Y, w, rowptr, col, n_dst, inv_denom, gout, ldo, dW, dxe, lddx, dY, dy_stride, stream)
 *   int sevenn_conv_info_<sig>(&x_dim, &mid_dim, &wn, &sh_dim, &units)
 * tabulated-radial variants (w interpolated from tab[n_int][wn] float4 nodes; Yt / dY rows as for sevenn_edge_*_tab):
 *   int sevenn_conv_fwd_tab_<sig>(x, ldx, Yt, tab, rowptr, col, n_dst, inv_denom, out, ldo, stream)
 *   int sevenn_conv_bwd_tab_<sig>(x, ldx, Yt, tab, rowptr, col, n_dst, inv_denom, gout, ldo, dxe, lddx, dY, dy_stride, stream)
 * dxe == NULL in the backward calls: the per-edge dE/dx rows are not written (first layer: x does not depend on positions).
 * replaces IrrepsConvolution.forward's TensorProduct + message_gather (sevenn/nn/convolution.py:16-25,118-137). */

#ifdef __cplusplus
}
#endif
#endif

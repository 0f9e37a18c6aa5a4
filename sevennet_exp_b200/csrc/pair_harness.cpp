// sevenn_pair_harness: what a LAMMPS `pair_style e3gnn` process does with a deployed model, without LAMMPS.
//
// The reference pair style (sevenn/pair_e3gnn/pair_e3gnn.cpp) is a libtorch client: coeff() loads the TorchScript
// archive with its `_extra_files` (:306-335: chemical_symbols_to_index, cutoff, model_type, ...), compute() builds the
// serial input dict from LAMMPS' full neighbour list (:147-225: x, pos, edge_index, num_atoms, cell_lattice_vectors,
// cell_volume, pbc_shift), calls forward and reads inferred_total_energy / inferred_force / inferred_stress /
// atomic_energy (:230-264).  This program does exactly that sequence of libtorch calls on an archive written by
// sevennet_exp_b200.deploy.deploy(): the only addition a pair style needs is the dlopen of libsevenn_b200_torch.so
// (registers torch.ops.sevenn_b200.evaluate) before torch::jit::load -- INTEGRATION.md, seam 3.  tests/ run it on the
// reference's HfO2 frames and compare with the calculator (tests/test_gpu_parity.py).
//
//   sevenn_pair_harness <op library .so> <deployed .pt> <structure.txt> <result.txt>
//   structure.txt: n / 9 cell numbers (rows = lattice vectors) / n lines "symbol x y z"
//   result.txt   : energy / 6 stress (model order xx yy zz xy yz zx) / n lines "fx fy fz e_atom"
#include <dlfcn.h>
#include <torch/script.h>

#include <cmath>
#include <cstdio>
#include <fstream>
#include <iostream>
#include <sstream>
#include <string>
#include <unordered_map>
#include <vector>

int main(int argc, char** argv) {
  if (argc != 5) {
    std::fprintf(stderr, "usage: %s <libsevenn_b200_torch.so> <deployed.pt> <structure.txt> <result.txt>\n", argv[0]);
    return 2;
  }
  if (dlopen(argv[1], RTLD_NOW | RTLD_GLOBAL) == nullptr) {
    std::fprintf(stderr, "dlopen %s: %s\n", argv[1], dlerror());
    return 3;
  }
  // ---- coeff(): archive + metadata
  std::unordered_map<std::string, std::string> meta = {{"chemical_symbols_to_index", ""}, {"cutoff", ""},
                                                       {"num_species", ""}, {"model_type", ""}, {"version", ""},
                                                       {"dtype", ""}, {"time", ""}};
  const torch::Device device(torch::kCUDA, 0);
  torch::jit::Module model = torch::jit::load(argv[2], device, meta);
  model.eval();
  const double cutoff = std::stod(meta["cutoff"]);
  std::unordered_map<std::string, long> species;                 // "Hf O" -> {Hf: 0, O: 1}
  {
    std::istringstream ss(meta["chemical_symbols_to_index"]);
    std::string sym;
    long k = 0;
    while (ss >> sym) species[sym] = k++;
  }
  if (meta["model_type"] != "E3_equivariant_model" || species.empty()) {
    std::fprintf(stderr, "unexpected metadata: model_type '%s'\n", meta["model_type"].c_str());
    return 4;
  }
  // ---- the "LAMMPS side": atoms and box
  std::ifstream in(argv[3]);
  long n = 0;
  in >> n;
  double cell[3][3];
  for (auto& row : cell)
    for (double& v : row) in >> v;
  std::vector<long> type(n);
  std::vector<double> x(3 * n);
  for (long i = 0; i < n; ++i) {
    std::string sym;
    in >> sym >> x[3 * i] >> x[3 * i + 1] >> x[3 * i + 2];
    if (!species.count(sym)) { std::fprintf(stderr, "species %s not in the model\n", sym.c_str()); return 5; }
    type[i] = species[sym];
  }
  if (!in) { std::fprintf(stderr, "cannot parse %s\n", argv[3]); return 5; }
  // full neighbour list over periodic images, strict r^2 < rc^2 like the pair style (:184); the image offset of each
  // neighbour plays the role of LAMMPS' ghost-atom coordinates
  auto cross = [](const double* a, const double* b, double* c) {
    c[0] = a[1] * b[2] - a[2] * b[1]; c[1] = a[2] * b[0] - a[0] * b[2]; c[2] = a[0] * b[1] - a[1] * b[0];
  };
  double c12[3], c20[3], c01[3];
  cross(cell[1], cell[2], c12); cross(cell[2], cell[0], c20); cross(cell[0], cell[1], c01);
  const double vol = cell[0][0] * c12[0] + cell[0][1] * c12[1] + cell[0][2] * c12[2];
  const double* nrm[3] = {c12, c20, c01};
  int reach[3];
  for (int d = 0; d < 3; ++d) {
    const double h = std::fabs(vol) / std::sqrt(nrm[d][0] * nrm[d][0] + nrm[d][1] * nrm[d][1] + nrm[d][2] * nrm[d][2]);
    reach[d] = (int)std::ceil(cutoff / h);
  }
  std::vector<long> src, dst;
  std::vector<float> shift_cart;                                   // x_ghost - x_owner per edge (Cartesian)
  for (long i = 0; i < n; ++i)
    for (long j = 0; j < n; ++j)
      for (int a = -reach[0]; a <= reach[0]; ++a)
        for (int b = -reach[1]; b <= reach[1]; ++b)
          for (int c = -reach[2]; c <= reach[2]; ++c) {
            if (i == j && a == 0 && b == 0 && c == 0) continue;
            double s[3], d2 = 0.0;
            for (int k = 0; k < 3; ++k) {
              s[k] = a * cell[0][k] + b * cell[1][k] + c * cell[2][k];
              const double d = x[3 * j + k] + s[k] - x[3 * i + k];
              d2 += d * d;
            }
            if (d2 < cutoff * cutoff) {
              src.push_back(i); dst.push_back(j);
              for (double v : s) shift_cart.push_back((float)v);
            }
          }
  const long E = (long)src.size();
  // ---- compute(): the input dict of pair_e3gnn.cpp:197-225
  const auto f32 = torch::TensorOptions().dtype(torch::kFloat32);
  const auto i64 = torch::TensorOptions().dtype(torch::kInt64);
  torch::Tensor inp_cell = torch::zeros({3, 3}, f32);
  for (int r = 0; r < 3; ++r)
    for (int k = 0; k < 3; ++k) inp_cell[r][k] = (float)cell[r][k];
  torch::Tensor inp_pos = torch::zeros({n, 3}, f32);
  torch::Tensor inp_type = torch::from_blob(type.data(), {n}, i64).clone();
  {
    auto p = inp_pos.accessor<float, 2>();
    for (long i = 0; i < n; ++i)
      for (int k = 0; k < 3; ++k) p[i][k] = (float)x[3 * i + k];
  }
  torch::Tensor inp_volume = torch::dot(inp_cell[0], torch::cross(inp_cell[1], inp_cell[2], 0));
  torch::Tensor edge_index = torch::stack({torch::from_blob(src.data(), {E}, i64), torch::from_blob(dst.data(), {E}, i64)});
  // integer cell shifts from the Cartesian offsets: shift = (cell^-1)^T @ (r' - r)
  torch::Tensor cell_inv_t = inp_cell.inverse().transpose(0, 1).unsqueeze(0).to(device);
  torch::Tensor shift = torch::bmm(cell_inv_t.expand({E, 3, 3}),
                                   torch::from_blob(shift_cart.data(), {E, 3}, f32).view({E, 3, 1}).to(device)).view({E, 3});
  long num_atoms[1] = {n};
  inp_pos.set_requires_grad(true);
  c10::Dict<std::string, torch::Tensor> input_dict;
  input_dict.insert("x", inp_type.to(device));
  input_dict.insert("pos", inp_pos.to(device));
  input_dict.insert("edge_index", edge_index.to(device));
  input_dict.insert("num_atoms", torch::from_blob(num_atoms, {1}, i64).to(device));
  input_dict.insert("cell_lattice_vectors", inp_cell.to(device));
  input_dict.insert("cell_volume", inp_volume.to(device));
  input_dict.insert("pbc_shift", shift);
  std::vector<torch::IValue> input(1, input_dict);
  auto output = model.forward(input).toGenericDict();
  // ---- what the pair style reads back (:230-264)
  const float energy = output.at("inferred_total_energy").toTensor().cpu().item<float>();
  torch::Tensor force = output.at("inferred_force").toTensor().cpu().contiguous();
  torch::Tensor stress = output.at("inferred_stress").toTensor().cpu().contiguous();
  torch::Tensor e_atom = output.at("atomic_energy").toTensor().cpu().squeeze().contiguous();
  std::FILE* out = std::fopen(argv[4], "w");
  if (out == nullptr) return 6;
  std::fprintf(out, "%.9g\n", energy);
  for (int k = 0; k < 6; ++k) std::fprintf(out, "%.9g%c", stress.data_ptr<float>()[k], k == 5 ? '\n' : ' ');
  for (long i = 0; i < n; ++i)
    std::fprintf(out, "%.9g %.9g %.9g %.9g\n", force.data_ptr<float>()[3 * i], force.data_ptr<float>()[3 * i + 1],
                 force.data_ptr<float>()[3 * i + 2], e_atom.data_ptr<float>()[i]);
  std::fclose(out);
  std::printf("sevenn_pair_harness: %ld atoms, %ld edges, E = %.6f eV\n", n, E, energy);
  return 0;
}

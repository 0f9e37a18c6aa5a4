// A C++ host on the C ABI, without torch and without Python: what a LAMMPS pair style does per step
// (sevenn/pair_e3gnn/pair_e3gnn.cpp:71-274: gather positions and types, build the edge list, call the model, add forces,
// energy, virial) written against include/sevenn_b200.h, plus a velocity-Verlet loop around it so that it can be run and
// checked on its own (the image has no LAMMPS).  The pair-style version of the same calls is in INTEGRATION.md.
//
//   sevenn_md_host model.bin structure.bin n_steps dt_fs out.bin
//
// model.bin     : written by sevennet_exp_b200.host.save_model (magic, n_prog, n_data, int32 program, fp32 data)
// structure.bin : int64 n | double cell[9] | int32 pbc[3] | int32 types[n] | double pos[3n] | double mass[n] | double vel[3n]
// out.bin       : per step (double E_pot, double E_kin, double virial[6]); then the forces of the first step as double[3n]
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../../include/sevenn_b200.h"

#define CK(x)                                                                      \
  do {                                                                             \
    int rc_ = (int)(x);                                                            \
    if (rc_ != 0) { std::fprintf(stderr, "%s failed: %d (line %d)\n", #x, rc_, __LINE__); return 1; } \
  } while (0)

template <class T>
static bool read_vec(std::FILE* f, std::vector<T>& v, size_t n) {
  v.resize(n);
  return std::fread(v.data(), sizeof(T), n, f) == n;
}

int main(int argc, char** argv) {
  if (argc < 6) { std::fprintf(stderr, "usage: %s model.bin structure.bin n_steps dt_fs out.bin\n", argv[0]); return 2; }
  const int n_steps = std::atoi(argv[3]);
  const double dt = std::atof(argv[4]);
  // ---- model
  std::FILE* f = std::fopen(argv[1], "rb");
  if (!f) { std::perror(argv[1]); return 1; }
  long long hdr[3];
  if (std::fread(hdr, 8, 3, f) != 3 || hdr[0] != 0x37454E42) { std::fprintf(stderr, "bad model file\n"); return 1; }
  std::vector<int> prog; std::vector<float> data;
  if (!read_vec(f, prog, (size_t)hdr[1]) || !read_vec(f, data, (size_t)hdr[2])) { std::fprintf(stderr, "short model file\n"); return 1; }
  std::fclose(f);
  sevenn_model* model = nullptr;
  CK(sevenn_model_create(prog.data(), (long long)prog.size(), data.data(), (long long)data.size(), &model));
  // ---- structure
  f = std::fopen(argv[2], "rb");
  if (!f) { std::perror(argv[2]); return 1; }
  long long n = 0;
  std::vector<double> cell, pos, mass, vel; std::vector<int> pbc, types;
  if (std::fread(&n, 8, 1, f) != 1 || !read_vec(f, cell, 9) || !read_vec(f, pbc, 3) || !read_vec(f, types, (size_t)n) ||
      !read_vec(f, pos, (size_t)3 * n) || !read_vec(f, mass, (size_t)n) || !read_vec(f, vel, (size_t)3 * n)) {
    std::fprintf(stderr, "short structure file\n");
    return 1;
  }
  std::fclose(f);
  // ---- device buffers: allocated once; the step itself allocates nothing
  cudaStream_t st;
  CK(cudaStreamCreate(&st));
  double* d_pos; int* d_types; char* d_res; void* ws = nullptr; long long ws_bytes = 0;
  CK(cudaMalloc(&d_pos, (size_t)3 * n * 8));
  CK(cudaMalloc(&d_types, (size_t)n * 4));
  const size_t res_bytes = 64 + (size_t)n * 16;                 // E, virial[6], status[2] | forces | atomic energies
  CK(cudaMalloc(&d_res, res_bytes));
  CK(cudaMemcpy(d_types, types.data(), (size_t)n * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d_pos, pos.data(), (size_t)3 * n * 8, cudaMemcpyHostToDevice));
  {   // the one synchronous call: size the edge capacity from the first configuration (+20 %, like pair_e3gnn.cpp:266-273)
    const long long tmp_bytes = 256 * n + (1 << 20);
    void* tmp;
    CK(cudaMalloc(&tmp, (size_t)tmp_bytes));
    long long e0 = 0;
    CK(sevenn_model_count_edges(model, n, n, d_pos, pos.data(), cell.data(), pbc.data(), tmp, tmp_bytes, st, &e0));
    cudaFree(tmp);
    const long long cap = (long long)(e0 * 1.2) + 1024;
    ws_bytes = sevenn_model_workspace_bytes(model, n, n, cap, 1);
    if (ws_bytes < 0) { std::fprintf(stderr, "workspace query failed\n"); return 1; }
    CK(cudaMalloc(&ws, (size_t)ws_bytes));
    std::printf("atoms %lld edges %lld capacity %lld workspace %.1f MB\n", n, e0, cap, ws_bytes / 1e6);
    sevenn_outputs out;
    out.energy = (double*)d_res; out.virial6 = (double*)d_res + 1; out.status = (int*)(d_res + 56);
    out.forces = (float*)(d_res + 64); out.atomic_energy = (float*)(d_res + 64 + (size_t)n * 12); out.virial_atoms = nullptr;
    std::vector<char> res(res_bytes);
    std::vector<double> f0((size_t)3 * n), force((size_t)3 * n);
    std::FILE* fo = std::fopen(argv[5], "wb");
    if (!fo) { std::perror(argv[5]); return 1; }
    const double ACC = 9.648533212331e-3;     // (eV/A/amu) -> A/fs^2
    for (int step = 0; step <= n_steps; ++step) {
      CK(cudaMemcpyAsync(d_pos, pos.data(), (size_t)3 * n * 8, cudaMemcpyHostToDevice, st));
      CK(sevenn_model_evaluate_positions(model, n, n, d_pos, pos.data(), cell.data(), pbc.data(), d_types, cap, nullptr,
                                         ws, ws_bytes, st, &out));
      CK(cudaMemcpyAsync(res.data(), d_res, res_bytes, cudaMemcpyDeviceToHost, st));
      CK(cudaStreamSynchronize(st));
      const double* hd = (const double*)res.data();
      const int* status = (const int*)(res.data() + 56);
      if (status[0] & 1) { std::fprintf(stderr, "edge capacity exceeded (%d edges)\n", status[1]); return 3; }
      const float* hf = (const float*)(res.data() + 64);
      for (long long i = 0; i < 3 * n; ++i) force[i] = hf[i];
      if (step == 0) f0 = force;
      else
        for (long long i = 0; i < n; ++i)
          for (int c = 0; c < 3; ++c) vel[3 * i + c] += 0.5 * dt * ACC * force[3 * i + c] / mass[i];   // second half kick
      double ekin = 0.0;
      for (long long i = 0; i < n; ++i)
        for (int c = 0; c < 3; ++c) ekin += 0.5 * mass[i] * vel[3 * i + c] * vel[3 * i + c] / ACC;
      const double rec[8] = {hd[0], ekin, hd[1], hd[2], hd[3], hd[4], hd[5], hd[6]};
      std::fwrite(rec, 8, 8, fo);
      std::printf("step %d  E_pot %.8f  E_kin %.8f  E_tot %.8f eV  edges %d\n", step, hd[0], ekin, hd[0] + ekin, status[1]);
      if (step == n_steps) break;
      for (long long i = 0; i < n; ++i)
        for (int c = 0; c < 3; ++c) {
          vel[3 * i + c] += 0.5 * dt * ACC * force[3 * i + c] / mass[i];
          pos[3 * i + c] += dt * vel[3 * i + c];
        }
    }
    std::fwrite(f0.data(), 8, f0.size(), fo);
    std::fclose(fo);
    std::printf("library calls %lld\n", sevenn_model_launches(model));
  }
  sevenn_model_destroy(model);
  cudaFree(ws); cudaFree(d_res); cudaFree(d_types); cudaFree(d_pos);
  return 0;
}

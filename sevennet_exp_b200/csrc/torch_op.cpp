// torch.ops.sevenn_b200.evaluate: the TorchScript-visible face of the C++ whole-model host (csrc/model.cu).
//
// The reference deploys a model by scripting + freezing the e3nn graph (sevenn/scripts/deploy.py:47-69) and LAMMPS
// calls its forward(Dict[str, Tensor]) (sevenn/pair_e3gnn/pair_e3gnn.cpp:218-264).  Here deploy() scripts a thin
// module whose forward hands the same dict entries to this op; the module file carries the packed model (int32
// program + fp32 data) as buffers, so `torch::jit::load` + one `dlopen` of this library is all a C++ host needs.
// This file is the only translation unit that sees torch; everything below the op is the plain C ABI.
#include <torch/library.h>
#include <torch/torch.h>
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>

#include <mutex>
#include <tuple>
#include <unordered_map>

#include "../../include/sevenn_b200.h"

namespace {

struct Entry {
  sevenn_model* model = nullptr;
  int device = -1;
  torch::Tensor workspace;
};
std::mutex g_mutex;
std::unordered_map<const void*, Entry> g_models;     // keyed by the program buffer of the scripted module

Entry& model_for(const torch::Tensor& prog, const torch::Tensor& data, int device) {
  std::lock_guard<std::mutex> lock(g_mutex);
  Entry& e = g_models[prog.data_ptr()];
  if (e.model == nullptr || e.device != device) {
    if (e.model != nullptr) sevenn_model_destroy(e.model);
    torch::Tensor p = prog.to(torch::kCPU, torch::kInt32).contiguous();
    torch::Tensor d = data.to(torch::kCPU, torch::kFloat32).contiguous();
    int rc = sevenn_model_create(p.data_ptr<int>(), p.numel(), d.data_ptr<float>(), d.numel(), &e.model);
    TORCH_CHECK(rc == 0, "sevenn_model_create failed with status ", rc);
    e.device = device;
    e.workspace = torch::Tensor();
  }
  return e;
}

// (prog, data, x, pos, edge_index, cell_lattice_vectors, pbc_shift) -> (energy (), forces [n,3], virial [6], atomic_energy [n,1])
std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor> evaluate(
    const torch::Tensor& prog, const torch::Tensor& data, const torch::Tensor& x, const torch::Tensor& pos,
    const torch::Tensor& edge_index, const torch::Tensor& cell, const torch::Tensor& pbc_shift) {
  TORCH_CHECK(torch::cuda::is_available(), "sevenn_b200::evaluate needs a CUDA device (B200, sm_100a); there is no CPU fallback");
  const torch::Device dev = pos.is_cuda() ? pos.device() : torch::Device(torch::kCUDA, c10::cuda::current_device());
  c10::cuda::CUDAGuard guard(dev);
  TORCH_CHECK(edge_index.dim() == 2 && edge_index.size(0) == 2, "edge_index must be [2, E]");
  auto f32 = torch::TensorOptions().dtype(torch::kFloat32).device(dev);
  auto f64 = torch::TensorOptions().dtype(torch::kFloat64).device(dev);
  torch::Tensor xd = x.to(dev, torch::kInt64).contiguous();
  torch::Tensor pd = pos.detach().to(dev, torch::kFloat32).contiguous();
  torch::Tensor ed = edge_index.to(dev, torch::kInt64).contiguous();
  torch::Tensor cd = cell.to(dev, torch::kFloat32).contiguous();
  torch::Tensor sd = pbc_shift.to(dev, torch::kFloat32).contiguous();
  const long long n = xd.numel(), E = ed.size(1);
  TORCH_CHECK(pd.numel() == 3 * n && sd.numel() == 3 * E && cd.numel() == 9, "inconsistent input shapes");
  Entry& e = model_for(prog, data, dev.index());
  const long long need = sevenn_model_edges_workspace_bytes(e.model, n, E);
  TORCH_CHECK(need >= 0, "sevenn_model_edges_workspace_bytes failed");
  if (!e.workspace.defined() || e.workspace.numel() < need)
    e.workspace = torch::empty({(long long)(need * 1.2) + 4096}, torch::TensorOptions().dtype(torch::kUInt8).device(dev));
  torch::Tensor ev = torch::zeros({7}, f64);                  // energy, virial x 6
  torch::Tensor forces = torch::zeros({n, 3}, f32), ae = torch::zeros({n, 1}, f32);
  sevenn_outputs out;
  out.energy = ev.data_ptr<double>();
  out.virial6 = ev.data_ptr<double>() + 1;
  out.atomic_energy = ae.data_ptr<float>();
  out.forces = forces.data_ptr<float>();
  out.virial_atoms = nullptr;
  out.status = nullptr;
  int rc = sevenn_model_evaluate_edges(e.model, n, xd.data_ptr<int64_t>() ? (const long long*)xd.data_ptr<int64_t>() : nullptr,
                                       pd.data_ptr<float>(), (const long long*)ed.data_ptr<int64_t>(), E,
                                       cd.data_ptr<float>(), sd.data_ptr<float>(), e.workspace.data_ptr(),
                                       e.workspace.numel(), (void*)c10::cuda::getCurrentCUDAStream(dev.index()).stream(),
                                       &out);
  TORCH_CHECK(rc == 0, "sevenn_model_evaluate_edges failed with status ", rc);
  return std::make_tuple(ev[0].to(torch::kFloat32), forces, ev.slice(0, 1, 7).to(torch::kFloat32), ae);
}

}  // namespace

TORCH_LIBRARY(sevenn_b200, m) {
  m.def("evaluate(Tensor prog, Tensor data, Tensor x, Tensor pos, Tensor edge_index, Tensor cell, Tensor pbc_shift) -> (Tensor, Tensor, Tensor, Tensor)");
}
TORCH_LIBRARY_IMPL(sevenn_b200, CompositeExplicitAutograd, m) { m.impl("evaluate", &evaluate); }

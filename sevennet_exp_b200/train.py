"""Training step on the B200 kernels (BASELINE config 5).

Mirrors the reference's ``Trainer.run_one_epoch`` (sevenn/train/trainer.py:137-170) and loss definitions
(sevenn/train/loss.py:78-196): per-atom energy MSE + force MSE + stress MSE in kbar (x 1602.1766208), weights from the
config; the model is ``sevennet_exp_b200.nn.AtomGraphSequential`` in training mode, whose forward / reverse / double
reverse all run through the C ABI (nn/train_functional.py); Adam's update is ``sevenn_adam_step`` (csrc/train.cu).
Graph batching follows torch_geometric's Collater as the reference uses it (node offsets per graph, ``batch`` vector,
per-graph ``num_atoms`` / ``cell_volume``)."""
from typing import Dict, List, Optional

import torch

from .nn._backend import kernels, ptr

TO_KB = 1602.1766208           # eV/A^3 -> kbar, sevenn/train/loss.py:180
K_X, K_Z, K_POS, K_EI, K_EV, K_VOL, K_NATOMS, K_BATCH = ('x', 'atomic_numbers', 'pos', 'edge_index', 'edge_vec',
                                                         'cell_volume', 'num_atoms', 'batch')
K_E, K_F, K_S = 'total_energy', 'force_of_atoms', 'stress'
P_E, P_F, P_S = 'inferred_total_energy', 'inferred_force', 'inferred_stress'


def collate(graphs: List[Dict[str, torch.Tensor]], device='cuda') -> Dict[str, torch.Tensor]:
    """Block-diagonal batch of graph dicts (keys as in the reference's AtomGraphData): node indices offset per graph."""
    out, off, ei, batch = {}, 0, [], []
    for b, g in enumerate(graphs):
        n = int(g[K_NATOMS])
        ei.append(g[K_EI] + off)
        batch.append(torch.full((n,), b, dtype=torch.long))
        off += n
    for k in (K_X, K_Z, K_POS, K_EV, K_F):
        if all(k in g for g in graphs):
            out[k] = torch.cat([g[k].detach() for g in graphs])
    out[K_EI] = torch.cat(ei, 1)
    out[K_BATCH] = torch.cat(batch)
    for k in (K_VOL, K_NATOMS, K_E, K_S):
        if all(k in g for g in graphs):
            out[k] = torch.stack([torch.as_tensor(g[k]).detach() for g in graphs])
    return {k: v.to(device) for k, v in out.items()}


def loss_from_config(output: Dict[str, torch.Tensor], config: dict) -> torch.Tensor:
    """sevenn/train/loss.py:198-223 with criterion MSE (mean over elements)."""
    mse = torch.nn.functional.mse_loss
    n = output[K_NATOMS].to(output[P_E].dtype)
    loss = mse(output[P_E] / n, output[K_E].to(n.dtype) / n)
    loss = loss + float(config.get('force_loss_weight', 1.0)) * mse(output[P_F].reshape(-1),
                                                                    output[K_F].to(n.dtype).reshape(-1))
    if config.get('is_train_stress', True):
        loss = loss + float(config.get('stress_loss_weight', 0.1)) * mse(
            (output[P_S] * TO_KB).reshape(-1), (output[K_S].to(n.dtype) * TO_KB).reshape(-1))
    return loss


class Adam:
    """torch.optim.Adam semantics (sevenn/train/optim.py 'adam'; defaults lr 1e-3, betas (0.9, 0.999), eps 1e-8) with
    the update done by sevenn_adam_step on each parameter's storage."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, **_):
        self.params = [p for p in params if p.requires_grad]
        self.lr, self.betas, self.eps, self.wd = float(lr), (float(betas[0]), float(betas[1])), float(eps), float(weight_decay)
        self.state = {}
        self.t = 0
        self.t_dev = None

    def zero_grad(self):
        for p in self.params:
            p.grad = None

    def step(self):
        self.t += 1
        dev = self.params[0].device if self.params else None
        if dev is not None and dev.type == 'cuda':
            # step count on the device: the same launches stay valid when the step is replayed from a CUDA graph
            if self.t_dev is None:
                self.t_dev = torch.zeros(1, dtype=torch.int32, device=dev)
            self.t_dev.add_(1)
        for p in self.params:
            if p.grad is None:
                continue
            k = kernels(p.device)
            st = self.state.get(p)
            if st is None:
                st = self.state[p] = (torch.zeros_like(p.data), torch.zeros_like(p.data))
            g = p.grad.detach().float().contiguous()
            assert p.data.is_contiguous() and p.data.dtype == torch.float32
            if self.t_dev is not None:
                k.lib.call('sevenn_adam_step_dev', ptr(p.data), ptr(g), ptr(st[0]), ptr(st[1]), p.data.numel(),
                           ptr(self.t_dev), self.lr, self.betas[0], self.betas[1], self.eps, self.wd, k._stream())
            else:
                k.lib.call('sevenn_adam_step', ptr(p.data), ptr(g), ptr(st[0]), ptr(st[1]), p.data.numel(), self.t,
                           self.lr, self.betas[0], self.betas[1], self.eps, self.wd, k._stream())
            k.n_launch += 1
            # the parameter changed in place behind autograd's back: bump its version so cached device weights rebuild
            with torch.no_grad():
                p.add_(0)


class Trainer:
    """One optimiser step per batch, like Trainer.run_one_epoch(is_train=True) of the reference."""

    def __init__(self, model, config: dict, optimizer: Optional[Adam] = None):
        self.model, self.config = model, config
        self.model.set_is_batch_data(True)
        self.optimizer = optimizer or Adam(self.model.parameters(), **config.get('optim_param', {}))

    def forward(self, batch: Dict[str, torch.Tensor]):
        return self.model({k: (v.clone() if torch.is_tensor(v) else v) for k, v in batch.items()})

    def prepare(self, batch: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        """Adds what would otherwise cost a host synchronisation per step: the (atoms, graphs) sizes and the CSR / CSC
        graph of the batch (built once per batch, like the reference's dataset does for edge lists)."""
        from .nn import functional as F
        from .nn.modules import GRAPH, SIZES
        batch = dict(batch)
        n_tot, n_b = int(batch[K_NATOMS].sum()), int(batch[K_NATOMS].numel())
        batch[SIZES] = (n_tot, n_b)
        batch[GRAPH] = F.build_graph(batch[K_EI], n_tot, n_tot, batch[K_EI].device)
        return batch

    def _step(self, batch):
        self.optimizer.zero_grad()
        out = self.forward(batch)
        loss = loss_from_config(out, self.config)
        loss.backward()
        self.optimizer.step()
        return loss.detach()

    def train_step(self, batch: Dict[str, torch.Tensor]) -> float:
        self.model.train()
        return float(self._step(batch))

    def train_step_graphed(self, batch: Dict[str, torch.Tensor]) -> torch.Tensor:
        """The same step replayed from a CUDA graph: `batch` must come from prepare() and keep its tensors (new data is
        copied INTO them).  The first call warms up and captures; returns the loss as a device tensor (no sync)."""
        self.model.train()
        if getattr(self, '_cg', None) is None or self._cg[1] is not batch:
            cur = torch.cuda.current_stream()
            side = torch.cuda.Stream()
            side.wait_stream(cur)
            with torch.cuda.stream(side):
                for _ in range(2):
                    self._step(batch)
            cur.wait_stream(side)
            g = torch.cuda.CUDAGraph()
            self.optimizer.zero_grad()
            with torch.cuda.graph(g):
                loss = self._step(batch)
            self._cg = (g, batch, loss)
        self._cg[0].replay()
        return self._cg[2]

    @torch.no_grad()
    def _metrics(self, out):
        n = out[K_NATOMS].double()
        dE = (out[P_E].double() - out[K_E].double()) / n
        dF = out[P_F].double() - out[K_F].double()
        dS = out[P_S].double() - out[K_S].double()
        return {'Energy_RMSE': float(torch.sqrt((dE ** 2).mean())), 'Force_RMSE': float(torch.sqrt((dF ** 2).sum(1).mean())),
                'Stress_RMSE': float(torch.sqrt((dS ** 2).sum(1).mean())) * TO_KB}

    def evaluate(self, batch: Dict[str, torch.Tensor]) -> Dict[str, float]:
        """error_recorder.py:155-175 metrics (Energy per atom, vector-force RMSE, stress RMSE in kbar), eval mode."""
        self.model.eval()
        out = self.forward(batch)
        return self._metrics({k: (v.detach() if torch.is_tensor(v) else v) for k, v in out.items()})

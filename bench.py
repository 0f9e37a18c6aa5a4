#!/usr/bin/env python
"""bench.py -- atom-steps/s of one SevenNet-0 energy+forces+stress evaluation (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload si1m|cfg1|cfg2|cfg3]

Default workload for EVERY N (north_star's target, BASELINE config 4): ONE 1,000,000-atom bulk-Si cell (50^3 diamond
cells, a = 5.431 A, sigma = 0.05 A rattle), SevenNet-0 hyper-parameters, random-init weights (seed 777), strong scaling:
N = 1 evaluates the periodic cell on one GPU, N > 1 cuts it into the LAMMPS brick grid (2x1x1, 2x2x1, 2x2x2) with
ghost-atom feature exchange per layer (domain.py).  A step = device neighbour list + forward + analytic reverse pass.
Other workloads (logged under profiles/, not the driver's line): cfg1 = 64-atom Si, cfg2 = 4,096-atom HfO2 (N > 1: weak
scaling, the cell replicated over the brick grid), cfg3 = SevenNet-l3i5 on 32,768 Li-P-S atoms.
--impl reference : the CPU restatement oracle (oracle/, kind "port": the reference itself needs e3nn/ase, which are
         not installable here) on the host cores, on a bounded sample of the same workload (same sample as the
         `cpu_baseline` block of the default line).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from sevennet_exp_b200 import presets  # noqa: E402
from tests.structures import fluorite_hfo2  # noqa: E402

METRIC = 'atom-steps/s (E+F+S) SevenNet-0'
UNIT = 'atom-steps/s'
A_LAT = 4.80


def workload(reps=(7, 7, 7), n_remove=20):
    numbers, pos, cell = fluorite_hfo2(reps, a=A_LAT, sigma=0.05, seed=0, n_remove=n_remove)
    return numbers, pos, cell


def model_config():
    cfg = dict(presets.PRESETS['7net-0'])
    cfg['version'] = '0.11.0'
    return cfg


def random_state_dict(cfg):
    """Random-init weights at the real hyper-parameters, keyed like the reference state_dict."""
    from sevennet_exp_b200.spec import build_spec
    spec = build_spec(cfg)
    g = torch.Generator().manual_seed(777)
    sd = {}
    for k, shp in spec.expected_keys().items():
        sd[k] = torch.randn(*shp, generator=g)
    nb = spec.n_basis
    sd['edge_embedding.basis_function.coeffs'] = torch.arange(1, nb + 1, dtype=torch.float32) * np.pi / spec.cutoff
    for L in spec.layers:
        sd[f'{L.conv.name}.denominator'] = torch.tensor([L.denominator])
    sd['rescale_atomic_energy.shift'] = torch.randn(spec.n_species, generator=g)
    sd['rescale_atomic_energy.scale'] = torch.ones(spec.n_species) * 1.7
    return sd


class ClockSampler(threading.Thread):
    """SM clock and throttle reasons DURING the timed region, every 20 ms through NVML (no process spawn, so the
    launching thread is not disturbed); falls back to polling `nvidia-smi` when pynvml is unusable."""

    NAMES = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']

    def __init__(self, index=0):
        super().__init__(daemon=True)
        self.rows, self.stop_flag, self.index = [], False, index
        self.nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get('CUDA_VISIBLE_DEVICES')
            phys = int(vis.split(',')[index]) if vis and vis.split(',')[index].strip().isdigit() else index
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.masks = [pynvml.nvmlClocksEventReasonHwSlowdown, pynvml.nvmlClocksEventReasonHwThermalSlowdown,
                          pynvml.nvmlClocksEventReasonSwThermalSlowdown, pynvml.nvmlClocksEventReasonSwPowerCap]
            self.nvml = pynvml
        except Exception:  # noqa: BLE001
            self.nvml = None

    def _sample_smi(self):
        q = ('clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
             'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')
        o = subprocess.run(['nvidia-smi', f'--query-gpu={q}', '--format=csv,noheader,nounits', '-i',
                            str(self.index)], capture_output=True, text=True, timeout=5).stdout.strip()
        t = [x.strip() for x in o.split(',')]
        self.max_mhz = float(t[1])
        return float(t[0]), [n for k, n in enumerate(self.NAMES) if t[2 + k] == 'Active']

    def run(self):
        while not self.stop_flag:
            try:
                if self.nvml is not None:
                    mhz = float(self.nvml.nvmlDeviceGetClockInfo(self.handle, self.nvml.NVML_CLOCK_SM))
                    bits = int(self.nvml.nvmlDeviceGetCurrentClocksEventReasons(self.handle))
                    self.rows.append((mhz, [n for n, m in zip(self.NAMES, self.masks) if bits & m]))
                else:
                    self.rows.append(self._sample_smi())
            except Exception:  # noqa: BLE001
                pass
            time.sleep(0.02 if self.nvml is not None else 0.1)

    def summary(self):
        if not self.rows:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['clock sampling unavailable']}
        sm = sorted(r[0] for r in self.rows)
        reasons = sorted({n for r in self.rows for n in r[1]})
        return {'sm_mhz': sm[len(sm) // 2], 'sm_max_mhz': self.max_mhz, 'reasons': reasons, 'samples': len(self.rows),
                'source': 'nvml' if self.nvml is not None else 'nvidia-smi'}


def measured_peak():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        return float(json.load(open(p))['hbm_gbs']), 'measured'
    return 6650.0, 'fallback'


def conv_bwd_algorithmic_bytes(conv, n_dst, n_edges):
    """each tensor the launch group must touch, once; gathers per edge."""
    S = (conv.lmax_sh + 1) ** 2
    per_edge = 4 * (2 * conv.wn + 2 * S + 1 + 2 * conv.lx.dim)      # w, dW, Y, dY, col, x_j gather, dx_e
    return n_edges * per_edge + n_dst * (4 * conv.lmid.dim + 4)


def conv_fwd_algorithmic_bytes(conv, n_dst, n_edges):
    S = (conv.lmax_sh + 1) ** 2
    return n_edges * 4 * (conv.wn + S + 1 + conv.lx.dim) + n_dst * (4 * conv.lmid.dim + 4)


def kdtree_graph(numbers, pos, cell, cutoff):
    """CPU-baseline graph build for an orthorhombic periodic cell with scipy's cKDTree (C code, like the matscipy /
    ase neighbour lists the reference calls, sevenn/train/dataload.py:32-88) -> the oracle's graph dict.  The
    brute-force O(N^2 x images) list of oracle/ is an edge-set checker, not a fair timing stand-in."""
    from scipy.spatial import cKDTree
    from oracle import sevenn_ref as S
    L = np.diag(cell)
    assert np.allclose(cell, np.diag(L)) and (L > 2 * cutoff).all()
    w = pos - np.floor(pos / L) * L
    w = np.where(w >= L, w - L, w)
    tree = cKDTree(w, boxsize=L)
    pairs = tree.query_pairs(cutoff, output_type='ndarray')
    i = np.concatenate([pairs[:, 0], pairs[:, 1]])
    j = np.concatenate([pairs[:, 1], pairs[:, 0]])
    d = w[j] - w[i]
    D = d - np.round(d / L) * L
    keep = (D ** 2).sum(1) < cutoff * cutoff
    i, j, D = i[keep], j[keep], D[keep]
    order = np.lexsort((j, i))
    i, j, D = i[order], j[order], D[order]
    sh = np.round((D - (pos[j] - pos[i])) / L)          # D = pos_j + S.cell - pos_i for the unwrapped positions
    return {S.X: torch.as_tensor(numbers, dtype=torch.int64), S.Z: torch.as_tensor(numbers, dtype=torch.int64),
            S.POS: torch.as_tensor(pos, dtype=torch.float32), S.EDGE_IDX: torch.as_tensor(np.stack([i, j])),
            S.EDGE_VEC: torch.as_tensor(D, dtype=torch.float32), S.CELL: torch.as_tensor(cell, dtype=torch.float32),
            S.SHIFT: torch.as_tensor(sh, dtype=torch.float32),
            S.VOL: torch.tensor(float(np.prod(L)), dtype=torch.float32),
            S.NATOMS: torch.tensor(len(numbers), dtype=torch.int64)}


def cpu_sample_structure(workload_name):
    """Bounded sample of the workload for the CPU legs: same lattice, density and model shape, >= 1,000 atoms (host
    memory permitting: the oracle keeps ~16 KB per edge and layer alive for autograd, like e3nn does)."""
    from tests.structures import diamond_si, jittered_lips
    try:
        import psutil
        avail = psutil.virtual_memory().available / 2 ** 30
    except Exception:  # noqa: BLE001
        avail = 32.0
    big = avail >= 96.0          # the SevenNet-0 oracle holds ~27 GB at 1,728 Si atoms
    if workload_name in ('si1m', 'cfg1'):
        r = 6 if big else 4
        numbers, pos, cell = diamond_si((r, r, r))
        return model_config(), numbers, pos, cell, f'bulk Si {r}x{r}x{r} diamond cells'
    if workload_name == 'cfg3':
        ns = 10 if big else 8
        numbers, pos, cell = jittered_lips(n_side=ns)
        return dict(presets.PRESETS['7net-l3i5']), numbers, pos, cell, f'disordered Li-P-S, {ns}^3 sites'
    r = 5 if big else 3
    numbers, pos, cell = fluorite_hfo2((r, r, r), a=A_LAT, sigma=0.05, seed=0)
    return model_config(), numbers, pos, cell, f'fluorite HfO2 {r}x{r}x{r} cells'


def cpu_oracle_leg(workload_name, steps, warmup, device='cpu'):
    """The CPU restatement (oracle/) timed on the host cores: `warmup` untimed + `steps` timed full evaluations (graph
    build + forward + autograd forces/stress), MEDIAN.  device='cuda' runs the same PyTorch program on the GPU (ATen
    library kernels: what the reference's e3nn path would launch) as an informational comparator."""
    from oracle import sevenn_ref as S
    cfg, numbers, pos, cell, what = cpu_sample_structure(workload_name)
    torch.manual_seed(777)
    model = S.build_model(cfg).eval()
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    ts = []
    if device != 'cpu':
        torch.set_default_device(device)
        model = model.to(device)
        if hasattr(model, 'z_to_onehot_tensor'):
            model.z_to_onehot_tensor = model.z_to_onehot_tensor.to(device)
    try:
        for k in range(warmup + steps):
            t0 = time.perf_counter()
            g = kdtree_graph(numbers, pos, cell, cfg['cutoff'])
            if device != 'cpu':
                g = {k_: v.to(device) for k_, v in g.items()}
            out = model(g)
            float(out[S.E_TOT].detach()); out[S.FORCE].sum().item(); out[S.STRESS].sum().item()
            if k >= warmup:
                ts.append(time.perf_counter() - t0)
            n_edges = int(g[S.EDGE_IDX].shape[1])
            del out, g
    finally:
        if device != 'cpu':
            torch.set_default_device('cpu')
    t = float(np.median(ts))
    n = len(numbers)
    sample = (f'oracle/ (CPU restatement of the reference path, fp32) on {n} atoms of {what} at the workload density '
              f'({n_edges / n:.1f} neighbours/atom): cKDTree graph + forward + autograd forces/stress, '
              f'{warmup} warm-up + {steps} timed evaluations, median {t:.2f} s')
    return {'value': n / t, 'unit': UNIT, 'cores': cores if device == 'cpu' else 0, 'kind': 'port', 'sample': sample,
            'n_atoms': n, 'n_edges': n_edges, 's_per_eval': t}


WORKLOAD_TEXT = {
    'si1m': 'SevenNet-0 (lmax2, 5 layers, 128x0e+64x1e+32x2e, rc=5A, random-init weights) E+F+S, ONE bulk-Si cell of '
            '{n} atoms ({c}^3 diamond cells, a=5.431A, rattle 0.05A), {nb:.1f} neighbours/atom',
    'cfg1': 'SevenNet-0 E+F+S, 64-atom bulk Si (2x2x2 diamond cells), {nb:.1f} neighbours/atom (BASELINE config 1)',
    'cfg2': 'SevenNet-0 (lmax2, 5 layers, 128x0e+64x1e+32x2e, rc=5A, random-init weights) E+F+S, {n}-atom bulk HfO2 '
            '(fluorite 7x7x7, a=4.8A, 20 vacancies), {nb:.1f} neighbours/atom (BASELINE config 2)',
    'cfg3': 'SevenNet-l3i5 (lmax3, 5 layers, 128x0e+64x1e+32x2e+32x3e, rc=5A, random-init weights) E+F+S, {n}-atom '
            'disordered Li-P-S cell (3:1:4, 0.048 atoms/A^3, hard core 1.95 A), {nb:.1f} neighbours/atom (BASELINE config 3)',
}


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    r = cpu_oracle_leg(args.workload, args.steps, args.warmup)
    line = {'impl': 'reference', 'metric': METRIC, 'value': r['value'], 'unit': UNIT, 'n_gpus': args.gpus,
            'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': r['s_per_eval'] * 1e3, 'higher_is_better': True,
            'scaling': 'weak' if args.workload == 'cfg2' else 'strong', 'vs_baseline': None, 'dtype': 'f32',
            'data': 'synthetic',
            'config': {'workload': WORKLOAD_TEXT[args.workload].split('{')[0] + f'... CPU sample: {r["n_atoms"]} atoms '
                                   '(per-atom cost of this path does not depend on the cell size)',
                       'n_atoms': r['n_atoms'], 'n_edges': r['n_edges']},
            'cpu_baseline': {k: r[k] for k in ('value', 'unit', 'cores', 'kind', 'sample')},
            'e2e': {'value': r['value'], 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}}
    if args.workload == 'cfg3':
        line['metric'] = 'atom-steps/s (E+F+S) SevenNet-l3i5'
    print(json.dumps(line), flush=True)


def make_workload(name, si_cells=50):
    """(model config, numbers, positions, cell) of a BASELINE configuration (SURVEY 8d's generators and seeds)."""
    from tests.structures import diamond_si, jittered_lips
    if name == 'si1m':
        nc = int(si_cells)
        numbers, pos, cell = diamond_si((nc, nc, nc))
        pos = pos + np.random.default_rng(2).normal(0.0, 0.05, pos.shape)
        return model_config(), numbers, pos, cell
    if name == 'cfg1':
        return (model_config(),) + tuple(diamond_si((2, 2, 2)))
    if name == 'cfg3':
        return (dict(presets.PRESETS['7net-l3i5']),) + tuple(jittered_lips())
    return (model_config(),) + tuple(workload())


def ncu_capture(kernel, n_edges):
    """dram bytes / tensor-pipe utilisation of ONE launch of `kernel` on this workload from the committed ncu --set full
    capture (profiles/ncu_captures.json: written from the raw CSVs under profiles/), or None."""
    try:
        for row in json.load(open(os.path.join(ROOT, 'profiles', 'ncu_captures.json'))):
            if row['kernel'] == kernel and abs(row['n_edges'] - n_edges) <= 0.001 * n_edges:
                return row
            if n_edges == 0 and row['kernel'].startswith(kernel) and row.get('tensor_pipe_pct'):
                return row
    except Exception:  # noqa: BLE001
        pass
    return None


def gemm_roofline(spec, stage, n_atoms, peak_hbm):
    """north_star asks for tensor-pipe utilisation per GEMM: the dominant node linear is self_interaction_2 (mid ->
    gate input, per-l blocks, accumulating onto the self connection).  A + C bytes and 3xTF32 MMA flops of all layers
    over the CUDA-event time of the `si2_fwd` stage."""
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
    except Exception:  # noqa: BLE001
        pass
    tf32_peak = float(peaks.get('bf16_tflops_sustained', 1400.0)) / 2.0      # dense tf32 = half the bf16 rate
    bytes_, flops = 0.0, 0.0
    for L in spec.layers:
        bytes_ += n_atoms * 4.0 * (L.conv.lmid.dim + 2 * L.gate.lin.dim)          # read A, read + write C
        flops += sum(2.0 * n_atoms * (2 * b.l + 1) * b.K * b.N for b in L.si2.blocks)
    ms = stage.get('si2_fwd')
    if not ms:
        return None
    cap = ncu_capture('gemm_tc', 0)          # first captured launch: the TMEM-A kernel on the l = 1 block of si2
    return {'bound': 'hbm', 'kernel': 'gemm_tc_kernel (tcgen05 3xTF32), self_interaction_2 forward of all layers',
            'achieved': bytes_ / (ms * 1e-3) / 1e9, 'peak': peak_hbm, 'unit': 'GB/s',
            'frac': bytes_ / (ms * 1e-3) / 1e9 / peak_hbm, 'algorithmic_bytes': bytes_, 'ms': ms,
            'tensor': {'tflops_tf32_mma': 3.0 * flops / (ms * 1e-3) / 1e12, 'peak_tf32': tf32_peak,
                       'frac': 3.0 * flops / (ms * 1e-3) / 1e12 / tf32_peak,
                       'note': '3 tf32 MMAs per fp32 product (hi*hi + hi*lo + lo*hi); K <= 352 per block: these '
                               'GEMMs are bound by streaming A and C, not by the tensor pipe',
                       'ncu_tensor_pipe_pct': cap.get('tensor_pipe_pct') if cap else None,
                       'ncu_source': cap.get('source') if cap else None}}


def cfg5_structures(n_struct=32, seed=5):
    """BASELINE config 5: 32 MPtrj-shaped structures, 64-256 atoms each, random species out of the model's 89,
    density 0.03-0.08 atoms/A^3 (jittered simple-cubic sites: no overlapping atoms), labels E, F, S ~ N(0, 1)."""
    rng = np.random.default_rng(seed)
    zs = np.array(sorted(presets.PRESETS['7net-0']['_type_map']))
    out = []
    for _ in range(n_struct):
        n = int(rng.integers(64, 257))
        rho = rng.uniform(0.03, 0.08)
        side = int(np.ceil(n ** (1 / 3)))
        a = (1.0 / rho) ** (1 / 3) * (n / side ** 3) ** (1 / 3)
        sites = np.stack(np.meshgrid(*[np.arange(side)] * 3, indexing='ij'), -1).reshape(-1, 3)
        sites = sites[rng.permutation(len(sites))[:n]].astype(np.float64)
        pos = sites * a + rng.uniform(-0.12 * a, 0.12 * a, (n, 3))
        cell = np.eye(3) * side * a
        numbers = rng.choice(zs, n)
        out.append((numbers, pos, cell, rng.normal(), rng.normal(size=(n, 3)), rng.normal(size=6) * 1e-2))
    return out


def run_train(args):
    """--workload cfg5: one training step (forward, force/stress loss, backward through the reverse kernels, Adam) on a
    batch of 32 structures; graphs are built beforehand, as the reference's dataset does (train/dataload.py)."""
    from sevennet_exp_b200 import nn as bnn, train as TR
    from sevennet_exp_b200.neighbor import NeighborList
    dev = torch.device('cuda:0')
    torch.cuda.set_device(dev)
    cfg = model_config()
    model = bnn.build_E3_equivariant_model(cfg)
    sd = random_state_dict(cfg)
    missing, _ = model.load_state_dict(sd, strict=False)
    assert not missing, missing
    model = model.to(dev)
    structs = cfg5_structures()
    nl = NeighborList(device=dev)
    tmap = cfg['_type_map']
    graphs = []
    for numbers, pos, cell, e, f, st in structs:
        r = nl.build(pos, cell, [True] * 3, cfg['cutoff'])
        graphs.append({TR.K_X: torch.tensor([tmap[int(z)] for z in numbers]), TR.K_Z: torch.tensor(numbers),
                       TR.K_POS: torch.tensor(pos, dtype=torch.float32),
                       TR.K_EI: torch.stack([r['center'], r['col']]).long().cpu(), TR.K_EV: r['edge_vec'].cpu(),
                       TR.K_VOL: torch.tensor(r['volume'], dtype=torch.float32), TR.K_NATOMS: torch.tensor(len(numbers)),
                       TR.K_E: torch.tensor(e, dtype=torch.float32), TR.K_F: torch.tensor(f, dtype=torch.float32),
                       TR.K_S: torch.tensor(st, dtype=torch.float32)})
    for g in graphs:
        g[TR.K_X] = g[TR.K_Z]            # the model maps atomic numbers to type indices itself (sequential.py:156-168)
    config = {'force_loss_weight': 1.0, 'stress_loss_weight': 0.1, 'is_train_stress': True, 'optim_param': {}}
    trainer = TR.Trainer(model, config)
    host = {k: v.pin_memory() for k, v in TR.collate(graphs, device='cpu').items()}
    batch = {k: v.to(dev) for k, v in host.items()}
    n_atoms, n_edges = int(batch[TR.K_NATOMS].sum()), int(batch[TR.K_EI].shape[1])
    from sevennet_exp_b200.nn._backend import kernels
    kern = kernels(dev)
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)
    batch = trainer.prepare(batch)             # sizes + CSR / CSC of the batch: once per batch, like the dataset's edge lists
    loss0 = trainer.train_step(batch)
    for _ in range(args.warmup):
        trainer.train_step_graphed(batch)      # first call warms up and captures the step into a CUDA graph
    torch.cuda.synchronize()
    sampler = ClockSampler()
    sampler.start()
    l0, evs = kern.n_launch, []
    per_step = getattr(trainer, '_launches_per_step', None)
    for _ in range(args.steps):
        flush.fill_(1.0)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        loss_t = trainer.train_step_graphed(batch)
        b.record()
        evs.append((a, b))
    torch.cuda.synchronize()
    ms_step = sum(a.elapsed_time(b) for a, b in evs) / args.steps
    # launches of one (eager) step x timed steps: the graph replays exactly those
    n0 = kern.n_launch
    trainer.train_step(batch)
    launches = (kern.n_launch - n0) * args.steps
    tensor_keys = [k for k, v in host.items() if torch.is_tensor(v)]
    t0 = time.perf_counter()
    for _ in range(args.steps):
        flush.fill_(1.0)
        for k in tensor_keys:                   # the step's inputs from pinned host memory into the static batch
            batch[k].copy_(host[k], non_blocking=True)
        loss = float(trainer.train_step_graphed(batch))          # device -> host read of the loss
    torch.cuda.synchronize()
    e2e_t = (time.perf_counter() - t0) / args.steps
    sampler.stop_flag = True
    sampler.join(timeout=2)
    h2d = int(sum(v.numel() * v.element_size() for v in host.values()))
    cpu = None
    if not os.environ.get('SEVENN_BENCH_NO_CPU'):
        cpu = cpu_train_leg(cfg, graphs[:2], config, 3, 1)
    peak, how = measured_peak()
    line = {
        'metric': 'atom-steps/s (training step: forward + force/stress loss backward + Adam) SevenNet-0',
        'value': n_atoms / (ms_step * 1e-3), 'unit': UNIT, 'n_gpus': 1, 'steps': args.steps, 'warmup': args.warmup,
        'ms_per_step': ms_step, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32',
        'data': 'synthetic',
        'config': {'workload': f'SevenNet-0 training step, batch of {len(graphs)} synthetic MPtrj-shaped structures '
                               f'(64-256 atoms, 89 species, 0.03-0.08 atoms/A^3): {n_atoms} atoms, {n_edges} edges '
                               '(BASELINE config 5)', 'n_atoms': n_atoms, 'n_edges': n_edges, 'global_batch': len(graphs),
                   'step': 'forward + autograd.grad(E, edge_vec, create_graph=True) + MSE(E/N, F, S) backward + Adam, replayed '
                           'from a CUDA graph; edge lists prebuilt (dataset), as in the reference',
                   'final_loss': loss, 'first_loss': loss0,
                   'l2': '256 MiB write between timed steps (outside the CUDA-event brackets)'},
        'e2e': {'value': n_atoms / e2e_t, 'unit': UNIT, 'ms_per_step': e2e_t * 1e3, 'h2d_bytes_per_step': h2d,
                'd2h_bytes_per_step': 4},
        'gpu_launches': int(launches), 'clocks': sampler.summary(),
        'roofline': {'bound': 'hbm', 'kernel': 'whole step (second order composed from ~6 first-order conv launches per layer)',
                     'achieved': None, 'peak': peak, 'peak_source': how, 'unit': 'GB/s', 'frac': None, 'traffic': None,
                     'note': 'no per-kernel roofline for the training step yet: launch-bound at this batch size'},
        'cpu_baseline': cpu,
    }
    print(json.dumps(line), flush=True)


def cpu_train_leg(cfg, graphs, config, steps, warmup):
    """The oracle (CPU restatement) doing the same training step under torch autograd on a bounded sample."""
    from oracle import sevenn_ref as S
    from sevennet_exp_b200 import train as TR
    torch.manual_seed(777)
    om = S.build_model(cfg).train()
    om.set_is_batch_data(True)
    opt = torch.optim.Adam([p for p in om.parameters() if p.requires_grad])
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    ts = []
    n = int(sum(int(g[TR.K_NATOMS]) for g in graphs))
    for k in range(warmup + steps):
        t0 = time.perf_counter()
        opt.zero_grad()
        out = om(TR.collate(graphs, device='cpu'))
        loss = TR.loss_from_config(out, config)
        loss.backward()
        opt.step()
        if k >= warmup:
            ts.append(time.perf_counter() - t0)
    t = float(np.median(ts))
    return {'value': n / t, 'unit': UNIT, 'cores': cores, 'kind': 'port',
            'sample': f'oracle/ training step (autograd double backward + torch Adam) on {len(graphs)} of the structures '
                      f'({n} atoms), {warmup} warm-up + {steps} timed, median {t:.2f} s'}


def run_single(args):
    from sevennet_exp_b200.sevennet_calculator import SevenNetCalculator
    dev = torch.device('cuda:0')
    torch.cuda.set_device(dev)
    wl = args.workload
    cfg, numbers, pos, cell = make_workload(wl, args.si_cells)
    sd = random_state_dict(cfg)
    calc = SevenNetCalculator((cfg, sd), file_type='model_instance', device=dev, keep_weights=False)
    n = len(numbers)
    pbc = [True] * 3
    types_dev = torch.from_numpy(calc._lut[numbers].astype(np.int32)).to(dev)
    pos_dev = torch.from_numpy(pos).to(dev)
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)

    native = calc.native           # C++ whole-model host (csrc/model.cu): the product path of SevenNetCalculator

    def n_launches():
        return native.n_launch if native is not None else calc.engine.n_launch + calc.nl.n_launch

    def step():
        if native is not None:
            return None, native.evaluate_graphed(types_dev, pos_dev, cell, pbc)
        g = calc.nl.graph(types_dev, None, cell, pbc, calc.cutoff, pos_dev=pos_dev)
        return g, calc.engine.evaluate(g)

    for _ in range(args.warmup):
        g, out = step()
        del out
    torch.cuda.synchronize()
    sampler = ClockSampler()
    sampler.start()
    l0 = n_launches()
    evs = []
    torch.cuda.synchronize()
    for _ in range(args.steps):
        flush.fill_(1.0)                       # L2 flush, outside the timed events
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        g, out = step()
        b.record()
        evs.append((a, b))
    torch.cuda.synchronize()
    total_ms = sum(a.elapsed_time(b) for a, b in evs)
    launches = n_launches() - l0
    ms_step = total_ms / args.steps
    value = n / (ms_step * 1e-3)
    if native is not None:          # the status word of the last step: the edge capacity held
        _, _, overflow, found = native.parse_header(out['header'].cpu().numpy())
        assert not overflow, 'edge capacity overflow during the timed steps'
    del out

    # end to end through the public calculator call: host arrays in, host E/F/S out
    for _ in range(2):
        calc.compute(numbers, pos, cell, pbc)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        flush.fill_(1.0)
        res = calc.compute(numbers, pos, cell, pbc)
    torch.cuda.synchronize()
    e2e_t = (time.perf_counter() - t0) / args.steps
    sampler.stop_flag = True
    sampler.join(timeout=2)

    # per-stage device times (separate pass through the per-launch Python driver: the same kernels in the same order,
    # CUDA events between stages on the launch stream)
    stage = {}
    reps_prof = 3
    if native is not None:
        native._ws = None
        torch.cuda.empty_cache()
    g = calc.nl.graph(types_dev, None, cell, pbc, calc.cutoff, pos_dev=pos_dev)
    calc.engine.evaluate(g)          # untimed: the Python driver builds its tcgen05 weight images on first use
    torch.cuda.synchronize()
    for _ in range(reps_prof):
        flush.fill_(1.0)
        # head start for the host: the GPU spins ~10 ms first, so the per-launch Python driver stays ahead of it and the
        # events between stages bracket kernel time, not launch gaps (matters for the small workloads)
        torch.cuda._sleep(20_000_000)
        calc.engine.prof = []
        calc.engine.evaluate(g)
        torch.cuda.synchronize()
        for k, v in calc.engine.profile_summary().items():
            stage[k] = stage.get(k, 0.0) + v / reps_prof
    calc.engine.prof = None
    inner = calc.spec.layers[1].conv
    key_b, key_f = f'conv_bwd_{inner.signature()}', f'conv_fwd_{inner.signature()}'
    n_inner = sum(1 for L in calc.spec.layers if L.conv.signature() == inner.signature())
    peak, how = measured_peak()
    E = g.n_edges
    dom_ms = stage[key_b] / n_inner
    alg = conv_bwd_algorithmic_bytes(inner, n, E)
    achieved = alg / (dom_ms * 1e-3) / 1e9
    fwd_ach = conv_fwd_algorithmic_bytes(inner, n, E) / (stage[key_f] / n_inner * 1e-3) / 1e9
    stage_sorted = dict(sorted(((k, round(v, 4)) for k, v in stage.items()), key=lambda kv: -kv[1]))
    tab = calc.engine.radial == 'table'
    kname = f'conv_bwd_{"tab_" if tab else ""}{inner.signature()}'
    cap = ncu_capture(kname, E)
    # bytes the launch really has to move through HBM: the tabulated kernels neither read w nor write dW
    # ("say which tensors were elided"); `achieved` keeps SURVEY 8d's figure so rounds stay comparable
    touched = alg - (E * 8 * inner.wn if tab else 0)
    # SURVEY 8d whole-path algorithmic bytes per atom (forward + reverse, 5 layers): 3.16 MB at 28 neighbours (Si),
    # 5.39 MB at ~50 (HfO2), l3i5 5.06 MB at 26
    path_bytes_per_atom = {'si1m': 3.16e6, 'cfg1': 3.16e6, 'cfg2': 5.39e6, 'cfg3': 5.06e6}[wl]

    cpu = gpu_torch = None
    if not os.environ.get('SEVENN_BENCH_NO_CPU'):      # tuning runs skip the CPU legs
        del g
        torch.cuda.empty_cache()
        cpu = cpu_oracle_leg(wl, 5, 3)
        try:
            gpu_torch = cpu_oracle_leg(wl, 5, 3, device='cuda')
            gpu_torch['kind'] = 'oracle port on cuda:0 (ATen library kernels, autograd): informational'
            gpu_torch.pop('cores')
        except Exception as e:  # noqa: BLE001
            gpu_torch = {'unavailable': f'{type(e).__name__}: {e}'[:200]}
    line = {
        'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': 1, 'steps': args.steps, 'warmup': args.warmup,
        'ms_per_step': ms_step, 'higher_is_better': True, 'scaling': 'weak' if wl == 'cfg2' else 'strong',
        'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
        'config': {'workload': WORKLOAD_TEXT[wl].format(n=n, c=args.si_cells, nb=E / n),
                   'n_atoms': n, 'n_edges': E, 'step': 'device neighbour list + forward + analytic reverse pass',
                   'host': 'C++ (sevenn_model_evaluate_positions: no host sync, the repeated step replayed from a CUDA graph)' if native is not None else 'python per-launch driver',
                   'radial': calc.engine.radial, 'parallelism': 'single GPU, whole periodic cell',
                   'energy_per_atom_eV': res['energy'] / n,
                   'l2': '256 MiB write between timed steps (outside the CUDA-event brackets)'},
        'e2e': {'value': n / e2e_t, 'unit': UNIT, 'ms_per_step': e2e_t * 1e3,
                'h2d_bytes_per_step': int(pos.nbytes + n * 4), 'd2h_bytes_per_step': int((1 + 6 + 3 * n + n) * 8)},
        'gpu_launches': int(launches),
        'clocks': sampler.summary(),
        'roofline': {'bound': 'hbm',
                     'kernel': f'{kname} (inner layers, one launch per layer)',
                     'achieved': achieved, 'peak': peak, 'peak_source': how, 'unit': 'GB/s', 'frac': achieved / peak,
                     'algorithmic_bytes': alg, 'ms': dom_ms,
                     'traffic': cap['dram_bytes'] if cap else None,
                     'traffic_source': cap.get('source') if cap else None,
                     'elided': ('w and dW [E, wn] (radial weights interpolated from an L2-resident table inside the '
                                'kernel)' if tab else None),
                     'touched_bytes': touched, 'touched_frac': touched / (dom_ms * 1e-3) / 1e9 / peak,
                     'limiter': cap.get('limiter') if cap else None,
                     'conv_fwd_achieved': fwd_ach, 'conv_fwd_frac': fwd_ach / peak,
                     'path': {'bytes_per_atom': path_bytes_per_atom,
                              'ceiling_atom_steps_s': peak * 1e9 / path_bytes_per_atom,
                              'frac': value / (peak * 1e9 / path_bytes_per_atom)}},
        'roofline_gemm': gemm_roofline(calc.spec, stage, n, peak),
        'stage_ms': stage_sorted,
        'cpu_baseline': None if cpu is None else {k: cpu[k] for k in ('value', 'unit', 'cores', 'kind', 'sample')},
        'torch_gpu_baseline': gpu_torch,
    }
    if wl == 'cfg3':
        line['metric'] = 'atom-steps/s (E+F+S) SevenNet-l3i5'
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=10)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours')
    ap.add_argument('--workload', default='si1m', choices=['si1m', 'cfg1', 'cfg2', 'cfg3', 'cfg5'],
                    help="'si1m' (default, every N: BASELINE config 4, one 1M-atom bulk-Si cell, strong scaling); "
                         "'cfg1' (64-atom Si); 'cfg2' (4,096-atom HfO2; N > 1: weak scaling over the brick grid); "
                         "'cfg3' (N = 1: SevenNet-l3i5 on 32,768 Li-P-S atoms); 'cfg5' (N = 1: training step on 32 structures)")
    ap.add_argument('--si-cells', type=int, default=50)
    args = ap.parse_args()
    if args.impl == 'reference':
        return run_reference(args)
    args.warmup = max(args.warmup, 3)
    if args.workload == 'cfg5':
        return run_train(args)
    if args.gpus > 1 or int(os.environ.get('WORLD_SIZE', '1')) > 1:
        from sevennet_exp_b200 import bench_multi
        return bench_multi.run(args)
    run_single(args)


if __name__ == '__main__':
    main()

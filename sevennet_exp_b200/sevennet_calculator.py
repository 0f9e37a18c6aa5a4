"""ASE-style entry point: mirrors sevenn/sevennet_calculator.py:22-202.

``SevenNetCalculator(model, file_type, device, ...)`` keeps the reference's constructor and fills
``results`` with energy, free_energy, energies, forces, stress (ASE Voigt order, same sign
convention: ``-inferred_stress[[0,1,2,4,5,3]]``).  The neighbour list, the model and the reverse
pass all run on the GPU through the C ABI; only E/F/S come back to the host.  ``ase`` itself is
optional: without it the class still works on any object with ``get_positions/get_cell/get_pbc/
get_atomic_numbers`` (or through :meth:`compute` with plain arrays)."""
import os
from typing import Dict, Optional, Union

import numpy as np
import torch

from . import presets
from .checkpoint import load_checkpoint, patch_old_config, patch_state_dict
from .engine import Engine
from .neighbor import NeighborList
from .spec import build_spec

try:  # pragma: no cover - ase is not part of this image
    from ase.calculators.calculator import Calculator as _AseCalculator, all_changes
except Exception:  # noqa: BLE001
    all_changes = ['positions', 'numbers', 'cell', 'pbc']

    class _AseCalculator:  # minimal stand-in with the attributes SevenNetCalculator relies on
        implemented_properties = []

        def __init__(self, **kwargs):
            self.results, self.atoms = {}, None

        def calculate(self, atoms=None, properties=None, system_changes=all_changes):
            self.atoms = atoms

        def get_potential_energy(self, atoms=None):
            self.calculate(atoms)
            return self.results['energy']

        def get_forces(self, atoms=None):
            self.calculate(atoms)
            return self.results['forces']

        def get_stress(self, atoms=None):
            self.calculate(atoms)
            return self.results['stress']


class SevenNetCalculator(_AseCalculator):
    implemented_properties = ['free_energy', 'energy', 'forces', 'stress', 'energies']

    def __init__(self, model: Union[str, tuple] = '7net-0', file_type: str = 'checkpoint',
                 device: Union[torch.device, str] = 'auto', modal: Optional[str] = None,
                 enable_cueq: bool = False, sevennet_config: Optional[Dict] = None, keep_weights: bool = True,
                 **kwargs):
        # keep_weights=False: the per-edge tensor-product weights (the largest activation, 3.84 KB per edge and layer
        # for SevenNet-0) are recomputed in the reverse pass instead of being kept: ~3x more atoms per GPU
        super().__init__(**kwargs)
        if enable_cueq:
            raise ValueError('enable_cueq: the B200 backend replaces cuEquivariance, not combines with it')
        file_type = file_type.lower()
        if file_type not in ('checkpoint', 'model_instance', 'torchscript'):
            raise ValueError('file_type should be checkpoint, model_instance or torchscript')
        self._ts_model = None
        if file_type == 'torchscript':
            # the reference does torch.jit.load here and calls the module on the graph dict
            # (sevennet_calculator.py:103-125, 176-190); so do we -- the file is the TorchScript archive written by
            # sevennet_exp_b200.deploy.deploy(), whose forward runs torch.ops.sevenn_b200.evaluate (a file frozen by
            # the reference itself raises ValueError: it inlines the e3nn graph)
            from . import deploy as _deploy
            if not os.path.isfile(str(model)):
                raise FileNotFoundError(str(model))
            b, is_ts = _deploy._read_bundle(str(model))
            if not (isinstance(b, dict) and b.get('format') == _deploy.SERIAL_FORMAT):
                raise ValueError(f'{model}: not a serial model deployed by sevennet_exp_b200.deploy.deploy()')
            if is_ts:
                self._ts_model, self._ts_extra = _deploy.load_deployed(str(model))
            model, file_type = (b['config'], b['model_state_dict']), 'model_instance'
            modal = b.get('modal', modal)
        if device == 'auto':
            device = 'cuda'
        self.device = torch.device(device)
        if self.device.type != 'cuda':
            raise RuntimeError('sevennet_exp_b200 runs on CUDA devices only (no CPU fallback)')
        if file_type == 'checkpoint':
            if isinstance(model, str) and os.path.isfile(model):
                spec, sd, cfg = load_checkpoint(model)
            elif isinstance(model, str) and model in presets.PRESETS:
                raise FileNotFoundError(
                    f'pretrained weights of {model} are not shipped here (reference blobs are absent); '
                    'pass a checkpoint path or file_type="model_instance" with (config, state_dict)')
            else:
                raise FileNotFoundError(str(model))
        else:
            if isinstance(model, torch.nn.Module):
                # reference sevennet_calculator.py:111-124: an AtomGraphSequential instance.  Ours carries the config it
                # was built from (nn.build_E3_equivariant_model stores it), so the whole-model engine takes its weights
                cfg = getattr(model, '_sevenn_config', None)
                if cfg is None:
                    raise ValueError('model_instance: the module has no `_sevenn_config` (build it with '
                                     'sevennet_exp_b200.nn.build_E3_equivariant_model, or pass (config, state_dict))')
                model = (dict(cfg, version='0.11.0'), {k: v.detach().cpu() for k, v in model.state_dict().items()})
            cfg, sd = model
            cfg = patch_old_config(cfg, require_version=False)
            spec = build_spec(cfg)
            sd = patch_state_dict(spec, sd, cfg.get('version', '0.11.0'))
        self.sevennet_config = sevennet_config or cfg
        self.spec, self.cutoff = spec, spec.cutoff
        self.type_map = spec.type_map
        # `modal` (sevennet_calculator.py:36-45,84-95): modality name of a multi-modal checkpoint; the engine folds the
        # fixed one-hot into constant rows of the scalar blocks and picks the modal-wise shift / scale
        self.modal = modal
        self.engine = Engine(spec, sd, device=self.device, keep_weights=keep_weights, modal=modal)
        self.nl = NeighborList(device=self.device)
        # C++ whole-model host (csrc/model.cu): one call per step, no per-kernel Python, no host sync before the result
        # copy.  SEVENN_HOST=python keeps the per-launch Python driver (engine.py), which also serves the variants the
        # C++ host does not run (radial='mlp', readout_as_fcn).
        self.native = None
        from . import host
        if os.environ.get('SEVENN_HOST', 'native') != 'python' and host.supported(self.engine) is None:
            self.native = host.NativeModel(self.engine)
        lut = -np.ones(120, dtype=np.int64)
        for z, i in self.type_map.items():
            lut[int(z)] = i
        self._lut = lut

    @property
    def n_tc_launch(self):
        """launches on the tcgen05 GEMM arm so far (C++ host + per-launch driver)"""
        return self.engine.n_tc_launch + (self.native.n_tc_launch if self.native is not None else 0)

    @property
    def n_launch(self):
        return self.engine.n_launch + self.nl.n_launch + (self.native.n_launch if self.native is not None else 0)

    # ---- plain-array API (also what bench.py times end to end)
    def compute(self, numbers, positions, cell, pbc):
        numbers = np.asarray(numbers)
        # species rarely change between calls (MD): the type lookup of the previous call is reused when they did not
        last = getattr(self, '_last_types', None)
        if last is not None and last[0].shape == numbers.shape and np.array_equal(last[0], numbers):
            types = last[1]
        else:
            types = self._lut[numbers]
            if (types < 0).any():
                bad = sorted({int(z) for z in numbers[types < 0]})
                raise ValueError(f'Unknown atomic numbers {bad} for this model')   # sevennet_calculator.py:153-161
            types = types.astype(np.int32)
            self._last_types = (numbers.copy(), types)
        pos = np.ascontiguousarray(np.asarray(positions, dtype=np.float64))
        n = len(numbers)
        if self._ts_model is not None and n > 0:
            return self._compute_torchscript(types, pos, cell, pbc)
        if self.native is not None and n > 0:
            if self.native.lib.cdll.sevenn_model_set_tc_min_rows(self.native.handle, int(self.engine.tc_min_rows)):
                raise RuntimeError('sevenn_model_set_tc_min_rows failed')
            r = self.native.evaluate_host(types, pos, cell, pbc)
            vol = abs(float(np.linalg.det(np.asarray(cell, dtype=np.float64).reshape(3, 3))))
            vol = float(np.finfo(float).eps) if vol == 0 else vol
            stress_model = -r['virial6'] / vol
            # fp32 arrays like the reference's (`.detach().cpu().numpy()` of fp32 tensors, sevennet_calculator.py:186-201);
            # copies, because the host's pinned result buffer is reused by the next call
            return {'energy': r['energy'], 'free_energy': r['energy'], 'forces': r['forces'].copy(),
                    'stress': -stress_model[[0, 1, 2, 4, 5, 3]], 'energies': r['atomic_energy'].copy(),
                    'inferred_stress': stress_model, 'n_edges': r['n_edges']}
        types_dev = torch.from_numpy(types.astype(np.int32)).to(self.device, non_blocking=True)
        g = self.nl.graph(types_dev, pos, cell, pbc, self.cutoff)
        out = self.engine.evaluate(g)
        host = torch.cat([out['energy'].reshape(1), out['virial6'],
                          out['forces'][:n].reshape(-1).double(), out['atomic_energy'].double()]).cpu().numpy()
        energy, vir = float(host[0]), host[1:7]
        forces = host[7:7 + 3 * n].reshape(n, 3)
        energies = host[7 + 3 * n:]
        stress_model = -vir / g.volume                      # (xx,yy,zz,xy,yz,zx), force_output.py:194-222
        stress = -stress_model[[0, 1, 2, 4, 5, 3]]           # ASE Voigt, sevennet_calculator.py:196-201
        return {'energy': energy, 'free_energy': energy, 'forces': forces, 'stress': stress,
                'energies': energies, 'inferred_stress': stress_model, 'n_edges': g.n_edges}

    def _compute_torchscript(self, types, pos, cell, pbc):
        """file_type='torchscript': graph on the GPU, then the loaded ScriptModule with the deployed-model dict
        (reference sevennet_calculator.py:176-202)."""
        dev = self.device
        r = self.nl.build(pos, cell, pbc, self.cutoff)
        cell_t = torch.as_tensor(np.asarray(cell, dtype=np.float64).reshape(3, 3), dtype=torch.float32, device=dev)
        data = {'x': torch.as_tensor(types, dtype=torch.int64, device=dev),
                'pos': torch.as_tensor(pos, dtype=torch.float32, device=dev),
                'edge_index': torch.stack([r['center'], r['col']]).to(torch.int64),
                'num_atoms': torch.tensor([len(types)], device=dev),
                'cell_lattice_vectors': cell_t,
                'cell_volume': torch.tensor(r['volume'], dtype=torch.float32, device=dev),
                'pbc_shift': r['shift'].to(torch.float32)}
        out = self._ts_model(data)
        energy = float(out['inferred_total_energy'].item())
        stress_model = out['inferred_stress'].detach().cpu().numpy().astype(np.float64)
        return {'energy': energy, 'free_energy': energy,
                'forces': out['inferred_force'].detach().cpu().numpy().astype(np.float64),
                'stress': -stress_model[[0, 1, 2, 4, 5, 3]],
                'energies': out['atomic_energy'].detach().cpu().numpy().reshape(-1).astype(np.float64),
                'inferred_stress': stress_model, 'n_edges': r['n_edges']}

    def compute_batch(self, structures):
        """Several structures in one evaluation (the reference batches graphs with torch_geometric's Collater,
        sevenn/train/dataload.py + nn/force_output.py:210-219 per-graph stress): the graphs are concatenated
        block-diagonally (node indices offset per graph), the kernels run once over the merged edge list, and
        energy / virial are reduced per graph.  ``structures`` = iterable of (numbers, positions, cell, pbc)."""
        from .engine import Graph
        from .neighbor import csc_build
        parts, types_all, n_off, e_off = [], [], [0], [0]
        for numbers, positions, cell, pbc in structures:
            numbers = np.asarray(numbers)
            types = self._lut[numbers]
            if (types < 0).any():
                bad = sorted({int(z) for z in numbers[types < 0]})
                raise ValueError(f'Unknown atomic numbers {bad} for this model')
            pos = np.ascontiguousarray(np.asarray(positions, dtype=np.float64))
            r = self.nl.build(pos, cell, pbc, self.cutoff)
            parts.append(r)
            types_all.append(types.astype(np.int32))
            n_off.append(n_off[-1] + len(numbers))
            e_off.append(e_off[-1] + r['n_edges'])
        nB, n_tot = len(parts), n_off[-1]
        if nB == 0:
            return []
        dev = self.device
        types_dev = torch.from_numpy(np.concatenate(types_all)).to(dev)
        col = torch.cat([r['col'] + n_off[b] for b, r in enumerate(parts)]).to(torch.int32).contiguous()
        rowptr = torch.cat([parts[0]['rowptr'][:1]] + [r['rowptr'][1:] + e_off[b] for b, r in enumerate(parts)]
                           ).to(torch.int32).contiguous()
        edge_vec = torch.cat([r['edge_vec'] for r in parts]).contiguous()
        cptr, cedge = csc_build(self.nl.lib, col, n_tot, dev, self.nl._stream())
        g = Graph(types_dev, edge_vec, rowptr, col, cptr, cedge, n_tot, 1.0)
        out = self.engine.evaluate(g)
        seg = torch.repeat_interleave(torch.arange(nB, device=dev),
                                      torch.tensor(np.diff(n_off), device=dev))
        e_g = torch.zeros(nB, dtype=torch.float64, device=dev).index_add_(0, seg, out['atomic_energy'].double())
        v_g = torch.zeros(nB, 6, dtype=torch.float64, device=dev).index_add_(0, seg, out['virial_per_atom'].double())
        host_e, host_v = e_g.cpu().numpy(), v_g.cpu().numpy()
        F = out['forces'].double().cpu().numpy()
        ae = out['atomic_energy'].double().cpu().numpy()
        res = []
        for b, r in enumerate(parts):
            stress_model = -host_v[b] / r['volume']
            res.append({'energy': float(host_e[b]), 'free_energy': float(host_e[b]),
                        'forces': F[n_off[b]:n_off[b + 1]], 'stress': -stress_model[[0, 1, 2, 4, 5, 3]],
                        'energies': ae[n_off[b]:n_off[b + 1]], 'inferred_stress': stress_model,
                        'n_edges': r['n_edges']})
        return res

    def set_atoms(self, atoms):
        """sevennet_calculator.py:153-161: ASE calls this on `atoms.calc = calc`; unknown species fail early."""
        for z in sorted(set(int(z) for z in atoms.get_atomic_numbers())):
            if z not in self.type_map:
                raise ValueError(f'Model do not know atomic number: {z}, (knows: {list(self.type_map.keys())})')

    def calculate(self, atoms=None, properties=None, system_changes=all_changes):
        super().calculate(atoms, properties, system_changes)
        if atoms is None:
            raise ValueError('No atoms to evaluate')
        res = self.compute(atoms.get_atomic_numbers(), atoms.get_positions(), np.asarray(atoms.get_cell()),
                           atoms.get_pbc())
        self.results = {k: res[k] for k in ('free_energy', 'energy', 'energies', 'forces', 'stress')}

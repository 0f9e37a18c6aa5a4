"""Deployed-model entry points: what LAMMPS `pair_style e3gnn` / `e3gnn/parallel` load and call.

The reference (sevenn/scripts/deploy.py:20-148) TorchScript-compiles the model and freezes it, so a deployed file
is e3nn's generated graph; there is nothing in it a kernel could be swapped into.  The B200 equivalent keeps the
same *calling contract* -- `forward(Dict[str, Tensor]) -> Dict[str, Tensor]` with the reference's keys, plus the
`_extra_files` metadata the pair styles read (pair_e3gnn.cpp:306-335) -- on top of the C-ABI kernels:

* `deploy()` / `DeployedSerial`: serial inputs `x` (type index), `pos`, `edge_index`, `num_atoms`,
  `cell_lattice_vectors`, `cell_volume`, `pbc_shift` (pair_e3gnn.cpp:218-225); outputs `inferred_total_energy`,
  `inferred_force`, `inferred_stress` (xx,yy,zz,xy,yz,zx), `atomic_energy` (pair_e3gnn.cpp:230-264).  One call = one
  pass of the whole-model engine (forward + analytic reverse kernels, no autograd).
* `deploy_parallel()` / `DeployedParallel`: one segment per interaction layer with the inter-segment keys `x`,
  `x_ghost`, `edge_vec`, `self_cont_tmp`, `nlocal` (pair_e3gnn_parallel.cpp:331-384); segments are autograd-capable
  module lists (sevennet_exp_b200.nn) because the pair style differentiates them one by one, exchanging ghost
  features / ghost gradients in between.  `DeployedParallel.compute()` is that loop (pair_e3gnn_parallel.cpp:
  331-437) with the two communications as callbacks.

File format (serial): a TorchScript archive written by `torch.jit.save` whose scripted forward calls the custom op
`torch.ops.sevenn_b200.evaluate` (csrc/torch_op.cpp -> the C++ whole-model host, csrc/model.cu) and whose buffers hold
the packed model; `_extra_files` carries the reference's metadata keys plus the weights-only bundle {format, config,
model_state_dict, extra_files} for the Python driver.  Parallel segments stay weights-only bundles (their consumer
differentiates them with autograd segment by segment).
"""
import datetime
import os
from typing import Callable, Dict, List, Optional, Tuple

import numpy as np
import torch

from . import __version__ as _VERSION  # noqa: F401
from .checkpoint import load_checkpoint, patch_old_config, patch_state_dict
from .spec import build_spec

SERIAL_FORMAT = 'sevenn_b200_deployed_serial'
PARALLEL_FORMAT = 'sevenn_b200_deployed_parallel'
_SYMBOLS = ('X H He Li Be B C N O F Ne Na Mg Al Si P S Cl Ar K Ca Sc Ti V Cr Mn Fe Co Ni Cu Zn Ga Ge As Se Br Kr Rb '
            'Sr Y Zr Nb Mo Tc Ru Rh Pd Ag Cd In Sn Sb Te I Xe Cs Ba La Ce Pr Nd Pm Sm Eu Gd Tb Dy Ho Er Tm Yb Lu Hf '
            'Ta W Re Os Ir Pt Au Hg Tl Pb Bi Po At Rn Fr Ra Ac Th Pa U Np Pu Am Cm Bk Cf Es Fm Md No Lr').split()


def _resolve(model):
    """checkpoint path | (config, state_dict) -> (spec, engine-ready state_dict, config)."""
    if isinstance(model, (tuple, list)):
        cfg, sd = model
        cfg = patch_old_config(dict(cfg), require_version=False)
        spec = build_spec(cfg)
        return spec, patch_state_dict(spec, sd, cfg.get('version', '0.11.0')), cfg
    return load_checkpoint(model)


def extra_files(cfg: dict, spec, comm_size: Optional[int] = None) -> Dict[str, str]:
    """The `_extra_files` of a deployed model (deploy.py:51-69,128-137)."""
    md = {'chemical_symbols_to_index': ''.join(_SYMBOLS[int(z)] + ' ' for z in spec.type_map),
          'cutoff': str(spec.cutoff), 'num_species': str(spec.n_species),
          'model_type': cfg.get('model_type', 'E3_equivariant_model'), 'version': _VERSION,
          'dtype': cfg.get('dtype', 'single'), 'time': datetime.datetime.now().strftime('%Y-%m-%d')}
    if comm_size is not None:
        md['comm_size'] = str(comm_size)
    return md


def _plain_sd(sd):
    return {k: (v.detach().cpu().clone() if torch.is_tensor(v) else torch.as_tensor(v)) for k, v in sd.items()}


TORCH_OP_LIB = os.path.join(os.path.dirname(os.path.abspath(__file__)), '_build', 'libsevenn_b200_torch.so')
_op_loaded = False


def load_torch_op():
    """Registers torch.ops.sevenn_b200.evaluate (csrc/torch_op.cpp).  A C++ host does the same with one dlopen of
    libsevenn_b200_torch.so before torch::jit::load."""
    global _op_loaded
    if not _op_loaded:
        if not os.path.exists(TORCH_OP_LIB):
            raise RuntimeError(f'{TORCH_OP_LIB} is missing: run `python -m sevennet_exp_b200.build`')
        torch.ops.load_library(TORCH_OP_LIB)
        _op_loaded = True


class _SerialScriptModule(torch.nn.Module):
    """What gets scripted and saved: the packed model as buffers + a forward with the reference's dict contract
    (inputs pair_e3gnn.cpp:218-225, outputs :230-264) that hands the tensors to torch.ops.sevenn_b200.evaluate."""

    def __init__(self, prog: torch.Tensor, data: torch.Tensor):
        super().__init__()
        self.register_buffer('prog', prog)
        self.register_buffer('data', data)

    def forward(self, data: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        energy, force, virial, atomic = torch.ops.sevenn_b200.evaluate(
            self.prog, self.data, data['x'], data['pos'], data['edge_index'], data['cell_lattice_vectors'],
            data['pbc_shift'])
        vol = data['cell_volume'].to(virial.device).to(virial.dtype)
        # forces and stress are analytic, not autograd: nothing here is part of a graph (pos arrives with requires_grad)
        data['inferred_total_energy'] = energy.detach()
        data['inferred_force'] = force.detach()
        data['inferred_stress'] = (-virial / vol).detach()    # (xx, yy, zz, xy, yz, zx), force_output.py:120-134
        data['atomic_energy'] = atomic.detach()
        return data


def deploy(checkpoint, fname: str = 'deployed_serial.pt', modal: Optional[str] = None, **kwargs) -> str:
    """sevenn/scripts/deploy.py:20-70: a TorchScript file (`torch.jit.load` / `torch::jit::load`) with the reference's
    `_extra_files` metadata.  Its forward calls torch.ops.sevenn_b200.evaluate, so the loading process registers the op
    first (`load_torch_op()` / dlopen of libsevenn_b200_torch.so).  ``modal`` (:31-36): the modality a multi-modal
    checkpoint is frozen to.  ``format='bundle'`` writes the weights-only torch.save bundle instead (models the C++ host
    does not run: radial='mlp'-only activations, readout_as_fcn); the TorchScript file carries the same bundle as the
    extra file `sevenn_b200_bundle`, which is what `load_deployed(..., lib=...)` and the Python driver read."""
    import io
    format = kwargs.pop('format', 'torchscript')
    if kwargs:
        raise TypeError(f'unexpected arguments {list(kwargs)}')
    spec, sd, cfg = _resolve(checkpoint)
    if spec.n_modal:
        if modal is None or modal not in (spec.modal_map or {}):
            raise ValueError(f'modal must be one of {list(spec.modal_map or {})} for this checkpoint')
    elif modal is not None:
        raise ValueError('Modality given, but model has no modal_map')
    if not fname.endswith('.pt'):
        fname += '.pt'
    bundle = {'format': SERIAL_FORMAT, 'config': {**cfg, 'version': '0.11.0'}, 'model_state_dict': _plain_sd(sd),
              'extra_files': extra_files(cfg, spec), 'modal': modal}
    if format == 'bundle':
        torch.save(bundle, fname)
        return fname
    if format != 'torchscript':
        raise ValueError("format must be 'torchscript' or 'bundle'")
    from . import _lib, host
    from .engine import Engine
    # weights are packed on the CPU (no GPU needed to deploy): the library is only asked for its GEMM tile geometry
    eng = Engine(spec, sd, lib=_lib.Lib(_lib.load_for_symbol_check()), device='cpu', radial='table', modal=modal)
    why = host.supported(eng)
    if why:
        raise NotImplementedError(f"{why}: deploy(..., format='bundle') keeps this model on the Python driver")
    prog, data = host.pack_engine(eng)
    load_torch_op()
    module = torch.jit.script(_SerialScriptModule(torch.from_numpy(prog.copy()), torch.from_numpy(data.copy())))
    buf = io.BytesIO()
    torch.save(bundle, buf)
    md = dict(bundle['extra_files'])
    md['sevenn_b200_bundle'] = buf.getvalue()
    torch.jit.save(module, fname, _extra_files=md)
    return fname


def deploy_parallel(checkpoint, fname: str = 'deployed_parallel', modal: Optional[str] = None) -> str:
    """sevenn/scripts/deploy.py:73-148: a directory of `deployed_parallel_{i}.pt`, one per segment."""
    if modal is not None:
        raise NotImplementedError('multi-modal checkpoints: serial deployment only (deploy(..., modal=...))')
    spec, sd, cfg = _resolve(checkpoint)
    comm = max(L.conv.lx.dim for L in spec.layers)          # widest ghost row exchanged (`_comm_size`)
    md = extra_files(cfg, spec, comm_size=comm)
    os.makedirs(fname, exist_ok=True)
    for i in range(len(spec.layers)):
        torch.save({'format': PARALLEL_FORMAT, 'segment': i, 'n_segments': len(spec.layers),
                    'config': {**cfg, 'version': '0.11.0'}, 'model_state_dict': _plain_sd(sd), 'extra_files': md},
                   os.path.join(fname, f'deployed_parallel_{i}.pt'))
    return fname


# ------------------------------------------------------------------------------------------------ serial --
class DeployedSerial:
    """Callable with the serial deployed-model contract; also usable as SevenNetCalculator(file_type='torchscript')."""

    def __init__(self, cfg: dict, sd: dict, device='cuda', extra: Optional[Dict[str, str]] = None, lib=None,
                 radial: Optional[str] = None, modal: Optional[str] = None):
        from .engine import Engine
        cfg = patch_old_config(dict(cfg), require_version=False)
        self.spec = build_spec(cfg)
        self.config = cfg
        self.device = torch.device(device)
        self.engine = Engine(self.spec, sd, lib=lib, device=self.device, radial=radial, modal=modal)
        self.extra_files = dict(extra) if extra is not None else extra_files(cfg, self.spec)
        self.cutoff, self.type_map = self.spec.cutoff, self.spec.type_map

    def eval(self):
        return self

    def to(self, device):
        if torch.device(device) != self.device:
            raise RuntimeError('a deployed B200 model is bound to the device it was loaded on')
        return self

    def forward(self, data: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        from .engine import Graph
        from .nn import functional as F
        dev = self.device
        x = data['x'].to(dev)
        pos = data['pos'].detach().to(dev)
        ei = data['edge_index'].to(dev)
        n = int(x.shape[0])
        cell = data['cell_lattice_vectors'].detach().to(dev).reshape(3, 3).to(pos.dtype)
        # EdgePreprocess (edge_embedding.py:13-78) without the strain tensor: stress is the analytic virial
        ev = (pos[ei[1]] - pos[ei[0]] + data['pbc_shift'].to(dev).to(pos.dtype) @ cell).float()
        graph, perm = F.build_graph(ei, n, n, dev)
        if perm is not None:
            ev = ev[perm]
        vol = float(data['cell_volume'])
        g = Graph(x.to(torch.int32).contiguous(), ev.contiguous(), graph['rowptr'], graph['col'], graph['cptr'],
                  graph['cedge'], n, vol if vol > 0 else float(np.finfo(float).eps))
        out = self.engine.evaluate(g)
        data['edge_vec'] = ev if perm is None else ev[torch.argsort(perm)]
        data['atomic_energy'] = out['atomic_energy'].float().reshape(n, 1)
        data['inferred_total_energy'] = out['energy'].float().reshape(())
        data['inferred_force'] = out['forces'][:n].float()
        data['inferred_stress'] = (-(out['virial6'] / g.volume)).float()
        return data

    __call__ = forward


# ---------------------------------------------------------------------------------------------- parallel --
class DeployedParallel:
    """The list of segment models of `pair_style e3gnn/parallel` + the pair style's compute loop."""

    def __init__(self, cfg: dict, sd: dict, device='cuda', extra: Optional[Dict[str, str]] = None):
        from . import nn as bnn
        cfg = patch_old_config(dict(cfg), require_version=False)
        self.spec, self.config, self.device = build_spec(cfg), cfg, torch.device(device)
        self.segments = bnn.build_E3_equivariant_model(cfg, parallel=True)
        full = dict(sd)
        for k, v in sd.items():                      # ghost twins share the weights (deploy.py:77-95)
            if k.startswith('onehot_to_feature_x') or k.startswith('0_self_interaction_1'):
                full[f'ghost_{k}'] = v
        for seg in self.segments:
            missing, _ = seg.load_state_dict(full, strict=False)
            assert not missing, missing
            seg.to(self.device).eval()
        comm = max(L.conv.lx.dim for L in self.spec.layers)
        self.extra_files = dict(extra) if extra is not None else extra_files(cfg, self.spec, comm_size=comm)
        self.cutoff = self.spec.cutoff

    def __len__(self):
        return len(self.segments)

    def __getitem__(self, i):
        return self.segments[i]

    def compute(self, x, x_ghost, edge_index, edge_vec, forward_comm, reverse_comm, record: Optional[dict] = None):
        """The pair style's evaluation of one rank; see `run_parallel_segments`."""
        return run_parallel_segments(self.segments, x, x_ghost, edge_index, edge_vec, forward_comm, reverse_comm,
                                     device=self.device, record=record)


def run_parallel_segments(segments, x: torch.Tensor, x_ghost: torch.Tensor, edge_index: torch.Tensor,
                          edge_vec: torch.Tensor, forward_comm: Callable[[torch.Tensor], torch.Tensor],
                          reverse_comm: Callable[[torch.Tensor, torch.Tensor], torch.Tensor], device='cuda',
                          record: Optional[dict] = None) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """pair_e3gnn_parallel.cpp:331-437 for any list of segment callables with the deployed-parallel dict contract
    (ours, or the reference's own `deployed_parallel_{i}.pt` loaded with torch.jit -- tests/golden uses that).
    x / x_ghost: type indices of the owned / ghost atoms; edge_index rows index [owned | ghost];
    `forward_comm(x_local) -> x_ghost` fills ghost rows from their owners; `reverse_comm(dx_local, dx_ghost) ->
    dx_local` adds ghost gradients to their owners.  Returns (energy of the owned atoms, dE/d edge_vec [E,3],
    atomic energies [nlocal,1]); `record` (optional) receives the ghost payload `x` sent before each segment."""
    dev = torch.device(device)
    nlocal = int(x.shape[0])
    ev0 = edge_vec.detach().to(dev).float().requires_grad_(True)
    data = {'x': x.to(dev), 'x_ghost': x_ghost.to(dev), 'edge_index': edge_index.to(dev), 'edge_vec': ev0,
            'num_atoms': torch.tensor([nlocal]).to(dev), 'nlocal': torch.tensor([nlocal])}    # 1-D CPU tensor, as pair_e3gnn_parallel.cpp:337 sends it
    wrt: List[List[torch.Tensor]] = [[ev0]]
    out = dict(segments[0](data))
    for k, seg in enumerate(segments[1:], start=1):
        if record is not None:
            record[f'x_{k}'] = out['x'].detach().cpu().clone()
        xg = forward_comm(out['x'].detach()).detach().requires_grad_(True)
        out['x_ghost'] = xg
        out['edge_vec'] = out['edge_vec'].detach().clone().requires_grad_(True)
        wrt.append([out['edge_vec'], out['x'], out['self_cont_tmp'], xg])
        out = dict(seg(out))
    energy = out['inferred_total_energy'].squeeze()
    dE_dr = torch.zeros_like(ev0)
    of: List[torch.Tensor] = []
    gouts: Optional[List[torch.Tensor]] = None
    for k in range(len(wrt) - 1, -1, -1):
        w = wrt[k]
        if gouts is None:
            grads = torch.autograd.grad([energy], w, allow_unused=True)
        else:
            grads = torch.autograd.grad(of, w, gouts, allow_unused=True)
        if grads[0] is not None:
            dE_dr = dE_dr + grads[0]
        if k == 0:
            break
        of = [w[1], w[2]]
        dx_ghost = grads[3] if grads[3] is not None else torch.zeros_like(w[3])
        dx_own = grads[1] if grads[1] is not None else torch.zeros_like(w[1])
        dx_local = reverse_comm(dx_own.detach(), dx_ghost.detach())
        dsc = grads[2] if grads[2] is not None else torch.zeros_like(w[2])
        gouts = [dx_local, dsc]
    return energy.detach(), dE_dr.detach(), out['atomic_energy'].detach()


EXTRA_KEYS = ('chemical_symbols_to_index', 'cutoff', 'num_species', 'model_type', 'version', 'dtype', 'time')


def _read_bundle(path: str):
    """(bundle dict, is_torchscript) of a file written by deploy() / deploy_parallel()."""
    import io
    import zipfile
    if zipfile.is_zipfile(path):
        with zipfile.ZipFile(path) as z:
            names = z.namelist()
            extra = [n for n in names if n.endswith('extra/sevenn_b200_bundle')]
            if extra:
                return torch.load(io.BytesIO(z.read(extra[0])), weights_only=False, map_location='cpu'), True
            if any(n.endswith('constants.pkl') or '/code/' in n for n in names):
                raise ValueError(f'{path}: a TorchScript file written by the reference freezes the e3nn graph and cannot '
                                 'run on the B200 kernels; deploy the checkpoint with sevennet_exp_b200.deploy.deploy()')
    try:
        b = torch.load(path, weights_only=False, map_location='cpu')
    except Exception as e:  # noqa: BLE001
        raise ValueError(f'{path}: not a sevennet_exp_b200 deployed model ({e})') from e
    return b, False


def load_deployed(path: str, device='cuda', lib=None, scripted: Optional[bool] = None):
    """`torch.jit.load(path, _extra_files=...)` of the reference -> (model, extra_files).

    A serial model deployed as TorchScript comes back as the `torch.jit.ScriptModule` itself (its forward runs the
    C++ host through torch.ops.sevenn_b200.evaluate) unless ``scripted=False`` or an explicit ``lib`` asks for the
    Python driver (`DeployedSerial`) on the bundle stored inside the file."""
    if os.path.isdir(path):
        path = os.path.join(path, 'deployed_parallel_0.pt')
    b, is_ts = _read_bundle(path)
    fmt = b.get('format') if isinstance(b, dict) else None
    if fmt == SERIAL_FORMAT and is_ts and lib is None and scripted is not False:
        load_torch_op()
        md = {k: '' for k in EXTRA_KEYS}
        m = torch.jit.load(path, map_location='cpu', _extra_files=md)
        return m, {k: (v.decode() if isinstance(v, bytes) else v) for k, v in md.items()}
    if fmt == SERIAL_FORMAT:
        m = DeployedSerial(b['config'], b['model_state_dict'], device=device, extra=b['extra_files'], lib=lib,
                           modal=b.get('modal'))
    elif fmt == PARALLEL_FORMAT:
        m = DeployedParallel(b['config'], b['model_state_dict'], device=device, extra=b['extra_files'])
    else:
        raise ValueError(f'{path}: a TorchScript file written by the reference freezes the e3nn graph and cannot run '
                         'on the B200 kernels; deploy the checkpoint with sevennet_exp_b200.deploy.deploy()')
    return m, dict(m.extra_files)

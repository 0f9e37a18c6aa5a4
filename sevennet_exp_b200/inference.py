"""Batch inference over labelled structures with the reference's on-disk result format.

Mirrors ``sevenn/scripts/inference.py:119-220`` (``inference``) and ``:16-103`` (``write_inference_csv``): every target
structure is evaluated (E, F, S) through :class:`SevenNetCalculator.compute_batch`, the six default error metrics of
``sevenn/util.py:240-260`` / ``sevenn/error_recorder.py:14-57,155-217`` are written to ``errors.txt`` and the
per-structure / per-atom tables to ``per_graph.csv`` / ``per_atom.csv`` with the reference's column names and order
(``tests/data/inferences/snet0_on_hfo2/*.csv`` in the reference tree).  Conventions kept from the reference:

* the stress label is ``-stress_ase[[0, 1, 2, 5, 3, 4]]`` (xx, yy, zz, xy, yz, zx; ``sevenn/train/dataload.py``), both
  label and prediction are written in kbar (x 1602.1766208);
* Energy errors are per atom, Force RMSE is the norm of the 3-vector error, Stress RMSE the norm of the 6-vector
  error, MAE averages over all components;
* a missing label is written as NaN and left out of the metrics (``ignore_unlabeled``).

Targets are extended-xyz files (read by the small parser below: ASE is not a dependency of this package) or
in-memory structures.  There is no CPU path: the calculator raises without the CUDA library.
"""
import csv
import math
import os
import re
from typing import Dict, Iterable, List, Optional, Sequence, Union

import numpy as np

KBAR = 1602.1766208          # eV/A^3 -> kbar, sevenn/train/loss.py:180
_STRESS_SFX = ('xx', 'yy', 'zz', 'xy', 'yz', 'zx')

_SYMBOLS = ('X H He Li Be B C N O F Ne Na Mg Al Si P S Cl Ar K Ca Sc Ti V Cr Mn Fe Co Ni Cu Zn Ga Ge As Se Br Kr Rb '
            'Sr Y Zr Nb Mo Tc Ru Rh Pd Ag Cd In Sn Sb Te I Xe Cs Ba La Ce Pr Nd Pm Sm Eu Gd Tb Dy Ho Er Tm Yb Lu Hf Ta '
            'W Re Os Ir Pt Au Hg Tl Pb Bi Po At Rn Fr Ra Ac Th Pa U Np Pu Am Cm Bk Cf Es Fm Md No Lr Rf Db Sg Bh Hs Mt '
            'Ds Rg Cn Nh Fl Mc Lv Ts Og').split()
_Z = {s: z for z, s in enumerate(_SYMBOLS)}


def _comment_fields(line: str) -> Dict[str, str]:
    out = {}
    for m in re.finditer(r'([A-Za-z_][\w.-]*)\s*=\s*(?:"([^"]*)"|\'([^\']*)\'|(\S+))', line):
        out[m.group(1)] = next(v for v in m.groups()[1:] if v is not None)
    return out


def read_extxyz(path: str) -> List[dict]:
    """Frames of an extended-xyz file as dicts: numbers, positions, cell, pbc and -- where present -- energy
    (``free_energy`` preferred, like ASE's calculator results), forces, stress (3x3, ASE sign) and ``info``."""
    frames = []
    with open(path) as f:
        lines = f.read().splitlines()
    k = 0
    while k < len(lines):
        if not lines[k].strip():
            k += 1
            continue
        n = int(lines[k].split()[0])
        kv = _comment_fields(lines[k + 1])
        spans, c = {}, 0
        spec = kv.get('Properties', 'species:S:1:pos:R:3').split(':')
        for a in range(0, len(spec), 3):
            w = int(spec[a + 2])
            spans[spec[a]] = (c, c + w)
            c += w
        body = [ln.split() for ln in lines[k + 2:k + 2 + n]]
        if len(body) != n:
            raise ValueError(f'{path}: frame at line {k + 1} is truncated')

        def column(name):
            a, b = spans[name]
            return np.array([[float(v) for v in row[a:b]] for row in body], dtype=np.float64)

        sp = spans['species'][0]
        syms = [row[sp] for row in body]
        fr = {
            'numbers': np.array([_Z[s] if s in _Z else int(s) for s in syms], dtype=np.int64),
            'positions': column('pos'),
            'cell': (np.array(kv['Lattice'].split(), dtype=np.float64).reshape(3, 3) if 'Lattice' in kv
                     else np.zeros((3, 3))),
            'pbc': np.array([t.upper().startswith('T') for t in kv.get('pbc', 'T T T' if 'Lattice' in kv
                                                                        else 'F F F').split()]),
        }
        for name in ('forces', 'force'):
            if name in spans:
                fr['forces'] = column(name)
                break
        e = kv.get('free_energy', kv.get('energy'))
        if e is not None:
            fr['energy'] = float(e)
        if 'stress' in kv:
            fr['stress'] = np.array(kv['stress'].split(), dtype=np.float64).reshape(3, 3)
        elif 'virial' in kv:
            vol = abs(np.linalg.det(fr['cell']))
            fr['stress'] = -np.array(kv['virial'].split(), dtype=np.float64).reshape(3, 3) / vol
        used = {'Lattice', 'Properties', 'pbc', 'energy', 'free_energy', 'stress', 'virial'}
        fr['info'] = {a: b for a, b in kv.items() if a not in used}
        frames.append(fr)
        k += 2 + n
    return frames


def _stress_label(fr: dict) -> np.ndarray:
    """(xx, yy, zz, xy, yz, zx) with the reference's sign, or NaNs."""
    s = fr.get('stress')
    if s is None:
        return np.full(6, np.nan)
    s = np.asarray(s, dtype=np.float64)
    if s.shape == (3, 3):
        voigt = np.array([s[0, 0], s[1, 1], s[2, 2], s[1, 2], s[0, 2], s[0, 1]])
    else:
        voigt = s.reshape(6)
    return -voigt[[0, 1, 2, 5, 3, 4]]


class _Mean:
    def __init__(self):
        self.s, self.n = 0.0, 0

    def add(self, values):
        v = np.asarray(values, dtype=np.float64).reshape(-1)
        self.s += float(v.sum())
        self.n += v.size

    def get(self):
        return self.s / self.n if self.n else float('nan')


def error_metrics(records: Sequence[dict]) -> Dict[str, float]:
    """Energy / Force / Stress RMSE and MAE over evaluated records (see :func:`evaluate_targets`), in the order and
    with the names of the reference's default recorder."""
    sq = {k: _Mean() for k in ('E', 'F', 'S')}
    ab = {k: _Mean() for k in ('E', 'F', 'S')}
    for r in records:
        n = len(r['numbers'])
        if not math.isnan(r['total_energy']):
            d = (r['inferred_total_energy'] - r['total_energy']) / n
            sq['E'].add([d * d])
            ab['E'].add([abs(d)])
        if not np.isnan(r['force_of_atoms']).any():
            d = r['inferred_force'] - r['force_of_atoms']
            sq['F'].add((d * d).sum(axis=1))
            ab['F'].add(np.abs(d))
        if not np.isnan(r['stress']).any():
            d = (r['inferred_stress'] - r['stress']) * KBAR
            sq['S'].add([(d * d).sum()])
            ab['S'].add(np.abs(d))
    return {
        'Energy_RMSE (eV/atom)': sq['E'].get() ** 0.5, 'Force_RMSE (eV/Å)': sq['F'].get() ** 0.5,
        'Stress_RMSE (kbar)': sq['S'].get() ** 0.5, 'Energy_MAE (eV/atom)': ab['E'].get(),
        'Force_MAE (eV/Å)': ab['F'].get(), 'Stress_MAE (kbar)': ab['S'].get(),
    }


def evaluate_targets(calc, frames: Iterable[dict], batch_size: int = 4) -> List[dict]:
    """Run the model over ``frames`` (dicts as returned by :func:`read_extxyz`) ``batch_size`` structures per
    evaluation; returns one record per structure holding labels and predictions under the reference's key names."""
    frames = list(frames)
    records = []
    for b0 in range(0, len(frames), max(1, int(batch_size))):
        chunk = frames[b0:b0 + max(1, int(batch_size))]
        results = calc.compute_batch([(fr['numbers'], fr['positions'], fr['cell'], fr['pbc']) for fr in chunk])
        for fr, res in zip(chunk, results):
            n = len(fr['numbers'])
            records.append({
                'numbers': np.asarray(fr['numbers']), 'pos': np.asarray(fr['positions'], dtype=np.float64),
                'cell_lattice_vectors': np.asarray(fr['cell'], dtype=np.float64),
                'user_label': fr.get('user_label', '-'),
                'total_energy': float(fr['energy']) if 'energy' in fr else float('nan'),
                'force_of_atoms': (np.asarray(fr['forces'], dtype=np.float64) if 'forces' in fr
                                   else np.full((n, 3), np.nan)),
                'stress': _stress_label(fr),
                'inferred_total_energy': res['energy'], 'inferred_force': np.asarray(res['forces']),
                'inferred_stress': np.asarray(res['inferred_stress']), 'atomic_energy': np.asarray(res['energies']),
                'info': dict(fr.get('info', {})),
            })
    return records


def _f32(v):
    return np.float32(v)


def write_inference_csv(records: Sequence[dict], out: str) -> None:
    """``info.csv``, ``per_graph.csv`` and ``per_atom.csv`` in ``out`` (sevenn/scripts/inference.py:16-103)."""
    os.makedirs(out, exist_ok=True)
    info_keys = []
    for r in records:
        for key in r['info']:
            if key not in info_keys:
                info_keys.append(key)
    if info_keys:
        with open(os.path.join(out, 'info.csv'), 'w', newline='') as f:
            w = csv.DictWriter(f, fieldnames=info_keys)
            w.writeheader()
            for r in records:
                w.writerow({key: r['info'].get(key, '') for key in info_keys})

    with open(os.path.join(out, 'per_graph.csv'), 'w', newline='') as f:
        head = (['num_atoms', 'user_label', 'total_energy', 'inferred_total_energy'] +
                [f'stress_{s}' for s in _STRESS_SFX] + [f'inferred_stress_{s}' for s in _STRESS_SFX] +
                [f'cell_lattice_vectors_{s}' for s in 'abc'])
        w = csv.writer(f)
        w.writerow(head)
        for r in records:
            cell = r['cell_lattice_vectors'].astype(np.float32)
            w.writerow([len(r['numbers']), r['user_label'], _f32(r['total_energy']), _f32(r['inferred_total_energy'])] +
                       [_f32(v * KBAR) for v in r['stress']] + [_f32(v * KBAR) for v in r['inferred_stress']] +
                       [str(cell[i]) for i in range(3)])

    with open(os.path.join(out, 'per_atom.csv'), 'w', newline='') as f:
        head = (['stct_id', 'atom_id', 'atomic_numbers', 'atomic_energy'] + [f'pos_{s}' for s in 'xyz'] +
                [f'force_of_atoms_{s}' for s in 'xyz'] + [f'inferred_force_{s}' for s in 'xyz'])
        w = csv.writer(f)
        w.writerow(head)
        for i, r in enumerate(records):
            for j in range(len(r['numbers'])):
                w.writerow([i, j, int(r['numbers'][j]), _f32(r['atomic_energy'][j])] +
                           [_f32(v) for v in r['pos'][j]] + [_f32(v) for v in r['force_of_atoms'][j]] +
                           [_f32(v) for v in r['inferred_force'][j]])


def inference(checkpoint, targets: Union[str, List[str], List[dict]], output_dir: str, num_workers: int = 1,
              device: str = 'auto', batch_size: int = 4, save_graph: bool = False, allow_unlabeled: bool = False,
              modal: Optional[str] = None, **data_kwargs) -> Dict[str, float]:
    """Evaluate ``checkpoint`` on ``targets`` and write errors.txt / per_graph.csv / per_atom.csv to ``output_dir``.

    ``checkpoint``: a path accepted by :class:`SevenNetCalculator`, a ``(config, state_dict)`` pair, or a calculator.
    ``targets``: extended-xyz path(s) or a list of frame dicts.  ``num_workers`` is accepted for signature
    compatibility (the neighbour list is built on the GPU).  Returns the error dict that errors.txt holds."""
    from .sevennet_calculator import SevenNetCalculator
    if save_graph:
        raise NotImplementedError('save_graph: the torch_geometric graph cache (sevenn_data/*.pt) is not part of this package')
    # `modal` (sevenn/scripts/inference.py:158-163): unknown / superfluous modality names raise ValueError in the engine
    if isinstance(checkpoint, SevenNetCalculator) or hasattr(checkpoint, 'compute_batch'):
        calc = checkpoint
        if modal and getattr(calc, 'modal', None) != modal:
            raise ValueError(f'calculator was built for modal {getattr(calc, "modal", None)!r}, not {modal!r}')
    elif isinstance(checkpoint, tuple):
        calc = SevenNetCalculator(checkpoint, file_type='model_instance', device=device, modal=modal)
    else:
        calc = SevenNetCalculator(checkpoint, file_type='checkpoint', device=device, modal=modal)
    if isinstance(targets, str):
        targets = [targets]
    index = data_kwargs.pop('index', None)
    if data_kwargs:
        raise TypeError(f'unsupported data arguments {sorted(data_kwargs)}')
    frames = []
    for t in targets:
        if isinstance(t, dict):
            frames.append(t)
            continue
        got = read_extxyz(t)
        if index is not None:
            if str(index) == ':':
                pass
            elif ':' in str(index):
                a, b = (int(v) if v else None for v in str(index).split(':')[:2])
                got = got[a:b]
            else:
                got = [got[int(index)]]
        for fr in got:
            fr['info'] = {**fr.get('info', {}), 'file': os.path.abspath(t)}
        frames.extend(got)
    if not allow_unlabeled:
        for fr in frames:
            if 'energy' not in fr or 'forces' not in fr:
                raise ValueError('target without energy / force labels (pass allow_unlabeled=True)')
    records = evaluate_targets(calc, frames, batch_size)
    errors = error_metrics(records)
    os.makedirs(output_dir, exist_ok=True)
    with open(os.path.join(output_dir, 'errors.txt'), 'w', encoding='utf-8') as f:
        for key, val in errors.items():
            f.write(f'{key}: {val}\n')
    write_inference_csv(records, output_dir)
    return errors

"""CPU calibration asked for by SURVEY 8d: is the oracle port (the `cpu_baseline` / `--impl reference` arm of bench.py)
at least as fast as REAL reference code at equal shapes?  Times the reference's frozen TorchScript artefact
(example_inputs/md_serial_example/deployed_serial.pt: e3nn's generated code, autograd forces and strain stress) and the
oracle restatement with the same (transplanted) weights on the same replicated HfO2 cell, same threads.
Runs only where /root/reference exists (this container); prints one JSON line.  Test infrastructure, not product."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import sevenn_ref as S                                  # noqa: E402
from tests.test_oracle import deployed_serial_case                  # noqa: E402

REF = '/root/reference'
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
cfg, sd, numbers, pos, cell, gold = deployed_serial_case(os.path.join(ROOT, 'tests', 'golden'))
shifts = np.stack(np.meshgrid(*[np.arange(reps)] * 3, indexing='ij'), -1).reshape(-1, 3)
pos = np.concatenate([pos + s @ cell for s in shifts])
numbers = np.tile(numbers, len(shifts))
cell = cell * reps
rc = float(cfg['cutoff'])
i, j, D, Sh = S.neighbor_list_bruteforce(pos, cell, [True] * 3, rc)
tmap = cfg['_type_map']
extra = {'cutoff': b''}
m = torch.jit.load(f'{REF}/example_inputs/md_serial_example/deployed_serial.pt', _extra_files=extra, map_location='cpu')
model = S.build_model(dict(cfg))
sd = S.sort_old_convolution(model, sd)                    # artefact is sevenn 0.8.6: pre-0.11 instruction order
own = model.state_dict()
model.load_state_dict({k: v for k, v in sd.items() if k in own}, strict=False)
model = model.eval()
ei = torch.tensor(np.stack([i, j]), dtype=torch.int64)


def run_ref():
    p = torch.tensor(pos, dtype=torch.float32, requires_grad=True)
    data = {'x': torch.tensor([tmap[int(z)] for z in numbers], dtype=torch.int64), 'pos': p,
            'edge_index': torch.tensor(np.stack([i, j]), dtype=torch.int64),
            'num_atoms': torch.tensor([len(numbers)], dtype=torch.int64),
            'cell_lattice_vectors': torch.tensor(cell, dtype=torch.float32),
            'cell_volume': torch.tensor(abs(np.linalg.det(cell)), dtype=torch.float32),
            'pbc_shift': torch.tensor(Sh, dtype=torch.float32)}
    r = m(data)
    return float(r['inferred_total_energy'].detach()), r['inferred_force'].detach().numpy()


def run_port():
    # same work as the artefact: edge vectors from positions + shifts, model, autograd forces / stress
    post, cellt = torch.tensor(pos, dtype=torch.float32), torch.tensor(cell, dtype=torch.float32)
    ev = (post[ei[1]] - post[ei[0]] + torch.tensor(Sh, dtype=torch.float32) @ cellt).requires_grad_(True)
    r = model({S.Z: torch.tensor(numbers), S.EDGE_IDX: ei, S.EDGE_VEC: ev, S.NATOMS: torch.tensor([len(numbers)]),
               S.VOL: torch.tensor(abs(np.linalg.det(cell)), dtype=torch.float32)})
    return float(r[S.E_TOT].detach()), r[S.FORCE].detach().numpy()


def timed(f, n=5):
    f(); f()
    ts = []
    for _ in range(n):
        t0 = time.perf_counter(); out = f(); ts.append(time.perf_counter() - t0)
    return float(np.median(ts)), out


t_ref, (e_ref, f_ref) = timed(run_ref)
t_port, (e_port, f_port) = timed(run_port)
print(json.dumps({'atoms': len(numbers), 'edges': len(i), 'threads': torch.get_num_threads(),
                  'reference_artifact_s': t_ref, 'oracle_port_s': t_port, 'port_over_reference': t_port / t_ref,
                  'dE_per_atom': abs(e_ref - e_port) / len(numbers), 'max_dF': float(np.abs(f_ref - f_port).max()),
                  'note': 'neighbour list outside both timed regions'}))

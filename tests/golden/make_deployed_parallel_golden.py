"""Runs the reference's own parallel deployed artefacts example_inputs/md_parallel_example/deployed_parallel/
deployed_parallel_{0..3}.pt (sevenn 0.8.6; same checkpoint as deployed_serial.pt) on frame 0 of hfo2.extxyz the way
pair_e3gnn_parallel.cpp drives them (segment by segment, ghost features forward, ghost gradients backward), on one
process with the periodic images as ghosts, and stores what crossed the wire + the results:

  tests/golden/deployed_parallel_hfo2.npz : edge_index over [owned | ghost], edge_vec, ghost owner map,
      x_1..x_3 (the per-layer ghost payloads), energy, dE/d edge_vec, atomic energies

Run in the authoring container only (needs /root/reference)."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, '..', '..'))
from sevennet_exp_b200.deploy import run_parallel_segments  # noqa: E402
from tests.test_deploy_emu import _periodic_ghost_graph  # noqa: E402

REF = '/root/reference/example_inputs/md_parallel_example/deployed_parallel'


def main():
    segs = [torch.jit.load(f'{REF}/deployed_parallel_{k}.pt', map_location='cpu') for k in range(4)]
    fr = np.load(f'{HERE}/hfo2_frames.npz')
    numbers, pos, cell = fr['f0_numbers'], fr['f0_pos'], fr['f0_cell']
    ei, ev, owner = _periodic_ghost_graph(numbers, pos, cell, 4.0)
    tmap = {72: 0, 8: 1}
    types = torch.tensor([tmap[int(z)] for z in numbers], dtype=torch.int64)
    own = torch.tensor(owner)
    rec = {}
    E, dE_dr, ae = run_parallel_segments(
        segs, types, types[own], torch.tensor(ei), torch.tensor(ev), lambda x: x[own],
        lambda dx, dg: dx.clone().index_add_(0, own, dg), device='cpu', record=rec)
    serial = np.load(f'{HERE}/deployed_serial_hfo2.npz')
    n = len(numbers)
    F = torch.zeros(n + len(owner), 3).index_add_(0, torch.tensor(ei[0]), dE_dr).index_add_(0, torch.tensor(ei[1]), -dE_dr)
    F = F[:n].index_add_(0, own, F[n:])
    print('E', E.item(), 'serial artefact', float(serial['energy']), 'max|dF| vs serial artefact',
          np.abs(F.numpy() - serial['force']).max())
    np.savez(f'{HERE}/deployed_parallel_hfo2.npz', edge_index=ei, edge_vec=ev, owner=owner, energy=E.numpy(),
             dE_dr=dE_dr.numpy(), atomic_energy=ae.numpy(), **{k: v.numpy() for k, v in rec.items()})
    print({k: tuple(v.shape) for k, v in rec.items()})


if __name__ == '__main__':
    main()

"""Spatial domain decomposition + halo exchange (sevennet_exp_b200/domain.py) on world_size 2 and 4 with
the gloo backend on CPU (kernels through the emulation build): decomposed E/F/virial == single-domain oracle,
like the reference's test_lammps.py:278-316 (1-4 MPI ranks == serial calculator)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from sevennet_exp_b200 import presets, spec as sp


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        from oracle import sevenn_ref as S
        from sevennet_exp_b200.domain import Decomposition, Halo, brick_grid, evaluate_decomposed
        from sevennet_exp_b200.engine import Engine
        from sevennet_exp_b200.neighbor import NeighborList
        from tests.emu_util import build_full_emu
        from tests.structures import random_cell
        torch.set_num_threads(1)
        cfg = dict(presets.PRESETS['test-so3-l2'])
        cfg['version'] = '0.11.0'
        torch.manual_seed(777)
        model = S.build_model(cfg).eval()
        spec = sp.build_spec(cfg)
        lib = build_full_emu([L.conv for L in spec.layers])
        numbers, pos, cell = random_cell(110, list(cfg['_type_map']), density=0.05, seed=11, triclinic=False)

        class Calc:   # the pieces of SevenNetCalculator that evaluate_decomposed uses, on the emulation lib
            pass
        calc = Calc()
        calc.device = torch.device('cpu')
        calc.engine = Engine(spec, model.state_dict(), lib=lib, device='cpu')
        calc.nl = NeighborList(lib=lib, device='cpu')
        calc.cutoff = spec.cutoff
        lut = -np.ones(120, dtype=np.int64)
        for z, i in spec.type_map.items():
            lut[int(z)] = i
        calc._lut = lut
        dec = Decomposition(numbers, pos, cell, spec.cutoff, brick_grid(world), rank)
        halo = Halo(dec, lib, 'cpu')
        E, F, vir, g = evaluate_decomposed(calc, dec, halo)
        ref = model(S.atoms_to_graph(numbers, pos, cell, [True] * 3, cfg['cutoff']))
        n = len(numbers)
        counts = [None] * world
        dist.all_gather_object(counts, dec.n_local)
        assert sum(counts) == n
        assert abs(E.item() - ref[S.E_TOT].item()) / n < 1e-5
        Fref = ref[S.FORCE].detach()[torch.from_numpy(dec.local_ids)]
        assert (F - Fref).abs().max().item() < 1e-4
        stress = -(vir / dec.volume).float()
        assert (stress - ref[S.STRESS].detach()).abs().max().item() * 1602.1766208 < 1e-3 * 10
        # atoms move across brick faces and the periodic boundary (pair_e3gnn_parallel.cpp:684-785 relies on
        # LAMMPS re-assigning owners and ghosts): a rigid shift of 0.37 of the cell along x and y plus a random walk,
        # positions left unwrapped; a new Decomposition / Halo from the moved positions must again equal the oracle
        rng = np.random.default_rng(5)
        pos2 = pos + 0.37 * (cell[0] + cell[1]) + rng.normal(scale=0.08, size=pos.shape)
        dec2 = Decomposition(numbers, pos2, cell, spec.cutoff, brick_grid(world), rank)
        moved = len(set(dec.local_ids.tolist()) ^ set(dec2.local_ids.tolist()))
        halo2 = Halo(dec2, lib, 'cpu')
        E2, F2, vir2, _ = evaluate_decomposed(calc, dec2, halo2)
        ref2 = model(S.atoms_to_graph(numbers, pos2, cell, [True] * 3, cfg['cutoff']))
        assert abs(E2.item() - ref2[S.E_TOT].item()) / n < 1e-5
        assert (F2 - ref2[S.FORCE].detach()[torch.from_numpy(dec2.local_ids)]).abs().max().item() < 1e-4
        assert ((-(vir2 / dec2.volume).float()) - ref2[S.STRESS].detach()).abs().max().item() * 1602.1766208 < 1e-3 * 10
        ret[rank] = (dec.n_local, dec.n_ghost, moved)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('world', [2, 4])
def test_decomposed_equals_single_domain(world):
    # compile the emulation library once, in the parent
    from tests.emu_util import build_full_emu
    cfg = dict(presets.PRESETS['test-so3-l2'])
    build_full_emu([L.conv for L in sp.build_spec(cfg).layers])
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), ret), nprocs=world, join=True)
    assert len(ret) == world and all(v[1] > 0 for v in ret.values())
    assert all(v[2] > 10 for v in ret.values())            # ownership really changed on every rank


def test_brick_grid():
    from sevennet_exp_b200.domain import brick_grid
    assert brick_grid(1) == (1, 1, 1) and brick_grid(2) == (2, 1, 1)
    assert brick_grid(4) == (2, 2, 1) and brick_grid(8) == (2, 2, 2)


def test_interior_atoms_have_no_ghost_neighbours():
    """domain.Decomposition puts the owned atoms farther than the cutoff from every brick face first (n_interior): a host
    may convolve those rows while the ghost rows are still in flight (SURVEY 8e "Overlap")."""
    from sevennet_exp_b200.domain import Decomposition
    from tests.structures import random_cell
    numbers, pos, cell = random_cell(600, [72, 8], density=0.05, seed=3, triclinic=False)
    rc = 4.0
    for grid, world in (((2, 1, 1), 2), ((2, 2, 1), 4)):
        for rank in range(world):
            dec = Decomposition(numbers, pos, cell, rc, grid, rank)
            assert 0 < dec.n_interior < dec.n_local
            owned, ghosts = dec.pos[:dec.n_local], dec.pos[dec.n_local:]
            d = np.linalg.norm(owned[:dec.n_interior, None, :] - ghosts[None, :, :], axis=2)
            assert d.min() >= rc                    # no ghost within the cutoff of an interior atom
            d2 = np.linalg.norm(owned[dec.n_interior:, None, :] - ghosts[None, :, :], axis=2)
            assert (d2.min(axis=1) < rc).mean() > 0.5     # most boundary atoms do have one

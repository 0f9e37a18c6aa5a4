"""Deployed-model entry points (sevennet_exp_b200/deploy.py) through the emulation build of the kernels:

* serial contract of `pair_style e3gnn` (pair_e3gnn.cpp:218-264) fed with the very input dict the reference's
  own deployed TorchScript artefact received, checked against that artefact's outputs (tests/golden);
* parallel contract of `pair_style e3gnn/parallel` (pair_e3gnn_parallel.cpp:331-437): segment models + the
  forward / reverse ghost communication, one process, ghosts = periodic images, against the serial oracle."""
import os

import numpy as np
import pytest
import torch

from oracle import sevenn_ref as S
from sevennet_exp_b200 import deploy as D
from sevennet_exp_b200 import nn as bnn
from sevennet_exp_b200 import presets, spec as sp
from sevennet_exp_b200.checkpoint import patch_old_config
from tests.emu_util import build_full_emu
from tests.structures import random_cell
from tests.test_oracle import deployed_serial_case

KBAR = 1602.1766208


@pytest.fixture(autouse=True)
def _reset_lib():
    yield
    bnn.set_lib(None)


def _emu_for(cfg):
    lib = build_full_emu([L.conv for L in sp.build_spec(patch_old_config(dict(cfg), require_version=False)).layers])
    bnn.set_lib(lib)
    return lib


def test_deployed_serial_contract_against_reference_artefact(golden_dir, tmp_path):
    cfg, sd, numbers, pos, cell, gold = deployed_serial_case(golden_dir)
    lib = _emu_for(cfg)
    path = D.deploy((cfg, sd), str(tmp_path / 'deployed_serial'))
    assert path.endswith('deployed_serial.pt')
    model, extra = D.load_deployed(path, device='cpu', lib=lib)
    # metadata the pair style reads (pair_e3gnn.cpp:306-335); the reference artefact says 'Hf O ', '4.0', '2'
    assert extra['chemical_symbols_to_index'] == 'Hf O ' and float(extra['cutoff']) == 4.0
    assert extra['num_species'] == '2' and extra['model_type'] == 'E3_equivariant_model' and extra['dtype'] == 'single'
    E = gold['edge_index'].shape[1]
    p = np.random.default_rng(0).permutation(E)            # LAMMPS gives no ordering guarantee
    data = {'x': torch.tensor([cfg['_type_map'][int(z)] for z in numbers], dtype=torch.int64),
            'pos': torch.tensor(pos, dtype=torch.float32, requires_grad=True),
            'edge_index': torch.tensor(gold['edge_index'][:, p], dtype=torch.int64),
            'num_atoms': torch.tensor([len(numbers)], dtype=torch.int64),
            'cell_lattice_vectors': torch.tensor(cell, dtype=torch.float32),
            'cell_volume': torch.tensor(abs(np.linalg.det(cell)), dtype=torch.float32),
            'pbc_shift': torch.tensor(gold['pbc_shift'][p], dtype=torch.float32)}
    out = model(data)
    n = len(numbers)
    assert abs(out['inferred_total_energy'].item() - float(gold['energy'])) / n < 1e-5
    assert np.abs(out['atomic_energy'].numpy() - gold['atomic_energy'].reshape(n, 1)).max() < 1e-5
    assert np.abs(out['inferred_force'].numpy() - gold['force']).max() < 1e-4
    assert np.abs(out['inferred_stress'].numpy() - gold['stress']).max() * KBAR < 1e-2
    with pytest.raises(ValueError):
        D.load_deployed(os.path.join(golden_dir, 'cp0_model.pt'), device='cpu', lib=lib)


def _periodic_ghost_graph(numbers, pos, cell, rc):
    """Owned atoms = the cell, ghosts = the periodic images that appear as neighbours (what LAMMPS hands to
    pair_e3gnn_parallel on one rank).  Returns edge_index over [owned | ghost], edge_vec, ghost -> owner map."""
    i, j, Dv, Sh = S.neighbor_list_bruteforce(pos, cell, [True] * 3, rc)
    n = len(numbers)
    ghost_of, owner, col = {}, [], np.empty_like(j)
    for e in range(len(i)):
        if not Sh[e].any():
            col[e] = j[e]
            continue
        key = (int(j[e]),) + tuple(int(s) for s in Sh[e])
        if key not in ghost_of:
            ghost_of[key] = n + len(owner)
            owner.append(int(j[e]))
        col[e] = ghost_of[key]
    return np.stack([i, col]), Dv.astype(np.float32), np.array(owner, dtype=np.int64)


@pytest.mark.parametrize('name', ['test-so3-l2', 'cp0'])
def test_deployed_parallel_segments_and_comm_loop(name, golden_dir, tmp_path):
    if name == 'cp0':
        om, cfg = S.load_checkpoint(os.path.join(golden_dir, 'cp0_model.pt'))
        cp = torch.load(os.path.join(golden_dir, 'cp0_model.pt'))
        model_in = (cp['config'], cp['model_state_dict'])
    else:
        cfg = dict(presets.PRESETS[name])
        torch.manual_seed(11)
        om = S.build_model(cfg)
        model_in = (cfg, om.state_dict())
    om.eval()
    _emu_for(cfg)
    numbers, pos, cell = random_cell(14, list(cfg['_type_map']), density=0.05, seed=2)
    ref = om(S.atoms_to_graph(numbers, pos, cell, [True] * 3, cfg['cutoff']))
    d = D.deploy_parallel(model_in, str(tmp_path / 'deployed_parallel'))
    n_layers = cfg['num_convolution_layer']
    assert sorted(os.listdir(d)) == [f'deployed_parallel_{k}.pt' for k in range(n_layers)]
    par, extra = D.load_deployed(d, device='cpu')
    assert len(par) == n_layers and int(extra['comm_size']) == max(c._comm_size for seg in par.segments
                                                                   for c in seg if hasattr(c, '_comm_size'))
    # segment structure of model_build.py:299-367
    names0 = [k for k, _ in par[0].named_children()]
    assert names0[:5] == ['edge_embedding', 'onehot_idx_to_onehot', 'onehot_to_feature_x', 'one_hot_ghost',
                          'ghost_onehot_to_feature_x']
    assert 'ghost_0_self_interaction_1' in names0 and names0[-1] == '1_self_interaction_1'
    names1 = [k for k, _ in par[1].named_children()]
    assert names1[0] == 'edge_embedding' and names1[1] == '1_convolution'
    assert names1[-1] == ('2_self_interaction_1' if n_layers > 2 else 'reduce_total_enegy')
    assert 'force_output' not in [k for k, _ in par[-1].named_children()]

    ei, ev, owner = _periodic_ghost_graph(numbers, pos, cell, cfg['cutoff'])
    n = len(numbers)
    tmap = cfg['_type_map']
    types = torch.tensor([tmap[int(z)] for z in numbers], dtype=torch.int64)
    own_t = torch.tensor(owner)
    calls = {'fwd': 0, 'rev': 0}

    def forward_comm(x_local):
        calls['fwd'] += 1
        return x_local[own_t]

    def reverse_comm(dx_local, dx_ghost):
        calls['rev'] += 1
        return dx_local.clone().index_add_(0, own_t, dx_ghost)

    E, dE_dr, ae = par.compute(types, types[own_t], torch.tensor(ei), torch.tensor(ev), forward_comm, reverse_comm)
    assert calls == {'fwd': n_layers - 1, 'rev': n_layers - 1}
    assert abs(E.item() - ref[S.E_TOT].item()) / n < 1e-5
    assert (ae.reshape(-1) - ref[S.AE].detach().reshape(-1)).abs().max() < 1e-4
    # forces the way the pair style assembles them (pair_e3gnn_parallel.cpp:456-470), ghosts folded onto owners
    F = torch.zeros(n + len(owner), 3).index_add_(0, torch.tensor(ei[0]), dE_dr).index_add_(
        0, torch.tensor(ei[1]), -dE_dr)
    F = F[:n].index_add_(0, own_t, F[n:])
    assert (F - ref[S.FORCE].detach()).abs().max() < 1e-4


def test_deployed_parallel_against_reference_parallel_artefacts(golden_dir):
    """tests/golden/deployed_parallel_hfo2.npz holds what the reference's own deployed_parallel_{0..3}.pt produced
    under the pair style's loop (make_deployed_parallel_golden.py): the ghost payload before every segment, the
    energy, dE/d edge_vec and the atomic energies.  Same weights (transplant), same graph -> same numbers."""
    cfg, sd, numbers, pos, cell, _ = deployed_serial_case(golden_dir)
    gold = np.load(os.path.join(golden_dir, 'deployed_parallel_hfo2.npz'))
    _emu_for(cfg)
    _, sd_now, cfg_now = D._resolve((cfg, sd))       # 0.8.6 column order -> sorted instruction order
    par = D.DeployedParallel(cfg_now, sd_now, device='cpu')
    assert int(par.extra_files['comm_size']) == 32                 # the artefact's own `comm_size` extra file
    own = torch.tensor(gold['owner'])
    types = torch.tensor([cfg['_type_map'][int(z)] for z in numbers], dtype=torch.int64)
    rec = {}
    E, dE_dr, ae = par.compute(types, types[own], torch.tensor(gold['edge_index']), torch.tensor(gold['edge_vec']),
                               lambda x: x[own], lambda dx, dg: dx.clone().index_add_(0, own, dg), record=rec)
    for k in (1, 2, 3):
        assert rec[f'x_{k}'].shape == gold[f'x_{k}'].shape
        assert np.abs(rec[f'x_{k}'].numpy() - gold[f'x_{k}']).max() < 2e-5, k
    assert abs(E.item() - float(gold['energy'])) / len(numbers) < 1e-5
    assert np.abs(ae.numpy() - gold['atomic_energy']).max() < 1e-5
    assert np.abs(dE_dr.numpy() - gold['dE_dr']).max() < 1e-4


def test_deploy_serial_freezes_a_modality(tmp_path):
    """scripts/deploy.py:31-36: a multi-modal checkpoint is deployed for ONE modality; the bundle carries it and the
    deployed model reproduces the oracle evaluated with that modality."""
    cfg = dict(presets.PRESETS['test-so3-l2'])
    cfg.update(use_modality=True, _number_of_modalities=2, _modal_map={'pbe': 0, 'scan': 1},
               use_modal_node_embedding=True, use_modal_self_inter_intro=True, use_modal_self_inter_outro=True,
               use_modal_output_block=True, use_modal_wise_shift=True, use_modal_wise_scale=True,
               shift=[[0.1, -0.2, 0.3], [0.5, 0.4, -0.1]], scale=[[1.0, 1.5, 0.7], [0.9, 1.1, 1.3]])
    torch.manual_seed(3)
    ref = S.build_model(cfg).eval()
    lib = _emu_for(cfg)
    with pytest.raises(ValueError):
        D.deploy((cfg, ref.state_dict()), str(tmp_path / 'x'))                    # modality missing
    with pytest.raises(ValueError):
        D.deploy((dict(presets.PRESETS['test-so3-l2']), S.build_model(dict(presets.PRESETS['test-so3-l2'])).state_dict()),
                 str(tmp_path / 'y'), modal='pbe')                                 # model has no modal_map
    path = D.deploy((cfg, ref.state_dict()), str(tmp_path / 'mf'), modal='scan')
    model, _ = D.load_deployed(path, device='cpu', lib=lib)
    assert model.engine.modal == 1
    numbers, pos, cell = random_cell(16, list(cfg['_type_map']), density=0.05, seed=8)
    g = S.atoms_to_graph(numbers, pos, cell, [True] * 3, cfg['cutoff'])
    g[S.MODAL_TYPE] = torch.tensor(1)
    want = ref(dict(g))
    ei, sh = g[S.EDGE_IDX], g[S.SHIFT]
    out = model({'x': torch.tensor([cfg['_type_map'][int(z)] for z in numbers], dtype=torch.int64),
                 'pos': torch.tensor(pos, dtype=torch.float32, requires_grad=True), 'edge_index': ei,
                 'num_atoms': torch.tensor([len(numbers)], dtype=torch.int64),
                 'cell_lattice_vectors': torch.tensor(cell, dtype=torch.float32),
                 'cell_volume': torch.tensor(abs(np.linalg.det(cell)), dtype=torch.float32),
                 'pbc_shift': sh.to(torch.float32)})
    assert abs(out['inferred_total_energy'].item() - float(want[S.E_TOT].detach())) / len(numbers) < 1e-5
    assert (out['inferred_force'] - want[S.FORCE].detach()).abs().max().item() < 1e-4

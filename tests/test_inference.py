"""sevennet_exp_b200.inference: the reference's on-disk inference format (sevenn/scripts/inference.py:16-220,
tests/data/inferences/snet0_on_hfo2/{per_graph,per_atom}.csv, errors.txt) and its error metrics
(sevenn/error_recorder.py:14-57,155-217), pinned on the test_train.py:284-316 goldens of cp_0.pth on hfo2.extxyz."""
import csv
import os

import numpy as np
import pytest
import torch

from sevennet_exp_b200 import inference as inf

# header lines of the reference's golden files tests/data/inferences/snet0_on_hfo2/per_graph.csv / per_atom.csv
PER_GRAPH_HEADER = ('num_atoms,user_label,total_energy,inferred_total_energy,stress_xx,stress_yy,stress_zz,stress_xy,'
                    'stress_yz,stress_zx,inferred_stress_xx,inferred_stress_yy,inferred_stress_zz,inferred_stress_xy,'
                    'inferred_stress_yz,inferred_stress_zx,cell_lattice_vectors_a,cell_lattice_vectors_b,'
                    'cell_lattice_vectors_c')
PER_ATOM_HEADER = ('stct_id,atom_id,atomic_numbers,atomic_energy,pos_x,pos_y,pos_z,force_of_atoms_x,force_of_atoms_y,'
                   'force_of_atoms_z,inferred_force_x,inferred_force_y,inferred_force_z')
ERROR_KEYS = ['Energy_RMSE (eV/atom)', 'Force_RMSE (eV/Å)', 'Stress_RMSE (kbar)', 'Energy_MAE (eV/atom)',
              'Force_MAE (eV/Å)', 'Stress_MAE (kbar)']
SYM = {72: 'Hf', 8: 'O'}


def write_hfo2_extxyz(golden_dir, path):
    """The two labelled frames of the reference's hfo2.extxyz, re-serialised from tests/golden/hfo2_frames.npz."""
    fr = np.load(os.path.join(golden_dir, 'hfo2_frames.npz'))
    with open(path, 'w') as f:
        for k in range(2):
            num, pos, cell, frc = (fr[f'f{k}_{n}'] for n in ('numbers', 'pos', 'cell', 'forces'))
            lat = ' '.join(repr(float(v)) for v in cell.reshape(-1))
            st = ' '.join(repr(float(v)) for v in fr[f'f{k}_stress3x3'].reshape(-1))
            f.write(f'{len(num)}\n')
            f.write(f'Lattice="{lat}" Properties=species:S:1:pos:R:3:forces:R:3 energy={float(fr[f"f{k}_energy"])!r} '
                    f'stress="{st}" free_energy={float(fr[f"f{k}_energy"])!r} config_type=test{k} pbc="T T T"\n')
            for z, p, g in zip(num, pos, frc):
                f.write(f'{SYM[int(z)]} ' + ' '.join(repr(float(v)) for v in list(p) + list(g)) + '\n')
    return fr


class OracleCalculator:
    """Stand-in with the calculator's compute_batch contract, answered by the oracle (host logic test only)."""

    def __init__(self, golden_dir):
        from oracle import sevenn_ref as S
        self.S = S
        self.model, self.cfg = S.load_checkpoint(os.path.join(golden_dir, 'cp0_model.pt'))
        self.model.eval()

    def compute_batch(self, structures):
        S, out = self.S, []
        for numbers, pos, cell, pbc in structures:
            g = S.atoms_to_graph(numbers, pos, cell, pbc, self.cfg['cutoff'])
            r = self.model(dict(g))
            sm = r[S.STRESS].detach().numpy().astype(np.float64).reshape(6)
            out.append({'energy': float(r[S.E_TOT].detach()), 'forces': r[S.FORCE].detach().numpy().astype(np.float64),
                        'inferred_stress': sm, 'stress': -sm[[0, 1, 2, 4, 5, 3]],
                        'energies': r[S.AE].detach().numpy().reshape(-1).astype(np.float64)})
        return out


def check_outputs(out_dir, errors, fr):
    assert list(errors) == ERROR_KEYS
    assert abs(errors['Energy_RMSE (eV/atom)'] - 28.977758) < 2e-5          # test_train.py:284-316
    assert abs(errors['Force_RMSE (eV/Å)'] - 0.214107) < 1e-5
    assert abs(errors['Stress_RMSE (kbar)'] - 190.014237) < 1e-3
    lines = open(os.path.join(out_dir, 'errors.txt'), encoding='utf-8').read().splitlines()
    assert [ln.split(': ')[0] for ln in lines] == ERROR_KEYS
    assert abs(float(lines[0].split(': ')[1]) - errors['Energy_RMSE (eV/atom)']) < 1e-12
    assert open(os.path.join(out_dir, 'per_graph.csv')).readline().strip() == PER_GRAPH_HEADER
    assert open(os.path.join(out_dir, 'per_atom.csv')).readline().strip() == PER_ATOM_HEADER
    pg = list(csv.DictReader(open(os.path.join(out_dir, 'per_graph.csv'))))
    pa = list(csv.DictReader(open(os.path.join(out_dir, 'per_atom.csv'))))
    assert len(pg) == 2 and len(pa) == sum(len(fr[f'f{k}_numbers']) for k in range(2))
    for k in range(2):
        assert int(pg[k]['num_atoms']) == len(fr[f'f{k}_numbers']) and pg[k]['user_label'] == '-'
        assert abs(float(pg[k]['total_energy']) - float(fr[f'f{k}_energy'])) < 1e-4
        s = fr[f'f{k}_stress3x3']
        assert abs(float(pg[k]['stress_xx']) - (-s[0, 0] * inf.KBAR)) < 1e-3          # label sign / unit of the reference
        assert abs(float(pg[k]['stress_zx']) - (-s[0, 2] * inf.KBAR)) < 1e-3
        assert pg[k]['cell_lattice_vectors_a'].startswith('[')
    # per-atom rows: forces label and prediction reproduce the Force metrics
    lab = np.array([[float(r[f'force_of_atoms_{c}']) for c in 'xyz'] for r in pa])
    pred = np.array([[float(r[f'inferred_force_{c}']) for c in 'xyz'] for r in pa])
    assert abs(np.sqrt(((pred - lab) ** 2).sum(1).mean()) - errors['Force_RMSE (eV/Å)']) < 1e-5
    assert abs(np.abs(pred - lab).mean() - errors['Force_MAE (eV/Å)']) < 1e-5
    assert [int(r['stct_id']) for r in pa] == [0] * len(fr['f0_numbers']) + [1] * len(fr['f1_numbers'])
    info = list(csv.DictReader(open(os.path.join(out_dir, 'info.csv'))))
    assert [r['config_type'] for r in info] == ['test0', 'test1'] and info[0]['file'].endswith('hfo2.extxyz')


def test_read_extxyz_round_trip(golden_dir, tmp_path):
    fr = write_hfo2_extxyz(golden_dir, str(tmp_path / 'hfo2.extxyz'))
    frames = inf.read_extxyz(str(tmp_path / 'hfo2.extxyz'))
    assert len(frames) == 2
    for k, f in enumerate(frames):
        assert np.array_equal(f['numbers'], fr[f'f{k}_numbers'])
        assert np.array_equal(f['positions'], fr[f'f{k}_pos']) and np.array_equal(f['cell'], fr[f'f{k}_cell'])
        assert np.array_equal(f['forces'], fr[f'f{k}_forces']) and f['energy'] == float(fr[f'f{k}_energy'])
        assert np.array_equal(f['stress'], fr[f'f{k}_stress3x3']) and f['pbc'].all()
        assert f['info'] == {'config_type': f'test{k}'}


def test_inference_outputs_with_oracle_backend(golden_dir, tmp_path):
    fr = write_hfo2_extxyz(golden_dir, str(tmp_path / 'hfo2.extxyz'))
    out = str(tmp_path / 'out')
    errors = inf.inference(OracleCalculator(golden_dir), str(tmp_path / 'hfo2.extxyz'), out, batch_size=1)
    check_outputs(out, errors, fr)
    # unlabeled targets: rejected by default, NaN labels and metrics over the labelled ones otherwise
    frames = inf.read_extxyz(str(tmp_path / 'hfo2.extxyz'))
    del frames[1]['energy'], frames[1]['forces'], frames[1]['stress']
    with pytest.raises(ValueError):
        inf.inference(OracleCalculator(golden_dir), frames, str(tmp_path / 'out2'))
    e2 = inf.inference(OracleCalculator(golden_dir), frames, str(tmp_path / 'out2'), allow_unlabeled=True)
    rec = inf.evaluate_targets(OracleCalculator(golden_dir), frames[:1], 4)
    assert abs(e2['Force_RMSE (eV/Å)'] - inf.error_metrics(rec)['Force_RMSE (eV/Å)']) < 1e-12
    with pytest.raises(ValueError):
        inf.inference(OracleCalculator(golden_dir), frames, out, modal='x', allow_unlabeled=True)


@pytest.mark.gpu
def test_inference_cp0_hfo2_on_gpu(golden_dir, tmp_path):
    """The same files through the CUDA path (SevenNetCalculator.compute_batch, both frames in one batch)."""
    fr = write_hfo2_extxyz(golden_dir, str(tmp_path / 'hfo2.extxyz'))
    cp = torch.load(os.path.join(golden_dir, 'cp0_model.pt'))
    out = str(tmp_path / 'out')
    errors = inf.inference((cp['config'], cp['model_state_dict']), str(tmp_path / 'hfo2.extxyz'), out, batch_size=4)
    check_outputs(out, errors, fr)

"""Training step on the GPU (SURVEY 8f-1 / BASELINE config 5) through the C ABI: parameter gradients vs the oracle under
autograd, and the reference's one-Adam-step known answer (reference tests/unit_tests/test_train.py:284-316)."""
import os

import numpy as np
import pytest
import torch

from oracle import sevenn_ref as S
from sevennet_exp_b200 import nn as bnn
from sevennet_exp_b200 import presets, train as TR
from tests.structures import random_cell
from tests.test_train_emu import compare_grads, labelled_graphs, oracle_grads, ours_grads

pytestmark = pytest.mark.gpu
CONFIG = {'force_loss_weight': 1.0, 'stress_loss_weight': 0.1, 'is_train_stress': True, 'optim_param': {}}


@pytest.mark.parametrize('name', ['test-so3-l2', 'test-so3-l3', '7net-0'])
def test_parameter_gradients_match_oracle_on_gpu(name):
    cfg = dict(presets.PRESETS[name])
    cfg.update(version='0.11.0', train_shift_scale=True, train_denominator=True)
    torch.manual_seed(5)
    om = S.build_model(cfg)
    rng = np.random.default_rng(0)
    if name == '7net-0':
        structs = [random_cell(n, [14, 8, 72], density=0.05, seed=40 + k) for k, n in enumerate((24, 16))]
    else:
        structs = [random_cell(n, list(cfg['_type_map']), density=0.05, seed=30 + k) for k, n in enumerate((40, 23))]
    labels = [(rng.normal(), rng.normal(size=(len(s[0]), 3)), rng.normal(size=6) * 1e-3) for s in structs]
    graphs = labelled_graphs(cfg, structs, labels)
    lo, go = oracle_grads(cfg, om, graphs, CONFIG)
    lb, gb, model = ours_grads(cfg, om.state_dict(), graphs, CONFIG, device='cuda')
    assert abs(lo - lb) < 1e-4 * max(1.0, abs(lo))
    worst = compare_grads(go, gb)
    print(name, 'worst relative gradient error', max(worst.values()))


def test_cp0_one_adam_step_known_answer_on_gpu(golden_dir):
    om, cfg = S.load_checkpoint(os.path.join(golden_dir, 'cp0_model.pt'))
    fr = np.load(os.path.join(golden_dir, 'hfo2_frames.npz'))
    structs, labels = [], []
    for k in range(2):
        num, pos, cell = (fr[f'f{k}_{n}'] for n in ('numbers', 'pos', 'cell'))
        s = fr[f'f{k}_stress3x3']
        voigt = np.array([s[0, 0], s[1, 1], s[2, 2], s[1, 2], s[0, 2], s[0, 1]])
        structs.append((num, pos, cell))
        labels.append((float(fr[f'f{k}_energy']), fr[f'f{k}_forces'], -voigt[[0, 1, 2, 5, 3, 4]]))
    graphs = labelled_graphs(cfg, structs, labels)
    model = bnn.build_E3_equivariant_model(cfg)
    assert not model.load_state_dict(om.state_dict(), strict=False)[0]
    tr = TR.Trainer(model.cuda(), CONFIG)
    batch = TR.collate(graphs, device='cuda')
    m1 = tr.evaluate(batch)
    assert np.allclose([m1['Energy_RMSE'], m1['Force_RMSE'], m1['Stress_RMSE']], [28.977758, 0.214107, 190.014237], rtol=1e-5)
    tr.train_step(batch)
    m2 = tr.evaluate(batch)
    print(m1, m2)
    assert np.allclose([m2['Energy_RMSE'], m2['Force_RMSE'], m2['Stress_RMSE']], [28.977878, 0.213105, 188.772557], rtol=1e-5)
    # a second step keeps decreasing the loss it optimises
    l1 = tr.train_step(batch)
    l2 = tr.train_step(batch)
    assert l2 < l1

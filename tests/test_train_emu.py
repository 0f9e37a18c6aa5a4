"""Training step through the emulation build: parameter gradients of the force/stress loss -- i.e. the
double backward composed from the first-order kernels (nn/train_functional.py) plus csrc/train.cu -- against the
oracle under torch autograd, and the reference's own known answer for one Adam step
(reference tests/unit_tests/test_train.py:284-316, second half: 28.977878 / 0.213105 / 188.772557)."""
import os

import numpy as np
import pytest
import torch

from oracle import sevenn_ref as S
from sevennet_exp_b200 import nn as bnn
from sevennet_exp_b200 import presets, spec as sp, train as TR
from tests.emu_util import build_full_emu
from tests.structures import random_cell


@pytest.fixture(autouse=True)
def _reset_lib():
    yield
    bnn.set_lib(None)


def labelled_graphs(cfg, structures, labels):
    gs = []
    for (numbers, pos, cell), (e, f, s) in zip(structures, labels):
        g = S.atoms_to_graph(numbers, pos, cell, [True] * 3, cfg['cutoff'])
        g[TR.K_E] = torch.tensor(float(e), dtype=torch.float32)
        g[TR.K_F] = torch.as_tensor(f, dtype=torch.float32)
        g[TR.K_S] = torch.as_tensor(s, dtype=torch.float32)
        gs.append(g)
    return gs


def oracle_grads(cfg, om, graphs, config):
    om = om.train()
    om.set_is_batch_data(True)
    for p in om.parameters():
        p.grad = None
    out = om(TR.collate(graphs, device='cpu'))
    loss = TR.loss_from_config(out, config)
    loss.backward()
    return float(loss), {n: p.grad.clone() for n, p in om.named_parameters() if p.requires_grad and p.grad is not None}


def ours_grads(cfg, sd, graphs, config, device='cpu'):
    model = bnn.build_E3_equivariant_model(cfg)
    missing, _ = model.load_state_dict(sd, strict=False)
    assert not missing
    model = model.to(device).train()
    model.set_is_batch_data(True)
    out = model(TR.collate(graphs, device=device))
    loss = TR.loss_from_config(out, config)
    loss.backward()
    return float(loss), {n: p.grad.detach().cpu().clone() for n, p in model.named_parameters()
                         if p.requires_grad and p.grad is not None}, model


def compare_grads(go, gb, rtol=2e-3):
    assert set(go) == set(gb), (sorted(set(go) ^ set(gb)))
    worst = {}
    for n in go:
        scale = go[n].abs().max().item()
        err = (go[n] - gb[n]).abs().max().item()
        worst[n] = err / max(scale, 1e-12)
        assert err <= rtol * scale + 1e-7, (n, err, scale)
    return worst


@pytest.mark.parametrize('name', ['test-so3-l2'])
def test_parameter_gradients_match_oracle_autograd(name):
    cfg = dict(presets.PRESETS[name])
    cfg.update(version='0.11.0', train_shift_scale=True, train_denominator=True)
    torch.manual_seed(5)
    om = S.build_model(cfg)
    spec = sp.build_spec(cfg)
    bnn.set_lib(build_full_emu([L.conv for L in spec.layers]))
    rng = np.random.default_rng(0)
    structs = [random_cell(n, list(cfg['_type_map']), density=0.05, seed=30 + k) for k, n in enumerate((10, 7))]
    labels = [(rng.normal(), rng.normal(size=(len(s[0]), 3)), rng.normal(size=6) * 1e-3) for s in structs]
    graphs = labelled_graphs(cfg, structs, labels)
    config = {'force_loss_weight': 1.0, 'stress_loss_weight': 0.1, 'is_train_stress': True}
    lo, go = oracle_grads(cfg, om, graphs, config)
    lb, gb, _ = ours_grads(cfg, om.state_dict(), graphs, config)
    assert abs(lo - lb) < 1e-4 * max(1.0, abs(lo))
    worst = compare_grads(go, gb)
    print({k: f'{v:.1e}' for k, v in worst.items()})
    assert 'edge_embedding.basis_function.coeffs' in gb and '1_convolution.denominator' in gb


def test_cp0_one_adam_step_known_answer(golden_dir):
    """reference test_train.py:284-316: cp_0.pth on both hfo2.extxyz frames (one batch of 2), eval -> train one step
    with Adam (checkpoint config: lr default, MSE, force weight 1.0, stress weight 0.1) -> eval."""
    om, cfg = S.load_checkpoint(os.path.join(golden_dir, 'cp0_model.pt'))
    spec = sp.build_spec(cfg)
    bnn.set_lib(build_full_emu([L.conv for L in spec.layers]))
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
    config = {'force_loss_weight': 1.0, 'stress_loss_weight': 0.1, 'is_train_stress': True, 'optim_param': {}}
    tr = TR.Trainer(model, config)
    batch = TR.collate(graphs, device='cpu')
    m1 = tr.evaluate(batch)
    assert abs(m1['Energy_RMSE'] - 28.977758) < 3e-5 and abs(m1['Force_RMSE'] - 0.214107) < 1e-5
    assert abs(m1['Stress_RMSE'] - 190.014237) < 2e-3
    tr.train_step(batch)
    m2 = tr.evaluate(batch)
    print(m1, m2)
    assert np.allclose(m2['Energy_RMSE'], 28.977878, rtol=1e-5)       # the reference's own comparison (np.allclose)
    assert np.allclose(m2['Force_RMSE'], 0.213105, rtol=1e-5)
    assert np.allclose(m2['Stress_RMSE'], 188.772557, rtol=1e-5)

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


def test_device_built_tc_weight_image_equals_host_image():
    """sevenn_gemm_tc_image (weights that change every step) builds exactly the image engine.tc_image_for builds on the host."""
    import ctypes
    from sevennet_exp_b200 import _lib
    from sevennet_exp_b200.engine import tc_image_for
    lib = _lib.load()
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    torch.manual_seed(3)
    for K, N in [(8, 64), (64, 960), (352, 32), (224, 224), (64, 1), (12, 20)]:
        W = torch.randn(K, N, device='cuda')
        host = torch.from_numpy(tc_image_for(lib, W.cpu().numpy()))
        img = torch.full((host.numel(),), float('nan'), device='cuda')
        lib.call('sevenn_gemm_tc_image', ctypes.c_void_p(W.data_ptr()), K, N, ctypes.c_void_p(img.data_ptr()), st)
        assert torch.equal(img.cpu(), host), (K, N)


def test_graphed_training_step_equals_eager():
    """Trainer.train_step_graphed (the whole step replayed from a CUDA graph, Adam's step count on the device) walks the
    same parameter trajectory as the eager step."""
    cfg = dict(presets.PRESETS['test-so3-l2'])
    cfg.update(version='0.11.0')
    rng = np.random.default_rng(1)
    structs = [random_cell(n, list(cfg['_type_map']), density=0.05, seed=50 + k) for k, n in enumerate((30, 21, 12))]
    labels = [(rng.normal(), rng.normal(size=(len(s[0]), 3)), rng.normal(size=6) * 1e-3) for s in structs]
    graphs = labelled_graphs(cfg, structs, labels)
    torch.manual_seed(9)
    om = S.build_model(cfg)
    finals = {}
    for mode in ('eager', 'graph'):
        model = bnn.build_E3_equivariant_model(cfg)
        assert not model.load_state_dict(om.state_dict(), strict=False)[0]
        tr = TR.Trainer(model.cuda(), CONFIG)
        batch = tr.prepare(TR.collate(graphs, device='cuda'))
        losses = []
        for _ in range(5):
            losses.append(float(tr.train_step(batch) if mode == 'eager' else tr.train_step_graphed(batch)))
        finals[mode] = (losses, {n: p.detach().cpu().clone() for n, p in model.named_parameters()})
    # the graphed trainer runs two extra warm-up steps before capturing: compare step k of eager with the same parameters
    le, lg = finals['eager'][0], finals['graph'][0]
    assert lg[0] < le[0] or abs(lg[0] - le[2]) < 1e-4 * abs(le[2])       # graph's first replayed step = eager's third
    assert all(b <= a + 1e-6 for a, b in zip(lg, lg[1:]))                # and the loss keeps going down
    assert abs(lg[0] - le[2]) < 1e-4 * abs(le[2]) and abs(lg[1] - le[3]) < 1e-4 * abs(le[3])

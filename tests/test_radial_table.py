"""Tabulated radial function (sevennet_exp_b200/radial_table.py) against the oracle's radial MLP: value and
derivative of the cubic-Hermite interpolant at random lengths, both envelopes."""
import math

import numpy as np
import pytest
import torch

from oracle import sevenn_ref as S
from sevennet_exp_b200 import irreps as ir, presets, radial_table as rt


@pytest.mark.parametrize('cutoff_cfg', [{'cutoff_function_name': 'poly_cut', 'poly_cut_p_value': 6},
                                        {'cutoff_function_name': 'XPLOR', 'cutoff_on': 4.5}])
def test_table_matches_oracle_radial_mlp(cutoff_cfg):
    cfg = dict(presets.PRESETS['test-so3-l2'])
    cfg['cutoff'] = 5.0
    cfg['cutoff_function'] = cutoff_cfg
    torch.manual_seed(5)
    model = S.build_model(cfg).double().eval()
    with torch.no_grad():
        model.edge_embedding.basis_function.coeffs.mul_(1.02)
    name = '1_convolution'
    conv = dict(model.named_modules())[name]
    kind = cutoff_cfg['cutoff_function_name']
    r_on = float(cutoff_cfg.get('cutoff_on', 0.0))
    n_int = rt.choose_intervals(5.0, kind, r_on)
    if kind == 'XPLOR':
        assert abs(r_on * n_int / 5.0 - round(r_on * n_int / 5.0)) < 1e-9      # cutoff_on is a node
    sd = {k: v.detach().numpy() for k, v in model.state_dict().items()}
    hid = [cfg['radial_basis']['bessel_basis_num']] + list(cfg['weight_nn_hidden_neurons'])
    Ws = [sd[f'{name}.weight_nn.layer{k}.weight'] / math.sqrt(hid[k]) for k in range(len(hid))]
    tab = rt.build_table(sd['edge_embedding.basis_function.coeffs'], kind, 5.0, 6.0, r_on, Ws, cfg['act_radial'],
                         ir.normalize2mom_const(cfg['act_radial']), n_int)
    assert tab.dtype == np.float32 and tab.shape[0] == n_int and tab.shape[2] == 4
    rng = np.random.default_rng(0)
    r = np.concatenate([rng.uniform(0.6, 5.0 - 1e-9, 4000), [4.5 - 1e-7, 4.5, 4.5 + 1e-7, 4.999999]])
    w, dw = rt.hermite_eval(tab, r, 5.0)
    rt_ = torch.tensor(r, dtype=torch.float64, requires_grad=True)
    emb = model.edge_embedding.basis_function(rt_) * model.edge_embedding.cutoff_function(rt_).unsqueeze(-1)
    w_ref = conv.weight_nn(emb)
    assert np.abs(w - w_ref.detach().numpy()).max() < 2e-7 * max(1.0, float(w_ref.abs().max()))
    cols = rng.choice(w_ref.shape[1], 6, replace=False)
    for c in cols:
        (g,) = torch.autograd.grad(w_ref[:, c].sum(), rt_, retain_graph=True)
        assert np.abs(dw[:, c] - g.numpy()).max() < 2e-5 * max(1.0, float(g.abs().max()))

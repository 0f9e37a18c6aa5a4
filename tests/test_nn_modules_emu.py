"""sevennet_exp_b200.nn (the reference's module seam) against the oracle modules of the same name:
same state_dict, same data dict, arbitrary edge order.  Kernels run through the emulation build."""
import os

import numpy as np
import pytest
import torch

from oracle import sevenn_ref as S
from sevennet_exp_b200 import nn as bnn
from sevennet_exp_b200 import presets, spec as sp
from sevennet_exp_b200.checkpoint import patch_state_dict
from tests.emu_util import build_full_emu
from tests.structures import random_cell

KBAR = 1602.1766208


@pytest.fixture(autouse=True)
def _reset_lib():
    yield
    bnn.set_lib(None)


def _shuffled(g, seed=0):
    """The reference accepts edges in any order; so must the drop-in."""
    E = g[S.EDGE_IDX].shape[1]
    p = torch.from_numpy(np.random.default_rng(seed).permutation(E))
    g = dict(g)
    g[S.EDGE_IDX] = g[S.EDGE_IDX][:, p].contiguous()
    g[S.EDGE_VEC] = g[S.EDGE_VEC].detach()[p].contiguous()
    return g


def _run_pair(cfg, sd, oracle_model, numbers, pos, cell, shuffle=True):
    spec = sp.build_spec(cfg)
    bnn.set_lib(build_full_emu([L.conv for L in spec.layers]))
    model = bnn.build_E3_equivariant_model(cfg).eval()
    missing, unexpected = model.load_state_dict(sd, strict=False)
    assert not missing, missing                    # sevenn/checkpoint.py:306-310 contract
    g = S.atoms_to_graph(numbers, pos, cell, [True] * 3, cfg['cutoff'])
    if shuffle:
        g = _shuffled(g)
    ref = oracle_model({k: (v.clone() if torch.is_tensor(v) else v) for k, v in g.items()})
    out = model({k: (v.clone() if torch.is_tensor(v) else v) for k, v in g.items()})
    n = len(numbers)
    dE = abs(out[S.E_TOT].item() - ref[S.E_TOT].item()) / n
    dF = (out[S.FORCE] - ref[S.FORCE].detach()).abs().max().item()
    dS = (out[S.STRESS] - ref[S.STRESS].detach()).abs().max().item() * KBAR
    dA = (out[S.AE] - ref[S.AE].detach()).abs().max().item()
    assert dE < 1e-5 and dF < 1e-4 and dA < 1e-4, (dE, dF, dA)
    assert dS < 1e-3 * max(1.0, ref[S.STRESS].abs().max().item() * KBAR * 1e-3), dS
    assert '_b200_graph' not in out
    return model


@pytest.mark.parametrize('name', ['test-so3-l2'])
def test_sequential_matches_oracle_random_weights(name):
    cfg = dict(presets.PRESETS[name])
    torch.manual_seed(777)
    om = S.build_model(cfg).eval()
    numbers, pos, cell = random_cell(20, list(cfg['_type_map']), density=0.05, seed=5)
    model = _run_pair(cfg, om.state_dict(), om, numbers, pos, cell)
    # module names and order are the reference's (state_dict prefixes)
    assert [n for n, _ in model.named_children()] == [n for n, _ in om.named_children()]


def test_sequential_cp0_checkpoint_o3_nequip_sc(golden_dir):
    om, cfg = S.load_checkpoint(os.path.join(golden_dir, 'cp0_model.pt'))
    om.eval()
    cp = torch.load(os.path.join(golden_dir, 'cp0_model.pt'))
    sd = patch_state_dict(sp.build_spec(cfg), cp['model_state_dict'], cfg.get('version', '0.0.0'))
    fr = np.load(os.path.join(golden_dir, 'hfo2_frames.npz'))
    num, pos, cell = (fr[f'f0_{n}'] for n in ('numbers', 'pos', 'cell'))
    _run_pair(cfg, sd, om, num, pos, cell)


def test_single_modules_match_oracle():
    """Module by module (forward value and gradient w.r.t. the input) on e3nn-layout tensors."""
    cfg = dict(presets.PRESETS['test-so3-l2'])
    spec = sp.build_spec(cfg)
    bnn.set_lib(build_full_emu([L.conv for L in spec.layers]))
    torch.manual_seed(3)
    om = S.build_model(cfg).eval()
    bm = bnn.build_E3_equivariant_model(cfg).eval()
    bm.load_state_dict(om.state_dict(), strict=False)
    numbers, pos, cell = random_cell(12, list(cfg['_type_map']), density=0.05, seed=1)
    g = S.atoms_to_graph(numbers, pos, cell, [True] * 3, cfg['cutoff'])

    def both(name, data_o, data_b, key_in, key_out):
        xo = data_o[key_in].detach().clone().requires_grad_(True)
        xb = data_b[key_in].detach().clone().requires_grad_(True)
        data_o, data_b = dict(data_o), dict(data_b)
        data_o[key_in], data_b[key_in] = xo, xb
        yo = getattr(om, name)(data_o)[key_out]
        yb = getattr(bm, name)(data_b)[key_out]
        assert torch.allclose(yo, yb, atol=2e-5, rtol=1e-4), (name, (yo - yb).abs().max())
        seed = torch.randn_like(yo)
        go, = torch.autograd.grad((yo * seed).sum(), xo)
        gb, = torch.autograd.grad((yb * seed).sum(), xb)
        assert torch.allclose(go, gb, atol=5e-5, rtol=1e-4), (name, (go - gb).abs().max())
        return data_o, data_b

    do = {k: v for k, v in g.items()}
    do[S.X] = om.z_to_onehot_tensor[do[S.Z]]
    db = dict(do)
    # edge embedding: value + d/d(edge_vec)
    for key in (S.EDGE_ATTR, S.EDGE_EMB):
        both('edge_embedding', do, db, S.EDGE_VEC, key)
    do, db = om.edge_embedding(do), bm.edge_embedding(db)
    do, db = om.onehot_idx_to_onehot(do), bm.onehot_idx_to_onehot(db)
    do, db = om.onehot_to_feature_x(do), bm.onehot_to_feature_x(db)
    for t in range(2):
        do, db = both(f'{t}_self_connection_intro', do, db, S.X, S.SC_TMP)
        do, db = both(f'{t}_self_interaction_1', do, db, S.X, S.X)
        do, db = both(f'{t}_convolution', do, db, S.X, S.X)
        do, db = both(f'{t}_self_interaction_2', do, db, S.X, S.X)
        do = getattr(om, f'{t}_self_connection_outro')(do)
        db = getattr(bm, f'{t}_self_connection_outro')(db)
        do, db = both(f'{t}_equivariant_gate', do, db, S.X, S.X)


@pytest.mark.parametrize('cutoff_fn', [{'cutoff_function_name': 'XPLOR', 'cutoff_on': 4.5},
                                       {'cutoff_function_name': 'poly_cut', 'poly_cut_p_value': 6}])
def test_multi_cutoff_edge_embedding_matches_oracle(cutoff_fn):
    """The fork's per-layer cutoffs (edge_embedding.py:220-257, interaction_blocks.py:54-60): every convolution
    runs over the sub-graph of its own cutoff."""
    cfg = dict(presets.PRESETS['test-so3-l2'])
    cfg['cutoff_function'] = cutoff_fn
    cfg['multi_cutoff'] = [5.0, 3.5, 5.0]
    torch.manual_seed(21)
    om = S.build_model(cfg).eval()
    assert [k for k in om.state_dict() if 'basis_functions' in k] == [
        'edge_embedding.basis_functions.0.coeffs', 'edge_embedding.basis_functions.1.coeffs']
    numbers, pos, cell = random_cell(20, list(cfg['_type_map']), density=0.05, seed=7)
    model = _run_pair(cfg, om.state_dict(), om, numbers, pos, cell)
    assert isinstance(model.edge_embedding, bnn.MultiCutoffEdgeEmbedding)


def test_irreps_linear_modality_bias_and_fcn_modules():
    """nn.IrrepsLinear with num_modalities / biases (sevenn/nn/linear.py:14-100) and nn.FCN_e3nn (:145-180) against the
    oracle modules of the same name: same flat weights, values and input gradients."""
    bnn.set_lib(build_full_emu([]))
    torch.manual_seed(4)
    irr_in, irr_out, M = '24x0e+8x1e', '16x0e+8x1e', 3
    ref = S.IrrepsLinear(S.Irreps(irr_in), S.Irreps(irr_out), S.X, biases=True, num_modalities=M)
    with torch.no_grad():
        ref.linear.bias.normal_(0.0, 0.5)
    mod = bnn.IrrepsLinear(irr_in, irr_out, S.X, num_modalities=M, biases=True)
    assert mod.num_modalities == M and mod.linear.weight.shape == ref.linear.weight.shape
    with pytest.raises(ValueError):
        mod.set_num_modalities(2)                       # already instantiated (linear.py:66-67)
    mod.load_state_dict(ref.state_dict())
    n = 40
    x = torch.randn(n, 24 + 24)
    batch = torch.randint(0, 2, (n,))
    onehot = torch.nn.functional.one_hot(torch.tensor([2, 0]), M).float()          # two graphs, two modalities
    for is_batch in (True, False):
        mod._is_batch_data = is_batch
        oh = onehot if is_batch else onehot[:1]
        xa, xb = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
        # the oracle restates the single-structure form: evaluate it per graph
        want = torch.zeros(n, 16 + 24)
        for gidx in range(2 if is_batch else 1):
            sel = (batch == gidx) if is_batch else torch.ones(n, dtype=torch.bool)
            want[sel] = ref({S.X: xa[sel], S.MODAL_ATTR: oh[gidx]})[S.X]
        data = {S.X: xb, 'modal_attr': oh}
        if is_batch:
            data['batch'] = batch
        got = mod(data)[S.X]
        assert (got - want).abs().max().item() < 2e-5
        gsel = torch.randn(n, 16 + 24)
        (ga,) = torch.autograd.grad((want * gsel).sum(), xa)
        (gb,) = torch.autograd.grad((got * gsel).sum(), xb)
        assert (ga - gb).abs().max().item() < 2e-5
    # FullyConnectedNet readout
    for act in ('relu', 'silu'):
        rf = S.FCN_e3nn(S.Irreps('32x0e'), 1, [16, 8], S.ACT[act], S.X, S.SCALED_AE)
        mf = bnn.FCN_e3nn('32x0e', 1, [16, 8], act, S.X, S.SCALED_AE)
        mf.load_state_dict(rf.state_dict())
        xa, xb = (torch.randn(50, 32, generator=torch.Generator().manual_seed(1)).requires_grad_(True) for _ in range(2))
        ya, yb = rf({S.X: xa})[S.SCALED_AE], mf({S.X: xb})[S.SCALED_AE]
        assert (ya - yb).abs().max().item() < 2e-5
        (ga,) = torch.autograd.grad(ya.sum(), xa)
        (gb,) = torch.autograd.grad(yb.sum(), xb)
        assert (ga - gb).abs().max().item() < 2e-5


@pytest.mark.parametrize('variant', ['bias', 'fcn'])
def test_sequential_bias_and_fcn_readout_variants(variant):
    """build_E3_equivariant_model with use_bias_in_linear / readout_as_fcn: same state_dict as the oracle model, same
    E / F / S."""
    cfg = dict(presets.PRESETS['test-so3-l2'])
    if variant == 'bias':
        cfg['use_bias_in_linear'] = True
    else:
        cfg.update(readout_as_fcn=True, readout_fcn_hidden_neurons=[16, 8], readout_fcn_activation='silu')
    torch.manual_seed(9)
    model = S.build_model(cfg).eval()
    with torch.no_grad():
        for n_, p_ in model.named_parameters():
            if n_.endswith('linear.bias'):
                p_.normal_(0.0, 0.3)
    numbers, pos, cell = random_cell(18, list(cfg['_type_map']), density=0.05, seed=13)
    ours = _run_pair(cfg, model.state_dict(), model, numbers, pos, cell)
    assert (set(k for k, v in ours.named_parameters() if v.numel()) ==
            set(k for k, v in model.named_parameters() if v.numel()))


def test_sequential_multi_modal_matches_oracle():
    """build_E3_equivariant_model on a multi-modal config: one_hot_modality + IrrepsLinear(num_modalities) +
    ModalWiseRescale, data['modal_type'] selects the modality (single-structure data, as the calculator feeds it)."""
    cfg = dict(presets.PRESETS['test-so3-l2'])
    cfg.update(use_modality=True, _number_of_modalities=2, _modal_map={'pbe': 0, 'scan': 1},
               use_modal_node_embedding=True, use_modal_self_inter_intro=True, use_modal_self_inter_outro=True,
               use_modal_output_block=True, use_modal_wise_shift=True, use_modal_wise_scale=True,
               shift=[[0.1, -0.2, 0.3], [0.5, 0.4, -0.1]], scale=[[1.0, 1.5, 0.7], [0.9, 1.1, 1.3]])
    torch.manual_seed(17)
    ref = S.build_model(cfg).eval()
    spec = sp.build_spec(cfg)
    bnn.set_lib(build_full_emu([L.conv for L in spec.layers]))
    model = bnn.build_E3_equivariant_model(cfg).eval()
    missing, _ = model.load_state_dict(ref.state_dict(), strict=False)
    assert not missing, missing
    model.set_is_batch_data(False)
    resc = model.rescale_atomic_energy
    assert resc.get_shift(cfg['_type_map'], cfg['_modal_map'])['scan'][15] == pytest.approx(0.4)
    numbers, pos, cell = random_cell(18, list(cfg['_type_map']), density=0.05, seed=14)
    energies = []
    for m in (0, 1):
        g = S.atoms_to_graph(numbers, pos, cell, [True] * 3, cfg['cutoff'])
        g[S.MODAL_TYPE] = torch.tensor(m)
        want = ref({k: (v.clone() if torch.is_tensor(v) else v) for k, v in g.items()})
        got = model({k: (v.clone() if torch.is_tensor(v) else v) for k, v in g.items()})
        assert abs(got[S.E_TOT].item() - want[S.E_TOT].item()) / len(numbers) < 1e-5
        assert (got[S.FORCE] - want[S.FORCE].detach()).abs().max().item() < 1e-4
        assert (got[S.STRESS] - want[S.STRESS].detach()).abs().max().item() * KBAR < 1e-2
        energies.append(got[S.E_TOT].item())
    assert abs(energies[0] - energies[1]) > 1e-3
    with pytest.raises(NotImplementedError):
        bnn.build_E3_equivariant_model(cfg, parallel=True)
    # sequential.py:139-176: modality by name from the data, then frozen for deployment
    assert model.modal_map == cfg['_modal_map']
    model.eval_modal_map = True
    g = S.atoms_to_graph(numbers, pos, cell, [True] * 3, cfg['cutoff'])
    g['data_modality'] = 'scan'
    assert abs(model(g)[S.E_TOT].item() - energies[1]) < 1e-4
    model.prepare_modal_deploy('pbe')
    assert not model.eval_modal_map and next(iter(model._modules)) == 'modal_input_prepare'
    g = S.atoms_to_graph(numbers, pos, cell, [True] * 3, cfg['cutoff'])
    g[S.X] = model._atomic_numbers_to_onehot(g[S.Z])          # _ModalInputPrepare reads data['x'].device
    assert abs(model(g)[S.E_TOT].item() - energies[0]) < 1e-4


def test_grouped_fp32_gemm_equals_single_calls_emulation():
    """sevenn_gemm_f32_grouped (small systems: the blocks of one or two o3.Linear in one launch) against one
    sevenn_gemm_f32 call per problem and against numpy, through the emulation build; more than SEVENN_GEMM_GROUP_MAX
    problems are refused."""
    import ctypes
    from sevennet_exp_b200 import _lib as L
    from tests.emu_util import build_full_emu
    lib = build_full_emu([])
    rng = np.random.default_rng(0)
    n = 11
    probs = []
    for (R, K, N, epi) in [(1, 12, 20, 0), (3, 8, 8, 1), (5, 4, 6, 0)]:
        A = rng.normal(size=(n, R * K + 2)).astype(np.float32)
        B = rng.normal(size=(K, N)).astype(np.float32)
        probs.append((R, K, N, epi, A, B))
    outs = []
    for grouped in (False, True):
        Cs = [np.full((n, R * N), 0.5, dtype=np.float32) for (R, K, N, *_r) in probs]
        gs = [L.GemmArgs(A.ctypes.data, R * K + 2, K, B.ctypes.data, N, C.ctypes.data, R * N, N, n * R, R, N, K, epi, 0, 1.0, None, None)
              for (R, K, N, epi, A, B), C in zip(probs, Cs)]
        if grouped:
            lib.call('sevenn_gemm_f32_grouped', len(gs), (L.GemmArgs * len(gs))(*gs), None)
        else:
            for g in gs:
                lib.call('sevenn_gemm_f32', ctypes.byref(g), None)
        outs.append(Cs)
    for (R, K, N, epi, A, B), a, b in zip(probs, *outs):
        assert np.array_equal(a, b)
        ref = np.einsum('nrk,kc->nrc', A[:, :R * K].reshape(n, R, K).astype(np.float64), B.astype(np.float64)).reshape(n, R * N)
        assert np.abs(a - (ref + (0.5 if epi else 0.0))).max() < 1e-5
    assert lib.cdll.sevenn_gemm_f32_grouped(9, (L.GemmArgs * 9)(), None) != 0

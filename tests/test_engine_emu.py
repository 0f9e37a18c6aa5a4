"""Whole E/F/S pipeline of sevennet_exp_b200.engine (weight packing, layouts, launch order, analytic
reverse pass) against the oracle, run through the EMULATION build of the kernel sources (g++, no GPU).
The GPU parity tests (test_gpu_*.py) repeat this through the real CUDA library."""
import os

import numpy as np
import pytest
import torch

from oracle import sevenn_ref as S
from sevennet_exp_b200 import presets, spec as sp
from sevennet_exp_b200.engine import Engine
from tests.emu_util import build_full_emu, graph_from_edges
from tests.structures import random_cell

KBAR = 1602.1766208


def _compare(cfg, sd, numbers, pos, cell, pbc, oracle_model, tol_e=1e-5, tol_f=1e-4, tol_s=1e-3, modal=None):
    spec = sp.build_spec(cfg)
    lib = build_full_emu([L.conv for L in spec.layers])
    g = S.atoms_to_graph(numbers, pos, cell, pbc, cfg['cutoff'])
    if modal is not None:
        g[S.MODAL_TYPE] = torch.tensor(modal)
    ref = oracle_model(dict(g))
    tmap = cfg['_type_map']
    types = np.array([tmap[int(z)] for z in numbers])
    vol = float(g[S.VOL])
    gr = graph_from_edges(types, g[S.EDGE_IDX].numpy(), g[S.EDGE_VEC].detach().numpy(), len(numbers), vol)
    n = len(numbers)
    res = {}
    for radial in ('mlp', 'table'):     # per-edge radial MLP (GEMM kernels) and the tabulated radial function
        eng = Engine(spec, sd, lib=lib, device='cpu', radial=radial, modal=modal)
        assert eng.radial == radial
        out = eng.evaluate(gr)
        dE = abs(out['energy'].item() - ref[S.E_TOT].item()) / n
        dF = (out['forces'] - ref[S.FORCE].detach()).abs().max().item()
        stress = -(out['virial6'] / vol).float()
        dS = (stress - ref[S.STRESS].detach()).abs().max().item() * KBAR
        assert dE < tol_e, (radial, dE)
        assert dF < tol_f, (radial, dF)
        assert dS < tol_s * max(1.0, ref[S.STRESS].abs().max().item() * KBAR * 1e-3), (radial, dS)
        res[radial] = (dE, dF, dS)
    return res


def test_engine_emu_cp0_hfo2(golden_dir):
    model, cfg = S.load_checkpoint(os.path.join(golden_dir, 'cp0_model.pt'))
    model.eval()
    cp = torch.load(os.path.join(golden_dir, 'cp0_model.pt'))
    from sevennet_exp_b200.checkpoint import patch_state_dict
    spec = sp.build_spec(cfg)
    sd = patch_state_dict(spec, cp['model_state_dict'], cfg.get('version', '0.0.0'))
    fr = np.load(os.path.join(golden_dir, 'hfo2_frames.npz'))
    for k in range(2):
        num, pos, cell, pbc = (fr[f'f{k}_{n}'] for n in ('numbers', 'pos', 'cell', 'pbc'))
        print(_compare(cfg, sd, num, pos, cell, pbc, model))


@pytest.mark.parametrize('name', ['test-so3-l2', 'test-so3-l3'])
def test_engine_emu_so3_random_weights(name):
    cfg = dict(presets.PRESETS[name])
    torch.manual_seed(777)
    model = S.build_model(cfg).eval()
    with torch.no_grad():
        model.edge_embedding.basis_function.coeffs.mul_(1.03)
    numbers, pos, cell = random_cell(24, list(cfg['_type_map']), density=0.05, seed=3)
    print(_compare(cfg, model.state_dict(), numbers, pos, cell, [True] * 3, model))


@pytest.mark.parametrize('packed', [False, True])
def test_engine_emu_two_chunk_units(monkeypatch, packed):
    """64 scalar channels: the tabulated kernels give one warp both 32-channel chunks of the l = 0 group
    (conv_codegen.TAB_CPW; this is the SevenNet-0 / l3i5 situation, the small presets have one chunk per group).
    packed = SEVENN_TAB_F2: the chunk pair's per-channel arithmetic as packed pairs (__ffma2_rn / __fmul2_rn;
    component-wise fmaf in the emulation build).  Same parity gates as everything else."""
    from sevennet_exp_b200.codegen import conv_codegen
    monkeypatch.setattr(conv_codegen, 'TAB_F2', packed)
    monkeypatch.setattr(conv_codegen, 'TAB_POLY', not packed)      # the packed variant reads Hermite rows, the default monomial ones
    cfg = dict(presets.PRESETS['test-so3-l2'])
    cfg['irreps_manual'] = ['64x0e', '64x0e+32x1e+32x2e', '64x0e+32x1e+32x2e', '32x0e']
    cfg['channel'] = 64
    conv = sp.build_spec(cfg).layers[1].conv
    assert [G.cpw for G in conv_codegen.make_groups(conv, True)] == [2, 1, 1]
    assert ('= make_float2(g' in conv_codegen.generate(conv)) == packed      # channel pairs of the output gradient
    torch.manual_seed(777)
    model = S.build_model(cfg).eval()
    numbers, pos, cell = random_cell(20, list(cfg['_type_map']), density=0.05, seed=4)
    print(_compare(cfg, model.state_dict(), numbers, pos, cell, [True] * 3, model))


def test_engine_emu_carried_radial_weights(monkeypatch):
    """conv_codegen.TAB_WCARRY: the tabulated kernels evaluate w (and h dw/dr) of the next edge at the end of a loop trip
    instead of carrying the table rows across it.  The emulation build shares the generated algebra and the weight
    evaluation with the device code; the pipelined loads themselves only exist on the device (GPU parity tests)."""
    from sevennet_exp_b200.codegen import conv_codegen
    monkeypatch.setattr(conv_codegen, 'TAB_WCARRY_DEFAULT', (True, True))
    monkeypatch.setattr(conv_codegen, 'TAB_WCARRY_BY_SIG', {})
    cfg = dict(presets.PRESETS['test-so3-l2'])
    conv = sp.build_spec(cfg).layers[1].conv
    src = conv_codegen.generate(conv)
    assert 'tn_' in src and 'T0_' in src
    torch.manual_seed(777)
    model = S.build_model(cfg).eval()
    numbers, pos, cell = random_cell(20, list(cfg['_type_map']), density=0.05, seed=4)
    print(_compare(cfg, model.state_dict(), numbers, pos, cell, [True] * 3, model))


@pytest.mark.parametrize('bias', [False, True])
def test_engine_emu_modality_and_bias(bias):
    """SevenNet-MF style model (model_build.py:198-244, nn/linear.py:65-92, nn/scale.py:165-330): modality one-hot
    appended to the inputs of the embedding / self-interaction / readout linears, modal-wise shift; optionally
    `use_bias_in_linear`.  The engine folds the fixed modality and the biases into constant rows of the scalar blocks;
    every modality must reproduce the oracle, and modalities must differ from each other."""
    cfg = dict(presets.PRESETS['test-so3-l2'])
    nm = 3
    cfg.update(use_modality=True, _number_of_modalities=nm, _modal_map={'pbe': 0, 'scan': 1, 'r2scan': 2},
               use_modal_node_embedding=True, use_modal_self_inter_intro=True, use_modal_self_inter_outro=True,
               use_modal_output_block=True, use_modal_wise_shift=True, use_modal_wise_scale=False,
               use_bias_in_linear=bias,
               shift=[[0.1, -0.2, 0.3], [0.5, 0.4, -0.1], [-0.3, 0.2, 0.0]], scale=[1.0, 1.5, 0.7])
    torch.manual_seed(11)
    model = S.build_model(cfg).eval()
    if bias:
        with torch.no_grad():
            for n_, p_ in model.named_parameters():
                if n_.endswith('linear.bias'):
                    p_.normal_(0.0, 0.3)
    sd = model.state_dict()
    spec = sp.build_spec(cfg)
    assert spec.n_modal == nm and spec.embed.n_modal == nm and (spec.read2.n_bias == 1) == bias
    assert {k: tuple(v.shape) for k, v in sd.items() if k in spec.expected_keys()} == spec.expected_keys()
    numbers, pos, cell = random_cell(20, list(cfg['_type_map']), density=0.05, seed=6)
    energies = []
    for m in range(nm):
        _compare(cfg, sd, numbers, pos, cell, [True] * 3, model, modal=m)
        g = S.atoms_to_graph(numbers, pos, cell, [True] * 3, cfg['cutoff'])
        g[S.MODAL_TYPE] = torch.tensor(m)
        energies.append(float(model(dict(g))[S.E_TOT].detach()))
    assert abs(energies[0] - energies[1]) > 1e-3 and abs(energies[1] - energies[2]) > 1e-3
    lib = build_full_emu([L.conv for L in spec.layers])
    with pytest.raises(ValueError):
        Engine(spec, sd, lib=lib, device='cpu')                      # modal missing
    with pytest.raises(ValueError):
        Engine(spec, sd, lib=lib, device='cpu', modal='hse')          # unknown modal
    assert Engine(spec, sd, lib=lib, device='cpu', modal='scan').modal == 1
    plain = sp.build_spec(dict(presets.PRESETS['test-so3-l2']))
    with pytest.raises(ValueError):
        Engine(plain, S.build_model(dict(presets.PRESETS['test-so3-l2'])).state_dict(), lib=lib, device='cpu', modal='pbe')


@pytest.mark.parametrize('act', ['relu', 'silu'])
def test_engine_emu_readout_as_fcn(act):
    """readout_as_fcn (model_build.py:141-155, nn/linear.py:145-180): FullyConnectedNet readout on the scalar features,
    forward and analytic reverse through the MLP launch path."""
    cfg = dict(presets.PRESETS['test-so3-l2'])
    cfg.update(readout_as_fcn=True, readout_fcn_hidden_neurons=[16, 8], readout_fcn_activation=act)
    torch.manual_seed(21)
    model = S.build_model(cfg).eval()
    spec = sp.build_spec(cfg)
    assert spec.readout_fcn == [32, 16, 8, 1]
    sd = model.state_dict()
    assert {k: tuple(v.shape) for k, v in sd.items() if k in spec.expected_keys()} == spec.expected_keys()
    numbers, pos, cell = random_cell(20, list(cfg['_type_map']), density=0.05, seed=12)
    print(_compare(cfg, sd, numbers, pos, cell, [True] * 3, model))


def test_engine_emu_reference_deployed_torchscript(golden_dir):
    """The reference's frozen TorchScript artefact (deployed_serial.pt): its weights, its frame -> our engine
    (emulation build) must give ITS outputs (tests/golden/deployed_serial_hfo2.npz), no oracle in between."""
    from sevennet_exp_b200.checkpoint import patch_old_config, patch_state_dict
    from tests.test_oracle import deployed_serial_case
    cfg, sd, numbers, pos, cell, gold = deployed_serial_case(golden_dir)
    cfg = patch_old_config(dict(cfg))
    spec = sp.build_spec(cfg)
    sd = patch_state_dict(spec, sd, cfg['version'])
    lib = build_full_emu([L.conv for L in spec.layers])
    ei = gold['edge_index']
    ev = pos[ei[1]] - pos[ei[0]] + gold['pbc_shift'] @ cell
    types = np.array([cfg['_type_map'][int(z)] for z in numbers])
    vol = abs(np.linalg.det(cell))
    gr = graph_from_edges(types, ei, ev.astype(np.float32), len(numbers), vol)
    for radial in ('mlp', 'table'):
        out = Engine(spec, sd, lib=lib, device='cpu', radial=radial).evaluate(gr)
        assert abs(out['energy'].item() - float(gold['energy'])) / len(numbers) < 1e-5, radial
        assert np.abs(out['forces'].numpy() - gold['force']).max() < 1e-4, radial
        stress = -(out['virial6'].numpy() / vol)
        assert np.abs(stress - gold['stress']).max() * KBAR < 1e-2, radial


@pytest.mark.parametrize('cutoff_fn', [{'cutoff_function_name': 'XPLOR', 'cutoff_on': 4.5},
                                       {'cutoff_function_name': 'poly_cut', 'poly_cut_p_value': 6}])
def test_engine_emu_multi_cutoff(cutoff_fn):
    """The fork's MultiCutoffEdgeEmbedding through the whole-model engine: per-layer radial tables on one grid
    (zero beyond the layer's cutoff) against the oracle's per-cutoff sub-graphs."""
    cfg = dict(presets.PRESETS['test-so3-l2'])
    cfg['cutoff_function'] = cutoff_fn
    cfg['multi_cutoff'] = [5.0, 3.5, 5.0]
    torch.manual_seed(31)
    model = S.build_model(cfg).eval()
    numbers, pos, cell = random_cell(24, list(cfg['_type_map']), density=0.05, seed=9)
    spec = sp.build_spec(cfg)
    assert spec.coeff_key(1) == 'edge_embedding.basis_functions.0.coeffs'
    lib = build_full_emu([L.conv for L in spec.layers])
    g = S.atoms_to_graph(numbers, pos, cell, [True] * 3, cfg['cutoff'])
    ref = model(dict(g))
    types = np.array([cfg['_type_map'][int(z)] for z in numbers])
    vol = float(g[S.VOL])
    gr = graph_from_edges(types, g[S.EDGE_IDX].numpy(), g[S.EDGE_VEC].detach().numpy(), len(numbers), vol)
    with pytest.raises(NotImplementedError):
        Engine(spec, model.state_dict(), lib=lib, device='cpu', radial='mlp')
    out = Engine(spec, model.state_dict(), lib=lib, device='cpu', radial='table').evaluate(gr)
    assert abs(out['energy'].item() - ref[S.E_TOT].item()) / len(numbers) < 1e-5
    assert (out['forces'] - ref[S.FORCE].detach()).abs().max().item() < 1e-4
    assert ((-(out['virial6'] / vol)).float() - ref[S.STRESS].detach()).abs().max().item() * KBAR < 1e-2

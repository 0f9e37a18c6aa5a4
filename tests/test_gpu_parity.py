"""GPU parity tests proper: the CUDA library (through the C ABI) against the CPU oracle.
Tolerances are north_star's: edge set bit-exact; E 1e-5 eV/atom; F 1e-4 eV/A; stress 1e-3 kbar (fp32)."""
import os

import numpy as np
import pytest
import torch

from oracle import sevenn_ref as S
from sevennet_exp_b200 import presets
from tests.structures import diamond_si, fluorite_hfo2, random_cell

pytestmark = pytest.mark.gpu
KBAR = 1602.1766208


@pytest.fixture(scope='module')
def lib():
    from sevennet_exp_b200 import _lib
    return _lib.load()


def _nl_check(pos, cell, pbc, rc):
    from sevennet_exp_b200.neighbor import NeighborList
    r = NeighborList().build(pos, cell, pbc, rc)
    i, j, D, Sh = S.neighbor_list_bruteforce(pos, cell, pbc, rc)
    assert r['n_edges'] == len(i)
    assert np.array_equal(r['center'].cpu().numpy(), i)
    assert np.array_equal(r['col'].cpu().numpy(), j)
    assert np.array_equal(r['shift'].cpu().numpy(), Sh.astype(np.int64))
    assert np.array_equal(r['edge_vec'].cpu().numpy(), D.astype(np.float32))
    cp, ce = r['cptr'].cpu().numpy(), r['cedge'].cpu().numpy()
    assert np.array_equal(np.sort(ce), np.arange(len(i)))
    assert np.array_equal(j[ce], np.sort(j, kind='stable'))
    assert np.array_equal(np.diff(cp), np.bincount(j, minlength=len(pos)))
    return r


def test_nl_bit_exact_edge_sets(lib):
    num, pos, cell = random_cell(200, [1, 2], density=0.05, seed=1)
    _nl_check(pos, cell, [True] * 3, 4.0)
    num, pos, cell = random_cell(30, [1], density=0.04, seed=2)
    pos = pos + np.array([2.3, -1.7, 0.4]) @ cell
    _nl_check(pos, cell, [True] * 3, 4.5)
    c = np.array([[0, 2.815, 2.815], [2.815, 0, 2.815], [2.815, 2.815, 0]])
    assert _nl_check(np.array([[0, 0, 0], [2.815] * 3]), c, [True] * 3, 4.0)['n_edges'] == 36
    _nl_check(np.zeros((1, 3)), np.eye(3) * 1.7, [True] * 3, 5.0)
    num, pos, cell = diamond_si((3, 3, 3))
    _nl_check(pos, cell, [True] * 3, 5.0)
    num, pos, cell = random_cell(40, [1], density=0.05, seed=5, triclinic=False)
    for pbc in ([False] * 3, [True, True, False]):
        _nl_check(pos, cell, pbc, 4.0)
    assert _nl_check(np.zeros((1, 3)), np.zeros((3, 3)), [False] * 3, 4.0)['n_edges'] == 0


def test_nl_large_property(lib):
    """4,116-atom HfO2 (cfg2 size): symmetric edge set, every |D| < rc, count equals the O(N^2)-free check."""
    from sevennet_exp_b200.neighbor import NeighborList
    num, pos, cell = fluorite_hfo2((7, 7, 7), a=5.0)
    r = NeighborList().build(pos, cell, [True] * 3, 5.0)
    i, j = r['center'].cpu().numpy().astype(np.int64), r['col'].cpu().numpy().astype(np.int64)
    sh, D = r['shift'].cpu().numpy(), r['edge_vec'].cpu().numpy().astype(np.float64)
    assert (np.linalg.norm(D, axis=1) < 5.0).all()
    assert np.all(np.diff(i) >= 0)
    fwd = set(zip(i.tolist(), j.tolist(), map(tuple, sh.tolist())))
    assert len(fwd) == len(i)
    assert all((b, a, (-s[0], -s[1], -s[2])) in fwd for a, b, s in list(fwd)[::37])
    ref = pos[j] + sh @ cell - pos[i]
    assert np.abs(ref - D).max() < 1e-5
    # count against a KD-free brute force on a slab of atoms
    sub = np.arange(0, len(pos), 97)
    for a in sub:
        d = pos - pos[a]
        f = d @ np.linalg.inv(cell)
        f -= np.round(f)
        cnt = (np.linalg.norm(f @ cell, axis=1) < 5.0).sum() - 1
        assert cnt == (i == a).sum()


def _engine_vs_oracle(cfg, sd, model, numbers, pos, cell, pbc):
    from sevennet_exp_b200.sevennet_calculator import SevenNetCalculator
    calc = SevenNetCalculator((cfg, sd), file_type='model_instance')
    res = calc.compute(numbers, pos, cell, pbc)
    # oracle evaluated in fp64 (it is fp64-capable): the difference below is the fp32 kernels' own error
    import copy
    torch.set_default_dtype(torch.float64)
    try:
        g = S.atoms_to_graph(numbers, pos, cell, pbc, cfg['cutoff'], dtype=torch.float64)
        ref = copy.deepcopy(model).double()(dict(g))
    finally:
        torch.set_default_dtype(torch.float32)
    n = len(numbers)
    assert res['n_edges'] == g[S.EDGE_IDX].shape[1]
    dE = abs(res['energy'] - ref[S.E_TOT].item()) / n
    dF = np.abs(res['forces'] - ref[S.FORCE].detach().numpy()).max()
    dS = np.abs(res['inferred_stress'] - ref[S.STRESS].detach().numpy()).max() * KBAR
    dAE = np.abs(res['energies'] - ref[S.AE].detach().numpy().ravel()).max()
    assert dE < 1e-5, f'energy/atom {dE}'
    assert dAE < 1e-4, f'atomic energies {dAE}'
    assert dF < 1e-4, f'forces {dF}'
    assert dS < 1e-3 * max(1.0, np.abs(ref[S.STRESS].detach().numpy()).max() * KBAR * 1e-3), f'stress {dS} kbar'
    # ASE Voigt convention of the calculator (sevennet_calculator.py:196-201)
    assert np.allclose(res['stress'], -res['inferred_stress'][[0, 1, 2, 4, 5, 3]])
    return dE, dF, dS


def test_cp0_checkpoint_hfo2_known_answer(lib, golden_dir):
    """reference test_train.py:284-316: cp_0.pth on hfo2.extxyz -> Energy/Force/Stress RMSE goldens, on the GPU."""
    from sevennet_exp_b200.sevennet_calculator import SevenNetCalculator
    cp = torch.load(os.path.join(golden_dir, 'cp0_model.pt'))
    calc = SevenNetCalculator((cp['config'], cp['model_state_dict']), file_type='model_instance')
    fr = np.load(os.path.join(golden_dir, 'hfo2_frames.npz'))
    dE, dF, dS = [], [], []
    for k in range(2):
        num, pos, cell, pbc = (fr[f'f{k}_{n}'] for n in ('numbers', 'pos', 'cell', 'pbc'))
        res = calc.compute(num, pos, cell, pbc)
        s = fr[f'f{k}_stress3x3']
        voigt = np.array([s[0, 0], s[1, 1], s[2, 2], s[1, 2], s[0, 2], s[0, 1]])
        dE.append((res['energy'] - float(fr[f'f{k}_energy'])) / len(num))
        dF.append(res['forces'] - fr[f'f{k}_forces'])
        dS.append(res['inferred_stress'] - (-voigt[[0, 1, 2, 5, 3, 4]]))
    dF = np.concatenate(dF)
    assert abs(np.sqrt(np.mean(np.square(dE))) - 28.977758) < 2e-5
    assert abs(np.sqrt(np.mean((dF ** 2).sum(1))) - 0.214107) < 1e-5
    assert abs(np.sqrt(np.mean((np.array(dS) ** 2).sum(1))) * KBAR - 190.014237) < 1e-3


def test_reference_deployed_torchscript_outputs_on_gpu(lib, golden_dir):
    """example_inputs/md_serial_example/deployed_serial.pt (the reference's frozen e3nn code path, sevenn 0.8.6):
    its weights (tests/golden/make_deployed_transplant.py) + its input frame through SevenNetCalculator on the GPU
    must reproduce ITS energy / atomic energies / forces / stress (tests/golden/deployed_serial_hfo2.npz) --
    the CUDA path against the real reference, no oracle in between.  Both radial modes."""
    from sevennet_exp_b200.sevennet_calculator import SevenNetCalculator
    from tests.test_oracle import deployed_serial_case
    cfg, sd, numbers, pos, cell, gold = deployed_serial_case(golden_dir)
    for radial in ('table', 'mlp'):
        os.environ['SEVENN_RADIAL'] = radial
        try:
            calc = SevenNetCalculator((dict(cfg), sd), file_type='model_instance')
        finally:
            os.environ.pop('SEVENN_RADIAL')
        assert calc.engine.radial == radial
        res = calc.compute(numbers, pos, cell, [True] * 3)
        assert res['n_edges'] == gold['edge_index'].shape[1]
        assert abs(res['energy'] - float(gold['energy'])) / len(numbers) < 1e-5, radial
        assert np.abs(res['energies'] - gold['atomic_energy'].reshape(-1)).max() < 1e-5, radial
        assert np.abs(res['forces'] - gold['force']).max() < 1e-4, radial
        assert np.abs(res['inferred_stress'] - gold['stress']).max() * KBAR < 1e-2, radial   # reference ran in fp32


def test_cp0_vs_oracle(lib, golden_dir):
    model, cfg = S.load_checkpoint(os.path.join(golden_dir, 'cp0_model.pt'))
    cp = torch.load(os.path.join(golden_dir, 'cp0_model.pt'))
    fr = np.load(os.path.join(golden_dir, 'hfo2_frames.npz'))
    num, pos, cell, pbc = (fr[f'f0_{n}'] for n in ('numbers', 'pos', 'cell', 'pbc'))
    print(_engine_vs_oracle(cp['config'], cp['model_state_dict'], model.eval(), num, pos, cell, pbc))


def _random_model(name, seed=777):
    cfg = dict(presets.PRESETS[name])
    cfg['version'] = '0.11.0'
    cfg['scale'] = [10.0] * cfg['_number_of_species']      # forces of O(0.1 eV/A) with random weights
    torch.manual_seed(seed)
    model = S.build_model(cfg).eval()
    with torch.no_grad():
        model.edge_embedding.basis_function.coeffs.mul_(1.03)
    return cfg, model


@pytest.mark.parametrize('name', ['test-so3-l2', 'test-so3-l3'])
def test_small_so3_models(lib, name):
    cfg, model = _random_model(name)
    numbers, pos, cell = random_cell(48, list(cfg['_type_map']), density=0.05, seed=3)
    print(_engine_vs_oracle(cfg, model.state_dict(), model, numbers, pos, cell, [True] * 3))


def test_sevennet0_shape_si64(lib):
    """BASELINE config 1: SevenNet-0 hyper-parameters (random weights), 64-atom bulk Si."""
    cfg, model = _random_model('7net-0')
    numbers, pos, cell = diamond_si((2, 2, 2))
    print(_engine_vs_oracle(cfg, model.state_dict(), model, numbers, pos, cell, [True] * 3))


def test_l3i5_shape_small(lib):
    """SevenNet-l3i5 hyper-parameters (lmax 3, random weights) on a 40-atom Li-P-S-like cell."""
    cfg, model = _random_model('7net-l3i5')
    numbers, pos, cell = random_cell(40, [3, 15, 16], density=0.048, seed=1, rmin=1.9)
    print(_engine_vs_oracle(cfg, model.state_dict(), model, numbers, pos, cell, [True] * 3))


@pytest.mark.parametrize('cutoff_fn', [{'cutoff_function_name': 'XPLOR', 'cutoff_on': 4.5},
                                       {'cutoff_function_name': 'poly_cut', 'poly_cut_p_value': 6}])
def test_multi_cutoff_model_on_gpu(lib, cutoff_fn):
    """MultiCutoffEdgeEmbedding (sevenn/nn/edge_embedding.py:220-257): per-layer cutoffs on one neighbour list, through
    the calculator's default (tabulated) radial mode."""
    cfg = dict(presets.PRESETS['test-so3-l2'])
    cfg.update(version='0.11.0', scale=[10.0] * cfg['_number_of_species'], cutoff_function=cutoff_fn,
               multi_cutoff=[5.0, 3.5, 5.0])
    torch.manual_seed(31)
    model = S.build_model(cfg).eval()
    numbers, pos, cell = random_cell(40, list(cfg['_type_map']), density=0.05, seed=9)
    print(_engine_vs_oracle(cfg, model.state_dict(), model, numbers, pos, cell, [True] * 3))


@pytest.mark.parametrize('pbc', [[False] * 3, [True, True, False]])
def test_open_and_slab_boundaries_on_gpu(lib, pbc):
    """Non-periodic directions (the reference inflates the cell, train/dataload.py:40-48): a cluster / slab cut out of
    a random cell; energy, atomic energies and forces against the fp64 oracle on the same edge set."""
    from sevennet_exp_b200.sevennet_calculator import SevenNetCalculator
    cfg, model = _random_model('test-so3-l2')
    numbers, pos, cell = random_cell(36, list(cfg['_type_map']), density=0.05, seed=5, triclinic=False)
    calc = SevenNetCalculator((cfg, model.state_dict()), file_type='model_instance')
    res = calc.compute(numbers, pos, cell, pbc)
    import copy
    torch.set_default_dtype(torch.float64)
    try:
        g = S.atoms_to_graph(numbers, pos, cell, pbc, cfg['cutoff'], dtype=torch.float64)
        ref = copy.deepcopy(model).double()(dict(g))
    finally:
        torch.set_default_dtype(torch.float32)
    assert res['n_edges'] == g[S.EDGE_IDX].shape[1]
    assert abs(res['energy'] - ref[S.E_TOT].item()) / len(numbers) < 1e-5
    assert np.abs(res['energies'] - ref[S.AE].detach().numpy().ravel()).max() < 1e-4
    assert np.abs(res['forces'] - ref[S.FORCE].detach().numpy()).max() < 1e-4
    assert np.abs(res['forces'].sum(0)).max() < 1e-4          # no net force on an isolated system


def test_bitwise_reproducible_and_edge_cases(lib):
    from sevennet_exp_b200.sevennet_calculator import SevenNetCalculator
    cfg, model = _random_model('test-so3-l2')
    calc = SevenNetCalculator((cfg, model.state_dict()), file_type='model_instance')
    numbers, pos, cell = random_cell(64, list(cfg['_type_map']), density=0.05, seed=4)
    a = calc.compute(numbers, pos, cell, [True] * 3)
    b = calc.compute(numbers, pos, cell, [True] * 3)
    assert a['energy'] == b['energy'] and np.array_equal(a['forces'], b['forces'])   # no atomics on the path
    assert np.abs(a['forces'].sum(0)).max() < 1e-3
    # isolated atom: no edges -> energy is the species shift, zero force
    one = calc.compute(numbers[:1], np.zeros((1, 3)), np.eye(3) * 30.0, [True] * 3)
    assert one['n_edges'] == 0 and np.all(one['forces'] == 0)
    ref = model(S.atoms_to_graph(numbers[:1], np.zeros((1, 3)), np.eye(3) * 30.0, [True] * 3, 5.0))
    assert abs(one['energy'] - ref[S.AE].sum().item()) < 1e-5
    with pytest.raises(ValueError):
        calc.compute(np.array([99]), np.zeros((1, 3)), np.eye(3) * 30.0, [True] * 3)


def test_gemm_against_torch(lib):
    import ctypes
    from sevennet_exp_b200 import _lib as L
    torch.manual_seed(0)
    for (n, R, K, N) in [(1000, 1, 8, 64), (777, 3, 64, 96), (513, 5, 352, 32), (4096, 1, 64, 960)]:
        ldn = R * K + 7
        A = torch.randn(n, ldn, device='cuda')
        B = torch.randn(K, N, device='cuda')
        C = torch.zeros(n, R * N, device='cuda')
        g = L.GemmArgs(A.data_ptr(), ldn, K, B.data_ptr(), N, C.data_ptr(), R * N, N, n * R, R, N, K, 0, 0, 1.0, None, None)
        lib.call('sevenn_gemm_f32', ctypes.byref(g), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
        ref = torch.einsum('nrk,kc->nrc', A[:, :R * K].reshape(n, R, K).double(), B.double()).reshape(n, R * N)
        assert (C.double() - ref).abs().max() < 1e-4 * max(1.0, ref.abs().max().item())


def test_gemm_tc_3xtf32_against_fp64(lib):
    """tcgen05 arm: all four epilogues, (node, m)-strided rows, ragged M/N/K; fp32-level accuracy."""
    import ctypes
    from sevennet_exp_b200 import _lib as L
    from sevennet_exp_b200.engine import tc_image_for
    torch.manual_seed(1)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    cst = 1.6791767923989418
    for (n, R, K, N, epi) in [(1000, 1, 8, 64, 2), (4133, 1, 64, 960, 0), (3000, 1, 960, 64, 3), (777, 3, 64, 96, 0),
                              (513, 5, 352, 32, 1), (2049, 1, 64, 64, 2), (300, 3, 384, 64, 1), (5000, 1, 64, 8, 1),
                              (1500, 1, 64, 1760, 0)]:
        ldn = R * K + 8
        A = torch.randn(n, ldn, device='cuda')
        B = torch.randn(K, N, device='cuda') / K ** 0.5
        C = torch.randn(n, R * N, device='cuda')
        C0 = C.clone()
        P = torch.randn(n, R * N, device='cuda')
        C2 = torch.zeros(n, R * N, device='cuda')
        img = torch.from_numpy(tc_image_for(lib, B.cpu().numpy())).cuda()
        g = L.GemmArgs(A.data_ptr(), ldn, K, B.data_ptr(), N, C.data_ptr(), R * N, N, n * R, R, N, K, epi, 0, cst,
                       C2.data_ptr(), P.data_ptr())
        lib.call('sevenn_gemm_tc', ctypes.byref(g), ctypes.c_void_p(img.data_ptr()), st)
        torch.cuda.synchronize()
        acc = torch.einsum('nrk,kc->nrc', A[:, :R * K].reshape(n, R, K).double(), B.double()).reshape(n, R * N)
        if epi == 0:
            ref = acc
        elif epi == 1:
            ref = C0.double() + acc
        elif epi == 2:
            ref = acc
            assert (C2.double() - cst * torch.nn.functional.silu(acc)).abs().max() < 2e-5
        else:
            p = P.double()
            s = torch.sigmoid(p)
            ref = acc * cst * s * (1 + p * (1 - s))
        err = (C.double() - ref).abs().max().item()
        # tensor-core fp32 accumulation truncates (round-toward-zero) at every K=8 step: the bound grows with K
        assert err < 2e-5 * max(1.0, K / 128), (n, R, K, N, epi, err)


def test_engine_tc_and_simt_arms_agree(lib):
    """Same SevenNet-0-shaped evaluation through the tcgen05 arm and the exact-fp32 SIMT arm."""
    from sevennet_exp_b200.sevennet_calculator import SevenNetCalculator
    cfg, model = _random_model('7net-0')
    numbers, pos, cell = fluorite_hfo2((3, 3, 3), a=4.8)
    out = {}
    for mode in ('simt', 'tc'):
        calc = SevenNetCalculator((cfg, model.state_dict()), file_type='model_instance')
        calc.engine.gemm_mode, calc.engine.tc_min_rows = mode, 1
        out[mode] = calc.compute(numbers, pos, cell, [True] * 3)
    n = len(numbers)
    assert abs(out['tc']['energy'] - out['simt']['energy']) / n < 5e-6
    assert np.abs(out['tc']['forces'] - out['simt']['forces']).max() < 5e-5
    assert np.abs(out['tc']['inferred_stress'] - out['simt']['inferred_stress']).max() * KBAR < 1e-3


@pytest.mark.parametrize('name,n_atoms', [('test-so3-l2', 48), ('7net-0', 64)])
def test_nn_module_seam_on_gpu(lib, name, n_atoms):
    """sevennet_exp_b200.nn.AtomGraphSequential (reference module names, reference state_dict, shuffled
    edge order, forces through torch.autograd.grad over our reverse kernels) vs the fp64 oracle."""
    import copy
    from sevennet_exp_b200 import nn as bnn
    cfg, om = _random_model(name)
    if name == '7net-0':
        numbers, pos, cell = diamond_si((2, 2, 2))
    else:
        numbers, pos, cell = random_cell(n_atoms, list(cfg['_type_map']), density=0.05, seed=3)
    model = bnn.build_E3_equivariant_model(cfg).eval()
    missing, _ = model.load_state_dict(om.state_dict(), strict=False)
    assert not missing
    model = model.cuda()
    torch.set_default_dtype(torch.float64)
    try:
        g64 = S.atoms_to_graph(numbers, pos, cell, [True] * 3, cfg['cutoff'], dtype=torch.float64)
        ref = copy.deepcopy(om).double()(dict(g64))
    finally:
        torch.set_default_dtype(torch.float32)
    g = S.atoms_to_graph(numbers, pos, cell, [True] * 3, cfg['cutoff'])
    p = torch.from_numpy(np.random.default_rng(0).permutation(g[S.EDGE_IDX].shape[1]))
    g[S.EDGE_IDX], g[S.EDGE_VEC] = g[S.EDGE_IDX][:, p].contiguous(), g[S.EDGE_VEC].detach()[p].contiguous()
    data = {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in g.items()}
    out = model(data)
    n = len(numbers)
    dE = abs(out[S.E_TOT].item() - ref[S.E_TOT].item()) / n
    dF = (out[S.FORCE].cpu().double() - ref[S.FORCE].detach()).abs().max().item()
    dS = (out[S.STRESS].cpu().double() - ref[S.STRESS].detach()).abs().max().item() * KBAR
    assert dE < 1e-5 and dF < 1e-4, (dE, dF)
    assert dS < 1e-3 * max(1.0, ref[S.STRESS].abs().max().item() * KBAR * 1e-3), dS


def test_batched_structures_match_single(lib):
    """compute_batch (block-diagonal graph, per-graph energy / stress reduction; reference: PyG batching +
    force_output.py:210-219) == one compute() per structure, incl. a non-periodic and an empty-neighbour structure."""
    from sevennet_exp_b200.sevennet_calculator import SevenNetCalculator
    cfg, model = _random_model('test-so3-l2')
    calc = SevenNetCalculator((cfg, model.state_dict()), file_type='model_instance')
    zs = list(cfg['_type_map'])
    structs = []
    for k, n in enumerate((17, 64, 5, 40)):
        numbers, pos, cell = random_cell(n, zs, density=0.05, seed=10 + k)
        structs.append((numbers, pos, cell, [True] * 3 if k != 2 else [False] * 3))
    structs.append((np.array([zs[0]]), np.zeros((1, 3)), np.eye(3) * 20.0, [True] * 3))      # no neighbours at all
    batch = calc.compute_batch(structs)
    assert len(batch) == len(structs)
    for s, b in zip(structs, batch):
        one = calc.compute(*s)
        assert b['n_edges'] == one['n_edges']
        # the merged graph is large enough for the tcgen05 (3xTF32) GEMM arm, the single small structures run the
        # exact-fp32 SIMT arm: agreement is at the level test_engine_tc_and_simt_arms_agree allows
        assert abs(b['energy'] - one['energy']) < 5e-6 * max(1, len(s[0]))
        assert np.abs(b['forces'] - one['forces']).max() < 5e-5
        assert np.abs(b['stress'] - one['stress']).max() * KBAR < 1e-3
        assert np.abs(b['energies'] - one['energies']).max() < 5e-5


def test_deployed_entry_points_on_gpu(lib, golden_dir, tmp_path):
    """deploy.py on the real library: serial bundle through the deployed-model dict contract and through
    SevenNetCalculator(file_type='torchscript'); parallel segments + comm loop; all against the reference
    artefact's outputs / the serial evaluation."""
    from sevennet_exp_b200 import deploy as D
    from sevennet_exp_b200.sevennet_calculator import SevenNetCalculator
    from tests.test_deploy_emu import _periodic_ghost_graph
    from tests.test_oracle import deployed_serial_case
    cfg, sd, numbers, pos, cell, gold = deployed_serial_case(golden_dir)
    path = D.deploy((cfg, sd), str(tmp_path / 'serial.pt'))
    model, extra = D.load_deployed(path)
    assert isinstance(model, torch.jit.ScriptModule)           # a real TorchScript archive (torch.jit.load)
    assert extra['chemical_symbols_to_index'] == 'Hf O ' and float(extra['cutoff']) == 4.0
    n = len(numbers)
    data = {'x': torch.tensor([cfg['_type_map'][int(z)] for z in numbers]),
            'pos': torch.tensor(pos, dtype=torch.float32), 'edge_index': torch.tensor(gold['edge_index']),
            'num_atoms': torch.tensor([n]), 'cell_lattice_vectors': torch.tensor(cell, dtype=torch.float32),
            'cell_volume': torch.tensor(abs(np.linalg.det(cell)), dtype=torch.float32),
            'pbc_shift': torch.tensor(gold['pbc_shift'], dtype=torch.float32)}
    out = model(data)
    assert out['inferred_force'].is_cuda
    assert abs(out['inferred_total_energy'].item() - float(gold['energy'])) / n < 1e-5
    assert np.abs(out['inferred_force'].cpu().numpy() - gold['force']).max() < 1e-4
    assert np.abs(out['inferred_stress'].cpu().numpy() - gold['stress']).max() * KBAR < 1e-2
    calc = SevenNetCalculator(path, file_type='torchscript')
    res = calc.compute(numbers, pos, cell, [True] * 3)
    assert abs(res['energy'] - float(gold['energy'])) / n < 1e-5
    assert np.abs(res['forces'] - gold['force']).max() < 1e-4
    with pytest.raises(ValueError):
        SevenNetCalculator(os.path.join(golden_dir, 'cp0_model.pt'), file_type='torchscript')

    par, extra = D.load_deployed(D.deploy_parallel((cfg, sd), str(tmp_path / 'par')))
    ei, ev, owner = _periodic_ghost_graph(numbers, pos, cell, float(extra['cutoff']))
    own_t = torch.tensor(owner, device='cuda')
    types = data['x'].cuda()
    E, dE_dr, ae = par.compute(types, types[own_t], torch.tensor(ei), torch.tensor(ev),
                               lambda x: x[own_t], lambda dx, dg: dx.clone().index_add_(0, own_t, dg))
    assert abs(E.item() - float(gold['energy'])) / n < 1e-5
    eit = torch.tensor(ei, device='cuda')
    F = torch.zeros(n + len(owner), 3, device='cuda').index_add_(0, eit[0], dE_dr).index_add_(0, eit[1], -dE_dr)
    F = F[:n].index_add_(0, own_t, F[n:])
    assert np.abs(F.cpu().numpy() - gold['force']).max() < 1e-4


# ---- the path the bench times: tcgen05 3xTF32 GEMM arm + tabulated radial weights at >= 1,000 atoms, against the fp64
# oracle (outputs committed by tests/golden/make_large_oracle_fixtures.py: the oracle needs minutes / tens of GB there)
def _large_case(golden_dir, tag):
    from tests.golden.make_large_oracle_fixtures import state_sha1
    fx = np.load(os.path.join(golden_dir, f'large_oracle_{tag}.npz'))
    cfg, model = _random_model(str(fx['preset']))
    assert state_sha1(model.state_dict()) == str(fx['state_sha1']), 'seeded weights differ from the fixture\'s'
    return cfg, model, fx


@pytest.mark.parametrize('tag', ['7net0_hfo2_324', '7net0_hfo2_1500', 'l3i5_lips_1000'])
def test_bench_path_vs_fp64_oracle_fixture(lib, golden_dir, tag):
    """SevenNet-0 shape on 324- and 1,500-atom fluorite HfO2 at the bench density, SevenNet-l3i5 shape on a 1,000-atom
    Li-P-S cell; weights scaled so forces are O(0.1 eV/A).  Every combination of radial mode (table = bench default,
    mlp) and GEMM arm selection (default threshold, everything forced onto the tcgen05 arm) against the fp64 oracle
    at north_star's tolerances -- not against another arm of this library."""
    from sevennet_exp_b200.sevennet_calculator import SevenNetCalculator
    cfg, model, fx = _large_case(golden_dir, tag)
    numbers, pos, cell = fx['numbers'], fx['pos'], fx['cell']
    n = len(numbers)
    worst = {}
    for radial in ('table', 'mlp'):
        for tc_min_rows in (None, 1):
            os.environ['SEVENN_RADIAL'] = radial
            try:
                calc = SevenNetCalculator((dict(cfg), model.state_dict()), file_type='model_instance')
            finally:
                os.environ.pop('SEVENN_RADIAL')
            if tc_min_rows is not None:
                calc.engine.tc_min_rows = tc_min_rows
            res = calc.compute(numbers, pos, cell, [True] * 3)
            assert calc.n_tc_launch > 0, 'the tcgen05 arm did not run'
            assert res['n_edges'] == int(fx['n_edges'])
            dE = abs(res['energy'] - float(fx['energy'])) / n
            dAE = np.abs(res['energies'] - fx['atomic_energy']).max()
            dF = np.abs(res['forces'] - fx['forces']).max()
            dS = np.abs(res['inferred_stress'] - fx['stress']).max() * KBAR
            worst[(radial, tc_min_rows)] = (dE, dAE, dF, dS, calc.n_tc_launch, 'native' if calc.native else 'python')
            assert dE < 1e-5, (tag, radial, tc_min_rows, dE)
            assert dAE < 1e-4, (tag, radial, tc_min_rows, dAE)
            assert dF < 1e-4, (tag, radial, tc_min_rows, dF)
            assert dS < 1e-3 * max(1.0, np.abs(fx['stress']).max() * KBAR * 1e-3), (tag, radial, tc_min_rows, dS)
    print(tag, 'max|F| %.3f eV/A' % np.abs(fx['forces']).max(), worst)


def test_tc_weight_image_follows_the_weights(lib):
    """Two models of one architecture built back to back (the caching allocator hands the second one the first one's
    addresses) and an in-place weight update: the tcgen05 arm must use the CURRENT weights (its pre-swizzled image is
    tied to the weight tensor and its version, not to data_ptr)."""
    import copy
    from sevennet_exp_b200 import nn as bnn
    from sevennet_exp_b200.nn import _backend
    numbers, pos, cell = random_cell(64, list(presets.PRESETS['test-so3-l2']['_type_map']), density=0.05, seed=8)
    kern = _backend.kernels('cuda')
    outs = {}
    for seed in (1, 2, 3):
        cfg, om = _random_model('test-so3-l2', seed=seed)
        g = S.atoms_to_graph(numbers, pos, cell, [True] * 3, cfg['cutoff'])
        assert g[S.EDGE_IDX].shape[1] >= 1024          # radial-MLP GEMMs are over the tc threshold
        for mode in ('tc', 'simt'):
            kern.gemm_mode, n0 = ('auto' if mode == 'tc' else 'simt'), kern.n_tc_launch
            model = bnn.build_E3_equivariant_model(cfg).eval()
            assert not model.load_state_dict(om.state_dict(), strict=False)[0]
            model = model.cuda()
            out = model({k: (v.cuda() if torch.is_tensor(v) else v) for k, v in copy.deepcopy(g).items()})
            outs[(seed, mode)] = (out[S.E_TOT].item(), out[S.FORCE].detach().cpu().numpy())
            assert (kern.n_tc_launch > n0) == (mode == 'tc')
            if seed == 3 and mode == 'tc':           # in-place update of live parameters, same addresses
                with torch.no_grad():
                    for p_ in model.parameters():
                        p_.mul_(1.05)
                out = model({k: (v.cuda() if torch.is_tensor(v) else v) for k, v in copy.deepcopy(g).items()})
                outs['inplace'] = (out[S.E_TOT].item(), out[S.FORCE].detach().cpu().numpy())
            del model, out
    kern.gemm_mode = 'auto'
    for seed in (1, 2, 3):
        (e_tc, f_tc), (e_s, f_s) = outs[(seed, 'tc')], outs[(seed, 'simt')]
        assert abs(e_tc - e_s) / len(numbers) < 5e-6 and np.abs(f_tc - f_s).max() < 5e-5, seed
    assert abs(outs[(1, 'tc')][0] - outs[(2, 'tc')][0]) > 1e-3       # the models really differ
    assert np.abs(outs['inplace'][1] - outs[(3, 'tc')][1]).max() > 1e-4


# ---- SURVEY 8f-4 variants on hardware: multi-modal checkpoints, biases in the linears, FCN readout
def _modal_cfg(bias):
    cfg = dict(presets.PRESETS['test-so3-l2'])
    zs = list(cfg['_type_map'])
    assert len(zs) == 3
    cfg.update(version='0.11.0', use_modality=True, _number_of_modalities=3, _modal_map={'pbe': 0, 'scan': 1, 'r2scan': 2},
               use_modal_node_embedding=True, use_modal_self_inter_intro=True, use_modal_self_inter_outro=True,
               use_modal_output_block=True, use_modal_wise_shift=True, use_modal_wise_scale=False,
               use_bias_in_linear=bias,
               shift=[[0.1, -0.2, 0.3], [0.5, 0.4, -0.1], [-0.3, 0.2, 0.0]], scale=[10.0, 15.0, 7.0])
    return cfg


@pytest.mark.parametrize('bias', [False, True])
def test_multi_modal_and_bias_on_gpu(lib, bias):
    """SevenNet-MF style checkpoint (modality one-hot into embedding / self-interaction / readout linears, modal-wise
    shift) with and without o3.Linear biases: every modality through SevenNetCalculator(modal=...) vs the fp64 oracle."""
    import copy
    from sevennet_exp_b200.sevennet_calculator import SevenNetCalculator
    cfg = _modal_cfg(bias)
    torch.manual_seed(11)
    model = S.build_model(cfg).eval()
    if bias:
        with torch.no_grad():
            for n_, p_ in model.named_parameters():
                if n_.endswith('linear.bias'):
                    p_.normal_(0.0, 0.3)
    numbers, pos, cell = random_cell(40, list(cfg['_type_map']), density=0.05, seed=6)
    energies = []
    for name, m in cfg['_modal_map'].items():
        calc = SevenNetCalculator((dict(cfg), model.state_dict()), file_type='model_instance', modal=name)
        res = calc.compute(numbers, pos, cell, [True] * 3)
        torch.set_default_dtype(torch.float64)
        try:
            g = S.atoms_to_graph(numbers, pos, cell, [True] * 3, cfg['cutoff'], dtype=torch.float64)
            g[S.MODAL_TYPE] = torch.tensor(m)
            ref = copy.deepcopy(model).double()(dict(g))
        finally:
            torch.set_default_dtype(torch.float32)
        n = len(numbers)
        assert abs(res['energy'] - ref[S.E_TOT].item()) / n < 1e-5, (name, bias)
        assert np.abs(res['forces'] - ref[S.FORCE].detach().numpy()).max() < 1e-4, (name, bias)
        sref = ref[S.STRESS].detach().numpy()
        assert np.abs(res['inferred_stress'] - sref).max() * KBAR < 1e-3 * max(1.0, np.abs(sref).max() * KBAR * 1e-3)
        energies.append(res['energy'])
    assert abs(energies[0] - energies[1]) > 1e-3 and abs(energies[1] - energies[2]) > 1e-3
    with pytest.raises(ValueError):
        SevenNetCalculator((dict(cfg), model.state_dict()), file_type='model_instance')          # modal missing


@pytest.mark.parametrize('act', ['relu', 'silu'])
def test_readout_as_fcn_on_gpu(lib, act):
    """readout_as_fcn (model_build.py:141-155, nn/linear.py:145-180) on the GPU kernels vs the fp64 oracle."""
    cfg = dict(presets.PRESETS['test-so3-l2'])
    cfg.update(version='0.11.0', readout_as_fcn=True, readout_fcn_hidden_neurons=[16, 8], readout_fcn_activation=act,
               scale=[10.0] * cfg['_number_of_species'])
    torch.manual_seed(21)
    model = S.build_model(cfg).eval()
    numbers, pos, cell = random_cell(40, list(cfg['_type_map']), density=0.05, seed=12)
    print(_engine_vs_oracle(cfg, model.state_dict(), model, numbers, pos, cell, [True] * 3))


def test_torchscript_artifact_in_a_fresh_process(lib, golden_dir, tmp_path):
    """What a C++ host does (sevenn/pair_e3gnn/pair_e3gnn.cpp:306-335, 218-264): load the op library, torch.jit.load
    the deployed file with its `_extra_files`, call forward on the input dict -- in a new interpreter that never imports
    this package -- and get the outputs of the REFERENCE's own deployed model on the same frame
    (tests/golden/deployed_serial_hfo2.npz).  The edge list is handed over in shuffled order."""
    import subprocess
    import sys
    from sevennet_exp_b200 import deploy as D
    from tests.test_oracle import deployed_serial_case
    cfg, sd, numbers, pos, cell, gold = deployed_serial_case(golden_dir)
    path = D.deploy((cfg, sd), str(tmp_path / 'deployed_serial.pt'))
    np.savez(tmp_path / 'inp.npz', x=np.array([cfg['_type_map'][int(z)] for z in numbers]), pos=pos, cell=cell,
             edge_index=gold['edge_index'], pbc_shift=gold['pbc_shift'])
    code = f"""
import numpy as np, torch
torch.ops.load_library({D.TORCH_OP_LIB!r})
md = {{'chemical_symbols_to_index': '', 'cutoff': '', 'num_species': '', 'model_type': '', 'version': '', 'dtype': '', 'time': ''}}
m = torch.jit.load({path!r}, _extra_files=md)
assert md['model_type'].decode() == 'E3_equivariant_model' and md['chemical_symbols_to_index'].decode() == 'Hf O '
d = np.load({str(tmp_path / 'inp.npz')!r})
p = np.random.default_rng(0).permutation(d['edge_index'].shape[1])
dev = 'cuda'
cell = torch.tensor(d['cell'], dtype=torch.float32)
data = {{'x': torch.tensor(d['x']).to(dev), 'pos': torch.tensor(d['pos'], dtype=torch.float32).to(dev).requires_grad_(True),
        'edge_index': torch.tensor(d['edge_index'][:, p]).to(dev), 'num_atoms': torch.tensor([len(d['x'])]).to(dev),
        'cell_lattice_vectors': cell.to(dev), 'cell_volume': torch.dot(cell[0], torch.cross(cell[1], cell[2], dim=0)).to(dev),
        'pbc_shift': torch.tensor(d['pbc_shift'][p], dtype=torch.float32).to(dev)}}
out = m(data)
np.savez({str(tmp_path / 'out.npz')!r}, energy=out['inferred_total_energy'].item(), force=out['inferred_force'].cpu().numpy(),
         stress=out['inferred_stress'].cpu().numpy(), atomic_energy=out['atomic_energy'].cpu().numpy())
import sys
assert 'sevennet_exp_b200' not in sys.modules
"""
    r = subprocess.run([sys.executable, '-c', code], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    out = np.load(tmp_path / 'out.npz')
    n = len(numbers)
    assert abs(float(out['energy']) - float(gold['energy'])) / n < 1e-5
    assert np.abs(out['atomic_energy'].reshape(-1) - gold['atomic_energy'].reshape(-1)).max() < 1e-5
    assert np.abs(out['force'] - gold['force']).max() < 1e-4
    assert np.abs(out['stress'] - gold['stress']).max() * KBAR < 1e-2            # the reference ran in fp32


def test_cpp_libtorch_client_like_the_lammps_pair_style(lib, golden_dir, tmp_path):
    """csrc/pair_harness.cpp is a libtorch program that does what `pair_style e3gnn` does (pair_e3gnn.cpp:306-335 coeff,
    :147-264 compute): dlopen of the op library, torch::jit::load with `_extra_files`, full neighbour list with strict
    r < rc, the serial input dict, forward, the four outputs.  Run on the frame of the REFERENCE's own deployed model
    (tests/golden/deployed_serial_hfo2.npz) with our archive of the same weights."""
    import subprocess
    from sevennet_exp_b200 import build as B, deploy as D
    from tests.test_oracle import deployed_serial_case
    cfg, sd, numbers, pos, cell, gold = deployed_serial_case(golden_dir)
    path = D.deploy((cfg, sd), str(tmp_path / 'deployed_serial.pt'))
    exe = B.build_pair_harness(verbose=False)
    with open(tmp_path / 'structure.txt', 'w') as f:
        f.write(f'{len(numbers)}\n' + ' '.join(repr(float(v)) for v in np.asarray(cell).reshape(-1)) + '\n')
        for z, r in zip(numbers, pos):
            f.write(f'{D._SYMBOLS[int(z)]} {float(r[0])!r} {float(r[1])!r} {float(r[2])!r}\n')
    r = subprocess.run([exe, D.TORCH_OP_LIB, path, str(tmp_path / 'structure.txt'), str(tmp_path / 'result.txt')],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout[-500:], r.stderr[-2000:])
    assert f'{gold["edge_index"].shape[1]} edges' in r.stdout          # same edge count as the reference's graph
    rows = [list(map(float, ln.split())) for ln in open(tmp_path / 'result.txt')]
    n = len(numbers)
    energy, stress, per_atom = rows[0][0], np.array(rows[1]), np.array(rows[2:])
    assert abs(energy - float(gold['energy'])) / n < 1e-5
    assert np.abs(per_atom[:, 3] - gold['atomic_energy'].reshape(-1)).max() < 1e-5
    assert np.abs(per_atom[:, :3] - gold['force']).max() < 1e-4
    assert np.abs(stress - gold['stress']).max() * KBAR < 1e-2             # the reference ran in fp32


def test_calculator_takes_an_nn_module_instance(lib):
    """reference sevennet_calculator.py:111-124: SevenNetCalculator(model=<AtomGraphSequential>, file_type='model_instance')."""
    from sevennet_exp_b200 import nn as bnn
    from sevennet_exp_b200.sevennet_calculator import SevenNetCalculator
    cfg, om = _random_model('test-so3-l2')
    model = bnn.build_E3_equivariant_model(cfg)
    assert not model.load_state_dict(om.state_dict(), strict=False)[0]
    numbers, pos, cell = random_cell(48, list(cfg['_type_map']), density=0.05, seed=3)
    a = SevenNetCalculator(model, file_type='model_instance').compute(numbers, pos, cell, [True] * 3)
    b = SevenNetCalculator((cfg, om.state_dict()), file_type='model_instance').compute(numbers, pos, cell, [True] * 3)
    assert a['energy'] == b['energy'] and np.array_equal(a['forces'], b['forces'])
    with pytest.raises(ValueError):
        SevenNetCalculator(torch.nn.Linear(2, 2), file_type='model_instance')

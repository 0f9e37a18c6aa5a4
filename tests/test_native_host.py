"""C++ whole-model host (csrc/model.cu through sevenn_model_* of the C ABI) against the per-launch Python driver
(engine.py) on the same kernels -- results must be identical -- and, through the calculator, against the oracle
(tests/test_gpu_parity.py runs with the C++ host by default)."""
import os

import numpy as np
import pytest
import torch

from oracle import sevenn_ref as S
from tests.structures import diamond_si, fluorite_hfo2, random_cell
from tests.test_gpu_parity import _random_model

pytestmark = pytest.mark.gpu


def _both(cfg, sd, structures, modal=None, tc_min_rows=None):
    from sevennet_exp_b200.sevennet_calculator import SevenNetCalculator
    res = {}
    for host in ('native', 'python'):
        os.environ['SEVENN_HOST'] = host
        try:
            calc = SevenNetCalculator((dict(cfg), sd), file_type='model_instance', modal=modal)
        finally:
            os.environ.pop('SEVENN_HOST')
        assert (calc.native is not None) == (host == 'native')
        if tc_min_rows is not None:
            calc.engine.tc_min_rows = tc_min_rows
        res[host] = [calc.compute(*s) for s in structures]
        if host == 'native':
            assert calc.native.n_launch > 0
    for a, b in zip(res['native'], res['python']):
        assert a['n_edges'] == b['n_edges']
        # same kernels, same arguments, same order of additions: bit-identical
        assert a['energy'] == b['energy']
        assert np.array_equal(a['forces'], b['forces'])
        assert np.array_equal(a['energies'], b['energies'])
        assert np.array_equal(a['inferred_stress'], b['inferred_stress'])
    return res['native']


def test_native_equals_python_driver_small_models(golden_dir):
    cp = torch.load(os.path.join(golden_dir, 'cp0_model.pt'))
    fr = np.load(os.path.join(golden_dir, 'hfo2_frames.npz'))
    frames = [tuple(fr[f'f{k}_{n}'] for n in ('numbers', 'pos', 'cell', 'pbc')) for k in range(2)]
    _both(cp['config'], cp['model_state_dict'], frames)           # O3 parity, nequip (species-indexed) self connection
    for name in ('test-so3-l2', 'test-so3-l3'):
        cfg, model = _random_model(name)
        zs = list(cfg['_type_map'])
        structs = []
        for k, n in enumerate((48, 17, 5)):
            numbers, pos, cell = random_cell(n, zs, density=0.05, seed=20 + k)
            structs.append((numbers, pos, cell, [True] * 3))
        numbers, pos, cell = random_cell(36, zs, density=0.05, seed=5, triclinic=False)
        structs.append((numbers, pos, cell, [False] * 3))                         # cluster
        structs.append((numbers, pos, cell, [True, True, False]))                 # slab
        structs.append((np.array([zs[0]]), np.zeros((1, 3)), np.eye(3) * 30.0, [True] * 3))   # no neighbours
        out = _both(cfg, model.state_dict(), structs)
        assert out[-1]['n_edges'] == 0 and np.all(out[-1]['forces'] == 0)


def test_native_equals_python_driver_bench_shapes():
    cfg, model = _random_model('7net-0')
    s1 = diamond_si((2, 2, 2)) + ([True] * 3,)
    s2 = fluorite_hfo2((5, 5, 5), a=4.8) + ([True] * 3,)
    _both(cfg, model.state_dict(), [s1, s2])
    _both(cfg, model.state_dict(), [s1], tc_min_rows=1)
    cfg, model = _random_model('7net-l3i5')
    numbers, pos, cell = random_cell(40, [3, 15, 16], density=0.048, seed=1, rmin=1.9)
    _both(cfg, model.state_dict(), [(numbers, pos, cell, [True] * 3)])


def test_native_modal_bias():
    from tests.test_gpu_parity import _modal_cfg
    cfg = _modal_cfg(True)
    torch.manual_seed(11)
    model = S.build_model(cfg).eval()
    with torch.no_grad():
        for n_, p_ in model.named_parameters():
            if n_.endswith('linear.bias'):
                p_.normal_(0.0, 0.3)
    numbers, pos, cell = random_cell(40, list(cfg['_type_map']), density=0.05, seed=6)
    _both(cfg, model.state_dict(), [(numbers, pos, cell, [True] * 3)], modal='scan')


def test_native_edge_capacity_overflow_recovers():
    """A structure whose neighbour list outgrows the capacity sized from the first call: the status word reports it,
    the host grows the workspace and repeats the step (no out-of-bounds write, right answer)."""
    from sevennet_exp_b200.sevennet_calculator import SevenNetCalculator
    cfg, model = _random_model('test-so3-l2')
    zs = list(cfg['_type_map'])
    calc = SevenNetCalculator((cfg, model.state_dict()), file_type='model_instance')
    numbers, pos, cell = random_cell(64, zs, density=0.03, seed=2)
    a = calc.compute(numbers, pos, cell, [True] * 3)
    dense = 0.45 * cell                                       # same atom count, 11x the density
    b = calc.compute(numbers, pos * 0.45, dense, [True] * 3)
    assert b['n_edges'] > 2 * a['n_edges'] + 2048
    os.environ['SEVENN_HOST'] = 'python'
    try:
        ref = SevenNetCalculator((cfg, model.state_dict()), file_type='model_instance').compute(
            numbers, pos * 0.45, dense, [True] * 3)
    finally:
        os.environ.pop('SEVENN_HOST')
    assert b['n_edges'] == ref['n_edges'] and b['energy'] == ref['energy'] and np.array_equal(b['forces'], ref['forces'])


def test_cpp_md_host_without_torch_or_python(tmp_path):
    """sevenn_md_host (csrc/md_host.cpp): a C++ program that links libsevenn_b200.so only -- the calls a LAMMPS pair style
    would make (sevenn/pair_e3gnn/pair_e3gnn.cpp:71-274) -- runs velocity-Verlet steps on SevenNet-0-shaped Si-64.
    Step-0 energy / forces / virial must be the calculator's (same kernels: bit-identical in fp32), and the total
    energy must be conserved along the trajectory (forces are the exact gradient of the energy)."""
    import subprocess
    from sevennet_exp_b200 import build, host
    from sevennet_exp_b200.sevennet_calculator import SevenNetCalculator
    exe = build.MD_HOST
    assert os.path.exists(exe), 'run python -m sevennet_exp_b200.build'
    cfg, model = _random_model('7net-0')
    calc = SevenNetCalculator((cfg, model.state_dict()), file_type='model_instance')
    numbers, pos, cell = diamond_si((2, 2, 2))
    n = len(numbers)
    host.save_model(str(tmp_path / 'model.bin'), calc.engine)
    rng = np.random.default_rng(0)
    vel = rng.normal(0.0, 0.004, (n, 3))                     # A/fs, ~ 400 K for Si
    vel -= vel.mean(0)
    with open(tmp_path / 'structure.bin', 'wb') as f:
        np.asarray([n], dtype=np.int64).tofile(f)
        np.asarray(cell, dtype=np.float64).reshape(9).tofile(f)
        np.asarray([1, 1, 1], dtype=np.int32).tofile(f)
        calc._lut[numbers].astype(np.int32).tofile(f)
        np.asarray(pos, dtype=np.float64).tofile(f)
        np.full(n, 28.0855).tofile(f)
        vel.tofile(f)
    steps = 40
    r = subprocess.run([exe, str(tmp_path / 'model.bin'), str(tmp_path / 'structure.bin'), str(steps), '1.0',
                        str(tmp_path / 'out.bin')], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    raw = np.fromfile(tmp_path / 'out.bin', dtype=np.float64)
    rec, f0 = raw[:8 * (steps + 1)].reshape(steps + 1, 8), raw[8 * (steps + 1):].reshape(n, 3)
    ref = calc.compute(numbers, pos, cell, [True] * 3)
    assert rec[0, 0] == ref['energy']
    assert np.array_equal(f0, ref['forces'])
    vol = abs(np.linalg.det(cell))
    assert np.array_equal(-rec[0, 2:] / vol, ref['inferred_stress'])
    etot = rec[:, 0] + rec[:, 1]
    swing = np.ptp(rec[:, 1])
    print(r.stdout[-400:], 'kinetic swing', swing, 'drift', np.abs(etot - etot[0]).max())
    assert swing > 1e-4                                        # atoms really moved
    assert np.abs(etot - etot[0]).max() < 0.02 * swing + 1e-6

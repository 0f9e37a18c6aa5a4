"""tests/properties.py checked on CPU with the oracle as backend (small cell: logic and sign conventions), and on the
GPU at BASELINE's full config-2 size (4,096-atom HfO2, SevenNet-0 shape) through SevenNetCalculator."""
import numpy as np
import pytest
import torch

from oracle import sevenn_ref as S
from sevennet_exp_b200 import presets
from tests.properties import check_properties
from tests.structures import random_cell


def test_properties_with_oracle_backend():
    cfg = dict(presets.PRESETS['test-so3-l2'])
    cfg['scale'] = [10.0] * cfg['_number_of_species']
    torch.manual_seed(5)
    torch.set_default_dtype(torch.float64)
    try:
        model = S.build_model(cfg).double().eval()

        def compute(numbers, pos, cell):
            g = S.atoms_to_graph(numbers, pos, cell, [True] * 3, cfg['cutoff'], dtype=torch.float64)
            r = model(dict(g))
            return {'energy': float(r[S.E_TOT].detach()), 'forces': r[S.FORCE].detach().numpy(),
                    'inferred_stress': r[S.STRESS].detach().numpy().reshape(6), 'n_edges': g[S.EDGE_IDX].shape[1]}

        numbers, pos, cell = random_cell(16, list(cfg['_type_map']), density=0.05, seed=2)
        out = check_properties(compute, numbers, pos, cell, tol_e=1e-9, tol_f=1e-9, tol_s_kbar=1e-6, fd_rel=1e-3)
    finally:
        torch.set_default_dtype(torch.float32)
    print(out)


def test_properties_fp32_emulation_bench_model():
    """The bench model (SevenNet-0 shape, bench.random_state_dict) on a 96-atom HfO2 cell through the fp32 EMULATION build
    of the kernel sources: the same tolerances the full-size GPU test uses hold with a wide margin."""
    import bench as B
    from sevennet_exp_b200 import spec as sp
    from sevennet_exp_b200.engine import Engine
    from tests.emu_util import build_full_emu, graph_from_edges
    cfg = B.model_config()
    spec = sp.build_spec(cfg)
    eng = Engine(spec, B.random_state_dict(cfg), lib=build_full_emu([L.conv for L in spec.layers]), device='cpu',
                 radial='table')
    tmap = cfg['_type_map']

    def compute(numbers, pos, cell):
        g = S.atoms_to_graph(numbers, pos, cell, [True] * 3, cfg['cutoff'])
        types, vol = np.array([tmap[int(z)] for z in numbers]), float(g[S.VOL])
        out = eng.evaluate(graph_from_edges(types, g[S.EDGE_IDX].numpy(), g[S.EDGE_VEC].detach().numpy(),
                                            len(numbers), vol))
        return {'energy': out['energy'].item(), 'forces': out['forces'].double().numpy(),
                'inferred_stress': -(out['virial6'].numpy() / vol), 'n_edges': g[S.EDGE_IDX].shape[1]}

    numbers, pos, cell = B.workload(reps=(2, 2, 2), n_remove=0)
    print(check_properties(compute, numbers, pos, cell, tol_e=2e-6, tol_f=2e-5, tol_s_kbar=2e-3))


@pytest.mark.gpu
def test_properties_full_size_cfg2_on_gpu():
    """BASELINE config 2 at full size (the bench workload): 4,096-atom HfO2, SevenNet-0 hyper-parameters."""
    import bench as B
    from sevennet_exp_b200.sevennet_calculator import SevenNetCalculator
    cfg = B.model_config()
    calc = SevenNetCalculator((cfg, B.random_state_dict(cfg)), file_type='model_instance')
    numbers, pos, cell = B.workload()
    assert len(numbers) == 4096
    # random-init weights give |E| ~ 1.2 eV/atom, forces of O(1e-3 eV/A), stress ~ 1 kbar on this cell: the fp32 kernels
    # reproduce a rotated / translated copy far inside the north_star tolerances (1e-5 eV/atom, 1e-4 eV/A, 1e-3 kbar)
    out = check_properties(lambda z, x, c: calc.compute(z, x, c, [True] * 3), numbers, pos, cell,
                           tol_e=2e-6, tol_f=2e-5, tol_s_kbar=2e-3)
    print(out)


@pytest.mark.gpu
def test_properties_full_size_cfg3_on_gpu():
    """BASELINE config 3 at full size: SevenNet-l3i5 hyper-parameters (lmax 3), 32,768-atom disordered Li-P-S cell."""
    import bench as B
    from sevennet_exp_b200.sevennet_calculator import SevenNetCalculator
    from tests.structures import jittered_lips
    cfg = dict(presets.PRESETS['7net-l3i5'])
    calc = SevenNetCalculator((cfg, B.random_state_dict(cfg)), file_type='model_instance')
    numbers, pos, cell = jittered_lips()
    assert len(numbers) == 32768
    out = check_properties(lambda z, x, c: calc.compute(z, x, c, [True] * 3), numbers, pos, cell,
                           tol_e=2e-6, tol_f=2e-5, tol_s_kbar=2e-3)
    print(out)

"""Single-GPU scale check: SevenNet-0 (random weights) on an n x n x n diamond-Si supercell (8 n^3 atoms)."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench as B
from sevennet_exp_b200.sevennet_calculator import SevenNetCalculator
from tests.structures import diamond_si

n = int(os.environ.get('CELLS', '25'))
model = os.environ.get('MODEL', '7net-0')
cfg = B.model_config() if model == '7net-0' else dict(__import__('sevennet_exp_b200.presets', fromlist=['PRESETS']).PRESETS[model])
calc = SevenNetCalculator((cfg, B.random_state_dict(cfg)), file_type='model_instance',
                          **({'keep_weights': False} if os.environ.get('RECOMPUTE') else {}))
num, pos, cell = diamond_si((n, n, n))
rng = np.random.default_rng(2)
pos = pos + rng.normal(0, 0.05, pos.shape)
if model != '7net-0':
    num = np.array(list(cfg['_type_map']))[rng.integers(0, len(cfg['_type_map']), len(num))]
print('atoms', len(num), flush=True)
for it in range(3):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    res = calc.compute(num, pos, cell, [True] * 3)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print(f'iter {it}: {dt * 1e3:.1f} ms, {len(num) / dt:.0f} atom-steps/s, edges {res["n_edges"]}, '
          f'peak mem {torch.cuda.max_memory_allocated() / 2**30:.1f} GiB, E/atom {res["energy"] / len(num):.6f}', flush=True)

"""Top kernels of one cfg5 training step (torch.profiler, CUDA activities)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench as B
from torch.profiler import profile, ProfilerActivity

class A: pass
def main():
    from sevennet_exp_b200 import nn as bnn, train as TR
    from sevennet_exp_b200.neighbor import NeighborList
    dev = torch.device('cuda:0')
    cfg = B.model_config()
    model = bnn.build_E3_equivariant_model(cfg)
    model.load_state_dict(B.random_state_dict(cfg), strict=False)
    model = model.to(dev)
    nl = NeighborList(device=dev)
    tmap = cfg['_type_map']
    graphs = []
    for numbers, pos, cell, e, f, st in B.cfg5_structures():
        r = nl.build(pos, cell, [True] * 3, cfg['cutoff'])
        graphs.append({TR.K_X: torch.tensor(numbers), TR.K_Z: torch.tensor(numbers), TR.K_POS: torch.tensor(pos, dtype=torch.float32),
                       TR.K_EI: torch.stack([r['center'], r['col']]).long().cpu(), TR.K_EV: r['edge_vec'].cpu(),
                       TR.K_VOL: torch.tensor(r['volume'], dtype=torch.float32), TR.K_NATOMS: torch.tensor(len(numbers)),
                       TR.K_E: torch.tensor(e, dtype=torch.float32), TR.K_F: torch.tensor(f, dtype=torch.float32),
                       TR.K_S: torch.tensor(st, dtype=torch.float32)})
    tr = TR.Trainer(model, {'optim_param': {}})
    batch = TR.collate(graphs, device=dev)
    for _ in range(3):
        tr.train_step(batch)
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        tr.train_step(batch)
        torch.cuda.synchronize()
    print(prof.key_averages().table(sort_by='cuda_time_total', row_limit=25, max_name_column_width=70))
main()

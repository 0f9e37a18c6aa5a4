"""bench.py --gpus N (N > 1): weak-scaling domain-decomposed E+F+S (one rank per GPU, NCCL over NVLink).

The 4,096-atom HfO2 workload cell is replicated over the brick grid (2x1x1, 2x2x1, 2x2x2), each rank owns one
replica (~4,096 atoms) plus its ghost shell; per step every rank rebuilds the neighbour list of its owned +
ghost atoms, runs the layers with a ghost-feature exchange before each convolution >= 1, the reverse pass
with the mirror exchange, folds ghost forces back to their owners and all-reduces energy and virial."""
import json
import os
import time

import numpy as np
import torch
import torch.distributed as dist


def run(args):
    import bench as B
    from .domain import Decomposition, Halo, brick_grid, evaluate_decomposed
    from .sevennet_calculator import SevenNetCalculator
    rank, world = int(os.environ.get('RANK', '0')), int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    dev = torch.device(f'cuda:{local}')
    torch.cuda.set_device(dev)
    if not dist.is_initialized():
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')          # plain `python bench.py --workload si1m` (N = 1)
        os.environ.setdefault('MASTER_PORT', '29555')
        dist.init_process_group('nccl', rank=rank, world_size=world, device_id=dev)
    grid = brick_grid(world)
    strong = getattr(args, 'workload', 'si1m') == 'si1m'
    if strong:
        # BASELINE config 4: one bulk-Si cell (50^3 conventional cells = 1,000,000 atoms, rattle 0.05 A, seed 2),
        # strong scaling: the same cell is cut into `world` bricks
        _, numbers, pos, cell = B.make_workload('si1m', args.si_cells)
    else:
        numbers0, pos0, cell0 = B.workload()
        reps = np.array(grid)
        shifts = np.stack(np.meshgrid(*[np.arange(r) for r in reps], indexing='ij'), -1).reshape(-1, 3)
        pos = np.concatenate([pos0 + s @ cell0 for s in shifts])
        numbers = np.tile(numbers0, len(shifts))
        cell = cell0 * reps[:, None]
    n_total = len(numbers)

    cfg = B.model_config()
    calc = SevenNetCalculator((cfg, B.random_state_dict(cfg)), file_type='model_instance', device=dev,
                              keep_weights=not strong)
    # decomposed == single-domain evaluation of the whole cell on one GPU (the gate of the reference's
    # tests/lammps_tests/test_lammps.py:278-316), always on, outside the timed region: every rank evaluates the whole
    # periodic cell by itself first (it fits one B200), keeps E, the forces of the atoms it will own and the stress
    ref = calc.compute(numbers, pos, cell, [True] * 3)
    ref_E, ref_S = ref['energy'], ref['inferred_stress']
    dec = Decomposition(numbers, pos, cell, calc.cutoff, grid, rank)
    ref_F = ref['forces'][dec.local_ids]
    del ref
    torch.cuda.empty_cache()
    halo = Halo(dec, calc.engine.lib, dev)
    types_dev = torch.from_numpy(calc._lut[dec.numbers].astype(np.int32)).to(dev)
    pos_dev = torch.from_numpy(dec.pos).to(dev)
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)

    def step():
        return evaluate_decomposed(calc, dec, halo, types_dev=types_dev, pos_dev=pos_dev)

    for _ in range(max(args.warmup, 3)):
        E, F, vir, g = step()
    chk = torch.tensor([abs(E.item() - ref_E) / n_total, float(np.abs(F.cpu().numpy() - ref_F).max()),
                        float(np.abs(-(vir.cpu().numpy() / dec.volume) - ref_S).max() * 1602.1766208)],
                       dtype=torch.float64, device=dev)
    dist.all_reduce(chk, op=dist.ReduceOp.MAX)
    check = {'dE_eV_per_atom': chk[0].item(), 'dF_eV_per_A': chk[1].item(), 'dS_kbar': chk[2].item(),
             'against': 'single-GPU evaluation of the whole periodic cell, max over ranks',
             'ok': bool(chk[0].item() < 1e-5 and chk[1].item() < 1e-4 and chk[2].item() < 1e-3)}
    del ref_F
    sampler = B.ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    l0 = calc.n_launch
    torch.cuda.synchronize()
    dist.barrier()
    evs = []
    for _ in range(args.steps):
        flush.fill_(1.0)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        E, F, vir, g = step()
        b.record()
        evs.append((a, b))
    torch.cuda.synchronize()
    dist.barrier()
    t = torch.tensor([sum(a.elapsed_time(b) for a, b in evs)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = t.item() / args.steps
    launches = calc.n_launch - l0

    # end to end: host positions in, owned forces + energy + virial out, every step
    pos_host = torch.from_numpy(dec.pos).pin_memory()
    host_F = torch.empty(dec.n_local, 3, dtype=torch.float32).pin_memory()      # results land in pinned buffers
    host_EV = torch.empty(7, dtype=torch.float64).pin_memory()
    pd = torch.empty_like(pos_dev)
    torch.cuda.synchronize()
    dist.barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        flush.fill_(1.0)
        pd.copy_(pos_host, non_blocking=True)
        E, F, vir, g = evaluate_decomposed(calc, dec, halo, types_dev=types_dev, pos_dev=pd)
        host_F.copy_(F, non_blocking=True)
        host_EV.copy_(torch.cat([E.reshape(1), vir]), non_blocking=True)
        torch.cuda.current_stream().synchronize()
    torch.cuda.synchronize()
    dist.barrier()
    te = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_t = te.item() / args.steps
    if os.environ.get('SEVENN_E2E_DEBUG'):       # where one end-to-end step spends its wall time (printed by rank 0)
        import sys
        for _ in range(3):
            marks = [time.perf_counter()]
            pd.copy_(pos_host, non_blocking=True); torch.cuda.synchronize(); marks.append(time.perf_counter())
            E, F, vir, g = evaluate_decomposed(calc, dec, halo, types_dev=types_dev, pos_dev=pd); marks.append(time.perf_counter())
            torch.cuda.synchronize(); marks.append(time.perf_counter())
            host_F.copy_(F, non_blocking=True); host_EV.copy_(torch.cat([E.reshape(1), vir]), non_blocking=True)
            torch.cuda.synchronize(); marks.append(time.perf_counter())
            if rank == 0:
                print('[e2e debug] h2d %.2f ms, host side of the step %.2f ms, wait for the GPU %.2f ms, d2h %.2f ms' % tuple(
                    1e3 * (b - a) for a, b in zip(marks[:-1], marks[1:])), file=sys.stderr, flush=True)
    stats = torch.tensor([dec.n_local, dec.n_ghost, g.n_edges], dtype=torch.float64, device=dev)
    smax = stats.clone()
    dist.all_reduce(stats)
    dist.all_reduce(smax, op=dist.ReduceOp.MAX)
    if sampler:
        sampler.stop_flag = True
        sampler.join(timeout=2)
    # per-stage device times of this rank (separate pass, every stage alone on its launch stream) -> roofline of the
    # dominant kernel on rank 0's sub-domain, same conventions as the single-GPU line
    stage, reps_prof = {}, 2
    for _ in range(reps_prof):
        flush.fill_(1.0)
        calc.engine.prof = []
        E, F, vir, g = evaluate_decomposed(calc, dec, halo, types_dev=types_dev, pos_dev=pos_dev, host='python')
        torch.cuda.synchronize()
        for k, v in calc.engine.profile_summary().items():
            stage[k] = stage.get(k, 0.0) + v / reps_prof
    calc.engine.prof = None
    dist.barrier()
    if rank == 0:
        inner = calc.spec.layers[1].conv
        n_inner = sum(1 for L in calc.spec.layers if L.conv.signature() == inner.signature())
        tab = calc.engine.radial == 'table'
        dom_ms = stage[f'conv_bwd_{inner.signature()}'] / n_inner
        alg = B.conv_bwd_algorithmic_bytes(inner, dec.n_local, g.n_edges)
        peak, how = B.measured_peak()
        touched = alg - (g.n_edges * 8 * inner.wn if tab else 0)
        roofline = {'bound': 'hbm', 'kernel': f'conv_bwd_{"tab_" if tab else ""}{inner.signature()} on rank 0 '
                                              f'({dec.n_local} owned atoms, {g.n_edges} edges)',
                    'achieved': alg / (dom_ms * 1e-3) / 1e9, 'peak': peak, 'peak_source': how, 'unit': 'GB/s',
                    'frac': alg / (dom_ms * 1e-3) / 1e9 / peak, 'algorithmic_bytes': alg, 'ms': dom_ms, 'traffic': None,
                    'elided': 'w and dW [E, wn] (tabulated radial weights)' if tab else None,
                    'touched_bytes': touched, 'touched_frac': touched / (dom_ms * 1e-3) / 1e9 / peak}
        stage_sorted = dict(sorted(((k, round(v, 4)) for k, v in stage.items()), key=lambda kv: -kv[1])[:12])
        # ghost exchanges of one step on rank 0 (CUDA events around the exchange calls of the profiling pass: pack + NVLink
        # write + device-side barrier, i.e. including the wait for the slowest neighbour) and the bytes this rank sends
        n_send = int(halo.send_off[-1])
        halo_info = {'fwd_ms': round(stage.get('halo_fwd', 0.0), 4), 'bwd_ms': round(stage.get('halo_bwd', 0.0), 4),
                     'exchanges_per_step': 2 * (len(calc.spec.layers) - 1) + 1,
                     'rows_sent_rank0': n_send, 'rows_received_rank0': int(halo.recv_off[-1]),
                     'bytes_sent_per_exchange_rank0': n_send * calc.spec.layers[1].conv.lx.dim * 4,
                     'share_of_step': round((stage.get('halo_fwd', 0.0) + stage.get('halo_bwd', 0.0)) / ms_step, 4)}
        x_dim = calc.spec.layers[1].conv.lx.dim
        line = {
            'metric': B.METRIC, 'value': n_total / (ms_step * 1e-3), 'unit': B.UNIT, 'n_gpus': world,
            'steps': args.steps, 'warmup': max(args.warmup, 3), 'ms_per_step': ms_step, 'higher_is_better': True,
            'scaling': 'strong' if strong else 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
            'config': {'workload': (B.WORKLOAD_TEXT['si1m'].format(n=n_total, c=args.si_cells, nb=stats[2].item() / n_total)
                                    + f', cut into {grid[0]}x{grid[1]}x{grid[2]} bricks, ghost feature exchange per layer'
                                    if strong else
                                    f'SevenNet-0 E+F+S, bulk HfO2 cfg2 cell replicated {grid[0]}x{grid[1]}x{grid[2]} '
                                    f'({n_total} atoms), brick domain decomposition, ghost feature exchange per layer'),
                       'n_atoms': n_total, 'atoms_per_gpu': n_total // world, 'parallelism': f'dd{world} (spatial bricks {grid[0]}x{grid[1]}x{grid[2]})', 'energy_per_atom_eV': float(E.item()) / n_total,
                       'ghosts_per_gpu_max': int(smax[1].item()), 'edges_total': int(stats[2].item()),
                       'halo_row_bytes': x_dim * 4,
                       'halo': ('peer memory: pack + NVLink write into the neighbour\'s slot in one kernel, device-side '
                                'barrier' if halo.peer is not None else 'NCCL all-to-all-v per exchange'),
                       'l2': '256 MiB write between timed steps (outside the CUDA-event brackets)',
                       'step': 'device neighbour list (owned+ghost) + forward + reverse pass + halo exchanges'},
            'e2e': {'value': n_total / e2e_t, 'unit': B.UNIT, 'ms_per_step': e2e_t * 1e3,
                    'h2d_bytes_per_step': int(pos_host.numel() * 8),
                    'd2h_bytes_per_step': int(7 * 8 + 3 * dec.n_local * 4)},
            'gpu_launches': int(launches), 'clocks': sampler.summary() if sampler else None,
            'check': check, 'roofline': roofline, 'stage_ms_rank0': stage_sorted, 'halo_rank0': halo_info,
            'cpu_baseline': None,
        }
        print(json.dumps(line), flush=True)
    dist.barrier()
    dist.destroy_process_group()

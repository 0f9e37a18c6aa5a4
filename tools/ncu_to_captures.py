"""profiles/ncu_captures.json from the committed `ncu --page raw --csv` files: per kernel and workload the DRAM bytes of
one launch, the counters that name its limiter and (GEMMs) the tensor-pipe utilisation -- what bench.py prints as
roofline.traffic / limiter / roofline_gemm.tensor.ncu_tensor_pipe_pct."""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EDGES = {'si1m': 28000042, 'cfg2': 198356, 'cfg3': 794000}
UNIT = {'Gbyte': 1e9, 'Mbyte': 1e6, 'Kbyte': 1e3, 'byte': 1.0}


def main(files):
    out = []
    for f in files:
        rows = list(csv.reader(open(f)))
        h, units, body = rows[0], rows[1], rows[2:]
        wl = next((w for w in EDGES if f'_{w}' in os.path.basename(f)), None)
        col = {n: i for i, n in enumerate(h)}

        def val(r, name, scale=False):
            if name not in col:
                return None
            v = float(r[col[name]].replace(',', ''))
            return v * UNIT.get(units[col[name]], 1.0) if scale else v
        seen = set()
        for r in body:
            name = r[col['Kernel Name']].split('(')[0].replace('void ', '')
            if name in seen and 'gemm' not in name:
                continue
            seen.add(name)
            dram = val(r, 'dram__bytes_read.sum', True) + val(r, 'dram__bytes_write.sum', True)
            # tcgen05 (UTCHMMA) work shows in the tensor pipe's cycles-active counter, not in the legacy HMMA instruction counter
            tensor = val(r, 'TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed')
            if tensor is None:
                tensor = val(r, 'sm__inst_executed_pipe_tensor_subpipe_hmma.avg.pct_of_peak_sustained_active')
            lim = (f"ncu r02: issue slots {val(r, 'smsp__issue_active.avg.pct_of_peak_sustained_active'):.0f} %, fma pipe "
                   f"{val(r, 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active'):.0f} %, L1 {val(r, 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed'):.0f} %, "
                   f"L2 {val(r, 'lts__throughput.avg.pct_of_peak_sustained_elapsed'):.0f} %, warps active "
                   f"{val(r, 'sm__warps_active.avg.pct_of_peak_sustained_active'):.0f} %, {int(val(r, 'launch__registers_per_thread'))} registers")
            out.append({'kernel': name, 'workload': wl, 'n_edges': 0 if 'gemm' in name else EDGES.get(wl, 0),
                        'dram_bytes': dram, 'dram_read': val(r, 'dram__bytes_read.sum', True),
                        'dram_write': val(r, 'dram__bytes_write.sum', True),
                        'duration_us': val(r, 'gpu__time_duration.sum') * (1e3 if units[col['gpu__time_duration.sum']] == 'ms' else 1.0),
                        'instructions': val(r, 'smsp__inst_executed.sum'), 'tensor_pipe_pct': tensor, 'limiter': lim,
                        'source': 'profiles/' + os.path.basename(f)})
    json.dump(out, open(os.path.join(ROOT, 'profiles', 'ncu_captures.json'), 'w'), indent=1)
    for o in out:
        print(o['kernel'], o['workload'], f"{o['dram_bytes'] / 1e6:.1f} MB", f"{o['duration_us']:.1f} us", o['tensor_pipe_pct'])


if __name__ == '__main__':
    main(sys.argv[1:])

#!/bin/bash
# launch list (gpu__time_duration) of one workload: tools/gpu_ncu_list.sh <workload> [extra bench args]
mkdir -p gpurun_out
w=$1; shift
SEVENN_BENCH_NO_CPU=1 timeout -s KILL 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_$w.csv python bench.py --workload $w --steps 2 --warmup 1 "$@" > gpurun_out/ncu_$w.log 2>&1
python - $w <<'PY'
import csv, sys, collections
lines = [l for l in open(f'gpurun_out/launches_{sys.argv[1]}.csv') if not l.startswith('==')]
rows = [d for d in csv.DictReader(lines) if d.get('Metric Name') == 'gpu__time_duration.sum']
agg = collections.OrderedDict()
for d in rows:
    k = d['Kernel Name'][:70]
    t = float(d['Metric Value'].replace(',', ''))
    a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += t
tot = sum(v[1] for v in agg.values())
print('launches', len(rows), 'total us', tot / 1e3)
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:28]:
    print(f'{v[0]:5d} {v[1] / 1e3:10.1f} us  {v[1] / v[0] / 1e3:8.2f} us/launch  {k}')
PY

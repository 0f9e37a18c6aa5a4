#!/bin/bash
# the three inference workloads on one box with the current library: prints value / ms_per_step / first stages
mkdir -p gpurun_out
for w in cfg2 cfg3 si1m cfg1; do
  SEVENN_BENCH_NO_CPU=1 timeout -s KILL 600 python bench.py --workload $w --steps 10 --warmup 3 > gpurun_out/b3_$w.json 2> gpurun_out/b3_$w.err || { echo "bench $w FAILED"; tail -3 gpurun_out/b3_$w.err; continue; }
  python - $w <<'PY'
import json, sys
d = json.loads(open(f'gpurun_out/b3_{sys.argv[1]}.json').read().strip().splitlines()[-1])
print(sys.argv[1], round(d['value']), round(d['ms_per_step'], 3), 'e2e', round(d['e2e']['value']), {k: round(v, 3) for k, v in list(d.get('stage_ms', {}).items())})
PY
done

#!/bin/bash
# small-system latency (cfg1) + cfg2 + GPU tests
mkdir -p gpurun_out
for w in cfg1 cfg2; do
  SEVENN_BENCH_NO_CPU=1 timeout -s KILL 600 python bench.py --workload $w --steps 20 --warmup 5 > gpurun_out/b3_$w.json 2> gpurun_out/b3_$w.err || { echo "bench $w FAILED"; tail -3 gpurun_out/b3_$w.err; continue; }
  python - $w <<'PY'
import json, sys
d = json.loads(open(f'gpurun_out/b3_{sys.argv[1]}.json').read().strip().splitlines()[-1])
print(sys.argv[1], round(d['value']), round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['value']), 'launches/step', d['gpu_launches'] / d['steps'], {k: round(v, 3) for k, v in list(d.get('stage_ms', {}).items())[:12]})
PY
done
timeout -s KILL 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -3

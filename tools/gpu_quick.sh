#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
for w in cfg2 cfg3; do
python bench.py --workload $w --steps 20 --warmup 3 > gpurun_out/bench_$w.json.log 2> gpurun_out/bench_$w.err; echo "bench $w exit $?"
python - $w <<'PY'
import json, sys
d = json.loads(open(f'gpurun_out/bench_{sys.argv[1]}.json.log').read())
print(round(d['value']), round(d['ms_per_step'], 3), 'e2e', round(d['e2e']['value']), {k: v for k, v in list(d['stage_ms'].items())[:8]})
PY
done

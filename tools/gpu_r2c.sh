#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/smoke.log
for w in cfg1 cfg2 cfg3 si1m; do
SEVENN_BENCH_NO_CPU=1 python bench.py --workload $w --steps 20 --warmup 3 > gpurun_out/bench_$w.json.log 2> gpurun_out/bench_$w.err; echo "bench $w exit $?"
done
python - <<'PY'
import json
for w in ('cfg1','cfg2','cfg3','si1m'):
    try:
        d = json.loads(open(f'gpurun_out/bench_{w}.json.log').read())
        print(w, round(d['value']), round(d['ms_per_step'], 3), 'e2e', round(d['e2e']['value']), round(d['e2e']['ms_per_step'],3), 'launches', d['gpu_launches'])
    except Exception as e:
        print(w, 'FAILED', e); print(open(f'gpurun_out/bench_{w}.err').read()[-1500:])
PY

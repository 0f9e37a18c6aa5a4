#!/bin/bash
# r02 bench lines: default (si1m, N=1, with CPU legs), cfg1/cfg2/cfg3 lines, reference arm
mkdir -p gpurun_out
( time python bench.py > gpurun_out/bench_default.json.log 2> gpurun_out/bench_default.err ) 2>&1 | grep real; echo "bench default exit $?"
for w in cfg1 cfg2 cfg3; do
SEVENN_BENCH_NO_CPU=1 python bench.py --workload $w --steps 20 --warmup 3 > gpurun_out/bench_$w.json.log 2> gpurun_out/bench_$w.err; echo "bench $w exit $?"
done
( time python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json.log 2> gpurun_out/bench_reference.err ) 2>&1 | grep real
python - <<'PY'
import json
for w in ('default','cfg1','cfg2','cfg3','reference'):
    try:
        d = json.loads(open(f'gpurun_out/bench_{w}.json.log').read())
        print(w, round(d['value']), round(d['ms_per_step'], 3), 'e2e', round(d['e2e']['value']), d.get('cpu_baseline'), d.get('torch_gpu_baseline'))
        print('   ', {k: v for k, v in list(d.get('stage_ms', {}).items())[:10]})
        if d.get('roofline_gemm'): print('   gemm', d['roofline_gemm']['frac'], d['roofline_gemm']['tensor'])
    except Exception as e:
        print(w, 'FAILED', e); print(open(f'gpurun_out/bench_{w}.err').read()[-1500:])
PY

#!/bin/bash
# final bench lines of the round (driver-style invocations), one JSON line each under gpurun_out/
mkdir -p gpurun_out
timeout -s KILL 1500 python bench.py > gpurun_out/final_si1m_n1.json 2> gpurun_out/final_si1m_n1.err; echo "default exit $?"; tail -c 600 gpurun_out/final_si1m_n1.json | head -c 600; echo
for w in cfg1 cfg2 cfg3 cfg5; do
  timeout -s KILL 1500 python bench.py --workload $w > gpurun_out/final_$w.json 2> gpurun_out/final_$w.err; echo "$w exit $?"
  python - $w <<'PY'
import json, sys
d = json.loads(open(f'gpurun_out/final_{sys.argv[1]}.json').read().strip().splitlines()[-1])
print(sys.argv[1], d['metric'], round(d['value']), round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['value']), 'cpu', d.get('cpu_baseline', {}).get('value'), 'roofline', round(d['roofline']['frac'], 3) if d.get('roofline') else None)
PY
done
timeout -s KILL 1500 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/final_reference.json 2> gpurun_out/final_reference.err; echo "reference exit $?"; tail -c 700 gpurun_out/final_reference.json

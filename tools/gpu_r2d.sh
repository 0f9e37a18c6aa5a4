#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log; tail -3 gpurun_out/pytest_gpu.log
for w in cfg2 si1m cfg3; do
  SEVENN_BENCH_NO_CPU=1 python bench.py --workload $w --steps 10 --warmup 3 --si-cells 30 > gpurun_out/gate_$w.json 2> gpurun_out/gate_$w.err
  python -c "
import json;d=json.loads(open('gpurun_out/gate_$w.json').read().strip().splitlines()[-1]);print('newgate $w',round(d['value']),round(d['ms_per_step'],3),{k:v for k,v in d['stage_ms'].items() if 'gate' in k})"
done
bash tools/gpu_ab2.sh ref s0_64 s0_65 s0_55 l3_44 l3_33
SEVENN_BENCH_NO_CPU=1 python bench.py --workload cfg5 --steps 10 --warmup 3 | python -c "import json,sys;d=json.loads(sys.stdin.read().strip().splitlines()[-1]);print('cfg5',round(d['value']),round(d['ms_per_step'],2),'e2e',round(d['e2e']['ms_per_step'],2),d['gpu_launches'])"

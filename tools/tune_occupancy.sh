#!/bin/bash
# Occupancy-target sweep of the tabulated conv kernels on the GPU box: rebuild with a different
# __launch_bounds__ minimum, run the bench, print the per-kernel stage times.
mkdir -p gpurun_out
run() {  # name, env value, workload
  SEVENN_MIN_CTAS_TAB="$2" python -m sevennet_exp_b200.build > gpurun_out/tune_build_$1.log 2>&1 || { echo "build $1 failed"; tail -3 gpurun_out/tune_build_$1.log; return; }
  grep -i "spill" gpurun_out/tune_build_$1.log | head -3
  python bench.py --workload $3 --steps 5 --warmup 3 2>/dev/null > gpurun_out/tune_$1.json.log
  python - "$1" <<'PY'
import json, sys
d = json.loads(open(f'gpurun_out/tune_{sys.argv[1]}.json.log').read())
conv = {k: v for k, v in d['stage_ms'].items() if k.startswith('conv_')}
print(sys.argv[1], round(d['value']), 'ms/step', round(d['ms_per_step'], 3), conv, flush=True)
PY
}
run l3_base  "c703e622066:3,2" cfg3
run l3_b3    "c703e622066:3,3" cfg3
run l3_f4b3  "c703e622066:4,3" cfg3
run l3_f4b2  "c703e622066:4,2" cfg3
run l3_b4    "c703e622066:3,4" cfg3
run s0_base  "c377d4386a8:6,5" cfg2
run s0_b6    "c377d4386a8:6,6" cfg2
run s0_f8    "c377d4386a8:8,5" cfg2
run s0_b4    "c377d4386a8:6,4" cfg2
run s0_f5    "c377d4386a8:5,5" cfg2
# restore the defaults
python -m sevennet_exp_b200.build > /dev/null 2>&1

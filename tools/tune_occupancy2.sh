#!/bin/bash
mkdir -p gpurun_out
run() {
  SEVENN_MIN_CTAS_TAB="$2" python -m sevennet_exp_b200.build > gpurun_out/tune_build_$1.log 2>&1 || { echo "build $1 failed"; return; }
  python bench.py --workload $3 --steps 10 --warmup 3 2>/dev/null > gpurun_out/tune_$1.json.log
  python - "$1" "$2" <<'PY'
import json, sys
d = json.loads(open(f'gpurun_out/tune_{sys.argv[1]}.json.log').read())
conv = {k: v for k, v in d['stage_ms'].items() if k.startswith('conv_')}
print(sys.argv[1], sys.argv[2], round(d['value']), 'ms/step', round(d['ms_per_step'], 3), conv, flush=True)
PY
}
S0=c377d4386a8; L3=c703e622066
run s0_54 "$S0:5,4" cfg2
run s0_55 "$S0:5,5" cfg2
run s0_64 "$S0:6,4" cfg2
run s0_44 "$S0:4,4" cfg2
run s0_53 "$S0:5,3" cfg2
run l3_34 "$L3:3,4" cfg3
run l3_33 "$L3:3,3" cfg3
run l3_44 "$L3:4,4" cfg3
run l3_24 "$L3:2,4" cfg3
python -m sevennet_exp_b200.build > /dev/null 2>&1

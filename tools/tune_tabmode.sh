#!/bin/bash
# Where the radial-table rows wait for their edge (shared-memory ring vs registers), per direction, with the
# occupancy target that goes with it.  Rebuilds on the GPU box and prints the conv stage times.
mkdir -p gpurun_out
run() {  # name, SEVENN_TAB_T, SEVENN_MIN_CTAS_TAB, workload
  SEVENN_TAB_T="$2" SEVENN_MIN_CTAS_TAB="$3" python -m sevennet_exp_b200.build > gpurun_out/tt_build_$1.log 2>&1 || { echo "build $1 failed"; tail -3 gpurun_out/tt_build_$1.log; return; }
  SEVENN_TAB_T="$2" python bench.py --workload $4 --steps 5 --warmup 3 2>/dev/null > gpurun_out/tt_$1.json.log
  python - "$1" "$2" "$3" <<'PY'
import json, sys
d = json.loads(open(f'gpurun_out/tt_{sys.argv[1]}.json.log').read())
conv = {k: v for k, v in d['stage_ms'].items() if k.startswith('conv_')}
print(sys.argv[1], sys.argv[2], sys.argv[3], round(d['value']), 'ms/step', round(d['ms_per_step'], 3), conv, flush=True)
PY
}
S0=c377d4386a8; L3=c703e622066
run s0_base   smem,smem "$S0:5,5" cfg2
run s0_rf5    regs,smem "$S0:5,5" cfg2
run s0_rf6    regs,smem "$S0:6,5" cfg2
run s0_rf4    regs,smem "$S0:4,5" cfg2
run s0_rb5    smem,regs "$S0:5,5" cfg2
run s0_rb4    smem,regs "$S0:5,4" cfg2
run s0_rr64   regs,regs "$S0:6,4" cfg2
run l3_base   smem,smem "$L3:3,4" cfg3
run l3_rf3    regs,smem "$L3:3,4" cfg3
run l3_rf4    regs,smem "$L3:4,4" cfg3
run l3_rb3    smem,regs "$L3:3,3" cfg3
run l3_rb4    smem,regs "$L3:3,4" cfg3
python -m sevennet_exp_b200.build > /dev/null 2>&1

#!/bin/bash
# Everything this repo runs on the GPU box, as subcommands of ONE script (run under `gpurun -- 'bash tools/gpu.sh <cmd> ...'`).
# Every python call sits under `timeout -s KILL` (a hung kernel must not hold the box) and writes into gpurun_out/.
#
#   check                 pytest -m gpu, smoke(), the default bench line          (what the driver does at round end)
#   lines                 driver-style `python bench.py` for every workload + `--impl reference`   -> profiles/*_bench_*.json.log
#   scale N               the default bench line on N GPUs through torchrun                           -> ..._si1m_nN.json.log
#   bench [w ...]         quick lines (no CPU legs) of the given workloads (default: cfg2 cfg3 si1m cfg1)
#   ab TAG [TAG ...]      same-box A/B of prebuilt library variants tools/ab/libsevenn_b200_TAG.so (tools/build_ab.sh), twice each
#   node                  tools/bench_node.py (per-shape node GEMMs, gate kernels) + tcgen05 per-role clock traces
#   launches W [args]     ncu launch list (gpu__time_duration) of one workload, summarised per kernel
#   ncu W REGEX SKIP CNT  ncu --set full of the kernels matching REGEX on workload W -> gpurun_out/ncu_W_raw.csv
#   sweep-occupancy       __launch_bounds__ targets of the tabulated conv kernels (rebuilds on the box)
#   sweep-tabmode         where the radial-table rows wait for their edge: ring vs registers (rebuilds on the box)
mkdir -p gpurun_out
PY="timeout -s KILL 1500 python"
short() {  # one line per bench JSON: value, ms per step, end to end, first stages
  python - "$1" "$2" <<'EOF'
import json, sys
d = json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
print(sys.argv[1], round(d['value']), round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['value']),
      'launches/step', d['gpu_launches'] / d['steps'], {k: round(v, 3) for k, v in list(d.get('stage_ms', {}).items())[:8]})
EOF
}

cmd=$1; shift
case "$cmd" in
check)
  $PY -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_gpu.log
  $PY -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"
  $PY bench.py > gpurun_out/bench.json.log 2> gpurun_out/bench.err; echo "bench exit $?"; cut -c1-400 gpurun_out/bench.json.log
  ;;
lines)
  $PY bench.py > gpurun_out/final_si1m_n1.json 2> gpurun_out/final_si1m_n1.err; echo "default exit $?"
  short si1m gpurun_out/final_si1m_n1.json
  for w in cfg1 cfg2 cfg3 cfg5; do
    $PY bench.py --workload $w > gpurun_out/final_$w.json 2> gpurun_out/final_$w.err; echo "$w exit $?"
    short $w gpurun_out/final_$w.json
  done
  $PY bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/final_reference.json 2> gpurun_out/final_reference.err
  echo "reference exit $?"; tail -c 500 gpurun_out/final_reference.json
  ;;
scale)
  N=$1
  SEVENN_BENCH_NO_CPU=1 $PY -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/scale_n$N.log 2> gpurun_out/scale_n$N.err
  echo "exit $?"; tail -1 gpurun_out/scale_n$N.log | cut -c1-1500; tail -3 gpurun_out/scale_n$N.err
  ;;
bench)
  ws="$*"; [ -z "$ws" ] && ws="cfg2 cfg3 si1m cfg1"
  for w in $ws; do
    SEVENN_BENCH_NO_CPU=1 $PY bench.py --workload $w --steps 10 --warmup 3 > gpurun_out/b_$w.json 2> gpurun_out/b_$w.err \
      || { echo "bench $w FAILED"; tail -3 gpurun_out/b_$w.err; continue; }
    short $w gpurun_out/b_$w.json
  done
  ;;
ab)
  cp sevennet_exp_b200/_build/libsevenn_b200.so /tmp/keep.so
  for pass in 1 2; do for t in "$@"; do
    cp tools/ab/libsevenn_b200_$t.so sevennet_exp_b200/_build/libsevenn_b200.so
    for w in cfg2 si1m cfg3; do
      SEVENN_BENCH_NO_CPU=1 $PY bench.py --workload $w --steps 10 --warmup 3 --si-cells 30 > gpurun_out/ab_${t}_$w.json 2> gpurun_out/ab_${t}_$w.err \
        || { echo "bench $t $w FAILED"; tail -3 gpurun_out/ab_${t}_$w.err; continue; }
      short "$t $w" gpurun_out/ab_${t}_$w.json
    done
  done; done
  cp /tmp/keep.so sevennet_exp_b200/_build/libsevenn_b200.so
  ;;
node)
  timeout -s KILL 120 python tools/bench_node.py 262144 > gpurun_out/node_bench.log 2>&1; tail -30 gpurun_out/node_bench.log
  for s in "si1 l2" "si1 l1" "si1 l0" "si2T l2"; do
    echo "--- DBG $s"; SEVENN_TC_DBG=1 NODE_ONLY="$s" timeout -s KILL 60 python tools/bench_node.py 262144 2>&1 | grep -A19 TCDBG | tail -20
  done > gpurun_out/node_dbg.log 2>&1
  tail -80 gpurun_out/node_dbg.log
  ;;
launches)
  w=$1; shift
  SEVENN_BENCH_NO_CPU=1 timeout -s KILL 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv \
    --log-file gpurun_out/launches_$w.csv python bench.py --workload $w --steps 2 --warmup 1 "$@" > gpurun_out/ncu_$w.log 2>&1
  python - $w <<'EOF'
import collections, csv, sys
lines = [l for l in open(f'gpurun_out/launches_{sys.argv[1]}.csv') if not l.startswith('==')]
rows = [d for d in csv.DictReader(lines) if d.get('Metric Name') == 'gpu__time_duration.sum']
agg = collections.OrderedDict()
for d in rows:
    a = agg.setdefault(d['Kernel Name'][:70], [0, 0.0])
    a[0] += 1
    a[1] += float(d['Metric Value'].replace(',', ''))
tot = sum(v[1] for v in agg.values())
print('launches', len(rows), 'total us', tot / 1e3)
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:28]:
    print(f'{v[0]:5d} {v[1] / 1e3:10.1f} us {100 * v[1] / tot:5.1f} %  {v[1] / v[0] / 1e3:8.2f} us/launch  {k}')
EOF
  ;;
ncu)
  w=$1; re=$2; skip=${3:-0}; cnt=${4:-2}
  SEVENN_BENCH_NO_CPU=1 $PY bench.py --workload $w --steps 2 --warmup 2 > gpurun_out/ncu_plain_$w.log 2>&1 || { echo "plain run failed"; exit 1; }
  SEVENN_BENCH_NO_CPU=1 timeout -s KILL 1500 ncu --set full --clock-control none --import-source on -k regex:"$re" -s $skip -c $cnt \
    -o gpurun_out/ncu_$w -f python bench.py --workload $w --steps 1 --warmup 1 > gpurun_out/ncu_full_$w.log 2>&1; echo "ncu exit $?"
  ncu -i gpurun_out/ncu_$w.ncu-rep --page raw --csv > gpurun_out/ncu_${w}_raw.csv 2>/dev/null
  rm -f gpurun_out/ncu_$w.ncu-rep; ls -la gpurun_out/ncu_${w}_raw.csv
  ;;
sweep-occupancy|sweep-tabmode)
  run() {  # name, SEVENN_TAB_T ('' = default), SEVENN_MIN_CTAS_TAB, workload
    env ${2:+SEVENN_TAB_T=$2} SEVENN_MIN_CTAS_TAB="$3" python -m sevennet_exp_b200.build > gpurun_out/sweep_build_$1.log 2>&1 \
      || { echo "build $1 failed"; tail -3 gpurun_out/sweep_build_$1.log; return; }
    env ${2:+SEVENN_TAB_T=$2} SEVENN_BENCH_NO_CPU=1 $PY bench.py --workload $4 --steps 5 --warmup 3 2> /dev/null > gpurun_out/sweep_$1.json
    python - "$1" "$2" "$3" <<'EOF'
import json, sys
d = json.loads(open(f'gpurun_out/sweep_{sys.argv[1]}.json').read().strip().splitlines()[-1])
print(*sys.argv[1:], round(d['value']), 'ms/step', round(d['ms_per_step'], 3), {k: v for k, v in d['stage_ms'].items() if k.startswith('conv_')}, flush=True)
EOF
  }
  S0=c377d4386a8; L3=c703e622066
  if [ "$cmd" = sweep-occupancy ]; then
    for v in 3,2 3,3 4,3 4,2 3,4; do run l3_$v "" "$L3:$v" cfg3; done
    for v in 6,5 6,6 8,5 6,4 5,5 5,4; do run s0_$v "" "$S0:$v" cfg2; done
  else
    for v in smem,smem regs,smem smem,regs regs,regs; do run s0_$v $v "$S0:5,4" cfg2; run l3_$v $v "$L3:3,4" cfg3; done
  fi
  python -m sevennet_exp_b200.build > /dev/null 2>&1      # restore the defaults
  ;;
*)
  sed -n 2,16p "$0"; exit 2
  ;;
esac

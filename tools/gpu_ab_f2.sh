#!/bin/bash
# A/B on one box: scalar vs packed-fp32 (SEVENN_TAB_F2 build) two-chunk conv units; the engine reads the table layout
# from the library (sevenn_conv_tab_packed_<sig>), so no environment is needed at run time
mkdir -p gpurun_out
run() {  # tag workload steps
  cp tools/ab/libsevenn_b200_$1.so sevennet_exp_b200/_lib/libsevenn_b200.so
  SEVENN_BENCH_NO_CPU=1 python bench.py --workload $2 --steps $3 --warmup 3 > gpurun_out/f2ab_$1_$2.json 2> gpurun_out/f2ab_$1_$2.err; echo "bench $1 $2 exit $?"
  python - $1 $2 <<'PY'
import json, sys
d = json.loads(open(f'gpurun_out/f2ab_{sys.argv[1]}_{sys.argv[2]}.json').read())
print(sys.argv[1], sys.argv[2], round(d['value']), round(d['ms_per_step'], 3), {k: v for k, v in list(d['stage_ms'].items())[:8] if k.startswith('conv')})
PY
}
run ref cfg2 20; run f2 cfg2 20; run ref cfg2 20; run f2 cfg2 20
run ref cfg3 8; run f2 cfg3 8
python -m pytest tests -m gpu -x -q 2>&1 | tail -2      # with the packed library in place
cp tools/ab/libsevenn_b200_ref.so sevennet_exp_b200/_lib/libsevenn_b200.so

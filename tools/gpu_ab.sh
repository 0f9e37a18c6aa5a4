#!/bin/bash
# A/B of two prebuilt libraries on the same box: tools/ab/libsevenn_b200_<tag>.so, env per tag
mkdir -p gpurun_out
run() {  # tag, env assignment
  cp tools/ab/libsevenn_b200_$1.so sevennet_exp_b200/_lib/libsevenn_b200.so
  for w in cfg2 cfg3; do
    env $2 SEVENN_BENCH_NO_CPU=1 python bench.py --workload $w --steps 20 --warmup 3 > gpurun_out/ab_$1_$w.json 2> gpurun_out/ab_$1_$w.err; echo "bench $1 $w exit $?"
    python - $1 $w <<'PY'
import json, sys
d = json.loads(open(f'gpurun_out/ab_{sys.argv[1]}_{sys.argv[2]}.json').read())
print(sys.argv[1], sys.argv[2], round(d['value']), round(d['ms_per_step'], 3), 'e2e', round(d['e2e']['value']), {k: v for k, v in list(d['stage_ms'].items())[:6]})
PY
  done
}
run cpw1 SEVENN_TAB_CPW=1
run cpw2 SEVENN_TAB_CPW=2
run cpw1 SEVENN_TAB_CPW=1
run cpw2 SEVENN_TAB_CPW=2
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log

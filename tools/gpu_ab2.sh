#!/bin/bash
# A/B of prebuilt library variants on one box (tools/build_ab.sh <tag> ENV=...): usage tools/gpu_ab2.sh tag1 tag2 ...
mkdir -p gpurun_out
cp sevennet_exp_b200/_build/libsevenn_b200.so /tmp/keep.so
run() {
  cp tools/ab/libsevenn_b200_$1.so sevennet_exp_b200/_build/libsevenn_b200.so
  for w in cfg2 si1m cfg3; do
    SEVENN_BENCH_NO_CPU=1 python bench.py --workload $w --steps 10 --warmup 3 --si-cells 30 > gpurun_out/ab_$1_$w.json 2> gpurun_out/ab_$1_$w.err || { echo "bench $1 $w FAILED"; tail -3 gpurun_out/ab_$1_$w.err; continue; }
    python - $1 $w <<'PY'
import json, sys
d = json.loads(open(f'gpurun_out/ab_{sys.argv[1]}_{sys.argv[2]}.json').read().strip().splitlines()[-1])
print(sys.argv[1], sys.argv[2], round(d['value']), round(d['ms_per_step'], 3), {k: v for k, v in list(d['stage_ms'].items())[:5]})
PY
  done
}
for t in "$@"; do run $t; done
for t in "$@"; do run $t; done
cp /tmp/keep.so sevennet_exp_b200/_build/libsevenn_b200.so

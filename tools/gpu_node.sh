#!/bin/bash
# node-stage microbenchmarks on one box: tools/bench_node.py, plan variants, tcgen05 pipeline traces
mkdir -p gpurun_out
timeout -s KILL 120 python tools/bench_node.py 262144 > gpurun_out/node_bench.log 2>&1; tail -30 gpurun_out/node_bench.log
timeout -s KILL 400 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "gemm or tc or fixture" 2>&1 | tail -3
for t in 2 3; do echo "--- SEVENN_TC_TILES=$t"; SEVENN_TC_TILES=$t NODE_ONLY="l0" timeout -s KILL 120 python tools/bench_node.py 262144 2>&1 | grep " l0"; done
for s in "si1 l2" "si1 l1"; do
  echo "--- DBG $s"; SEVENN_TC_DBG=1 NODE_ONLY="$s" timeout -s KILL 60 python tools/bench_node.py 262144 2>&1 | grep -A19 TCDBG | tail -20
done > gpurun_out/node_dbg.log 2>&1
python - <<'PY'
import re
for line in open('gpurun_out/node_dbg.log'):
    nums = re.findall(r'-?\d+', line)
    if line.startswith('  ') and len(nums) >= 8:
        vals = [int(x) for x in nums[-8:]]
        base = 188000000000000
        print(line.split(':')[0] + ':', ' '.join(f'{(v % 1000000):7d}' if v > 0 else '     -1' for v in vals))
    else:
        print(line.rstrip())
PY

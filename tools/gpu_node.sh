#!/bin/bash
# node-stage microbenchmarks on one box: tools/bench_node.py, tcgen05 pipeline traces of the skinny shapes
mkdir -p gpurun_out
timeout -s KILL 120 python tools/bench_node.py 262144 > gpurun_out/node_bench.log 2>&1; tail -30 gpurun_out/node_bench.log
timeout -s KILL 400 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "gemm or tc or fixture" 2>&1 | tail -5
for s in "si1 l2" "si1 l0" "si2T l2"; do
  echo "--- DBG $s"; SEVENN_TC_DBG=1 NODE_ONLY="$s" timeout -s KILL 60 python tools/bench_node.py 262144 2>&1 | grep -A19 TCDBG | tail -20
done > gpurun_out/node_dbg.log 2>&1
tail -85 gpurun_out/node_dbg.log

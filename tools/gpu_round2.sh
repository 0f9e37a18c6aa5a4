#!/bin/bash
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/smoke.log
python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json.log 2> gpurun_out/bench.err; echo "bench exit $?"
python bench.py --workload cfg3 --steps 10 --warmup 3 > gpurun_out/bench_cfg3.json.log 2> gpurun_out/bench_cfg3.err; echo "bench cfg3 exit $?"
SEVENN_RADIAL=mlp python bench.py --workload cfg3 --steps 10 --warmup 3 > gpurun_out/bench_cfg3_mlp.json.log 2> gpurun_out/bench_cfg3_mlp.err; echo "bench cfg3 mlp exit $?"
cut -c1-400 gpurun_out/bench.json.log gpurun_out/bench_cfg3.json.log gpurun_out/bench_cfg3_mlp.json.log

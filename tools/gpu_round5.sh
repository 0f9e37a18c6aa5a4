#!/bin/bash
# r01h evidence: GPU tests, smoke, cfg2 + cfg3 bench lines (both arms for cfg2), launch list, one ncu --set full capture
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"
python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json.log 2> gpurun_out/bench.err; echo "bench exit $?"
python bench.py --workload cfg3 --steps 10 --warmup 3 > gpurun_out/bench_cfg3.json.log 2> gpurun_out/bench_cfg3.err; echo "bench cfg3 exit $?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json.log 2> gpurun_out/bench_reference.err; echo "bench reference exit $?"
cut -c1-330 gpurun_out/bench.json.log gpurun_out/bench_cfg3.json.log gpurun_out/bench_reference.json.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 400 -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'conv_(fwd|bwd)_tab_c377d4386a8' -s 8 -c 2 -o gpurun_out/conv_tab -f python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_full.log 2>&1; echo "ncu full exit $?"

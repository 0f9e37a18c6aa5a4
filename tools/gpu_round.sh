#!/bin/bash
# One GPU-box pass: parity tests, bench (both radial modes), ncu launch list, ncu --set full of the conv kernels.
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json.log 2> gpurun_out/bench.err; echo "bench exit $?"
SEVENN_RADIAL=mlp python bench.py --steps 20 --warmup 3 > gpurun_out/bench_mlp.json.log 2> gpurun_out/bench_mlp.err; echo "bench mlp exit $?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json.log 2>&1; echo "ref exit $?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 400 -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'conv_(fwd|bwd)_tab_c377d4386a8' -s 8 -c 2 -o gpurun_out/conv_tab -f python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_full.log 2>&1; echo "ncu full exit $?"

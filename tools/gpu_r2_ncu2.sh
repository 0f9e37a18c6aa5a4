#!/bin/bash
mkdir -p gpurun_out
export SEVENN_BENCH_NO_CPU=1
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log; tail -3 gpurun_out/pytest_gpu.log
for w in si1m cfg2; do
  python bench.py --workload $w --steps 5 --warmup 3 > gpurun_out/ncu_plain_$w.log 2>&1 || { echo "plain $w failed"; continue; }
  python -c "import json;d=json.loads(open('gpurun_out/ncu_plain_$w.log').read().strip().splitlines()[-1]);print('$w',d['value'],d['ms_per_step'])"
  for k in fwd bwd; do
    timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"conv_${k}_tab_c377d4386a8" -s 3 -c 1 \
      -o gpurun_out/r02_conv_${k}_tab_$w -f python bench.py --workload $w --steps 1 --warmup 3 > gpurun_out/ncu_full_${k}_$w.log 2>&1; echo "ncu $k $w exit $?"
    ncu -i gpurun_out/r02_conv_${k}_tab_$w.ncu-rep --page raw --csv > gpurun_out/r02_conv_${k}_tab_${w}_raw.csv 2>/dev/null
  done
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r02_launches_$w.csv \
    python bench.py --workload $w --steps 1 --warmup 3 > gpurun_out/ncu_launch_$w.log 2>&1; echo "launch list $w exit $?"
done

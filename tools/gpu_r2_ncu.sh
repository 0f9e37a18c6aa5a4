#!/bin/bash
# r02 evidence: launch lists + ncu --set full captures of the dominant kernels on the bench workloads
mkdir -p gpurun_out
export SEVENN_BENCH_NO_CPU=1
for w in si1m cfg2; do
  python bench.py --workload $w --steps 3 --warmup 3 > gpurun_out/ncu_plain_$w.log 2>&1 || { echo "plain $w failed"; continue; }
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/r02_launches_$w.csv \
    python bench.py --workload $w --steps 1 --warmup 3 > gpurun_out/ncu_launch_$w.log 2>&1; echo "launch list $w exit $?"
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'conv_(fwd|bwd)_tab_c377d4386a8' -s 6 -c 2 \
    -o gpurun_out/r02_conv_tab_$w -f python bench.py --workload $w --steps 1 --warmup 3 > gpurun_out/ncu_full_$w.log 2>&1; echo "ncu full $w exit $?"
  ncu -i gpurun_out/r02_conv_tab_$w.ncu-rep --page raw --csv > gpurun_out/r02_conv_tab_${w}_raw.csv 2>/dev/null
done
timeout 900 ncu --set full --clock-control none -k regex:'gemm_tc_kernel' -s 20 -c 3 -o gpurun_out/r02_gemm_tc_cfg2 -f \
  python bench.py --workload cfg2 --steps 1 --warmup 3 > gpurun_out/ncu_full_gemm.log 2>&1; echo "ncu gemm exit $?"
ncu -i gpurun_out/r02_gemm_tc_cfg2.ncu-rep --page raw --csv > gpurun_out/r02_gemm_tc_cfg2_raw.csv 2>/dev/null
ls -la gpurun_out | head -30

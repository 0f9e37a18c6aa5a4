#!/bin/bash
# r02 (late) evidence: launch lists of the final build + ncu --set full of the node-stage kernels on the 1M-atom workload
mkdir -p gpurun_out
export SEVENN_BENCH_NO_CPU=1
for w in si1m cfg2; do
  timeout -s KILL 900 python bench.py --workload $w --steps 3 --warmup 3 > gpurun_out/ncu_plain_$w.log 2>&1 || { echo "plain $w failed"; continue; }
  timeout -s KILL 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/r02b_launches_$w.csv \
    python bench.py --workload $w --steps 1 --warmup 3 > gpurun_out/ncu_launch_$w.log 2>&1; echo "launch list $w exit $?"
done
timeout -s KILL 1200 ncu --set full --clock-control none --import-source on -k regex:'gemm_tc_kernel|gemm_tca_kernel' -s 30 -c 12 \
  -o gpurun_out/r02b_gemm_si1m -f python bench.py --workload si1m --steps 1 --warmup 1 > gpurun_out/ncu_full_gemm.log 2>&1; echo "ncu gemm exit $?"
ncu -i gpurun_out/r02b_gemm_si1m.ncu-rep --page raw --csv > gpurun_out/r02b_gemm_tc_si1m_raw.csv 2>/dev/null
timeout -s KILL 1200 ncu --set full --clock-control none --import-source on -k regex:'gate_(fwd|bwd)_v4_kernel' -s 4 -c 2 \
  -o gpurun_out/r02b_gate_si1m -f python bench.py --workload si1m --steps 1 --warmup 1 > gpurun_out/ncu_full_gate.log 2>&1; echo "ncu gate exit $?"
ncu -i gpurun_out/r02b_gate_si1m.ncu-rep --page raw --csv > gpurun_out/r02b_gate_v4_si1m_raw.csv 2>/dev/null
rm -f gpurun_out/*.ncu-rep
ls -la gpurun_out | head -40

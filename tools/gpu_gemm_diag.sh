#!/bin/bash
mkdir -p gpurun_out
python tools/bench_gemm.py tc cfg3 > gpurun_out/gemm_cfg3.log 2>&1; echo "exit $?"; cat gpurun_out/gemm_cfg3.log
timeout 300 ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 2 -c 1 -o gpurun_out/gemm_k32 -f python tools/bench_gemm.py tc one > gpurun_out/ncu_gemm.log 2>&1; echo "ncu exit $?"

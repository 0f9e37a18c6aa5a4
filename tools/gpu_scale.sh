#!/bin/bash
# strong-scaling line on N GPUs of one box (default workload: the 1M-atom Si cell): tools/gpu_scale.sh N
mkdir -p gpurun_out
N=$1
SEVENN_BENCH_NO_CPU=1 timeout -s KILL 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/scale_n$N.log 2> gpurun_out/scale_n$N.err
echo "exit $?"; tail -1 gpurun_out/scale_n$N.log | cut -c1-1500; tail -3 gpurun_out/scale_n$N.err

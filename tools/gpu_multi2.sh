#!/bin/bash
# BASELINE config 4 strong scaling: the 1M-atom Si cell on 2, 4 and 8 GPUs of one box.
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for n in 8 4 2; do
  timeout 900 $TR --nproc-per-node $n --master-port $((29600+n)) bench.py --gpus $n --workload si1m --steps 4 --warmup 3 > gpurun_out/bench_si1m_n$n.log 2> gpurun_out/bench_si1m_n$n.err; echo "si1m n=$n exit $?"
  grep -h '"metric"' gpurun_out/bench_si1m_n$n.log | cut -c1-330
  tail -3 gpurun_out/bench_si1m_n$n.err | cut -c1-300
done

#!/bin/bash
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for n in 2 4 8; do
  SEVENN_CHECK=1 timeout 600 $TR --nproc-per-node $n --master-port $((29500+n)) bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/bench_n$n.log 2> gpurun_out/bench_n$n.err; echo "n=$n exit $?"
  grep -h "check rank 0\|metric" gpurun_out/bench_n$n.log | cut -c1-260
done
for n in 8 4 2; do
  timeout 900 $TR --nproc-per-node $n --master-port $((29600+n)) bench.py --gpus $n --workload si1m --steps 4 --warmup 3 > gpurun_out/bench_si1m_n$n.log 2> gpurun_out/bench_si1m_n$n.err; echo "si1m n=$n exit $?"
  grep -h '"metric"' gpurun_out/bench_si1m_n$n.log | cut -c1-260
done
timeout 900 python bench.py --workload si1m --steps 3 --warmup 3 > gpurun_out/bench_si1m_n1.log 2> gpurun_out/bench_si1m_n1.err; echo "si1m n=1 exit $?"
grep -h '"metric"' gpurun_out/bench_si1m_n1.log | cut -c1-260

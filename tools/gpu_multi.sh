#!/bin/bash
# One 8-GPU box pass: weak scaling cfg2 at 2/4/8 GPUs (with the decomposed == single-domain check), BASELINE config 4 (si1m) at 8.
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for n in 2 4 8; do
  SEVENN_CHECK=1 timeout 600 $TR --nproc-per-node $n --master-port $((29500+n)) bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/bench_n$n.log 2> gpurun_out/bench_n$n.err; echo "n=$n exit $?"
  grep -h "check rank\|metric" gpurun_out/bench_n$n.log | cut -c1-300
done
timeout 900 $TR --nproc-per-node 8 --master-port 29611 bench.py --gpus 8 --workload si1m --steps 5 --warmup 3 > gpurun_out/bench_si1m_n8.log 2> gpurun_out/bench_si1m_n8.err; echo "si1m exit $?"
grep -h metric gpurun_out/bench_si1m_n8.log | cut -c1-400

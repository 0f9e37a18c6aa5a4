#!/bin/bash
# NCCL all-to-all halo vs peer-memory halo on the same box, weak-scaling workload, with the decomposed == single check
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
N=$(nvidia-smi -L | wc -l)
port=29511
for mode in nccl p2p; do
  SEVENN_HALO=$mode SEVENN_CHECK=1 timeout 300 $TR --nproc-per-node $N --master-port $port bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/halo_${mode}_n$N.log 2> gpurun_out/halo_${mode}_n$N.err; echo "$mode n=$N exit $?"
  grep -h "check rank 0\|halo rank" gpurun_out/halo_${mode}_n$N.log | cut -c1-200
  python - $mode $N <<'PY'
import json, sys
for ln in open(f'gpurun_out/halo_{sys.argv[1]}_n{sys.argv[2]}.log'):
    if ln.startswith('{'):
        d = json.loads(ln); print(sys.argv[1], round(d['value']), round(d['ms_per_step'], 3), 'e2e', round(d['e2e']['value']))
PY
  tail -2 gpurun_out/halo_${mode}_n$N.err | cut -c1-300
  port=$((port+1))
done

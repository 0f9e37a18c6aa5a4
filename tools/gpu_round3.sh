#!/bin/bash
set -x
mkdir -p gpurun_out
python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json.log 2> gpurun_out/bench.err; echo "bench exit $?"
python - <<'PY'
import json
d = json.loads(open('gpurun_out/bench.json.log').read())
print(round(d['value']), d['ms_per_step'], 'e2e', round(d['e2e']['value']), d['clocks'], {k: v for k, v in list(d['stage_ms'].items())[:6]})
PY
SEVENN_CHECK=1 timeout 600 python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 2 --master-port 29502 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_n2.log 2> gpurun_out/bench_n2.err; echo "n=2 exit $?"
grep -h "check rank\|metric" gpurun_out/bench_n2.log | cut -c1-250; tail -3 gpurun_out/bench_n2.err
timeout 900 python bench.py --workload si1m --steps 3 --warmup 3 > gpurun_out/bench_si1m_n1.log 2> gpurun_out/bench_si1m_n1.err; echo "si1m n=1 exit $?"
grep -h metric gpurun_out/bench_si1m_n1.log | cut -c1-300; tail -3 gpurun_out/bench_si1m_n1.err | cut -c1-300
nvidia-smi --query-gpu=memory.used --format=csv | head -3

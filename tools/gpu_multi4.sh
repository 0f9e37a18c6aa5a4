#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python bench.py --steps 20 --warmup 3 > gpurun_out/bench_cfg2.json.log 2> gpurun_out/bench_cfg2.err; echo "bench exit $?"
python - <<'PY'
import json
d = json.loads(open('gpurun_out/bench_cfg2.json.log').read())
print(round(d['value']), round(d['ms_per_step'], 3), 'e2e', round(d['e2e']['value']), {k: v for k, v in list(d['stage_ms'].items())[:6]})
PY
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
N=$(nvidia-smi -L | wc -l)
SEVENN_CHECK=1 timeout 600 $TR --nproc-per-node $N --master-port 29507 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_nN.log 2> gpurun_out/bench_nN.err; echo "n=$N exit $?"
grep -h "check rank 0\|metric" gpurun_out/bench_nN.log | cut -c1-260; tail -3 gpurun_out/bench_nN.err | cut -c1-200

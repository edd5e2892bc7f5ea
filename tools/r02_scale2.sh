#!/bin/bash
# 2 GPUs: the driver's scaling command (data parallel + the tensor-parallel extra record) and the multi-GPU tests
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
nvidia-smi -L | head -4
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/scale2_bench.json 2> gpurun_out/scale2_bench.err ) 2> gpurun_out/scale2.time
tail -3 gpurun_out/scale2.time; tail -5 gpurun_out/scale2_bench.err | cut -c1-300
python - <<'PY'
import json
for l in open('gpurun_out/scale2_bench.json'):
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l)
        print('value', d['value'], 'ms', d['ms_per_step'], 'n_gpus', d['n_gpus'], 'e2e', d['e2e']['value'] if d.get('e2e') else None)
        print('extra', d.get('extra'))
PY
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 2>&1 | tail -2 | cut -c1-400

#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
for args in "--batch 256" "--batch 64" "--batch 8" "--model 7b --quant q8 --batch 1 --seq-len 512" "--model 7b --quant q8 --batch 64 --seq-len 512"; do
timeout 300 python bench.py --no-cpu-baseline --no-e2e --no-extra --no-shares --steps 20 --warmup 5 $args 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$args', round(d['value'],1), 'tok/s', round(d['ms_per_step'],4), 'ms', 'frac', round(d['roofline']['frac'],3))"
done
python tools/step_timeline.py --model 7b --quant q8 --seq-len 512 --batch 1 --pos 10 --show 16 > gpurun_out/timeline_7bq8_1.txt 2>&1
python tools/step_timeline.py --batch 256 --pos 10 --show 16 > gpurun_out/timeline_256.txt 2>&1
timeout 1500 python -m pytest tests/test_gpu_ops.py tests/test_gpu_baseline_shapes.py tests/test_gpu_forward.py tests/test_gpu_sampler.py -m gpu -q -x 2>&1 | tail -3

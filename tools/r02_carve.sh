#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
for C in 1 0; do
for args in "--batch 256" "--model 7b --quant q8 --batch 1 --seq-len 512"; do
THA_CARVEOUT=$C timeout 300 python bench.py --no-cpu-baseline --no-e2e --no-extra --no-shares --steps 20 --warmup 5 $args 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('CARVEOUT=$C $args', round(d['value'],1), 'tok/s', round(d['ms_per_step'],4), 'ms', 'frac', round(d['roofline']['frac'],3))"
done; done
THA_CARVEOUT=1 python tools/step_timeline.py --model 7b --quant q8 --seq-len 512 --batch 1 --pos 10 --show 16 > gpurun_out/timeline_7bq8_1_carve.txt 2>&1
THA_CARVEOUT=1 python tools/step_timeline.py --batch 256 --pos 10 --show 16 > gpurun_out/timeline_256_carve.txt 2>&1

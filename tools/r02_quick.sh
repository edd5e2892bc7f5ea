#!/bin/bash
# quick loop: matmul + forward parity, the step timeline, one bench line without the side records
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ops.py tests/test_gpu_baseline_shapes.py tests/test_gpu_forward.py -m gpu -q -x 2>&1 | tail -3
python tools/step_timeline.py --batch 256 --pos 10 --show 12 > gpurun_out/timeline_256.txt 2>&1
timeout 300 python bench.py --no-cpu-baseline --no-e2e --no-extra --steps 20 --warmup 5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value'],1), 'tok/s', round(d['ms_per_step'],4), 'ms', 'frac', round(d['roofline']['frac'],3)); print(d['roofline'].get('live_shares_us'))"
THA_DUAL=16 timeout 300 python bench.py --no-cpu-baseline --no-e2e --no-extra --no-shares --steps 20 --warmup 5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('dual', round(d['value'],1), 'tok/s', round(d['ms_per_step'],4), 'ms')"

#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
THA_SAMPLE_RESIDENT=0 python tools/sampler_time.py
python tools/sampler_time.py
python tools/sampler_time.py --batch 64
timeout 900 python -m pytest tests/test_gpu_sampler.py tests/test_host_cpp.py -m gpu -q -x 2>&1 | tail -3
timeout 300 python bench.py --no-cpu-baseline --no-extra --no-shares --steps 20 --warmup 5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'greedy', round(d['e2e']['greedy_call']['value'],1))"

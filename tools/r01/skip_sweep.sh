#!/bin/bash
# live share of each kernel class: step time with the class left out (THA_SKIP), graph + PDL on
cd "$GRAFT_REPO_ROOT" || exit 1
for S in 0 1 2 4 8 16 32 64 127; do
  THA_SKIP=$S timeout 300 python bench.py --no-cpu-baseline --no-e2e "$@" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('SKIP=$S', round(d['ms_per_step']*1000,1), 'us/step', d['gpu_launches']//d['steps'], 'launches')"
done

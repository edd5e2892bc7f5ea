#!/bin/bash
# split-K override sweep (THA_TC_Z="qkv,wo,w13,w2"): tools/z_sweep.sh "<bench args>" combo...
cd "$GRAFT_REPO_ROOT" || exit 1
ARGS="$1"; shift
for Z in "$@"; do
  THA_TC_Z=$Z timeout 300 python bench.py --no-cpu-baseline --no-e2e $ARGS 2>/dev/null | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('Z=$Z', round(d['value'],1), 'tok/s', round(d['ms_per_step']*1000,1), 'us')
except Exception as e: print('Z=$Z failed')"
done

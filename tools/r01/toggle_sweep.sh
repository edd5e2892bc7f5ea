#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
run() {
  local E="$1"; shift
  env $E timeout 300 python bench.py --no-cpu-baseline --no-e2e "$@" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$E', '$*', round(d['value'],1), 'tok/s', round(d['ms_per_step'],4), 'ms')"
}
for E in "THA_PDL=1 THA_GRAPH=1" "THA_PDL=0 THA_GRAPH=1" "THA_PDL=1 THA_GRAPH=0" "THA_PDL=0 THA_GRAPH=0"; do
  run "$E" --batch 256 --steps 48 --warmup 4
  run "$E" --model 7b --quant q8 --batch 1 --seq-len 512 --steps 48 --warmup 4
done

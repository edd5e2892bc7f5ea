#!/bin/bash
# THA_ATTN_J sweep of the fused decode attention (segment width = 128*J floats per CTA)
cd "$GRAFT_REPO_ROOT" || exit 1
run() { # label, env J, args
  local J=$1; shift
  THA_ATTN_J=$J timeout 300 python bench.py --no-cpu-baseline --no-e2e "$@" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('J=$J', '$*', round(d['value'],1), 'tok/s', round(d['ms_per_step'],4), 'ms', 'frac', round(d['roofline']['frac'],3))"
}
for J in 0 6 3 2 1; do run $J --batch 256 --pos0 128 --steps 32 --warmup 4; done
for J in 0 6 3 2 1; do run $J --batch 64 --pos0 128 --steps 32 --warmup 4; done
for J in 0 6 3 2 1; do run $J --batch 256 --pos0 512 --steps 32 --warmup 4; done
for J in 0 8 4 2 1; do run $J --model 7b --quant q8 --batch 64 --seq-len 512 --pos0 256 --steps 16 --warmup 3; done

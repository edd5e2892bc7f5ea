#!/bin/bash
# THA_DUAL sweep: rows from which a step runs as two concurrent half-batch branches (0 = never)
cd "$GRAFT_REPO_ROOT" || exit 1
run() {
  local D=$1; shift
  THA_DUAL=$D timeout 300 python bench.py --no-cpu-baseline --no-e2e "$@" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('DUAL=$D', '$*', round(d['value'],1), 'tok/s', round(d['ms_per_step'],4), 'ms', 'frac', round(d['roofline']['frac'],3))"
}
for D in 0 16; do
  run $D --batch 256 --steps 256 --warmup 8
  run $D --batch 256 --steps 48 --warmup 4
  run $D --batch 128 --steps 48 --warmup 4
  run $D --batch 64 --steps 48 --warmup 4
  run $D --batch 32 --steps 48 --warmup 4
  run $D --batch 16 --steps 48 --warmup 4
  run $D --model 7b --quant q8 --batch 64 --seq-len 512 --steps 48 --warmup 4
  run $D --model 7b --quant q8 --batch 16 --seq-len 512 --steps 48 --warmup 4
  run $D --model 7b --quant f32 --batch 16 --seq-len 512 --steps 32 --warmup 4
  run $D --model s110m --quant q8 --batch 256 --steps 48 --warmup 4
done

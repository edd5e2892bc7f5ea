#!/bin/bash
# THA_DUAL (two concurrent half-batch branches per step) with each branch's matmuls limited to THA_DUAL_SMS SMs
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
run() {
  local D=$1 S=$2; shift; shift
  THA_DUAL=$D THA_DUAL_SMS=$S timeout 300 python bench.py --no-cpu-baseline --no-e2e --no-extra --no-shares "$@" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('DUAL=$D SMS=$S', '$*', round(d['value'],1), 'tok/s', round(d['ms_per_step'],4), 'ms', 'frac', round(d['roofline']['frac'],3), 'parity', d.get('parity_sample',{}).get('max_abs_err'))"
}
run 0 0 --batch 256 --steps 20 --warmup 5
for S in 74 60 90 110 148; do run 16 $S --batch 256 --steps 20 --warmup 5; done
run 0 0 --batch 128 --steps 20 --warmup 5
run 16 74 --batch 128 --steps 20 --warmup 5
run 0 0 --batch 64 --steps 20 --warmup 5
run 16 74 --batch 64 --steps 20 --warmup 5
run 0 0 --model s110m --quant q8 --batch 256 --steps 20 --warmup 5
run 16 74 --model s110m --quant q8 --batch 256 --steps 20 --warmup 5
run 0 0 --model 7b --quant q8 --batch 64 --seq-len 512 --steps 20 --warmup 5
run 16 74 --model 7b --quant q8 --batch 64 --seq-len 512 --steps 20 --warmup 5
THA_DUAL=16 timeout 900 python -m pytest tests/test_gpu_baseline_shapes.py tests/test_gpu_forward.py -m gpu -q -x 2>&1 | tail -3

#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
python -m pytest tests/test_gpu_forward.py tests/test_gpu_ops.py -x -q -m gpu -k "q8 or prefill or tensor_core" 2>&1 | tail -3
run() {
  python bench.py --model $1 --quant $2 --batch $3 --steps 48 --warmup 4 --no-cpu-baseline --no-e2e 2>gpurun_out/r1f_$1_$2_b$3.err | python -c "
import sys,json
for l in sys.stdin:
    try: j=json.loads(l)
    except Exception: continue
    print('$1 $2 B=$3 tc=$THA_TC', round(j['value'],1),'tok/s', round(j['ms_per_step'],3),'ms frac',round(j['roofline']['frac'],3),'launches/step',j['gpu_launches']//j['steps'])
"
}
for cfgs in "7b q8 1" "7b q8 8" "s110m q8 1" "s110m q8 8"; do
  set -- $cfgs
  THA_TC=1 run $1 $2 $3
done
for cfgs in "7b q8 16" "7b q8 32" "7b q8 64" "s110m q8 64" "s110m q8 256"; do
  set -- $cfgs
  THA_TC=-1 run $1 $2 $3
done

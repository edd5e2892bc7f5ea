#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
run() {
  python bench.py --model $1 --quant $2 --batch $3 --steps 48 --warmup 4 --no-cpu-baseline --no-e2e 2>gpurun_out/r1f_$1_$2_b$3.err | python -c "
import sys,json
for l in sys.stdin:
    try: j=json.loads(l)
    except Exception: continue
    print('$1 $2 B=$3 tc=$THA_OPT_TC', round(j['value'],1),'tok/s', round(j['ms_per_step'],3),'ms frac',round(j['roofline']['frac'],3),'launches/step',j['gpu_launches']//j['steps'])
"
}
for cfgs in "7b q8 1" "7b q8 8" "s110m q8 1" "s110m q8 8"; do
  set -- $cfgs
  THA_TC=-1 THA_OPT_TC=-1 run $1 $2 $3
  THA_TC=1 THA_OPT_TC=1 run $1 $2 $3
done
for cfgs in "7b q8 16" "7b q8 32" "7b q8 64" "7b q8 256" "s110m q8 64" "s110m q8 256"; do
  set -- $cfgs
  THA_TC=-1 THA_OPT_TC=-1 run $1 $2 $3
done

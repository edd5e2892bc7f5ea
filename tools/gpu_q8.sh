#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
for cfgs in "7b q8 16" "7b q8 64" "7b q8 8" "s110m q8 64" "s110m q8 256" "s110m q8 8"; do
  set -- $cfgs
  python bench.py --model $1 --quant $2 --batch $3 --steps 48 --warmup 4 --no-cpu-baseline --no-e2e 2>gpurun_out/r1f_$1_$2_b$3.err | python -c "
import sys,json
for l in sys.stdin:
    try: j=json.loads(l)
    except Exception: continue
    print('$1 $2 B=$3', round(j['value'],1),'tok/s', round(j['ms_per_step'],3),'ms frac',round(j['roofline']['frac'],3),'launches/step',j['gpu_launches']//j['steps'])
"
done

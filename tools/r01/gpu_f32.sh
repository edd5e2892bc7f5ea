#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
run() {
  python bench.py --model $1 --quant $2 --batch $3 --steps 48 --warmup 4 --no-cpu-baseline --no-e2e 2>gpurun_out/r1j_$1_$2_b$3.err | python -c "
import sys,json
for l in sys.stdin:
    try: j=json.loads(l)
    except Exception: continue
    print('$1 $2 B=$3', round(j['value'],1),'tok/s', round(j['ms_per_step'],3),'ms frac',round(j['roofline']['frac'],3),'launches/step',j['gpu_launches']//j['steps'])
"
}
for cfgs in "s110m f32 16" "s110m f32 64" "s110m f32 256" "7b f32 16" "7b f32 64"; do
  set -- $cfgs
  run $1 $2 $3
done

#!/bin/bash
# chain kernel: sensitivity of the step time to the split-K factors (wo,w13,w2,qkv,cls) and to attention
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
for z in "1,1,1,1,1" "2,2,2,2,1" "4,2,4,4,1" "8,2,8,4,1" "0,0,0,0,0"; do
  THA_CHAIN_Z=$z timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/c_bench.json 2> gpurun_out/c_bench.err
  python -c "import json;d=json.load(open('gpurun_out/c_bench.json'));print('z=$z', round(d['value']), d['ms_per_step'], d['gpu_launches'])" | tee -a gpurun_out/c_sweep.txt
done
THA_SKIP=2 timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/c_bench.json 2> gpurun_out/c_bench.err
python -c "import json;d=json.load(open('gpurun_out/c_bench.json'));print('no attention', round(d['value']), d['ms_per_step'], d['gpu_launches'])" | tee -a gpurun_out/c_sweep.txt
THA_SKIP=1 timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/c_bench.json 2> gpurun_out/c_bench.err
python -c "import json;d=json.load(open('gpurun_out/c_bench.json'));print('no chains', round(d['value']), d['ms_per_step'], d['gpu_launches'])" | tee -a gpurun_out/c_sweep.txt

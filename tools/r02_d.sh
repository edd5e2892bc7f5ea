#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_forward.py -m gpu -x -q -k "fp32 or modes or greedy" > gpurun_out/d_tests1.log 2>&1; echo "rc=$?" >> gpurun_out/d_tests1.log
tail -3 gpurun_out/d_tests1.log
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/d_bench.json 2> gpurun_out/d_bench.err
python -c "import json;d=json.load(open('gpurun_out/d_bench.json'));print('default', round(d['value']), d['ms_per_step'], d['gpu_launches'])"
THA_GRAPH=0 timeout 300 python tools/chain_trace.py --launch 5 > gpurun_out/d_trace5.txt 2>&1
THA_GRAPH=0 THA_CHAIN_Z=1,1,1,1,1 timeout 300 python tools/chain_trace.py --launch 5 > gpurun_out/d_trace5_z1.txt 2>&1
head -30 gpurun_out/d_trace5.txt

#!/bin/bash
# round 2, call B: the chained matmuls (chain_tc.cu): parity first, then timing
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_forward.py -m gpu -x -q -k "fp32 or modes or greedy" > gpurun_out/b_tests1.log 2>&1; echo "rc=$?" >> gpurun_out/b_tests1.log
tail -4 gpurun_out/b_tests1.log
timeout 900 python -m pytest tests/test_gpu_baseline_shapes.py -m gpu -q -s > gpurun_out/b_tests2.log 2>&1; echo "rc=$?" >> gpurun_out/b_tests2.log
tail -4 gpurun_out/b_tests2.log
for ch in 1 0; do
  THA_CHAIN=$ch timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/b_bench_chain$ch.json 2> gpurun_out/b_bench_chain$ch.err
  python -c "import json;d=json.load(open('gpurun_out/b_bench_chain$ch.json'));print('chain=$ch', d['value'], d['ms_per_step'], d['gpu_launches'])"
done

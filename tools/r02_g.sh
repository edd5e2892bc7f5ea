#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ops.py -m gpu -q -k "bf16" > gpurun_out/g_tests1.log 2>&1; echo "rc=$?" >> gpurun_out/g_tests1.log
tail -15 gpurun_out/g_tests1.log | cut -c1-200
timeout 900 python -m pytest tests/test_gpu_bf16.py -m gpu -q > gpurun_out/g_tests2.log 2>&1; echo "rc=$?" >> gpurun_out/g_tests2.log
grep -E "^(FAILED|PASSED)|passed|failed|ratio" gpurun_out/g_tests2.log | cut -c1-260 | head -60

#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_sampler.py tests/test_host_cpp.py -m gpu -q -s > gpurun_out/e_tests.log 2>&1; echo "rc=$?" >> gpurun_out/e_tests.log
grep -E "identical|passed|failed|rc=|Error" gpurun_out/e_tests.log | head -30

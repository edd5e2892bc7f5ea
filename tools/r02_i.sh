#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_paged_kv.py -m gpu -q > gpurun_out/i_tests.log 2>&1; echo "rc=$?" >> gpurun_out/i_tests.log
grep -E "passed|failed|Error|rc=" gpurun_out/i_tests.log | cut -c1-300 | head -30

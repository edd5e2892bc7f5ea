#!/bin/bash
# round 2, call A: the new parity tests at the BASELINE shapes + baseline timings of the round-1 kernels on this pool's GPU
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
nproc > gpurun_out/a_nproc.txt
timeout 1500 python -m pytest tests/test_gpu_baseline_shapes.py tests/test_host_cpp.py -m gpu -q -s > gpurun_out/a_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/a_tests.log
tail -5 gpurun_out/a_tests.log


#!/bin/bash
# the no-allocation entry point's test alone in a fresh process, then the whole round-end sequence (tools/r02_all.sh)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/noalloc_alone.log 2>&1; echo "rc=$?" >> gpurun_out/noalloc_alone.log
tail -15 gpurun_out/noalloc_alone.log | cut -c1-400
bash tools/r02_all.sh

#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
bash tools/r02_j.sh
bash tools/r02_k.sh 2>&1 | tail -9
timeout 600 python -m pytest tests/test_gpu_forward.py tests/test_gpu_bf16.py tests/test_gpu_paged_kv.py -m gpu -q -x 2>&1 | tail -3

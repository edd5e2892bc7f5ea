#!/bin/bash
# launch list of our kernels for one configuration: tools/gpu_ll.sh <model> <quant> <batch> <launches per step> <tag>
cd "$GRAFT_REPO_ROOT" || exit 1
CMD="python bench.py --model $1 --quant $2 --batch $3 --steps 4 --warmup 3 --no-cpu-baseline --no-e2e"
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"gemm_|tc_|attn_|argmax|advance|gemv|q8_prep|mega|rmsnorm|rope|swiglu" -s $((3 * $4 + 8)) -c $4 --csv --log-file gpurun_out/launches_$5.csv $CMD > gpurun_out/ncu_$5.log 2>&1
tail -1 gpurun_out/ncu_$5.log | cut -c1-120

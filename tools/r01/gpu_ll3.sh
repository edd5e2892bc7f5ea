#!/bin/bash
# launch list (times + DRAM bytes) of one decode step: tools/gpu_ll3.sh <tag> <launches per step> [bench args...]
cd "$GRAFT_REPO_ROOT" || exit 1
T=$1; LPS=$2; shift; shift
CMD="python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-e2e $*"
timeout 300 $CMD > gpurun_out/${T}_plain.json 2> gpurun_out/${T}_plain.err || { tail -5 gpurun_out/${T}_plain.err; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
  -k regex:"gemm_|tc_|attn_|argmax|advance|gemv|q8_prep|mega|rmsnorm|rope|swiglu|tp_" -s $((4 * LPS)) -c $LPS --csv \
  --log-file gpurun_out/${T}_launches.csv $CMD > gpurun_out/${T}_ncu.log 2>&1
tail -1 gpurun_out/${T}_ncu.log | cut -c1-200

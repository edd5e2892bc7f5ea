#!/bin/bash
# launch list (times + DRAM bytes) of ONE late decode step of the default bench: tools/gpu_ll2.sh <tag> [bench args...]
cd "$GRAFT_REPO_ROOT" || exit 1
T=$1; shift
CMD="python bench.py --steps 128 --warmup 4 --no-cpu-baseline --no-e2e $*"
$CMD > gpurun_out/${T}_plain.json 2> gpurun_out/${T}_plain.err || { tail -5 gpurun_out/${T}_plain.err; exit 1; }
cat gpurun_out/${T}_plain.json | cut -c1-400
LPS=${LPS:-112}
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
  -k regex:"gemm_|tc_|attn_|argmax|advance|gemv|q8_prep|mega|rmsnorm|rope|swiglu|tp_" -s $((100 * LPS)) -c $LPS --csv \
  --log-file gpurun_out/${T}_launches.csv $CMD > gpurun_out/${T}_ncu.log 2>&1
tail -2 gpurun_out/${T}_ncu.log | cut -c1-200

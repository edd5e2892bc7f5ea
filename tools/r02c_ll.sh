#!/bin/bash
# launch list (time + DRAM bytes) of the timed positions of the default bench as shipped at the end of round 2 (two concurrent
# half-batch branches: 201 launches per step), and the step timeline of the same configuration
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
K="gemm_|tc_|attn_|argmax|advance|gemv|q8_prep|mega|rmsnorm|rope|swiglu|tp_|chain_|sample_|bf_"
CMD="python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-extra --no-shares"
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
  -k regex:"$K" -s $((5 * 201)) -c $((20 * 201)) --csv --log-file gpurun_out/r02c_launches_s110m_f32_B256_pos5-24.csv $CMD > gpurun_out/r02c_ll.log 2>&1
tail -1 gpurun_out/r02c_ll.log | cut -c1-200
python tools/step_timeline.py --batch 256 --pos 10 --show 40 > gpurun_out/timeline_256_dual.txt 2>&1

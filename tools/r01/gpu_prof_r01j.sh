#!/bin/bash
# r01j: launch list (time + DRAM bytes) of one decode step of the default bench at mid-run context (position ~136), and one
# full capture of the dominant kernel there (attn_fused_kernel<1>)
cd "$GRAFT_REPO_ROOT" || exit 1
bash tools/gpu_ll3.sh r01j_s110m_f32_B256_pos132 113 --pos0 132
CMD="python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-e2e --pos0 132"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_fused -s 40 -c 2 -o gpurun_out/r01j_attn_fused_B256_pos132 -f $CMD > gpurun_out/r01j_ncu_full.log 2>&1
tail -2 gpurun_out/r01j_ncu_full.log | cut -c1-200
ls -la gpurun_out/ | grep r01j

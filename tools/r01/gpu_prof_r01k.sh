#!/bin/bash
# r01k (end of round 1): launch lists (time + DRAM bytes per launch) of one decode step of the default bench at mid-run
# context and of Llama-2-7B Q8_0 at 1 slot; one full capture of the dominant kernel of the default bench (attention).
cd "$GRAFT_REPO_ROOT" || exit 1
bash tools/gpu_ll3.sh r01k_s110m_f32_B256_pos132 113 --pos0 132
bash tools/gpu_ll3.sh r01k_7b_q8_B1 292 --model 7b --quant q8 --batch 1 --seq-len 512
CMD="python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-e2e --pos0 132"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_fused -s 40 -c 2 -o gpurun_out/r01k_attn_fused_B256_pos132 -f $CMD > gpurun_out/r01k_ncu_full.log 2>&1
tail -1 gpurun_out/r01k_ncu_full.log | cut -c1-200

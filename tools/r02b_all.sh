#!/bin/bash
# the round-end sequence (tests, smoke, bench) plus what the second-pass profiles still need (Q8_0 64-slot timeline and capture)
cd "$GRAFT_REPO_ROOT" || exit 1
bash tools/r02_all.sh
python tools/step_timeline.py --model 7b --quant q8 --seq-len 512 --batch 64 --pos 10 --show 16 > gpurun_out/timeline_7bq8_64.txt 2>&1
B="--no-cpu-baseline --no-e2e --no-extra --no-shares"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_q8i -s 200 -c 4 -o gpurun_out/r02b_gemm_q8i_7b_B64 -f python bench.py --steps 4 --warmup 3 $B --model 7b --quant q8 --batch 64 --seq-len 512 > gpurun_out/r02b_full3.log 2>&1; tail -1 gpurun_out/r02b_full3.log | cut -c1-160
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 2>/dev/null | tail -1 | cut -c1-400
# the bf16 prompt pass: full capture of its matmuls and attention, launch list (time + DRAM bytes) of one layer and a bit
P="python tools/bench_prefill.py --model 7b --len 2048 --reps 1 --bf16 1"
timeout 300 $P 2>&1 | tail -1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"gemm_bf16|attn_prefill_tc" -s 170 -c 6 -o gpurun_out/r02b_prefill_7b_bf16_2048 -f $P > gpurun_out/r02b_full4.log 2>&1; tail -1 gpurun_out/r02b_full4.log | cut -c1-160
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"gemm_|tc_|attn_|argmax|bf_|rope" -s 340 -c 44 --csv --log-file gpurun_out/r02b_launches_prefill_7b_bf16_2048.csv $P > gpurun_out/pf_ll.log 2>&1; tail -1 gpurun_out/pf_ll.log
